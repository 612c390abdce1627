#!/usr/bin/env python
"""Regenerates tests/golden/ (run in the build container, where /root/reference exists).

1. Copies the reference's own known-answer fixtures for the hot path
   (/root/reference/tests/data/{single_chrom,mult_chroms}.{fa,bwt,sa}); the text .sa files
   (one decimal per line, scripts/generate_truth_set.py:92-100) are stored gzip-compressed.
2. Runs the UNMODIFIED reference binaries (oracle/_ref, built by oracle/Makefile from the
   sources where they lie) on seeded synthetic inputs and records sha256[:16] of every file
   they write in ref_hashes.json, so that parity can be checked where the reference is absent.
"""
import gzip
import json
import os
import shutil
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from pfp_util import (REF_DIR, write_fasta, random_bases, pangenome_records, nrun_records, run, sha16, exts_for)

REF_DATA = "/root/reference/tests/data"

CASES = {
    # name: (generator spec, bits, flags)
    "single_chrom": (("file", "single_chrom.fa"), 32, ["-s", "-r"]),
    "mult_chroms": (("file", "mult_chroms.fa"), 32, ["-s", "-r"]),
    "rand300k": (("rand", 300_000, 12345), 32, ["-s", "-r"]),
    "rand300k_64": (("rand", 300_000, 12345), 64, ["-s"]),
    "pan6x100k": (("pan", 100_000, 6, 100, 2024), 64, ["-s", "-r"]),
    "pan6x100k_nosa": (("pan", 100_000, 6, 100, 2024), 32, []),
    "nrun": (("nrun", 3, 60_000, 38), 64, ["-s", "-r", "--non-acgt-to-a"]),
    "w5p13": (("pan", 50_000, 3, 30, 5), 32, ["-s", "-r", "-w", "5", "-p", "13"]),
}


def make_input(spec, path):
    if spec[0] == "file":
        shutil.copy(os.path.join(HERE, spec[1]), path)
    elif spec[0] == "rand":
        write_fasta(path, [("rand", random_bases(spec[1], spec[2]))], line_width=80)
    elif spec[0] == "pan":
        write_fasta(path, pangenome_records(spec[1], spec[2], spec[3], spec[4]), line_width=60)
    elif spec[0] == "nrun":
        write_fasta(path, nrun_records(spec[1], spec[2], spec[3]), line_width=60)


def main():
    for stem in ["single_chrom", "mult_chroms"]:
        shutil.copy(os.path.join(REF_DATA, stem + ".fa"), os.path.join(HERE, stem + ".fa"))
        shutil.copy(os.path.join(REF_DATA, stem + ".bwt"), os.path.join(HERE, stem + ".bwt"))
        with open(os.path.join(REF_DATA, stem + ".sa"), "rb") as f, \
                gzip.GzipFile(os.path.join(HERE, stem + ".sa.gz"), "wb", mtime=0) as g:
            g.write(f.read())
        for e in ["fa", "bwt"]:
            os.chmod(os.path.join(HERE, stem + "." + e), 0o644)
    hashes = {}
    with tempfile.TemporaryDirectory() as td:
        for name, (spec, bits, flags) in CASES.items():
            fa = os.path.join(td, name + ".fa")
            make_input(spec, fa)
            exe = os.path.join(REF_DIR, "pfbwt-f64" if bits == 64 else "pfbwt-f")
            out = os.path.join(td, name + ".out")
            r = run([exe] + flags + ["-o", out, fa])
            log = r.stderr.decode()
            h = {e: sha16(f"{out}.{e}") for e in exts_for("-s" in flags, "-r" in flags)}
            h["input"] = sha16(fa)
            for line in log.splitlines():
                if line.startswith("# easy cases"):
                    h["easy_hard"] = line.strip()
                if line.startswith("r: "):
                    h["r"] = int(line[3:])
            hashes[name] = {"spec": list(spec), "bits": bits, "flags": flags, "files": h}
    with open(os.path.join(HERE, "ref_hashes.json"), "w") as f:
        json.dump(hashes, f, indent=1, sort_keys=True)
    print("wrote", len(hashes), "cases")


if __name__ == "__main__":
    main()
