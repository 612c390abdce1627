"""CPU tests (-m "not gpu"): pin the oracle.
  * pfp_oracle (oracle/pfp_oracle.c) reproduces the hashes the UNMODIFIED reference produced
    (tests/golden/ref_hashes.json, made by tests/golden/make_golden.py) for every file;
  * where oracle/_ref holds the reference binaries, they are run live and compared byte for byte;
  * both reproduce the reference's own golden vectors tests/data/{single_chrom,mult_chroms}.{bwt,sa}
    (the FASTA is recovered by inverting the golden);
  * the merge restatement equals reference merge_pfp and the monolithic parse;
  * a PFP-independent numpy suffix array agrees.
"""
import gzip
import json
import os
import numpy as np
import pytest
from pfp_util import (GOLDEN, write_fasta, random_bases, pangenome_records, nrun_records, run, sha16, exts_for,
                      assert_same_files, invert_bwt_sa, naive_sa_bwt)

HASHES = json.load(open(os.path.join(GOLDEN, "ref_hashes.json")))


def make_input(spec, path):
    if spec[0] == "file":
        with open(os.path.join(GOLDEN, spec[1]), "rb") as f, open(path, "wb") as g:
            g.write(f.read())
    elif spec[0] == "rand":
        write_fasta(path, [("rand", random_bases(spec[1], spec[2]))], line_width=80)
    elif spec[0] == "pan":
        write_fasta(path, pangenome_records(spec[1], spec[2], spec[3], spec[4]), line_width=60)
    elif spec[0] == "nrun":
        write_fasta(path, nrun_records(spec[1], spec[2], spec[3]), line_width=60)


@pytest.mark.parametrize("name", sorted(HASHES))
def test_port_matches_reference_hashes(name, ref_bins, datadir):
    case = HASHES[name]
    fa = os.path.join(datadir, name + ".fa")
    make_input(case["spec"], fa)
    assert sha16(fa) == case["files"]["input"], "generator drifted; regenerate goldens"
    out = os.path.join(datadir, name + ".port")
    r = run([ref_bins["pfp_oracle"], "--bits", str(case["bits"])] + case["flags"] + ["-o", out, fa])
    for e in exts_for("-s" in case["flags"], "-r" in case["flags"]):
        assert sha16(f"{out}.{e}") == case["files"][e], f".{e} differs from the reference"
    log = r.stderr.decode()
    assert case["files"]["easy_hard"] in log
    assert f"r: {case['files']['r']}" in log


@pytest.mark.parametrize("name", ["rand300k", "pan6x100k", "nrun"])
def test_port_matches_reference_live(name, ref_bins, datadir):
    case = HASHES[name]
    exe = ref_bins["pfbwt-f64" if case["bits"] == 64 else "pfbwt-f"]
    if not exe:
        pytest.skip("reference binaries not built (oracle/_ref)")
    fa = os.path.join(datadir, name + ".live.fa")
    make_input(case["spec"], fa)
    a, b = os.path.join(datadir, name + ".liveref"), os.path.join(datadir, name + ".liveport")
    run([exe] + case["flags"] + ["-o", a, fa])
    run([ref_bins["pfp_oracle"], "--bits", str(case["bits"])] + case["flags"] + ["-o", b, fa])
    assert_same_files(a, b, exts_for("-s" in case["flags"], "-r" in case["flags"]))


@pytest.mark.parametrize("stem,reclen", [("single_chrom", 10010), ("mult_chroms", 10010)])
def test_reference_golden_vectors(stem, reclen, ref_bins, datadir):
    """tests/data/<stem>.{bwt,sa}: independent (non-PFP) truth shipped with the reference."""
    bwt = np.frombuffer(open(os.path.join(GOLDEN, stem + ".bwt"), "rb").read(), dtype=np.uint8)
    sa = np.array(gzip.open(os.path.join(GOLDEN, stem + ".sa.gz"), "rt").read().split(), dtype=np.uint64)
    n = len(bwt) - 1
    assert len(sa) == n + 1 and sa[0] == n
    S = invert_bwt_sa(bwt, sa)
    assert n % reclen == 0
    recs = []
    for i in range(n // reclen):
        rec = S[i * reclen:(i + 1) * reclen]
        assert bytes(rec[-10:]) == b"A" * 10
        recs.append((f"s{i}", rec[:-10]))
    fa = os.path.join(datadir, stem + ".inv.fa")
    write_fasta(fa, recs, line_width=60)
    out = os.path.join(datadir, stem + ".inv.port")
    run([ref_bins["pfp_oracle"], "--bits", "64", "-s", "-o", out, fa])
    assert open(out + ".bwt", "rb").read() == bwt.tobytes()
    got = np.fromfile(out + ".sa", dtype=np.uint64)
    assert np.array_equal(got, sa)
    if ref_bins["pfbwt-f64"]:
        out2 = os.path.join(datadir, stem + ".inv.ref")
        run([ref_bins["pfbwt-f64"], "-s", "-o", out2, fa])
        assert open(out2 + ".bwt", "rb").read() == bwt.tobytes()
        assert np.array_equal(np.fromfile(out2 + ".sa", dtype=np.uint64), sa)


def test_port_against_naive_suffix_array(ref_bins, datadir):
    recs = pangenome_records(20_000, 4, 40, 99)
    fa = os.path.join(datadir, "naive.fa")
    write_fasta(fa, recs)
    out = os.path.join(datadir, "naive.port")
    run([ref_bins["pfp_oracle"], "-s", "-o", out, fa])
    S = np.concatenate([np.concatenate([r[1], np.frombuffer(b"A" * 10, dtype=np.uint8)]) for r in recs])
    sa, bwt = naive_sa_bwt(S)
    assert np.array_equal(np.fromfile(out + ".sa", dtype=np.uint32).astype(np.uint64), sa)
    assert open(out + ".bwt", "rb").read() == bwt.tobytes()


def test_merge_restatement(ref_bins, datadir):
    recs = pangenome_records(80_000, 8, 80, 31)
    parts = []
    for g in range(4):
        p = os.path.join(datadir, f"opart{g}.fa")
        write_fasta(p, recs[2 * g:2 * g + 2], line_width=60)
        run([ref_bins["pfp_oracle"], "--bits", "64", "--parse-only", "-s", p])
        parts.append(p)
    allfa = os.path.join(datadir, "oall.fa")
    write_fasta(allfa, recs, line_width=60)
    mono = os.path.join(datadir, "omono")
    run([ref_bins["pfp_oracle"], "--bits", "64", "--parse-only", "-s", "-o", mono, allfa])
    merged = os.path.join(datadir, "omerged")
    run([ref_bins["pfp_oracle"], "--merge", "--bits", "64", "-s", "-o", merged] + parts)
    exts = ["dict", "occ", "parse", "n", "bwlast", "ilist", "bwsai"]
    assert_same_files(mono, merged, exts)
    if ref_bins["merge_pfp"]:
        rm = os.path.join(datadir, "orefmerged")
        run([ref_bins["merge_pfp"], "-s", "--parse-bwt", "-o", rm, "-t", "1"] + parts)
        assert_same_files(rm, merged, exts)
