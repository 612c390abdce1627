"""FASTA ingestion on the device (pfbwt-f_b200/csrc/ingest.cuh): record splitting, header
skipping, line ends and the A^w padding must reproduce kseq as driven by PfParser::add_fasta
(include/kseq.h:178-216, include/pfparser.hpp:299-337) on awkward files.

CPU part (-m "not gpu"): the logic emulator runs the three ingestion kernels as serial loops; its files must equal
the reference's and those of the host reader (PFB_INGEST=host).  GPU part: the real binaries, incl. --print-docs."""
import os
import subprocess
import pytest
from pfp_util import ROOT, random_bases, run, exts_for, assert_same_files

EMU = os.path.join(ROOT, "tests", "emu", "pfb_emu")
BIN = os.path.join(ROOT, "bin")


def wrap(seq, width):
    return b"\n".join(seq[i:i + width] for i in range(0, len(seq), width)) if len(seq) else b""


def case_files():
    a, b, c, d = (random_bases(n, s).tobytes() for n, s in ((40_000, 1), (25_000, 2), (7, 3), (30_001, 4)))
    low = bytes(ch + 32 if i % 7 < 3 else ch for i, ch in enumerate(d))          # soft-masked stretches
    withn = d[:9000] + b"N" * 700 + d[9700:20000] + b"n" * 3 + d[20003:]
    cases = {}
    # comments in headers, blank lines inside and between records, no newline at the end of the file
    cases["blank_lines_no_eol"] = b">r1 some comment\tmore\n" + wrap(a, 60) + b"\n\n\n>r2\n\n" + wrap(b, 61) + b"\n>r3 x\n" + wrap(c, 3)
    # text before the first header (no '>' or '@' in it), also with line ends
    cases["garbage_first"] = b"this is not fasta\nACGT\n\n  >h1 c\n" + wrap(a, 80) + b"\n>h2\n" + wrap(d, 80) + b"\n"
    # records without sequence: header directly followed by a header, and a header as the last line without newline
    cases["empty_records"] = b">e0\n>r1\n" + wrap(a, 70) + b"\n>e1 c\n>e2\n\n>r2\n" + wrap(b, 70) + b"\n>last"
    # line lengths around the 64-byte tile of the kernels, one very long line, one-character lines
    cases["tile_edges"] = (b">t63\n" + wrap(a, 63) + b"\n>t64\n" + wrap(b, 64) + b"\n>t65\n" + wrap(d, 65) + b"\n>long\n" + a + b"\n>one\n" + wrap(a[:300], 1) + b"\n")
    # 62 + newline + header start exactly at tile boundaries
    cases["aligned"] = b">" + b"x" * 62 + b"\n" + wrap(a, 63) + b"\n" + b">" + b"y" * 30 + b"\n" + wrap(b, 127) + b"\n"
    # a header marker as the very last byte of the file: kseq's name read hits EOF, no record (kseq.h:188)
    cases["lone_marker_at_eof"] = b">r1 c\n" + wrap(a, 60) + b"\n>r2\n" + wrap(b, 60) + b"\n>"
    cases["lower_and_n"] = b">m1\n" + wrap(low, 60) + b"\n>m2 has N\n" + wrap(withn, 60) + b"\n"
    # handled by the host reader (detected on the device): CR LF line ends, FASTQ
    cases["crlf"] = b">c1 x\r\n" + wrap(a, 60).replace(b"\n", b"\r\n") + b"\r\n>c2\r\n" + wrap(b, 60).replace(b"\n", b"\r\n") + b"\r\n"
    cases["fastq"] = b"@q1\n" + a[:20000] + b"\n+\n" + b"I" * 20000 + b"\n@q2 c\n" + b[:15000] + b"\n+q2\n" + b"#" * 15000 + b"\n"
    # FASTQ corners (kseq.h:205-213): CR LF line ends with multi-line quality, quality characters that look like markers,
    # a truncated quality string (the record and everything behind it are dropped), EOF inside the '+' line
    cases["fastq_crlf_multiline"] = (b"@q1\r\n" + wrap(a[:3000], 70).replace(b"\n", b"\r\n") + b"\r\n+\r\n" +
                                     wrap(b"@>I#" * 750, 70).replace(b"\n", b"\r\n") + b"\r\n@q2\r\n" + b[:500] + b"\r\n+\r\n" + b"I" * 500 + b"\r\n")
    cases["fastq_truncated_qual"] = b"@q1\n" + a[:2000] + b"\n+\n" + b"I" * 2000 + b"\n@q2\n" + b[:900] + b"\n+\n" + b"I" * 400 + b"\n"
    cases["fastq_eof_in_plus"] = b"@q1\n" + a[:2000] + b"\n+\n" + b"I" * 2000 + b"\n@q2\n" + b[:900] + b"\n+"
    return cases


CASES = case_files()
BAD = {"gt_inside_line": b">b1\n" + random_bases(5000, 9).tobytes() + b">notaheader\n" + b"ACGT" * 1000 + b"\n"}


def write(datadir, name, data):
    p = os.path.join(datadir, "ing_" + name + ".fa")
    with open(p, "wb") as f:
        f.write(data)
    return p


@pytest.fixture(scope="module")
def emu():
    subprocess.run(["make", "-C", ROOT, "emu"], check=True, stdout=subprocess.DEVNULL)
    return EMU


@pytest.mark.parametrize("name", sorted(CASES))
def test_ingest_logic_matches_reference_and_host_reader(name, emu, ref_bins, datadir):
    fa = write(datadir, name, CASES[name])
    flags = ["-s", "-r", "--print-docs"]
    dev, host = os.path.join(datadir, name + ".idev"), os.path.join(datadir, name + ".ihost")
    r = subprocess.run([emu, "--bits", "32"] + flags + ["-o", dev, fa], check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    took = b"ingest: host" if name == "crlf" or name.startswith("fastq") else b"ingest: device"     # which reader the file must have taken
    assert took in r.stderr, r.stderr.decode()[-500:]
    env = dict(os.environ, PFB_INGEST="host")
    subprocess.run([emu, "--bits", "32"] + flags + ["-o", host, fa], check=True, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    assert_same_files(host, dev, exts_for(True, True) + ["docs"])
    if ref_bins["pfbwt-f"]:
        ref = os.path.join(datadir, name + ".iref")
        run([ref_bins["pfbwt-f"]] + flags + ["-o", ref, fa])
        assert_same_files(ref, dev, exts_for(True, True) + ["docs"])


@pytest.mark.parametrize("name", ["blank_lines_no_eol", "tile_edges", "garbage_first", "crlf", "fastq"])
@pytest.mark.parametrize("how", ["gz", "stdin", "gz_stdin"])
def test_ingest_logic_streams(name, how, emu, ref_bins, datadir):
    """gzip files and stdin: zlib inflates / drains the stream on the host, the records are split by the same kernels;
    a stream that turns out to need the host reader (FASTQ, CR LF) is handed back to it from memory"""
    import gzip
    data = CASES[name]
    if "gz" in how:
        data = gzip.compress(data)
    fa = write(datadir, name + "_" + how, data) + (".gz" if how == "gz" else "")
    if how == "gz":
        os.replace(fa[:-3], fa)
    src, stdin = (fa, None) if how == "gz" else ("-", data)
    dev = os.path.join(datadir, f"{name}_{how}.sdev")
    r = subprocess.run([emu, "--bits", "32", "-s", "-r", "-o", dev, src], input=stdin, check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    took = b"ingest: host" if name == "crlf" or name.startswith("fastq") else b"ingest: device"
    assert took in r.stderr, r.stderr.decode()[-500:]
    if ref_bins["pfbwt-f"]:
        ref = os.path.join(datadir, f"{name}_{how}.sref")
        subprocess.run([ref_bins["pfbwt-f"], "-s", "-r", "-o", ref, src], input=stdin, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        assert_same_files(ref, dev, exts_for(True, True))
    else:
        host = os.path.join(datadir, f"{name}_{how}.shost")
        subprocess.run([emu, "--bits", "32", "-s", "-r", "-o", host, src], input=stdin, check=True, env=dict(os.environ, PFB_INGEST="host"),
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        assert_same_files(host, dev, exts_for(True, True))


@pytest.mark.parametrize("chunk", ["1", "7", "64"])
def test_host_reader_across_chunk_ends(chunk, emu, ref_bins, datadir):
    """the host reader (FASTQ, CR LF, and the checker of the device reader) is a state machine fed in chunks: with chunks of
    1, 7 and 64 bytes (PFB_READ_CHUNK) every state is cut at a chunk end somewhere in these files"""
    if not ref_bins["pfbwt-f"]:
        pytest.skip("reference binaries not built")
    for name in ("blank_lines_no_eol", "empty_records", "crlf", "fastq", "garbage_first"):
        fa = write(datadir, f"hc_{name}", CASES[name])
        ref, host = os.path.join(datadir, f"{name}.hcref"), os.path.join(datadir, f"{name}.hc{chunk}")
        run([ref_bins["pfbwt-f"], "-s", "-o", ref, fa])
        subprocess.run([emu, "--bits", "32", "-s", "-o", host, fa], check=True, env=dict(os.environ, PFB_INGEST="host", PFB_READ_CHUNK=chunk),
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        assert_same_files(ref, host, exts_for(True, False))


def random_fasta(seed):
    """a structurally random FASTA: record count, header text, line widths (also per line), blank lines, missing last newline,
    text before the first header, lower case / N"""
    import numpy as np
    rng = np.random.default_rng(seed)
    out = bytearray()
    if rng.random() < 0.3:
        out += bytes(rng.choice(list(b"acgt xyz\n"), int(rng.integers(1, 200))).astype(np.uint8))
    nrec = int(rng.integers(1, 6))
    alphabet = np.frombuffer(b"ACGTACGTACGTACGTacgtN", dtype=np.uint8)
    for r in range(nrec):
        out += b">rec%d" % r
        if rng.random() < 0.5:
            out += b" " + bytes(rng.choice(list(b"abc \t>@+"), int(rng.integers(0, 90))).astype(np.uint8))
        out += b"\n"
        n = int(rng.integers(0, 30_000)) if rng.random() < 0.9 else 0
        seq = alphabet[rng.integers(0, len(alphabet), n)].tobytes()
        fixed = int(rng.integers(1, 200))
        pos = 0
        while pos < n:
            wdt = fixed if rng.random() < 0.8 else int(rng.integers(1, 300))
            out += seq[pos:pos + wdt] + b"\n"
            pos += wdt
            if rng.random() < 0.02:
                out += b"\n" * int(rng.integers(1, 4))
    if rng.random() < 0.3 and out.endswith(b"\n"):
        out = out[:-1]
    return bytes(out)


@pytest.mark.parametrize("block", range(4))
def test_ingest_logic_random_structures(block, emu, datadir):
    """device reader logic == host reader on 40 structurally random files (-w 4 -p 11 so that short inputs still parse);
    when the host route refuses an input (too few phrases) the device route must refuse it with the same message"""
    for seed in range(block * 10, block * 10 + 10):
        fa = write(datadir, f"rnd{seed}", random_fasta(1000 + seed))
        res = {}
        for mode in ("device", "host"):
            env = dict(os.environ)
            if mode == "host":
                env["PFB_INGEST"] = "host"
            o = os.path.join(datadir, f"rnd{seed}.{mode}")
            r = subprocess.run([emu, "--bits", "64", "-s", "-r", "--print-docs", "--non-acgt-to-a", "-w", "4", "-p", "11", "-o", o, fa],
                               env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            res[mode] = (r.returncode, o, r.stderr)
        assert res["device"][0] == res["host"][0], (seed, res["device"][2][-300:], res["host"][2][-300:])
        if res["host"][0] == 0:
            assert b"ingest: device" in res["device"][2], seed
            assert_same_files(res["host"][1], res["device"][1], exts_for(True, True) + ["docs"])
        else:
            assert res["device"][2].splitlines()[-1:] == res["host"][2].splitlines()[-1:], seed


def test_ingest_logic_several_files_mixed_readers(emu, ref_bins, datadir):
    """add_fasta once per file: a plain file (device reader), a CR LF file (host reader) and a gzip file (device reader after
    inflate) in one context give the parse of the concatenated record list"""
    import gzip
    if not ref_bins["pfbwt-f"]:
        pytest.skip("reference binaries not built")
    f1 = write(datadir, "mix1", CASES["tile_edges"])
    f2 = write(datadir, "mix2", CASES["crlf"])
    f3 = write(datadir, "mix3", gzip.compress(CASES["empty_records"] + b"\n")) + ".gz"
    os.replace(f3[:-3], f3)
    whole = write(datadir, "mix_all", CASES["tile_edges"] + CASES["crlf"] + CASES["empty_records"] + b"\n")
    ref, dev = os.path.join(datadir, "mix.ref"), os.path.join(datadir, "mix.dev")
    run([ref_bins["pfbwt-f"], "-s", "-r", "-o", ref, whole])
    r = subprocess.run([emu, "--bits", "32", "-s", "-r", "-o", dev, f1, f2, f3], check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert b"ingest: mixed" in r.stderr
    assert_same_files(ref, dev, exts_for(True, True))


def test_ingest_logic_invalid_character(emu, datadir):
    """a '>' that is not the first character of a line is sequence: the reference dies in the hasher (include/hash.hpp:31)"""
    fa = write(datadir, "gt_inside", BAD["gt_inside_line"])
    r = subprocess.run([emu, "--bits", "32", "-o", os.path.join(datadir, "gt_inside.out"), fa], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode != 0 and b"invalid character" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_ingest_gpu_cli(name, ref_bins, datadir):
    if not ref_bins["pfbwt-f"]:
        pytest.skip("reference binaries not built")
    fa = write(datadir, name, CASES[name])
    a, b = os.path.join(datadir, name + ".gref"), os.path.join(datadir, name + ".ggpu")
    run([ref_bins["pfbwt-f"], "-s", "-r", "--print-docs", "-o", a, fa])
    run([os.path.join(BIN, "pfbwt-f"), "-s", "-r", "--print-docs", "-o", b, fa])
    assert_same_files(a, b, exts_for(True, True) + ["docs"])


@pytest.mark.gpu
@pytest.mark.parametrize("how", ["gz", "stdin"])
def test_ingest_gpu_streams(how, ref_bins, datadir):
    import gzip
    if not ref_bins["pfbwt-f"]:
        pytest.skip("reference binaries not built")
    for name in ("empty_records", "tile_edges", "fastq"):
        data = CASES[name]
        fa = write(datadir, f"g_{name}_{how}", gzip.compress(data) if how == "gz" else data)
        src, stdin = (fa, None) if how == "gz" else ("-", data)
        a, b = os.path.join(datadir, f"{name}_{how}.gsref"), os.path.join(datadir, f"{name}_{how}.gsgpu")
        subprocess.run([ref_bins["pfbwt-f"], "-s", "-r", "--print-docs", "-o", a, src], input=stdin, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        subprocess.run([os.path.join(BIN, "pfbwt-f"), "-s", "-r", "--print-docs", "-o", b, src], input=stdin, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        assert_same_files(a, b, exts_for(True, True) + ["docs"])


@pytest.mark.gpu
def test_ingest_gpu_multi_file_and_host_reader_agree(ref_bins, datadir):
    """64-bit build, larger file (several 32 MB read chunks are not needed for the logic, but many tiles are):
    device reader == host reader == reference"""
    recs = [(f"rec{i} len", random_bases(700_000 + 13 * i, 100 + i)) for i in range(4)]
    data = b"".join(b">" + n.encode() + b"\n" + wrap(s.tobytes(), 60) + b"\n" for n, s in recs)
    fa = write(datadir, "big", data)
    ours = os.path.join(BIN, "pfbwt-f64")
    b, c = os.path.join(datadir, "big.gdev"), os.path.join(datadir, "big.ghost")
    run([ours, "-s", "--print-docs", "-o", b, fa])
    subprocess.run([ours, "-s", "--print-docs", "-o", c, fa], check=True, env=dict(os.environ, PFB_INGEST="host"),
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    assert_same_files(c, b, exts_for(True, False) + ["docs"])
    if ref_bins["pfbwt-f64"]:
        a = os.path.join(datadir, "big.gref")
        run([ref_bins["pfbwt-f64"], "-s", "--print-docs", "-o", a, fa])
        assert_same_files(a, b, exts_for(True, False) + ["docs"])


@pytest.mark.gpu
def test_ingest_gpu_invalid_character(datadir):
    fa = write(datadir, "gt_inside_gpu", BAD["gt_inside_line"])
    r = subprocess.run([os.path.join(BIN, "pfbwt-f"), "-o", os.path.join(datadir, "gt_inside_gpu.out"), fa], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 1 and b"invalid character" in r.stderr
