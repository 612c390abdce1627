"""GPU parity through the drop-in binaries (bin/pfbwt-f, bin/pfbwt-f64, bin/merge_pfp, all thin
drivers over the C ABI) against the oracle: every file the reference writes must be byte-identical.
Oracle = unmodified reference binaries (oracle/_ref) when present, else the CPU restatement."""
import os
import pytest
import numpy as np
from pfp_util import (ROOT, write_fasta, random_bases, pangenome_records, nrun_records, run, exts_for,
                      assert_same_files, GOLDEN)

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "bin")


def oracle_cmd(ref_bins, bits, args, out, fa):
    name = "pfbwt-f64" if bits == 64 else "pfbwt-f"
    if ref_bins[name]:
        return [ref_bins[name]] + args + ["-o", out, fa]
    return [ref_bins["pfp_oracle"], "--bits", str(bits)] + args + ["-o", out, fa]


CASES = [
    # name, generator, bits, flags
    ("rand200k_sr", lambda: [("r", random_bases(200_000, 11))], 32, ["-s", "-r"]),
    ("rand1m", lambda: [("r", random_bases(1_000_000, 1))], 32, []),
    ("rand1m_s64", lambda: [("r", random_bases(1_000_000, 1))], 64, ["-s"]),
    ("pan8x200k", lambda: pangenome_records(200_000, 8, 200, 7), 64, ["-s", "-r"]),
    ("pan8x200k_nosa", lambda: pangenome_records(200_000, 8, 200, 7), 32, []),
    ("nrun", lambda: nrun_records(3, 100_000, 38), 64, ["-s", "-r", "--non-acgt-to-a"]),
    ("nrun_keepN", lambda: nrun_records(3, 100_000, 38), 32, ["-s"]),
    ("w4p7", lambda: [("a", random_bases(50_000, 3)), ("b", random_bases(1, 4)), ("c", random_bases(30_000, 5))], 32,
     ["-s", "-r", "-w", "4", "-p", "7"]),
    ("w20p300", lambda: pangenome_records(300_000, 4, 100, 9), 64, ["-s", "-w", "20", "-p", "300"]),
]


@pytest.mark.parametrize("name,gen,bits,flags", CASES, ids=[c[0] for c in CASES])
def test_cli_matches_oracle(name, gen, bits, flags, ref_bins, datadir):
    fa = os.path.join(datadir, name + ".fa")
    write_fasta(fa, gen(), line_width=70)
    a, b = os.path.join(datadir, name + ".ref"), os.path.join(datadir, name + ".gpu")
    run(oracle_cmd(ref_bins, bits, flags, a, fa))
    run([os.path.join(BIN, "pfbwt-f64" if bits == 64 else "pfbwt-f")] + flags + ["-o", b, fa])
    assert_same_files(a, b, exts_for("-s" in flags, "-r" in flags))


@pytest.mark.parametrize("fa", ["single_chrom.fa", "mult_chroms.fa"])
def test_cli_reference_fixtures(fa, ref_bins, datadir):
    """config 1: the reference's own tests/data FASTA, pfbwt-f -w 10 -p 100 -s -r"""
    src = os.path.join(GOLDEN, fa)
    a, b = os.path.join(datadir, fa + ".ref"), os.path.join(datadir, fa + ".gpu")
    run(oracle_cmd(ref_bins, 32, ["-s", "-r"], a, src))
    run([os.path.join(BIN, "pfbwt-f"), "-s", "-r", "-o", b, src])
    assert_same_files(a, b, exts_for(True, True))


def test_cli_merge_matches_monolithic(ref_bins, datadir):
    """config 5 (scaled): per-part --parse-only, merge_pfp -s --parse-bwt, --pfbwt-only == monolithic"""
    recs = pangenome_records(150_000, 8, 150, 21)
    parts = []
    for g in range(4):
        p = os.path.join(datadir, f"mpart{g}.fa")
        write_fasta(p, recs[2 * g:2 * g + 2], line_width=60)
        run([os.path.join(BIN, "pfbwt-f64"), "--parse-only", "-s", p])
        parts.append(p)
    allfa = os.path.join(datadir, "mall.fa")
    write_fasta(allfa, recs, line_width=60)
    mono = os.path.join(datadir, "mono")
    run(oracle_cmd(ref_bins, 64, ["-s"], mono, allfa))
    merged = os.path.join(datadir, "merged")
    run([os.path.join(BIN, "merge_pfp"), "-s", "--parse-bwt", "-o", merged, "-t", "4"] + parts)
    assert_same_files(mono, merged, ["dict", "occ", "parse", "n", "bwlast", "ilist", "bwsai"])
    run([os.path.join(BIN, "pfbwt-f64"), "--pfbwt-only", "-s", "-o", merged])
    assert_same_files(mono, merged, ["bwt", "sa"])


def test_cli_invalid_char_exits_1(datadir):
    fa = os.path.join(datadir, "bad.fa")
    with open(fa, "w") as f:
        f.write(">x\nACGTRACGT\n")
    r = run([os.path.join(BIN, "pfbwt-f"), fa], check=False)
    assert r.returncode == 1
    assert b"error, invalid character 82/R -> 5" in r.stderr


@pytest.mark.parametrize("env", [{"PFB_PIVOT": "1"}, {"PFB_PIVOT": "1", "PFB_KEY_SYMS": "3"}, {"PFB_PIVOT": "0"}],
                         ids=["pivot", "pivot_keysyms3", "nopivot"])
def test_cli_pivot_rounds(env, ref_bins, datadir):
    """tie refinement by pivot rounds (build.cuh): forced from the first round / on nearly all suffixes / off"""
    fa = os.path.join(datadir, "pivot.fa")
    recs = pangenome_records(200_000, 10, 600, 19) + nrun_records(1, 60_000, 12, big_run=15000)
    write_fasta(fa, recs, line_width=70)
    flags = ["-s", "-r", "--non-acgt-to-a"]
    a, b = os.path.join(datadir, "pivot.ref"), os.path.join(datadir, "pivot.gpu" + "".join(env.values()))
    if not os.path.exists(a + ".bwt"):
        run(oracle_cmd(ref_bins, 64, flags, a, fa))
    import subprocess
    r = subprocess.run([os.path.join(BIN, "pfbwt-f64")] + flags + ["-o", b, fa], env=dict(os.environ, **env),
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    assert_same_files(a, b, exts_for(True, True))


def test_cli_surface_gzip_stdin_stdout_docs(ref_bins, datadir):
    """the flags vcf_to_bwt.py drives the binaries with (vcf_to_bwt.py:118-129,174-181) - gzip input,
    stdin `-` with -o, --stdout bwt|sa, --print-docs (<o>.docs), --parse-only then --pfbwt-only - against the reference."""
    import gzip, subprocess
    if not ref_bins["pfbwt-f"]:
        pytest.skip("reference binaries not built")
    recs = pangenome_records(90_000, 4, 90, 55) + [("last one", random_bases(12_345, 56))]
    fa = os.path.join(datadir, "surf.fa")
    write_fasta(fa, recs, line_width=60)
    raw = open(fa, "rb").read()
    gz = fa + ".gz"
    with gzip.open(gz, "wb") as f:
        f.write(raw)
    ours, ref = os.path.join(BIN, "pfbwt-f"), ref_bins["pfbwt-f"]
    # gzip input + --print-docs
    a, b = os.path.join(datadir, "surf_gz.ref"), os.path.join(datadir, "surf_gz.gpu")
    run([ref, "-s", "--print-docs", "-o", a, gz])
    run([ours, "-s", "--print-docs", "-o", b, gz])
    assert_same_files(a, b, exts_for(True, False) + ["docs"])
    # stdin
    a2, b2 = os.path.join(datadir, "surf_in.ref"), os.path.join(datadir, "surf_in.gpu")
    run([ref, "-r", "-o", a2, "-"], stdin=raw)
    run([ours, "-r", "-o", b2, "-"], stdin=raw)
    assert_same_files(a2, b2, exts_for(False, True))
    # --stdout bwt / sa
    for what, flags in (("bwt", []), ("sa", ["-s"])):
        ra = subprocess.run([ref] + flags + ["--stdout", what, "-o", os.path.join(datadir, "surf_so.ref"), fa], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        rb = subprocess.run([ours] + flags + ["--stdout", what, "-o", os.path.join(datadir, "surf_so.gpu"), fa], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert ra.returncode == 0 and rb.returncode == 0, rb.stderr.decode()[-1000:]
        assert ra.stdout == rb.stdout and len(rb.stdout) > 0, f"--stdout {what} differs"
    # two-step: --parse-only, then --pfbwt-only reading <o>.n
    a3, b3 = os.path.join(datadir, "surf_2s.ref"), os.path.join(datadir, "surf_2s.gpu")
    for exe, o in ((ref, a3), (ours, b3)):
        run([exe, "--parse-only", "-s", "-o", o, fa])
        run([exe, "--pfbwt-only", "-s", "-r", "-o", o])
    assert_same_files(a3, b3, exts_for(True, True))


def test_config2_full_size_bwt_hash(datadir):
    """BASELINE.json configs[1] at FULL size: one record of 10^9 i.i.d. uniform bases (NumPy default_rng(1), SURVEY.md 8(d)
    recipe), `pfbwt-f -w 10 -p 100`.  SURVEY.md 8(d) records what the unmodified reference wrote for this input when it was
    run on it (n = 1 000 000 010, P = 10 153 944, D = 1 105 921 870, r = 749 996 069, sha256[:16] of .bwt = 4ab639c13f486fd2);
    the drop-in binary must reproduce all of it.  The FASTA goes through the device record reader (60-column lines)."""
    import hashlib, subprocess
    import numpy as np
    n, lw = 1_000_000_000, 60
    d = None
    for cand in ("/dev/shm", datadir):
        try:
            sv = os.statvfs(cand)
            if os.access(cand, os.W_OK) and sv.f_bavail * sv.f_frsize > 4 * n + (1 << 30):
                d = cand
                break
        except OSError:
            pass
    if d is None:
        pytest.skip("no scratch directory with 5 GB free")
    fa, out = os.path.join(d, "pfb_t_config2.fa"), os.path.join(d, "pfb_t_config2")
    rng = np.random.default_rng(1)
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    try:
        with open(fa, "wb") as f:
            f.write(b">rand1g\n")
            for s in range(0, n, 100_000_000):
                x = lut[rng.integers(0, 4, 100_000_000, dtype=np.uint8)]
                full = x.shape[0] // lw * lw
                a = np.empty((full // lw, lw + 1), dtype=np.uint8)
                a[:, :lw] = x[:full].reshape(-1, lw)
                a[:, lw] = 10
                f.write(a.tobytes())
                f.write(x[full:].tobytes())       # 100 000 000 = 1 666 666 * 60 + 40: the line is continued by the next chunk
            f.write(b"\n")
        r = subprocess.run([os.path.join(BIN, "pfbwt-f"), "-w", "10", "-p", "100", "-o", out, fa], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert r.returncode == 0, r.stderr.decode()[-800:]
        err = r.stderr.decode()
        assert "n: 1000000010" in err and "r: 749996069" in err, err[-800:]
        assert os.path.getsize(out + ".dict") == 1_105_921_870
        assert os.path.getsize(out + ".parse") == 4 * 10_153_944
        h = hashlib.sha256()
        with open(out + ".bwt", "rb") as f:
            for blk in iter(lambda: f.read(1 << 24), b""):
                h.update(blk)
        assert h.hexdigest()[:16] == "4ab639c13f486fd2"
    finally:
        for p in [fa] + [out + "." + e for e in ("bwt", "dict", "occ", "parse", "n", "bwlast", "ilist", "bwsai")]:
            try:
                os.remove(p)
            except OSError:
                pass


@pytest.mark.parametrize("bits", [32, 64])
def test_cli_write_last_sai(bits, ref_bins, datadir):
    """--write-last / --write-sai: <o>.last = PfParser::last_ (P bytes), <o>.sai = sai_ (P x U); the reference defines the
    extensions (include/utils.h:22,24) but never writes them, so the checker is the CPU restatement's arrays"""
    fa = os.path.join(datadir, f"wl{bits}.fa")
    write_fasta(fa, pangenome_records(120_000, 3, 100, 91) + [("x", random_bases(5, 3))], line_width=70)
    a, b = os.path.join(datadir, f"wl{bits}.ref"), os.path.join(datadir, f"wl{bits}.gpu")
    run([ref_bins["pfp_oracle"], "--bits", str(bits), "--write-last", "--parse-only", "-o", a, fa])
    run([os.path.join(BIN, "pfbwt-f64" if bits == 64 else "pfbwt-f"), "--write-last", "--write-sai", "--parse-only", "-o", b, fa])
    assert_same_files(a, b, ["last", "sai", "parse", "dict"])
    P = os.path.getsize(a + ".parse") // 4
    assert os.path.getsize(b + ".last") == P and os.path.getsize(b + ".sai") == P * bits // 8


def test_cli_merge_docs_and_log(ref_bins, datadir):
    """merge_pfp --docs over pre-parsed prefixes: the loaded .docs are offset and merged (include/pfbwt_io.hpp:216-218,
    include/pfparser.hpp:194-263), and the serial mode writes <out>.pfbwt.log (src/merge_pfp.cpp:157-160)"""
    if not ref_bins["merge_pfp"]:
        pytest.skip("reference binaries not built")
    recs = pangenome_records(60_000, 4, 60, 33)
    outs = {}
    for who, parse_exe, merge_exe in (("ref", ref_bins["pfbwt-f64"], ref_bins["merge_pfp"]),
                                      ("gpu", os.path.join(BIN, "pfbwt-f64"), os.path.join(BIN, "merge_pfp"))):
        parts = []
        for g in range(2):
            p = os.path.join(datadir, f"md_{who}{g}.fa")
            write_fasta(p, [(f"{n}_{g}", s) for n, s in recs[2 * g:2 * g + 2]], line_width=60)
            run([parse_exe, "--parse-only", "--print-docs", p])
            parts.append(p)
        out = os.path.join(datadir, f"md_{who}_merged")
        run([merge_exe, "--docs", "-o", out, "-t", "1"] + parts)
        outs[who] = out
    assert_same_files(outs["ref"], outs["gpu"], ["dict", "occ", "parse", "n", "docs"])
    rl = open(outs["ref"] + ".pfbwt.log").read().replace("md_ref", "md_X")
    gl = open(outs["gpu"] + ".pfbwt.log").read().replace("md_gpu", "md_X")
    assert rl == gl and "not using threads (2 files, 1 threads specified)" in gl


def test_cli_one_word_leaves_parse_files(ref_bins, datadir):
    """the reference saves .dict/.occ/.parse/.n before bwt_of_parse (src/pfbwt-f.cpp:221-228): on the 'only one dict word'
    exit those files exist, the parse-BWT files do not"""
    fa = os.path.join(datadir, "oneword.fa")
    write_fasta(fa, [("r", random_bases(3000, 5))])
    a, b = os.path.join(datadir, "oneword.ref"), os.path.join(datadir, "oneword.gpu")
    rb = run([os.path.join(BIN, "pfbwt-f"), "-p", "1000000", "-o", b, fa], check=False)
    assert rb.returncode == 1 and b"only one dict word" in rb.stderr
    for e in ("dict", "occ", "parse", "n"):
        assert os.path.exists(b + "." + e), e
    assert not os.path.exists(b + ".bwlast") and not os.path.exists(b + ".ilist")
    if ref_bins["pfbwt-f"]:
        ra = run([ref_bins["pfbwt-f"], "-p", "1000000", "-o", a, fa], check=False)
        assert ra.returncode == 1
        assert_same_files(a, b, ["dict", "occ", "parse", "n"])


@pytest.mark.parametrize("env", [{"PFB_BUCKETS": "4"}, {"PFB_HBM_LIMIT": "40M"}, {"PFB_BUCKETS": "1"}], ids=["buckets4", "limit40M", "one"])
def test_cli_bucketed_build(env, ref_bins, datadir):
    """-m: workspace cut into key-range buckets, every slice of .bwt/.sa/.ssa/.esa streamed to its offset -
    byte-identical to the reference whatever the number of buckets"""
    import re, subprocess
    fa = os.path.join(datadir, "bucketed.fa")
    flags = ["-s", "-r", "--non-acgt-to-a"]
    a = os.path.join(datadir, "bucketed.ref")
    if not os.path.exists(a + ".bwt"):
        write_fasta(fa, pangenome_records(400_000, 6, 400, 5) + nrun_records(2, 150_000, 9, big_run=20_000), line_width=70)
        run(oracle_cmd(ref_bins, 64, flags, a, fa))
    b = os.path.join(datadir, "bucketed.gpu" + "".join(env.values()))
    r = subprocess.run([os.path.join(BIN, "pfbwt-f64"), "-m"] + flags + ["-o", b, fa], env=dict(os.environ, **env),
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    m = re.search(r"buckets: (\d+), peak device bytes: (\d+)", r.stderr.decode())
    assert m, r.stderr.decode()[-800:]
    if "PFB_HBM_LIMIT" in env:
        assert int(m.group(1)) > 1 and int(m.group(2)) <= 40 << 20, m.group(0)
    if "PFB_BUCKETS" in env:
        assert int(m.group(1)) == int(env["PFB_BUCKETS"])
    assert_same_files(a, b, exts_for(True, True))
    # the counts the driver prints come from the carried totals
    ra = run(oracle_cmd(ref_bins, 64, flags, a + "2", fa))
    for key in (b"n: ", b"r: ", b"# easy cases"):
        la = [l for l in ra.stderr.splitlines() if l.startswith(key)]
        lb = [l for l in r.stderr.splitlines() if l.startswith(key)]
        if la:
            assert la == lb, (la, lb)


def test_cli_doc_array(ref_bins, datadir):
    """--doc-array: <o>.da[i] = index of the record holding SA[i]; checked against the reference's own .sa and
    .docs of the same input (numpy searchsorted), through --pfbwt-only too (docs loaded from <o>.docs)"""
    if not ref_bins["pfbwt-f"]:
        pytest.skip("reference binaries not built")
    recs = pangenome_records(50_000, 5, 50, 61) + [("tiny", random_bases(3, 9)), ("tail", random_bases(20_000, 10))]
    fa = os.path.join(datadir, "da.fa")
    write_fasta(fa, recs, line_width=60)
    a, b = os.path.join(datadir, "da.ref"), os.path.join(datadir, "da.gpu")
    run([ref_bins["pfbwt-f"], "-s", "--print-docs", "-o", a, fa])
    run([os.path.join(BIN, "pfbwt-f"), "-s", "--doc-array", "-o", b, fa])
    assert_same_files(a, b, ["sa", "docs", "bwt"])
    sa = np.fromfile(a + ".sa", dtype=np.uint32).astype(np.int64)
    starts = np.array([int(l.split()[-1]) for l in open(a + ".docs")], dtype=np.int64)
    want = (np.searchsorted(starts, sa, side="right") - 1).astype(np.uint32)
    got = np.fromfile(b + ".da", dtype=np.uint32)
    assert np.array_equal(want, got)
    assert got.max() == len(recs) - 1 and got[0] == len(recs) - 1
