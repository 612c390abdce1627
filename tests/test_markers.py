"""The marker-array join: src/mps_to_ma.cpp / write_marker_array (include/marker_array.hpp:137-174) over the
run-length window array of include/rle_window_array.hpp.  The reference's golden `.markers` files cannot be regenerated here
(their `.mps` inputs come from vcf_scan, which needs htslib), so the join is pinned on the UNMODIFIED reference tool
(oracle/_ref/mps_to_ma, compiled from the reference's sources) run on synthetic marker runs and on real suffix arrays.

CPU: the C restatement (oracle) == the reference tool; the host logic of the device join (emulator) == the reference, for
every chunk size (runs crossing chunk borders), from a file and from stdin, and from the SA a build leaves in memory.
GPU: bin/mps_to_ma and `pfbwt-f[64] -s --marker-array` write the reference's bytes."""
import os
import subprocess

import numpy as np
import pytest

from pfp_util import ROOT, write_fasta, random_bases, pangenome_records, run, ref_bin

EMU = os.path.join(ROOT, "tests", "emu", "pfb_emu")
BIN = os.path.join(ROOT, "bin")
DELIM = 0xFFFFFFFFFFFFFFFF


def marker(pos, allele, seq):
    """create_marker_t (include/marker.hpp:41-50): allele:4 | seq:16 | pos:44"""
    return ((allele & 0xF) << 60) | ((seq & 0xFFFF) << 46) | (pos & ((1 << 44) - 1))


def random_mps(path, n, seed, first=3, max_gap=3, max_len=40):
    """ascending, disjoint runs of text positions below n with 1-5 markers each; lists repeat (also on neighbouring runs: the
    join compares lists by content) and runs may touch"""
    rng = np.random.default_rng(seed)
    lists = [[marker(int(rng.integers(1, 1 << 30)), int(rng.integers(0, 4)), int(rng.integers(0, 50))) for _ in range(int(rng.integers(1, 6)))]
             for _ in range(12)]
    words, pos, runs = [], first, 0
    while pos < n - max_len:
        ln = int(rng.integers(1, max_len))
        words += [pos, pos + ln - 1] + lists[int(rng.integers(0, len(lists)))] + [DELIM]
        pos += ln + int(rng.integers(0, max_gap))
        runs += 1
    np.array(words, dtype=np.uint64).tofile(path)
    return runs


def reference_join(ref_bins, mps, sa, out):
    tool = ref_bin("mps_to_ma")
    if tool:
        run([tool, "-o", out, mps, sa])
    else:
        run([ref_bins["pfp_oracle"], "--marker-join", mps, sa, out])


@pytest.fixture(scope="module")
def emu():
    subprocess.run(["make", "-C", ROOT, "emu"], check=True, stdout=subprocess.DEVNULL)
    return EMU


@pytest.fixture(scope="module")
def join_case(datadir):
    n = 60_000
    mps, sa = os.path.join(datadir, "mj.mps"), os.path.join(datadir, "mj.sa")
    random_mps(mps, n, 5)
    np.random.default_rng(6).permutation(n + 1).astype(np.uint64).tofile(sa)
    # a suffix array with long stretches of consecutive text positions (repetitive input): long output runs
    sa2 = os.path.join(datadir, "mj_blocks.sa")
    blocks = np.random.default_rng(7).permutation(60)
    np.concatenate([np.arange(b * 1000, (b + 1) * 1000, dtype=np.uint64) for b in blocks]).tofile(sa2)
    return mps, sa, sa2


def test_restatement_matches_reference_tool(join_case, ref_bins, datadir):
    if ref_bin("mps_to_ma") is None:
        pytest.skip("reference mps_to_ma not built")
    mps, sa, sa2 = join_case
    for k, s in enumerate((sa, sa2)):
        a, b = os.path.join(datadir, f"mj_ref{k}.ma"), os.path.join(datadir, f"mj_port{k}.ma")
        run([ref_bin("mps_to_ma"), "-o", a, mps, s])
        run([ref_bins["pfp_oracle"], "--marker-join", mps, s, b])
        assert open(a, "rb").read() == open(b, "rb").read()
        assert os.path.getsize(a) > 0


@pytest.mark.parametrize("chunk", ["1", "7", "1000", "0"])
def test_emu_join_matches_reference(chunk, join_case, emu, ref_bins, datadir):
    mps, sa, sa2 = join_case
    for k, s in enumerate((sa, sa2)):
        a, b = os.path.join(datadir, f"mj_e_ref{k}.ma"), os.path.join(datadir, f"mj_e{k}_{chunk}.ma")
        if not os.path.exists(a):
            reference_join(ref_bins, mps, s, a)
        env = dict(os.environ, **({"PFB_MA_CHUNK": chunk} if chunk != "0" else {}))
        subprocess.run([emu, "--marker-join", mps, s, b], check=True, env=env)
        assert open(a, "rb").read() == open(b, "rb").read(), (k, chunk)
    # the SA as a stream (pfbwt-f64 --stdout sa | mps_to_ma ... -)
    b = os.path.join(datadir, f"mj_e_stdin_{chunk}.ma")
    with open(sa, "rb") as f:
        subprocess.run([emu, "--marker-join", mps, "-", b], check=True, stdin=f, env=dict(os.environ, PFB_MA_CHUNK="513"))
    assert open(os.path.join(datadir, "mj_e_ref0.ma"), "rb").read() == open(b, "rb").read()


def test_emu_join_edge_cases(emu, ref_bins, datadir):
    sa = os.path.join(datadir, "edge.sa")
    np.arange(50, dtype=np.uint64)[::-1].copy().tofile(sa)

    def both(words, name, tool=None):
        # tiny files: checked against the C restatement (pinned on the reference tool by the tests above) - the reference
        # sizes a bit vector as file words / 13 * 10 (rle_window_array.hpp:22) and asserts on files this small
        mps = os.path.join(datadir, name + ".mps")
        np.array(words, dtype=np.uint64).tofile(mps)
        a, b = os.path.join(datadir, name + "_ref.ma"), os.path.join(datadir, name + "_emu.ma")
        ra = subprocess.run([tool, "-o", a, mps, sa] if tool else [ref_bins["pfp_oracle"], "--marker-join", mps, sa, a],
                            stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        rb = subprocess.run([emu, "--marker-join", mps, sa, b], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        return ra, rb, a, b

    # a run that starts at text position 0: the reference refuses it (include/rle_window_array.hpp:28-31), so do we
    ra, rb, _, _ = both([0, 4, marker(9, 1, 1), DELIM, 10, 12, marker(5, 2, 1), DELIM], "edge0", ref_bin("mps_to_ma"))
    assert ra.returncode != 0 and rb.returncode != 0
    assert b"warning equal run start at 0 or run end at 0" in ra.stderr and b"warning equal run start at 0 or run end at 0" in rb.stderr
    # a run one position wide at the very end of the text; positions outside every run carry nothing
    ra, rb, a, b = both([7, 8, marker(3, 1, 1), DELIM, 49, 49, marker(1, 1, 1), marker(2, 0, 1), DELIM], "edge1")
    assert ra.returncode == 0 and rb.returncode == 0 and open(a, "rb").read() == open(b, "rb").read()
    assert np.fromfile(b, dtype=np.uint64).tolist() == [0, 0, marker(1, 1, 1), marker(2, 0, 1), DELIM, 41, 42, marker(3, 1, 1), DELIM]
    # two touching runs with the SAME list make one output run; a third with another list splits it
    m = marker(77, 3, 2)
    ra, rb, a, b = both([5, 9, m, DELIM, 10, 14, m, DELIM, 15, 15, marker(78, 3, 2), DELIM], "edge2")
    assert ra.returncode == 0 and rb.returncode == 0 and open(a, "rb").read() == open(b, "rb").read()
    assert np.fromfile(b, dtype=np.uint64).tolist() == [34, 34, marker(78, 3, 2), DELIM, 35, 44, m, DELIM]
    # runs out of order are refused (rank/select over the starts assumes ascending input; the reference reads garbage)
    mps = os.path.join(datadir, "edge3.mps")
    np.array([20, 24, m, DELIM, 10, 14, m, DELIM], dtype=np.uint64).tofile(mps)
    r = subprocess.run([emu, "--marker-join", mps, sa, os.path.join(datadir, "edge3.ma")], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode != 0 and b"ascending" in r.stderr


@pytest.mark.parametrize("bits", [32, 64])
def test_emu_marker_array_after_streamed_build(bits, emu, ref_bins, datadir):
    """-m (key-range buckets streamed to the files) with --marker-array: the join reads the <prefix>.sa file the build wrote,
    4-byte entries in the 32-bit binary"""
    fa = os.path.join(datadir, "ma_stream.fa")
    recs = pangenome_records(20_000, 4, 30, 71)
    write_fasta(fa, recs, line_width=60)
    n = sum(len(s) + 10 for _, s in recs)
    mps = os.path.join(datadir, "ma_stream.mps")
    random_mps(mps, n, 72)
    out = os.path.join(datadir, f"ma_stream{bits}")
    subprocess.run([emu, "--bits", str(bits), "-m", "-s", "--marker-array", mps, "-o", out, fa], check=True,
                   env=dict(os.environ, PFB_BUCKETS="3", PFB_MA_CHUNK="1001"), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    sa64 = out + ".sa64"
    np.fromfile(out + ".sa", dtype=np.uint32 if bits == 32 else np.uint64).astype(np.uint64).tofile(sa64)
    ref = out + ".ref.ma"
    reference_join(ref_bins, mps, sa64, ref)
    assert os.path.getsize(ref) > 0
    assert open(ref, "rb").read() == open(out + ".ma", "rb").read()


@pytest.mark.parametrize("bits", [32, 64])
def test_emu_marker_array_from_the_build(bits, emu, ref_bins, datadir):
    """pfbwt-f[64] -s --marker-array <mps>: the join with the SA the build left in memory == the reference tool on the .sa file"""
    fa = os.path.join(datadir, "ma_build.fa")
    recs = pangenome_records(20_000, 4, 30, 51)
    write_fasta(fa, recs, line_width=60)
    n = sum(len(s) + 10 for _, s in recs)
    mps = os.path.join(datadir, "ma_build.mps")
    random_mps(mps, n, 52)
    out = os.path.join(datadir, f"ma_build{bits}")
    subprocess.run([emu, "--bits", str(bits), "-s", "--marker-array", mps, "-o", out, fa], check=True, env=dict(os.environ, PFB_MA_CHUNK="4099"),
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    sa64 = out + ".sa64"
    np.fromfile(out + ".sa", dtype=np.uint32 if bits == 32 else np.uint64).astype(np.uint64).tofile(sa64)
    ref = out + ".ref.ma"
    reference_join(ref_bins, mps, sa64, ref)
    assert os.path.getsize(ref) > 0
    assert open(ref, "rb").read() == open(out + ".ma", "rb").read()


# ---- GPU: the drop-in tool and the flag ------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("chunk", ["0", "4097"])
def test_gpu_mps_to_ma_matches_reference(chunk, join_case, ref_bins, datadir):
    mps, sa, sa2 = join_case
    env = dict(os.environ, **({"PFB_MA_CHUNK": chunk} if chunk != "0" else {}))
    for k, s in enumerate((sa, sa2)):
        a, b = os.path.join(datadir, f"mj_g_ref{k}.ma"), os.path.join(datadir, f"mj_g{k}_{chunk}.ma")
        reference_join(ref_bins, mps, s, a)
        subprocess.run([os.path.join(BIN, "mps_to_ma"), "-o", b, mps, s], check=True, env=env)
        assert open(a, "rb").read() == open(b, "rb").read(), (k, chunk)
    b = os.path.join(datadir, f"mj_g_stdin_{chunk}.ma")
    with open(sa, "rb") as f:
        subprocess.run([os.path.join(BIN, "mps_to_ma"), "-m", "-o", b, mps, "-"], check=True, stdin=f, env=env)
    assert open(os.path.join(datadir, "mj_g_ref0.ma"), "rb").read() == open(b, "rb").read()


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [32, 64])
def test_gpu_marker_array_flag(bits, ref_bins, datadir):
    """the pipeline of vcf_to_bwt.py (`pfbwt-f64 --pfbwt-only --stdout sa | mps_to_ma`, vcf_to_bwt.py:174-181,242) in one
    process: the SA never leaves the device"""
    fa = os.path.join(datadir, "ma_gpu.fa")
    recs = pangenome_records(300_000, 6, 300, 61) + [("r", random_bases(200_000, 62))]
    write_fasta(fa, recs, line_width=60)
    n = sum(len(s) + 10 for _, s in recs)
    mps = os.path.join(datadir, "ma_gpu.mps")
    assert random_mps(mps, n, 63, max_gap=200, max_len=60) > 1000
    out = os.path.join(datadir, f"ma_gpu{bits}")
    run([os.path.join(BIN, "pfbwt-f64" if bits == 64 else "pfbwt-f"), "-s", "--marker-array", mps, "-o", out, fa])
    if bits == 32:            # the same through a streamed build (-m): the join reads the 4-byte .sa file
        subprocess.run([os.path.join(BIN, "pfbwt-f"), "-m", "-s", "--marker-array", mps, "-o", out + "_m", fa], check=True,
                       env=dict(os.environ, PFB_BUCKETS="3"), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        assert open(out + ".ma", "rb").read() == open(out + "_m.ma", "rb").read()
    sa64 = out + ".sa64"
    np.fromfile(out + ".sa", dtype=np.uint32 if bits == 32 else np.uint64).astype(np.uint64).tofile(sa64)
    ref = out + ".ref.ma"
    reference_join(ref_bins, mps, sa64, ref)
    assert os.path.getsize(ref) > 0
    assert open(ref, "rb").read() == open(out + ".ma", "rb").read()
    if bits == 64:            # and the two-process form with the stream
        p1 = subprocess.Popen([os.path.join(BIN, "pfbwt-f64"), "--pfbwt-only", "-s", "--stdout", "sa", "-o", out], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL)
        p2 = subprocess.run([os.path.join(BIN, "mps_to_ma"), "-o", out + ".piped.ma", mps, "-"], stdin=p1.stdout)
        p1.wait()
        assert p1.returncode == 0 and p2.returncode == 0
        assert open(ref, "rb").read() == open(out + ".piped.ma", "rb").read()
