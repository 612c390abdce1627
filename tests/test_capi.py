"""C-ABI tests.  CPU part: the shared library loads and exports every symbol include/pfbwt_b200.h
declares, and refuses to run without a GPU (no fallback).  GPU part: parity of every getter with the
oracle's files, calling through the C ABI with host buffers."""
import os
import re
import numpy as np
import pytest
from pfp_util import ROOT, write_fasta, random_bases, pangenome_records, nrun_records, run, naive_sa_bwt
from pfbwt_b200_loader import binding


def header_symbols():
    h = open(os.path.join(ROOT, "include", "pfbwt_b200.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(pfb_[a-z_0-9]+)\s*\(", h)))


def test_library_exports_every_declared_symbol():
    B = binding()
    lib = B.load()
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/pfbwt_b200.h but not exported"
    assert sorted(B.EXPORTS) == syms


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    B = binding()
    with pytest.raises(B.PfbError) as e:
        B.Context()
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)


def oracle_files(ref_bins, bits, flags, out, fa):
    exe = ref_bins["pfbwt-f64" if bits == 64 else "pfbwt-f"]
    if exe:
        run([exe] + flags + ["-o", out, fa])
    else:
        run([ref_bins["pfp_oracle"], "--bits", str(bits)] + flags + ["-o", out, fa])


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [32, 64])
def test_capi_getters_match_oracle(bits, ref_bins, datadir):
    B = binding()
    recs = pangenome_records(120_000, 5, 120, 17) + nrun_records(1, 50_000, 3)
    fa = os.path.join(datadir, f"capi{bits}.fa")
    write_fasta(fa, recs, line_width=61)
    ref = os.path.join(datadir, f"capi{bits}.ref")
    oracle_files(ref_bins, bits, ["-s", "-r"], ref, fa)
    # records through host buffers (no file)
    seq = np.concatenate([r[1] for r in recs])
    lens = np.array([len(r[1]) for r in recs], dtype=np.uint64)
    ctx = B.Context(device=0, bits=bits, w=10, p=100, get_sai=True)
    ctx.add_records(seq, lens)
    ctx.finalize()
    ctx.bwt_of_parse()
    ctx.generate_bwt(sa=True, rssa=True)
    got = {"dict": ctx.dict(), "occ": ctx.occ(), "parse": ctx.parse(), "bwlast": ctx.bwlast(), "ilist": ctx.ilist(),
           "bwsai": ctx.bwsai(), "bwt": ctx.bwt(), "sa": ctx.sa(), "ssa": ctx.ssa(), "esa": ctx.esa()}
    for ext, arr in got.items():
        assert arr.tobytes() == open(f"{ref}.{ext}", "rb").read(), f".{ext} differs"
    sz = ctx.sizes()
    assert sz.n == int(open(ref + ".n").read())
    assert sz.r * 2 == len(got["ssa"])
    # last_/sai_ (include/pfparser.hpp:599-600): sai ascending, last byte = T'[sai - w]
    sai, last = ctx.sai(), ctx.last()
    assert np.all(np.diff(sai.astype(np.int64)) > 0) and int(sai[-1]) == sz.n + 10
    ctx.close()


@pytest.mark.gpu
def test_capi_resident_rerun_is_idempotent(datadir):
    """pfb_upload_text + pfb_run_resident (the benchmark's timed region) twice gives the same BWT,
    and the BWT inverts back to the input text (size-independent round-trip property)."""
    B = binding()
    S = random_bases(3_000_000, 123)
    ctx = B.Context(device=0, bits=32, get_sai=True)
    ctx.add_records(S, np.array([len(S)], dtype=np.uint64))
    ctx.upload_text()
    ctx.run_resident(True, False)
    bwt1, sa1 = ctx.bwt().copy(), ctx.sa().copy()
    ctx.run_resident(True, False)
    assert np.array_equal(bwt1, ctx.bwt()) and np.array_equal(sa1, ctx.sa())
    n = len(S) + 10
    assert sa1[0] == n and len(bwt1) == n + 1
    # SA is a permutation; BWT[i] = T[SA[i]-1]
    T = np.concatenate([S, np.frombuffer(b"A" * 10, dtype=np.uint8), np.zeros(1, dtype=np.uint8)])
    assert np.array_equal(np.sort(sa1), np.arange(n + 1, dtype=sa1.dtype))
    assert np.array_equal(bwt1, T[(sa1.astype(np.int64) - 1) % (n + 1)])
    # sortedness of sampled adjacent suffix pairs
    rng = np.random.default_rng(0)
    for i in rng.integers(1, n, 2000):
        a, b = int(sa1[i - 1]), int(sa1[i])
        assert T[a:a + 64].tobytes() + b"" <= T[b:b + 64].tobytes() or T[a:a + 64].tobytes() == T[b:b + 64].tobytes()
    ctx.close()


@pytest.mark.gpu
def test_capi_errors(datadir):
    B = binding()
    ctx = B.Context(device=0, bits=32)
    with pytest.raises(B.PfbError) as e:
        ctx.add_records(np.frombuffer(b"ACGTXACGT" * 10, dtype=np.uint8), np.array([90], dtype=np.uint64))
        ctx.finalize()
    assert "invalid character 88/X" in str(e.value)
    ctx.clear()
    # a text that parses into a single phrase (include/pfparser.hpp:390-392)
    ctx.add_records(np.frombuffer(b"A" * 500, dtype=np.uint8), np.array([500], dtype=np.uint64))
    ctx.finalize()
    with pytest.raises(B.PfbError) as e:
        ctx.bwt_of_parse()
    assert "only one dict word total" in str(e.value)
    ctx.close()
    with pytest.raises(B.PfbError):
        B.Context(device=0, bits=32, w=40)


@pytest.mark.gpu
def test_trigger_known_answers():
    """a1/a2 of SURVEY 8: the device wang_hash equals the reference's (KAT wang_hash(0) = 0x77cfa1eef01bca90, probe of
    include/hash.hpp:12-21; the C restatement for the rest) and the division-free divisibility test of the trigger
    kernel equals `hash % p == 0` (include/pfparser.hpp:347) for odd, even and power-of-two moduli"""
    import ctypes
    B = binding()
    ctx = B.Context(device=0, bits=32, w=10, p=100)
    rng = np.random.default_rng(7)
    kmers = np.concatenate([np.array([0, 1, 2, (1 << 20) - 1, (1 << 62) - 1, (1 << 64) - 1], dtype=np.uint64),
                            rng.integers(0, 1 << 20, 50_000, dtype=np.uint64), rng.integers(0, 1 << 63, 50_000, dtype=np.uint64)])
    lib = os.path.join(ROOT, "oracle", "_ref", "libpfp_oracle.so")
    ora = ctypes.CDLL(lib) if os.path.exists(lib) else None
    if ora:
        ora.pfo_wang_hash.restype = ctypes.c_uint64
        ora.pfo_wang_hash.argtypes = [ctypes.c_uint64]
    for p in (100, 1, 2, 3, 5, 7, 64, 96, 97, 1000, 4096, 1_000_003, (1 << 31) - 1, 6_000_000_000):
        h, d = ctx.kat_trigger(kmers, p)
        assert int(h[0]) == 0x77cfa1eef01bca90
        if ora:
            exp = np.array([ora.pfo_wang_hash(int(k)) for k in kmers[:2000]], dtype=np.uint64)
            assert np.array_equal(h[:2000], exp)
        assert np.array_equal(d.astype(bool), (h % np.uint64(p)) == 0), f"p={p}"
        if p <= 100:
            assert d.sum() > 0
    ctx.close()


@pytest.mark.gpu
def test_release_text_drops_the_staged_input(datadir):
    """pfb_release_text after the parse: the BWT build still works (it needs dict/parse only), parsing again is refused
    with a message until pfb_clear"""
    B = binding()
    fa = os.path.join(datadir, "rel.fa")
    seq = random_bases(200_000, 21)
    write_fasta(fa, [("r", seq)], line_width=60)
    ctx = B.Context(device=0, bits=32, w=10, p=100)
    ctx.add_fasta(fa)
    ctx.finalize(); ctx.bwt_of_parse()
    ctx.release_text()
    ctx.generate_bwt(True, False)
    bwt1 = ctx.bwt().copy()
    with pytest.raises(Exception) as ei:
        ctx.finalize()
    assert "released" in str(ei.value)
    ctx.clear()
    ctx.add_fasta(fa)
    ctx.finalize(); ctx.bwt_of_parse(); ctx.generate_bwt(True, False)
    assert np.array_equal(bwt1, ctx.bwt())
    ctx.close()


@pytest.mark.gpu
def test_gpu_fuzz_against_restatement(ref_bins, datadir):
    """the 50 random jobs of tests/test_emu_logic.py (window, modulus, record mix incl. tandem repeats, one-symbol records, N runs,
    records shorter than the window; flags) through the C ABI on the device: every file equals the C restatement's, and a
    job the restatement refuses is refused"""
    import subprocess
    from test_emu_logic import fuzz_case
    from pfp_util import exts_for, assert_same_files
    B = binding()
    accepted = 0
    njobs = int(os.environ.get("PFB_FUZZ_N", "50"))          # a longer hunt: PFB_FUZZ_N=500
    for seed in range(njobs):
        recs, bits, flags = fuzz_case(7000 + seed)
        w, p = int(flags[flags.index("-w") + 1]), int(flags[flags.index("-p") + 1])
        sa, rssa, ntoa = "-s" in flags, "-r" in flags, "--non-acgt-to-a" in flags
        fa = os.path.join(datadir, f"gfz{seed}.fa")
        write_fasta(fa, recs, line_width=50 + seed)
        a, b = os.path.join(datadir, f"gfz{seed}.port"), os.path.join(datadir, f"gfz{seed}.gpu")
        ra = subprocess.run([ref_bins["pfp_oracle"], "--bits", str(bits)] + flags + ["-o", a, fa], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        ctx = B.Context(device=0, bits=bits, w=w, p=p, get_sai=sa or rssa, non_acgt_to_a=ntoa)
        try:
            ctx.add_fasta(fa)
            ctx.finalize()
            ctx.bwt_of_parse()
            ctx.save_parser(b, True, sa or rssa)
            ctx.generate_bwt(sa, rssa)
            ctx.write_bwt(b, sa, rssa)
            ok = True
        except Exception as e:
            ok = False
            msg = str(e)
        finally:
            ctx.close()
        assert ok == (ra.returncode == 0), (seed, flags, ra.stderr[-300:], "" if ok else msg)
        if ok:
            accepted += 1
            try:
                assert_same_files(a, b, exts_for(sa, rssa))
            except AssertionError as e:
                raise AssertionError(f"seed {seed} flags {flags} bits {bits}: {e}")
    assert accepted >= njobs * 4 // 5
