"""BASELINE.json configs 2-4 at (or near) full size on one B200, where the reference cannot be run in a test: size-independent
properties of the returned BWT + SA checked on the GPU with the checker of bench.py (every (SA[i], BWT[i]) pair against the text
= inversion, SA is a permutation, sampled adjacent suffixes in order with ties followed to their end - no pair is skipped -,
run count = r, .ssa/.esa = the run starts / ends).  Default sizes keep the suite within minutes; PFB_FULL_SIZE=1 runs
configs[2] (64 haplotypes x 100 Mbp) and configs[3] (3.1 Gbp assembly) as BASELINE.json states them (tools/check_large.py and
tools/check_config4.py are the stand-alone forms)."""
import os

import numpy as np
import pytest

import bench
from pfp_util import nrun_records

pytestmark = pytest.mark.gpu
FULL = os.environ.get("PFB_FULL_SIZE") == "1"


def _torch():
    import torch
    return torch


def _need_hbm(gb):
    torch = _torch()
    free, _ = torch.cuda.mem_get_info()
    if free < gb * (1 << 30):
        pytest.skip(f"needs {gb} GB of free HBM")


def _check_rssa(bwt, sa, ssa, esa, n, r):
    starts = np.flatnonzero(np.concatenate([[True], bwt[1:] != bwt[:-1]]))
    ends = np.concatenate([starts[1:] - 1, [n]])
    assert len(ssa) == r and len(esa) == r
    assert np.array_equal(ssa[:, 0], starts.astype(ssa.dtype)) and np.array_equal(ssa[:, 1], sa[starts])
    assert np.array_equal(esa[:, 0], ends.astype(esa.dtype)) and np.array_equal(esa[:, 1], sa[ends])


def test_config2_full_size_sa_properties():
    """configs[1] with -s at full size (10^9 bases, 32-bit): the .bwt hash test covers the BWT, this covers the 4 GB .sa"""
    torch = _torch()
    _need_hbm(70)
    from pfbwt_b200_loader import binding
    B = binding()
    nb = 1_000_000_000
    host = torch.empty(nb, dtype=torch.uint8).pin_memory()
    bench.gen_bases(host.numpy(), 1)
    ctx = B.Context(device=0, bits=32, w=10, p=100, get_sai=True)
    ctx.add_records(host.data_ptr(), np.array([nb], dtype=np.uint64))
    ctx.finalize(); ctx.bwt_of_parse(); ctx.release_text(); ctx.generate_bwt(sa=True, rssa=False)
    sz = ctx.sizes()
    assert (sz.n, sz.r, sz.n_phrases, sz.dict_bytes) == (bench.CONFIG2["n"], bench.CONFIG2["r"], bench.CONFIG2["phrases"], bench.CONFIG2["dict_bytes"])
    bwt, sa = ctx.bwt(), ctx.sa()
    ctx.close()
    assert bench.sha16(bwt) == bench.CONFIG2["bwt_sha16"]
    tt = host.cuda()
    fill = torch.full((), 65, dtype=torch.uint8, device="cuda")
    res = bench.verify_slices(torch, None, 0, 1, int(sz.n), bwt, sa.astype(np.uint64), 0, int(sz.r),
                              lambda pos: torch.where(pos >= nb, fill, tt[pos.clamp(max=nb - 1)]))
    assert res["ok"], res


def test_config3_family_properties():
    """configs[2] family: haplotypes of one 100 Mbp base with 0.1 % substitutions, pfbwt-f64 -s -r (16 haplotypes; 64 with
    PFB_FULL_SIZE=1)"""
    torch = _torch()
    haps, base = (64, 100_000_000) if FULL else (16, 100_000_000)
    _need_hbm(150 if FULL else 60)
    from pfbwt_b200_loader import binding
    B = binding()
    codes = bench.job_codes(base)
    host = torch.empty(haps * base, dtype=torch.uint8).pin_memory()
    bench.gen_job_text(host.numpy(), 0, haps, base, codes)
    ctx = B.Context(device=0, bits=64, w=10, p=100, get_sai=True)
    ctx.add_records(host.data_ptr(), np.full(haps, base, dtype=np.uint64))
    del host
    ctx.finalize(); ctx.bwt_of_parse(); ctx.release_text(); ctx.generate_bwt(sa=True, rssa=True)
    sz = ctx.sizes()
    n, r = int(sz.n), int(sz.r)
    assert n == haps * (base + 10)
    bwt, sa = ctx.bwt(), ctx.sa()
    ssa, esa = ctx.ssa().reshape(-1, 2), ctx.esa().reshape(-1, 2)
    ctx.close()
    res = bench.verify_slices(torch, None, 0, 1, n, bwt, sa, 0, r, bench.job_text_evaluator(torch, codes, haps, base))
    assert res["ok"], res
    _check_rssa(bwt, sa, ssa, esa, n, r)


def test_config4_family_properties():
    """configs[3] family: an assembly with N runs (telomeres, Mbp gaps, soft-masked lower case) under --non-acgt-to-a: the
    poly-A ties of millions of symbols are closed by run-aware pivot rounds.  24 x 12.5 Mbp (the 3.1 Gbp form: tools/check_config4.py)"""
    torch = _torch()
    _need_hbm(40)
    from pfbwt_b200_loader import binding
    B = binding()
    nrec, rec_len = 24, 12_500_000
    recs = nrun_records(nrec, rec_len, 17, big_run=2_000_000, small_runs=20)
    raw = np.concatenate([s for _, s in recs])
    ctx = B.Context(device=0, bits=64, w=10, p=100, get_sai=True, non_acgt_to_a=True)
    ctx.add_records(raw, np.full(nrec, rec_len, dtype=np.uint64))
    ctx.finalize(); ctx.bwt_of_parse(); ctx.release_text(); ctx.generate_bwt(sa=True, rssa=True)
    sz = ctx.sizes()
    n, r = int(sz.n), int(sz.r)
    bwt, sa = ctx.bwt(), ctx.sa()
    ssa, esa = ctx.ssa().reshape(-1, 2), ctx.esa().reshape(-1, 2)
    ctx.close()
    # S: upper case, everything but ACGT -> A, every record followed by A^w
    up = raw & 0xDF
    up[~np.isin(up, np.frombuffer(b"ACGT", dtype=np.uint8))] = ord("A")
    S = np.full(nrec * (rec_len + 10), ord("A"), dtype=np.uint8)
    S.reshape(nrec, rec_len + 10)[:, :rec_len] = up.reshape(nrec, rec_len)
    assert n == S.shape[0]
    tt = torch.from_numpy(S).cuda()
    res = bench.verify_slices(torch, None, 0, 1, n, bwt, sa, 0, r, lambda pos: tt[pos])
    assert res["ok"], res
    _check_rssa(bwt, sa, ssa, esa, n, r)
