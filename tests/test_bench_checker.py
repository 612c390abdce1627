"""The output checker of bench.py (verify_slices / job_text_evaluator / gen_bases_range) is itself checked here on CPU
against an independent suffix array (numpy prefix doubling): a correct BWT+SA passes, single-byte / single-swap
corruptions are caught.  The checker runs outside every timed region of the bench."""
import numpy as np
import pytest
import torch

import bench
from pfp_util import naive_sa_bwt


def test_gen_bases_range_is_a_window_of_gen_bases():
    a = np.empty(1000, dtype=np.uint8)
    bench.gen_bases(a, 1)
    b = np.empty(300, dtype=np.uint8)
    bench.gen_bases_range(1, 200, 500, b)
    assert (a[200:500] == b).all()


def test_checker_accepts_correct_and_rejects_corrupt(monkeypatch):
    L = 5000
    monkeypatch.setattr(bench, "JOB_SNP_RATE", 0.01)
    codes = bench.job_codes(L)
    out = np.empty(4 * L, dtype=np.uint8)
    for r in range(2):
        bench.gen_job_text(out[r * 2 * L:(r + 1) * 2 * L], r, 2, L, codes)
    S = np.concatenate([np.concatenate([out[h * L:(h + 1) * L], np.full(bench.W, 65, np.uint8)]) for h in range(4)])
    ev = bench.job_text_evaluator(torch, codes, 4, L)
    got = ev(torch.arange(len(S), dtype=torch.int64)).cpu().numpy()
    assert (got == S).all()
    sa, bwt = naive_sa_bwt(S)
    n = len(S)
    r = int((bwt[1:] != bwt[:-1]).sum()) + 1
    res = bench.verify_slices(torch, None, 0, 1, n, bwt, sa.astype(np.uint64), 0, r, ev, pairs=500)
    assert res["ok"] and res["positions"] == n + 1 and res["sampled_pairs"] == 500, res
    bw2 = bwt.copy()
    bw2[100] ^= 2
    assert not bench.verify_slices(torch, None, 0, 1, n, bw2, sa.astype(np.uint64), 0, r, ev, pairs=50)["ok"]
    i = int(np.flatnonzero(bwt[1:] != bwt[:-1])[10])   # SA entries swapped under different BWT bytes: the inversion breaks
    sa2 = sa.copy()
    sa2[[i, i + 1]] = sa2[[i + 1, i]]
    res2 = bench.verify_slices(torch, None, 0, 1, n, bwt, sa2.astype(np.uint64), 0, r, ev, pairs=50)
    assert not res2["ok"] and res2["inversion_mismatches"] > 0
    sa4, bw4 = sa.copy(), bwt.copy()                    # (SA, BWT) pairs swapped together: still inverts, but out of order
    for j in range(100, n - 2, 100):
        sa4[[j, j + 1]] = sa4[[j + 1, j]]
        bw4[[j, j + 1]] = bw4[[j + 1, j]]
    res4 = bench.verify_slices(torch, None, 0, 1, n, bw4, sa4.astype(np.uint64), 0, r, ev, pairs=4000)
    assert res4["inversion_mismatches"] == 0 and res4["order_violations"] > 0 and not res4["ok"]
    sa3 = sa.copy()
    sa3[7] = sa3[8]                                     # not a permutation
    assert not bench.verify_slices(torch, None, 0, 1, n, bwt, sa3.astype(np.uint64), 0, r, ev, pairs=50)["sa_is_permutation"]
    assert not bench.verify_slices(torch, None, 0, 1, n, bwt, sa.astype(np.uint64), 0, r + 1, ev, pairs=50)["ok"]
