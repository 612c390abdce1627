"""Dictionaries of 2^32 bytes or more (the reference's pfbwt-f64 indexes the dictionary with 64-bit uint_t,
gsa/gsacak.h:44-58, include/pfbwt.hpp:206-223).  The build then runs on 64-bit dictionary positions - (u64, u64) sort pairs,
two-array word lookup, ilist bases from the classify pass - one key range of fewer than 2^32 suffixes after the other.

  * PFB_DICT64=1 / PFB_MEM_BUCKETS force that path at sizes the reference binary runs: byte-identical files through the
    drop-in binaries, in memory, streamed (-m) and over a 2-rank job that shares cuda:0
  * a REAL dictionary beyond 2^32 bytes: 1.1 Gbp of uniform ACGT with w = 31, p = 10 (every phrase carries 32 extra bytes:
    dictionary = 4.6 GB), pfbwt-f64 -s semantics, checked with the size-independent checker of bench.py"""
import os
import subprocess

import numpy as np
import pytest

import bench
from pfp_util import ROOT, write_fasta, random_bases, pangenome_records, nrun_records, run, exts_for, assert_same_files
from test_gpu_cli_parity import oracle_cmd

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "bin")

KNOBS = [{"PFB_DICT64": "1"}, {"PFB_DICT64": "1", "PFB_MEM_BUCKETS": "5"}, {"PFB_DICT64": "1", "PFB_BUCKETS": "3", "-m": "1"},
         {"PFB_DICT64": "1", "PFB_PIVOT": "1", "PFB_KEY_SYMS": "3", "PFB_MEM_BUCKETS": "2"}, {"PFB_DICT64": "1", "PFB_EXT_ROUNDS": "0"},
         {"PFB_MEM_BUCKETS": "3"}]


@pytest.mark.parametrize("knobs", KNOBS, ids=["-".join(f"{k[4:] if k.startswith('PFB_') else k}{v}" for k, v in kn.items()) for kn in KNOBS])
@pytest.mark.parametrize("flags,bits", [(["-s", "-r"], 64), ([], 32)])
def test_cli_forced_wide_positions_match_reference(knobs, flags, bits, ref_bins, datadir):
    fa = os.path.join(datadir, "widegpu.fa")
    if not os.path.exists(fa):
        write_fasta(fa, pangenome_records(150_000, 6, 150, 23) + nrun_records(2, 60_000, 29, big_run=9000) + [("r", random_bases(300_000, 31))], line_width=60)
    fl = flags + ["--non-acgt-to-a"]
    a = os.path.join(datadir, f"widegpu{bits}{''.join(flags)}.ref")
    b = os.path.join(datadir, f"widegpu{bits}{''.join(flags)}_{'_'.join(knobs.values())}.gpu")
    if not os.path.exists(a + ".bwt"):
        run(oracle_cmd(ref_bins, bits, fl, a, fa))
    env = dict(os.environ, PFB_TRACE="1", **{k: v for k, v in knobs.items() if k.startswith("PFB_")})
    r = subprocess.run([os.path.join(BIN, "pfbwt-f64" if bits == 64 else "pfbwt-f")] + (["-m"] if "-m" in knobs else []) + fl + ["-o", b, fa],
                       env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-1500:]
    if "PFB_DICT64" in knobs:
        ranges = int(knobs.get("PFB_MEM_BUCKETS", knobs.get("PFB_BUCKETS", "1")))
        assert r.stderr.decode().count("K7 word tables (64-bit dictionary positions)") == ranges
    assert_same_files(a, b, exts_for("-s" in flags, "-r" in flags))


@pytest.mark.parametrize("mode,world", [("cuda", 2), ("nccl", 2)])
def test_ranks_forced_wide_positions(mode, world, tmp_path, monkeypatch):
    """the key router (position-sliced key exchange, peer stores / all-to-all) with 64-bit positions: ranks sharing cuda:0
    over the host transport, and one GPU per rank over NCCL (skipped on a one-GPU box)"""
    import torch
    from pfp_util import ref_bin
    from test_shard_gloo import run_job, split_records
    if ref_bin("pfbwt-f") is None:
        pytest.skip("reference binaries not built")
    if mode == "nccl" and torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("PFB_DICT64", "1")
    recs = pangenome_records(300_000, 6, 300, seed=43)
    out, refp = run_job(mode, str(tmp_path), split_records(recs, world), 10, 100, 64, True, True)
    assert_same_files(out, refp, exts_for(True, True))


def test_real_dictionary_beyond_2_32_bytes():
    torch = pytest.importorskip("torch")
    free, _ = torch.cuda.mem_get_info()
    if free < 150 * (1 << 30):
        pytest.skip("needs 150 GB of free HBM")
    from pfbwt_b200_loader import binding
    B = binding()
    nb, w, p = 1_100_000_000, 31, 10
    host = torch.empty(nb, dtype=torch.uint8).pin_memory()
    bench.gen_bases(host.numpy(), 5)
    ctx = B.Context(device=0, bits=64, w=w, p=p, get_sai=True)
    ctx.add_records(host.data_ptr(), np.array([nb], dtype=np.uint64))
    ctx.finalize(); ctx.bwt_of_parse(); ctx.release_text(); ctx.generate_bwt(sa=True, rssa=True)
    sz = ctx.sizes()
    n, r = int(sz.n), int(sz.r)
    assert n == nb + w
    assert int(sz.dict_bytes) > (1 << 32), sz.dict_bytes            # the point of the test
    bwt, sa = ctx.bwt(), ctx.sa()
    ssa, esa = ctx.ssa().reshape(-1, 2), ctx.esa().reshape(-1, 2)
    ctx.close()
    tt = host.cuda()
    fill = torch.full((), 65, dtype=torch.uint8, device="cuda")
    res = bench.verify_slices(torch, None, 0, 1, n, bwt, sa, 0, r, lambda pos: torch.where(pos >= nb, fill, tt[pos.clamp(max=nb - 1)]))
    assert res["ok"], res
    starts = np.flatnonzero(np.concatenate([[True], bwt[1:] != bwt[:-1]]))
    ends = np.concatenate([starts[1:] - 1, [n]])
    assert len(ssa) == r and len(esa) == r
    assert np.array_equal(ssa[:, 0], starts.astype(ssa.dtype)) and np.array_equal(ssa[:, 1], sa[starts])
    assert np.array_equal(esa[:, 0], ends.astype(esa.dtype)) and np.array_equal(esa[:, 1], sa[ends])
