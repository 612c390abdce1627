#!/usr/bin/env python
"""Marker-array join on one GPU, timed next to the reference's tool:

  python tools/bench_markers.py [--bases 1000000000] [--runs 2000000] [--ref-sample 40000000]

Builds the BWT + 64-bit SA of a config-2 style text (uniform ACGT, seed 1) on the GPU, joins a synthetic marker file with
the SA in HBM (pfb_write_marker_array; output in /dev/shm) and times it; then the unmodified reference mps_to_ma
(oracle/_ref, one host core) and bin/mps_to_ma on the first `ref-sample` SA entries written to a file, outputs compared
byte for byte.  One JSON line."""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bases", type=int, default=1_000_000_000)
    ap.add_argument("--runs", type=int, default=2_000_000)
    ap.add_argument("--ref-sample", type=int, default=40_000_000)
    ap.add_argument("--dir", default="/dev/shm")
    a = ap.parse_args()
    import torch
    import bench
    from pfbwt_b200_loader import binding
    from test_markers import marker, DELIM
    B = binding()
    nb = a.bases
    host = torch.empty(nb, dtype=torch.uint8).pin_memory()
    bench.gen_bases(host.numpy(), 1)
    # marker runs: a.runs runs of 1-60 positions spread over the text, 1-4 markers each from a pool of lists
    rng = np.random.default_rng(3)
    R = a.runs
    gap = nb // R
    starts = (np.arange(R, dtype=np.uint64) * np.uint64(gap) + np.uint64(5) + rng.integers(0, max(gap // 4, 1), R).astype(np.uint64))
    lens = rng.integers(1, min(60, max(gap // 2, 2)), R).astype(np.uint64)
    nm = rng.integers(1, 5, R)
    pool = np.array([marker(int(p), int(p) % 4, int(p) % 40) for p in rng.integers(1, 1 << 30, 4096)], dtype=np.uint64)
    words = np.empty(int(3 * R + nm.sum()), dtype=np.uint64)
    off = np.concatenate([[0], np.cumsum(3 + nm)]).astype(np.int64)
    words[off[:-1]] = starts
    words[off[:-1] + 1] = starts + lens - np.uint64(1)
    words[off[1:] - 1] = np.uint64(DELIM)
    pick = rng.integers(0, 4096 - 4, R)
    for k in range(4):
        sel = nm > k
        words[off[:-1][sel] + 2 + k] = pool[pick[sel] + k]
    mps = os.path.join(a.dir, "pfb_bench.mps")
    words.tofile(mps)
    ctx = B.Context(device=0, bits=64, w=10, p=100, get_sai=True)
    ctx.add_records(host.data_ptr(), np.array([nb], dtype=np.uint64))
    ctx.finalize(); ctx.bwt_of_parse(); ctx.release_text(); ctx.generate_bwt(sa=True, rssa=False)
    n = int(ctx.sizes().n)
    out = os.path.join(a.dir, "pfb_bench.ma")
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        ctx.write_marker_array(mps, out)
        ts.append(time.perf_counter() - t0)
    ma_bytes = os.path.getsize(out)
    # reference and the drop-in tool on a prefix of the SA (file -> file)
    m = min(a.ref_sample, n + 1)
    sa = ctx.sa()[:m]
    ctx.close()
    saf = os.path.join(a.dir, "pfb_bench.sa")
    sa.tofile(saf)
    ref_tool = os.path.join(ROOT, "oracle", "_ref", "mps_to_ma")
    line = {"what": "marker-array join (mps_to_ma): SA in HBM -> .ma file", "bases": nb, "sa_entries": n + 1, "marker_runs": R, "mps_bytes": int(words.nbytes),
            "ma_bytes": int(ma_bytes), "join_s_best_of_3": min(ts), "join_s_all": ts, "entries_per_s": (n + 1) / min(ts)}
    if os.path.exists(ref_tool):
        t0 = time.perf_counter()
        subprocess.run([ref_tool, "-o", out + ".ref", mps, saf], check=True)
        tr = time.perf_counter() - t0
        t0 = time.perf_counter()
        subprocess.run([os.path.join(ROOT, "bin", "mps_to_ma"), "-o", out + ".gpu", mps, saf], check=True)
        tg = time.perf_counter() - t0
        same = open(out + ".ref", "rb").read() == open(out + ".gpu", "rb").read()
        line.update({"sample_entries": m, "reference_tool_s": tr, "reference_entries_per_s": m / tr, "bin_mps_to_ma_s_incl_cuda_start": tg, "sample_outputs_identical": bool(same)})
        for e in (".ref", ".gpu"):
            os.unlink(out + e)
    print(json.dumps(line))
    for f in (mps, out, saf):
        if os.path.exists(f):
            os.unlink(f)
    return 0


if __name__ == "__main__":
    sys.exit(main())
