#!/usr/bin/env python
"""The north-star run (BASELINE.json: "a byte-identical .bwt/.sa for an 8 Gbp synthetic multi-haplotype fasta produced
end-to-end on 8xB200"), FROM FASTA FILES TO OUTPUT FILES, timed and verified:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
      tools/north_star.py [--haps 10] [--base 100000000] [--dir /dev/shm] [--split-one-file]

Rank g's input is a 60-column FASTA file with its `haps` haplotypes (config-5 family: N x haps haplotypes of one base
sequence with 0.1 % substitutions); with --split-one-file there is ONE FASTA file and rank g takes a byte range of it
(pfb_add_fasta_range: cuts fall inside records).  The timed region is what pfbwt-f_b200/multi.py does - pfb_add_fasta on
GPU g, collective finalize / parse BWT / BWT build with full 64-bit SA, every rank writes its slice of OUT.bwt / OUT.sa at
its offset - between two barriers, wall clock, max over ranks.  Afterwards (untimed) the OUTPUT FILES are read back slice by
slice and checked with the checker of bench.py: every (SA[i], BWT[i]) pair against the text, SA a permutation, sampled
adjacent suffixes in order, run count = r.  One JSON line on rank 0."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def write_fasta_records(path, recs, names, mode="wb", lw=60):
    with open(path, mode) as f:
        for nm, x in zip(names, recs):
            f.write(b">" + nm.encode() + b"\n")
            full = x.shape[0] // lw * lw
            a = np.empty((full // lw, lw + 1), dtype=np.uint8)
            a[:, :lw] = x[:full].reshape(-1, lw)
            a[:, lw] = 10
            f.write(a.tobytes())
            if x.shape[0] > full:
                f.write(x[full:].tobytes() + b"\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--haps", type=int, default=10)
    ap.add_argument("--base", type=int, default=100_000_000)
    ap.add_argument("--dir", default="/dev/shm")
    ap.add_argument("--split-one-file", action="store_true")
    ap.add_argument("--keep", action="store_true")
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    import bench
    from pfbwt_b200_loader import binding
    B = binding()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- input files (untimed) -------------------------------------------------------------------------
    codes = bench.job_codes(a.base)
    mine = np.empty(a.haps * a.base, dtype=np.uint8)
    bench.gen_job_text(mine, rank, a.haps, a.base, codes)
    recs = [mine[h * a.base:(h + 1) * a.base] for h in range(a.haps)]
    names = [f"hap{rank * a.haps + h}" for h in range(a.haps)]
    out = os.path.join(a.dir, "pfb_north_star_out")
    one = os.path.join(a.dir, "pfb_north_star_all.fa")
    part = os.path.join(a.dir, f"pfb_north_star_part{rank}.fa")
    if a.split_one_file:
        for r in range(world):                      # ranks append their records in turn
            if r == rank:
                write_fasta_records(one, recs, names, mode="wb" if r == 0 else "ab")
            barrier()
        fsize = os.path.getsize(one)
        in_bytes = fsize
    else:
        write_fasta_records(part, recs, names)
        in_bytes = os.path.getsize(part) * world
    del mine, recs
    for ext in ("bwt", "sa"):
        if rank == 0 and os.path.exists(out + "." + ext):
            os.unlink(out + "." + ext)
    ctx = B.Context(device=local_rank, bits=64, w=bench.W, p=bench.PMOD, get_sai=True)
    if world > 1:
        ids = [B.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.shard_init_nccl(rank, world, ids[0])
    barrier()

    # ---- timed: files -> files --------------------------------------------------------------------------
    t0 = time.perf_counter()
    if a.split_one_file:
        ctx.add_fasta_range(one, fsize * rank // world, fsize * (rank + 1) // world)
    else:
        ctx.add_fasta(part)
    t_in = time.perf_counter() - t0
    ctx.finalize()
    ctx.bwt_of_parse()
    ctx.release_text()
    ctx.generate_bwt(sa=True, rssa=False)
    torch.cuda.synchronize()
    t_dev = time.perf_counter() - t0
    ctx.write_bwt(out, True, False)
    barrier()
    t_all = time.perf_counter() - t0
    sz, info = ctx.sizes(), ctx.shard_info()
    ctx.close()
    tt = torch.tensor([t_in, t_dev, t_all], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_in, t_dev, t_all = (float(x) for x in tt.tolist())

    # ---- verification of the FILES (untimed) ---------------------------------------------------------------
    n = int(sz.n)
    off, ln = int(info.bwt_off), int(info.bwt_len)
    with open(out + ".bwt", "rb") as f:
        f.seek(off)
        bw = np.frombuffer(f.read(ln), dtype=np.uint8)
    with open(out + ".sa", "rb") as f:
        f.seek(off * 8)
        sa = np.frombuffer(f.read(ln * 8), dtype=np.uint64)
    sizes_ok = os.path.getsize(out + ".bwt") == n + 1 and os.path.getsize(out + ".sa") == 8 * (n + 1)
    text_at = bench.job_text_evaluator(torch, codes, world * a.haps, a.base)
    res = bench.verify_slices(torch, dist, rank, world, n, bw, sa, off, int(sz.r), text_at)
    if rank == 0:
        line = {"what": "north star: FASTA files -> .bwt + 64-bit .sa files, ONE job over all GPUs (pfbwt-f64 -s semantics)",
                "n_gpus": world, "input": ("one FASTA split by byte ranges (mid-record cuts)" if a.split_one_file else "one FASTA per rank"),
                "haplotypes": world * a.haps, "bases": world * a.haps * a.base, "n": n, "r": int(sz.r), "phrases": int(sz.n_phrases),
                "dict_bytes": int(sz.dict_bytes), "input_bytes": int(in_bytes), "output_bytes": int(9 * (n + 1)),
                "wall_s": t_all, "bases_per_s": world * a.haps * a.base / t_all,
                "phases_s_max_over_ranks": {"read + split records on device": t_in, "until BWT + SA in HBM": t_dev, "until all files written": t_all},
                "files_verified": dict(res, output_sizes_ok=bool(sizes_ok)), "ok": bool(res["ok"] and sizes_ok), "dir": a.dir}
        print(json.dumps(line))
    barrier()
    if not a.keep and rank == 0:
        for p_ in [out + ".bwt", out + ".sa", one]:
            if os.path.exists(p_):
                os.unlink(p_)
    if not a.keep and os.path.exists(part):
        os.unlink(part)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
