#!/usr/bin/env python
"""One PFP-BWT job over the GPUs of a box: the multi-GPU counterpart of
`merge_pfp -o out part0 .. partN-1 && pfbwt-f64 --pfbwt-only -o out` (reference src/merge_pfp.cpp:97-170,
src/pfbwt-f.cpp:351-367), without the per-part files in between.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
      pfbwt-f_b200/multi.py [-w W] [-p P] [-s] [-r] [--bits 32|64] [--non-acgt-to-a] [--save-parse] -o OUT in0.fa [in1.fa ...]

The input files are dealt to the ranks in order, as contiguous groups (rank order = text order).  With fewer files than
ranks (typically ONE multi-record or single-record FASTA) every file is instead cut into byte ranges, one per rank of its
share of the ranks (pfb_add_fasta_range: the cuts fall anywhere, also inside a record - the w-overlapping chunk split).  Rank g parses its files on GPU g; the parses are merged over NCCL; every rank builds and writes its
slice of OUT.bwt (/.sa/.ssa/.esa) at its offset.  An input that names a PRE-PARSED prefix (<in>.dict and <in>.parse exist, written
by `pfbwt-f64 --parse-only`) is loaded instead of parsed - exactly merge_pfp's rule (src/merge_pfp.cpp:100-113 load_or_generate_parser):
`merge_pfp -t N` then is this launcher on N GPUs, each rank holding only ITS parts.  Host code only plumbs: all work is in
libpfbwt_b200.so (include/pfbwt_b200.h, "one job over several GPUs")."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-w", type=int, default=10)
    ap.add_argument("-p", type=int, default=100)
    ap.add_argument("-s", action="store_true", help="write OUT.sa")
    ap.add_argument("-r", action="store_true", help="write OUT.ssa / OUT.esa")
    ap.add_argument("--bits", type=int, default=64, choices=[32, 64])
    ap.add_argument("--non-acgt-to-a", action="store_true")
    ap.add_argument("--save-parse", action="store_true", help="rank 0 also writes OUT.dict/.occ/.parse/.bwlast/.ilist(/.bwsai)")
    ap.add_argument("-o", required=True)
    ap.add_argument("inputs", nargs="+")
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    from pfbwt_b200_loader import binding
    B = binding()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    preparsed = [os.path.exists(f + ".dict") and os.path.exists(f + ".parse") for f in a.inputs]
    split = len(a.inputs) < world
    if split and any(preparsed):
        if rank == 0:
            sys.stderr.write("pre-parsed prefixes cannot be cut into byte ranges: give at least one input per rank\n")
        return 1
    if split and world % len(a.inputs) != 0:
        if rank == 0:
            sys.stderr.write(f"{len(a.inputs)} input files cannot be split evenly over {world} ranks (use 1 file, or a divisor of the ranks)\n")
        return 1
    torch.cuda.set_device(local_rank)
    ctx = B.Context(device=local_rank, bits=a.bits, w=a.w, p=a.p, get_sai=a.s or a.r, non_acgt_to_a=a.non_acgt_to_a)
    if world > 1:
        dist.init_process_group("gloo")          # only carries the NCCL id; the data path is the library's own communicator
        ids = [B.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.shard_init_nccl(rank, world, ids[0])
    # contiguous, balanced, never empty (len(inputs) >= world): rank g takes files [n*g/world, n*(g+1)/world)
    nin = len(a.inputs)
    mine = a.inputs[nin * rank // world:nin * (rank + 1) // world]
    rc = 0
    try:
        if split:
            per = world // nin                         # ranks per file
            f, k = a.inputs[rank // per], rank % per
            size = os.path.getsize(f)
            ctx.add_fasta_range(f, size * k // per, size * (k + 1) // per)
        else:
            for f in mine:
                if os.path.exists(f + ".dict") and os.path.exists(f + ".parse"):
                    part = B.Context(device=local_rank, bits=a.bits, w=a.w, p=a.p, get_sai=a.s or a.r, non_acgt_to_a=a.non_acgt_to_a)
                    part.load_parser(f, False, False)
                    ctx.append(part)                       # PfParser::operator+= (include/pfparser.hpp:194-263)
                    part.close()
                else:
                    ctx.add_fasta(f)
    except B.PfbError as e:                        # keep going: the failure is agreed on inside finalize()
        sys.stderr.write(f"[rank {rank}] {e.msg}\n")
    try:
        ctx.finalize()
        ctx.bwt_of_parse()
        if a.save_parse and rank == 0:
            ctx.save_parser(a.o, True, a.s or a.r)
        ctx.generate_bwt(sa=a.s, rssa=a.r)
        ctx.write_bwt(a.o, a.s, a.r)
        sz, info = ctx.sizes(), ctx.shard_info()
        sys.stderr.write(f"[rank {rank}] n: {sz.n} r: {sz.r} slice: [{info.bwt_off}, {info.bwt_off + info.bwt_len})\n")
    except B.PfbError as e:
        sys.stderr.write(f"[rank {rank}] {e.msg}\n")
        rc = 1
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
