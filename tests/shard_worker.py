"""One rank of a world_size-N job whose exchanges go through torch.distributed (gloo) -- the host
all-gather the library's HostCallComm calls back into (pfbwt-f_b200/csrc/comm.hpp).

  python shard_worker.py emu  <rank> <world> <port> <out> <w> <p> <bits> <sa> <rssa> <fasta_r>
      serial logic emulator (tests/emu/libpfb_emu.so): CPU only, `-m "not gpu"` tests
  python shard_worker.py cuda <rank> <world> <port> <out> <w> <p> <bits> <sa> <rssa> <fasta_r>
      libpfbwt_b200.so on cuda:0 (all ranks share the one GPU of the test box), `-m gpu` tests
  python shard_worker.py nccl ...
      libpfbwt_b200.so on cuda:<rank>, exchanges over NCCL (needs <world> GPUs)
"""
import ctypes
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

ALLGATHERV = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                              ctypes.POINTER(ctypes.c_uint64), ctypes.c_int)


def make_allgatherv():
    """host all-gather-v over the default process group: pad to the longest contribution"""
    def fn(user, send, recv, counts, world):
        try:
            cs = [int(counts[r]) for r in range(world)]
            rank = dist.get_rank()
            mx = max(max(cs), 1)
            mine = torch.zeros(mx, dtype=torch.uint8)
            if cs[rank]:
                src = np.ctypeslib.as_array(ctypes.cast(send, ctypes.POINTER(ctypes.c_uint8)), shape=(cs[rank],))
                mine[:cs[rank]] = torch.from_numpy(src.copy())
            parts = [torch.zeros(mx, dtype=torch.uint8) for _ in range(world)]
            dist.all_gather(parts, mine)
            tot = sum(cs)
            if tot:
                dst = np.ctypeslib.as_array(ctypes.cast(recv, ctypes.POINTER(ctypes.c_uint8)), shape=(tot,))
                o = 0
                for r in range(world):
                    dst[o:o + cs[r]] = parts[r][:cs[r]].numpy()
                    o += cs[r]
            return 0
        except Exception as e:  # noqa: BLE001 -- must not unwind through the C frame
            sys.stderr.write(f"allgatherv callback failed: {e!r}\n")
            return 1
    return ALLGATHERV(fn)


def main():
    mode, rank, world, port, out, w, p, bits, sa, rssa, fasta = sys.argv[1:12]
    rank, world, w, p, bits, sa, rssa = int(rank), int(world), int(w), int(p), int(bits), int(sa), int(rssa)
    rng = None                     # "<path>@<begin>:<end>": this rank takes a byte range of a common FASTA file (mid-record cuts)
    if "@" in fasta:
        fasta, r_ = fasta.rsplit("@", 1)
        rng = tuple(int(x) for x in r_.split(":"))
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    cb = make_allgatherv()
    if mode == "emu":
        lib = ctypes.CDLL(os.path.join(ROOT, "tests", "emu", "libpfb_emu.so"))
        info = (ctypes.c_ulonglong * 8)()
        lib.pfb_emu_shard_run.argtypes = [ctypes.c_int, ctypes.c_int, ALLGATHERV, ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int,
                                          ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_char_p,
                                          ctypes.POINTER(ctypes.c_ulonglong), ctypes.c_ulonglong, ctypes.c_ulonglong]
        rc = lib.pfb_emu_shard_run(rank, world, cb, None, fasta.encode(), w, p, bits, sa, rssa, out.encode(), info,
                                   rng[0] if rng else 0, rng[1] if rng else 0)
    else:
        from pfbwt_b200_loader import binding
        b = binding()
        ctx = b.Context(device=rank if mode == "nccl" else 0, bits=bits, w=w, p=p, get_sai=bool(sa or rssa))
        if mode == "nccl":      # one GPU per rank, exchanges over NCCL; the id travels through the gloo group
            ids = [b.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            ctx.shard_init_nccl(rank, world, ids[0])
        else:
            ctx.shard_init_host(rank, world, cb)
        rc = 0
        try:
            if rng:
                ctx.add_fasta_range(fasta, rng[0], rng[1])
            elif os.path.exists(fasta + ".dict") and os.path.exists(fasta + ".parse"):      # a pre-parsed part: merge_pfp's rule
                part = b.Context(device=rank if mode == "nccl" else 0, bits=bits, w=w, p=p, get_sai=bool(sa or rssa))
                part.load_parser(fasta, False, False)
                ctx.append(part)
                part.close()
            else:
                ctx.add_fasta(fasta)
            ctx.finalize()
            ctx.bwt_of_parse()
            if rank == 0:
                ctx.save_parser(out, True, bool(sa or rssa))
            ctx.generate_bwt(sa=bool(sa), rssa=bool(rssa))
            ctx.write_bwt(out, bool(sa), bool(rssa))
        except Exception as e:  # noqa: BLE001
            sys.stderr.write(f"[rank {rank}] {e}\n")
            rc = 1
        ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if rc else 0)


if __name__ == "__main__":
    main()
