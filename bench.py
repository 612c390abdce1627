#!/usr/bin/env python
"""bench.py -- headline benchmark of the fasta -> .bwt hot path (BASELINE.json metric:
input bases/sec end-to-end at 1/2/4/8 B200 + fraction of the HBM roofline).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--bases B]

Workload (config.workload): BASELINE.json configs[1] = synthetic uniform-random ACGT single
sequence, -w 10 -p 100, 32-bit mode, .bwt only; 1 Gbp per GPU by default (--bases overrides, e.g.
for quick checks).  A "step" is one pass of the whole hot path (parse -> parse BWT -> dictionary
suffix sort -> BWT assembly) over one text.

  value      bases/s with the text already resident in HBM (pfb_run_resident), all ranks summed
  e2e        bases/s through the C ABI with HOST buffers: H2D of the record bytes from pinned
             memory and D2H of the .bwt image inside the timed region
  roofline   dominant kernel k_radix_pass (one launch per 8-bit digit of the dictionary suffix sort):
             algorithmic bytes D*(1+U) per launch divided by the average CUDA-event time
             of a launch, against MEASURED_PEAKS.json hbm_gbs; moved_gbs = the 24 B/pair a pass really moves
  cpu_baseline  unmodified reference binary (oracle/_ref/pfbwt-f) on a bounded sample, 1 core

  single_job ONE job of the BASELINE.json configs[4] family over all ranks (N x 10 haplotypes of 100 Mbp, 0.1 %
             substitutions, pfbwt-f64 -s): piece-wise parse, NCCL all-gather merge, key-range partitioned suffix sort,
             every rank emits one slice of .bwt/.sa.  Weak scaling; at N=1 the same workload without a communicator.
             `value` = text resident in HBM, `e2e` = from pinned host buffers incl. D2H of the BWT and SA slices.

  file_to_file (N=1) the same text as a 60-column FASTA file on tmpfs through bin/pfbwt-f until every output file is closed
             (process wall clock; SURVEY.md 8(d)'s user-level metric).  --no-file-leg skips it.

--impl reference times the reference's own CPU implementation (oracle/_ref/pfbwt-f) on a bounded
sample of the same workload (rank 0 only).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W, PMOD = 10, 100
LUT = np.frombuffer(b"ACGT", dtype=np.uint8)


def gen_bases(out, seed):
    """uniform ACGT (config 2 recipe: default_rng(seed), lut[rng.integers(0,4,...)]) into `out`."""
    rng = np.random.default_rng(seed)
    n = len(out)
    step = 50_000_000
    for s in range(0, n, step):
        m = min(step, n - s)
        out[s:s + m] = LUT[rng.integers(0, 4, m, dtype=np.uint8)]


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.stop_flag = False
        self.max_mhz = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                   stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=5).stdout.decode().strip()
                f = [x.strip() for x in o.split(",")]
                self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for nm, v in zip(names, f[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def run_reference_sample(bases, sample_n):
    """time the unmodified reference pfbwt-f on the first sample_n bases; returns (bases/s, seconds)"""
    exe = os.path.join(ROOT, "oracle", "_ref", "pfbwt-f")
    kind = "reference"
    if not os.path.exists(exe):
        exe = os.path.join(ROOT, "oracle", "_ref", "pfp_oracle")
        kind = "port"
        if not os.path.exists(exe):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, stdout=subprocess.DEVNULL)
    with tempfile.TemporaryDirectory() as td:
        fa = os.path.join(td, "sample.fa")
        with open(fa, "wb") as f:
            f.write(b">sample\n")
            f.write(bases[:sample_n].tobytes())
            f.write(b"\n")
        t0 = time.perf_counter()
        subprocess.run([exe, "-w", str(W), "-p", str(PMOD), "-o", os.path.join(td, "out"), fa], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        dt = time.perf_counter() - t0
    return sample_n / dt, dt, kind


def file_to_file_leg(bases, runs=2, line_width=60):
    """the metric as a user of the binaries sees it - config 2 as a FASTA file with 60-column lines on
    tmpfs -> bin/pfbwt-f -> every output file closed; wall clock of the whole process (CUDA start-up and exit included).
    One untimed run first (page cache, driver), then `runs` timed ones."""
    exe = os.path.join(ROOT, "bin", "pfbwt-f")
    if not os.path.exists(exe):
        return {"unavailable": "bin/pfbwt-f not built"}
    n = int(bases.shape[0])
    d = None
    for cand in ("/dev/shm", tempfile.gettempdir()):
        try:
            sv = os.statvfs(cand)
            if os.access(cand, os.W_OK) and sv.f_bavail * sv.f_frsize > 4 * n + (1 << 30):
                d = cand
                break
        except OSError:
            pass
    if d is None:
        return {"unavailable": "no scratch directory with room for the FASTA and the output files"}
    with tempfile.TemporaryDirectory(dir=d) as td:
        fa = os.path.join(td, "config2.fa")
        with open(fa, "wb") as f:
            f.write(b">rand1g\n")
            step = 60_000_000 // line_width * line_width
            for s0 in range(0, n, step):
                x = bases[s0:s0 + step]
                full = x.shape[0] // line_width * line_width
                if full:
                    a = np.empty((full // line_width, line_width + 1), dtype=np.uint8)
                    a[:, :line_width] = x[:full].reshape(-1, line_width)
                    a[:, line_width] = 10
                    f.write(a.tobytes())
                if x.shape[0] > full:
                    f.write(x[full:].tobytes() + b"\n")
        fbytes = os.path.getsize(fa)
        walls, phases = [], {}
        env = dict(os.environ, PFB_CLI_TIMES="1")
        for it in range(runs + 1):
            t0 = time.perf_counter()
            r = subprocess.run([exe, "-w", str(W), "-p", str(PMOD), "-o", os.path.join(td, "out"), fa], env=env,
                               stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
            dt = time.perf_counter() - t0
            if r.returncode != 0:
                return {"unavailable": "bin/pfbwt-f failed: " + r.stderr.decode()[-200:]}
            if it:
                walls.append(dt)
            phases = {}
            for l in r.stderr.decode().splitlines():
                f3 = l.split("\t")
                if len(f3) >= 3 and f3[0] in ("TASK", "TIME"):
                    try:
                        phases[f3[1]] = float(f3[2].rstrip("s"))
                    except ValueError:
                        pass
    wall = sum(walls) / len(walls)
    return {"value": n / wall, "unit": "bases/s", "wall_s": wall, "runs": len(walls), "fasta_bytes": fbytes,
            "what": f"bin/pfbwt-f -w {W} -p {PMOD} on the config-2 text as a {line_width}-column FASTA in {d}; process wall clock until exit, "
                    ".dict .occ .parse .bwlast .ilist .bwt .n written",
            "phases_s_last_run": phases}


# ---- single job over all ranks (BASELINE.json configs[4] family) ----------------------
JOB_HAPS, JOB_BASE, JOB_SNP_RATE = 10, 100_000_000, 0.001


def gen_job_text(out, rank, haps, base_len):
    """rank's records of the pangenome job: `haps` haplotypes of one base sequence (default_rng(2024)),
    global haplotype H>0 carries 0.1 % substitutions at default_rng(5000+H) positions (drawn with replacement)."""
    codes = np.random.default_rng(2024).integers(0, 4, base_len, dtype=np.uint8)
    nsnp = int(base_len * JOB_SNP_RATE)
    for h in range(haps):
        H = rank * haps + h
        c = codes
        if H > 0:
            rng = np.random.default_rng(5000 + H)
            pos = rng.integers(0, base_len, nsnp)
            c = codes.copy()
            c[pos] = (c[pos] + rng.integers(1, 4, nsnp, dtype=np.uint8)) % 4
        out[h * base_len:(h + 1) * base_len] = LUT[c]


def single_job_leg(B, torch, dist, rank, local_rank, world, steps, warmup, haps, base_len, with_sa, do_e2e):
    """One job over `world` GPUs: rank g holds haplotypes [g*haps, (g+1)*haps); pfbwt-f64 semantics
    (-s when with_sa).  Returns the dict reported as "single_job" (rank 0) or None."""
    n_rank = haps * base_len
    host = torch.empty(n_rank, dtype=torch.uint8).pin_memory()
    gen_job_text(host.numpy(), rank, haps, base_len)
    rec_len = np.full(haps, base_len, dtype=np.uint64)
    ctx = B.Context(device=local_rank, bits=64, w=W, p=PMOD, get_sai=with_sa)
    if world > 1:
        g = dist.new_group(backend="gloo")
        ids = [B.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0, group=g)
        ctx.shard_init_nccl(rank, world, ids[0])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx.add_records(host.data_ptr(), rec_len)
    ctx.upload_text()
    for _ in range(warmup):
        ctx.run_resident(with_sa, False)
    st0 = ctx.stats()
    barrier()
    dev_ms, acc = 0.0, {}
    t0 = time.perf_counter()
    for _ in range(steps):
        ctx.run_resident(with_sa, False)
        st = ctx.stats()
        dev_ms += st["run_ms"]
        for k, v in st.items():
            if k.endswith("_ms") and k not in ("run_ms", "radix_pass_ms"):
                acc[k] = acc.get(k, 0.0) + v
    barrier()
    t_wall = time.perf_counter() - t0
    st1 = ctx.stats()
    sz, info = ctx.sizes(), ctx.shard_info()
    t_e2e = float("nan")
    h2d = d2h = 0
    if do_e2e:
        out_bwt = torch.empty(info.bwt_len + (info.bwt_len >> 2) + 4096, dtype=torch.uint8).pin_memory()
        out_sa = torch.empty((info.bwt_len + (info.bwt_len >> 2) + 4096) if with_sa else 1, dtype=torch.int64).pin_memory()

        def e2e_step():
            ctx.clear()
            ctx.add_records(host.data_ptr(), rec_len)
            ctx.finalize()
            ctx.bwt_of_parse()
            ctx.generate_bwt(with_sa, False)
            m = ctx.shard_info().bwt_len
            ctx.bwt(out=out_bwt.numpy()[:m])
            if with_sa:
                ctx.sa(out=out_sa.numpy()[:m].view(np.uint64))
            return m

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            m = e2e_step()
        barrier()
        t_e2e = time.perf_counter() - t0
        h2d, d2h = n_rank, m * (9 if with_sa else 1)
    t = torch.tensor([dev_ms * 1e-3, t_wall, t_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t_dev, t_wall, t_e2e = float(t[0]), float(t[1]), float(t[2])
    ctx.close()
    if rank != 0:
        return None
    n_tot = world * n_rank
    return {
        "workload": f"config5 family: ONE job, {world * haps} haplotypes x {base_len / 1e6:g} Mbp (0.1% substitutions), {haps} per GPU, "
                    f"pfbwt-f64 -w {W} -p {PMOD}" + (" -s" if with_sa else ""),
        "scaling": "weak", "value": n_tot * steps / t_dev, "unit": "bases/s", "ms_per_step": 1e3 * t_dev / steps,
        "wall_ms_per_step": 1e3 * t_wall / steps,
        "e2e": {"value": n_tot * steps / t_e2e if do_e2e else None, "unit": "bases/s", "ms_per_step": 1e3 * t_e2e / steps if do_e2e else None,
                "h2d_bytes_per_step_rank0": int(h2d), "d2h_bytes_per_step_rank0": int(d2h)},
        "n": int(sz.n), "phrases": int(sz.n_phrases), "dict_words": int(sz.n_words), "dict_bytes": int(sz.dict_bytes), "r": int(sz.r),
        "rank0_slice": {"bwt_off": int(info.bwt_off), "bwt_len": int(info.bwt_len)},
        "exchange": {"transport": "NCCL all-gather / grouped broadcast over NVLink" if world > 1 else "none",
                     "bytes_received_rank0_per_step": int((st1["shard_bytes_moved"] - st0["shard_bytes_moved"]) / max(1, steps)),
                     "exchange_ms_rank0": acc.get("shard_exchange_ms", 0.0) / steps, "merge_ms_rank0": acc.get("shard_merge_ms", 0.0) / steps},
        "stages_ms_rank0": {k: v / steps for k, v in acc.items()},
        "gpu_launches_rank0": int(st1["launches_own"] - st0["launches_own"]),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bases", type=int, default=1_000_000_000)
    ap.add_argument("--ref-sample", type=int, default=40_000_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling aid: skip the host-buffer leg")
    ap.add_argument("--no-single-job", action="store_true", help="skip the one-job-over-all-GPUs leg")
    ap.add_argument("--no-file-leg", action="store_true", help="skip the file-to-file run of bin/pfbwt-f (N=1 only)")
    ap.add_argument("--job-haps", type=int, default=JOB_HAPS)
    ap.add_argument("--job-base", type=int, default=JOB_BASE)
    ap.add_argument("--job-no-sa", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n = args.bases
    workload = f"config2: synthetic {n / 1e9:g} Gbp uniform-random ACGT single sequence per GPU, -w {W} -p {PMOD}, 32-bit mode, .bwt only"

    if args.impl == "reference":
        if rank != 0:
            return 0
        sample_n = min(args.ref_sample, n)
        bases = np.empty(sample_n, dtype=np.uint8)
        gen_bases(bases, 1)
        for _ in range(min(args.warmup, 1)):
            run_reference_sample(bases, min(sample_n, 2_000_000))
        t_tot, kind = 0.0, "reference"
        for _ in range(args.steps):
            _, dt, kind = run_reference_sample(bases, sample_n)
            t_tot += dt
        v = sample_n * args.steps / t_tot
        line = {"impl": "reference", "metric": "input bases/sec end-to-end fasta->.bwt", "value": v, "unit": "bases/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/u32/u64 integer",
                "data": "synthetic", "config": {"workload": workload},
                "cpu_baseline": {"value": v, "unit": "bases/s", "cores": 1, "kind": kind,
                                 "sample": f"first {sample_n} bases of the workload text, whole pfbwt-f run incl. file I/O"},
                "e2e": {"value": v, "unit": "bases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist
    from pfbwt_b200_loader import binding
    B = binding()

    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)

    # synthetic text of this rank in pinned host memory (rank r uses seed 1+r)
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    bases = host.numpy()
    gen_bases(bases, 1 + rank)
    out_host = torch.empty(n + W + 1, dtype=torch.uint8).pin_memory()
    rec_len = np.array([n], dtype=np.uint64)

    ctx = B.Context(device=local_rank, bits=32, w=W, p=PMOD)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- value: text resident in HBM ---------------------------------------------
    ctx.add_records(host.data_ptr(), rec_len)
    ctx.upload_text()
    for _ in range(args.warmup):
        ctx.run_resident(False, False)
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = ctx.stats()
    barrier()
    t0 = time.perf_counter()
    stage_acc = {}
    dev_ms = 0.0
    for _ in range(args.steps):
        ctx.run_resident(False, False)
        st = ctx.stats()
        dev_ms += st["run_ms"]                      # CUDA events on the library's own stream
        for k, v in st.items():
            if k.endswith("_ms") and k not in ("run_ms", "radix_pass_ms"):
                stage_acc[k] = stage_acc.get(k, 0.0) + v
    barrier()
    t_res = time.perf_counter() - t0
    st1 = ctx.stats()
    sz = ctx.sizes()
    D = sz.dict_bytes
    launches = (st1["launches_own"] - launches0["launches_own"])
    launches_lib = (st1["launches_lib"] - launches0["launches_lib"])

    # ---------------- e2e: host buffers, H2D + D2H inside the timed region -----------------------
    def e2e_step():
        ctx.clear()
        ctx.add_records(host.data_ptr(), rec_len)      # H2D from pinned memory
        ctx.finalize()
        ctx.bwt_of_parse()
        ctx.generate_bwt(False, False)
        ctx.bwt(out=out_host.numpy()[:n + W + 1])         # D2H of the .bwt image (n+w+1 bytes)

    t_e2e = float("nan")
    if not args.no_e2e:
        for _ in range(max(1, min(args.warmup, 2))):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        barrier()
        t_e2e = time.perf_counter() - t0
    sampler.stop_flag = True
    sampler.join()

    # max over ranks
    t_dev = dev_ms * 1e-3
    if world > 1:
        t = torch.tensor([t_res, t_e2e, t_dev], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_res, t_e2e, t_dev = float(t[0]), float(t[1]), float(t[2])

    ctx.close()
    del host, out_host
    job = None
    if not args.no_single_job:
        job = single_job_leg(B, torch, dist, rank, local_rank, world, max(1, min(args.steps, 3)), max(1, min(args.warmup, 2)),
                             args.job_haps, args.job_base, not args.job_no_sa, not args.no_e2e)

    if rank == 0:
        peak, peak_kind = hbm_peak()
        stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
        # dominant kernel: k_radix_pass (5 launches of the dictionary suffix sort per step).  Algorithmic
        # bytes per launch = SURVEY 8(d) per-unit figure of the dictionary suffix sort (1+U bytes per
        # suffix) x the D suffixes one launch processes; duration = CUDA events around every launch.
        U = 4
        npass = st1["radix_passes"] - launches0["radix_passes"]
        pass_ms = (st1["radix_pass_ms"] - launches0["radix_pass_ms"]) / max(1, npass)
        pairs = (st1["radix_pairs"] - launches0["radix_pairs"]) / max(1, npass)
        alg_bytes = pairs * (1 + U)
        achieved = alg_bytes / (pass_ms * 1e-3) / 1e9 if pass_ms else 0.0
        traffic = None
        tp = os.path.join(ROOT, "profiles", "radix_pass_traffic.json")   # dram bytes per launch from the committed ncu --set full capture
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp))["dram_bytes_per_launch"]
            except Exception:
                traffic = None
        cpu = None
        if not args.no_cpu_baseline and world == 1:        # reported at N=1 only (rank 0's host cores)
            sample_n = min(args.ref_sample, n)
            v, dt, kind = run_reference_sample(bases, sample_n)
            cpu = {"value": v, "unit": "bases/s", "cores": 1, "kind": kind,
                   "sample": f"first {sample_n} bases of rank 0's text, whole pfbwt-f -w {W} -p {PMOD} run incl. file I/O ({dt:.1f} s)"}
        f2f = None
        if not args.no_file_leg and world == 1:
            try:
                f2f = file_to_file_leg(bases)
            except Exception as e:      # the leg is an extra: never lose the bench line over a full disk
                f2f = {"unavailable": repr(e)[:200]}
        line = {
            "metric": "input bases/sec end-to-end fasta->.bwt",
            "value": world * n * args.steps / t_dev, "unit": "bases/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t_dev / args.steps, "wall_ms_per_step": 1e3 * t_res / args.steps,
            "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8/u32/u64 integer", "data": "synthetic",
            "config": {"workload": workload, "n": int(sz.n), "phrases": int(sz.n_phrases), "dict_words": int(sz.n_words),
                       "dict_bytes": int(D), "r": int(sz.r), "l2": "inputs larger than L2 (text 1 GB, dictionary 1.1 GB per step)",
                       "multi_gpu": ("value/e2e: independent config-2 text per rank, no data-path collective; single_job: ONE job over all "
                                     "ranks (piece-wise parse, NCCL all-gather merge, key-range partitioned suffix sort)") if world > 1 else "single GPU"},
            "e2e": {"value": world * n * args.steps / t_e2e, "unit": "bases/s", "h2d_bytes_per_step": int(n),
                    "d2h_bytes_per_step": int(n + W + 1), "ms_per_step": 1e3 * t_e2e / args.steps},
            "gpu_launches": int(launches), "gpu_launches_library_cub": int(launches_lib),
            "roofline": {"bound": "hbm", "kernel": "k_radix_pass (dictionary suffix sort, one 8-bit digit per launch)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs, burst copy)", "algorithmic_bytes": int(alg_bytes),
                         "ms": pass_ms, "launches_timed": int(npass),
                         "moved_bytes": int(pairs * 24), "moved_gbs": pairs * 24 / (pass_ms * 1e-3) / 1e9 if pass_ms else 0.0,
                         "share_of_step": pass_ms * npass / args.steps / (1e3 * t_dev / args.steps) if t_dev else None},
            "stages_ms": stage_ms,
            "clocks": sampler.summary(),
            "dev_bytes_peak": int(st1["dev_bytes_peak"]),
        }
        if cpu:
            line["cpu_baseline"] = cpu
        if job:
            line["single_job"] = job
        if f2f:
            line["file_to_file"] = f2f
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
