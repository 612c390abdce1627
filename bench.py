#!/usr/bin/env python
"""bench.py -- headline benchmark of the fasta -> .bwt hot path (BASELINE.json metric:
input bases/sec end-to-end at 1/2/4/8 B200 + fraction of the HBM roofline).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--bases B]

Workload (config.workload): BASELINE.json configs[1] = ONE synthetic 1 Gbp uniform-random ACGT single
sequence (NumPy default_rng(1)), -w 10 -p 100, 32-bit mode, .bwt only (--bases overrides the size for quick
checks).  A "step" is one pass of the whole hot path (parse -> parse BWT -> dictionary suffix sort -> BWT
assembly) over that text.  With N > 1 the SAME text is ONE job strong-scaled over the N ranks: rank g holds
bases [g n/N, (g+1) n/N) (a mid-record cut, pfb_add_chunk), the parses are merged over NCCL, every rank
sorts one key range of the dictionary suffixes and emits one slice of the .bwt.

  value      bases/s with the text already resident in HBM (pfb_run_resident; device time, max over ranks)
  e2e        bases/s through the C ABI with HOST buffers: H2D of the rank's bases from pinned memory and D2H
             of its .bwt slice inside the timed region
  parity     computed OUTSIDE the timed region: sha256[:16] of the concatenated .bwt slices against the hash the
             unmodified reference wrote for this text (4ab639c13f486fd2) plus n, r, P, D; for other
             --bases sizes an inversion / permutation / sampled-order / run-count check of a BWT+SA build instead
  roofline   dominant kernel k_radix_pass (one launch per 8-bit digit of the dictionary suffix sort):
             algorithmic bytes D*(1+U) per launch divided by the average CUDA-event time
             of a launch, against MEASURED_PEAKS.json hbm_gbs; moved_gbs = the 24 B/pair a pass really moves;
             stage = the same floor over the whole dictionary-suffix-sort stage
  cpu_baseline  unmodified reference binary (oracle/_ref/pfbwt-f) on a bounded sample, 1 core

  single_job ONE job of the BASELINE.json configs[4] family over all ranks (N x 10 haplotypes of 100 Mbp, 0.1 %
             substitutions, pfbwt-f64 -s).  Weak scaling; at N=1 the same workload without a communicator.
             `value` = text resident in HBM, `e2e` = from pinned host buffers incl. D2H of the BWT and SA slices.
             `parity`: every (SA[i], BWT[i]) of every slice checked against the text (inversion), SA is a permutation
             (mixing-hash multiset sum), sampled adjacent suffixes in order (ties followed to the end), run count = r,
             and a 100 Mbp two-record sub-job cut mid-record over the N ranks writes the same .bwt/.sa bytes as the
             same records on ONE GPU without a communicator (sha256 per slice).

  file_to_file (N=1) the same text as a 60-column FASTA file on tmpfs through bin/pfbwt-f until every output file is closed
             (process wall clock; SURVEY.md 8(d)'s user-level metric), without and with -s.  --no-file-leg skips it.

--impl reference times the reference's own CPU implementation (oracle/_ref/pfbwt-f) on a bounded
sample of the same workload (rank 0 only); config.sample_bases says how many bases that was.
"""
import argparse
import hashlib
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
# what the unmodified reference wrote for the full-size config-2 text
CONFIG2 = {"bases": 1_000_000_000, "n": 1_000_000_010, "phrases": 10_153_944, "dict_bytes": 1_105_921_870, "r": 749_996_069,
           "bwt_sha16": "4ab639c13f486fd2"}


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


def file_to_file_leg(bases, runs=2, line_width=60, with_sa=False):
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
            if os.access(cand, os.W_OK) and sv.f_bavail * sv.f_frsize > (8 if with_sa else 4) * n + (1 << 30):
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
            r = subprocess.run([exe, "-w", str(W), "-p", str(PMOD)] + (["-s"] if with_sa else []) + ["-o", os.path.join(td, "out"), fa], env=env,
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
            "what": f"bin/pfbwt-f -w {W} -p {PMOD}{' -s' if with_sa else ''} on the config-2 text as a {line_width}-column FASTA in {d}; process wall clock until exit, "
                    ".dict .occ .parse .bwlast .ilist .bwt .n" + (" .bwsai .sa" if with_sa else "") + " written",
            "phases_s_last_run": phases}




def gen_bases_range(seed, lo, hi, out):
    """bases [lo, hi) of the text gen_bases(., seed) writes (the generator's stream is sequential: earlier chunks are
    drawn and dropped)"""
    rng = np.random.default_rng(seed)
    step = 50_000_000
    s = 0
    while s < hi:
        m = step
        x = rng.integers(0, 4, m, dtype=np.uint8)
        a, b = max(lo, s), min(hi, s + m)
        if a < b:
            out[a - lo:b - lo] = LUT[x[a - s:b - s]]
        s += m


def sha16(buf):
    h = hashlib.sha256()
    mv = memoryview(buf)
    for o in range(0, len(mv), 1 << 26):
        h.update(mv[o:o + (1 << 26)])
    return h.hexdigest()[:16]


# ---- checker of a distributed BWT(+SA): size-independent properties, computed with torch on the GPUs, outside every
#      timed region.  text_at(pos) -> uint8 tensor of S at int64 positions pos in [0, n) (cuda). ------------------------
def _i64(c):
    return c - (1 << 64) if c >= (1 << 63) else c


def _mix(torch, x):
    x = x * _i64(0x9E3779B97F4A7C15)
    x = x ^ ((x >> 29) & 0x7FFFFFFFF)
    x = x * _i64(0xBF58476D1CE4E5B9)
    x = x ^ ((x >> 32) & 0xFFFFFFFF)
    return x


def verify_slices(torch, dist, rank, world, n, bwt_np, sa_np, off, r_reported, text_at, pairs=4096, seed=7):
    """bwt_np / sa_np: this rank's slice [off, off+len) of the BWT / SA of S.$ (|S| = n).  Returns a dict (same on all ranks)."""
    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")     # cpu: the checker's own unit test
    m = int(bwt_np.shape[0])
    bad_inv = 0
    hsum = torch.zeros(1, dtype=torch.int64, device=dev)
    psum = torch.zeros(1, dtype=torch.int64, device=dev)
    nruns = 0
    step = 1 << 26
    first_b = int(bwt_np[0]) if m else -1
    last_b = int(bwt_np[m - 1]) if m else -1
    first_sa = int(sa_np[0]) if m else -1
    last_sa = int(sa_np[m - 1]) if m else -1
    for s in range(0, m, step):
        e = min(m, s + step)
        sa = torch.from_numpy(sa_np[s:e].view(np.int64)).to(dev)
        bw = torch.from_numpy(bwt_np[s:e]).to(dev)
        pos = sa - 1
        term = pos < 0
        exp = text_at(torch.where(term, torch.zeros_like(pos), pos))
        exp = torch.where(term, torch.zeros_like(exp), exp)
        bad_inv += int((exp != bw).sum().item()) + int(((sa < 0) | (sa > n)).sum().item())
        hsum += _mix(torch, sa).sum()
        psum += _mix(torch, torch.arange(off + s, off + e, dtype=torch.int64, device=dev)).sum()
        nruns += int((bw[1:] != bw[:-1]).sum().item())
        if s:
            nruns += int(bwt_np[s] != bwt_np[s - 1])
        del sa, bw, pos, exp
    # neighbours' edge values (run boundaries and the adjacent pair across a slice border)
    edges = [None] * world
    mine = (m, first_b, last_b, first_sa, last_sa)
    if world > 1:
        dist.all_gather_object(edges, mine)
    else:
        edges = [mine]
    prev = None
    for g in range(rank - 1, -1, -1):
        if edges[g][0]:
            prev = edges[g]
            break
    if m:
        nruns += 1 if (prev is None and first_b != 0) or (prev is not None and prev[2] != first_b) else 0
    # sampled adjacent suffixes in order: S[SA[i]:] < S[SA[i+1]:], ties followed window by window to the end of the text
    rng = np.random.default_rng(seed + rank)
    K = min(pairs, max(0, m - 1))
    a_l, b_l = [], []
    if K:
        ii = rng.integers(0, m - 1, K)
        a_l.append(sa_np[ii].astype(np.int64))
        b_l.append(sa_np[ii + 1].astype(np.int64))
    if prev is not None and m:
        a_l.append(np.array([prev[4]], dtype=np.int64))
        b_l.append(np.array([first_sa], dtype=np.int64))
    bad_ord = unresolved = npairs = 0
    if a_l:
        a = torch.from_numpy(np.concatenate(a_l)).to(dev)
        b = torch.from_numpy(np.concatenate(b_l)).to(dev)
        npairs = int(a.shape[0])
        WN = 2048
        k = torch.arange(WN, dtype=torch.int64, device=dev)[None, :]
        for _round in range(40000):
            if a.shape[0] == 0:
                break
            pa, pb = a[:, None] + k, b[:, None] + k
            ca = torch.where(pa >= n, torch.full_like(pa, -1), text_at(pa.clamp(max=n - 1)).to(torch.int64))
            cb = torch.where(pb >= n, torch.full_like(pb, -1), text_at(pb.clamp(max=n - 1)).to(torch.int64))
            diff = ca != cb
            anyd = diff.any(dim=1)
            fi = diff.to(torch.int8).argmax(dim=1)
            ra = ca.gather(1, fi[:, None])[:, 0]
            rb = cb.gather(1, fi[:, None])[:, 0]
            bad_ord += int((anyd & (ra > rb)).sum().item())
            keep = ~anyd
            a, b = a[keep] + WN, b[keep] + WN
        unresolved = int(a.shape[0])
    tot = torch.tensor([bad_inv, bad_ord, unresolved, nruns, npairs, m], dtype=torch.int64, device=dev)
    hs = torch.cat([hsum, psum])
    if world > 1:
        dist.all_reduce(tot)
        # the mixing sums wrap mod 2^64: add them on the host from every rank's value
        parts = [None] * world
        dist.all_gather_object(parts, (int(hsum.item()), int(psum.item())))
        h_all = sum(p[0] for p in parts) & ((1 << 64) - 1)
        p_all = sum(p[1] for p in parts) & ((1 << 64) - 1)
    else:
        h_all, p_all = int(hs[0].item()) & ((1 << 64) - 1), int(hs[1].item()) & ((1 << 64) - 1)
    bad_inv, bad_ord, unresolved, nruns, npairs, mtot = (int(x) for x in tot.tolist())
    res = {"positions": mtot, "inversion_mismatches": bad_inv, "sa_is_permutation": h_all == p_all and mtot == n + 1,
           "sampled_pairs": npairs, "order_violations": bad_ord, "pairs_unresolved": unresolved,
           "runs_counted": nruns, "runs_reported": int(r_reported)}
    res["ok"] = (bad_inv == 0 and res["sa_is_permutation"] and bad_ord == 0 and unresolved == 0 and nruns == int(r_reported))
    return res


# ---- single job over all ranks (BASELINE.json configs[4] family) ----------------------
JOB_HAPS, JOB_BASE, JOB_SNP_RATE = 10, 100_000_000, 0.001


def job_codes(base_len):
    return np.random.default_rng(2024).integers(0, 4, base_len, dtype=np.uint8)


def job_snps(H, base_len):
    """(positions, new 2-bit codes' deltas) of global haplotype H > 0: 0.1 % substitutions at default_rng(5000+H)
    positions drawn with replacement (a position drawn twice keeps the later draw, as numpy's fancy assignment does)"""
    nsnp = int(base_len * JOB_SNP_RATE)
    rng = np.random.default_rng(5000 + H)
    pos = rng.integers(0, base_len, nsnp)
    delta = rng.integers(1, 4, nsnp, dtype=np.uint8)
    return pos, delta


def gen_job_text(out, rank, haps, base_len, codes=None):
    """rank's records of the pangenome job: `haps` haplotypes of one base sequence (default_rng(2024)),
    global haplotype H>0 carries 0.1 % substitutions."""
    if codes is None:
        codes = job_codes(base_len)
    for h in range(haps):
        H = rank * haps + h
        c = codes
        if H > 0:
            pos, delta = job_snps(H, base_len)
            c = codes.copy()
            c[pos] = (c[pos] + delta) % 4
        out[h * base_len:(h + 1) * base_len] = LUT[c]


def job_text_evaluator(torch, codes, nhaps, base_len):
    """text_at(pos) for S of the whole job (nhaps records of base_len bases, each followed by A^W) from the base sequence and
    the substitution table: no rank needs the other ranks' text."""
    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    base_t = torch.from_numpy(LUT[codes]).to(dev)
    keys, vals = [], []
    for H in range(1, nhaps):
        pos, delta = job_snps(H, base_len)
        rev = pos[::-1]
        up, first = np.unique(rev, return_index=True)              # first in the reversed order = last draw
        d = delta[::-1][first]
        keys.append(H * base_len + up.astype(np.int64))
        vals.append(LUT[(codes[up] + d) % 4])
    if keys:
        kt = torch.from_numpy(np.concatenate(keys)).to(dev)
        vt = torch.from_numpy(np.concatenate(vals)).to(dev)
    else:
        kt = torch.zeros(1, dtype=torch.int64, device=dev) - 1
        vt = torch.zeros(1, dtype=torch.uint8, device=dev)
    rec = base_len + W

    def text_at(pos):
        hap = torch.div(pos, rec, rounding_mode="floor")
        o = pos - hap * rec
        pad = o >= base_len
        c = base_t[o.clamp(max=base_len - 1)]
        key = hap * base_len + o
        ix = torch.searchsorted(kt, key).clamp(max=kt.shape[0] - 1)
        hit = (kt[ix] == key) & ~pad
        c = torch.where(hit, vt[ix], c)
        return torch.where(pad, torch.full_like(c, 65), c)
    return text_at


def sub_job_equality(B, torch, dist, rank, local_rank, world, ctx, codes, base_len, sub_len):
    """two records (haplotypes 0 and 1 cut to sub_len bases) as ONE job: cut anywhere - mid-record - over the `world` ranks
    through the communicator of `ctx`, and on rank 0 alone without a communicator; the slices must be the same bytes."""
    pos, delta = job_snps(1, base_len)
    c1 = codes.copy()
    c1[pos] = (c1[pos] + delta) % 4
    recs = [LUT[codes[:sub_len]], LUT[c1[:sub_len]]]
    tot = 2 * sub_len
    lo, hi = tot * rank // world, tot * (rank + 1) // world
    ctx.clear()
    for r in range(2):
        a, b = max(lo, r * sub_len), min(hi, (r + 1) * sub_len)
        if a < b:
            piece = np.ascontiguousarray(recs[r][a - r * sub_len:b - r * sub_len])
            ctx.add_chunk(piece, ends_record=(b == (r + 1) * sub_len))
    ctx.finalize()
    ctx.bwt_of_parse()
    ctx.generate_bwt(True, False)
    info = ctx.shard_info()
    bw, sa = ctx.bwt(), ctx.sa()
    mine = (int(info.bwt_off), int(info.bwt_len), sha16(bw), sha16(sa.view(np.uint8)))
    allv = [None] * world
    if world > 1:
        dist.all_gather_object(allv, mine)
    else:
        allv = [mine]
    res = None
    if rank == 0:
        one = B.Context(device=local_rank, bits=64, w=W, p=PMOD, get_sai=True)
        one.add_records(np.concatenate(recs), np.array([sub_len, sub_len], dtype=np.uint64))
        one.finalize()
        one.bwt_of_parse()
        one.generate_bwt(True, False)
        fb, fs = one.bwt(), one.sa()
        one.close()
        ok = sum(v[1] for v in allv) == fb.shape[0]
        for (o, l, hb, hs_) in allv:
            ok = ok and sha16(fb[o:o + l]) == hb and sha16(fs[o:o + l].view(np.uint8)) == hs_
        res = {"ok": bool(ok), "bases": int(tot), "records": 2, "ranks": world, "bwt_sha16_one_gpu": sha16(fb), "sa_sha16_one_gpu": sha16(fs.view(np.uint8)),
               "what": "same two records: mid-record cuts over all ranks vs ONE GPU without a communicator, sha256 of every .bwt/.sa slice"}
    return res


def single_job_leg(B, torch, dist, rank, local_rank, world, steps, warmup, haps, base_len, with_sa, do_e2e, do_parity):
    """One job over `world` GPUs: rank g holds haplotypes [g*haps, (g+1)*haps); pfbwt-f64 semantics
    (-s when with_sa).  Returns the dict reported as "single_job" (rank 0) or None."""
    n_rank = haps * base_len
    host = torch.empty(n_rank, dtype=torch.uint8).pin_memory()
    codes = job_codes(base_len)
    gen_job_text(host.numpy(), rank, haps, base_len, codes)
    rec_len = np.full(haps, base_len, dtype=np.uint64)
    ctx = B.Context(device=local_rank, bits=64, w=W, p=PMOD, get_sai=with_sa)
    if world > 1:
        ids = [B.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
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
    parity = None
    if do_parity and with_sa:
        bw, sa = ctx.bwt(), ctx.sa()
        text_at = job_text_evaluator(torch, codes, world * haps, base_len)
        parity = verify_slices(torch, dist, rank, world, int(sz.n), bw, sa, int(info.bwt_off), int(sz.r), text_at)
        del bw, sa, text_at
        torch.cuda.empty_cache()
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
        del out_bwt, out_sa
    del host
    sub = None
    if do_parity:
        sub = sub_job_equality(B, torch, dist, rank, local_rank, world, ctx, codes, base_len, min(base_len, 50_000_000))
    t = torch.tensor([dev_ms * 1e-3, t_wall, t_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t_dev, t_wall, t_e2e = float(t[0]), float(t[1]), float(t[2])
    ctx.close()
    if rank != 0:
        return None
    n_tot = world * n_rank
    if parity is not None or sub is not None:
        pr = {"ok": bool((parity is None or parity["ok"]) and (sub is None or sub["ok"]))}
        if parity is not None:
            pr["properties"] = parity
        if sub is not None:
            pr["sub_job_vs_one_gpu"] = sub
    else:
        pr = None
    return {
        "workload": f"config5 family: ONE job, {world * haps} haplotypes x {base_len / 1e6:g} Mbp (0.1% substitutions), {haps} per GPU, "
                    f"pfbwt-f64 -w {W} -p {PMOD}" + (" -s" if with_sa else ""),
        "scaling": "weak", "value": n_tot * steps / t_dev, "unit": "bases/s", "ms_per_step": 1e3 * t_dev / steps,
        "wall_ms_per_step": 1e3 * t_wall / steps,
        "e2e": {"value": n_tot * steps / t_e2e if do_e2e else None, "unit": "bases/s", "ms_per_step": 1e3 * t_e2e / steps if do_e2e else None,
                "h2d_bytes_per_step_rank0": int(h2d), "d2h_bytes_per_step_rank0": int(d2h)},
        "parity": pr,
        "n": int(sz.n), "phrases": int(sz.n_phrases), "dict_words": int(sz.n_words), "dict_bytes": int(sz.dict_bytes), "r": int(sz.r),
        "rank0_slice": {"bwt_off": int(info.bwt_off), "bwt_len": int(info.bwt_len)},
        "exchange": {"transport": "NCCL over NVLink" if world > 1 else "none",
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
    ap.add_argument("--bases", type=int, default=CONFIG2["bases"])
    ap.add_argument("--ref-sample", type=int, default=40_000_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling aid: skip the host-buffer leg")
    ap.add_argument("--no-single-job", action="store_true", help="skip the one-job-over-all-GPUs leg of the config-5 family")
    ap.add_argument("--no-file-leg", action="store_true", help="skip the file-to-file runs of bin/pfbwt-f (N=1 only)")
    ap.add_argument("--no-parity", action="store_true", help="profiling aid: skip the output checks")
    ap.add_argument("--job-haps", type=int, default=JOB_HAPS)
    ap.add_argument("--job-base", type=int, default=JOB_BASE)
    ap.add_argument("--job-no-sa", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n = args.bases
    workload = (f"config2: ONE synthetic {n / 1e9:g} Gbp uniform-random ACGT single sequence (default_rng(1)), -w {W} -p {PMOD}, 32-bit mode, "
                f".bwt only" + (f"; one job strong-scaled over {world} GPUs (rank g holds bases [g n/N, (g+1) n/N), mid-record cuts)" if world > 1 else ""))

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
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8/u32/u64 integer",
                "data": "synthetic",
                "config": {"workload": workload, "sample_bases": int(sample_n),
                           "sample_note": f"each step = the unmodified reference on the first {sample_n} bases of the {n}-base text (bounded sample; the "
                                          "reference's throughput falls with size, so this flatters the reference); FASTA read and file writes included; "
                                          "built with -O3 -march=x86-64-v3 (oracle/Makefile; BASELINE.md says -march=native)"},
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

    # this rank's share of THE text (seed 1) in pinned host memory
    lo, hi = n * rank // world, n * (rank + 1) // world
    host = torch.empty(hi - lo, dtype=torch.uint8).pin_memory()
    bases = host.numpy()
    gen_bases_range(1, lo, hi, bases)
    last = rank == world - 1
    out_host = torch.empty(int((n + W + 1) / world * 1.5) + 4096, dtype=torch.uint8).pin_memory()

    ctx = B.Context(device=local_rank, bits=32, w=W, p=PMOD)
    if world > 1:
        ids = [B.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.shard_init_nccl(rank, world, ids[0])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def add_text():
        if world == 1:
            ctx.add_records(host.data_ptr(), np.array([n], dtype=np.uint64))
        else:
            ctx.add_chunk(host.data_ptr(), hi - lo, ends_record=last)        # H2D from pinned memory

    # ---------------- value: text resident in HBM ---------------------------------------------
    add_text()
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
        add_text()
        ctx.finalize()
        ctx.bwt_of_parse()
        ctx.generate_bwt(False, False)
        m = ctx.shard_info().bwt_len
        ctx.bwt(out=out_host.numpy()[:m])                 # D2H of this rank's slice of the .bwt image
        return m

    t_e2e = float("nan")
    m_slice = 0
    if not args.no_e2e:
        for _ in range(max(1, min(args.warmup, 2))):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            m_slice = e2e_step()
        barrier()
        t_e2e = time.perf_counter() - t0
    sampler.stop_flag = True
    sampler.join()

    # ---------------- parity (outside the timed regions) -------------------------------------------
    parity = None
    if not args.no_parity:
        if args.no_e2e:
            m_slice = e2e_step()
        info = ctx.shard_info()
        sz = ctx.sizes()
        if n == CONFIG2["bases"]:
            # the slices of the last e2e step, brought together on rank 0 and hashed
            sl = torch.from_numpy(out_host.numpy()[:m_slice]).cuda()
            if world > 1:
                lens = [None] * world
                dist.all_gather_object(lens, int(m_slice))
                if rank == 0:
                    full = torch.empty(sum(lens), dtype=torch.uint8, device="cuda")
                    full[:lens[0]] = sl
                    o = lens[0]
                    for r in range(1, world):
                        if lens[r]:
                            dist.recv(full[o:o + lens[r]], src=r)
                        o += lens[r]
                elif m_slice:
                    dist.send(sl, dst=0)
            else:
                full = sl
            if rank == 0:
                got = sha16(full.cpu().numpy())
                parity = {"bwt_sha16": got, "expected_bwt_sha16": CONFIG2["bwt_sha16"], "expected_from": "unmodified reference on this text (SURVEY.md 8(d))",
                          "bwt_bytes": int(full.shape[0]), "n": int(sz.n), "r": int(sz.r), "phrases": int(sz.n_phrases), "dict_bytes": int(sz.dict_bytes),
                          "slices": world}
                parity["ok"] = bool(got == CONFIG2["bwt_sha16"] and sz.n == CONFIG2["n"] and sz.r == CONFIG2["r"] and
                                    sz.n_phrases == CONFIG2["phrases"] and sz.dict_bytes == CONFIG2["dict_bytes"] and full.shape[0] == CONFIG2["n"] + 1)
                del full
            del sl
        else:
            # another size: no reference hash on record - property check of a BWT+SA build of the same job
            whole = np.empty(n, dtype=np.uint8)
            gen_bases(whole, 1)
            tt = torch.from_numpy(whole).cuda()
            ctx.clear()
            add_text()
            ctx.finalize()
            ctx.bwt_of_parse()
            ctx.generate_bwt(True, False)
            bw, sa = ctx.bwt(), ctx.sa()
            info = ctx.shard_info()
            fill = torch.full((), 65, dtype=torch.uint8, device="cuda")        # the A^w padding behind the record

            def text_at(pos):
                return torch.where(pos >= n, fill, tt[pos.clamp(max=n - 1)])
            res = verify_slices(torch, dist, rank, world, int(sz.n), bw, sa.astype(np.uint64), int(info.bwt_off), int(ctx.sizes().r), text_at)
            parity = dict(res, kind="properties (no reference hash recorded for this --bases)")
            del tt, whole, bw, sa
        torch.cuda.empty_cache()

    # max over ranks
    t_dev = dev_ms * 1e-3
    if world > 1:
        t = torch.tensor([t_res, t_e2e, t_dev], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_res, t_e2e, t_dev = float(t[0]), float(t[1]), float(t[2])

    moved0 = ctx.stats()
    ctx.close()
    bases0 = bases.copy() if (rank == 0 and world == 1) else None
    del host, out_host, bases
    job = None
    if not args.no_single_job:
        job = single_job_leg(B, torch, dist, rank, local_rank, world, max(1, min(args.steps, 3)), max(1, min(args.warmup, 2)),
                             args.job_haps, args.job_base, not args.job_no_sa, not args.no_e2e, not args.no_parity)

    if rank == 0:
        peak, peak_kind = hbm_peak()
        stage_ms = {k: v / args.steps for k, v in stage_acc.items()}
        # dominant kernel: k_radix_pass (5 launches of the dictionary suffix sort per step).  Algorithmic
        # bytes per launch = SURVEY 8(d) per-unit figure of the dictionary suffix sort (1+U bytes per
        # suffix) x the suffixes one launch processes; duration = CUDA events around every launch.
        U = 4
        npass = st1["radix_passes"] - launches0["radix_passes"]
        pass_ms = (st1["radix_pass_ms"] - launches0["radix_pass_ms"]) / max(1, npass)
        pairs = (st1["radix_pairs"] - launches0["radix_pairs"]) / max(1, npass)
        alg_bytes = pairs * (1 + U)
        achieved = alg_bytes / (pass_ms * 1e-3) / 1e9 if pass_ms else 0.0
        traffic = None
        tp = os.path.join(ROOT, "profiles", "radix_pass_traffic.json")   # dram bytes per launch from the committed ncu --set full capture
        if os.path.exists(tp) and world == 1:
            try:
                traffic = json.load(open(tp))["dram_bytes_per_launch"]
            except Exception:
                traffic = None
        cpu = None
        if not args.no_cpu_baseline and world == 1:        # reported at N=1 only (rank 0's host cores)
            sample_n = min(args.ref_sample, n)
            v, dt, kind = run_reference_sample(bases0, sample_n)
            cpu = {"value": v, "unit": "bases/s", "cores": 1, "kind": kind,
                   "sample": f"first {sample_n} bases of the text, whole pfbwt-f -w {W} -p {PMOD} run incl. file I/O ({dt:.1f} s)"}
        f2f = None
        if not args.no_file_leg and world == 1:
            try:
                f2f = file_to_file_leg(bases0)
                f2f["with_sa"] = file_to_file_leg(bases0, runs=1, with_sa=True)
            except Exception as e:      # the leg is an extra: never lose the bench line over a full disk
                f2f = f2f or {"unavailable": repr(e)[:200]}
        sort_ms = stage_ms.get("dict_sort_ms", 0.0)
        step_ms = 1e3 * t_dev / args.steps
        line = {
            "metric": "input bases/sec end-to-end fasta->.bwt",
            "value": n * args.steps / t_dev, "unit": "bases/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": step_ms, "wall_ms_per_step": 1e3 * t_res / args.steps,
            "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8/u32/u64 integer", "data": "synthetic",
            "config": {"workload": workload, "n": int(sz.n), "phrases": int(sz.n_phrases), "dict_words": int(sz.n_words),
                       "dict_bytes": int(D), "r": int(sz.r), "l2": "inputs larger than L2 (text 1 GB, dictionary 1.1 GB per step)",
                       "multi_gpu": ("ONE job over all ranks: chunk-wise parse with w-byte halos, NCCL merge of the parses, key-range partitioned "
                                     "dictionary suffix sort, one .bwt slice per rank; times are max over ranks") if world > 1 else "single GPU"},
            "e2e": {"value": n * args.steps / t_e2e, "unit": "bases/s", "h2d_bytes_per_step": int(n),
                    "d2h_bytes_per_step": int(n + W + 1), "ms_per_step": 1e3 * t_e2e / args.steps},
            "parity": parity,
            "gpu_launches": int(launches), "gpu_launches_library_cub": int(launches_lib),
            "roofline": {"bound": "hbm", "kernel": "k_radix_pass (dictionary suffix sort, one 8-bit digit per launch" + (", rank 0" if world > 1 else "") + ")",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs, burst copy)", "algorithmic_bytes": int(alg_bytes),
                         "ms": pass_ms, "launches_timed": int(npass),
                         "moved_bytes": int(pairs * 24), "moved_gbs": pairs * 24 / (pass_ms * 1e-3) / 1e9 if pass_ms else 0.0,
                         "share_of_step": pass_ms * npass / args.steps / step_ms if t_dev else None,
                         "stage": {"what": "the same floor D*(1+U) over the whole dictionary-suffix-sort stage (keys + all passes + refinement)",
                                   "ms": sort_ms, "achieved": D * (1 + U) / world / (sort_ms * 1e-3) / 1e9 if sort_ms else None,
                                   "frac": D * (1 + U) / world / (sort_ms * 1e-3) / 1e9 / peak if sort_ms else None}},
            "stages_ms": stage_ms,
            "clocks": sampler.summary(),
            "dev_bytes_peak": int(st1["dev_bytes_peak"]),
        }
        if world > 1:
            line["exchange"] = {"transport": "NCCL over NVLink", "bytes_received_rank0": int(moved0["shard_bytes_moved"]),
                                "exchange_ms_rank0": stage_ms.get("shard_exchange_ms"), "merge_ms_rank0": stage_ms.get("shard_merge_ms")}
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
