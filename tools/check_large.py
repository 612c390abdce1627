"""Full-size property check past 2^32 text positions (configs 3-5 are all larger than 4 Gbp; the reference
cannot be run at that size here).  Builds BWT + SA of `haps` haplotypes x `base` bases on one GPU (pfbwt-f64 -s)
and verifies, size-independently:
  (a) inversion: T[SA[i]-1] = BWT[i] for every i reproduces S exactly, the 0x00 terminator sits at SA[i] = 0,
      SA[0] = n (so SA is a permutation and BWT is the BWT of S under that SA);
  (b) sortedness on sampled adjacent pairs: S[SA[i]:] < S[SA[i+1]:] (first 64 KiB compared, ties skipped);
  (c) r = number of runs of the returned BWT.
  (d) with a third argument `r`: .ssa / .esa equal the run starts / ends of the returned BWT with their SA values.
  python tools/check_large.py [haps] [base] [r]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from bench import gen_job_text
from pfbwt_b200_loader import binding

haps = int(sys.argv[1]) if len(sys.argv) > 1 else 45
base = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000
W = 10
B = binding()
t0 = time.time()
host = torch.empty(haps * base, dtype=torch.uint8).pin_memory()
gen_job_text(host.numpy(), 0, haps, base)
print(f"generated {haps} x {base} in {time.time() - t0:.0f}s", flush=True)
ctx = B.Context(device=0, bits=64, w=W, p=100, get_sai=True)
ctx.add_records(host.data_ptr(), np.full(haps, base, dtype=np.uint64))
t0 = time.time()
RSSA = len(sys.argv) > 3 and sys.argv[3] == "r"          # also build and check the run-length SA samples (-r)
ctx.finalize(); ctx.bwt_of_parse(); ctx.release_text(); ctx.generate_bwt(sa=True, rssa=RSSA)
sz = ctx.sizes()
n = int(sz.n)
print(f"built in {time.time() - t0:.2f}s wall: n={n} (> 2^32: {n > 2**32}) P={sz.n_phrases} Dw={sz.n_words} D={sz.dict_bytes} r={sz.r} "
      f"stages={ {k: round(v, 1) for k, v in ctx.stats().items() if k.endswith('_ms')} }", flush=True)
bwt = ctx.bwt(); sa = ctx.sa()
if RSSA:
    ssa = ctx.ssa().reshape(-1, 2); esa = ctx.esa().reshape(-1, 2)
ctx.close()
assert len(bwt) == n + 1 and len(sa) == n + 1
assert int(sa[0]) == n, "SA[0] != n"
# S = records each followed by A^w
S = np.empty(n, dtype=np.uint8)
src = host.numpy()
for h in range(haps):
    S[h * (base + W):h * (base + W) + base] = src[h * base:(h + 1) * base]
    S[h * (base + W) + base:(h + 1) * (base + W)] = ord("A")
del host, src
T = np.full(n + 1, 255, dtype=np.uint8)
step = 1 << 28
for s in range(0, n + 1, step):
    e = min(n + 1, s + step)
    idx = sa[s:e].astype(np.int64) - 1
    idx[idx < 0] = n
    T[idx] = bwt[s:e]
assert T[n] == 0, "terminator misplaced"
bad = np.flatnonzero(T[:n] != S)
assert bad.size == 0, f"inversion differs at {bad[:5]} ({bad.size} positions)"
print("inversion ok", flush=True)
rng = np.random.default_rng(7)
K = 1 << 16
ties = 0
for i in rng.integers(1, n, 20000):
    a, b = int(sa[i]), int(sa[i + 1])
    x, y = S[a:a + K], S[b:b + K]
    m = min(len(x), len(y))
    d = np.flatnonzero(x[:m] != y[:m])
    if d.size:
        assert x[d[0]] < y[d[0]], f"order violated at i={i}"
    elif len(x) < K or len(y) < K:
        assert len(x) < len(y), f"order violated (prefix) at i={i}"
    else:
        ties += 1
print(f"sampled order ok ({ties} pairs tied beyond {K} bytes)", flush=True)
r = int(np.count_nonzero(bwt[1:] != bwt[:-1])) + 1
assert r == int(sz.r), (r, int(sz.r))
if RSSA:
    # src/pfbwt-f.cpp:304-328: .ssa = (i, SA[i]) at every run start, .esa = (i, SA[i]) at every run end
    starts = np.flatnonzero(np.concatenate([[True], bwt[1:] != bwt[:-1]]))
    ends = np.concatenate([starts[1:] - 1, [n]])
    assert len(ssa) == r and len(esa) == r
    assert np.array_equal(ssa[:, 0], starts.astype(np.uint64)) and np.array_equal(ssa[:, 1], sa[starts])
    assert np.array_equal(esa[:, 0], ends.astype(np.uint64)) and np.array_equal(esa[:, 1], sa[ends])
    print("ssa/esa ok", flush=True)
print("runs ok; ALL OK", flush=True)
