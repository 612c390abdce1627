"""BASELINE.json configs[3] at (near) full size on one B200: a human-scale assembly with N runs,
`pfbwt-f64 --non-acgt-to-a -w 10 -p 100` - 24 records, ~3.1 Gbp in total, 60-column lines are irrelevant here
(records are handed over as buffers), 10 % lower-case soft-masked stretches, per record one centromere-like N run
of 1-30 Mbp, 50 runs of 1-50 kbp and 10 kbp telomeric Ns at both ends.  Exercises toupper, N->A,
multi-Mbp phrases (poly-A after the substitution: prefix doubling), a dictionary of ~3.4e9 bytes (32-bit offsets near
their limit) and 64-bit text positions.  Verified by BWT/SA inversion when the SA fits, else by run count + symbol counts.
  python tools/check_config4.py [total_bases] [sa 0|1]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from pfbwt_b200_loader import binding

total = int(sys.argv[1]) if len(sys.argv) > 1 else 3_100_000_000
want_sa = bool(int(sys.argv[2])) if len(sys.argv) > 2 else True
NREC, W = 24, 10
B = binding()
rng = np.random.default_rng(38)
rec_len = total // NREC
host = torch.empty(NREC * rec_len, dtype=torch.uint8).pin_memory()
buf = host.numpy()
LUT = np.frombuffer(b"ACGT", dtype=np.uint8)
t0 = time.time()
for r in range(NREC):
    s = buf[r * rec_len:(r + 1) * rec_len]
    for o in range(0, rec_len, 50_000_000):
        m = min(50_000_000, rec_len - o)
        s[o:o + m] = LUT[rng.integers(0, 4, m, dtype=np.uint8)]
    a = int(rng.integers(0, rec_len - rec_len // 10))
    s[a:a + rec_len // 10] |= 0x20                                   # soft-masked
    tel = min(10_000, rec_len // 20)
    s[:tel] = ord("N"); s[rec_len - tel:] = ord("N")
    big = int(rng.integers(1_000_000, 30_000_000)) if rec_len > 80_000_000 else rec_len // 8
    b = int(rng.integers(tel, rec_len - big - tel))
    s[b:b + big] = ord("N")
    for _ in range(50):
        L = int(rng.integers(1_000, 50_000))
        c = int(rng.integers(tel, rec_len - L - tel))
        s[c:c + L] = ord("n") if rng.integers(0, 2) else ord("N")
print(f"generated {NREC} x {rec_len} in {time.time() - t0:.0f}s", flush=True)
ctx = B.Context(device=0, bits=64, w=W, p=100, get_sai=want_sa, non_acgt_to_a=True)
ctx.add_records(host.data_ptr(), np.full(NREC, rec_len, dtype=np.uint64))
t0 = time.time()
ctx.finalize(); ctx.bwt_of_parse(); ctx.release_text()
sz = ctx.sizes()
print(f"parsed in {time.time() - t0:.2f}s: n={sz.n} P={sz.n_phrases} Dw={sz.n_words} D={sz.dict_bytes}", flush=True)
t1 = time.time()
ctx.generate_bwt(sa=want_sa, rssa=False)
sz = ctx.sizes()
st = ctx.stats()
print(f"built in {time.time() - t1:.2f}s (total {time.time() - t0:.2f}s wall): r={sz.r} rounds={st['dict_sort_rounds']} peak={st['dev_bytes_peak'] / 1e9:.1f} GB "
      f"stages={ {k: round(v, 1) for k, v in st.items() if k.endswith('_ms')} }", flush=True)
n = int(sz.n)
bwt = ctx.bwt()
sa = ctx.sa() if want_sa else None
ctx.close()
# expected S: upper-cased, non-ACGT -> A, A^w after every record
S = np.empty(n, dtype=np.uint8)
for r in range(NREC):
    x = buf[r * rec_len:(r + 1) * rec_len] & 0xDF
    x[(x != ord("A")) & (x != ord("C")) & (x != ord("G")) & (x != ord("T"))] = ord("A")
    S[r * (rec_len + W):r * (rec_len + W) + rec_len] = x
    S[r * (rec_len + W) + rec_len:(r + 1) * (rec_len + W)] = ord("A")
assert len(bwt) == n + 1
cb = np.bincount(bwt, minlength=256); cs = np.bincount(S, minlength=256)
assert cb[0] == 1 and all(cb[c] == cs[c] for c in b"ACGT"), "symbol counts differ"
r = int(np.count_nonzero(bwt[1:] != bwt[:-1])) + 1
assert r == int(sz.r), (r, int(sz.r))
print("symbol counts + runs ok", flush=True)
if want_sa:
    assert int(sa[0]) == n
    T = np.full(n + 1, 255, dtype=np.uint8)
    step = 1 << 28
    for s0 in range(0, n + 1, step):
        e = min(n + 1, s0 + step)
        idx = sa[s0:e].astype(np.int64) - 1
        idx[idx < 0] = n
        T[idx] = bwt[s0:e]
    assert T[n] == 0
    bad = np.flatnonzero(T[:n] != S)
    assert bad.size == 0, f"inversion differs at {bad[:5]} ({bad.size})"
    print("inversion ok", flush=True)
    rng2 = np.random.default_rng(9)
    K = 1 << 16
    ties = 0
    for i in rng2.integers(1, n, 20000):
        a, b = int(sa[i]), int(sa[i + 1])
        x, y = S[a:a + K], S[b:b + K]
        m = min(len(x), len(y))
        d = np.flatnonzero(x[:m] != y[:m])
        if d.size: assert x[d[0]] < y[d[0]], f"order violated at i={i}"
        elif len(x) < K or len(y) < K: assert len(x) < len(y), f"order violated (prefix) at i={i}"
        else: ties += 1
    print(f"sampled order ok ({ties} pairs tied beyond {K} bytes: poly-A runs)", flush=True)
print("ALL OK", flush=True)
