"""Shared helpers for the test-suite: seeded synthetic FASTA generators (SURVEY.md
§8(d) recipes, scaled), runners for the oracle binaries, file comparison.

Nothing here reads /root/reference at run time (it does not exist on the GPU box);
the reference binaries are the prebuilt oracle/_ref/{pfbwt-f,pfbwt-f64,merge_pfp}.
"""
import os
import subprocess
import hashlib
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")
PARSE_EXTS = ["dict", "occ", "parse", "n", "bwlast", "ilist"]
LUT = np.frombuffer(b"ACGT", dtype=np.uint8)


def ref_bin(name):
    p = os.path.join(REF_DIR, name)
    return p if os.path.exists(p) else None


def random_bases(n, seed):
    """i.i.d. uniform ACGT, NumPy default_rng(seed) (config 2 recipe)."""
    rng = np.random.default_rng(seed)
    out = np.empty(n, dtype=np.uint8)
    step = 100_000_000
    for s in range(0, n, step):
        m = min(step, n - s)
        out[s:s + m] = LUT[rng.integers(0, 4, m, dtype=np.uint8)]
    return out


def write_fasta(path, records, line_width=None):
    """records: list of (name, uint8 array)."""
    with open(path, "wb") as f:
        for name, seq in records:
            f.write(b">" + name.encode() + b"\n")
            if line_width:
                b = seq.tobytes()
                for i in range(0, len(b), line_width):
                    f.write(b[i:i + line_width] + b"\n")
            else:
                f.write(seq.tobytes() + b"\n")


def pangenome_records(base_len, nhap, nsnp, seed):
    """config 3 recipe: hap0 = base, hap h>0 = base with nsnp SNPs."""
    rng = np.random.default_rng(seed)
    codes = rng.integers(0, 4, base_len, dtype=np.uint8)
    recs = [("hap0", LUT[codes])]
    for h in range(1, nhap):
        c = codes.copy()
        pos = rng.choice(base_len, nsnp, replace=False)
        c[pos] = (c[pos] + rng.integers(1, 4, nsnp, dtype=np.uint8)) % 4
        recs.append((f"hap{h}", LUT[c]))
    return recs


def nrun_records(nrec, rec_len, seed, big_run=20000, small_runs=5):
    """config 4 recipe (scaled): N runs, soft-masked lower case, telomeric Ns."""
    rng = np.random.default_rng(seed)
    recs = []
    for r in range(nrec):
        s = LUT[rng.integers(0, 4, rec_len, dtype=np.uint8)].copy()
        # lower-case 10 %
        a = int(rng.integers(0, rec_len - rec_len // 10))
        s[a:a + rec_len // 10] |= 0x20
        # N runs
        tel = min(100, rec_len // 20)
        s[:tel] = ord("N")
        s[rec_len - tel:] = ord("N")
        b = int(rng.integers(tel, max(tel + 1, rec_len - big_run - tel)))
        s[b:b + min(big_run, rec_len // 3)] = ord("N")
        for _ in range(small_runs):
            L = int(rng.integers(1, 500))
            c = int(rng.integers(tel, max(tel + 1, rec_len - L - tel)))
            s[c:c + L] = ord("n") if rng.integers(0, 2) else ord("N")
        recs.append((f"chr{r + 1}", s))
    return recs


def sha16(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        while True:
            b = f.read(1 << 24)
            if not b:
                break
            h.update(b)
    return h.hexdigest()[:16]


def run(cmd, cwd=None, check=True, stdin=None):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, input=stdin)
    if check and r.returncode != 0:
        raise RuntimeError(f"{cmd} failed ({r.returncode}):\n{r.stderr.decode(errors='replace')[-2000:]}")
    return r


def exts_for(sa, rssa, parse_only=False):
    e = list(PARSE_EXTS)
    if sa or rssa:
        e.append("bwsai")
    if not parse_only:
        e.append("bwt")
        if sa:
            e.append("sa")
        if rssa:
            e += ["ssa", "esa"]
    return e


def assert_same_files(pfx_a, pfx_b, exts):
    for e in exts:
        a, b = f"{pfx_a}.{e}", f"{pfx_b}.{e}"
        assert os.path.exists(a), a
        assert os.path.exists(b), b
        da, db = open(a, "rb").read(), open(b, "rb").read()
        if da != db:
            n = min(len(da), len(db))
            first = next((i for i in range(n) if da[i] != db[i]), n)
            raise AssertionError(f".{e} differs: sizes {len(da)} vs {len(db)}, first diff at byte {first}")


def invert_bwt_sa(bwt, sa):
    """Recover S from BWT/SA of S.$ : T[SA[i]-1] = BWT[i]."""
    n = len(bwt) - 1
    T = np.zeros(n + 1, dtype=np.uint8)
    idx = (sa.astype(np.int64) - 1) % (n + 1)
    T[idx] = bwt
    assert T[n] == 0
    return T[:n]


def naive_sa_bwt(S):
    """Independent SA/BWT of S.$ by numpy prefix doubling ($ = smallest, written 0x00)."""
    n = len(S)
    T = np.concatenate([S.astype(np.int64) + 1, np.zeros(1, dtype=np.int64)])
    m = n + 1
    rank = T.copy()
    h = 1
    while True:
        r1 = np.zeros(m, dtype=np.int64)
        r1[:m - h] = rank[h:] + 1 if h < m else 0
        order = np.lexsort((r1, rank))
        k0, k1 = rank[order], r1[order]
        new = np.empty(m, dtype=np.int64)
        diff = np.ones(m, dtype=bool)
        diff[1:] = (k0[1:] != k0[:-1]) | (k1[1:] != k1[:-1])
        new[order] = np.cumsum(diff) - 1
        rank = new
        if diff.all() or h >= m:
            break
        h *= 2
    sa = order.astype(np.uint64)
    full = np.concatenate([S, np.zeros(1, dtype=np.uint8)])
    bwt = full[(order - 1) % m]
    return sa, bwt
