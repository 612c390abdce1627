"""File-to-file timing of the drop-in binary (FASTA on tmpfs -> all output files closed), with the
FASTA records split on the device (default) and by the host reader (PFB_INGEST=host).
usage: [BENCH_GZ=1] [BENCH_HAPS=64] python tools/bench_cli.py [bases=1000000000] [line_width=60] [extra flags ...]"""
import os, subprocess, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000_000
lw = int(sys.argv[2]) if len(sys.argv) > 2 else 60
flags = sys.argv[3:]
d = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
fa = os.path.join(d, "pfb_cli_bench.fa")
rng = np.random.default_rng(1)
lut = np.frombuffer(b"ACGT", dtype=np.uint8)
haps = int(os.environ.get("BENCH_HAPS", "0"))     # >0: that many haplotypes of ONE base sequence of `bases` bases, 0.1 % substitutions
t0 = time.time()


def write_record(f, name, x):
    f.write(b">" + name + b"\n")
    m = x.shape[0]
    full = m // lw * lw
    if full:
        a = np.empty((full // lw, lw + 1), dtype=np.uint8)
        a[:, :lw] = x[:full].reshape(-1, lw)
        a[:, lw] = 10
        f.write(a.tobytes())
    if m > full:
        f.write(x[full:].tobytes() + b"\n")


with open(fa, "wb") as f:
    if haps:
        codes = np.random.default_rng(2024).integers(0, 4, n, dtype=np.uint8)
        for h in range(haps):
            c = codes
            if h:
                r2 = np.random.default_rng(5000 + h)
                pos = r2.integers(0, n, n // 1000)
                c = codes.copy()
                c[pos] = (c[pos] + r2.integers(1, 4, n // 1000, dtype=np.uint8)) % 4
            write_record(f, b"hap%d some comment" % h, lut[c])
        n = n * haps
    else:
        f.write(b">rand\n")
        step = 60_000_000 // lw * lw
        for s in range(0, n, step):
            m = min(step, n - s)
            x = lut[rng.integers(0, 4, m, dtype=np.uint8)]
            full = m // lw * lw
            if full:
                a = np.empty((full // lw, lw + 1), dtype=np.uint8)
                a[:, :lw] = x[:full].reshape(-1, lw)
                a[:, lw] = 10
                f.write(a.tobytes())
            if m > full:
                f.write(x[full:].tobytes() + b"\n")
if os.environ.get("BENCH_GZ"):           # gzip input: zlib inflate on the host, records split on the device
    import gzip, shutil
    t1 = time.time()
    with open(fa, "rb") as fi, gzip.open(fa + ".gz", "wb", compresslevel=1) as fo:
        shutil.copyfileobj(fi, fo, 1 << 24)
    os.remove(fa)
    fa = fa + ".gz"
    print(f"# gzip -1 in {time.time() - t1:.1f} s", flush=True)
print(f"# {fa}: {os.path.getsize(fa)} bytes, {n} bases, {lw}-column lines, written in {time.time() - t0:.1f} s", flush=True)
exe = os.path.join(ROOT, "bin", "pfbwt-f64" if n > 4_000_000_000 else "pfbwt-f")
for mode in (("device", "host") if haps else ("device", "host", "device")):
    env = dict(os.environ, PFB_CLI_TIMES="1")
    if mode == "host":
        env["PFB_INGEST"] = "host"
    out = os.path.join(d, "pfb_cli_bench_out")
    t0 = time.time()
    r = subprocess.run([exe] + flags + ["-o", out, fa], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    wall = time.time() - t0
    tasks = [l.split("\t") for l in r.stderr.decode().splitlines() if l.startswith("TASK") or l.startswith("TIME")]
    print(f"ingest={mode:6s} rc={r.returncode} wall {wall:.3f} s ({n / wall / 1e9:.2f} Gbases/s file to file)  " +
          "  ".join(f"[{t[1]}: {t[2]}]" for t in tasks if len(t) > 2), flush=True)
    import hashlib
    def sha16(path):
        h = hashlib.sha256()
        with open(path, "rb") as f:
            for blk in iter(lambda: f.read(1 << 24), b""):
                h.update(blk)
        return h.hexdigest()[:16]
    if r.returncode == 0:
        print("    sha256[:16]  " + "  ".join(f".{e} {sha16(out + '.' + e)}" for e in ("bwt", "dict", "parse", "docs") if os.path.exists(out + "." + e)), flush=True)
    for ext in ("bwt", "sa", "dict", "occ", "parse", "n", "bwlast", "ilist", "bwsai", "ssa", "esa", "docs"):
        try:
            os.remove(out + "." + ext)
        except OSError:
            pass
os.remove(fa)
