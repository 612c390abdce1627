#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and shares.
usage: summarize_launches.py launches.csv [steps_in_capture]"""
import csv, re, sys, collections
path = sys.argv[1]
steps = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rows = list(csv.DictReader(lines))
tot = collections.defaultdict(float); cnt = collections.Counter()
scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
for x in rows:
    v = float(x["Metric Value"].replace(",", "")) * scale[x["Metric Unit"]]
    nm = x["Kernel Name"]
    m = re.search(r"pfb::(\w+)(?:.*?pfb::(\w+))?", nm)
    if "cub::" in nm and not nm.startswith("void pfb"):
        short = "cub::" + re.search(r"cub::(?:CUB_\w+::)?(\w+)", nm).group(1) + "<" + re.sub(r".*policy_hub<([^>]*)>.*", r"\1", nm)[:48] + ">"
    elif "for_kernel" in nm:
        inner = re.search(r"for_kernel<(.*)>\(unsigned long", nm)
        body = inner.group(1) if inner else nm
        fn = re.findall(r"pfb::(\w+)", body)
        inst = re.search(r"instance (\d+)", body)
        short = "for_kernel[" + ("/".join(dict.fromkeys(fn)) if fn else body[:40]) + (" #" + inst.group(1) if inst else "") + "]"
    else:
        short = re.sub(r"\(.*", "", nm)[:100]
    tot[short] += v; cnt[short] += 1
T = sum(tot.values())
print(f"# {path}: {len(rows)} launches, {T:.2f} ms total device time ({T/steps:.2f} ms per step over {steps:g} steps); cold-cache serialised ncu times: compare shares")
print(f"{'ms/step':>10} {'share':>7} {'launches/step':>14}  kernel")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    if v / T < 0.002: continue
    print(f"{v/steps:10.3f} {100*v/T:6.1f}% {cnt[k]/steps:14.1f}  {k}")
