"""Stage trace (PFB_TRACE=1) of the single-GPU path on the config-3/5 family: `haps` haplotypes of one base sequence.
  PFB_TRACE=1 python tools/trace_pangenome.py [haps] [base_len] [sa]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from bench import gen_job_text
from pfbwt_b200_loader import binding
haps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
base = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000
sa = bool(int(sys.argv[3])) if len(sys.argv) > 3 else True
B = binding()
host = torch.empty(haps * base, dtype=torch.uint8).pin_memory()
gen_job_text(host.numpy(), 0, haps, base)
ctx = B.Context(device=0, bits=64, w=10, p=100, get_sai=sa)
ctx.add_records(host.data_ptr(), np.full(haps, base, dtype=np.uint64))
ctx.upload_text()
for i in range(3):
    t0 = time.perf_counter()
    ctx.run_resident(sa, False)
    st = ctx.stats()
    print(f"run {i}: {st['run_ms']:.1f} ms  " + " ".join(f"{k[:-3]}={v:.1f}" for k, v in st.items() if k.endswith('_ms') and k not in ('run_ms', 'radix_pass_ms')), file=sys.stderr)
sz = ctx.sizes()
print(f"n={sz.n} P={sz.n_phrases} Dw={sz.n_words} D={sz.dict_bytes} r={sz.r} rounds={st['dict_sort_rounds']}", file=sys.stderr)
