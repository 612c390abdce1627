"""config 2 text (10^9-base recipe of SURVEY.md 8(d): NumPy default_rng(1), uniform ACGT) as a FASTA file with 60-column lines.
usage: python tools/make_fasta.py <bases> <out.fa>"""
import numpy as np, sys
n=int(sys.argv[1]); lw=60
rng=np.random.default_rng(1); lut=np.frombuffer(b"ACGT",dtype=np.uint8)
with open(sys.argv[2],"wb") as f:
    f.write(b">rand\n")
    step=60_000_000
    for s in range(0,n,step):
        m=min(step,n-s); x=lut[rng.integers(0,4,m,dtype=np.uint8)]
        full=m//lw*lw
        a=np.empty((full//lw,lw+1),dtype=np.uint8); a[:,:lw]=x[:full].reshape(-1,lw); a[:,lw]=10
        f.write(a.tobytes())
        if m>full: f.write(x[full:].tobytes()+b"\n")
