#!/bin/bash
# profiles/ncu_top.sh TAG KERNEL_REGEX [KERNEL_REGEX...]   (run under gpurun, one GPU)
# 1. plain bench run (must exit 0), 2. ncu launch list of the same command, 3. one `ncu --set full`
# capture of the longest launch of every kernel regex given (regexes see DEMANGLED names, as the launch list prints them:
# `for_kernel<...k_fingerprint...>` lambdas can be picked).  Outputs under gpurun_out/.
set -u
TAG=$1; shift
CMD=${NCU_CMD:-"python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-single-job --no-file-leg --no-parity"}
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1
for RX in "$@"; do
  IDX=$(python - "$RX" gpurun_out/${TAG}_launches.csv <<'PY'
import csv, re, sys
rx = re.compile(sys.argv[1]); best = (-1.0, 1); count = 0
rows = list(csv.reader(l for l in open(sys.argv[2]) if not l.startswith("==")))
hdr = rows[0]; kn = hdr.index("Kernel Name"); mv = hdr.index("Metric Value"); mu = hdr.index("Metric Unit")
for r in rows[1:]:
    if len(r) <= mv or not rx.search(r[kn]): continue
    count += 1
    v = float(r[mv].replace(",", ""))
    if v > best[0]: best = (v, count)
print(best[1])
PY
)
  NAME=$(echo "$RX" | tr -c 'A-Za-z0-9_' '_')
  echo "profiling $RX invocation $IDX"
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled --kernel-id ::regex:"$RX":$IDX -o gpurun_out/${TAG}_${NAME} -f $CMD > gpurun_out/${TAG}_${NAME}.log 2>&1
done
ls -la gpurun_out/ | grep ${TAG}
