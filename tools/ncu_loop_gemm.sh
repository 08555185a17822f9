#!/bin/bash
# Durations and DRAM traffic of the in-loop kernels in their REAL cache state (no cache flush, one pass per kernel).
# usage: tools/ncu_loop_gemm.sh <tag>   (env SAT_STATIC_W etc. are passed through)
mkdir -p gpurun_out
timeout 600 ncu --cache-control none --clock-control none --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
  --launch-skip 600 -c 260 --csv --log-file gpurun_out/ncu_loop_$1.csv python bench.py --steps 1 --warmup 3 > gpurun_out/ncu_loop_$1.log 2>&1
python - "$1" <<'PY'
import csv, sys, collections
tag = sys.argv[1]
lines = [l for l in open(f"gpurun_out/ncu_loop_{tag}.csv") if l.startswith('"')]
agg = collections.OrderedDict()
for r in csv.DictReader(lines):
    name = r["Kernel Name"].split("(")[0][-60:] + " " + r["Grid Size"]
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    if u == "ns": v /= 1e3
    if u == "ms": v *= 1e3
    if u == "byte": v /= 1e6
    if u == "Kbyte": v /= 1e3
    if u == "Gbyte": v *= 1e3
    a = agg.setdefault(name, collections.defaultdict(list))
    a[r["Metric Name"]].append(v)
print(f"# {tag}: kernel | n | avg us | avg dram read MB | avg dram write MB")
for k, a in agg.items():
    n = len(a["gpu__time_duration.sum"])
    f = lambda m: sum(a[m]) / max(1, len(a[m]))
    print(f"{k:90s} {n:4d} {f('gpu__time_duration.sum'):8.1f} {f('dram__bytes_read.sum'):9.2f} {f('dram__bytes_write.sum'):9.2f}")
PY
