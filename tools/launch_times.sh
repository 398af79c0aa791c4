#!/bin/sh
# Per-launch device times of one bench step (cold-cache, serialised: compare SHARES):  tools/launch_times.sh SKIP COUNT bench.py args...
SKIP=$1; COUNT=$2; shift 2
CMD="python bench.py --no-cpu --no-e2e --steps 1 --warmup 3 $*"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -s "$SKIP" -c "$COUNT" --csv --log-file gpurun_out/launches_tmp.csv $CMD > gpurun_out/ncu_tmp.log 2>&1
python - <<PY
import csv, collections
rows = list(csv.reader(open("gpurun_out/launches_tmp.csv")))
h = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[h]
ik, im, iv, iid = (hdr.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "ID"))
d = collections.OrderedDict()
for r in rows[h + 1:]:
    if len(r) > iv:
        d.setdefault((r[iid], r[ik][:44]), {})[r[im]] = r[iv]
for (i, k), v in d.items():
    print("%3s %-44s %8.3f ms  %6.0f Minstr  thr/instr %5.1f  issue %5.1f%%" % (i, k, float(v["gpu__time_duration.sum"]) / 1e6,
          float(v["smsp__inst_executed.sum"]) / 1e6, float(v["smsp__thread_inst_executed_per_inst_executed.ratio"]),
          float(v["smsp__issue_active.avg.pct_of_peak_sustained_active"])))
PY
