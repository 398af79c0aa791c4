# Per-kernel device times of ONE per-image load (config C1): latency probe, then an ncu launch list of the same command.
python tools/latency_probe.py 2>&1 | cut -c1-330
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 14 --csv --log-file gpurun_out/c1_launches.csv python tools/latency_probe.py > /dev/null 2>&1
python - <<'PY'
import csv
rows = list(csv.reader(open("gpurun_out/c1_launches.csv")))
h = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[h]
ik, iv, ig, ib = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
for r in rows[h+1:]:
    if len(r) > iv: print("%-50s %10s ns grid %s block %s" % (r[ik][:50], r[iv], r[ig], r[ib]))
PY
