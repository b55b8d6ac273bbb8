#!/bin/bash
# usage (under gpurun): tools/side_launches.sh <cfg key> -- ncu launch list (kernel names + durations) of one side config of bench.py
key=${1:-cfg5_dextreme_adr}
ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 160 --csv --log-file gpurun_out/side_$key.csv \
  python -c "import bench, json; print(json.dumps(bench.measure_other('$key', 'cuda:0', 20)))" > gpurun_out/side_$key.log 2>&1
python - "$key" <<'PY'
import csv, sys, collections, statistics
rows = list(csv.reader(l for l in open(f"gpurun_out/side_{sys.argv[1]}.csv") if l.startswith('"')))
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value")
seq = [(r[ki][:60], float(r[vi])) for r in rows[1:]]
d = collections.OrderedDict()
for k, v in seq: d.setdefault(k, []).append(v)
for k, v in d.items(): print(f"{len(v):4d} x {statistics.median(v)/1e3:7.2f} us  {k}")
print("order:", [k[:24] for k, _ in seq[:14]])
PY
