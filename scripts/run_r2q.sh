#!/bin/bash
# round-2: ncu launch list of the default bench command (per-launch durations, serialised, cold cache)
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/plain_r2q.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/ncu_r2q.log 2>&1; tail -2 gpurun_out/ncu_r2q.log | cut -c1-200
python - <<'PY'
import csv, collections, re
rows = [r for r in csv.reader(open('gpurun_out/r02_launches_bench.csv')) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value')
t = collections.Counter(); n = collections.Counter()
for r in rows[1:]:
    name = re.sub(r'\(.*', '', r[ki]).replace('unnamed>::', '').replace('void ', '')
    try: t[name] += float(r[vi].replace(',', '')); n[name] += 1
    except ValueError: pass
tot = sum(t.values())
for k, v in t.most_common(14): print(f"{k[:60]:60s} {n[k]:4d} launches {v/1e6:9.3f} ms {100*v/tot:5.1f}%")
PY
