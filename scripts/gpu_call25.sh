#!/bin/bash
# GPU call 26 (1 GPU): mixed-dual element Hessians, two-block sparse product: parity, bench config 6, launch list
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_implicit.py -m gpu -x -q > gpurun_out/pytest26.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest26.log; tail -3 gpurun_out/pytest26.log
timeout 600 python bench.py --config 6 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b26_c6.json 2> gpurun_out/b26_c6.err; python -c "
import json
d=json.loads(open('gpurun_out/b26_c6.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['phases_ms_per_step'], d['roofline']['frac'], d['roofline']['avg_launch_us'])"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches26_c6.csv \
   python bench.py --config 6 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu26.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/launches26_c6.csv')) if len(r) > 10]
hdr = rows[0]; ix = {h: i for i, h in enumerate(hdr)}
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for r in rows[1:]:
    name = r[ix['Kernel Name']][:60]; m = r[ix['Metric Name']]; v = float(r[ix['Metric Value']].replace(',', ''))
    if m == 'gpu__time_duration.sum': agg[name][0] += 1; agg[name][1] += v
    elif m.startswith('dram__bytes'): agg[name][2] += v
for k, (n, t, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]): print(k, n, round(t / n / 1e3, 2), 'us avg', round(b / max(n, 1) / 1e6, 1), 'MB/launch')
PY
