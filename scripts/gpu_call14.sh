#!/bin/bash
# GPU call 14 (1 GPU): full parity suite with the four rank-1 "next" energies; persistent kernel at 16M vertices
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest14.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest14.log
tail -15 gpurun_out/pytest14.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke14.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke14.log; tail -2 gpurun_out/smoke14.log
SPGX_VBD_PERSIST=0 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --config 5 > gpurun_out/b14_c5_p0.json 2>&1
SPGX_VBD_PERSIST=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --config 5 > gpurun_out/b14_c5_p1.json 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b14_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, d['config']['vertices'], round(d['ms_per_step'], 3), round(d['value'] / 1e9, 3), round(d['roofline']['avg_launch_us'], 2), round(d['roofline']['frac'], 3))
        elif 'rror' in ln: print(f, ln[:200])
PY
