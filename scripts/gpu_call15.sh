#!/bin/bash
# GPU call 15 (1 GPU): fused cooperative XPBD frame: parity (whole XPBD + golden suites), config 2 A/B
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "xpbd or golden or facade" > gpurun_out/pytest15.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest15.log
tail -12 gpurun_out/pytest15.log
SPGX_XPBD_FUSED=0 timeout 300 python bench.py --config 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b15_c2_passes.json 2>&1
SPGX_XPBD_FUSED=1 timeout 300 python bench.py --config 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b15_c2_fused.json 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b15_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, round(d['ms_per_step'], 3), round(d['value'] / 1e9, 3), d['gpu_launches'], round(d['e2e']['value']/1e9, 3))
        elif 'rror' in ln: print(f, ln[:300])
PY
