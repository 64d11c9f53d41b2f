#!/bin/bash
# GPU call 23 (1 GPU): programmatic dependent launch for the XPBD passes: parity + config 2 A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -k "xpbd or golden or facade or smoke" > gpurun_out/pytest23.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest23.log; tail -3 gpurun_out/pytest23.log
for P in 0 1; do
SPGX_XPBD_PDL=$P timeout 300 python bench.py --config 2 --steps 20 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('pdl=$P', round(d['ms_per_step'],3), round(d['value']/1e9,3), round(d['roofline']['avg_launch_us'],2), round(d['e2e']['ms_per_step'],3), d['gpu_launches'])
" | tee -a gpurun_out/sweep23.txt
done
