#!/bin/bash
# GPU call 29 (1 GPU): full parity suite, smoke, benches of configs 3 and 6
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest29.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest29.log; tail -4 gpurun_out/pytest29.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke29.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke29.log; tail -2 gpurun_out/smoke29.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/b29_c3.json 2> gpurun_out/b29_c3.err
timeout 600 python bench.py --config 6 --steps 5 --warmup 3 > gpurun_out/b29_c6.json 2> gpurun_out/b29_c6.err; tail -2 gpurun_out/b29_c6.err
python -c "
import json
for f in ('b29_c3','b29_c6'):
    d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
    print(f, d['ms_per_step'], d.get('phases_ms_per_step'), d['roofline']['frac'], d['roofline']['avg_launch_us'], d['e2e']['ms_per_step'])"
