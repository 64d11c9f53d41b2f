#!/bin/bash
# GPU call 28 (1 GPU): bench config 6 with per-frame reset
mkdir -p gpurun_out
timeout 600 python bench.py --config 6 --steps 5 --warmup 3 > gpurun_out/b28_c6.json 2> gpurun_out/b28_c6.err; tail -2 gpurun_out/b28_c6.err; python -c "
import json
d=json.loads(open('gpurun_out/b28_c6.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['phases_ms_per_step'], d['roofline']['frac'], d['roofline']['avg_launch_us'], d['e2e']['ms_per_step'], d['cpu_baseline'])"
