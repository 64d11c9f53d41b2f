#!/bin/bash
# GPU call 45 (1 GPU): CG vector kernels with the first trip requested above the reductions (config 6) + implicit parity
mkdir -p gpurun_out; rm -f gpurun_out/c45.log
for v in 1 2; do
  timeout 300 python bench.py --config 6 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c45_c6.json 2>/dev/null
  grep -o '"phases_ms_per_step": {[^}]*}' gpurun_out/c45_c6.json >> gpurun_out/c45.log
done
cat gpurun_out/c45.log
timeout 600 python -m pytest tests/test_implicit.py tests/test_live_edits.py -m gpu -q 2>&1 | tail -3
