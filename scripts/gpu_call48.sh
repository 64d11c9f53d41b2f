#!/bin/bash
# GPU call 48 (1 GPU): VBD streaming kernel with the opt-in L1 prefetch compiled out of the incidence loop: config 3 x2 + step parity
mkdir -p gpurun_out; rm -f gpurun_out/c48.log
for i in 1 2; do
  timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | grep -o '"ms_per_step": [0-9.]*' | head -1 >> gpurun_out/c48.log
done
cat gpurun_out/c48.log
timeout 600 python -m pytest tests/test_gpu_step.py -m gpu -q 2>&1 | tail -3
