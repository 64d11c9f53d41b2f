#!/bin/bash
# GPU call 43 (1 GPU): pipelined sparse product (variant 3) against variants 0 and 2 (config 6) + parity
mkdir -p gpurun_out; rm -f gpurun_out/c43.log
for v in 0 2 3 2 3; do
  echo "== SPGX_SPMV_VARIANT=$v" >> gpurun_out/c43.log
  SPGX_SPMV_VARIANT=$v timeout 300 python bench.py --config 6 --steps 4 --warmup 2 --no-cpu-baseline 2>/dev/null | grep -o '"phases_ms_per_step": {[^}]*}' >> gpurun_out/c43.log
done
cat gpurun_out/c43.log
SPGX_SPMV_VARIANT=3 timeout 300 python -m pytest tests/test_implicit.py -m gpu -q 2>&1 | tail -3
