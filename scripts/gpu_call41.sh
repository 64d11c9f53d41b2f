#!/bin/bash
# GPU call 41 (1 GPU): sparse-product variants in the CG loop (config 6) + parity with the variants on
mkdir -p gpurun_out; rm -f gpurun_out/c41.log
for v in 0 1 2 0 1 2; do
  echo "== SPGX_SPMV_VARIANT=$v" >> gpurun_out/c41.log
  SPGX_SPMV_VARIANT=$v timeout 300 python bench.py --config 6 --steps 4 --warmup 2 --no-cpu-baseline 2>/dev/null | grep -o '"phases_ms_per_step": {[^}]*}' >> gpurun_out/c41.log
done
cat gpurun_out/c41.log
SPGX_SPMV_VARIANT=1 timeout 300 python -m pytest tests/test_implicit.py -m gpu -q 2>&1 | tail -3
SPGX_SPMV_VARIANT=2 timeout 300 python -m pytest tests/test_implicit.py -m gpu -q 2>&1 | tail -3
