#!/bin/bash
# GPU call 44 (1 GPU): pipelined sparse product with the row's own entries hoisted (default) + full implicit parity + linear-solve users
mkdir -p gpurun_out; rm -f gpurun_out/c44.log
for v in 3 3; do
  echo "== SPGX_SPMV_VARIANT=$v" >> gpurun_out/c44.log
  SPGX_SPMV_VARIANT=$v timeout 300 python bench.py --config 6 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c44_c6.json 2>/dev/null
  grep -o '"phases_ms_per_step": {[^}]*}' gpurun_out/c44_c6.json >> gpurun_out/c44.log
done
cat gpurun_out/c44.log
timeout 600 python -m pytest tests/test_implicit.py tests/test_live_edits.py tests/test_meshio.py -m gpu -q 2>&1 | tail -3
