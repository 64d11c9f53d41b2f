#!/bin/bash
# GPU call 46 (1 GPU): hand-derived hinge Hessian in the block-sparse assembly: config 6 + implicit parity + generic fallback
mkdir -p gpurun_out; rm -f gpurun_out/c46.log
for h in 1 0; do
  echo "== SPGX_HINGE_HAND=$h" >> gpurun_out/c46.log
  SPGX_HINGE_HAND=$h timeout 300 python bench.py --config 6 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c46_c6_$h.json 2>/dev/null
  grep -o '"phases_ms_per_step": {[^}]*}' gpurun_out/c46_c6_$h.json >> gpurun_out/c46.log
done
cat gpurun_out/c46.log
timeout 600 python -m pytest tests/test_implicit.py tests/test_live_edits.py tests/test_meshio.py -m gpu -q 2>&1 | tail -3
