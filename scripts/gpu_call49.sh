#!/bin/bash
# GPU call 49 (1 GPU): gather-list assembly of the block-sparse Hessian: config 6 with and without + implicit parity
mkdir -p gpurun_out; rm -f gpurun_out/c49.log
for g in 1 0; do
  echo "== SPGX_ASM_GATHER=$g" >> gpurun_out/c49.log
  SPGX_ASM_GATHER=$g timeout 300 python bench.py --config 6 --steps 4 --warmup 2 --no-cpu-baseline > gpurun_out/c49_c6_$g.json 2>/dev/null
  grep -o '"phases_ms_per_step": {[^}]*}' gpurun_out/c49_c6_$g.json >> gpurun_out/c49.log
done
cat gpurun_out/c49.log
timeout 600 python -m pytest tests/test_implicit.py tests/test_live_edits.py -m gpu -q 2>&1 | tail -3
