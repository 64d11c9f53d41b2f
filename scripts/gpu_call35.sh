#!/bin/bash
# GPU call 35 (1 GPU): persistent streaming variant re-measured with the resident-block carveout (config 3), next to the default
mkdir -p gpurun_out
for p in 0 1 0 1; do
  echo "== SPGX_VBD_PERSIST=$p" >> gpurun_out/c35.log
  SPGX_VBD_PERSIST=$p timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline >> gpurun_out/c35.log 2>&1
done
echo "== persist carveout 100" >> gpurun_out/c35.log
SPGX_VBD_PERSIST=1 SPGX_VBD_CARVEOUT=100 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline >> gpurun_out/c35.log 2>&1
grep -o '== .*\|"ms_per_step": [0-9.]*' gpurun_out/c35.log
