#!/bin/bash
# config 2 (cloth 256^2 springs + quadratic bending, XPBD 20 x 1) under the measurement knobs of xpbd.cu: one line per variant
cd "$(dirname "$0")/.."
run() {
  echo "$*"
  env "$@" python bench.py --config 2 --no-cpu-baseline --no-other-configs --steps 30 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('  ms/step', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4), 'e2e_ms', round(d['e2e'].get('ms_per_step',0),4), 'launches', d['gpu_launches'])"
}
run SPGX_XPBD_HAND=0
run SPGX_XPBD_HAND=1
run SPGX_XPBD_HAND=1 SPGX_XPBD_PDL=1
run SPGX_XPBD_HAND=0 SPGX_XPBD_PDL=1
run SPGX_XPBD_HAND=1 SPGX_XPBD_THREADS=64
run SPGX_XPBD_HAND=1 SPGX_XPBD_THREADS=32
run SPGX_XPBD_HAND=1 SPGX_XPBD_THREADS=64 SPGX_XPBD_PDL=1
run SPGX_XPBD_DEBUG_EMPTY=1
run SPGX_XPBD_DEBUG_EMPTY=1 SPGX_XPBD_THREADS=32
