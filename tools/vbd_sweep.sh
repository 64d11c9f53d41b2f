#!/bin/bash
# config 3 under the knobs of the streaming colour pass: one line per variant (args: env assignments, one variant per argument, ',' separated)
cd "$(dirname "$0")/.."
run() {
  echo "$*"
  env "$@" python bench.py --config ${CONFIG:-3} --no-cpu-baseline --no-other-configs --steps 20 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('  ms/step', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4), 'launch_us', round(d['roofline']['avg_launch_us'],2), 'e2e_ms', round(d['e2e'].get('ms_per_step',0),4))"
}
for v in "$@"; do run ${v//,/ }; done
