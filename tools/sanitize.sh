#!/bin/bash
# compute-sanitizer over small runs of every kernel family (the reference has no race detection; the
# new build carries it).  Usage (on a GPU box): bash tools/sanitize.sh <out-dir>   -- one log per (tool, run), a summary line each.
out=${1:-gpurun_out/sanitizer}
mkdir -p "$out"
cd "$(dirname "$0")/.."
for tool in memcheck racecheck; do
  for run in vbd vbd_generic vbd_tets xpbd newton; do
    log="$out/${tool}_${run}.log"
    timeout 600 compute-sanitizer --tool $tool --error-exitcode 9 python tools/sanitizer_driver.py $run > "$log" 2>&1
    echo "$tool $run rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' "$log" | tail -1)"
  done
done | tee "$out/summary.txt"
# the 2-rank peer-memory exchange (system-scope flags, remote stores) under memcheck, when the box has 2 GPUs
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
  log="$out/memcheck_multi_2gpu.log"
  timeout 900 compute-sanitizer --tool memcheck --target-processes all --error-exitcode 9 python -m torch.distributed.run --nnodes=1 \
    --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29541 tests/multi_worker.py cloth /tmp/sanit_multi.npz > "$log" 2>&1
  echo "memcheck multi_2gpu rc=$? : $(grep -E 'ERROR SUMMARY' "$log" | tail -2 | tr '\n' ' ')" | tee -a "$out/summary.txt"
fi
