#!/bin/bash
# GPU call 31 (2 GPUs): k-GPU == 1-GPU identity with the current kernel, 2-GPU weak scaling + config 4 / 5
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/pytest31.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest31.log; tail -3 gpurun_out/pytest31.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b31_w1.json 2> gpurun_out/b31_w1.err
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/b31_w2_p2p.json 2> gpurun_out/b31_w2_p2p.err
timeout 300 $TR bench.py --gpus 2 --steps 10 --warmup 3 --config 4 > gpurun_out/b31_c4_n2.json 2> gpurun_out/b31_c4_n2.err
timeout 400 $TR bench.py --gpus 2 --steps 5 --warmup 3 --config 5 > gpurun_out/b31_c5_n2.json 2> gpurun_out/b31_c5_n2.err
timeout 300 $TR bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/b31_ref_n2.json 2> gpurun_out/b31_ref_n2.err
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b31_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, d['n_gpus'], d['config'].get('vertices'), round(d['ms_per_step'], 3), round(d['value'] / 1e9, 4))
    
PY
tail -3 gpurun_out/b31_w2_p2p.err
