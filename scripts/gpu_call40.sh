#!/bin/bash
# GPU call 40 (8 GPUs): scaling refresh with the current kernel: config 5 (16.8M-vertex cloth) at 1, 4 and 8 GPUs, weak scaling at 8
mkdir -p gpurun_out
export SPGX_HALO_TIMEOUT_S=5
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29601"
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29602"
CUDA_VISIBLE_DEVICES=4 timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --config 5 > gpurun_out/b40_c5_n1.json 2> gpurun_out/b40_c5_n1.err &
P1=$!
CUDA_VISIBLE_DEVICES=5 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b40_c3_n1.json 2> gpurun_out/b40_c3_n1.err &
P2=$!
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 500 $TR4 bench.py --gpus 4 --steps 5 --warmup 3 --config 5 > gpurun_out/b40_c5_n4.json 2> gpurun_out/b40_c5_n4.err
wait $P1 $P2
timeout 500 $TR8 bench.py --gpus 8 --steps 5 --warmup 3 --config 5 > gpurun_out/b40_c5_n8.json 2> gpurun_out/b40_c5_n8.err
timeout 300 $TR8 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/b40_w8.json 2> gpurun_out/b40_w8.err
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b40_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, d['n_gpus'], d['config']['vertices'], round(d['ms_per_step'], 3), round(d['value'] / 1e9, 3), round(d['roofline']['avg_launch_us'], 2), round(d['e2e']['value']/1e9, 3))
PY
tail -n 2 gpurun_out/b40_*.err | cut -c1-200 | tail -20
