#!/bin/bash
# round-2 GPU session 39 (two GPUs): end-of-round build -- multi-device test, the bench under torch.distributed.run with two ranks
mkdir -p gpurun_out
P=gpurun_out/r02end2
free -g | head -2; nproc
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 8 --warmup 3 > ${P}_bench_2gpu.json 2> ${P}_bench_2gpu.err
echo "torchrun rc=$?"; tail -c 300 ${P}_bench_2gpu.err; python -c "
import json; d=json.loads(open('${P}_bench_2gpu.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('n_gpus','value','ms_per_step','e2e','e2e_pageable','cr','scaling')}); print({k: v for k, v in d['decompress'].items() if k != 'kernels'}); print(d['modes']); print(d['violations'])"
EBCC_PIPELINES=16 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 EBCC_CUDA_DEVICES=0,1 timeout 600 python tools/time_modes.py max 0.1 768 3 2>&1 | grep "rep" | tail -4
echo done
