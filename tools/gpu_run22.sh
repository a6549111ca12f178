#!/bin/bash
# round-2 GPU session 22 (eight GPUs): the bench under torch.distributed.run with eight ranks, as the driver's scaling run does
mkdir -p gpurun_out
P=gpurun_out/r02v
nproc; free -g | head -2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 8 --warmup 3 > ${P}_bench_8gpu.json 2> ${P}_bench_8gpu.err
echo "torchrun rc=$?"; tail -c 600 ${P}_bench_8gpu.err; python -c "
import json; d=json.loads(open('${P}_bench_8gpu.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('n_gpus','value','ms_per_step','e2e','e2e_pageable','cr','scaling','clocks')}); print({k: v for k, v in d['decompress'].items() if k != 'kernels'}); print(d['modes']); print(d['violations'])"
echo done
