#!/bin/bash
# round-2 GPU session 16 (two GPUs): frame-fastest tile order (capped probes), huge-page decode outputs, the bench under
# torch.distributed.run with two ranks
mkdir -p gpurun_out
P=gpurun_out/r02p
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
EBCC_CUDA_DEVICE=0 timeout 300 python bench.py --only-value --steps 6 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']; print('value %.2f GB/s ms/step %.1f |' % (d['value'], d['ms_per_step']), ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])[:8]))"
EBCC_CUDA_DEVICE=0 EBCC_PIPELINES=12 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 600 python tools/time_modes.py max 0.1 288 3 2>&1 | tail -4
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 8 --warmup 3 > ${P}_bench_2gpu.json 2> ${P}_bench_2gpu.err
echo "torchrun rc=$?"; tail -c 400 ${P}_bench_2gpu.err; python -c "
import json; d=json.loads(open('${P}_bench_2gpu.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('n_gpus','value','ms_per_step','e2e','e2e_pageable','cr','scaling')}); print({k: v for k, v in d['decompress'].items() if k != 'kernels'}); print(d['modes']); print(d['violations'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 | cut -c1-300
echo done
