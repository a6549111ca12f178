#!/bin/bash
# round-2 GPU session 21: pinned staging ring for pageable caller memory (uploads and decoded frames)
mkdir -p gpurun_out
P=gpurun_out/r02u
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
EBCC_PIPELINES=16 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 EBCC_SPIN_SYNC=1 timeout 600 python tools/time_modes.py max 0.1 384 3 2>&1 | tail -5
EBCC_PIPELINES=16 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 EBCC_SPIN_SYNC=1 timeout 600 python tools/time_modes.py none 0 384 3 2>&1 | tail -5
timeout 1200 python bench.py --steps 12 --warmup 3 --no-cpu > ${P}_bench.json 2> ${P}_bench.err
echo "bench rc=$?"; tail -c 400 ${P}_bench.err; python -c "
import json; d=json.loads(open('${P}_bench.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','e2e','e2e_pageable','cr')}); print({k: v for k, v in d['decompress'].items() if k != 'kernels'}); print(d['modes'])"
echo done
