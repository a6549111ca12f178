#!/bin/bash
# round-2 GPU session 1: parity suite, error-bounded timings, launch list
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02a_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02a_pytest.log
tail -5 gpurun_out/r02a_pytest.log
for e in 0.5 0.1; do
  EBCC_PROFILE=1 timeout 300 python tools/time_modes.py max $e 24 3 > gpurun_out/r02a_time_max${e}_24.log 2>&1
  tail -8 gpurun_out/r02a_time_max${e}_24.log
done
EBCC_PIPELINES=16 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 600 python tools/time_modes.py max 0.1 384 3 > gpurun_out/r02a_time_max0.1_384.log 2>&1
tail -6 gpurun_out/r02a_time_max0.1_384.log
EBCC_PIPELINES=16 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 600 python tools/time_modes.py max 0.5 384 3 > gpurun_out/r02a_time_max0.5_384.log 2>&1
tail -6 gpurun_out/r02a_time_max0.5_384.log
timeout 300 python tools/time_modes.py max 0.1 24 1 > gpurun_out/r02a_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02a_launches_max01_24frames.csv python tools/time_modes.py max 0.1 24 1 > gpurun_out/r02a_ncu.log 2>&1
echo done
