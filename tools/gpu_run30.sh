#!/bin/bash
# round-2 GPU session 30: branch-free MQ decision, k_t1_dec<1> capped at 64 registers (8 CTAs per SM)
mkdir -p gpurun_out
P=gpurun_out/r02dec
timeout 900 python -m pytest tests/test_gpu_none.py tests/test_gpu_residual.py tests/test_gpu_pipelines.py tests/test_gpu_zstd.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/time_modes.py max 0.1 24 3 2>&1 | tail -4
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.1_24frames.csv python tools/time_modes.py max 0.1 24 1 > ${P}_ncu1.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.1_24frames.csv 4
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file ${P}_launches_none_1frame.csv python tools/time_modes.py none 0 1 1 > ${P}_ncu2.log 2>&1
python tools/launch_summary.py ${P}_launches_none_1frame.csv 4
export EBCC_BATCH_FRAMES=24 EBCC_PIPELINES=16
timeout 300 python tools/time_modes.py max 0.1 384 3 2>&1 | grep "rep" | tail -4
timeout 300 python tools/time_modes.py none 0 384 3 2>&1 | grep "decode rep" | tail -2
echo done
