#!/bin/bash
# round-2 GPU session 20: where a batch's wall time goes with 16 pipelines in flight (EBCC_PROFILE per batch)
mkdir -p gpurun_out
P=gpurun_out/r02t
EBCC_PROFILE=1 EBCC_PIPELINES=16 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 EBCC_SPIN_SYNC=1 timeout 600 python tools/time_modes.py max 0.1 384 3 > ${P}_profile16.log 2>&1
grep "ebcc profile" ${P}_profile16.log | tail -16 | cut -c1-330; grep "encode rep\|decode rep" ${P}_profile16.log
EBCC_PROFILE=1 timeout 300 python tools/time_modes.py max 0.1 24 2 2>&1 | grep "ebcc profile\|rep" | cut -c1-330
echo done
