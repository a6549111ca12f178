#!/bin/bash
# round-2 GPU session 43: call size (batches per pipeline) and pipelines per GPU
mkdir -p gpurun_out
export EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32
for n in 384 768 1536; do
  echo "== $n frames, 16 pipelines (spin)"
  EBCC_PIPELINES=16 EBCC_SPIN_SYNC=1 timeout 600 python tools/time_modes.py max 0.1 $n 3 2>&1 | grep "rep 2"
done
echo "== 768 frames, 16 pipelines (blocking)"
EBCC_PIPELINES=16 EBCC_SPIN_SYNC=0 timeout 600 python tools/time_modes.py max 0.1 768 3 2>&1 | grep "rep 2"
for p in 20 24 32; do
  echo "== 768 frames, $p pipelines (blocking)"
  EBCC_PIPELINES=$p EBCC_SPIN_SYNC=0 timeout 600 python tools/time_modes.py max 0.1 768 3 2>&1 | grep "rep 2"
done
echo done
