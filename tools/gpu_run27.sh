#!/bin/bash
# round-2 GPU session 27: code-blocks per warp of the tier-1 decoder at 32 blocks per SM (one 384-frame call, 16 pipelines)
mkdir -p gpurun_out
export EBCC_BATCH_FRAMES=24 EBCC_PIPELINES=16
for l in 2 1 4; do
  echo "== EBCC_DEC_LANES=$l"
  EBCC_DEC_LANES=$l timeout 300 python tools/time_modes.py max 0.1 384 3 2>&1 | grep "decode rep" | tail -2
done
echo "== NONE"
for l in 2 1; do
  EBCC_DEC_LANES=$l timeout 300 python tools/time_modes.py none 0 384 3 2>&1 | grep "decode rep" | tail -1
done
echo done
