#!/bin/bash
# round-2 GPU session 10 (two GPUs): SPIHT decoder with run skipping, pipeline pool spanning devices
mkdir -p gpurun_out
P=gpurun_out/r02j
nvidia-smi -L > ${P}_gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
timeout 300 python tools/time_modes.py max 0.5 24 1 > ${P}_plain.log 2>&1 &&
EBCC_CUDA_DEVICE=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.5_24frames.csv python tools/time_modes.py max 0.5 24 1 > ${P}_ncu.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.5_24frames.csv 12 > ${P}_launches_max0.5_24frames_summary.txt 2>&1; cat ${P}_launches_max0.5_24frames_summary.txt
# one call over both GPUs against the same call on one (host buffers in, streams out)
for devs in 0 0,1; do
  EBCC_CUDA_DEVICES=$devs EBCC_PIPELINES=12 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 600 python tools/time_modes.py max 0.1 576 3 > ${P}_time_max0.1_576_dev$devs.log 2>&1
  echo "devices $devs"; tail -4 ${P}_time_max0.1_576_dev$devs.log
done
du -sm gpurun_out
echo done
