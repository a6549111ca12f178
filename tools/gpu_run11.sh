#!/bin/bash
# round-2 GPU session 11: tier-1 decoder with candidate words -- parity suite, decode timings, launch list
mkdir -p gpurun_out
P=gpurun_out/r02k
timeout 1500 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
for m in "max 0.1" "max 0.5" "none 0"; do
  timeout 300 python tools/time_modes.py $m 24 3 2>&1 | tail -4
done
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.1_24frames.csv python tools/time_modes.py max 0.1 24 1 > ${P}_ncu.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.1_24frames.csv 10 > ${P}_launches_max0.1_24frames_summary.txt 2>&1; cat ${P}_launches_max0.1_24frames_summary.txt
EBCC_PIPELINES=12 EBCC_BATCH_FRAMES=24 CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 600 python tools/time_modes.py max 0.1 288 3 2>&1 | tail -4
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:^k_t1_dec$' -c 1 -o ${P}_t1dec python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_t1dec.log 2>&1
ncu -i ${P}_t1dec.ncu-rep --page source --csv --print-source cuda,sass > ${P}_t1dec_source.csv 2>/dev/null
python tools/ncu_extract.py ${P}_t1dec.ncu-rep ${P}_ncu_t1dec.md "ncu --set full, round 2: k_t1_dec<1> with candidate words (24 frames 721x1440, MAX_ERROR 0.1 decode)" > ${P}_extract2.log 2>&1; tail -4 ${P}_extract2.log
rm -f ${P}_t1dec.ncu-rep
echo done
