#!/bin/bash
# round-2 GPU session 29: end-of-round evidence -- launch lists (24-frame call, one frame, bench value), ncu full of the decoder
mkdir -p gpurun_out
P=gpurun_out/r02final
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.1_24frames.csv python tools/time_modes.py max 0.1 24 1 > ${P}_ncu1.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.1_24frames.csv 16 > ${P}_launches_max0.1_24frames_summary.txt 2>&1; cat ${P}_launches_max0.1_24frames_summary.txt
timeout 300 python tools/time_modes.py none 0 1 1 > ${P}_plain1.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file ${P}_launches_none_1frame.csv python tools/time_modes.py none 0 1 1 > ${P}_ncu2.log 2>&1
python tools/launch_summary.py ${P}_launches_none_1frame.csv 30 > ${P}_launches_none_1frame_summary.txt 2>&1; cat ${P}_launches_none_1frame_summary.txt
timeout 300 python bench.py --only-value --steps 2 --warmup 1 > ${P}_bv.json 2> ${P}_bv.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file ${P}_launches_bench_value.csv python bench.py --only-value --steps 2 --warmup 1 > ${P}_ncu3.log 2>&1
python tools/launch_summary.py ${P}_launches_bench_value.csv 16 > ${P}_launches_bench_value_summary.txt 2>&1; cat ${P}_launches_bench_value_summary.txt
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:^k_t1_dec$' -s 0 -c 1 -o ${P}_t1dec python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_t1dec.log 2>&1
python tools/ncu_extract.py ${P}_t1dec.ncu-rep ${P}_ncu_t1dec.md "ncu --set full, end of round 2: k_t1_dec<1> with one-byte candidate words (24 frames 721x1440, MAX_ERROR 0.1 decode)" > ${P}_extract.log 2>&1; tail -4 ${P}_extract.log
rm -f ${P}_t1dec.ncu-rep
du -sm gpurun_out
echo done
