#!/bin/bash
# round-2 GPU session 51: smoke(), launch lists of the final build (24-frame call; one bench step of one batch per pipeline)
mkdir -p gpurun_out
P=gpurun_out/r02fin
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.1_24frames.csv python tools/time_modes.py max 0.1 24 1 > ${P}_ncu1.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.1_24frames.csv 16 | tee ${P}_launches_max0.1_24frames_summary.txt
EBCC_BENCH_ROUNDS=1 timeout 300 python bench.py --only-value --steps 2 --warmup 1 > ${P}_bv.json 2> ${P}_bv.err &&
EBCC_BENCH_ROUNDS=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file ${P}_launches_bench_value.csv python bench.py --only-value --steps 2 --warmup 1 > ${P}_ncu3.log 2>&1
python tools/launch_summary.py ${P}_launches_bench_value.csv 16 | tee ${P}_launches_bench_value_summary.txt
echo done
