#!/bin/bash
# round-2 GPU session 4: whole parity suite, both bench arms, launch list of the bench's value region, ncu --set full of the
# HBM-bound stages and of the serial kernels (24 frames MAX_ERROR 0.1, one pipeline)
mkdir -p gpurun_out
P=gpurun_out/r02d
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > ${P}_gpu.txt 2>&1
nproc >> ${P}_gpu.txt
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log
tail -5 ${P}_pytest.log
for e in 0.5 0.1; do
  EBCC_PROFILE=1 timeout 300 python tools/time_modes.py max $e 24 3 > ${P}_time_max${e}_24.log 2>&1
  tail -7 ${P}_time_max${e}_24.log
done
timeout 300 python tools/time_modes.py max 0.5 1 4 > ${P}_time_max0.5_1.log 2>&1; tail -8 ${P}_time_max0.5_1.log
timeout 300 python tools/time_modes.py none 0 1 4 > ${P}_time_none_1.log 2>&1; tail -8 ${P}_time_none_1.log
timeout 900 python bench.py --impl reference --steps 6 --warmup 1 > ${P}_bench_ref.json 2> ${P}_bench_ref.err
echo "ref rc=$?"; cat ${P}_bench_ref.json | head -c 1200
timeout 1200 python bench.py --steps 20 --warmup 3 > ${P}_bench.json 2> ${P}_bench.err
echo "bench rc=$?"; tail -c 1500 ${P}_bench.err; head -c 7000 ${P}_bench.json
timeout 300 python bench.py --steps 1 --warmup 1 --only-value > ${P}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_bench_value.csv python bench.py --steps 1 --warmup 1 --only-value > ${P}_ncu.log 2>&1
python tools/launch_summary.py ${P}_launches_bench_value.csv > ${P}_launches_bench_value_summary.txt 2>&1; head -40 ${P}_launches_bench_value_summary.txt
# --set full: memory-bound stages
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_minmax$|k_dwt_|k_idwt_|k_finalize|k_dequant|k_recon_dequant|k_sp_load|k_sp_dwt_|k_sp_idwt_|k_sp_recon|k_sp_errstat|k_sp_desc|k_sp_normalize' -c 60 -o ${P}_hbm python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_hbm.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_t1_dec|k_pcrd_t2|k_sp_encode|k_sp_decode|k_t2_parse|k_t1_cm|k_t1_mq|k_zstd' -c 70 -o ${P}_serial python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_serial.log 2>&1
ls -la gpurun_out
echo done
