#!/bin/bash
# round-2 GPU session 5: the allocation-free entry point (alone, then after the whole suite), both bench arms, launch lists
# (bench value region; one 24-frame MAX_ERROR call, encode + decode), ncu --set full of every stage, summarised on the box
mkdir -p gpurun_out
P=gpurun_out/r02e
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > ${P}_gpu.txt 2>&1
nproc >> ${P}_gpu.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > ${P}_pytest_noalloc_alone.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest_noalloc_alone.log; tail -3 ${P}_pytest_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -4 ${P}_pytest.log
timeout 900 python bench.py --impl reference --steps 4 --warmup 1 > ${P}_bench_ref.json 2> ${P}_bench_ref.err
echo "ref rc=$?"
timeout 1200 python bench.py --steps 20 --warmup 3 > ${P}_bench.json 2> ${P}_bench.err
echo "bench rc=$?"; tail -c 600 ${P}_bench.err; python -c "
import json; d=json.loads(open('${P}_bench.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','e2e','e2e_pageable','decompress','cr','gpu_launches')}); print(d['modes']); print(d.get('cpu_baseline'))"
timeout 300 python bench.py --steps 1 --warmup 1 --only-value > ${P}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_bench_value.csv python bench.py --steps 1 --warmup 1 --only-value > ${P}_ncu.log 2>&1
python tools/launch_summary.py ${P}_launches_bench_value.csv 40 > ${P}_launches_bench_value_summary.txt 2>&1
for e in 0.1 0.5; do
  timeout 300 python tools/time_modes.py max $e 24 1 > ${P}_plain_$e.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max${e}_24frames.csv python tools/time_modes.py max $e 24 1 > ${P}_ncu_$e.log 2>&1
  python tools/launch_summary.py ${P}_launches_max${e}_24frames.csv 50 > ${P}_launches_max${e}_24frames_summary.txt 2>&1; cat ${P}_launches_max${e}_24frames_summary.txt
done
NCU="timeout 900 ncu --set full --clock-control none --import-source on"
$NCU -k 'regex:k_minmax$|k_dwt_|k_idwt_|k_finalize|k_recon_dequant' -c 24 -o ${P}_hbm python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_hbm.log 2>&1
$NCU -k 'regex:k_sp_load|k_sp_subdc|k_sp_dwt_|k_sp_normalize|k_sp_desc|k_sp_recon|k_sp_idwt_|k_sp_errstat' -c 20 -o ${P}_spiht python tools/time_modes.py max 0.5 24 1 > ${P}_ncu_spiht.log 2>&1
$NCU -k 'regex:k_t1_cm|k_t1_mq|k_t1_dec_cuts|k_pcrd_t2|k_t2_bodies' -c 6 -o ${P}_enc_serial python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_enc_serial.log 2>&1
$NCU -k 'regex:k_sp_encode|k_zstd_compress|k_t2_parse|k_t1_dec$|k_t1_dec<|k_sp_decode|k_zstd_decompress|k_sp_add_residual|k_dequant' -c 8 -o ${P}_dec_serial python tools/time_modes.py max 0.5 24 1 > ${P}_ncu_dec_serial.log 2>&1
for n in hbm spiht enc_serial dec_serial; do
  python tools/ncu_extract.py ${P}_$n.ncu-rep ${P}_ncu_$n.md "ncu --set full, round 2: $n kernels (24 frames 721x1440, MAX_ERROR, one pipeline)" > ${P}_extract_$n.log 2>&1
  tail -12 ${P}_extract_$n.log
done
ls -la gpurun_out
# the reports are large: keep what fits into the 64 MiB that travel back (serial kernels first: their source pages are read offline)
for n in hbm spiht dec_serial enc_serial; do
  tot=$(du -sm gpurun_out | cut -f1)
  if [ "$tot" -gt 58 ]; then rm -f ${P}_$n.ncu-rep; fi
done
du -sm gpurun_out
echo done
