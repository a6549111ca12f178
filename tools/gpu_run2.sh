#!/bin/bash
# round-2 GPU session 2: changed tests, the new bench (both arms), launch list of the bench's value region
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_residual.py tests/test_dropin.py tests/test_gpu_tiles.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/r02b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02b_pytest.log
tail -4 gpurun_out/r02b_pytest.log
for e in 0.5 0.1; do
  EBCC_PROFILE=1 timeout 300 python tools/time_modes.py max $e 24 3 > gpurun_out/r02b_time_max${e}_24.log 2>&1
  tail -7 gpurun_out/r02b_time_max${e}_24.log
done
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err
echo "bench rc=$?"; tail -c 1500 gpurun_out/r02b_bench.err; head -c 6000 gpurun_out/r02b_bench.json
timeout 300 python bench.py --steps 2 --warmup 1 --only-value > gpurun_out/r02b_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02b_launches_bench_value.csv python bench.py --steps 2 --warmup 1 --only-value > gpurun_out/r02b_ncu.log 2>&1
echo done
