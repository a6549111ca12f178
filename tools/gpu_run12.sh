#!/bin/bash
# round-2 GPU session 12: everything in -- whole parity suite, both bench arms, launch list of the bench's value region
mkdir -p gpurun_out
P=gpurun_out/r02l
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > ${P}_gpu.txt 2>&1; nproc >> ${P}_gpu.txt
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
timeout 900 python bench.py --impl reference --steps 4 --warmup 1 > ${P}_bench_ref.json 2> ${P}_bench_ref.err
echo "ref rc=$?"
timeout 1200 python bench.py --steps 20 --warmup 3 > ${P}_bench.json 2> ${P}_bench.err
echo "bench rc=$?"; tail -c 600 ${P}_bench.err; python -c "
import json; d=json.loads(open('${P}_bench.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','e2e','e2e_pageable','decompress','cr','gpu_launches','violations','parity')}); print(d['modes']); print(d.get('cpu_baseline')); print(d['roofline'])
k=d['kernels']; print(' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])))"
for pp in 12 24; do
EBCC_BENCH_PIPES=$pp timeout 300 python bench.py --only-value --steps 6 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('pipes $pp value %.2f GB/s ms/step %.1f' % (d['value'], d['ms_per_step']))"
done
timeout 300 python bench.py --steps 1 --warmup 1 --only-value > ${P}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file ${P}_launches_bench_value.csv python bench.py --steps 1 --warmup 1 --only-value > ${P}_ncu.log 2>&1
python tools/launch_summary.py ${P}_launches_bench_value.csv 40 > ${P}_launches_bench_value_summary.txt 2>&1; head -24 ${P}_launches_bench_value_summary.txt
du -sm gpurun_out
echo done
