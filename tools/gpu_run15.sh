#!/bin/bash
# round-2 GPU session 15: capped search probes, division-free first DWT level, hoisted column indices
mkdir -p gpurun_out
P=gpurun_out/r02o
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
timeout 1200 python bench.py --steps 12 --warmup 3 --no-cpu > ${P}_bench.json 2> ${P}_bench.err
echo "bench rc=$?"; tail -c 600 ${P}_bench.err; python -c "
import json; d=json.loads(open('${P}_bench.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value','ms_per_step','e2e','e2e_pageable','cr','gpu_launches')}); print({k: v for k, v in d['decompress'].items() if k != 'kernels'})
k=d['kernels']; print('enc:', ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])))
k=d['decompress']['kernels']; print('dec:', ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])))"
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:k_idwt_tile|k_dwt_tile' -s 0 -c 11 -o ${P}_fused python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_fused.log 2>&1
python tools/ncu_extract.py ${P}_fused.ncu-rep ${P}_ncu_fused.md "ncu --set full, round 2: fused forward / inverse 9/7 levels (24 frames 721x1440, MAX_ERROR 0.1, one pipeline)" > ${P}_extract.log 2>&1; tail -6 ${P}_extract.log
ncu -i ${P}_fused.ncu-rep --page source --csv --print-source cuda,sass > ${P}_fused_source.csv 2>/dev/null
rm -f ${P}_fused.ncu-rep
du -sm gpurun_out
echo done
