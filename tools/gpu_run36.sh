#!/bin/bash
# round-2 GPU session 36: source-level profile of k_t1_cm (recording variant, 3 CTAs per SM), the largest consumer of a pipelined step
mkdir -p gpurun_out
P=gpurun_out/r02cm
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:^k_t1_cm$' -s 0 -c 1 -o ${P}_cm python tools/time_modes.py max 0.1 24 1 > ${P}_ncu.log 2>&1
python tools/ncu_extract.py ${P}_cm.ncu-rep ${P}_ncu_cm.md "ncu --set full, end of round 2: k_t1_cm<recording, 3 CTAs per SM> (24 frames 721x1440, MAX_ERROR 0.1)" > ${P}_extract.log 2>&1; tail -4 ${P}_extract.log
ncu -i ${P}_cm.ncu-rep --page source --csv --print-source cuda,sass > ${P}_cm_source.csv 2>/dev/null
python tools/ncu_source_lines.py ${P}_cm_source.csv ${P}_source_lines_cm.md "k_t1_cm<recording>: hottest source lines (24 frames 721x1440, MAX_ERROR 0.1)" 70 > /dev/null 2>&1
rm -f ${P}_cm.ncu-rep; ls -la ${P}_*; du -sm gpurun_out
timeout 300 python bench.py --only-value --steps 8 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']; print('value %.2f GB/s ms/step %.1f |' % (d['value'], d['ms_per_step']), ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])[:7]))"
echo done
