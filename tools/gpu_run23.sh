#!/bin/bash
# round-2 GPU session 23: per-frame cache of PCRD evaluations across the probes of a search
mkdir -p gpurun_out
P=gpurun_out/r02w
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
EBCC_PROFILE=1 timeout 300 python tools/time_modes.py max 0.1 24 3 2>&1 | tail -5 | cut -c1-250
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.1_24frames.csv python tools/time_modes.py max 0.1 24 1 > ${P}_ncu.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.1_24frames.csv 12 > ${P}_launches_max0.1_24frames_summary.txt 2>&1; cat ${P}_launches_max0.1_24frames_summary.txt
grep k_pcrd_t2 ${P}_launches_max0.1_24frames.csv | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' '; echo
for i in 1 2; do
timeout 300 python bench.py --only-value --steps 8 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']; print('value %.2f GB/s ms/step %.1f |' % (d['value'], d['ms_per_step']), ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])[:7]))"
done
echo done
