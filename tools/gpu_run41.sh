#!/bin/bash
# round-2 GPU session 41: pool of unproven cuts decoded from the bit-plane boundary -- full suite, launch list, bench value, latencies
mkdir -p gpurun_out
P=gpurun_out/r02cut
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1; echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
timeout 300 python tools/time_modes.py max 0.1 24 3 2>&1 | grep "rep 2"
timeout 300 python tools/time_modes.py max 0.1 1 3 2>&1 | grep "rep 2"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.1_24frames.csv python tools/time_modes.py max 0.1 24 1 > ${P}_ncu1.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.1_24frames.csv 8 | tee ${P}_launches_max0.1_24frames_summary.txt
for i in 1 2; do
timeout 300 python bench.py --only-value --steps 8 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']; print('value %.2f GB/s ms/step %.1f |' % (d['value'], d['ms_per_step']), ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])[:7]))"
done
EBCC_CUTS_FROM_PASS0=1 timeout 300 python bench.py --only-value --steps 8 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']; print('from pass 0: value %.2f GB/s ms/step %.1f |' % (d['value'], d['ms_per_step']), ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])[:7]))"
echo done
