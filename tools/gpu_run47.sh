#!/bin/bash
# round-2 GPU session 47: no host copy of the first codestream (two pinned download buffers take turns)
mkdir -p gpurun_out
P=gpurun_out/r02h
timeout 900 python -m pytest tests/test_gpu_residual.py tests/test_gpu_pipelines.py tests/test_dropin.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -2
for i in 1 2 3; do
timeout 300 python bench.py --only-value --steps 8 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']; print('value %.2f GB/s ms/step %.1f |' % (d['value'], d['ms_per_step']), ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])[:7]))"
done
EBCC_PROFILE=1 timeout 300 python bench.py --only-value --steps 2 --warmup 1 2>&1 >/dev/null | grep "ebcc profile" | tail -6 | cut -c1-330
echo done
