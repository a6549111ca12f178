#!/bin/bash
# round-2 GPU session 19: one launch less per probe (statistics reset folded into k_jobs_from_enc)
mkdir -p gpurun_out
P=gpurun_out/r02s
timeout 900 python -m pytest tests/test_gpu_residual.py tests/test_gpu_tiles.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do
timeout 300 python bench.py --only-value --steps 8 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']; print('value %.2f GB/s ms/step %.1f | spans/batch %.0f |' % (d['value'], d['ms_per_step'], sum(v['ms_per_step'] for v in k.values())/16), ' '.join('%s %.0f' % (n, k[n]['ms_per_step']) for n in sorted(k, key=lambda n: -k[n]['ms_per_step'])[:7]))"
done
echo done
