#!/bin/bash
# round-2 GPU session 18: whole-probe spans against the sum of their launch groups
mkdir -p gpurun_out
P=gpurun_out/r02r
EBCC_KT_WHOLE_PROBE=1 timeout 300 python bench.py --only-value --steps 6 --warmup 2 > ${P}_whole.json 2> ${P}_whole.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02r_whole.json").read().strip().splitlines()[-1]); k = d["kernels"]
print("value %.2f ms/step %.1f" % (d["value"], d["ms_per_step"]))
for n in sorted(k, key=lambda n: -k[n]["ms_per_step"]): print("  %-10s %7.1f ms/step  %5.1f groups/step  %.2f ms/group" % (n, k[n]["ms_per_step"], k[n]["launch_groups_per_step"], k[n]["ms_per_launch_group"]))
PY
echo done
