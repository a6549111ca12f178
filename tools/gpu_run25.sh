#!/bin/bash
# round-2 GPU session 25: byte budgets / targets read from pinned host memory, per-batch host-side breakdown under 16 pipelines
mkdir -p gpurun_out
P=gpurun_out/r02y
timeout 900 python -m pytest tests/test_gpu_none.py tests/test_gpu_residual.py tests/test_gpu_pipelines.py -m gpu -x -q 2>&1 | tail -2
run() {
  label=$1; shift
  env "$@" timeout 300 python bench.py --only-value --steps 8 --warmup 2 > ${P}_$label.json 2> ${P}_$label.err
  python - "$label" <<'PY'
import json, sys
lab = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/r02y_%s.json" % lab).read().strip().splitlines()[-1])
    k = d["kernels"]
    print("%-14s value %6.2f GB/s  ms/step %7.1f |" % (lab, d["value"], d["ms_per_step"]), " ".join("%s %.0f" % (n, k[n]["ms_per_step"]) for n in sorted(k, key=lambda n: -k[n]["ms_per_step"])[:7]))
except Exception as e:
    print(lab, "FAILED", e)
PY
}
run zc1 A=1
run zc0 EBCC_ZERO_COPY=0
run zc1b A=1
EBCC_PROFILE=1 timeout 300 python bench.py --only-value --steps 3 --warmup 1 > ${P}_prof.json 2> ${P}_prof.err
grep "ebcc profile" ${P}_prof.err | tail -24
echo done
