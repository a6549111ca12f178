#!/bin/bash
# round-2 GPU session 17: how the pipeline threads wait (spin / blocking) against the number of pipelines and host cores
mkdir -p gpurun_out
P=gpurun_out/r02q
nproc
run() {
  label=$1; shift
  env "$@" timeout 300 python bench.py --only-value --steps 6 --warmup 2 > ${P}_$label.json 2> ${P}_$label.err
  python - "$label" <<'PY'
import json, sys
lab = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/r02q_%s.json" % lab).read().strip().splitlines()[-1])
    k = d["kernels"]
    print("%-14s value %6.2f GB/s  ms/step %7.1f | sum of kernel spans per batch %.0f ms |" % (lab, d["value"], d["ms_per_step"], sum(v["ms_per_step"] for v in k.values()) / d["pipelines_per_gpu"]),
          " ".join("%s %.0f" % (n, k[n]["ms_per_step"]) for n in sorted(k, key=lambda n: -k[n]["ms_per_step"])[:6]))
except Exception as e:
    print(lab, "FAILED", e)
PY
}
run spin16 EBCC_SPIN_SYNC=1
run block16 EBCC_SPIN_SYNC=0
run spin14 EBCC_SPIN_SYNC=1 EBCC_BENCH_PIPES=14
run spin12 EBCC_SPIN_SYNC=1 EBCC_BENCH_PIPES=12
run block20 EBCC_SPIN_SYNC=0 EBCC_BENCH_PIPES=20
run block24 EBCC_SPIN_SYNC=0 EBCC_BENCH_PIPES=24
echo done
