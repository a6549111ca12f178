#!/bin/bash
# round-2 GPU session 13: scheduling knobs on the bench's value region (16 pipelines)
mkdir -p gpurun_out
P=gpurun_out/r02m
run() {
  label=$1; shift
  env "$@" timeout 300 python bench.py --only-value --steps 6 --warmup 2 > ${P}_$label.json 2> ${P}_$label.err
  python - "$label" <<'PY'
import json, sys
lab = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/r02m_%s.json" % lab).read().strip().splitlines()[-1])
    k = d["kernels"]
    print("%-14s value %6.2f GB/s  ms/step %7.1f |" % (lab, d["value"], d["ms_per_step"]), " ".join("%s %.0f" % (n, k[n]["ms_per_step"]) for n in sorted(k, key=lambda n: -k[n]["ms_per_step"])[:8]))
except Exception as e:
    print(lab, "FAILED", e)
PY
}
run default A=1
run lanes16 EBCC_MQ_LANES=16
run lanes8 EBCC_MQ_LANES=8
run lanes4 EBCC_MQ_LANES=4
run nosplit EBCC_PRIO_SPLIT=0
run t1rec3 EBCC_T1REC_CTAS=3
run batch12x32 EBCC_BENCH_BATCH=12 EBCC_BENCH_PIPES=32
run batch48x8 EBCC_BENCH_BATCH=48 EBCC_BENCH_PIPES=8
echo done
