#!/bin/bash
# experiment: what one more launch group of a class costs the pipelined step (EBCC_DEBUG_REPEAT)
mkdir -p gpurun_out
P=gpurun_out/r02exp
run() {
  label=$1; shift
  env "$@" timeout 300 python bench.py --only-value --steps 8 --warmup 2 > ${P}_$label.json 2> ${P}_$label.err
  python - "$label" <<'PY'
import json, sys
lab = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/r02exp_%s.json" % lab).read().strip().splitlines()[-1])
    k = d["kernels"]
    print("%-14s value %6.2f GB/s  ms/step %7.1f |" % (lab, d["value"], d["ms_per_step"]), " ".join("%s %.0f" % (n, k[n]["ms_per_step"]) for n in sorted(k, key=lambda n: -k[n]["ms_per_step"])[:7]))
except Exception as e:
    print(lab, "FAILED", e)
PY
}
run base A=1
run mq16 EBCC_MQ_LANES=16
run mq8 EBCC_MQ_LANES=8
run p12 EBCC_BENCH_PIPES=12
run p20 EBCC_BENCH_PIPES=20
run r3 EBCC_BENCH_ROUNDS=3
run base2 A=1
echo done
