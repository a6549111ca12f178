#!/bin/bash
# round-2 GPU session 3: parity after the PCRD / stream-priority changes, then a sweep of the scheduling knobs on the bench's value region
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_none.py tests/test_gpu_residual.py tests/test_gpu_tiles.py tests/test_gpu_pipelines.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/r02c_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02c_pytest.log
tail -4 gpurun_out/r02c_pytest.log
for e in 0.5 0.1; do
  EBCC_PROFILE=1 timeout 300 python tools/time_modes.py max $e 24 3 > gpurun_out/r02c_time_max${e}_24.log 2>&1
  tail -6 gpurun_out/r02c_time_max${e}_24.log | head -2
done
run() {
  label=$1; shift
  env "$@" timeout 300 python bench.py --only-value --steps 4 --warmup 2 > gpurun_out/r02c_$label.json 2> gpurun_out/r02c_$label.err
  python - "$label" <<'PY'
import json, sys
lab = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/r02c_%s.json" % lab).read().strip().splitlines()[-1])
    k = d["kernels"]
    print("%-14s value %6.2f GB/s  ms/step %7.1f |" % (lab, d["value"], d["ms_per_step"]), " ".join("%s %.0f" % (n, k[n]["ms_per_step"]) for n in sorted(k, key=lambda n: -k[n]["ms_per_step"])[:8]))
except Exception as e:
    print(lab, "FAILED", e)
PY
}
run default A=1
run nosplit EBCC_PRIO_SPLIT=0
run lanes16 EBCC_MQ_LANES=16
run lanes8 EBCC_MQ_LANES=8
run pipes16 EBCC_BENCH_PIPES=16
run pipes8 EBCC_BENCH_PIPES=8
run pipes24 EBCC_BENCH_PIPES=24
run dwtcw8 EBCC_DWT_CW=8
run t1rec3 EBCC_T1REC_CTAS=3
run pipes16cw8 EBCC_BENCH_PIPES=16 EBCC_DWT_CW=8
echo done
