#!/bin/bash
# round-2 GPU session 6: fused inverse transform -- parity suite, timings, launch list, ncu of the fused kernels
mkdir -p gpurun_out
P=gpurun_out/r02f
timeout 1800 python -m pytest tests -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -4 ${P}_pytest.log
for e in 0.1 0.5; do
  EBCC_PROFILE=1 timeout 300 python tools/time_modes.py max $e 24 3 > ${P}_time_max${e}_24.log 2>&1; tail -5 ${P}_time_max${e}_24.log
done
timeout 300 python tools/time_modes.py none 0 24 3 > ${P}_time_none_24.log 2>&1; tail -4 ${P}_time_none_24.log
run() {
  label=$1; shift
  env "$@" timeout 300 python bench.py --only-value --steps 6 --warmup 2 > ${P}_$label.json 2> ${P}_$label.err
  python - "$label" <<'PY'
import json, sys
lab = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/r02f_%s.json" % lab).read().strip().splitlines()[-1])
    k = d["kernels"]
    print("%-14s value %6.2f GB/s  ms/step %7.1f |" % (lab, d["value"], d["ms_per_step"]), " ".join("%s %.0f" % (n, k[n]["ms_per_step"]) for n in sorted(k, key=lambda n: -k[n]["ms_per_step"])[:10]))
except Exception as e:
    print(lab, "FAILED", e)
PY
}
run fused A=1
run unfused EBCC_FUSED_IDWT=0
run fused_p16 EBCC_BENCH_PIPES=16
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.1_24frames.csv python tools/time_modes.py max 0.1 24 1 > ${P}_ncu.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.1_24frames.csv 30 > ${P}_launches_max0.1_24frames_summary.txt 2>&1; cat ${P}_launches_max0.1_24frames_summary.txt
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:k_idwt_fused' -c 11 -o ${P}_idwt_fused python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_fused.log 2>&1
python tools/ncu_extract.py ${P}_idwt_fused.ncu-rep ${P}_ncu_idwt_fused.md "ncu --set full, round 2: fused inverse 9/7 levels (24 frames 721x1440, MAX_ERROR 0.1, one pipeline)" > ${P}_extract.log 2>&1; tail -8 ${P}_extract.log
du -sm gpurun_out
echo done
