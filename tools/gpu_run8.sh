#!/bin/bash
# round-2 GPU session 8: fused transforms with all loads of a tile row in flight; source-level profile of the tier-1 decoder
mkdir -p gpurun_out
P=gpurun_out/r02h
timeout 900 python -m pytest tests/test_gpu_none.py tests/test_gpu_decode.py tests/test_gpu_residual.py tests/test_gpu_tiles.py -m gpu -x -q > ${P}_pytest.log 2>&1
echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
run() {
  label=$1; shift
  env "$@" timeout 300 python bench.py --only-value --steps 6 --warmup 2 > ${P}_$label.json 2> ${P}_$label.err
  python - "$label" <<'PY'
import json, sys
lab = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/r02h_%s.json" % lab).read().strip().splitlines()[-1])
    k = d["kernels"]
    print("%-14s value %6.2f GB/s  ms/step %7.1f |" % (lab, d["value"], d["ms_per_step"]), " ".join("%s %.0f" % (n, k[n]["ms_per_step"]) for n in sorted(k, key=lambda n: -k[n]["ms_per_step"])[:10]))
except Exception as e:
    print(lab, "FAILED", e)
PY
}
run fused A=1
run fused_p16 EBCC_BENCH_PIPES=16
run fused_p24 EBCC_BENCH_PIPES=24
timeout 300 python tools/time_modes.py max 0.1 24 1 > ${P}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file ${P}_launches_max0.1_24frames.csv python tools/time_modes.py max 0.1 24 1 > ${P}_ncu.log 2>&1
python tools/launch_summary.py ${P}_launches_max0.1_24frames.csv 14 > ${P}_launches_max0.1_24frames_summary.txt 2>&1; cat ${P}_launches_max0.1_24frames_summary.txt
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:k_idwt_tile|k_dwt_tile' -c 16 -o ${P}_fused python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_fused.log 2>&1
python tools/ncu_extract.py ${P}_fused.ncu-rep ${P}_ncu_fused.md "ncu --set full, round 2: fused forward / inverse 9/7 levels (24 frames 721x1440, MAX_ERROR 0.1, one pipeline)" > ${P}_extract.log 2>&1; tail -8 ${P}_extract.log
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:k_t1_dec<' -c 1 -o ${P}_t1dec python tools/time_modes.py max 0.1 24 1 > ${P}_ncu_t1dec.log 2>&1
ncu -i ${P}_t1dec.ncu-rep --page source --csv --print-source cuda,sass > ${P}_t1dec_source.csv 2>/dev/null
python tools/ncu_extract.py ${P}_t1dec.ncu-rep ${P}_ncu_t1dec.md "ncu --set full, round 2: k_t1_dec<1> (24 frames 721x1440, MAX_ERROR 0.1 decode)" > ${P}_extract2.log 2>&1
rm -f ${P}_t1dec.ncu-rep
du -sm gpurun_out
echo done
