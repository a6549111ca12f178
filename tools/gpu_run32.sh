#!/bin/bash
# round-2 GPU sessions 32, 38, 44: the bench line with both arms on the end-of-round build
mkdir -p gpurun_out
P=gpurun_out/r02end
free -g | head -2; nproc
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > ${P}_bench_reference.json 2> ${P}_bench_reference.err; tail -c 400 ${P}_bench_reference.json; echo
timeout 1500 python bench.py > ${P}_bench.json 2> ${P}_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02end_bench.json").read().strip().splitlines()[-1])
print({k: d.get(k) for k in ("value", "ms_per_step", "steps", "cr", "gpu_launches")}, d["e2e"], d.get("e2e_pageable"), d.get("cpu_baseline"))
print({k: v for k, v in d["decompress"].items() if k != "kernels"}); print(d.get("parity")); print(d["modes"]); print(d["clocks"])
PY
echo done
