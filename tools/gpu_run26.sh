#!/bin/bash
# round-2 GPU session 26: tier-1 decoder with one-byte candidate words (4 CTAs per SM): tests, full bench line
mkdir -p gpurun_out
P=gpurun_out/r02z
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --no-cpu --steps 8 --warmup 3 > ${P}_bench.json 2> ${P}_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02z_bench.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "pageable", d.get("e2e_pageable"), "decompress", d.get("decompress"))
PY
python tools/time_modes.py max 0.1 24 3 2>&1 | tail -6
python tools/time_modes.py none 0 1 4 2>&1 | tail -8
echo done
