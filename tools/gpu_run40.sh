#!/bin/bash
# round-2 GPU session 40: compute-sanitizer (memcheck, racecheck) over small error-bounded encodes + decodes; single-frame and
# MAX_ERROR 0.5 figures
mkdir -p gpurun_out
P=gpurun_out/r02san
cat > /tmp/san_case.py <<'PY'
import ctypes as C, os, sys
sys.path.insert(0, "tests")
import numpy as np, harness as H
lib = H.prod_lib()
lib.ebcc_encode_batch.restype = C.c_size_t
for (n, h, w, cr, mode, err) in ((3, 96, 160, 12.0, H.MAX_ERROR, 0.05), (2, 181, 263, 40.0, H.MAX_ERROR, 0.02), (2, 70, 130, 8.0, H.NONE, 0.0), (2, 96, 160, 10.0, H.RELATIVE_ERROR, 1e-3)):
    frames = H.synth_stack(n, h, w)
    cfg = H.CodecConfig(); cfg.dims[:] = [1, h, w]; cfg.base_cr = cr; cfg.mode = mode; cfg.error = err
    outs, sizes = (C.POINTER(C.c_uint8) * n)(), (C.c_size_t * n)()
    assert lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.byref(cfg), outs, sizes) == n
    for i in range(n):
        s = C.string_at(outs[i], sizes[i])
        d = H.decode(lib, s)
        e = float(np.abs(d - frames[i].ravel()).max())
        print(mode, h, w, i, len(s), "coeffs", int.from_bytes(s[16:24], "little"), "max err", e, flush=True)
PY
EBCC_INIT_BASE_ERROR_QUANTILE=1e-2 EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK=1 timeout 120 python /tmp/san_case.py > ${P}_plain.log 2>&1; tail -3 ${P}_plain.log
for tool in memcheck racecheck; do
  EBCC_INIT_BASE_ERROR_QUANTILE=1e-2 EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK=1 timeout 900 compute-sanitizer --tool $tool --print-limit 20 python /tmp/san_case.py > ${P}_$tool.log 2>&1
  echo "$tool rc=$?"; grep -c "coeffs" ${P}_$tool.log; grep "ERROR SUMMARY\|RACECHECK SUMMARY\|Error:\|Race reported" ${P}_$tool.log | sort | uniq -c | head -12
done
echo "== one frame, MAX_ERROR 0.1 / 0.5"
timeout 300 python tools/time_modes.py max 0.1 1 3 2>&1 | grep "rep 2"
timeout 300 python tools/time_modes.py max 0.5 1 3 2>&1 | grep "rep 2"
export EBCC_BATCH_FRAMES=24 EBCC_PIPELINES=16
echo "== 384 frames, MAX_ERROR 0.5"
timeout 300 python tools/time_modes.py max 0.5 384 3 2>&1 | grep "rep 2"
echo done
