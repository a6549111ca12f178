#!/bin/bash
# round-2 GPU sessions 52, 54: library defaults (16 pipelines x 24 frames, CUDA_DEVICE_MAX_CONNECTIONS set by the library): suite, plain calls
mkdir -p gpurun_out
P=gpurun_out/r02def
timeout 600 python -m pytest tests/test_gpu_none.py tests/test_gpu_residual.py tests/test_gpu_pipelines.py tests/test_dropin.py -m gpu -x -q > ${P}_pytest.log 2>&1; echo "pytest rc=$?" >> ${P}_pytest.log; tail -3 ${P}_pytest.log
echo "== library defaults, nothing set"
cat > /tmp/defcall.py <<'PY'
import ctypes as C, os, sys, time
sys.path.insert(0, "tests")
import numpy as np, harness as H
lib = H.prod_lib(); lib.ebcc_encode_batch.restype = C.c_size_t
n = 768
base = H.synth_stack(8)
frames = np.ascontiguousarray(np.concatenate([base + np.float32(0.37 * k) for k in range(n // 8)]))
cfg = H.CodecConfig(); cfg.dims[:] = [1, 721, 1440]; cfg.base_cr = 30.0; cfg.mode = H.MAX_ERROR; cfg.error = 0.1
outs, sizes = (C.POINTER(C.c_uint8) * n)(), (C.c_size_t * n)()
for r in range(3):
    t0 = time.perf_counter()
    assert lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.byref(cfg), outs, sizes) == n
    dt = time.perf_counter() - t0
    print("encode rep %d: %.1f ms, %.2f GB/s, CUDA_DEVICE_MAX_CONNECTIONS=%s" % (r, dt * 1e3, frames.nbytes / dt / 1e9, os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS")), flush=True)
    for i in range(n): lib.free_buffer(outs[i])
PY
for m in unset 1 0; do
  echo "-- EBCC_SPIN_SYNC $m"
  if [ $m = unset ]; then env -u EBCC_BATCH_FRAMES -u EBCC_PIPELINES -u CUDA_DEVICE_MAX_CONNECTIONS -u EBCC_SPIN_SYNC timeout 300 python /tmp/defcall.py | tail -2
  else env -u EBCC_BATCH_FRAMES -u EBCC_PIPELINES -u CUDA_DEVICE_MAX_CONNECTIONS EBCC_SPIN_SYNC=$m timeout 300 python /tmp/defcall.py | tail -1; fi
done
echo done
