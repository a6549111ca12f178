#!/bin/bash
# where does CUDA_DEVICE_MAX_CONNECTIONS have to be set?  (library setenv / shell export / os.environ before the first CUDA call)
cat > /tmp/defcall.py <<'PY'
import ctypes as C, os, sys, time
if len(sys.argv) > 1 and sys.argv[1] == "py": os.environ["CUDA_DEVICE_MAX_CONNECTIONS"] = "32"
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
    if r == 2: print("encode rep %d: %.1f ms, %.2f GB/s" % (r, dt * 1e3, frames.nbytes / dt / 1e9), flush=True)
    for i in range(n): lib.free_buffer(outs[i])
PY
E="env -u EBCC_BATCH_FRAMES -u EBCC_PIPELINES -u CUDA_DEVICE_MAX_CONNECTIONS -u EBCC_SPIN_SYNC"
echo -n "library setenv only:   "; $E timeout 200 python /tmp/defcall.py


echo done
