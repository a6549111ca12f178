"""Wall-clock of ebcc_encode_batch / ebcc_decode_batch through the public C ABI for one mode (host buffers in, streams out)."""
import ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H

def main():
    mode = {"none": H.NONE, "max": H.MAX_ERROR, "rel": H.RELATIVE_ERROR}[sys.argv[1]]
    err = float(sys.argv[2]); n = int(sys.argv[3]); reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
    os.environ["EBCC_BATCH_FRAMES"] = os.environ.get("EBCC_BATCH_FRAMES", str(n))
    lib = H.prod_lib()
    lib.ebcc_encode_batch.restype = C.c_size_t; lib.ebcc_decode_batch.restype = C.c_size_t
    base = H.synth_stack(min(n, 8))
    frames = np.ascontiguousarray(np.concatenate([base + np.float32(0.37 * k) for k in range((n + 7) // 8)])[:n])
    cfg = H.CodecConfig(); cfg.dims[:] = [1, 721, 1440]; cfg.base_cr = 30.0; cfg.mode = mode; cfg.error = err
    outs, sizes = (C.POINTER(C.c_uint8) * n)(), (C.c_size_t * n)()
    for r in range(reps):
        t0 = time.perf_counter()
        k = lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.byref(cfg), outs, sizes)
        dt = time.perf_counter() - t0
        assert k == n
        tot = sum(sizes)
        print("encode rep %d: %.1f ms (%.2f ms/frame, %.3f GB/s in), CR %.2f" % (r, dt * 1e3, dt * 1e3 / n, frames.nbytes / dt / 1e9, frames.nbytes / tot), flush=True)
        if r + 1 < reps:
            for i in range(n): lib.free_buffer(outs[i])
    ptrs = (C.c_void_p * n)(*[C.cast(outs[i], C.c_void_p).value for i in range(n)])
    douts, counts = (C.POINTER(C.c_float) * n)(), (C.c_size_t * n)()
    for r in range(reps):
        t0 = time.perf_counter()
        k = lib.ebcc_decode_batch(ptrs, sizes, C.c_size_t(n), douts, counts)
        dt = time.perf_counter() - t0
        assert k == n
        worst = max(float(np.abs(np.ctypeslib.as_array(douts[i], shape=(counts[i],)) - frames[i].ravel()).max()) for i in range(n))
        print("decode rep %d: %.1f ms (%.2f ms/frame, %.3f GB/s out), max err %.4f" % (r, dt * 1e3, dt * 1e3 / n, frames.nbytes / dt / 1e9, worst), flush=True)
        for i in range(n): lib.free_buffer(douts[i]); douts[i] = None

if __name__ == "__main__":
    main()
