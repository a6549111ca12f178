"""Development aid: end-to-end NONE compress from T caller threads (24-frame ebcc_encode_batch calls, pinned host input)."""
import ctypes as C, os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
import torch

def main():
    T = int(sys.argv[1]); calls = int(sys.argv[2]); reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    os.environ.setdefault("EBCC_PIPELINES", str(T))
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
    os.environ.setdefault("EBCC_SPIN_SYNC", "1" if T <= (os.cpu_count() or 1) else "0")
    lib = H.prod_lib(); lib.ebcc_encode_batch.restype = C.c_size_t
    base = H.synth_stack(8)
    frames = torch.from_numpy(np.ascontiguousarray(np.concatenate([base] * 3))).pin_memory().numpy()
    cfg = H.CodecConfig(); cfg.dims[:] = [1, 721, 1440]; cfg.base_cr = 30.0; cfg.mode = H.NONE
    def worker(n, gate):
        o, z = (C.POINTER(C.c_uint8) * 24)(), (C.c_size_t * 24)()
        gate.wait()
        for _ in range(n):
            assert lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(24), C.byref(cfg), o, z) == 24
            for i in range(24): lib.free_buffer(o[i])
    def round_(n):
        gate = threading.Barrier(T)
        th = [threading.Thread(target=worker, args=(n, gate)) for _ in range(T)]
        t0 = time.perf_counter()
        for t in th: t.start()
        for t in th: t.join()
        return time.perf_counter() - t0
    round_(2)
    for r in range(reps):
        dt = round_(calls)
        print("threads %d calls %d: %.1f ms  %.2f GB/s" % (T, calls, dt * 1e3, T * calls * frames.nbytes / dt / 1e9), flush=True)

if __name__ == "__main__":
    main()
