"""Print per-stage device times of the resident encode pass (CUDA events inside the library)."""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

STAGES = ["h2d", "minmax", "dwt", "t1_cm", "t1_mq", "pcrd_t2", "bodies", "d2h"]


def main():
    nframes = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    lib = C.CDLL(H.PROD_SO)
    lib.ebcc_cuda_ctx_create.restype = C.c_void_p
    ctx = C.c_void_p(lib.ebcc_cuda_ctx_create(0))
    assert ctx.value, "no CUDA device"
    t0 = time.time()
    frames = H.synth_stack(min(nframes, 8))
    frames = np.ascontiguousarray(np.concatenate([frames] * ((nframes + 7) // 8))[:nframes])
    print("synth %.1fs" % (time.time() - t0))
    assert lib.ebcc_cuda_reserve(ctx, C.c_size_t(721), C.c_size_t(1440), C.c_size_t(nframes)) == 0
    assert lib.ebcc_cuda_upload(ctx, frames.ctypes.data_as(C.c_void_p), C.c_size_t(nframes)) == 0
    total = C.c_float()
    st = (C.c_float * 8)()
    rc = lib.ebcc_cuda_bench_resident(ctx, C.c_size_t(nframes), C.c_float(30.0), steps, 2, C.c_size_t(256 << 20), C.byref(total), st)
    assert rc == 0
    ms = total.value / steps
    print("frames %d  ms/pass %.3f  GB/s(in) %.2f" % (nframes, ms, nframes * 721 * 1440 * 4 / ms / 1e6))
    for i, n in enumerate(STAGES):
        print("  %-8s %.3f ms" % (n, st[i] / steps))


if __name__ == "__main__":
    main()
