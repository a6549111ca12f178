"""Development aid: two resident batches on two contexts (two streams), run back to back vs concurrently."""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402


def main():
    nframes = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    nctx = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    lib = C.CDLL(H.PROD_SO)
    lib.ebcc_cuda_ctx_create.restype = C.c_void_p
    frames = H.synth_stack(min(nframes, 8))
    frames = np.ascontiguousarray(np.concatenate([frames] * ((nframes + 7) // 8))[:nframes])
    ctxs = []
    for _ in range(nctx):
        ctx = C.c_void_p(lib.ebcc_cuda_ctx_create(0))
        assert lib.ebcc_cuda_reserve(ctx, C.c_size_t(721), C.c_size_t(1440), C.c_size_t(nframes)) == 0
        assert lib.ebcc_cuda_upload(ctx, frames.ctypes.data_as(C.c_void_p), C.c_size_t(nframes)) == 0
        ctxs.append(ctx)

    def run(ctx, out):
        total = C.c_float(); st = (C.c_float * 8)()
        rc = lib.ebcc_cuda_bench_resident(ctx, C.c_size_t(nframes), C.c_float(30.0), steps, 2, C.c_size_t(0), C.byref(total), st)
        out.append((rc, total.value))

    # sequential
    t0 = time.perf_counter()
    for ctx in ctxs:
        o = []; run(ctx, o)
    seq = time.perf_counter() - t0
    # concurrent
    outs = [[] for _ in ctxs]
    th = [threading.Thread(target=run, args=(c, o)) for c, o in zip(ctxs, outs)]
    t0 = time.perf_counter()
    for t in th: t.start()
    for t in th: t.join()
    con = time.perf_counter() - t0
    nst = (steps + 2) * nctx
    print("ctxs %d frames %d: sequential %.2f ms/step, concurrent %.2f ms/step (wall, incl. warm-up steps)" % (nctx, nframes, seq * 1e3 / nst, con * 1e3 / nst))


if __name__ == "__main__":
    main()
