"""Development aid: per-code-block tier-1 output of one synthetic 721x1440 frame (bit-planes, passes, MQ bytes)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402


def main():
    lib = C.CDLL(H.PROD_SO)
    lib.ebcc_cuda_ctx_create.restype = C.c_void_p
    ctx = C.c_void_p(lib.ebcc_cuda_ctx_create(0))
    frames = H.synth_stack(2)
    assert lib.ebcc_cuda_reserve(ctx, C.c_size_t(721), C.c_size_t(1440), C.c_size_t(2)) == 0
    assert lib.ebcc_cuda_upload(ctx, frames.ctypes.data_as(C.c_void_p), C.c_size_t(2)) == 0
    assert lib.ebcc_cuda_analyze_resident(ctx, C.c_size_t(2)) == 0
    rows = []
    cb = 0
    while True:
        nb, tp, npz = C.c_int(), C.c_int(), C.c_int()
        rates = (C.c_int * 128)(); dists = (C.c_double * 128)()
        if lib.ebcc_cuda_debug_cblk(ctx, C.c_size_t(0), C.c_size_t(cb), C.byref(nb), C.byref(tp), rates, dists, C.byref(npz)) != 0:
            break
        rows.append((cb, nb.value, tp.value, rates[tp.value - 1] if tp.value > 0 else 0))
        cb += 1
    a = np.array(rows)
    print("blocks", len(a), "total bytes", a[:, 3].sum(), "max", a[:, 3].max(), "mean", a[:, 3].mean())
    order = np.argsort(-a[:, 3])
    for i in order[:12]:
        print("  cblk %4d numbps %2d passes %3d bytes %6d" % tuple(a[i]))
    print("bytes percentiles", np.percentile(a[:, 3], [10, 25, 50, 75, 90, 99]))
    print("passes percentiles", np.percentile(a[:, 2], [10, 25, 50, 75, 90, 99]))


if __name__ == "__main__":
    main()
