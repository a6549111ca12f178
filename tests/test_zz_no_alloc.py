"""GPU: the allocation-free entry point ebcc_cuda_frame_stats_noalloc (include/ebcc_cuda.h; no counterpart in the
reference).  Its scratch is a static __device__ array shared by every call,
so the test runs it twice in a row on different batches and compares each result, bit for bit, with what the existing
entry point ebcc_cuda_frame_stats reports after the encoder's analysis pass on the same frames.  Collected last on
purpose (pytest -x at the end of the round)."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from stage_api import cuda_api

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

NOALLOC_MAX_FRAMES = 4096


def _api():
    lib = cuda_api()
    lib.ebcc_cuda_frame_stats.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ebcc_cuda_frame_stats_noalloc.restype = C.c_int
    lib.ebcc_cuda_frame_stats_noalloc.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def _existing(lib, frames):
    """min / max / flags through the existing C ABI: context, upload, analysis pass, ebcc_cuda_frame_stats"""
    n, h, w = frames.shape
    ctx = C.c_void_p(lib.ebcc_cuda_ctx_create(0))
    assert ctx.value, "no CUDA device: the EBCC B200 codec has no CPU path"
    try:
        assert lib.ebcc_cuda_reserve(ctx, h, w, n) == 0
        assert lib.ebcc_cuda_upload(ctx, frames.ctypes.data, n) == 0
        assert lib.ebcc_cuda_analyze_resident(ctx, n) == 0
        mn, mx, fl = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
        assert lib.ebcc_cuda_frame_stats(ctx, n, mn.ctypes.data, mx.ctypes.data, fl.ctypes.data) == 0
    finally:
        lib.ebcc_cuda_ctx_destroy(ctx)
    return mn, mx, fl


def _batches():
    a = H.synth_stack(5, 96, 160)
    a[1] = -3.25                                 # constant frame
    a[2, 10, 7] = np.float32("nan")              # NaN -> flag 2
    a[3] = np.abs(a[3] - a[3].mean())            # minimum near zero
    a[3, 40, 9] = -0.0; a[3, 50, 11] = 0.0       # first occurrence decides the sign of a zero minimum
    b = H.synth_stack(9, 96, 160)[::-1].copy() * np.float32(1.7) - np.float32(40.0)
    b[0, 95, 159] = np.float32("inf")
    b[8] = 0.0
    return a, b


def test_noalloc_frame_stats_equals_existing_entry_point_twice_in_a_row(built):
    lib = _api()
    torch.cuda.set_device(0)
    batches = _batches()
    want = [_existing(lib, f) for f in batches]
    # every buffer the entry point touches is allocated by the caller, beforehand
    devs = [torch.from_numpy(f).cuda() for f in batches]
    outs = [(torch.full((len(f),), -7.0, device="cuda"), torch.full((len(f),), -7.0, device="cuda"),
             torch.full((len(f),), -7, dtype=torch.int32, device="cuda")) for f in batches]
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    free_before = None
    for rnd in range(2):                         # two rounds: the second call of each batch finds the other batch's leftovers
        for k, f in enumerate(batches):
            n, h, w = f.shape
            mn, mx, fl = outs[k]
            mn.fill_(-7.0); mx.fill_(-7.0); fl.fill_(-7)
            torch.cuda.synchronize()
            if rnd == 1:
                free_before = torch.cuda.mem_get_info()[0]
            rc = lib.ebcc_cuda_frame_stats_noalloc(devs[k].data_ptr(), n, h * w, mn.data_ptr(), mx.data_ptr(), fl.data_ptr(),
                                                   C.c_void_p(stream.cuda_stream))
            assert rc == 0
            stream.synchronize()
            if rnd == 1:
                assert torch.cuda.mem_get_info()[0] == free_before, "the allocation-free entry point changed the free device memory"
            wmn, wmx, wfl = want[k]
            assert np.array_equal(fl.cpu().numpy(), wfl), (rnd, k, fl.cpu().numpy(), wfl)
            ok = (wfl & 2) == 0                  # frames with NaN / Inf have no defined extrema in either path
            assert np.array_equal(mn.cpu().numpy().view(np.uint32)[ok], wmn.view(np.uint32)[ok])
            assert np.array_equal(mx.cpu().numpy().view(np.uint32)[ok], wmx.view(np.uint32)[ok])
            # and against the definition (findMinMaxf: first occurrence of the extremum, so -0.0 before +0.0 stays -0.0)
            for i in np.nonzero(ok)[0]:
                assert mn.cpu().numpy()[i] == f[i].min() and mx.cpu().numpy()[i] == f[i].max()
                assert bool(wfl[i] & 1) == (f[i].min() == f[i].max())
    assert np.signbit(outs[0][0].cpu().numpy()[3])          # the -0.0 came first


def test_noalloc_rejects_oversized_and_null_arguments(built):
    lib = _api()
    torch.cuda.set_device(0)
    d = torch.zeros(16, device="cuda")
    o = torch.zeros(4, device="cuda")
    fl = torch.zeros(4, dtype=torch.int32, device="cuda")
    assert lib.ebcc_cuda_frame_stats_noalloc(d.data_ptr(), NOALLOC_MAX_FRAMES + 1, 4, o.data_ptr(), o.data_ptr(), fl.data_ptr(), None) < 0
    assert lib.ebcc_cuda_frame_stats_noalloc(d.data_ptr(), 4, 1 << 31, o.data_ptr(), o.data_ptr(), fl.data_ptr(), None) < 0
    assert lib.ebcc_cuda_frame_stats_noalloc(None, 4, 4, o.data_ptr(), o.data_ptr(), fl.data_ptr(), None) < 0
    assert lib.ebcc_cuda_frame_stats_noalloc(d.data_ptr(), 0, 4, o.data_ptr(), o.data_ptr(), fl.data_ptr(), None) < 0
    # still usable afterwards, on the default stream
    mx = torch.zeros(4, device="cuda")
    d = torch.arange(16, dtype=torch.float32, device="cuda")
    assert lib.ebcc_cuda_frame_stats_noalloc(d.data_ptr(), 4, 4, o.data_ptr(), mx.data_ptr(), fl.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert o.cpu().tolist() == [0.0, 4.0, 8.0, 12.0] and mx.cpu().tolist() == [3.0, 7.0, 11.0, 15.0] and fl.cpu().tolist() == [0, 0, 0, 0]
