"""GPU, BASELINE.json's full-size configurations through size-independent properties (the reference needs ~6 s of CPU
per error-bounded 721x1440 frame, so it is only sampled): many ERA5-shaped frames, spread over several device batches
and pipelines by ebcc_encode_batch, must round-trip with ZERO bound violations (mean adjustment off, as the bound is
enforced before it), every stream must equal the single-frame call, and sampled frames must take the reference's
decisions (same J2K bytes, size within 1 %).
  configs[2]: RELATIVE_ERROR 1e-3 on pressure levels x hours (different value ranges per level)
  configs[3]: MAX_ERROR 0.1 K tight bound on an hourly stack (stresses the sparsification search)"""
import ctypes as C
import os
import struct

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu
HH, WW = 721, 1440


def _batch_round_trip(frames, mode, err, base_cr=30.0):
    lib = H.prod_lib()
    lib.ebcc_encode_batch.restype = C.c_size_t
    lib.ebcc_decode_batch.restype = C.c_size_t
    n = frames.shape[0]
    cfg = H.CodecConfig()
    cfg.dims[:] = [1, HH, WW]
    cfg.base_cr, cfg.mode, cfg.error = base_cr, mode, err
    outs, sizes = (C.POINTER(C.c_uint8) * n)(), (C.c_size_t * n)()
    assert lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.byref(cfg), outs, sizes) == n
    streams = [C.string_at(outs[i], sizes[i]) for i in range(n)]
    dec = np.empty_like(frames)
    ptrs = (C.c_void_p * n)(*[C.cast(outs[i], C.c_void_p).value for i in range(n)])
    douts, cnts = (C.POINTER(C.c_float) * n)(), (C.c_size_t * n)()
    for i in range(n):
        douts[i] = C.cast(dec[i].ctypes.data, C.POINTER(C.c_float))
    assert lib.ebcc_decode_batch(ptrs, sizes, C.c_size_t(n), douts, cnts) == n
    assert all(cnts[i] == HH * WW for i in range(n))
    for i in range(n):
        lib.free_buffer(outs[i])
    return streams, dec


@pytest.fixture
def strict_bound():
    os.environ["EBCC_DISABLE_MEAN_ADJUSTMENT"] = "1"
    os.environ["EBCC_BATCH_FRAMES"] = "12"          # several device batches per call, spread over the pipelines
    os.environ["EBCC_PIPELINES"] = "4"
    yield
    for k in ("EBCC_DISABLE_MEAN_ADJUSTMENT", "EBCC_BATCH_FRAMES", "EBCC_PIPELINES"):
        os.environ.pop(k, None)


def _j2k(stream):
    zsz = struct.unpack("<Q", stream[32:40])[0]
    return stream[48 + zsz:]


def test_config3_tight_max_error_hourly_stack(built, strict_bound):
    n, err = 40, 0.1
    frames = H.synth_stack(n)
    streams, dec = _batch_round_trip(frames, H.MAX_ERROR, err)
    worst = np.abs(dec - frames).reshape(n, -1).max(axis=1)
    assert worst.max() <= np.float32(err), "bound violated in frame %d: %g" % (int(worst.argmax()), float(worst.max()))
    total = sum(len(s) for s in streams)
    assert frames.nbytes / total > 3.0                                  # still a compressor at this bound
    lib = H.prod_lib()
    for i in (0, 17, n - 1):                                            # batching / pipelines change nothing
        assert streams[i] == H.encode(lib, frames[i], (1, HH, WW), 30.0, H.MAX_ERROR, err)
    if H.have_ref():
        for i in (5,):                                                  # ~6 s of CPU per frame
            want = H.encode(H.ref_lib(), frames[i], (1, HH, WW), 30.0, H.MAX_ERROR, err)
            assert _j2k(streams[i]) == _j2k(want)
            assert abs(len(streams[i]) - len(want)) <= 0.01 * len(want)
            assert np.array_equal(H.decode(H.ref_lib(), streams[i]).view(np.uint32), dec[i].ravel().view(np.uint32))


def test_config2_relative_error_pressure_levels(built, strict_bound):
    levels, hours, rel = 6, 6, 1e-3
    frames = np.concatenate([H.synth_stack(hours, level=3 * k) + np.float32(20.0 * k) for k in range(levels)]).astype(np.float32)
    n = frames.shape[0]
    streams, dec = _batch_round_trip(frames, H.RELATIVE_ERROR, rel)
    for i in range(n):
        rng = np.float32(frames[i].max()) - np.float32(frames[i].min())
        target = np.float32(rel) * rng                                  # get_data_range * error, src/ebcc_codec.c:724-726
        assert np.abs(dec[i] - frames[i]).max() <= target, "frame %d" % i
    lib = H.prod_lib()
    for i in (0, n // 2, n - 1):
        assert streams[i] == H.encode(lib, frames[i], (1, HH, WW), 30.0, H.RELATIVE_ERROR, rel)
    if H.have_ref():
        i = n - 2
        want = H.encode(H.ref_lib(), frames[i], (1, HH, WW), 30.0, H.RELATIVE_ERROR, rel)
        assert _j2k(streams[i]) == _j2k(want) and abs(len(streams[i]) - len(want)) <= 0.01 * len(want)
