"""GPU: scheduling knobs must not change a single byte -- code-blocks per warp of the MQ kernel, batches of one call
spread over several device pipelines, and concurrent callers (each owns one pipeline of the pool)."""
import ctypes as C
import os
import threading

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu


def _encode_batch(lib, frames, h, w, cr, mode=H.NONE, err=0.0):
    n = len(frames)
    lib.ebcc_encode_batch.restype = C.c_size_t
    cfg = H.CodecConfig()
    cfg.dims[:] = [1, h, w]
    cfg.base_cr = cr
    cfg.mode = mode
    cfg.error = err
    outs = (C.POINTER(C.c_uint8) * n)()
    sizes = (C.c_size_t * n)()
    fr = np.ascontiguousarray(frames, dtype=np.float32)
    got = lib.ebcc_encode_batch(fr.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.byref(cfg), outs, sizes)
    assert got == n
    res = [C.string_at(outs[t], sizes[t]) for t in range(n)]
    for t in range(n):
        lib.free_buffer(outs[t])
    return res


@pytest.fixture
def env_guard():
    keys = ("EBCC_MQ_LANES", "EBCC_DEC_LANES", "EBCC_BATCH_FRAMES", "EBCC_PIPELINES", "EBCC_FUSED_DWT", "EBCC_FUSED_IDWT",
            "EBCC_SP_DECODE_SERIAL", "EBCC_NO_PCRD_CACHE", "EBCC_CUTS_FROM_PASS0", "EBCC_DEBUG_ALL_CUTS_UNPROVEN",
            "EBCC_INIT_BASE_ERROR_QUANTILE", "EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK")
    saved = {k: os.environ.get(k) for k in keys}
    yield
    for k, v in saved.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v


def test_mq_lanes_are_byte_identical(built, env_guard):
    lib = H.prod_lib()
    frames = H.synth_stack(5, 200, 300)
    want = [H.orc_j2k_encode(H.scale_u16(f), 6.0) for f in frames]
    for lanes in (1, 2, 4, 8, 16, 32):
        os.environ["EBCC_MQ_LANES"] = str(lanes)
        got = _encode_batch(lib, frames, 200, 300, 6.0)
        for t in range(5):
            assert got[t][48:] == want[t], "lanes %d frame %d" % (lanes, t)


def test_decoder_lanes_are_byte_identical(built, env_guard):
    """The tier-1 decoder runs 1, 2, 4 or 8 code-blocks per warp; the CR search and the decode do not depend on it."""
    lib = H.prod_lib()
    lib.ebcc_decode_batch.restype = C.c_size_t
    frames = H.synth_stack(5, 200, 300)
    ref_streams, ref_dec = None, None
    for lanes in (1, 2, 4, 8):
        os.environ["EBCC_DEC_LANES"] = str(lanes)
        got = _encode_batch(lib, frames, 200, 300, 20.0, H.MAX_ERROR, 0.25)
        n = len(got)
        bufs = [C.create_string_buffer(s, len(s)) for s in got]
        ptrs = (C.c_void_p * n)(*[C.addressof(b) for b in bufs])
        sizes = (C.c_size_t * n)(*[len(s) for s in got])
        douts, cnts = (C.POINTER(C.c_float) * n)(), (C.c_size_t * n)()
        assert lib.ebcc_decode_batch(ptrs, sizes, C.c_size_t(n), douts, cnts) == n
        dec = [np.ctypeslib.as_array(douts[t], shape=(cnts[t],)).copy() for t in range(n)]
        for t in range(n):
            lib.free_buffer(douts[t])
            assert float(np.abs(dec[t] - frames[t].ravel()).max()) <= 0.25
        if ref_streams is None:
            ref_streams, ref_dec = got, dec
        else:
            assert got == ref_streams, "lanes %d" % lanes
            for t in range(n):
                assert np.array_equal(dec[t], ref_dec[t]), "lanes %d frame %d" % (lanes, t)


def test_multi_batch_call_uses_pipelines(built, env_guard):
    lib = H.prod_lib()
    frames = H.synth_stack(13, 160, 224)
    os.environ["EBCC_BATCH_FRAMES"] = "32"
    os.environ["EBCC_PIPELINES"] = "1"
    one = _encode_batch(lib, frames, 160, 224, 12.0, H.MAX_ERROR, 0.3)
    os.environ["EBCC_BATCH_FRAMES"] = "4"          # 4 batches (4 + 4 + 4 + 1 frames) on 3 pipelines
    os.environ["EBCC_PIPELINES"] = "3"
    many = _encode_batch(lib, frames, 160, 224, 12.0, H.MAX_ERROR, 0.3)
    assert one == many
    # one decode call over all 13 streams: 4 batches on 3 pipelines, results equal the single-stream decodes
    lib.ebcc_decode_batch.restype = C.c_size_t
    n = len(many)
    bufs = [C.create_string_buffer(s, len(s)) for s in many]
    ptrs = (C.c_void_p * n)(*[C.addressof(b) for b in bufs])
    sizes = (C.c_size_t * n)(*[len(s) for s in many])
    douts, cnts = (C.POINTER(C.c_float) * n)(), (C.c_size_t * n)()
    assert lib.ebcc_decode_batch(ptrs, sizes, C.c_size_t(n), douts, cnts) == n
    os.environ["EBCC_PIPELINES"] = "1"
    for t in range(n):
        got = np.ctypeslib.as_array(douts[t], shape=(cnts[t],)).copy()
        lib.free_buffer(douts[t])
        assert np.array_equal(got, H.decode(lib, many[t]))
        assert float(np.abs(got - frames[t].ravel()).max()) <= 0.3 * 1.01 + 1e-6


def test_concurrent_callers(built, env_guard):
    lib = H.prod_lib()
    os.environ["EBCC_PIPELINES"] = "3"
    stacks = [H.synth_stack(6, 128, 192, level=k) for k in range(4)]
    want = [_encode_batch(lib, s, 128, 192, 10.0) for s in stacks]
    got = [None] * 4

    def work(k):
        got[k] = _encode_batch(lib, stacks[k], 128, 192, 10.0)
    th = [threading.Thread(target=work, args=(k,)) for k in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert got == want
    # decode from several threads as well
    lib.ebcc_decode_batch.restype = C.c_size_t
    dec = [None] * 4

    def dwork(k):
        dec[k] = [H.decode(lib, s) for s in want[k]]
    th = [threading.Thread(target=dwork, args=(k,)) for k in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for k in range(4):
        for t in range(6):
            assert dec[k][t] is not None
            assert np.array_equal(dec[k][t], H.decode(lib, want[k][t]))


def test_cross_check_paths_are_byte_identical(built, env_guard):
    """The kernels the fused / warp-cooperative ones replaced stay in the library as cross-check paths (EBCC_FUSED_DWT=0,
    EBCC_FUSED_IDWT=0, EBCC_SP_DECODE_SERIAL=1, EBCC_NO_PCRD_CACHE=1: every probe bisects its threshold from scratch,
    EBCC_CUTS_FROM_PASS0=1: the pool of unproven cuts decoded from pass 0, EBCC_DEBUG_ALL_CUTS_UNPROVEN=1: every cut declared
    unproven, so the pool overflows and the probes fall back to real tier-1 decodes): forward transform, probe back half,
    decode back half, SPIHT list decoder, rate allocation and the handling of unproven cuts must give the same bytes and
    the same floats either way.  The residual layer is forced on these small frames."""
    lib = H.prod_lib()
    lib.ebcc_decode_batch.restype = C.c_size_t
    os.environ["EBCC_INIT_BASE_ERROR_QUANTILE"] = "1e-2"
    os.environ["EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK"] = "1"
    frames = H.synth_stack(4, 181, 263)

    def roundtrip():
        streams = _encode_batch(lib, frames, 181, 263, 40.0, H.MAX_ERROR, 0.05)
        n = len(streams)
        keep = [C.create_string_buffer(s, len(s)) for s in streams]
        ptrs = (C.c_void_p * n)(*[C.addressof(k) for k in keep])
        sizes = (C.c_size_t * n)(*[len(s) for s in streams])
        out = np.zeros((n, 181 * 263), dtype=np.float32)
        douts = (C.c_void_p * n)(*[out[i].ctypes.data for i in range(n)])
        cnts = (C.c_size_t * n)()
        assert lib.ebcc_decode_batch(ptrs, sizes, C.c_size_t(n), douts, cnts) == n
        return streams, out

    want_s, want_d = roundtrip()
    assert any(int.from_bytes(s[16:24], "little") > 0 for s in want_s), "no frame carries a residual layer"
    for key, val in (("EBCC_FUSED_DWT", "0"), ("EBCC_FUSED_IDWT", "0"), ("EBCC_SP_DECODE_SERIAL", "1"), ("EBCC_NO_PCRD_CACHE", "1"),
                     ("EBCC_CUTS_FROM_PASS0", "1"), ("EBCC_DEBUG_ALL_CUTS_UNPROVEN", "1")):
        os.environ[key] = val
        got_s, got_d = roundtrip()
        os.environ.pop(key)
        assert got_s == want_s, key
        assert np.array_equal(got_d.view(np.uint32), want_d.view(np.uint32)), key
