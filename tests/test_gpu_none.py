"""GPU parity, mode NONE: the CUDA path through the C ABI (ebcc_encode / ebcc_encode_batch) must emit
codestreams byte-identical to the oracle restatement and to the reference-derived checker."""
import ctypes as C

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu


def _hdr(stream):
    return stream[:48], stream[48:]


@pytest.mark.parametrize("shape,cr", [((64, 64), 10.0), ((33, 35), 5.0), ((128, 96), 3.0), ((100, 200), 1.5),
                                      ((257, 129), 20.0), ((200, 300), 100.0), ((32, 2047), 8.0), ((2047, 32), 8.0)])
def test_small_frames_vs_oracle(built, shape, cr):
    lib = H.prod_lib()
    f = H.synth_stack(1, shape[0], shape[1])[0]
    s = H.encode(lib, f, (1,) + shape, cr, H.NONE)
    assert s[:4] == b"EBCC"
    hdr, j2k = _hdr(s)
    assert j2k == H.orc_j2k_encode(H.scale_u16(f), cr)
    mn, mx = np.frombuffer(hdr[8:16], dtype=np.float32)
    assert mn == f.min() and mx == f.max()


def test_era5_frame_vs_reference(built):
    lib = H.prod_lib()
    frames = H.synth_stack(3)
    for t in range(3):
        s = H.encode(lib, frames[t], (1, 721, 1440), 30.0, H.NONE)
        want = H.orc_j2k_encode(H.scale_u16(frames[t]), 30.0)
        assert s[48:] == want
        if H.have_ref():
            assert s == H.encode(H.ref_lib(), frames[t], (1, 721, 1440), 30.0, H.NONE)


def test_batch_equals_single(built):
    lib = H.prod_lib()
    lib.ebcc_encode_batch.restype = C.c_size_t
    frames = H.synth_stack(5, 160, 224)
    cfg = H.CodecConfig()
    cfg.dims[:] = [1, 160, 224]
    cfg.base_cr = 12.0
    outs = (C.POINTER(C.c_uint8) * 5)()
    sizes = (C.c_size_t * 5)()
    n = lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(5), C.byref(cfg), outs, sizes)
    assert n == 5
    for t in range(5):
        got = C.string_at(outs[t], sizes[t])
        lib.free_buffer(outs[t])
        assert got == H.encode(lib, frames[t], (1, 160, 224), 12.0, H.NONE)
        assert got[48:] == H.orc_j2k_encode(H.scale_u16(frames[t]), 12.0)


def test_constant_field(built):
    lib = H.prod_lib()
    f = np.full((64, 64), 42.0, dtype=np.float32)
    s = H.encode(lib, f, (1, 64, 64), 10.0, H.NONE)
    assert len(s) == 56 and s[5] == 1
    if H.have_ref():
        assert s == H.encode(H.ref_lib(), f, (1, 64, 64), 10.0, H.NONE)


def test_invalid_dims(built):
    lib = H.prod_lib()
    f = np.zeros((16, 64), dtype=np.float32)
    assert H.encode(lib, f, (1, 16, 64), 10.0, H.NONE) == b""
