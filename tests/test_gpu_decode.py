"""GPU parity, decoder: ebcc_decode of the CUDA path must reproduce the reference decoder bit for bit
(integer tier-1 / tier-2 work and, by matching OpenJPEG's float operation order, the inverse DWT too), for
streams written by the reference and by the CUDA encoder (cross-decoding both ways)."""
import ctypes as C
import os

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("shape,cr", [((64, 64), 2.0), ((33, 35), 5.0), ((128, 96), 3.0), ((100, 200), 1.5), ((257, 129), 20.0),
                                      ((200, 300), 100.0), ((32, 2047), 8.0), ((2047, 32), 8.0), ((721, 1440), 30.0),
                                      ((721, 1440), 300.0)])
def test_decode_reference_streams_bit_exact(built, shape, cr):
    lib, ref = H.prod_lib(), H.ref_lib()
    f = H.synth_stack(1, shape[0], shape[1])[0]
    s = H.encode(ref, f, (1,) + shape, cr, H.NONE)
    got, want = H.decode(lib, s), H.decode(ref, s)
    assert got is not None and got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@needs_ref
def test_cross_decode_none(built):
    lib, ref = H.prod_lib(), H.ref_lib()
    f = H.synth_stack(2, 300, 420)[1]
    ours = H.encode(lib, f, (1, 300, 420), 25.0, H.NONE)
    a, b = H.decode(ref, ours), H.decode(lib, ours)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_decode_batch_and_caller_buffer(built):
    lib = H.prod_lib()
    lib.ebcc_decode_batch.restype = C.c_size_t
    frames = H.synth_stack(4, 96, 160)
    streams = [H.encode(lib, frames[t], (1, 96, 160), 8.0 + t, H.NONE) for t in range(4)]
    singles = [H.decode(lib, s) for s in streams]
    bufs = [(C.c_uint8 * len(s)).from_buffer_copy(s) for s in streams]
    ptrs = (C.c_void_p * 4)(*[C.addressof(b) for b in bufs])
    lens = (C.c_size_t * 4)(*[len(s) for s in streams])
    outs = (C.POINTER(C.c_float) * 4)()
    mine = np.zeros(96 * 160, dtype=np.float32)                 # frame 2 is decoded into a caller-owned buffer
    outs[2] = mine.ctypes.data_as(C.POINTER(C.c_float))
    counts = (C.c_size_t * 4)()
    assert lib.ebcc_decode_batch(ptrs, lens, C.c_size_t(4), outs, counts) == 4
    for t in range(4):
        got = mine if t == 2 else np.ctypeslib.as_array(outs[t], shape=(counts[t],)).copy()
        assert counts[t] == 96 * 160 and np.array_equal(got, singles[t])
        if t != 2:
            lib.free_buffer(outs[t])


def test_constant_field_round_trip(built):
    lib = H.prod_lib()
    f = np.full((64, 64), 42.0, dtype=np.float32)
    for mode, err in ((H.NONE, 0.0), (H.MAX_ERROR, 0.1)):
        d = H.decode(lib, H.encode(lib, f, (1, 64, 64), 10.0, mode, err))
        assert d is not None and np.all(d == np.float32(42.0))


def test_malformed_streams_fail_cleanly(built):
    lib = H.prod_lib()
    f = H.synth_stack(1, 64, 96)[0]
    s = H.encode(lib, f, (1, 64, 96), 10.0, H.NONE)
    assert H.decode(lib, s[:40]) is None                              # shorter than the header
    assert H.decode(lib, s[:-7]) is None                              # payload size mismatch
    bad = bytearray(s); bad[48 + 60] ^= 0xFF                          # corrupt COD/QCD in the main header
    assert H.decode(lib, bytes(bad)) is None


def test_chunked_container_round_trip(built):
    lib = H.prod_lib()
    k, i, j = np.meshgrid(np.arange(3), np.arange(33), np.arange(35), indexing="ij")
    data = (100.0 * k + 1.5 * i + 0.25 * j).astype(np.float32)
    s = H.encode(lib, data, (3, 33, 35), 2.0, H.MAX_ERROR, 0.01, fn="ebcc_encode_chunking", chunk_dims=(1, 32, 32))
    assert s[:4] == b"EBCK"
    d = H.decode(lib, s, fn="ebcc_decode_chunking")
    assert d is not None and d.size == data.size
    assert np.allclose(d.reshape(data.shape), data, atol=0.02)       # tolerance of the reference's tests/test_c_api.py:168-171


def _to_legacy(stream):
    """Re-serialise a current stream in the header-less layout of older EBCC releases (reference
    ebcc_decode_legacy, src/ebcc_codec.c:1147-1213)."""
    import struct
    magic, ver, flags, _r, mn, mx, coeffs, rmn, rmx, zsize, tsize = struct.unpack("<4sBBHIIQIIQQ", stream[:48])
    assert magic == b"EBCC"
    z, tail = stream[48:48 + zsize], stream[48 + zsize:48 + zsize + tsize]
    return struct.pack("<IIQIIQ", mn, mx, coeffs, rmn, rmx, zsize) + z + tail


@pytest.mark.parametrize("mode,err", [(H.NONE, 0.0), (H.MAX_ERROR, 0.2)])
def test_legacy_streams_decode(built, mode, err):
    lib = H.prod_lib()
    f = H.synth_stack(1, 150, 210)[0]
    s = H.encode(lib, f, (1, 150, 210), 15.0, mode, err)
    want = H.decode(lib, s)
    leg = _to_legacy(s)
    got = H.decode(lib, leg)
    assert got is not None and np.array_equal(got, want)
    if H.have_ref():
        assert np.array_equal(H.decode(H.ref_lib(), leg), got)
    # constant field in the legacy layout: min == max, then the element count
    c = np.full((40, 50), 3.25, dtype=np.float32)
    sc = H.encode(lib, c, (1, 40, 50), 10.0)
    gc = H.decode(lib, _to_legacy(sc))
    assert gc is not None and gc.size == 2000 and np.all(gc == np.float32(3.25))
    # truncated legacy header fails cleanly
    assert H.decode(lib, leg[:20]) is None
