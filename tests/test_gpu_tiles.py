"""GPU parity, multi-tile internal images (dims[0] > 1): the reference stacks dims[0] tiles of
dims[1] x dims[2] into one JPEG 2000 image (j2k_encode_internal with n_tiles > 1, src/ebcc_codec.c:105-180).  Tiles
below the first one start at a non-zero canvas row, which changes the DWT row parity, the code-block anchoring and the
per-tile rate split.  The CUDA path must write the same bytes as the reference and decode its streams bit-exactly;
checked against committed golden vectors (tests/golden/tiles_golden.npz, generated from the unmodified reference by
tests/golden/make_golden_tiles.py) and, where oracle/_ref is built, against the reference itself."""
import ctypes as C
import hashlib
import os
import struct

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tiles_golden.npz")
MODES = {"none": H.NONE, "max": H.MAX_ERROR, "rel": H.RELATIVE_ERROR}

from golden.make_golden_tiles import CASES, make_input  # noqa: E402


def parse(stream):
    magic, ver, flags, _r, mn, mx, coeffs, rmn, rmx, zsz, tail = struct.unpack("<4sBBHIIQIIQQ", stream[:48])
    return dict(mn=mn, mx=mx, coeffs=coeffs, rmn=rmn, rmx=rmx, zsz=zsz, tail=tail, j2k=stream[48 + zsz:])


def sha(b):
    return np.frombuffer(hashlib.sha256(b).digest(), dtype=np.uint8)


@pytest.mark.parametrize("case", [c for c in CASES if c[5] == "none"], ids=lambda c: c[0])
def test_none_streams_equal_golden(built, case):
    name, t, h, w, cr, mode, err = case
    g = np.load(GOLDEN)
    f = make_input(t, h, w)
    assert np.array_equal(sha(f.tobytes()), g[name + "_in_sha"]), "seeded input differs from the one the fixture was made from"
    s = H.encode(H.prod_lib(), f, (t, h, w), cr, H.NONE)
    assert len(s) == int(g[name + "_len"][0])
    assert np.array_equal(sha(s), g[name + "_sha"])          # whole EBCC stream, byte for byte
    if name + "_stream" in g:
        assert s == g[name + "_stream"].tobytes()
    assert s[48:] == H.orc_j2k_encode_tiled(H.scale_u16(f), t, cr)      # and the oracle restatement agrees


@pytest.mark.parametrize("case", CASES, ids=lambda c: c[0])
def test_decode_of_reference_streams(built, case):
    """streams written by the reference decode to the reference's own output, bit for bit"""
    name, t, h, w, cr, mode, err = case
    g = np.load(GOLDEN)
    if name + "_stream" not in g:
        pytest.skip("stream too large to be stored in the fixture (covered by test_cross_decode_against_reference)")
    d = H.decode(H.prod_lib(), g[name + "_stream"].tobytes())
    assert d is not None and d.size == t * h * w
    assert np.array_equal(sha(d.tobytes()), g[name + "_dec_sha"])


@pytest.mark.parametrize("case", [c for c in CASES if c[5] != "none"], ids=lambda c: c[0])
def test_error_bounded_equal_golden(built, case):
    name, t, h, w, cr, mode, err = case
    g = np.load(GOLDEN)
    f = make_input(t, h, w)
    s = H.encode(H.prod_lib(), f, (t, h, w), cr, MODES[mode], err)
    assert s
    a = parse(s)
    if name + "_stream" in g:
        b = parse(g[name + "_stream"].tobytes())
        # same decisions as the reference: base-layer bytes (=> ratio found by the search), truncation length, range
        assert a["j2k"] == b["j2k"] and a["coeffs"] == b["coeffs"] and a["rmn"] == b["rmn"] and a["rmx"] == b["rmx"]
        assert abs(int(a["mn"]) - int(b["mn"])) <= 1 and abs(int(a["mx"]) - int(b["mx"])) <= 1   # mean adjustment: <= 1 ulp
    assert abs(len(s) - int(g[name + "_len"][0])) <= 0.01 * int(g[name + "_len"][0])
    d = H.decode(H.prod_lib(), s)
    target = np.float32(err) if mode == "max" else np.float32(err) * (f.max() - f.min())
    shift = abs(float(np.frombuffer(struct.pack("<I", a["mn"]), dtype=np.float32)[0]) - float(f.min()))
    assert np.abs(d - f.ravel()).max() <= float(target) + shift + 1e-6 * float(np.abs(f).max())
    os.environ["EBCC_DISABLE_MEAN_ADJUSTMENT"] = "1"
    try:
        strict = H.encode(H.prod_lib(), f, (t, h, w), cr, MODES[mode], err)
        assert np.abs(H.decode(H.prod_lib(), strict) - f.ravel()).max() <= target      # zero violations
    finally:
        del os.environ["EBCC_DISABLE_MEAN_ADJUSTMENT"]


@needs_ref
@pytest.mark.parametrize("shape,cr,mode,err", [((2, 32, 32), 4.0, "none", 0), ((3, 45, 33), 7.0, "none", 0), ((2, 721, 1440), 30.0, "none", 0),
                                               ((6, 77, 300), 15.0, "none", 0), ((2, 64, 96), 12.0, "max", 0.2), ((5, 50, 50), 25.0, "rel", 0.005)])
def test_cross_decode_against_reference(built, shape, cr, mode, err):
    lib, ref = H.prod_lib(), H.ref_lib()
    t, h, w = shape
    f = H.synth_stack(1, t * h, w, level=1)[0]
    ours, want = H.encode(lib, f, shape, cr, MODES[mode], err), H.encode(ref, f, shape, cr, MODES[mode], err)
    assert ours and want
    assert parse(ours)["j2k"] == parse(want)["j2k"]
    if mode == "none":
        assert ours == want
    for st in (ours, want):
        a, b = H.decode(lib, st), H.decode(ref, st)
        assert a is not None and np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_batch_of_images_equals_singles(built):
    lib = H.prod_lib()
    lib.ebcc_encode_batch.restype = C.c_size_t
    t, h, w, n = 3, 40, 72, 5
    frames = H.synth_stack(n, t * h, w)
    cfg = H.CodecConfig()
    cfg.dims[:] = [t, h, w]
    cfg.base_cr = 9.0
    outs = (C.POINTER(C.c_uint8) * n)()
    sizes = (C.c_size_t * n)()
    assert lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.byref(cfg), outs, sizes) == n
    for i in range(n):
        got = C.string_at(outs[i], sizes[i])
        lib.free_buffer(outs[i])
        assert got == H.encode(lib, frames[i], (t, h, w), 9.0, H.NONE)


def test_single_tile_after_tiled_and_back(built):
    """a pipeline context is re-used across geometries: tiled -> plain -> tiled must not leak state"""
    lib = H.prod_lib()
    f = H.synth_stack(1, 128, 64)[0]
    a1 = H.encode(lib, f, (2, 64, 64), 6.0, H.NONE)
    b1 = H.encode(lib, f, (1, 128, 64), 6.0, H.NONE)
    c1 = H.encode(lib, f, (4, 32, 64), 6.0, H.NONE)
    assert a1 == H.encode(lib, f, (2, 64, 64), 6.0, H.NONE) and b1 == H.encode(lib, f, (1, 128, 64), 6.0, H.NONE)
    assert c1 == H.encode(lib, f, (4, 32, 64), 6.0, H.NONE)
    assert b1[48:] == H.orc_j2k_encode(H.scale_u16(f), 6.0)
    for s in (a1, b1, c1):
        assert H.decode(lib, s).size == 128 * 64


def test_constant_multi_tile_image(built):
    lib = H.prod_lib()
    f = np.full((3 * 32, 48), 7.25, dtype=np.float32)
    s = H.encode(lib, f, (3, 32, 48), 10.0, H.MAX_ERROR, 0.1)
    assert s and np.array_equal(H.decode(lib, s), f.ravel())
    if H.have_ref():
        assert s == H.encode(H.ref_lib(), f, (3, 32, 48), 10.0, H.MAX_ERROR, 0.1)


# the reference's own chunking tests use chunks with chunk_dims[0] > 1, i.e. multi-tile images
# (/root/reference/tests/test_c_api.py:174-300: same shapes, base_cr 2, MAX_ERROR 0.01)
@pytest.mark.parametrize("shape,chunk", [((2, 32, 32), (1, 32, 32)), ((3, 33, 35), (2, 32, 32)), ((2, 32, 32), (4, 32, 32)),
                                         ((2, 33, 35), (1, 64, 64)), ((2, 32, 32), (0, 0, 0))])
def test_reference_chunking_cases(built, shape, chunk):
    lib = H.prod_lib()
    idx = np.indices(shape, dtype=np.float32)
    data = (idx[0] * 0.5 + np.sin(idx[1] / 7.0) + np.cos(idx[2] / 5.0)).astype(np.float32)
    enc = H.encode(lib, data, shape, 2.0, H.MAX_ERROR, 0.01, fn="ebcc_encode_chunking", chunk_dims=chunk)
    assert enc[:4] == b"EBCK"
    dec = H.decode(lib, enc, fn="ebcc_decode_chunking")
    assert dec is not None and dec.size == data.size
    assert np.abs(dec.reshape(shape) - data).max() <= 0.01 + 1e-5
    if H.have_ref():
        want = H.encode(H.ref_lib(), data, shape, 2.0, H.MAX_ERROR, 0.01, fn="ebcc_encode_chunking", chunk_dims=chunk)
        assert enc[:80] == want[:80]                        # container header
        dref = H.decode(H.ref_lib(), enc, fn="ebcc_decode_chunking")
        assert np.array_equal(dref.view(np.uint32), dec.view(np.uint32))
        assert np.array_equal(H.decode(lib, want, fn="ebcc_decode_chunking").view(np.uint32),
                              H.decode(H.ref_lib(), want, fn="ebcc_decode_chunking").view(np.uint32))


def test_malformed_multi_tile_streams_fail_cleanly(built):
    """the tile-part walk of the parser (Psot chain, Isot match) runs on untrusted bytes: every corruption must end in a
    clean failure (ebcc_decode returns 0), never in a crash or a silently wrong image"""
    lib = H.prod_lib()
    f = H.synth_stack(1, 3 * 32, 48)[0]
    s = H.encode(lib, f, (3, 32, 48), 6.0, H.NONE)
    good = H.decode(lib, s)
    assert good is not None and good.size == 3 * 32 * 48
    j = 48                                   # start of the codestream inside the EBCC stream
    sot0 = s.index(b"\xff\x90", j)           # first SOT: marker(2) Lsot(2) Isot(2) Psot(4) TPsot TNsot
    assert s[sot0 + 4:sot0 + 6] == b"\x00\x00"
    psot0 = int.from_bytes(s[sot0 + 6:sot0 + 10], "big")
    sot1 = sot0 + psot0
    assert s[sot1:sot1 + 2] == b"\xff\x90" and s[sot1 + 4:sot1 + 6] == b"\x00\x01"

    def patched(off, val):
        b = bytearray(s); b[off:off + len(val)] = val; return bytes(b)

    cases = {
        "Psot of tile 0 runs past the end": patched(sot0 + 6, (0x7fffff00).to_bytes(4, "big")),
        "Psot of tile 0 shorter than a tile-part header": patched(sot0 + 6, (5).to_bytes(4, "big")),
        "Psot of tile 0 is zero although tiles follow": patched(sot0 + 6, (0).to_bytes(4, "big")),
        "Psot of tile 0 points into the middle of its packets": patched(sot0 + 6, (psot0 - 3).to_bytes(4, "big")),
        "tile 1 carries a wrong Isot": patched(sot1 + 4, b"\x00\x07"),
        "tile 0 appears twice, tile 1 never": patched(sot1 + 4, b"\x00\x00"),
        "image height is not a multiple of the tile height": patched(j + 12, (3 * 32 + 1).to_bytes(4, "big")),
        "tile height of zero": patched(j + 28, (0).to_bytes(4, "big")),
        "tile offset not at the origin": patched(j + 36, (1).to_bytes(4, "big")),
    }
    for what, bad in cases.items():
        assert H.decode(lib, bad) is None, what
    # the library is still usable afterwards and the good stream still decodes to the same frame
    assert np.array_equal(H.decode(lib, s), good)
