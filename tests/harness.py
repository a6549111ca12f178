"""ctypes harness shared by the tests: loads the reference-derived checker
(oracle/_ref/libh5z_ebcc_ref.so, the UNMODIFIED reference built by oracle/Makefile), the
plain-C oracle restatement (oracle/libebcc_oracle.so) and the product library
(ebcc_b200/libh5z_ebcc.so).  Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libh5z_ebcc_ref.so")
ORC_SO = os.path.join(ROOT, "oracle", "libebcc_oracle.so")
PROD_SO = os.environ.get("EBCC_PROD_SO", os.path.join(ROOT, "ebcc_b200", "libh5z_ebcc.so"))

NONE, MAX_ERROR, RELATIVE_ERROR = 0, 1, 2


class CodecConfig(C.Structure):
    """Mirror of codec_config_t (reference src/ebcc_codec.h:32-39)."""
    _fields_ = [("dims", C.c_size_t * 3), ("base_cr", C.c_float), ("mode", C.c_int),
                ("residual_cr", C.c_float), ("error", C.c_float), ("chunk_dims", C.c_size_t * 3)]


class CodecDataBuffer(C.Structure):
    """Mirror of codec_data_buffer_t (reference src/ebcc_codec.c:38-43)."""
    _fields_ = [("buffer", C.c_void_p), ("size", C.c_size_t), ("length", C.c_size_t), ("offset", C.c_size_t)]


def _bind_codec_api(lib):
    u8pp = C.POINTER(C.POINTER(C.c_uint8))
    f32pp = C.POINTER(C.POINTER(C.c_float))
    for name in ("ebcc_encode", "ebcc_encode_chunking", "ebcc_encode_chunking_compat"):
        fn = getattr(lib, name)
        fn.restype = C.c_size_t
        fn.argtypes = [C.c_void_p, C.POINTER(CodecConfig), u8pp]
    for name in ("ebcc_decode", "ebcc_decode_chunking"):
        fn = getattr(lib, name)
        fn.restype = C.c_size_t
        fn.argtypes = [C.c_void_p, C.c_size_t, f32pp]
    lib.free_buffer.restype = None
    lib.free_buffer.argtypes = [C.c_void_p]
    return lib


_cache = {}


def have_ref():
    return os.path.exists(REF_SO)


def ref_lib():
    if "ref" not in _cache:
        lib = _bind_codec_api(C.CDLL(REF_SO))
        lib.j2k_encode_internal.restype = None
        lib.j2k_encode_internal.argtypes = [C.c_void_p, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), C.c_float,
                                            C.POINTER(CodecDataBuffer)]
        lib.j2k_decode_internal.restype = None
        lib.j2k_decode_internal.argtypes = [C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_size_t),
                                            C.POINTER(C.c_size_t), C.c_float, C.c_float, C.POINTER(CodecDataBuffer)]
        lib.codec_data_buffer_init.argtypes = [C.POINTER(CodecDataBuffer)]
        lib.codec_data_buffer_destroy.argtypes = [C.POINTER(CodecDataBuffer)]
        _cache["ref"] = lib
    return _cache["ref"]


def build_oracle():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True)


def orc_lib():
    if "orc" not in _cache:
        if not os.path.exists(ORC_SO):
            build_oracle()
        lib = C.CDLL(ORC_SO)
        lib.orc_j2k_encode.restype = C.c_size_t
        lib.orc_j2k_encode.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.POINTER(C.POINTER(C.c_uint8))]
        lib.orc_j2k_encode_tiled.restype = C.c_size_t
        lib.orc_j2k_encode_tiled.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.POINTER(C.POINTER(C.c_uint8))]
        lib.orc_scale_u16.argtypes = [C.c_void_p, C.c_size_t, C.c_float, C.c_float, C.c_void_p]
        lib.orc_minmax.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        _cache["orc"] = lib
    return _cache["orc"]


def prod_lib():
    if "prod" not in _cache:
        _cache["prod"] = _bind_codec_api(C.CDLL(PROD_SO))
    return _cache["prod"]


_libc = C.CDLL(None)
_libc.free.argtypes = [C.c_void_p]


def ref_j2k_encode(u16, base_cr):
    """reference j2k_encode_internal (src/ebcc_codec.c:105-180) on an (H, W) uint16 image."""
    lib = ref_lib()
    u16 = np.ascontiguousarray(u16, dtype=np.uint16)
    h, w = u16.shape
    buf = CodecDataBuffer()
    lib.codec_data_buffer_init(C.byref(buf))
    img = (C.c_size_t * 2)(h, w)
    tile = (C.c_size_t * 2)(h, w)
    lib.j2k_encode_internal(u16.ctypes.data, img, tile, C.c_float(base_cr), C.byref(buf))
    out = C.string_at(buf.buffer, buf.length)
    lib.codec_data_buffer_destroy(C.byref(buf))
    return out


def orc_j2k_encode(u16, base_cr):
    lib = orc_lib()
    u16 = np.ascontiguousarray(u16, dtype=np.uint16)
    h, w = u16.shape
    out = C.POINTER(C.c_uint8)()
    n = lib.orc_j2k_encode(u16.ctypes.data, h, w, C.c_float(base_cr), C.byref(out))
    res = C.string_at(out, n)
    _libc.free(out)
    return res


def orc_j2k_encode_tiled(u16, ntiles, base_cr):
    """oracle restatement of a multi-tile image: u16 is (ntiles * h, w), tiles of h lines stacked vertically"""
    lib = orc_lib()
    u16 = np.ascontiguousarray(u16, dtype=np.uint16)
    hh, w = u16.shape
    assert hh % ntiles == 0
    out = C.POINTER(C.c_uint8)()
    n = lib.orc_j2k_encode_tiled(u16.ctypes.data, ntiles, hh // ntiles, w, C.c_float(base_cr), C.byref(out))
    res = C.string_at(out, n)
    _libc.free(out)
    return res


def scale_u16(frame):
    """numpy restatement of src/ebcc_codec.c:686-689 (f32 arithmetic, C truncation)."""
    f = np.ascontiguousarray(frame, dtype=np.float32)
    mn, mx = f.min(), f.max()
    v = ((f - mn) / np.float32(mx - mn)) * np.float32(65535.0)
    return v.astype(np.uint16)


def encode(lib, data, dims, base_cr, mode=NONE, error=0.0, fn="ebcc_encode", chunk_dims=(0, 0, 0)):
    data = np.ascontiguousarray(data, dtype=np.float32)
    cfg = CodecConfig()
    cfg.dims[:] = list(dims)
    cfg.base_cr = base_cr
    cfg.mode = mode
    cfg.error = error
    cfg.chunk_dims[:] = list(chunk_dims)
    out = C.POINTER(C.c_uint8)()
    n = getattr(lib, fn)(data.ctypes.data, C.byref(cfg), C.byref(out))
    if n == 0:
        return b""
    res = C.string_at(out, n)
    lib.free_buffer(out)
    return res


def decode(lib, stream, fn="ebcc_decode"):
    out = C.POINTER(C.c_float)()
    buf = (C.c_uint8 * len(stream)).from_buffer_copy(stream)
    n = getattr(lib, fn)(buf, len(stream), C.byref(out))
    if n == 0:
        return None
    res = np.ctypeslib.as_array(out, shape=(n,)).copy()
    lib.free_buffer(out)
    return res


def synth_frame(t, h=721, w=1440, level=0):
    """Seeded smooth ERA5-shaped field: k^-3 spectral noise on a
    latitudinal temperature profile.  Frames are generated independently (the AR(1) mixing of
    the survey text is applied by synth_stack)."""
    rng = np.random.default_rng(20241104 + t)
    z = rng.standard_normal((h, w))
    Z = np.fft.rfft2(z)
    ky = np.fft.fftfreq(h)[:, None] * h
    kx = np.fft.rfftfreq(w)[None, :] * w
    Z *= (1.0 + np.sqrt(kx * kx + ky * ky)) ** -1.5
    Z[0, 0] = 0
    f = np.fft.irfft2(Z, s=(h, w))
    f /= f.std()
    return f


def synth_stack(n, h=721, w=1440, level=0):
    lat = np.linspace(-np.pi / 2, np.pi / 2, h)[:, None]
    out = np.empty((n, h, w), dtype=np.float32)
    prev = None
    for t in range(n):
        f = synth_frame(t, h, w)
        if prev is not None:
            f = 0.9 * prev + np.sqrt(1 - 0.81) * f
        prev = f
        out[t] = (250.0 + 35.0 * np.cos(lat) + 8.0 * (1 + level / 12.0) * f).astype(np.float32)
    return out
