"""Call patterns of the reference's own bindings, restated for the tests (test infrastructure only):

* `ZarrDecodeLikeReference` -- the ctypes sequence of the reference's EBCCZarrFilter.encode / .decode
  (/root/reference/ebcc/zarr_filter.py:45-76), including `decode(buf, out=arr)`, which hands the array's own pointer to
  ebcc_decode and returns the array (the in-place contract of src/ebcc_codec.c:1126-1128).
* `h5z_filter` -- what libhdf5 does with a dynamically loaded filter: H5PLget_plugin_info() -> H5Z_class2_t -> call
  `filter` with a libc-malloc'ed chunk buffer; the callee frees it and hands back its own malloc'ed buffer
  (src/h5z_ebcc.c:124-148, struct src/hdf5_stub.h:29-43).
"""
import ctypes as C
import struct

import numpy as np

import harness as H

H5Z_FLAG_REVERSE = 0x0100

_libc = C.CDLL(None)
_libc.malloc.restype = C.c_void_p
_libc.malloc.argtypes = [C.c_size_t]
_libc.free.argtypes = [C.c_void_p]

H5Z_FUNC = C.CFUNCTYPE(C.c_size_t, C.c_uint, C.c_size_t, C.POINTER(C.c_uint), C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_void_p))


class H5ZClass2(C.Structure):
    """H5Z_class2_t (reference src/hdf5_stub.h:34-43)"""
    _fields_ = [("version", C.c_int), ("id", C.c_int), ("encoder_present", C.c_uint), ("decoder_present", C.c_uint),
                ("name", C.c_char_p), ("can_apply", C.c_void_p), ("set_local", C.c_void_p), ("filter", H5Z_FUNC)]


def plugin_class(lib):
    lib.H5PLget_plugin_info.restype = C.POINTER(H5ZClass2)
    lib.H5PLget_plugin_type.restype = C.c_int
    return lib.H5PLget_plugin_info().contents


def cd_values(h, w, base_cr, mode, error=None):
    v = [h, w, struct.unpack("<I", struct.pack("<f", base_cr))[0], mode]
    if mode != H.NONE:
        v.append(struct.unpack("<I", struct.pack("<f", error))[0])
    return v


def h5z_filter(lib, flags, cd, payload):
    """runs the plugin's filter callback on a malloc'ed copy of payload; returns (return value, *buf_size, bytes of the
    buffer the callee handed back interpreted with `nbytes_out(ret)` left to the caller)."""
    cls = plugin_class(lib)
    n = len(payload)
    p = _libc.malloc(n)
    C.memmove(p, payload, n)
    buf = C.c_void_p(p)
    buf_size = C.c_size_t(n)
    cdv = (C.c_uint * len(cd))(*cd)
    ret = cls.filter(flags, len(cd), cdv, n, C.byref(buf_size), C.byref(buf))
    return ret, buf_size.value, buf


def take(buf, nbytes):
    """copies nbytes out of the callee's buffer and frees it with libc free, as libhdf5 would"""
    data = C.string_at(buf, nbytes)
    _libc.free(buf)
    return data


class ZarrDecodeLikeReference:
    def __init__(self, lib, arglist):
        self.lib = lib
        self.arglist = (C.c_uint * len(arglist))(*arglist)
        self.n = len(arglist)
        lib.populate_config.argtypes = [C.POINTER(H.CodecConfig), C.c_size_t, C.POINTER(C.c_uint), C.c_size_t]
        lib.populate_config.restype = None

    def encode(self, buf):
        out = C.POINTER(C.c_ubyte)()
        cfg = H.CodecConfig()
        buf = np.ascontiguousarray(buf, dtype=np.float32).ravel()
        self.lib.populate_config(C.byref(cfg), self.n, self.arglist, buf.nbytes)
        size = self.lib.ebcc_encode(buf.ctypes.data, C.byref(cfg), C.cast(C.byref(out), C.POINTER(C.POINTER(C.c_uint8))))
        data = bytes(C.cast(out, C.POINTER(C.c_ubyte * size)).contents)
        self.lib.free_buffer(out)
        return data

    def decode(self, buf, out=None):
        if out is not None:
            out_buffer = out.ctypes.data_as(C.POINTER(C.c_float))
        else:
            out_buffer = C.POINTER(C.c_float)()
        buf_c = (C.c_ubyte * len(buf)).from_buffer_copy(buf)
        n = self.lib.ebcc_decode(C.byref(buf_c), len(buf), C.byref(out_buffer))
        if out is not None:
            return out
        arr = np.ctypeslib.as_array(out_buffer, shape=(n,)).copy()
        self.lib.free_buffer(out_buffer)
        return arr
