"""numcodecs/zarr codec over the C ABI; same interface as the reference's ebcc/zarr_filter.py:19-85
(codec id 'ebcc_filter', arglist = HDF5 cd_values).  numcodecs is optional: without it the class is a
plain object with the same encode/decode methods."""
import ctypes

import numpy as np

from . import EBCC_FILTER_PATH

try:  # pragma: no cover - numcodecs is not installed in the build container
    import numcodecs
    from numcodecs.abc import Codec
except Exception:  # noqa: BLE001
    numcodecs = None
    Codec = object


class CodecConfigT(ctypes.Structure):
    _fields_ = [("dims", ctypes.c_size_t * 3), ("base_cr", ctypes.c_float), ("residual_compression_type", ctypes.c_int),
                ("residual_cr", ctypes.c_float), ("error", ctypes.c_float), ("chunk_dims", ctypes.c_size_t * 3)]


class EBCCZarrFilter(Codec):
    codec_id = "ebcc_filter"

    def __init__(self, arglist):
        if EBCC_FILTER_PATH is None:
            raise RuntimeError("libh5z_ebcc.so not built: run __graft_entry__.build()")
        self.arglist = np.array(arglist, dtype=np.uint32)
        lib = ctypes.CDLL(EBCC_FILTER_PATH)
        lib.populate_config.argtypes = [ctypes.POINTER(CodecConfigT), ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t]
        lib.ebcc_encode.restype = ctypes.c_size_t
        lib.ebcc_encode.argtypes = [ctypes.c_void_p, ctypes.POINTER(CodecConfigT), ctypes.POINTER(ctypes.POINTER(ctypes.c_ubyte))]
        lib.ebcc_decode.restype = ctypes.c_size_t
        lib.ebcc_decode.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.POINTER(ctypes.c_float))]
        lib.free_buffer.argtypes = [ctypes.c_void_p]
        lib.ebcc_decode_batch.restype = ctypes.c_size_t
        lib.ebcc_decode_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
        self.lib = lib

    def encode(self, buf):
        cfg, out = CodecConfigT(), ctypes.POINTER(ctypes.c_ubyte)()
        cai = getattr(buf, "__cuda_array_interface__", None)
        if cai is not None and cai["typestr"].endswith("f4") and cai.get("strides") is None:
            # chunk already in GPU memory (torch / cupy): the C ABI takes the device pointer, no host round trip
            addr, nbytes = int(cai["data"][0]), 4 * int(np.prod(cai["shape"]))
            if hasattr(buf, "is_cuda"):
                import torch
                torch.cuda.current_stream(buf.device).synchronize()
        else:
            buf = np.ascontiguousarray(buf, dtype=np.float32).ravel()
            addr, nbytes = buf.ctypes.data, buf.nbytes
        self.lib.populate_config(ctypes.byref(cfg), len(self.arglist), self.arglist.ctypes.data, nbytes)
        n = self.lib.ebcc_encode(ctypes.c_void_p(addr), ctypes.byref(cfg), ctypes.byref(out))
        if n == 0:
            raise RuntimeError("ebcc_encode failed")
        data = ctypes.string_at(out, n)
        self.lib.free_buffer(out)
        return data

    def decode(self, buf, out=None):
        buf = bytes(buf)
        src = (ctypes.c_ubyte * len(buf)).from_buffer_copy(buf)
        if out is not None:
            # in place: host numpy array or a device array (torch / cupy, __cuda_array_interface__)
            cai = getattr(out, "__cuda_array_interface__", None)
            addr = int(cai["data"][0]) if cai is not None else out.ctypes.data
            srcs = (ctypes.c_void_p * 1)(ctypes.addressof(src))
            lens, cnts = (ctypes.c_size_t * 1)(len(buf)), (ctypes.c_size_t * 1)()
            outs = (ctypes.c_void_p * 1)(addr)
            if self.lib.ebcc_decode_batch(srcs, lens, ctypes.c_size_t(1), outs, cnts) != 1:
                raise RuntimeError("ebcc_decode failed")
            return out
        ptr = ctypes.POINTER(ctypes.c_float)()
        n = self.lib.ebcc_decode(src, len(buf), ctypes.byref(ptr))
        if n == 0:
            raise RuntimeError("ebcc_decode failed")
        arr = np.ctypeslib.as_array(ptr, shape=(n,)).copy()
        self.lib.free_buffer(ptr)
        return arr

    def get_config(self):
        return {"id": self.codec_id, "arglist": self.arglist.astype(int).tolist()}

    @classmethod
    def from_config(cls, config):
        return cls(config["arglist"])


if numcodecs is not None:  # pragma: no cover
    numcodecs.register_codec(EBCCZarrFilter)
