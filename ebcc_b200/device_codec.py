"""Device-resident batch codec: frames that already live in GPU memory (a torch / cupy array, anything
with ``__cuda_array_interface__``) are compressed through the same C ABI as host arrays -- ``ebcc_encode_batch`` /
``ebcc_decode_batch`` of ``include/ebcc_codec.h`` accept device pointers -- so no host round trip of the frame data
happens.  Host numpy arrays work too.  The semantics are the reference's ``ebcc_encode`` / ``ebcc_decode``
(src/ebcc_codec.c:607-918, :1215-1320) applied to every frame of the batch; the streams are ordinary EBCC streams.

    codec = EBCCDeviceCodec(base_cr=30.0, mode="max_error_target", error=0.5)
    streams = codec.encode(t2m)            # t2m: torch.float32 [n, 721, 1440] on cuda:0 -> list of bytes
    frames = codec.decode(streams, out=torch.empty_like(t2m))
"""
import ctypes as C

import numpy as np

from . import EBCC_FILTER_PATH
from .zarr_filter import CodecConfigT

_MODES = {"none": 0, "max_error_target": 1, "relative_error_target": 2}


def _pointer_of(arr):
    """(address, is_device, keep_alive) of a C-contiguous float32 array on the host or on a CUDA device"""
    cai = getattr(arr, "__cuda_array_interface__", None)
    if cai is not None:
        if cai["typestr"] not in ("<f4", "=f4", "f4") or cai.get("strides") is not None:
            raise ValueError("device frames must be C-contiguous float32")
        return int(cai["data"][0]), True, arr
    host = np.ascontiguousarray(arr, dtype=np.float32)
    return host.ctypes.data, False, host


def _sync_producer(arr):
    # the codec runs on its own streams: whatever produced a device array must have finished
    try:
        import torch
        if isinstance(arr, torch.Tensor) and arr.is_cuda:
            torch.cuda.current_stream(arr.device).synchronize()
    except ImportError:  # pragma: no cover
        pass


class EBCCDeviceCodec:
    def __init__(self, base_cr, mode="none", error=0.0):
        if EBCC_FILTER_PATH is None:
            raise RuntimeError("libh5z_ebcc.so not built: run __graft_entry__.build()")
        if mode not in _MODES:
            raise ValueError(f"Unknown mode {mode!r}: use 'none', 'max_error_target' or 'relative_error_target'")
        self.base_cr, self.mode, self.error = float(base_cr), mode, float(error)
        lib = C.CDLL(EBCC_FILTER_PATH)
        lib.ebcc_encode_batch.restype = C.c_size_t
        lib.ebcc_encode_batch.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(CodecConfigT), C.POINTER(C.POINTER(C.c_uint8)),
                                          C.POINTER(C.c_size_t)]
        lib.ebcc_decode_batch.restype = C.c_size_t
        lib.ebcc_decode_batch.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_size_t,
                                          C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_size_t)]
        lib.free_buffer.argtypes = [C.c_void_p]
        self.lib = lib

    def encode(self, frames):
        """frames: [n, H, W] or [n, T, H, W] float32 (T tiles per image, T * H <= 2047), host or device -> list of n EBCC streams"""
        shape = tuple(frames.shape)
        if len(shape) == 3:
            shape = (shape[0], 1) + shape[1:]
        if len(shape) != 4:
            raise ValueError("frames must be [n, H, W] or [n, T, H, W]")
        n = shape[0]
        _sync_producer(frames)
        ptr, _, keep = _pointer_of(frames)
        cfg = CodecConfigT()
        cfg.dims[:] = list(shape[1:])
        cfg.base_cr, cfg.residual_compression_type, cfg.error = self.base_cr, _MODES[self.mode], self.error
        outs, sizes = (C.POINTER(C.c_uint8) * n)(), (C.c_size_t * n)()
        done = self.lib.ebcc_encode_batch(ptr, n, C.byref(cfg), outs, sizes)
        streams = [C.string_at(outs[i], sizes[i]) if sizes[i] else None for i in range(n)]
        for i in range(n):
            if outs[i]:
                self.lib.free_buffer(outs[i])
        del keep
        if done != n:
            raise RuntimeError(f"ebcc_encode_batch encoded {done} of {n} frames")
        return streams

    def decode(self, streams, out):
        """streams: list of EBCC streams of equal frame size; out: [n, ...] float32 array (host numpy or device tensor)
        that receives the frames in place.  Returns out."""
        n = len(streams)
        if out.shape[0] != n:
            raise ValueError("out must have one leading entry per stream")
        _sync_producer(out)
        base, is_dev, keep = _pointer_of(out)
        if not is_dev and keep is not out:
            raise ValueError("host out must be a C-contiguous float32 numpy array")
        per = int(np.prod(out.shape[1:]))
        bufs = [(C.c_uint8 * len(s)).from_buffer_copy(s) for s in streams]
        ptrs = (C.c_void_p * n)(*[C.addressof(b) for b in bufs])
        lens = (C.c_size_t * n)(*[len(s) for s in streams])
        outs, counts = (C.POINTER(C.c_float) * n)(), (C.c_size_t * n)()
        for i in range(n):
            outs[i] = C.cast(C.c_void_p(base + 4 * per * i), C.POINTER(C.c_float))
        done = self.lib.ebcc_decode_batch(ptrs, lens, n, outs, counts)
        if done != n or any(counts[i] != per for i in range(n)):
            raise RuntimeError(f"ebcc_decode_batch decoded {done} of {n} frames")
        return out
