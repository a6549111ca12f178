"""h5py keyword builder for HDF5 filter 308; same interface as the reference's
ebcc/filter_wrapper.py:16-68 (cd_values = [H, W, f32bits(base_cr), mode, f32bits(error)])."""
import struct
from collections.abc import Mapping
from typing import Optional, Tuple


def float_to_uint32(f: float) -> int:
    return struct.unpack("I", struct.pack("f", f))[0]


class EBCC_Filter(Mapping):
    FILTER_ID = 308
    _MODES = {"none": 0, "max_error_target": 1, "relative_error_target": 2}

    def __init__(self, base_cr: float, height: int, width: int, residual_opt: Optional[Tuple[str, float]] = None,
                 data_dim: int = 2):
        assert height > 0 and width > 0
        self.base_cr = float(base_cr)
        self.height, self.width, self.data_dim = int(height), int(width), int(data_dim)
        self.residual_opt = residual_opt if residual_opt is not None else ("none", 0)
        kind, val = self.residual_opt
        if kind not in self._MODES:
            raise ValueError(f"Unknown residual_type {kind!r}: use 'none', 'max_error_target' or 'relative_error_target'")
        opts = [self.height, self.width, float_to_uint32(self.base_cr), self._MODES[kind]]
        if kind != "none":
            opts.append(float_to_uint32(float(val)))
        self.hdf_filter_opts = tuple(opts)
        self.chunks = (*[1] * (self.data_dim - 2), self.height, self.width)

    @property
    def _kwargs(self):
        return {"dtype": "float32", "chunks": self.chunks, "compression": self.FILTER_ID,
                "compression_opts": self.hdf_filter_opts}

    def __hash__(self):
        return hash((self.FILTER_ID, self.hdf_filter_opts))

    def __len__(self):
        return len(self._kwargs)

    def __iter__(self):
        return iter(self._kwargs)

    def __getitem__(self, item):
        return self._kwargs[item]
