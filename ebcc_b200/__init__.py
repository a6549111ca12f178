"""ebcc_b200 -- B200-native (CUDA sm_100a) implementation of EBCC's per-frame compress/decompress hot
path behind the reference's API surface: HDF5 filter 308, the ebcc_codec.h C API and the zarr codec.

The package mirrors the reference's ``ebcc`` Python package (ebcc/__init__.py:5-29): it locates the
filter library and exports ``EBCC_FILTER_PATH`` / ``EBCC_FILTER_DIR``.  The library itself
(``libh5z_ebcc.so`` in this directory) is built in-tree by ``__graft_entry__.build()``.
"""
import os

_here = os.path.dirname(os.path.abspath(__file__))
EBCC_FILTER_DIR = os.environ.get("EBCC_FILTER_DIR", _here)
EBCC_FILTER_PATH = os.path.join(EBCC_FILTER_DIR, "libh5z_ebcc.so")
if not os.path.exists(EBCC_FILTER_PATH):
    EBCC_FILTER_PATH = None

from .filter_wrapper import EBCC_Filter  # noqa: E402,F401
