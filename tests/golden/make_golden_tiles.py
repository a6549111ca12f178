"""Generates tests/golden/tiles_golden.npz from the UNMODIFIED reference (oracle/_ref/libh5z_ebcc_ref.so):
multi-tile internal images (dims[0] > 1; j2k_encode_internal with n_tiles > 1, src/ebcc_codec.c:105-180).
Run in the build container only:

    python tests/golden/make_golden_tiles.py

Each case stores the reference's EBCC stream (small ones verbatim, all as sha256 + length), the sha256 of
the seeded input, and the sha256 of what the reference decodes from its own stream."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as H  # noqa: E402

MODES = {"none": H.NONE, "max": H.MAX_ERROR, "rel": H.RELATIVE_ERROR}
CASES = [  # (name, tiles, tile height, width, base_cr, mode, error)
    ("t2_32", 2, 32, 32, 4.0, "none", 0.0),          # reference tests/test_c_api.py:174-192 geometry
    ("t4_32", 4, 32, 32, 2.0, "none", 0.0),          # tests/test_c_api.py:211-226 (padded chunk of 4)
    ("t3_odd", 3, 37, 45, 6.0, "none", 0.0),         # odd tile height: odd row parity, shifted code-block grid
    ("t2_100", 2, 100, 64, 8.0, "none", 0.0),        # level-1 bands of tile 1 straddle a multiple of 64
    ("t4_33", 4, 33, 70, 20.0, "none", 0.0),
    ("t5_129", 5, 129, 200, 12.0, "none", 0.0),
    ("t63_32", 63, 32, 40, 10.0, "none", 0.0),       # maximum tile count (63 * 32 = 2016 <= 2047)
    ("t2_lossless", 2, 48, 80, 1.5, "none", 0.0),    # rate <= 1: every pass kept
    ("t2_era5", 2, 721, 1440, 30.0, "none", 0.0),    # two ERA5 frames as one image (1442 lines)
    ("t2_max", 2, 37, 45, 10.0, "max", 0.3),
    ("t3_rel", 3, 40, 64, 20.0, "rel", 0.01),
    ("t4_max", 4, 96, 160, 30.0, "max", 0.25),
]


def make_input(tiles, h, w):
    return H.synth_stack(1, tiles * h, w)[0]


def main():
    out = {}
    lib = H.ref_lib()
    for name, t, h, w, cr, mode, err in CASES:
        f = make_input(t, h, w)
        s = H.encode(lib, f, (t, h, w), cr, MODES[mode], err)
        d = H.decode(lib, s)
        out[name + "_sha"] = np.frombuffer(hashlib.sha256(s).digest(), dtype=np.uint8)
        out[name + "_len"] = np.array([len(s)])
        zsz = int.from_bytes(s[32:40], "little")                       # ebcc_header_t.compressed_size
        out[name + "_j2k_sha"] = np.frombuffer(hashlib.sha256(s[48 + zsz:]).digest(), dtype=np.uint8)
        out[name + "_in_sha"] = np.frombuffer(hashlib.sha256(f.tobytes()).digest(), dtype=np.uint8)
        out[name + "_dec_sha"] = np.frombuffer(hashlib.sha256(d.tobytes()).digest(), dtype=np.uint8)
        out[name + "_maxerr"] = np.array([np.abs(d - f.ravel()).max()], dtype=np.float64)
        if len(s) < 40000:
            out[name + "_stream"] = np.frombuffer(s, dtype=np.uint8)
        print(name, len(s), float(out[name + "_maxerr"][0]))
    np.savez_compressed(os.path.join(HERE, "tiles_golden.npz"), **out)
    print("wrote", len(CASES), "cases")


if __name__ == "__main__":
    main()
