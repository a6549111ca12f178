"""Generates tests/golden/j2k_golden.npz from the UNMODIFIED reference (oracle/_ref/libh5z_ebcc_ref.so,
built by oracle/Makefile from /root/reference/src).  Run in the build container only:

    python tests/golden/make_golden.py

Each case stores the seeded input recipe, and the reference's output: full EBCC stream for small
frames, sha256 + length for large ones."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as H  # noqa: E402

CASES = [  # (name, h, w, base_cr, kind)
    ("s64", 64, 64, 10.0, "synth"), ("odd", 33, 35, 5.0, "synth"), ("wide", 32, 300, 8.0, "synth"),
    ("tall", 300, 32, 8.0, "synth"), ("mid", 128, 96, 3.0, "synth"), ("lossless", 100, 120, 1.5, "synth"),
    ("noise", 96, 96, 4.0, "noise"), ("hi_cr", 200, 300, 400.0, "synth"), ("era5", 721, 1440, 30.0, "synth"),
]


def make_input(kind, h, w):
    if kind == "noise":
        return np.random.default_rng(7).standard_normal((h, w)).astype(np.float32) * 50 + 3
    return H.synth_stack(1, h, w)[0]


def main():
    out = {}
    lib = H.ref_lib()
    for name, h, w, cr, kind in CASES:
        f = make_input(kind, h, w)
        s = H.encode(lib, f, (1, h, w), cr, H.NONE)
        out[name + "_sha"] = np.frombuffer(hashlib.sha256(s).digest(), dtype=np.uint8)
        out[name + "_len"] = np.array([len(s)])
        out[name + "_in_sha"] = np.frombuffer(hashlib.sha256(f.tobytes()).digest(), dtype=np.uint8)
        if len(s) < 20000:
            out[name + "_stream"] = np.frombuffer(s, dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "j2k_golden.npz"), **out)
    print("wrote", len(CASES), "cases")


if __name__ == "__main__":
    main()
