"""Generates tests/golden/era5_known_answers.json from the UNMODIFIED reference (oracle/_ref/libh5z_ebcc_ref.so)
on the reference's own bundled frame data/test_data.npy (committed here as era5_t2m_721x1440.npy): BASELINE.json
configs[0] and the three other de-facto known answers SURVEY.md 8(c) recorded for it.  Run in the build container:

    python tests/golden/make_golden_era5.py

For every case: total stream length, the header fields, sha256 of the JPEG 2000 codestream and of the SPIHT prefix
(the zstd-independent parts), and the zstd payload length of the reference (informative: any valid frame is allowed)."""
import ctypes as C
import hashlib
import json
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as H  # noqa: E402

CASES = [("none_30", H.NONE, 0.0, {}), ("max_0.5", H.MAX_ERROR, 0.5, {}), ("max_0.1", H.MAX_ERROR, 0.1, {}),
         ("rel_1e-3", H.RELATIVE_ERROR, 1e-3, {}),
         # environment variables that change the bytes (src/ebcc_codec.c:634-649; data/logs.txt:12-29)
         ("max_0.5_q0", H.MAX_ERROR, 0.5, {"EBCC_INIT_BASE_ERROR_QUANTILE": "0"}),
         ("max_0.5_q1e-5", H.MAX_ERROR, 0.5, {"EBCC_INIT_BASE_ERROR_QUANTILE": "1e-5"}),
         ("max_0.5_q1e-4", H.MAX_ERROR, 0.5, {"EBCC_INIT_BASE_ERROR_QUANTILE": "1e-4"}),
         ("max_0.5_nopure", H.MAX_ERROR, 0.5, {"EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK": "1"}),
         ("max_0.1_noconsist", H.MAX_ERROR, 0.1, {"EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK_CONSISTENCY": "1"}),
         ("max_0.5_nomean", H.MAX_ERROR, 0.5, {"EBCC_DISABLE_MEAN_ADJUSTMENT": "1"})]


def describe(stream, ref):
    magic, ver, flags, _r, mn, mx, coeffs, rmn, rmx, zsz, tail = struct.unpack("<4sBBHIIQIIQQ", stream[:48])
    z, j2k = stream[48:48 + zsz], stream[48 + zsz:]
    d = dict(length=len(stream), flags=flags, min_bits=mn, max_bits=mx, coeffs_size=coeffs, rmin_bits=rmn, rmax_bits=rmx,
             zstd_size_reference=zsz, j2k_size=tail, j2k_sha256=hashlib.sha256(j2k).hexdigest())
    if coeffs:
        ref.ZSTD_decompress.restype = C.c_size_t
        ref.ZSTD_decompress.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.c_size_t]
        back = C.create_string_buffer(coeffs)
        assert ref.ZSTD_decompress(back, coeffs, z, len(z)) == coeffs
        d["spiht_sha256"] = hashlib.sha256(back.raw).hexdigest()
    return d


def main():
    f = np.load(os.path.join(HERE, "era5_t2m_721x1440.npy"))
    ref = H.ref_lib()
    out = {"input_sha256": hashlib.sha256(f.tobytes()).hexdigest(), "shape": list(f.shape), "base_cr": 30.0, "cases": {}}
    for name, mode, err, env in CASES:
        os.environ.update(env)
        try:
            s = H.encode(ref, f, (1,) + f.shape, 30.0, mode, err)
        finally:
            for k in env:
                del os.environ[k]
        out["cases"][name] = dict(mode=mode, error=err, env=env, **describe(s, ref))
        print(name, out["cases"][name]["length"], out["cases"][name]["coeffs_size"])
    with open(os.path.join(HERE, "era5_known_answers.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
