"""Development aid: the SPIHT list decoder (k_sp_decode) on the cuemu build against the reference's spiht_decode, through
ebcc_decode of reference-written streams that carry a residual layer."""
import os, struct, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
H.PROD_SO = os.path.join(ROOT, "build", "emu", "libh5z_ebcc_emu.so")

def main():
    os.environ["EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK"] = "1"; os.environ["EBCC_INIT_BASE_ERROR_QUANTILE"] = sys.argv[1] if len(sys.argv) > 1 else "1e-3"
    lib = H.prod_lib(); ref = H.ref_lib()
    era = np.load(os.path.join(ROOT, "tests", "golden", "era5_t2m_crop_256x384.npy")).astype(np.float32)
    cases = [(era, 30.0, 0.5), (era[:200, :300].copy(), 60.0, 0.3), (era[3:131, 5:261].copy(), 100.0, 0.2), (H.synth_stack(1, 160, 224)[0], 200.0, 0.08)]
    bad = 0
    for f, cr, err in cases:
        h, w = f.shape
        s = H.encode(ref, f, (1, h, w), cr, H.MAX_ERROR, err)
        coeffs = struct.unpack("<Q", s[16:24])[0]
        t0 = time.time()
        d, dr = H.decode(lib, s), H.decode(ref, s)
        ok = d is not None and np.array_equal(d.view(np.uint32), dr.view(np.uint32))
        print((h, w), "cr", cr, "err", err, "coeffs_size", coeffs, "decode", "equal" if ok else "DIFF", "%.1fs" % (time.time() - t0), flush=True)
        bad += not ok
    print("FAILED" if bad else "all equal")

if __name__ == "__main__":
    main()
