"""Development aid: bit-exactness of the decoder back half (fused IDWT) on the cuemu build against the reference, over
shapes that exercise tile borders, odd sizes and multi-tile images (odd row parity)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
H.PROD_SO = os.path.join(ROOT, "build", "emu", "libh5z_ebcc_emu.so")

def main():
    lib = H.prod_lib(); ref = H.ref_lib()
    cases = [((1, 33, 37), 4.0), ((1, 181, 263), 6.0), ((1, 32, 300), 3.0), ((1, 290, 32), 3.0), ((3, 37, 131), 5.0), ((2, 65, 140), 8.0), ((1, 140, 515), 10.0)]
    if len(sys.argv) > 1:
        cases = [((int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])), float(sys.argv[4]))]
    bad = 0
    for dims, cr in cases:
        t, h, w = dims
        f = H.synth_stack(1, t * h, w)[0]
        t0 = time.time()
        r = H.encode(ref, f, dims, cr, H.NONE)
        e = H.encode(lib, f, dims, cr, H.NONE)
        ok0 = e == r
        d, dr = H.decode(lib, r), H.decode(ref, r)
        ok = d is not None and np.array_equal(d.view(np.uint32), dr.view(np.uint32))
        m = H.encode(lib, f, dims, cr, H.MAX_ERROR, 0.2)
        mr = H.encode(ref, f, dims, cr, H.MAX_ERROR, 0.2)
        ok2 = m[48:] == mr[48:] if len(m) == len(mr) else False
        print(dims, cr, "NONE stream", "equal" if ok0 else "DIFF", "decode", "equal" if ok else "DIFF", "| MAX_ERROR stream", "equal" if ok2 else "DIFF (%d vs %d)" % (len(m), len(mr)), "%.1fs" % (time.time() - t0), flush=True)
        bad += (not ok) + (not ok2) + (not ok0)
    print("FAILED" if bad else "all equal")

if __name__ == "__main__":
    main()
