"""Development aid: run parity checks against the cuemu build (build/emu) in a container without a GPU."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
H.PROD_SO = os.path.join(ROOT, "build", "emu", "libh5z_ebcc_emu.so")

def main():
    lib = H.prod_lib(); ref = H.ref_lib()
    what = sys.argv[1] if len(sys.argv) > 1 else "none"
    h, w = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (96, 160)
    f = H.synth_stack(1, h, w)[0]
    if what == "none":
        for cr in (10.0, 40.0):
            t0 = time.time(); s = H.encode(lib, f, (1, h, w), cr, H.NONE); t1 = time.time()
            r = H.encode(ref, f, (1, h, w), cr, H.NONE)
            print("encode cr", cr, len(s), len(r), "equal" if s == r else "DIFF", "%.2fs" % (t1 - t0))
            t0 = time.time(); d = H.decode(lib, r); t1 = time.time()
            dr = H.decode(ref, r)
            print("decode", None if d is None else (np.array_equal(d, dr), float(np.abs(d - dr).max())), "%.2fs" % (t1 - t0))
    else:
        mode = H.MAX_ERROR if what == "max" else H.RELATIVE_ERROR
        err = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
        cr = float(sys.argv[5]) if len(sys.argv) > 5 else 30.0
        t0 = time.time(); s = H.encode(lib, f, (1, h, w), cr, mode, err); t1 = time.time()
        r = H.encode(ref, f, (1, h, w), cr, mode, err)
        print("encode", len(s), len(r), "equal" if s == r else "DIFF", "%.2fs" % (t1 - t0))
        print(" ours hdr", s[:48].hex()); print(" ref  hdr", r[:48].hex())
        for name, st in (("ours", s), ("ref", r)):
            for dn, dl in (("ours", lib), ("ref", ref)):
                d = H.decode(dl, st)
                print("  stream", name, "decoded by", dn, None if d is None else float(np.abs(d - f.ravel()).max()))

if __name__ == "__main__":
    main()
