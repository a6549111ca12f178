"""Development aid: the pass recording must give the same probe statistics as real decodes of every probe."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
if os.environ.get("EBCC_EMU", "1") == "1":
    H.PROD_SO = os.path.join(ROOT, "build", "emu", "libh5z_ebcc_emu.so")
import stage_api as S
h, w, b = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
frames = H.synth_stack(b, h, w)
crs = [30.0, 60.0, 120.0, 90.0, 75.0, 15.0, 7.5, 20.0, 45.0, 3.0, 200.0]
a = S.probe_sequence(frames, crs, 0.2, True)
r = S.probe_sequence(frames, crs, 0.2, False)
bad = 0
for cr, x, y in zip(crs, a, r):
    same = x == y
    bad += not same
    print(cr, "same" if same else "DIFF", x[0], y[0] if not same else "")
print("OK" if not bad else "MISMATCH")
