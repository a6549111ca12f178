"""Development aid: SPIHT stage parity (stream bytes, truncation probes) of the cuemu build vs the reference."""
import os, sys, time, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
if os.environ.get("EBCC_EMU", "1") == "1":
    H.PROD_SO = os.path.join(ROOT, "build", "emu", "libh5z_ebcc_emu.so")
import stage_api as S

def main():
    h, w = int(sys.argv[1]), int(sys.argv[2])
    cr = float(sys.argv[3]) if len(sys.argv) > 3 else 30.0
    src = sys.argv[4] if len(sys.argv) > 4 else "synth"
    if src == "synth":
        f = H.synth_stack(1, h, w)[0]
    else:
        f = np.load("/root/reference/data/test_data.npy").astype(np.float32)[:h, :w].copy()
    ok = S.check_residual_stage(f, cr, verbose=True)
    print("OK" if ok else "MISMATCH")

if __name__ == "__main__":
    main()
