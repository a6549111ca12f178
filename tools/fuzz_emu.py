"""Fuzzing the decoders' parsers on the CPU: byte-flip, truncation and splice mutations of valid EBCC streams (header, zstd
frame, SPIHT prefix, JPEG 2000 main header / tile-parts / packet headers / MQ segments) through ebcc_decode of the cuemu
build compiled with AddressSanitizer + UBSan (tools/cuemu/build_san.sh).  compute-sanitizer is closed on the GPU pool, so
this is the memory-safety evidence behind the malformed-input row.  Run:
    LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)" ASAN_OPTIONS=detect_leaks=0 \
        python tools/fuzz_emu.py <cases> [seed]
A sanitizer report aborts the process (non-zero exit); the script prints how many mutants decoded, failed cleanly, and
decoded to something else than the original."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
H.PROD_SO = os.environ.get("EBCC_FUZZ_SO", os.path.join(ROOT, "build", "emu_san", "libh5z_ebcc_emu_san.so"))


def seeds(ref):
    os.environ["EBCC_INIT_BASE_ERROR_QUANTILE"] = "1e-2"            # reference encoder: force a residual layer on small frames
    os.environ["EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK"] = "1"
    out = []
    f = H.synth_stack(1, 64, 96)[0]
    out.append(("none", H.encode(ref, f, (1, 64, 96), 8.0, H.NONE)))
    out.append(("max+residual", H.encode(ref, f, (1, 64, 96), 60.0, H.MAX_ERROR, 0.05)))
    g = H.synth_stack(1, 3 * 40, 64)[0]
    out.append(("three tiles", H.encode(ref, g, (3, 40, 64), 6.0, H.NONE)))
    out.append(("const", H.encode(ref, np.full((32, 32), 4.5, np.float32), (1, 32, 32), 5.0, H.NONE)))
    return out


def mutate(rng, s):
    b = bytearray(s)
    kind = int(rng.integers(0, 6))
    if kind == 0:                                   # flips anywhere
        for pos in rng.integers(0, len(b), int(rng.integers(1, 5))):
            b[pos] ^= 1 << int(rng.integers(0, 8))
    elif kind == 1:                                 # flips in the first 200 bytes (EBCC header, zstd header, SIZ / COD / QCD)
        for pos in rng.integers(0, min(200, len(b)), int(rng.integers(1, 4))):
            b[pos] = int(rng.integers(0, 256))
    elif kind == 2:                                 # truncation
        b = b[:int(rng.integers(1, len(b)))]
    elif kind == 3:                                 # a field of the 48-byte header replaced by an extreme value
        off = int(rng.choice([8, 12, 16, 24, 28, 32, 40]))
        val = int(rng.choice([0, 1, 0xff, 0xffff, 0x7fffffff, 0xffffffff, 2 ** 62 + 1, len(b), len(b) + 1]))
        n = 8 if off in (16, 32, 40) else 4
        b[off:off + n] = (val % (1 << (8 * n))).to_bytes(n, "little")
    elif kind == 4:                                 # splice: a chunk copied elsewhere
        n = int(rng.integers(1, 64)); a = int(rng.integers(0, max(1, len(b) - n))); d = int(rng.integers(0, max(1, len(b) - n)))
        b[d:d + n] = b[a:a + n]
    else:                                           # marker bytes sprinkled in the codestream
        for pos in rng.integers(48, len(b), int(rng.integers(1, 4))):
            b[pos] = 0xff
            if pos + 1 < len(b): b[pos + 1] = int(rng.choice([0x90, 0x93, 0xd9, 0x4f, 0x51, 0x52, 0x5c, 0xff]))
    return bytes(b)


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 20241105
    lib = H.prod_lib(); ref = H.ref_lib()
    H.prod_lib().log_set_level_from_env
    os.environ["EBCC_LOG_LEVEL"] = "6"
    base = seeds(ref)
    want = {name: H.decode(lib, s) for name, s in base}
    assert all(w is not None for w in want.values()), "the unmutated streams must decode"
    rng = np.random.default_rng(seed)
    ok = failed = changed = 0
    t0 = time.time()
    for i in range(cases):
        name, s = base[i % len(base)]
        d = H.decode(lib, mutate(rng, s))
        if d is None: failed += 1
        elif d.shape == want[name].shape and np.array_equal(d.view(np.uint32), want[name].view(np.uint32)): ok += 1
        else: changed += 1
    again = {name: H.decode(lib, s) for name, s in base}
    assert all(a is not None and np.array_equal(a.view(np.uint32), want[n].view(np.uint32)) for n, a in again.items()), "library unusable after the fuzz run"
    print("fuzz: %d mutants (seed %d) in %.0f s: %d rejected cleanly, %d decoded to other data, %d unaffected; no sanitizer report; "
          "the originals still decode bit-identically" % (cases, seed, time.time() - t0, failed, changed, ok))


if __name__ == "__main__":
    main()
