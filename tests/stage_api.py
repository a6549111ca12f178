"""Stage-level drivers over the C ABI of include/ebcc_cuda.h, shared by the GPU parity tests and the
development scripts: each function runs one stage on the product library and the same stage on the
reference-derived checker (oracle/_ref) and compares.  Test infrastructure only."""
import ctypes as C

import numpy as np

import harness as H


class ProbeResult(C.Structure):
    _fields_ = [("n_over", C.c_ulonglong), ("max_err", C.c_float), ("rmin", C.c_float), ("rmax", C.c_float),
                ("sum_err", C.c_double), ("stream_bytes", C.c_size_t)]


def cuda_api():
    lib = C.CDLL(H.PROD_SO)
    lib.ebcc_cuda_ctx_create.restype = C.c_void_p
    lib.ebcc_cuda_ctx_create.argtypes = [C.c_int]
    lib.ebcc_cuda_ctx_destroy.argtypes = [C.c_void_p]
    lib.ebcc_cuda_reserve.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t]
    lib.ebcc_cuda_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    lib.ebcc_cuda_analyze_resident.argtypes = [C.c_void_p, C.c_size_t]
    lib.ebcc_cuda_rate_probe.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.ebcc_cuda_residual_encode.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ebcc_cuda_residual_probe.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ebcc_cuda_residual_pack.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ebcc_cuda_debug_spiht_stream.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.ebcc_cuda_download.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def ref_spiht_api():
    ref = H.ref_lib()
    ref.spiht_encode.restype = None
    ref.spiht_encode.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_size_t),
                                 C.c_size_t, C.c_size_t]
    ref.spiht_decode.restype = None
    ref.spiht_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t]
    ref.ZSTD_decompress.restype = C.c_size_t
    ref.ZSTD_decompress.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    ref.ZSTD_isError.restype = C.c_uint
    ref.ZSTD_isError.argtypes = [C.c_size_t]
    return ref


def check_residual_stage(frame, cr, verbose=False, probes=(1.0, 0.5, 0.23, 0.07)):
    """SPIHT stream bytes, truncation-probe errors and the zstd frame of one frame at ratio cr vs the reference."""
    f = np.ascontiguousarray(frame, dtype=np.float32)
    h, w = f.shape
    lib, ref = cuda_api(), ref_spiht_api()
    ctx = C.c_void_p(lib.ebcc_cuda_ctx_create(0))
    assert ctx.value, "no CUDA device: the EBCC B200 codec has no CPU path"
    ok = True
    try:
        assert lib.ebcc_cuda_reserve(ctx, h, w, 1) == 0
        assert lib.ebcc_cuda_upload(ctx, f.ctypes.data, 1) == 0
        assert lib.ebcc_cuda_analyze_resident(ctx, 1) == 0
        crs, tg, pr = (C.c_float * 1)(cr), (C.c_float * 1)(0.0), (ProbeResult * 1)()
        assert lib.ebcc_cuda_rate_probe(ctx, 1, crs, tg, 1, pr) == 0
        # reference: same base layer (bit-exact by the base-layer tests), residual, SPIHT
        s = H.encode(H.ref_lib(), f, (1, h, w), cr, H.NONE)
        dec = H.decode(H.ref_lib(), s).reshape(h, w)
        res = f - dec
        rmin, rmax = res.min(), res.max()
        ok &= (np.float32(pr[0].rmin) == rmin) and (np.float32(pr[0].rmax) == rmax) and pr[0].stream_bytes == len(s) - 48
        if verbose:
            print("base layer", pr[0].stream_bytes, len(s) - 48, pr[0].rmin, rmin, pr[0].rmax, rmax)
        rn = ((res - rmin) / np.float32(rmax - rmin)).astype(np.float32)
        tb = (len(s) - 48) * 8
        buf, n = C.POINTER(C.c_uint8)(), C.c_size_t()
        ref.spiht_encode(rn.ctypes.data, h, w, C.byref(buf), C.byref(n), tb, 3)
        want = C.string_at(buf, n.value)
        act, tbs, csz = (C.c_int * 1)(1), (C.c_size_t * 1)(tb), (C.c_size_t * 1)(0)
        assert lib.ebcc_cuda_residual_encode(ctx, 1, act, tbs, csz) == 0
        got = (C.c_uint8 * (csz[0] + 16))()
        ln = C.c_size_t()
        assert lib.ebcc_cuda_debug_spiht_stream(ctx, 0, got, csz[0] + 16, C.byref(ln)) == 0
        got = bytes(got[:ln.value])
        same = got == want
        ok &= same
        if verbose:
            print("spiht bytes", len(got), len(want), "equal" if same else "DIFF")
            if not same:
                k = next((i for i in range(min(len(got), len(want))) if got[i] != want[i]), None)
                print(" first diff at byte", k, got[max(0, (k or 0) - 4):(k or 0) + 8].hex(), want[max(0, (k or 0) - 4):(k or 0) + 8].hex())
        # truncation probes
        for frac in probes:
            nb = max(136, int(len(want) * frac) // 1 * 8) if frac < 1.0 else len(want) * 8
            nb = (nb // 8) * 8
            out = np.zeros((h, w), dtype=np.float32)
            ref.spiht_decode(want, nb // 8, out.ctypes.data, h, w, nb)
            r = out * np.float32(rmax - rmin) + rmin
            e = f - (dec + r)
            want_max, want_sum = np.abs(e).max(), float(e.astype(np.float64).sum())
            nbs, me, se = (C.c_size_t * 1)(nb), (C.c_float * 1)(), (C.c_double * 1)()
            assert lib.ebcc_cuda_residual_probe(ctx, 1, act, nbs, me, se) == 0
            good = np.float32(me[0]) == want_max and abs(se[0] - want_sum) <= 1e-6 * max(1.0, abs(want_sum))
            ok &= bool(good)
            if verbose:
                print(" probe bits", nb, "max_err", me[0], want_max, "sum", se[0], want_sum, "ok" if good else "DIFF")
        # zstd frame of a prefix must decode (with the reference's zstd) to that prefix
        for cut in (len(want), max(17, len(want) // 3)):
            cs, zf, zn = (C.c_size_t * 1)(cut), (C.c_void_p * 1)(), (C.c_size_t * 1)()
            assert lib.ebcc_cuda_residual_pack(ctx, 1, cs, zf, zn) == 0
            z = C.string_at(zf[0], zn[0])
            back = C.create_string_buffer(cut + 64)
            r = ref.ZSTD_decompress(back, cut, z, len(z))
            good = (not ref.ZSTD_isError(r)) and r == cut and back.raw[:cut] == want[:cut]
            ok &= bool(good)
            if verbose:
                print(" zstd", cut, "->", len(z), "round trip through reference ZSTD_decompress:", "ok" if good else "FAIL (%d)" % r)
    finally:
        lib.ebcc_cuda_ctx_destroy(ctx)
    return bool(ok)


def check_cuts(frames):
    """every truncation point of every code-block: recording + verdict of the encoder vs a real truncated decode.
    Returns (cuts checked, flagged unproven, flagged proven but different, flagged unproven but equal, unproven cuts found
    in the pre-decoded pool, pool entries that differ from the real decode)."""
    f = np.ascontiguousarray(frames, dtype=np.float32)
    b, h, w = f.shape
    lib = cuda_api()
    lib.ebcc_cuda_debug_check_cuts.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
    ctx = C.c_void_p(lib.ebcc_cuda_ctx_create(0))
    assert ctx.value
    try:
        assert lib.ebcc_cuda_reserve(ctx, h, w, b) == 0
        assert lib.ebcc_cuda_upload(ctx, f.ctypes.data, b) == 0
        assert lib.ebcc_cuda_analyze_resident(ctx, b) == 0
        cnt = (C.c_ulonglong * 6)()
        assert lib.ebcc_cuda_debug_check_cuts(ctx, b, cnt) == 0
        return tuple(int(x) for x in cnt)
    finally:
        lib.ebcc_cuda_ctx_destroy(ctx)


def probe_sequence(frames, crs, target, recording):
    """error statistics of a sequence of rate probes over a resident batch; recording=False forces a real tier-1
    decode of every block in every probe (EBCC_NO_RECORDING=1), recording="decoder" makes the recordings with the
    tier-1 decoder and its truncation shadows (EBCC_DECODER_RECORDING=1) instead of taking them from the encoder"""
    import os
    f = np.ascontiguousarray(frames, dtype=np.float32)
    b, h, w = f.shape
    os.environ.pop("EBCC_NO_RECORDING", None)
    os.environ.pop("EBCC_DECODER_RECORDING", None)
    if recording == "decoder":
        os.environ["EBCC_DECODER_RECORDING"] = "1"
    elif not recording:
        os.environ["EBCC_NO_RECORDING"] = "1"
    lib = cuda_api()
    ctx = C.c_void_p(lib.ebcc_cuda_ctx_create(0))
    assert ctx.value
    out = []
    try:
        assert lib.ebcc_cuda_reserve(ctx, h, w, b) == 0
        assert lib.ebcc_cuda_upload(ctx, f.ctypes.data, b) == 0
        assert lib.ebcc_cuda_analyze_resident(ctx, b) == 0
        for cr in crs:
            crs_a = (C.c_float * b)(*([cr] * b))
            tg = (C.c_float * b)(*([target] * b))
            pr = (ProbeResult * b)()
            assert lib.ebcc_cuda_rate_probe(ctx, b, crs_a, tg, 0, pr) == 0
            out.append([(p.n_over, p.max_err, p.rmin, p.rmax, p.stream_bytes) for p in pr])
    finally:
        lib.ebcc_cuda_ctx_destroy(ctx)
        os.environ.pop("EBCC_NO_RECORDING", None)
        os.environ.pop("EBCC_DECODER_RECORDING", None)
    return out
