"""Drop-in boundary: the product .so driven exactly as the reference's own callers drive the reference --
the zarr codec's in-place decode, libhdf5's filter callback with its buffer-ownership swap, the byte-changing
environment variables, ebcc_encode_chunking_compat, NaN/Inf -> exit(1) -- every case compared with the UNMODIFIED reference
(oracle/_ref).  The `ref`-only variants run on CPU (they pin the call patterns themselves); the product needs the GPU."""
import ctypes as C
import hashlib
import json
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

import dropin_api as D
import harness as H

needs_ref = pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
gpu = pytest.mark.gpu


def parse(stream):
    magic, ver, flags, _r, mn, mx, coeffs, rmn, rmx, zsz, tail = struct.unpack("<4sBBHIIQIIQQ", stream[:48])
    return dict(flags=flags, mn=mn, mx=mx, coeffs=coeffs, rmn=rmn, rmx=rmx, zsz=zsz, tail=tail, z=stream[48:48 + zsz], j2k=stream[48 + zsz:])


def spiht_prefix(p):
    if not p["coeffs"]:
        return b""
    ref = H.ref_lib()
    ref.ZSTD_decompress.restype = C.c_size_t
    ref.ZSTD_decompress.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.c_size_t]
    back = C.create_string_buffer(p["coeffs"])
    assert ref.ZSTD_decompress(back, p["coeffs"], p["z"], len(p["z"])) == p["coeffs"]
    return back.raw


def same_decisions(ours, want, ulps=1):
    """everything but the zstd bytes (any valid frame is allowed there) must match the reference"""
    a, b = parse(ours), parse(want)
    assert a["flags"] == b["flags"] and a["j2k"] == b["j2k"]
    assert a["coeffs"] == b["coeffs"] and a["rmn"] == b["rmn"] and a["rmx"] == b["rmx"]
    assert abs(a["mn"] - b["mn"]) <= ulps and abs(a["mx"] - b["mx"]) <= ulps
    assert spiht_prefix(a) == spiht_prefix(b)
    assert abs(len(ours) - len(want)) <= 0.01 * len(want)


# ---- (i) zarr codec: decode(buf, out=arr) writes into arr -------------------------------------------------------------------
def _zarr_roundtrip(lib, mode, err):
    f = H.synth_stack(1, 96, 128)[0]
    z = D.ZarrDecodeLikeReference(lib, D.cd_values(96, 128, 10.0, mode, err))
    enc = z.encode(f)
    fresh = z.decode(enc)
    arr = np.full(96 * 128, -1.0, dtype=np.float32)
    addr = arr.ctypes.data
    back = z.decode(enc, out=arr)
    assert back is arr and arr.ctypes.data == addr
    assert np.array_equal(arr.view(np.uint32), fresh.view(np.uint32))      # written in place, same floats as a fresh buffer
    return enc, fresh


@needs_ref
@pytest.mark.parametrize("mode,err", [(H.NONE, None), (H.MAX_ERROR, 0.05)])
def test_zarr_inplace_decode_reference_pattern(mode, err):
    _zarr_roundtrip(H.ref_lib(), mode, err)


@gpu
@needs_ref
@pytest.mark.parametrize("mode,err", [(H.NONE, None), (H.MAX_ERROR, 0.05)])
def test_zarr_inplace_decode_product(built, mode, err):
    enc, fresh = _zarr_roundtrip(H.prod_lib(), mode, err)
    renc, rfresh = _zarr_roundtrip(H.ref_lib(), mode, err)
    if mode == H.NONE:
        assert enc == renc
    else:
        same_decisions(enc, renc)
    assert np.array_equal(fresh.view(np.uint32), rfresh.view(np.uint32))
    # and across: the reference's codec object decodes product bytes into a caller array, and vice versa
    for lib, data in ((H.ref_lib(), enc), (H.prod_lib(), renc)):
        arr = np.zeros(96 * 128, dtype=np.float32)
        D.ZarrDecodeLikeReference(lib, D.cd_values(96, 128, 10.0, mode, err)).decode(data, out=arr)
        assert np.array_equal(arr.view(np.uint32), fresh.view(np.uint32))


@gpu
def test_const_field_always_fresh_buffer(built):
    """reference :1278: the const-field branch mallocs a new buffer and overwrites the caller's pointer"""
    lib = H.prod_lib()
    f = np.full((64, 64), 3.25, dtype=np.float32)
    s = H.encode(lib, f, (1, 64, 64), 5.0, H.NONE)
    assert parse(s)["flags"] & 1
    mine = np.zeros(64 * 64, dtype=np.float32)
    out = mine.ctypes.data_as(C.POINTER(C.c_float))
    buf = (C.c_uint8 * len(s)).from_buffer_copy(s)
    assert lib.ebcc_decode(buf, len(s), C.byref(out)) == 64 * 64
    assert C.cast(out, C.c_void_p).value != mine.ctypes.data
    assert np.all(np.ctypeslib.as_array(out, shape=(64 * 64,)) == np.float32(3.25)) and np.all(mine == 0)
    lib.free_buffer(out)


# ---- (ii) HDF5 filter callback --------------------------------------------------------------------------------------------------
def _h5z_roundtrip(lib, f, mode, err):
    h, w = f.shape[-2:]
    cd = D.cd_values(h, w, 12.0, mode, err)
    cls = D.plugin_class(lib)
    assert lib.H5PLget_plugin_type() == 0 and cls.id == 308 and cls.version == 1 and cls.encoder_present == 1 and cls.decoder_present == 1
    assert cls.name == b"HDF5 EBCC filter L&L" and not cls.can_apply and not cls.set_local
    raw = np.ascontiguousarray(f, dtype=np.float32).tobytes()
    ret, size, buf = D.h5z_filter(lib, 0, cd, raw)
    assert ret == size and ret > 48 and buf.value          # bytes written; the chunk buffer was replaced
    stream = D.take(buf, ret)
    assert stream[:4] == b"EBCC"
    ret2, size2, buf2 = D.h5z_filter(lib, D.H5Z_FLAG_REVERSE, cd, stream)
    # quirk kept from the reference (src/h5z_ebcc.c:129-134): the decode direction reports the ELEMENT count
    assert ret2 == size2 == f.size
    dec = np.frombuffer(D.take(buf2, 4 * ret2), dtype=np.float32)
    return stream, dec


@needs_ref
def test_h5z_callback_reference_pattern():
    f = H.synth_stack(1, 64, 96)[0]
    s, d = _h5z_roundtrip(H.ref_lib(), f, H.NONE, None)
    assert s == H.encode(H.ref_lib(), f, (1, 64, 96), 12.0, H.NONE)
    assert np.abs(d - f.ravel()).max() < 10.0


@gpu
@needs_ref
@pytest.mark.parametrize("mode,err,shape", [(H.NONE, None, (64, 96)), (H.MAX_ERROR, 0.05, (64, 96)), (H.RELATIVE_ERROR, 1e-3, (64, 96)),
                                            (H.NONE, None, (3, 40, 64))])
def test_h5z_callback_product(built, mode, err, shape):
    f = H.synth_stack(shape[0], shape[1], shape[2]) if len(shape) == 3 else H.synth_stack(1, shape[0], shape[1])[0]
    s, d = _h5z_roundtrip(H.prod_lib(), f, mode, err)
    rs, rd = _h5z_roundtrip(H.ref_lib(), f, mode, err)
    if mode == H.NONE:
        assert s == rs
    else:
        same_decisions(s, rs)
    assert np.array_equal(d.view(np.uint32), rd.view(np.uint32))
    # each plugin decodes the other's chunk
    cd = D.cd_values(f.shape[-2], f.shape[-1], 12.0, mode, err)
    for lib, data, want in ((H.ref_lib(), s, d), (H.prod_lib(), rs, rd)):
        n, _, buf = D.h5z_filter(lib, D.H5Z_FLAG_REVERSE, cd, data)
        assert n == f.size and np.array_equal(np.frombuffer(D.take(buf, 4 * n), dtype=np.float32).view(np.uint32), want.view(np.uint32))


@gpu
@pytest.mark.parametrize("mode,err", [(H.NONE, None), (H.MAX_ERROR, 0.05)])
def test_batch_route_writes_what_the_filter_callback_writes(built, mode, err):
    """libhdf5 calls a filter once per chunk and waits for it (src/h5z_ebcc.c:124-148), so chunk writes cannot be batched
    behind the callback.  The high-throughput route is on the caller's side: ebcc_encode_batch over all the chunks, then
    direct chunk writes (H5Dwrite_chunk / h5py write_direct_chunk with filter_mask 0).  What lands in the file must be what
    the callback would have produced, chunk by chunk."""
    lib = H.prod_lib()
    lib.ebcc_encode_batch.restype = C.c_size_t
    frames = H.synth_stack(7, 64, 96)
    frames[4] = -2.5                                                       # a constant chunk
    n = len(frames)
    cfg = H.CodecConfig(); cfg.dims[:] = [1, 64, 96]; cfg.base_cr = 12.0; cfg.mode = mode; cfg.error = err or 0.0
    outs, sizes = (C.POINTER(C.c_uint8) * n)(), (C.c_size_t * n)()
    assert lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.byref(cfg), outs, sizes) == n
    cd = D.cd_values(64, 96, 12.0, mode, err)
    for i in range(n):
        batch_chunk = C.string_at(outs[i], sizes[i])
        lib.free_buffer(outs[i])
        ret, _, buf = D.h5z_filter(lib, 0, cd, frames[i].tobytes())
        assert ret == len(batch_chunk) and D.take(buf, ret) == batch_chunk, i
        # and the read side: the callback decodes the chunk the batch call wrote
        m, _, back = D.h5z_filter(lib, D.H5Z_FLAG_REVERSE, cd, batch_chunk)
        assert m == frames[i].size
        dec = np.frombuffer(D.take(back, 4 * m), dtype=np.float32)
        # (the bound is enforced before the reference's mean adjustment shifts min / max, src/ebcc_codec.c:863-868: allow the shift)
        assert np.abs(dec - frames[i].ravel()).max() <= (1.5 * err if mode == H.MAX_ERROR else 5.0)


@gpu
def test_h5z_callback_failure_returns_zero(built):
    """HDF5 convention: 0 = failure; the input chunk is freed and replaced by NULL like the reference does"""
    ret, size, buf = D.h5z_filter(H.prod_lib(), D.H5Z_FLAG_REVERSE, D.cd_values(64, 64, 10.0, H.NONE), b"EBCC" + bytes(60))
    assert ret == 0 and size == 0 and not buf.value


# ---- (iii) environment variables that change the bytes (src/ebcc_codec.c:634-649) -----------------------------------------------
ENV_CASES = [({"EBCC_INIT_BASE_ERROR_QUANTILE": "0"}, 0.2), ({"EBCC_INIT_BASE_ERROR_QUANTILE": "1e-5"}, 0.2),
             ({"EBCC_INIT_BASE_ERROR_QUANTILE": "1e-4"}, 0.2), ({"EBCC_INIT_BASE_ERROR_QUANTILE": "1e-2"}, 0.05),
             ({"EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK": "1"}, 0.2), ({"EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK": "1"}, 0.01),
             ({"EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK_CONSISTENCY": "1"}, 0.2),
             ({"EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK_CONSISTENCY": "1"}, 0.02),
             ({"EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK": "1", "EBCC_INIT_BASE_ERROR_QUANTILE": "1e-3"}, 0.05),
             ({"EBCC_DISABLE_MEAN_ADJUSTMENT": "1", "EBCC_INIT_BASE_ERROR_QUANTILE": "0"}, 0.1)]


def with_env(env, fn):
    os.environ.update(env)
    try:
        return fn()
    finally:
        for k in env:
            del os.environ[k]


@gpu
@needs_ref
@pytest.mark.parametrize("env,err", ENV_CASES)
@pytest.mark.parametrize("shape", [(128, 192), (200, 300)])
def test_env_vars_match_reference(built, env, err, shape):
    f = H.synth_stack(1, shape[0], shape[1])[0]
    dims = (1,) + shape
    ours = with_env(env, lambda: H.encode(H.prod_lib(), f, dims, 30.0, H.MAX_ERROR, err))
    want = with_env(env, lambda: H.encode(H.ref_lib(), f, dims, 30.0, H.MAX_ERROR, err))
    base = H.encode(H.ref_lib(), f, dims, 30.0, H.MAX_ERROR, err)
    same_decisions(ours, want)
    assert np.array_equal(H.decode(H.prod_lib(), ours), H.decode(H.ref_lib(), ours))
    test_env_vars_match_reference.changed = getattr(test_env_vars_match_reference, "changed", 0) + (parse(want)["j2k"] != parse(base)["j2k"] or parse(want)["coeffs"] != parse(base)["coeffs"])


@gpu
@needs_ref
def test_residual_misses_target_keeps_full_residual(built):
    """When the SPIHT layer cannot reach the target the reference keeps the FULL residual in the stream unless the pure
    base-layer fallback replaces it (src/ebcc_codec.c:752-758, :838-846).  A frame with a few huge spikes at a loose
    quantile forces that path; with the fallback disabled the stream must still carry the residual."""
    rng = np.random.default_rng(5)
    f = H.synth_stack(1, 96, 128)[0].copy()
    idx = rng.choice(f.size, 40, replace=False)
    f.ravel()[idx] += rng.choice([-1.0, 1.0], 40).astype(np.float32) * 3000.0
    kept = 0
    # quantile 3e-2 leaves outliers the 8-bit residual cannot reach at target 8.0 ("Could not reach error target", reference :756)
    for env in ({"EBCC_INIT_BASE_ERROR_QUANTILE": "3e-2", "EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK": "1"}, {"EBCC_INIT_BASE_ERROR_QUANTILE": "3e-2"}):
        for err in (8.0, 2.0):
            ours = with_env(env, lambda: H.encode(H.prod_lib(), f, (1, 96, 128), 200.0, H.MAX_ERROR, err))
            want = with_env(env, lambda: H.encode(H.ref_lib(), f, (1, 96, 128), 200.0, H.MAX_ERROR, err))
            same_decisions(ours, want)
            assert np.array_equal(H.decode(H.prod_lib(), ours), H.decode(H.ref_lib(), ours))
            kept += len(env) == 2 and err == 8.0 and parse(want)["coeffs"] > 0
    assert kept == 1          # fallback disabled: the stream carries the full-length residual although it missed the target


# ---- (iv) ebcc_encode_chunking_compat (src/ebcc_codec.c:1054-1090; reference tests/test_c_api.py:260-289) ------------------------
def _chunk_header(b):
    magic, ver, nd, _r = struct.unpack("<4sIII", b[:16])
    v = struct.unpack("<8Q", b[16:80])
    return dict(magic=magic, version=ver, ndims=nd, dims=v[0:3], chunk_dims=v[3:6], num_chunks=v[6], chunk_size=v[7])


def _chunks(b):
    h, off, out = _chunk_header(b), 80, []
    for _ in range(h["num_chunks"]):
        n = struct.unpack("<Q", b[off:off + 8])[0]
        out.append(b[off + 8:off + 8 + n])
        off += 8 + n
    assert off == len(b)
    return out


def _compat_case(lib, shape, mode, err, chunk_dims=(0, 0, 0)):
    rng = np.random.default_rng(11)
    data = (np.cumsum(rng.standard_normal(shape), axis=-1) * 0.5 + 20.0).astype(np.float32)
    return data, H.encode(lib, data, shape, 8.0, mode, err, fn="ebcc_encode_chunking_compat", chunk_dims=chunk_dims)


@needs_ref
def test_chunking_compat_reference_pattern():
    data, enc = _compat_case(H.ref_lib(), (1, 2048, 32), H.NONE, 0.0)
    h = _chunk_header(enc)
    assert h["magic"] == b"EBCK" and h["chunk_dims"] == (1, 1024, 32) and h["num_chunks"] == 2 and h["chunk_size"] == 1024 * 32


@gpu
@needs_ref
@pytest.mark.parametrize("shape,mode,err,cd", [((1, 2048, 32), H.NONE, 0.0, (0, 0, 0)), ((2, 32, 32), H.RELATIVE_ERROR, 0.01, (0, 0, 0)),
                                               ((3, 48, 80), H.RELATIVE_ERROR, 0.005, (0, 0, 0)), ((2, 64, 64), H.MAX_ERROR, 0.05, (0, 0, 0)),
                                               ((1, 40, 2100), H.NONE, 0.0, (0, 0, 0)), ((2, 70, 50), H.RELATIVE_ERROR, 0.01, (1, 40, 40))])
def test_chunking_compat_matches_reference(built, shape, mode, err, cd):
    data, ours = _compat_case(H.prod_lib(), shape, mode, err, cd)
    _, want = _compat_case(H.ref_lib(), shape, mode, err, cd)
    assert ours[:80] == want[:80]                        # container header: dims, default chunk dims, counts
    a, b = _chunks(ours), _chunks(want)
    assert len(a) == len(b)
    for x, y in zip(a, b):
        if mode == H.NONE:
            assert x == y
        else:
            same_decisions(x, y)
    da, db = H.decode(H.prod_lib(), ours, "ebcc_decode_chunking"), H.decode(H.ref_lib(), ours, "ebcc_decode_chunking")
    assert da is not None and np.array_equal(da.view(np.uint32), db.view(np.uint32)) and da.size == data.size
    dc = H.decode(H.prod_lib(), want, "ebcc_decode_chunking")
    assert np.array_equal(dc.view(np.uint32), H.decode(H.ref_lib(), want, "ebcc_decode_chunking").view(np.uint32))
    if mode == H.RELATIVE_ERROR:
        # the global range turns the relative bound into an absolute one (:1077-1087); the mean adjustment may add its shift
        bound = err * float(data.max() - data.min())
        assert np.abs(da - data.ravel()).max() <= bound * 1.5


# ---- (v) NaN / Inf -> log_fatal + exit(1) (src/ebcc_codec.c:598-605) -----------------------------------------------------------
_NAN_SNIPPET = """
import sys, numpy as np
sys.path.insert(0, {tests!r})
import harness as H
lib = H.{which}()
f = H.synth_stack(1, 64, 64)[0].copy()
f[{pos}] = np.float32({val!r})
n = H.encode(lib, f, (1, 64, 64), 10.0, {mode}, 0.1, fn={fn!r})
print("returned", len(n))
sys.exit(7)
"""


def _nan_exit(which, val, mode=0, pos="10, 20", fn="ebcc_encode"):
    code = _NAN_SNIPPET.format(tests=os.path.dirname(os.path.abspath(__file__)), which=which, val=val, mode=mode, pos=pos, fn=fn)
    return subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)


@needs_ref
def test_nan_exit_reference_pattern():
    r = _nan_exit("ref_lib", "nan")
    assert r.returncode == 1 and "returned" not in r.stdout


@gpu
@pytest.mark.parametrize("val", ["nan", "inf", "-inf"])
@pytest.mark.parametrize("mode,fn,pos", [(H.NONE, "ebcc_encode", "10, 20"), (H.MAX_ERROR, "ebcc_encode", "63, 63"), (H.NONE, "ebcc_encode_chunking", "0, 0"),
                                         (H.RELATIVE_ERROR, "ebcc_encode_chunking_compat", "31, 5")])
def test_nan_inf_exits_with_status_1(built, val, mode, fn, pos):
    r = _nan_exit("prod_lib", val, mode, pos, fn)
    assert r.returncode == 1 and "returned" not in r.stdout, (r.returncode, r.stdout, r.stderr[-400:])
    assert "NaN or Inf" in r.stderr


# ---- (vi) the reference's bundled frame at full size: BASELINE.json configs[0] and SURVEY 8(c)'s known answers ------------------
def _known():
    with open(os.path.join(GOLDEN, "era5_known_answers.json")) as fh:
        return json.load(fh)


def test_known_answers_fixture_is_consistent():
    k = _known()
    f = np.load(os.path.join(GOLDEN, "era5_t2m_721x1440.npy"))
    assert list(f.shape) == k["shape"] == [721, 1440] and hashlib.sha256(f.tobytes()).hexdigest() == k["input_sha256"]
    c = k["cases"]
    # what the survey recorded from the reference: sizes, coeffs_size, header min/max bits
    assert c["none_30"]["length"] == 138455 and c["max_0.5"]["length"] == 47177 and c["max_0.5"]["coeffs_size"] == 4274
    assert c["max_0.5"]["zstd_size_reference"] == 867 and c["max_0.5"]["j2k_size"] == 46262
    assert c["max_0.1"]["length"] == 125344 and c["rel_1e-3"]["length"] == 192510
    assert struct.pack("<I", c["none_30"]["min_bits"]).hex() == "f3516843" and struct.pack("<I", c["none_30"]["max_bits"]).hex() == "2f7b8f43"


@gpu
@pytest.mark.parametrize("name", ["none_30", "max_0.5", "max_0.1", "rel_1e-3", "max_0.5_q0", "max_0.5_q1e-5", "max_0.5_q1e-4", "max_0.5_nopure",
                                  "max_0.1_noconsist", "max_0.5_nomean"])
def test_era5_frame_known_answers(built, name):
    k = _known()
    c = k["cases"][name]
    f = np.load(os.path.join(GOLDEN, "era5_t2m_721x1440.npy"))
    s = with_env(c["env"], lambda: H.encode(H.prod_lib(), f, (1, 721, 1440), k["base_cr"], c["mode"], c["error"]))
    p = parse(s)
    assert hashlib.sha256(p["j2k"]).hexdigest() == c["j2k_sha256"] and p["tail"] == c["j2k_size"]
    assert p["coeffs"] == c["coeffs_size"] and p["rmn"] == c["rmin_bits"] and p["rmx"] == c["rmax_bits"] and p["flags"] == c["flags"]
    assert abs(p["mn"] - c["min_bits"]) <= 1 and abs(p["mx"] - c["max_bits"]) <= 1
    if c["coeffs_size"]:
        d = H.decode(H.prod_lib(), s)          # the SPIHT prefix is checked through our own zstd decoder and the reference's
        assert d is not None
        if H.have_ref():
            assert hashlib.sha256(spiht_prefix(p)).hexdigest() == c["spiht_sha256"]
    if c["mode"] == H.NONE:
        assert len(s) == c["length"]
    else:
        assert abs(len(s) - c["length"]) <= 0.01 * c["length"]
    if H.have_ref():
        assert np.array_equal(H.decode(H.prod_lib(), s).view(np.uint32), H.decode(H.ref_lib(), s).view(np.uint32))
