"""GPU parity, error-bounded modes (MAX_ERROR / RELATIVE_ERROR): the CUDA path must take the same decisions as the
reference -- compression ratio of the base layer, SPIHT stream bytes, truncation length -- so that everything in the
EBCC stream except the zstd payload (any valid frame is allowed there) is byte-identical, the error bound holds with
zero violations, and the size is within 1 % of the reference."""
import os
import struct

import numpy as np
import pytest

import harness as H
import stage_api as S

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def parse(stream):
    magic, ver, flags, _r, mn, mx, coeffs, rmn, rmx, zsz, tail = struct.unpack("<4sBBHIIQIIQQ", stream[:48])
    return dict(mn=mn, mx=mx, coeffs=coeffs, rmn=rmn, rmx=rmx, zsz=zsz, tail=tail, z=stream[48:48 + zsz], j2k=stream[48 + zsz:])


def ulp_close(a_bits, b_bits, tol=1):
    return abs(int(a_bits) - int(b_bits)) <= tol


def compare_with_reference(f, cr, mode, err):
    lib, ref = H.prod_lib(), H.ref_lib()
    h, w = f.shape
    ours, want = H.encode(lib, f, (1, h, w), cr, mode, err), H.encode(ref, f, (1, h, w), cr, mode, err)
    assert ours and want
    a, b = parse(ours), parse(want)
    # identical decisions: base-layer codestream bytes (=> chosen ratio), truncation length, residual range
    assert a["j2k"] == b["j2k"]
    assert a["coeffs"] == b["coeffs"] and a["rmn"] == b["rmn"] and a["rmx"] == b["rmx"]
    # min/max carry the mean-error adjustment: a double sum whose order differs -> at most 1 ulp of float
    assert ulp_close(a["mn"], b["mn"]) and ulp_close(a["mx"], b["mx"])
    assert abs(len(ours) - len(want)) <= 0.01 * len(want)
    # The bound is enforced on the reconstruction BEFORE the reference's mean-error adjustment shifts min/max
    # (src/ebcc_codec.c:863-868), so a decoded frame may exceed it by at most that shift; identical in both codecs.
    target = np.float32(err) if mode == H.MAX_ERROR else np.float32(err) * (f.max() - f.min())
    shift = abs(float(np.frombuffer(struct.pack("<I", a["mn"]), dtype=np.float32)[0]) - float(f.min()))
    for st in (ours, want):
        for dl in (lib, ref):
            d = H.decode(dl, st)
            assert d is not None and np.abs(d - f.ravel()).max() <= float(target) + shift + 1e-6 * float(np.abs(f).max())
    os.environ["EBCC_DISABLE_MEAN_ADJUSTMENT"] = "1"
    try:
        strict = H.encode(lib, f, (1, h, w), cr, mode, err)
        assert strict == H.encode(ref, f, (1, h, w), cr, mode, err) or parse(strict)["j2k"] == parse(H.encode(ref, f, (1, h, w), cr, mode, err))["j2k"]
        assert np.abs(H.decode(lib, strict) - f.ravel()).max() <= target      # zero violations without the adjustment
    finally:
        del os.environ["EBCC_DISABLE_MEAN_ADJUSTMENT"]
    # both decoders agree bit for bit on both streams
    assert np.array_equal(H.decode(lib, ours), H.decode(ref, ours))
    assert np.array_equal(H.decode(lib, want), H.decode(ref, want))
    if a["coeffs"]:
        rf = S.ref_spiht_api()
        import ctypes as C
        back_a, back_b = C.create_string_buffer(a["coeffs"]), C.create_string_buffer(b["coeffs"])
        assert rf.ZSTD_decompress(back_a, a["coeffs"], a["z"], len(a["z"])) == a["coeffs"]
        assert rf.ZSTD_decompress(back_b, b["coeffs"], b["z"], len(b["z"])) == b["coeffs"]
        assert back_a.raw == back_b.raw                      # the SPIHT prefix inside the zstd frame is byte-identical
    return a, b


@needs_ref
@pytest.mark.parametrize("shape,cr,err", [((96, 160), 30.0, 0.5), ((96, 160), 30.0, 0.05), ((128, 256), 10.0, 0.01), ((200, 300), 50.0, 0.2)])
def test_max_error_small(built, shape, cr, err):
    compare_with_reference(H.synth_stack(1, shape[0], shape[1])[0], cr, H.MAX_ERROR, err)


@needs_ref
def test_relative_error(built):
    compare_with_reference(H.synth_stack(1, 128, 256)[0], 30.0, H.RELATIVE_ERROR, 1e-3)


@needs_ref
@pytest.mark.parametrize("err", [0.5, 0.1])
def test_era5_crop_real_data(built, err):
    f = np.load(os.path.join(GOLDEN, "era5_t2m_crop_256x384.npy"))
    compare_with_reference(f, 30.0, H.MAX_ERROR, err)


@needs_ref
@pytest.mark.parametrize("err", [0.5, 0.1])
def test_era5_full_frame(built, err):
    compare_with_reference(H.synth_stack(1)[0], 30.0, H.MAX_ERROR, err)


@needs_ref
@pytest.mark.parametrize("shape,cr", [((64, 96), 20.0), ((257, 129), 20.0), ((721, 1440), 90.0)])
def test_residual_stage(built, shape, cr):
    """SPIHT stream bytes, truncation probes (max / mean error) and the zstd frame vs the reference, stage by stage."""
    assert S.check_residual_stage(H.synth_stack(1, shape[0], shape[1])[0], cr)


def test_batch_equals_single_max_error(built):
    import ctypes as C
    lib = H.prod_lib()
    lib.ebcc_encode_batch.restype = C.c_size_t
    frames = H.synth_stack(3, 160, 224)
    cfg = H.CodecConfig()
    cfg.dims[:] = [1, 160, 224]
    cfg.base_cr = 20.0
    cfg.mode = H.MAX_ERROR
    cfg.error = 0.1
    outs, sizes = (C.POINTER(C.c_uint8) * 3)(), (C.c_size_t * 3)()
    assert lib.ebcc_encode_batch(frames.ctypes.data_as(C.c_void_p), C.c_size_t(3), C.byref(cfg), outs, sizes) == 3
    for t in range(3):
        got = C.string_at(outs[t], sizes[t])
        lib.free_buffer(outs[t])
        assert got == H.encode(lib, frames[t], (1, 160, 224), 20.0, H.MAX_ERROR, 0.1)
        assert np.abs(H.decode(lib, got) - frames[t].ravel()).max() <= 0.1


@pytest.mark.parametrize("shape,nf", [((96, 160), 2), ((64, 64), 3), ((200, 130), 1)])
def test_every_cut_verdict_small(built, shape, nf):
    """the encoder's recording + its verdict per truncation point vs a real truncated decode, for EVERY pass of EVERY block"""
    checked, unproven, wrong, pessimistic, pooled, pool_wrong = S.check_cuts(H.synth_stack(nf, shape[0], shape[1]))
    assert checked > 0 and wrong == 0 and pooled == unproven and pool_wrong == 0, (checked, unproven, wrong, pessimistic, pooled, pool_wrong)


def test_every_cut_verdict_full_frames(built):
    frames = np.concatenate([H.synth_stack(2), np.load(os.path.join(GOLDEN, "era5_t2m_721x1440.npy"))[None]], axis=0)
    checked, unproven, wrong, pessimistic, pooled, pool_wrong = S.check_cuts(frames)
    print("cuts checked %d, unproven %d, proven-but-different %d, unproven-but-equal %d, pooled %d, pool-differs %d" % (checked, unproven, wrong, pessimistic, pooled, pool_wrong))
    assert checked > 20000 and wrong == 0
    assert pessimistic == 0          # the criterion is exact: a cut it rejects really decodes differently
    assert pooled == unproven and pool_wrong == 0     # every unproven cut was decoded ahead of the search, correctly


def test_pool_entries_from_bitplane_boundary(built):
    """k_t1_dec_cuts starts an entry at the bit-plane boundary below its last pass (sample state from the recording, contexts
    from the encoder's snapshots, C register replayed from the final bytes) instead of pass 0.  With every cut declared
    unproven (test aid) the pool fills up with entries of every pass count; each must equal the real truncated decode,
    and the pool decoded from pass 0 (EBCC_CUTS_FROM_PASS0) must say the same."""
    frames = np.concatenate([H.synth_stack(1, 300, 460), H.synth_stack(1, 300, 460, level=3)], axis=0)
    keep = {k: os.environ.get(k) for k in ("EBCC_DEBUG_ALL_CUTS_UNPROVEN", "EBCC_CUTS_FROM_PASS0")}
    try:
        os.environ["EBCC_DEBUG_ALL_CUTS_UNPROVEN"] = "1"
        fast = S.check_cuts(frames)
        os.environ["EBCC_CUTS_FROM_PASS0"] = "1"
        slow = S.check_cuts(frames)
    finally:
        for k, v in keep.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    checked, unproven, wrong, pessimistic, pooled, pool_wrong = fast
    assert checked > 2000 and unproven == checked and pooled > 150 and pool_wrong == 0 and wrong == 0, fast
    assert slow == fast


def test_recording_equals_real_decodes(built):
    """CR-search probes served from the pass recording (one tier-1 decode + truncation shadows) must give exactly the
    statistics of probes that decode every block for real, for ratios below and above the recorded one."""
    frames = H.synth_stack(3)
    crs = [30.0, 60.0, 120.0, 90.0, 75.0, 82.5, 15.0, 7.5, 20.0, 45.0, 250.0, 2.0]
    a = S.probe_sequence(frames, crs, 0.3, True)
    b = S.probe_sequence(frames, crs, 0.3, False)
    assert a == b
    assert a == S.probe_sequence(frames, crs, 0.3, "decoder")       # round-1 recordings (tier-1 decoder + shadows) agree too
