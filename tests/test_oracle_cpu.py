"""CPU-only tests: the oracle restatement against the committed golden vectors (made from the unmodified
reference by tests/golden/make_golden.py) and, when present, against the reference-derived checker; the
host-side logic of the product library; and the exported C ABI.  No compute call needs a GPU here."""
import ctypes as C
import hashlib
import os
import re
import struct

import numpy as np
import pytest

import harness as H
from golden.make_golden import CASES, make_input

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "j2k_golden.npz"))


def _ebcc_none_stream(frame, cr):
    """EBCC mode-NONE stream assembled from the oracle pieces (header layout: src/ebcc_codec.c:190-202)."""
    f = np.ascontiguousarray(frame, dtype=np.float32)
    j2k = H.orc_j2k_encode(H.scale_u16(f), cr)
    hdr = b"EBCC" + bytes([1, 0, 0, 0]) + struct.pack("<ff", f.min(), f.max()) + struct.pack("<Q", 0) + \
        struct.pack("<ff", 0.0, 0.0) + struct.pack("<QQ", 0, len(j2k))
    return hdr + j2k


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_matches_golden(built, case):
    name, h, w, cr, kind = case
    f = make_input(kind, h, w)
    if hashlib.sha256(f.tobytes()).digest() != GOLD[name + "_in_sha"].tobytes():
        pytest.skip("seeded input generator is not bit-reproducible on this machine")
    s = _ebcc_none_stream(f, cr)
    assert len(s) == int(GOLD[name + "_len"][0])
    assert hashlib.sha256(s).digest() == GOLD[name + "_sha"].tobytes()
    if name + "_stream" in GOLD:
        assert s == GOLD[name + "_stream"].tobytes()


@pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed", range(6))
def test_oracle_matches_reference_random(built, seed):
    rng = np.random.default_rng(seed)
    h, w = int(rng.integers(32, 160)), int(rng.integers(32, 160))
    cr = float(rng.choice([1.2, 2.5, 6.0, 17.0, 90.0, 700.0]))
    kind = seed % 3
    if kind == 0:
        img = rng.integers(0, 65536, (h, w)).astype(np.uint16)
    elif kind == 1:
        yy, xx = np.mgrid[0:h, 0:w]
        img = (32768 + 25000 * np.sin(xx / 11.0) * np.cos(yy / 7.0)).astype(np.uint16)
    else:
        img = H.scale_u16(H.synth_stack(1, h, w)[0])
    assert H.orc_j2k_encode(img, cr) == H.ref_j2k_encode(img, cr)


def test_scale_u16_matches_c(built):
    f = H.synth_stack(1, 40, 50)[0]
    out = np.zeros(f.size, dtype=np.uint16)
    H.orc_lib().orc_scale_u16(f.ctypes.data, f.size, C.c_float(f.min()), C.c_float(f.max()), out.ctypes.data)
    assert np.array_equal(out.reshape(f.shape), H.scale_u16(f))


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference sources absent")
def test_tables_match_reference_sources(built):
    """context / distortion LUTs and the MQ state table, generated from their definitions, equal the
    tables in the reference's t1_luts.h and mqc.c"""
    lib = H.orc_lib()
    zc, sc, spb = (C.c_uint8 * 2048)(), (C.c_uint8 * 256)(), (C.c_uint8 * 256)()
    nm = [(C.c_int16 * 128)() for _ in range(4)]
    lib.orc_get_luts(zc, sc, spb, *nm)
    src = open("/root/reference/src/openjpeg/src/lib/openjp2/t1_luts.h").read()

    def table(name):
        body = re.search(name + r"\[[^\]]*\]\s*=\s*\{([^}]*)\}", src).group(1)
        return [int(x, 0) for x in re.findall(r"0x[0-9a-fA-F]+|\d+", body)]
    assert list(zc) == table("lut_ctxno_zc")
    assert list(sc) == table("lut_ctxno_sc")
    assert list(spb) == table("lut_spb")
    for arr, name in zip(nm, ["lut_nmsedec_sig", "lut_nmsedec_sig0", "lut_nmsedec_ref", "lut_nmsedec_ref0"]):
        assert list(arr) == table(name)
    qe, nmps, nlps, sw = (C.c_uint16 * 47)(), (C.c_uint8 * 47)(), (C.c_uint8 * 47)(), (C.c_uint8 * 47)()
    lib.orc_get_mq_table(qe, nmps, nlps, sw)
    mqc = open("/root/reference/src/openjpeg/src/lib/openjp2/mqc.c").read()
    ent = re.findall(r"\{(0x[0-9a-fA-F]+),\s*(\d),\s*&mqc_states\[(\d+)\],\s*&mqc_states\[(\d+)\]\}", mqc)
    assert len(ent) == 94
    for i in range(47):
        e0, e1 = ent[2 * i], ent[2 * i + 1]
        assert int(e0[0], 16) == qe[i] and int(e0[2]) // 2 == nmps[i] and int(e0[3]) // 2 == nlps[i]
        assert (int(e0[3]) % 2 != int(e0[1])) == bool(sw[i])   # NLPS entry flips the MPS sense iff SWITCH
        assert int(e1[0], 16) == qe[i]


# ---- product library: ABI + host logic (no GPU needed) ----------------------------------------------------
def _declared_symbols():
    inc = os.path.join(H.ROOT, "include")
    names = set()
    for fn in os.listdir(inc):
        txt = open(os.path.join(inc, fn)).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        names |= set(re.findall(r"\b((?:ebcc_|free_buffer|print_config|log_set_level_from_env|populate_config)\w*)\s*\(", txt))
    return names


def test_library_exports_declared_abi(built):
    lib = C.CDLL(H.PROD_SO)
    for name in sorted(_declared_symbols()) + ["H5PLget_plugin_type", "H5PLget_plugin_info"]:
        assert hasattr(lib, name), name


def test_library_asks_for_hardware_queues_at_load_time(built):
    """16 pipelines need more than the driver's default 8 hardware queues (CUDA_DEVICE_MAX_CONNECTIONS); the variable has to be in
    the environment before the CUDA runtime linked into the library initialises, i.e. before the library's other constructors
    run -- and a value the application chose must survive."""
    import subprocess
    import sys
    code = ("import ctypes as C, sys\n"
            "libc = C.CDLL(None); libc.getenv.restype = C.c_char_p\n"
            "before = libc.getenv(b'CUDA_DEVICE_MAX_CONNECTIONS')\n"
            "C.CDLL(sys.argv[1])\n"
            "print(before, libc.getenv(b'CUDA_DEVICE_MAX_CONNECTIONS'))\n")
    env = {k: v for k, v in os.environ.items() if k != "CUDA_DEVICE_MAX_CONNECTIONS"}
    r = subprocess.run([sys.executable, "-c", code, H.PROD_SO], capture_output=True, text=True, env=env, timeout=120)
    assert r.returncode == 0 and r.stdout.split() == ["None", "b'32'"], (r.stdout, r.stderr[-300:])
    r = subprocess.run([sys.executable, "-c", code, H.PROD_SO], capture_output=True, text=True, env=dict(env, CUDA_DEVICE_MAX_CONNECTIONS="12"), timeout=120)
    assert r.returncode == 0 and r.stdout.split() == ["b'12'", "b'12'"], (r.stdout, r.stderr[-300:])


def test_plugin_info_and_populate_config(built):
    lib = C.CDLL(H.PROD_SO)
    lib.H5PLget_plugin_type.restype = C.c_int
    assert lib.H5PLget_plugin_type() == 0

    class H5Z(C.Structure):
        _fields_ = [("version", C.c_int), ("id", C.c_int), ("enc", C.c_uint), ("dec", C.c_uint), ("name", C.c_char_p),
                    ("can_apply", C.c_void_p), ("set_local", C.c_void_p), ("filter", C.c_void_p)]
    lib.H5PLget_plugin_info.restype = C.POINTER(H5Z)
    info = lib.H5PLget_plugin_info().contents
    assert (info.version, info.id, info.enc, info.dec) == (1, 308, 1, 1) and info.filter
    assert info.name == b"HDF5 EBCC filter L&L"
    cfg = H.CodecConfig()
    cd = (C.c_uint * 5)(721, 1440, struct.unpack("I", struct.pack("f", 30.0))[0], 1, struct.unpack("I", struct.pack("f", 0.5))[0])
    lib.populate_config(C.byref(cfg), C.c_size_t(5), cd, C.c_size_t(721 * 1440 * 4))
    assert list(cfg.dims) == [1, 721, 1440] and cfg.base_cr == 30.0 and cfg.mode == 1 and cfg.error == 0.5
    assert list(cfg.chunk_dims) == [0, 0, 0]


def test_invalid_dims_rejected_without_device(built):
    lib = H.prod_lib()
    f = np.zeros((16, 64), dtype=np.float32)
    assert H.encode(lib, f, (1, 16, 64), 10.0, H.NONE) == b""
    assert H.encode(lib, np.zeros((1, 64, 4000), np.float32), (1, 64, 4000), 10.0, H.NONE) == b""


def test_no_cpu_fallback(built):
    """without a CUDA device the product must fail loudly (return 0), never compute on the host"""
    lib = C.CDLL(H.PROD_SO)
    lib.ebcc_cuda_device_count.restype = C.c_int
    if lib.ebcc_cuda_device_count() > 0:
        pytest.skip("a CUDA device is present")
    f = H.synth_stack(1, 64, 64)[0]
    assert H.encode(H.prod_lib(), f, (1, 64, 64), 10.0, H.NONE) == b""


# ---- multi-tile images (dims[0] > 1): the oracle restatement against golden vectors made from the reference -------
from golden.make_golden_tiles import CASES as TILE_CASES, make_input as make_tile_input  # noqa: E402

TILES_GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tiles_golden.npz")


@pytest.mark.parametrize("case", [c for c in TILE_CASES if c[5] == "none"], ids=lambda c: c[0])
def test_tiled_oracle_matches_golden(built, case):
    """orc_j2k_encode_tiled (tile geometry at a non-zero canvas row, odd row parity in the vertical 9/7 pass, canvas-
    anchored code-block grid, per-tile byte budget) writes the reference's multi-tile codestream byte for byte"""
    name, t, h, w, cr, _mode, _err = case
    g = np.load(TILES_GOLDEN)
    f = make_tile_input(t, h, w)
    assert np.array_equal(np.frombuffer(hashlib.sha256(f.tobytes()).digest(), dtype=np.uint8), g[name + "_in_sha"])
    j2k = H.orc_j2k_encode_tiled(H.scale_u16(f), t, cr)
    assert np.array_equal(np.frombuffer(hashlib.sha256(j2k).digest(), dtype=np.uint8), g[name + "_j2k_sha"])
    if name + "_stream" in g:
        assert j2k == g[name + "_stream"].tobytes()[48:]


@pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("shape,cr", [((2, 40, 56), 3.0), ((3, 51, 33), 9.0), ((7, 35, 64), 50.0), ((2, 97, 130), 1.8)])
def test_tiled_oracle_matches_reference(built, shape, cr):
    t, h, w = shape
    f = np.random.default_rng(t * 1000 + h).standard_normal((t * h, w)).astype(np.float32).cumsum(axis=1)
    want = H.encode(H.ref_lib(), f, shape, cr, H.NONE)
    assert H.orc_j2k_encode_tiled(H.scale_u16(f), t, cr) == want[48:]


def test_tiled_oracle_single_tile_is_plain_encode(built):
    f = H.synth_stack(1, 70, 90)[0]
    u = H.scale_u16(f)
    assert H.orc_j2k_encode_tiled(u, 1, 7.0) == H.orc_j2k_encode(u, 7.0)
