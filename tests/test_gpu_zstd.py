"""GPU: the device zstd decompressor (k_zstd_decompress) on its own, fed with frames the REFERENCE's zstd 1.5.6 writes
(ZSTD_compress exported by oracle/_ref) at levels 1 / 3 / 19 / 22 over byte strings chosen to reach every block and
table kind the decoder claims (RFC 8878): raw and RLE blocks, Huffman literals with direct and FSE-compressed weights,
treeless literals of a second block, predefined / RLE / FSE / repeat sequence tables, long matches, repeat offsets,
several blocks per frame.  ebcc_decode only ever feeds it level-22 frames of SPIHT prefixes, which is why this exists."""
import ctypes as C

import numpy as np
import pytest

import harness as H

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")]


def _ref_zstd():
    ref = H.ref_lib()
    ref.ZSTD_compress.restype = C.c_size_t
    ref.ZSTD_compress.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_int]
    ref.ZSTD_compressBound.restype = C.c_size_t
    ref.ZSTD_compressBound.argtypes = [C.c_size_t]
    ref.ZSTD_isError.restype = C.c_uint
    ref.ZSTD_isError.argtypes = [C.c_size_t]
    return ref


def _compress(ref, data, level):
    cap = ref.ZSTD_compressBound(len(data))
    buf = C.create_string_buffer(cap)
    n = ref.ZSTD_compress(buf, cap, data, len(data), level)
    assert not ref.ZSTD_isError(n)
    return buf.raw[:n]


def _gpu_decompress(lib, frame, cap):
    out = C.create_string_buffer(max(cap, 1))
    got = C.c_size_t(0)
    rc = lib.ebcc_cuda_debug_zstd_decompress(0, frame, len(frame), out, cap, C.byref(got))
    return rc, out.raw[:got.value]


def _corpus():
    rng = np.random.default_rng(308)
    words = [b"temperature", b"pressure", b"geopotential", b"era5", b"level", b"hour", b" ", b"\n", b"0.25", b"721x1440"]
    text = b"".join(words[i] for i in rng.integers(0, len(words), 60000))
    skew = rng.choice(np.arange(256, dtype=np.uint8), size=200000, p=np.r_[np.full(8, 0.1), np.full(248, 0.2 / 248)]).tobytes()
    block = rng.integers(0, 256, 5000, dtype=np.uint8).tobytes()
    spiht_like = np.unpackbits(rng.integers(0, 256, 20000, dtype=np.uint8))
    spiht_like = np.packbits(spiht_like & np.unpackbits(rng.integers(0, 256, 20000, dtype=np.uint8))).tobytes()   # mostly zero bits
    return {
        "empty-ish": b"x",
        "random (raw blocks)": rng.integers(0, 256, 70000, dtype=np.uint8).tobytes(),
        "one byte repeated (RLE block)": b"\x5a" * 300000,
        "short period (long matches, repeat offsets)": (b"abcdefghij" * 40000)[:333333],
        "text (FSE sequence tables, Huffman literals)": text,
        "skewed bytes (FSE-compressed Huffman weights)": skew,
        "far copy (offset near the window)": block + rng.integers(0, 256, 150000, dtype=np.uint8).tobytes() + block,
        "several blocks, second one treeless": (text[:100000] + text[:100000][::-1] + skew[:90000]),
        "sparse bits (SPIHT prefix shape)": spiht_like,
        "zeros then noise then zeros": b"\0" * 140000 + rng.integers(0, 256, 3000, dtype=np.uint8).tobytes() + b"\0" * 140000,
    }


@pytest.mark.parametrize("level", [1, 3, 19, 22])
def test_device_decoder_on_reference_frames(built, level):
    ref = _ref_zstd()
    lib = C.CDLL(H.PROD_SO)
    lib.ebcc_cuda_debug_zstd_decompress.argtypes = [C.c_int, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
    for name, data in _corpus().items():
        frame = _compress(ref, data, level)
        rc, back = _gpu_decompress(lib, frame, len(data))
        assert rc == 0, (name, level, rc, len(frame))
        assert back == data, (name, level, len(frame), len(back))


def test_device_decoder_rejects_damaged_frames(built):
    ref = _ref_zstd()
    lib = C.CDLL(H.PROD_SO)
    lib.ebcc_cuda_debug_zstd_decompress.argtypes = [C.c_int, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
    data = _corpus()["text (FSE sequence tables, Huffman literals)"]
    frame = _compress(ref, data, 19)
    # truncated, too small an output buffer, wrong magic: clean failures, and the decoder is still usable afterwards
    assert _gpu_decompress(lib, frame[:len(frame) // 2], len(data))[0] != 0
    assert _gpu_decompress(lib, frame, len(data) // 2)[0] != 0
    assert _gpu_decompress(lib, b"\x00" + frame[1:], len(data))[0] != 0
    rng = np.random.default_rng(1)
    for _ in range(40):                                   # byte flips: anything but a crash or a hang is acceptable
        bad = bytearray(frame)
        for pos in rng.integers(4, len(bad), 3):
            bad[pos] ^= 1 << int(rng.integers(0, 8))
        rc, back = _gpu_decompress(lib, bytes(bad), len(data))
        assert rc != 0 or len(back) <= len(data)
    rc, back = _gpu_decompress(lib, frame, len(data))
    assert rc == 0 and back == data
