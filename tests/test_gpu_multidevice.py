"""GPU (two or more devices): one ebcc_encode_chunking / ebcc_decode_chunking call spread over the GPUs of the box
(EBCC_CUDA_DEVICES, pipelines of the pool live on device k mod n) must write the container a single GPU writes, byte
for byte -- frames are independent (reference frame loop, src/ebcc_codec.c:1007-1046) and the EBCK layout is a prefix
sum over per-frame sizes on the host.  The device list is read once per process, so every configuration runs in a
subprocess.  Skipped on a one-GPU box."""
import hashlib
import json
import os
import subprocess
import sys

import pytest

import harness as H

pytestmark = pytest.mark.gpu
TESTS = os.path.dirname(os.path.abspath(__file__))

SNIPPET = r"""
import ctypes as C, hashlib, json, os, sys
sys.path.insert(0, {tests!r})
import numpy as np, harness as H
lib = H.prod_lib()
frames = H.synth_stack(10, 96, 160)
frames[3] = 7.25                                        # a constant frame in the middle
cfg = H.CodecConfig(); cfg.dims[:] = [10, 96, 160]; cfg.chunk_dims[:] = [1, 96, 160]; cfg.base_cr = 12.0; cfg.mode = {mode}; cfg.error = {err}
out = C.POINTER(C.c_uint8)()
n = lib.ebcc_encode_chunking(frames.ctypes.data, C.byref(cfg), C.byref(out))
stream = C.string_at(out, n)
back = C.POINTER(C.c_float)()
m = lib.ebcc_decode_chunking(C.cast(out, C.c_void_p), n, C.byref(back))
dec = np.ctypeslib.as_array(back, shape=(m,)).copy() if m else np.zeros(0, np.float32)
lib.ebcc_cuda_total_launches.restype = C.c_ulonglong
print(json.dumps({{"n": n, "sha": hashlib.sha256(stream).hexdigest(), "m": m, "dec_sha": hashlib.sha256(dec.tobytes()).hexdigest(),
                  "max_err": float(np.abs(dec - frames.ravel()).max()) if m else -1.0}}))
"""


def _run(env, mode, err):
    code = SNIPPET.format(tests=TESTS, mode=mode, err=err)
    e = dict(os.environ, EBCC_BATCH_FRAMES="2", EBCC_PIPELINES="2", **env)
    e.pop("EBCC_CUDA_DEVICE" if "EBCC_CUDA_DEVICES" in env else "EBCC_CUDA_DEVICES", None)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=900, env=e)
    assert r.returncode == 0, r.stderr[-800:]
    return json.loads(r.stdout.strip().splitlines()[-1])


def _device_count():
    import ctypes as C
    lib = C.CDLL(H.PROD_SO)
    lib.ebcc_cuda_device_count.restype = C.c_int
    return lib.ebcc_cuda_device_count()


@pytest.mark.parametrize("mode,err", [(H.NONE, 0.0), (H.MAX_ERROR, 0.05)])
def test_container_from_two_gpus_equals_one_gpu(built, mode, err):
    if _device_count() < 2:
        pytest.skip("needs two GPUs")
    one = _run({"EBCC_CUDA_DEVICE": "0"}, mode, err)
    two = _run({"EBCC_CUDA_DEVICES": "0,1"}, mode, err)
    other = _run({"EBCC_CUDA_DEVICE": "1"}, mode, err)
    every = _run({"EBCC_CUDA_DEVICES": "all"}, mode, err)
    assert one["n"] > 80 and one["m"] == 10 * 96 * 160
    assert two == one and other == one and every == one
    if mode == H.MAX_ERROR:
        assert one["max_err"] <= err + 1e-3
