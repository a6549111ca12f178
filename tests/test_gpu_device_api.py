"""GPU: device-resident frames through the public C ABI.  ebcc_encode_batch / ebcc_encode /
ebcc_decode_batch take CUDA device pointers (unified virtual addressing); the streams and the decoded frames must be
identical to the ones of the host-memory path, which the other GPU tests pin against the reference."""
import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.mark.parametrize("mode,err", [("none", 0.0), ("max_error_target", 0.3)])
def test_device_tensor_equals_host_array(built, mode, err):
    from ebcc_b200.device_codec import EBCCDeviceCodec
    codec = EBCCDeviceCodec(12.0, mode, err)
    frames = H.synth_stack(5, 120, 200)
    host_streams = codec.encode(frames)
    dev = torch.from_numpy(frames).cuda()
    dev_streams = codec.encode(dev)
    assert host_streams == dev_streams
    lib = H.prod_lib()
    hmode = H.NONE if mode == "none" else H.MAX_ERROR
    assert host_streams[2] == H.encode(lib, frames[2], (1, 120, 200), 12.0, hmode, err)
    out_dev = torch.zeros_like(dev)
    codec.decode(dev_streams, out_dev)
    out_host = np.zeros_like(frames)
    codec.decode(host_streams, out_host)
    assert np.array_equal(out_dev.cpu().numpy().view(np.uint32), out_host.view(np.uint32))
    assert np.array_equal(out_host[3].ravel().view(np.uint32), H.decode(lib, host_streams[3]).view(np.uint32))


def test_device_multi_tile_and_constant_frames(built):
    from ebcc_b200.device_codec import EBCCDeviceCodec
    codec = EBCCDeviceCodec(8.0, "max_error_target", 0.2)
    frames = H.synth_stack(3, 3 * 40, 64).reshape(3, 3, 40, 64).copy()
    frames[1] = 4.5                                            # constant image: no codestream, value + count only
    dev = torch.from_numpy(frames).cuda()
    streams = codec.encode(dev)
    assert streams == codec.encode(frames)
    out = torch.full_like(dev, -1.0)
    codec.decode(streams, out)
    got = out.cpu().numpy()
    assert np.all(got[1] == 4.5)
    assert np.abs(got - frames).max() <= 0.2 + 1e-4
    lib = H.prod_lib()
    assert np.array_equal(got[0].ravel().view(np.uint32), H.decode(lib, streams[0]).view(np.uint32))


def test_zarr_codec_accepts_device_arrays(built):
    from ebcc_b200.zarr_filter import EBCCZarrFilter
    from ebcc_b200.filter_wrapper import EBCC_Filter
    f = H.synth_stack(1, 96, 128)[0]
    filt = EBCCZarrFilter(EBCC_Filter(base_cr=10.0, height=96, width=128, residual_opt=("max_error_target", 0.25)).hdf_filter_opts)
    a = filt.encode(f)
    b = filt.encode(torch.from_numpy(f).cuda())
    assert a == b
    out = torch.empty(96 * 128, dtype=torch.float32, device="cuda")
    filt.decode(a, out=out)
    assert np.array_equal(out.cpu().numpy().view(np.uint32), filt.decode(a).view(np.uint32))
    host_out = np.empty(96 * 128, dtype=np.float32)
    filt.decode(a, out=host_out)
    assert np.array_equal(host_out.view(np.uint32), filt.decode(a).view(np.uint32))
