"""photon_b200_export_u16_static -- the export entry point that allocates nothing (include/photon_b200.h; no
counterpart in the reference) -- against photon_scene_export_u16,
through the C ABI, bit for bit (the bar of tests/test_export.py), on calls in a row that share the static scratch.

The file name keeps this module last in pytest's collection order."""
import numpy as np
import pytest

from photon_b200 import scenes

MAX_PIXELS = 2048 * 2048            # PHOTON_B200_EXPORT_STATIC_MAX_PIXELS


def _static_export(lib, torch, img, gain, bits, rescale=1, noise=0.0, seed=0, stream=None, check_free=True):
    """img: float32 host image -> (uint16 image, float32 image as the call left it), via caller-owned device buffers."""
    d_img = torch.from_numpy(np.ascontiguousarray(img, np.float32)).cuda()
    d_out = torch.full((img.size,), -1, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.photon_b200_export_u16_static(d_img.data_ptr(), img.size, float(gain), int(bits), int(rescale), float(noise),
                                           int(seed), d_out.data_ptr(), stream.cuda_stream if stream is not None else None)
    assert rc == 0, lib.photon_b200_last_error().decode()
    torch.cuda.synchronize()
    if check_free:                                      # (not on a process's first launch: the driver loads kernels lazily)
        assert torch.cuda.mem_get_info()[0] >= free0    # nothing was taken from the device
    return d_out.cpu().numpy().view(np.uint16).reshape(img.shape), d_img.cpu().numpy()


@pytest.mark.gpu
def test_export_u16_static_matches_scene_export(tmp_path, cuda_lib):
    import torch
    from photon_b200.scene import ResidentScene
    lib = cuda_lib
    call = scenes.c3_bos(str(tmp_path / "b.nrrd"), n=32, dots_per_side=10, points_per_dot=6, rays_per_point=64, nx=256, ny=256)
    s = ResidentScene(call)
    s.trace()
    raw = s.download_image()
    assert raw.max() > 0
    rng = np.random.default_rng(11)
    dim = (raw * np.float32(0.125)).astype(np.float32)          # a smaller maximum than the call before: a stale one would show
    dim[5, 7] = np.nan; dim[9, 1] = -np.inf; dim[3, 3] = -4.0
    wild = (rng.standard_normal(raw.shape) * 50).astype(np.float32)
    wild[0, 0] = np.inf
    side = torch.cuda.Stream()
    # calls in a row on the shared scratch: bright, dim, random; default stream and a stream of the caller's
    for k, (img, gain, bits, stream) in enumerate([(raw, 25.0, 10, None), (dim, 25.0, 10, side), (wild, 12.5, 16, None), (raw, 0.0, 8, side)]):
        s.upload_image(img)
        want = s.export_u16(gain, bits)
        got, after = _static_export(lib, torch, img, gain, bits, stream=stream, check_free=k > 0)
        np.testing.assert_array_equal(got, want, err_msg=f"call {k}")
        np.testing.assert_array_equal(after, img)                # no noise: the input is left alone
    # without rescaling (pr03.py:2231 off): truncation only
    s.upload_image(raw)
    np.testing.assert_array_equal(_static_export(lib, torch, raw, 25.0, 10, rescale=0)[0], s.export_u16(25.0, 10, intensity_rescaling=False))
    # image noise from the same Philox key: same noisy image, same 16-bit image, twice in a row
    for seed in (7, 8):
        s.upload_image(raw)
        want = s.export_u16(25.0, 12, image_noise=0.05, seed=seed)
        want_img = s.download_image()
        got, after = _static_export(lib, torch, raw, 25.0, 12, noise=0.05, seed=seed)
        np.testing.assert_array_equal(got, want)
        np.testing.assert_array_equal(after, want_img)
        assert not np.array_equal(after, raw)
    s.close()


@pytest.mark.gpu
def test_export_u16_static_limits(cuda_lib):
    import torch
    lib = cuda_lib
    d_img = torch.zeros(16, dtype=torch.float32, device="cuda")
    d_out = torch.zeros(16, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    # beyond the static scratch: refused before anything is touched
    assert lib.photon_b200_export_u16_static(d_img.data_ptr(), MAX_PIXELS + 1, 25.0, 10, 1, 0.0, 0, d_out.data_ptr(), None) == 1
    msg = lib.photon_b200_last_error().decode()
    assert "exceed" in msg and str(MAX_PIXELS) in msg
    # bad arguments and host pointers are errors, not faults
    assert lib.photon_b200_export_u16_static(None, 16, 25.0, 10, 1, 0.0, 0, d_out.data_ptr(), None) == 1
    assert lib.photon_b200_export_u16_static(d_img.data_ptr(), 16, 25.0, 17, 1, 0.0, 0, d_out.data_ptr(), None) == 1
    host = np.zeros(16, np.float32)
    assert lib.photon_b200_export_u16_static(host.ctypes.data, 16, 25.0, 10, 1, 0.0, 0, d_out.data_ptr(), None) == 1
    assert "device memory" in lib.photon_b200_last_error().decode()
    assert lib.photon_b200_export_u16_static(d_img.data_ptr(), 0, 25.0, 10, 1, 0.0, 0, d_out.data_ptr(), None) == 0
    assert lib.photon_b200_export_u16_static(d_img.data_ptr(), 16, 25.0, 10, 1, 0.0, 0, d_out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert not d_out.any()
    # the largest image it takes (2048 x 2048) against the numpy restatement of the export
    from oracle.postprocess import postprocess_u16
    rng = np.random.default_rng(5)
    big = (rng.random((2048, 2048), dtype=np.float32) * np.float32(3.0e4)).astype(np.float32)
    got, _ = _static_export(lib, torch, big, 25.0, 10)
    np.testing.assert_array_equal(got, postprocess_u16(big, 25.0, 10))
