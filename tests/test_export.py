"""Image export on the device (pr03.py:2189-2243) against the numpy restatement, bit for bit."""
import numpy as np
import pytest

from oracle.postprocess import postprocess_u16
from photon_b200 import scenes


def test_restatement_basics():
    I = np.array([[0.0, -1.0, 2.0], [np.nan, 8.0, 4.0]], np.float32)
    out = postprocess_u16(I, 25.0, 10)
    assert out.dtype == np.uint16 and out.max() == 65535 and out[0, 1] == 0 and out[1, 0] == 0
    assert out[1, 2] == np.uint16(np.float32(np.round(np.float32(1023 * 0.5))) * np.float32(65535.0 / 1023.0))


@pytest.mark.gpu
@pytest.mark.parametrize("gain,bits", [(25.0, 10), (0.0, 8), (12.5, 16), (25.0, 12)])
def test_device_export_bit_exact(tmp_path, gain, bits):
    from photon_b200.scene import ResidentScene
    call = scenes.c3_bos(str(tmp_path / "b.nrrd"), n=32, dots_per_side=10, points_per_dot=6, rays_per_point=64, nx=256, ny=256)
    s = ResidentScene(call)
    s.trace()
    raw = s.download_image()
    assert raw.max() > 0
    got = s.export_u16(gain, bits)
    np.testing.assert_array_equal(got, postprocess_u16(raw, gain, bits))
    # negatives and non-finite pixels
    rng = np.random.default_rng(3)
    img = (rng.standard_normal(raw.shape) * 50).astype(np.float32)
    img[5, 7] = np.nan; img[9, 1] = np.inf; img[0, 0] = -np.inf
    s.upload_image(img)
    np.testing.assert_array_equal(s.export_u16(gain, bits), postprocess_u16(img, gain, bits))
    s.zero_image()
    assert not s.export_u16(gain, bits).any()
    # noise: statistics only (the reference seeds from the clock); raw image carries it afterwards
    s.zero_image()
    s.export_u16(gain, bits, image_noise=0.05, seed=7)
    noisy = s.download_image()
    assert abs(noisy.std() - 5.0) < 0.1 and abs(noisy.mean()) < 0.1
    s.close()
