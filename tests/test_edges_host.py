"""The edge cases of tests/test_gpu_edges.py run on the CPU: the product's per-ray stage code compiled for
the host (tests/hostcheck) stands where the CUDA library stands in the GPU tests, against the same oracle.
Covers the host logic of those cases (batch loop over zero / one point, miss = unchanged, NaN protocol,
Philox noise streams) without a device."""
import pytest

from tests import test_gpu_edges as E


def test_empty_call(hostcheck_lib, tmp_path):
    E.test_empty_call_leaves_image_untouched(hostcheck_lib, tmp_path)


def test_single_ray(hostcheck_lib, oracle_lib, tmp_path):
    E.test_single_point_single_ray(hostcheck_lib, oracle_lib, tmp_path)


def test_miss_is_unchanged(hostcheck_lib, oracle_lib, tmp_path):
    E.test_rays_that_miss_the_field_are_unchanged(hostcheck_lib, oracle_lib, tmp_path)


def test_vignetted_nan(hostcheck_lib, oracle_lib, tmp_path):
    E.test_vignetted_rays_are_nan_in_both(hostcheck_lib, oracle_lib, tmp_path)


@pytest.mark.parametrize("scene", ["bos_diffraction", "piv_bilinear"])
def test_position_noise(scene, hostcheck_lib, oracle_lib, tmp_path):
    E.test_position_noise_matches_oracle_and_has_the_requested_spread(scene, hostcheck_lib, oracle_lib, tmp_path)


def test_gradient_noise(hostcheck_lib, oracle_lib, tmp_path):
    E.test_gradient_noise_matches_oracle(hostcheck_lib, oracle_lib, tmp_path)
