"""Ray-dump consumers: the host-side mirror of python_codes/light_ray_processing.py.

Two ways to the same numbers:
  * files -- ``load_light_ray_positions`` / ``load_light_ray_directions`` / ``load_light_ray_data``
    read the ``pos_%04d.bin`` / ``dir_%04d.bin`` dumps (raw float32 [n][3], prt.cu:3561-3611) exactly
    as the reference does (lrp.py:74-141), so scripts written against it keep working;
  * device -- ``dot_average_on_device`` reads the per-dot sums the trace kernel accumulated
    (``ResidentScene.enable_dot_stats``) and returns what ``calculate_dot_average`` would return for
    the dump of the same rays, without the 24 B/ray round trip through the file system.
The arithmetic on loaded dumps (deflections, dot averages) is array bookkeeping on the caller's
side of the boundary and stays numpy, with the reference's names, arguments and return types.
"""
import os

import numpy as np


def load_light_ray_positions(filename):
    """lrp.py:74-96: {'x','y','z': float32[n], 'num_rays': n} from a pos_*.bin dump."""
    pos = np.fromfile(filename, dtype="float32")
    num_rays = int(pos.size / 3)
    pos = np.reshape(pos, (num_rays, 3))
    return {"x": pos[:, 0], "y": pos[:, 1], "z": pos[:, 2], "num_rays": num_rays}


def load_light_ray_directions(filename):
    """lrp.py:121-141: direction cosines -> angles with arccos."""
    d = np.fromfile(filename, dtype="float32")
    num_rays = int(d.size / 3)
    d = np.reshape(d, (num_rays, 3))
    return {"x": np.arccos(d[:, 0]), "y": np.arccos(d[:, 1]), "z": np.arccos(d[:, 2]), "num_rays": num_rays}


def load_intermediate_light_ray_positions(filename, num_intermediate_positions):
    """lrp.py:47-71: intermediate_pos_*.bin = float32 [n][num_intermediate][3]."""
    pos = np.fromfile(filename, dtype="float32")
    num_rays = int(pos.size / (3 * num_intermediate_positions))
    pos = np.reshape(pos, (num_rays, num_intermediate_positions, 3))
    return {"x": pos[:, :, 0], "y": pos[:, :, 1], "z": pos[:, :, 2], "num_rays": num_rays}


def load_intermediate_light_ray_directions(filename, num_intermediate_positions):
    """lrp.py:98-119."""
    d = np.fromfile(filename, dtype="float32")
    num_rays = int(d.size / (3 * num_intermediate_positions))
    d = np.reshape(d, (num_rays, num_intermediate_positions, 3))
    return {"x": np.arccos(d[:, :, 0]), "y": np.arccos(d[:, :, 1]), "z": np.arccos(d[:, :, 2]), "num_rays": num_rays}


def load_light_ray_data(folder):
    """lrp.py:143-178: (pos1, pos2, dir1, dir2) of the reference (im1) and gradient (im2) images."""
    pos1 = load_light_ray_positions(os.path.join(folder, "light-ray-positions", "im1", "pos_0000.bin"))
    pos2 = load_light_ray_positions(os.path.join(folder, "light-ray-positions", "im2", "pos_0000.bin"))
    dir1 = load_light_ray_directions(os.path.join(folder, "light-ray-directions", "im1", "dir_0000.bin"))
    dir2 = load_light_ray_directions(os.path.join(folder, "light-ray-directions", "im2", "dir_0000.bin"))
    return pos1, pos2, dir1, dir2


def load_intermediate_light_ray_data(folder, num_intermediate_positions):
    """lrp.py:181-208."""
    ipos = load_intermediate_light_ray_positions(
        os.path.join(folder, "light-ray-positions", "im2", "intermediate_pos_0000.bin"), num_intermediate_positions)
    idir = load_intermediate_light_ray_directions(
        os.path.join(folder, "light-ray-directions", "im2", "intermediate_dir_0000.bin"), num_intermediate_positions)
    return ipos, idir


def calculate_lightray_deflections(pos1, pos2, dir1, dir2):
    """lrp.py:211-237: d_pos = pos1 - pos2, d_dir = dir2 - dir1, for every key of pos1 (the reference
    also differences 'num_rays'; kept)."""
    d_pos, d_dir = {}, {}
    for coord in pos1.keys():
        d_pos[coord] = pos1[coord] - pos2[coord]
        d_dir[coord] = dir2[coord] - dir1[coord]
    return d_pos, d_dir


def calculate_dot_average(a, num_rays_per_dot):
    """lrp.py:239-275: sums of consecutive segments of num_rays_per_dot rays / num_rays_per_dot."""
    num_rays_per_dot = int(num_rays_per_dot)
    coords = list(a.keys())
    num_rays = a[coords[0]].size
    a_dot = {}
    for coord in coords:
        if coord == "num_rays":
            continue
        a_dot[coord] = np.add.reduceat(a[coord], range(0, num_rays, num_rays_per_dot)) / num_rays_per_dot
    return a_dot


def convert_pos_to_pix(pos, pos0, pixel_pitch):
    """lrp.py:278-303."""
    return {c: (pos[c] - pos0[c]) / pixel_pitch for c in ("x", "y")}


def dot_average_on_device(scene, nan_policy="propagate"):
    """(pos_dot, dir_dot) from the sums the trace kernel accumulated on the GPU: equal (to float32
    summation error) to ``calculate_dot_average(load_light_ray_positions(dump), n)`` and the same of the
    directions, for the rays traced since ``scene.enable_dot_stats(n)``.

    ``nan_policy='propagate'`` reproduces numpy on the dump files: a dot with one culled ray (NaN in the
    dump) is NaN.  ``'omit'`` averages over the rays that have a value (and is NaN only for empty dots).
    Also returns the per-dot counts under the key 'count'."""
    st = scene.download_dot_stats()
    n = float(scene._rays_per_dot)
    total = scene.num_points * scene.call.lightray_number_per_particle
    full = np.minimum(n, total - np.arange(st.shape[0]) * n)        # the last segment may be short (reduceat)
    out = []
    for base, cnt in ((0, 6), (3, 7)):
        c = st[:, cnt]
        with np.errstate(invalid="ignore", divide="ignore"):
            if nan_policy == "propagate":
                vals = [np.where(c == full, st[:, base + k] / n, np.nan) for k in range(3)]
            elif nan_policy == "omit":
                vals = [np.where(c > 0, st[:, base + k] / c, np.nan) for k in range(3)]
            else:
                raise ValueError(nan_policy)
        out.append({"x": vals[0].astype(np.float32), "y": vals[1].astype(np.float32), "z": vals[2].astype(np.float32),
                    "count": c.astype(np.int64)})
    return out[0], out[1]
