"""Shared helpers for the parity tests."""
import glob
import os
import tempfile

import numpy as np

from photon_b200 import scenes


def run_with_dumps(lib, call, workdir=None):
    """Run start_ray_tracing with ray dumps on; returns (image, pos[n,3], dir[n,3]) over all batches."""
    d = workdir or tempfile.mkdtemp(prefix="photon_dump_")
    for f in glob.glob(os.path.join(d, "*.bin")):
        os.remove(f)
    call.save_lightrays = True
    call.lightray_position_save_path = d
    call.lightray_direction_save_path = d
    spn = min(call.source_point_number, call.num_particles)
    call.num_lightrays_save = spn * call.lightray_number_per_particle
    img = call.run(lib)
    pos = [np.fromfile(f, np.float32).reshape(-1, 3) for f in sorted(glob.glob(os.path.join(d, "pos_*.bin")))]
    dr = [np.fromfile(f, np.float32).reshape(-1, 3) for f in sorted(glob.glob(os.path.join(d, "dir_*.bin")))]
    call.save_lightrays = False
    call.num_lightrays_save = 0
    return img, np.concatenate(pos), np.concatenate(dr)


def compare(ref, got, pixel_pitch):
    """ref/got = (image, pos, dir).  Returns dict of parity metrics (north-star units)."""
    i0, p0, d0 = ref
    i1, p1, d1 = got
    nan_mismatch = int((np.isnan(p0[:, 0]) != np.isnan(p1[:, 0])).sum() + (np.isnan(d0[:, 0]) != np.isnan(d1[:, 0])).sum())
    both = np.isfinite(p0[:, 0]) & np.isfinite(p1[:, 0])
    dpos = float(np.abs(p0[both] - p1[both])[:, :2].max() / pixel_pitch) if both.any() else 0.0
    bd = np.isfinite(d0[:, 0]) & np.isfinite(d1[:, 0])
    ddir = float(np.abs(d0[bd] - d1[bd]).max()) if bd.any() else 0.0
    l2 = float(np.linalg.norm(i0.astype(np.float64) - i1) / max(np.linalg.norm(i0.astype(np.float64)), 1e-300))
    return {"nan_mismatch": nan_mismatch, "dpos_px": dpos, "ddir_rad": ddir, "image_rel_l2": l2,
            "hit_fraction": float(np.isfinite(p0[:, 0]).mean()), "rays": int(len(p0))}


# north-star tolerances (BASELINE.json): <=1e-3 px, <=1e-5 rad, image <=1e-4 relative L2
TOL_POS_PX = 1e-3
TOL_DIR_RAD = 1e-5
TOL_IMG_L2 = 1e-4


def parity_cases(tmpdir):
    """Small scenes the oracle finishes in about a second each; name -> (call, pixel_pitch)."""
    cases = {}
    nrrd = os.path.join(str(tmpdir), "blob64.nrrd")
    c = scenes.c3_bos(nrrd, n=64, dots_per_side=10, points_per_dot=10, rays_per_point=20)
    cases["bos_rk4_thin_diffraction"] = (c, 10.0)
    c = scenes.c3_bos(nrrd, n=64, dots_per_side=10, points_per_dot=10, rays_per_point=20, write_field=False, algorithm=1)
    cases["bos_euler_thin_diffraction"] = (c, 10.0)
    c = scenes.c3_bos(nrrd, n=64, dots_per_side=10, points_per_dot=10, rays_per_point=20, write_field=False)
    c.simulate_density_gradients = False
    cases["bos_nogradient"] = (c, 10.0)
    c = scenes.c3_bos(nrrd, n=64, dots_per_side=6, points_per_dot=4, rays_per_point=300, write_field=False,
                      source_point_number=50)
    cases["bos_rk4_multibatch_ragged"] = (c, 10.0)      # 144 points in batches of 50 (last one ragged)
    cases["piv_mie_thick_bilinear"] = (scenes.c1_piv(particles=300, rays_per_particle=200), 17.0)
    cases["piv_diffuse_thin_bilinear"] = (scenes.c1_piv(particles=300, rays_per_particle=200, mie=False,
                                                        lens_model="thin-lens"), 17.0)
    c = scenes.c1_piv(particles=300, rays_per_particle=200, lens_model="apparent")
    c.camera.diffraction_diameter = 3.0
    cases["piv_apparent"] = (c, 17.0)
    c = scenes.c1_piv(particles=300, rays_per_particle=200)
    c.camera.implement_diffraction = True
    c.camera.diffraction_diameter = 3.0
    cases["piv_mie_thick_diffraction"] = (c, 17.0)
    c = scenes.c1_piv(particles=40, rays_per_particle=1)
    cases["piv_chief_ray_only"] = (c, 17.0)             # rpp == 1 -> chief ray (prt.cu:111-116)
    return cases
