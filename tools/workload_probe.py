"""Resident-scene throughput of the non-C3 workloads (rays/s, CUDA-event time).
usage: python tools/workload_probe.py [piv|piv_diffraction|bos_euler|bos_nograd ...]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from photon_b200 import scenes
from photon_b200.scene import ResidentScene

KEEP = {}


def make(name, d):
    if name.startswith("c4"):                    # c4 or c4_<n>: BASELINE C4 (turbulent n^3 field, Mie, thick lens, 4 MP)
        import torch
        n = int(name.split("_")[1]) if "_" in name else 1024
        nrrd = os.path.join(d, "probe_turb_%d.nrrd" % n)
        c = scenes.c4_piv(nrrd, n=n, particles=100000, rays_per_particle=1000, write_field=not os.path.exists(nrrd),
                          device="cuda" if torch.cuda.is_available() else None)
        if name.endswith("morton"):              # c4_<n>_morton: source points in Morton order of (x, y)
            def spread(v):
                v = v.astype(np.uint64) & 0xffff
                v = (v | (v << 8)) & 0x00ff00ff; v = (v | (v << 4)) & 0x0f0f0f0f
                v = (v | (v << 2)) & 0x33333333; v = (v | (v << 1)) & 0x55555555
                return v
            qx = ((c.x - c.x.min()) / (np.ptp(c.x) + 1e-9) * 65535).astype(np.uint64)
            qy = ((c.y - c.y.min()) / (np.ptp(c.y) + 1e-9) * 65535).astype(np.uint64)
            o = np.argsort(spread(qx) | (spread(qy) << 1), kind="stable")
            c.x, c.y, c.z, c.radiance, c.diameter_index = c.x[o], c.y[o], c.z[o], c.radiance[o], c.diameter_index[o]
        return c
    if name.startswith("c5"):                    # c5 or c5_<n>: BASELINE C5 field (n x n x min(n, 1024) blob made on the GPU,
        import torch                             # handed over in memory), three optical elements; PHOTON_STEP_SCALE sweeps the step
        from photon_b200.lib import load_library
        from photon_b200.nrrd import register_field
        n = int(name.split("_")[1]) if "_" in name else 2048
        nz = min(n, 1024)
        extent = 511 * 176.0
        sx, sz = extent / (n - 1), extent / (nz - 1)
        tag = "c5_%d" % n
        if tag not in KEEP:
            dev = torch.device("cuda")
            cx = (torch.arange(n, dtype=torch.float64, device=dev) - (n - 1) / 2.0) * sx
            cz = (torch.arange(nz, dtype=torch.float64, device=dev) - (nz - 1) / 2.0) * sz
            gx = torch.exp(-cx ** 2 / (2.0 * 15e3 ** 2)).float(); gz = torch.exp(-cz ** 2 / (2.0 * 15e3 ** 2)).float()
            rho = gz[:, None, None] * gx[None, :, None]
            rho = rho * gx[None, None, :]
            rho.mul_(0.5).add_(1.225)
            torch.cuda.synchronize()
            register_field(load_library(require_gpu=True), tag, (rho.data_ptr(), (nz, n, n)), (sx, sx, sz),
                           (-extent / 2, -extent / 2, 750e3 - extent / 2))
            KEEP.clear()                          # one big volume at a time
            KEEP[tag] = rho
        return scenes.c5_stress("mem:" + tag, n=n, nz=nz, write_field=False)
    if name.startswith("piv"):
        c = scenes.c1_piv(particles=20000, rays_per_particle=10000)
        if name == "piv_diffraction":
            c.camera.implement_diffraction = True; c.camera.diffraction_diameter = 3.0
        if name == "piv_thin":
            c = scenes.c1_piv(particles=20000, rays_per_particle=10000, lens_model="thin-lens")
        if name == "piv_diffuse":
            c = scenes.c1_piv(particles=20000, rays_per_particle=10000, mie=False)
        return c
    nrrd = os.path.join(d, "probe_blob_256.nrrd")
    alg = 1 if name == "bos_euler" else 2
    c = scenes.c3_bos(nrrd, n=256, dots_per_side=100, points_per_dot=100, rays_per_point=500, algorithm=alg,
                      write_field=not os.path.exists(nrrd))
    if name == "bos_nograd":
        c.simulate_density_gradients = False
    return c

d = "/tmp/photon_bench"; os.makedirs(d, exist_ok=True)
for name in (sys.argv[1:] or ["piv", "piv_thin", "piv_diffuse", "piv_diffraction", "bos_nograd", "bos_euler"]):
    if "@" in name:                              # <workload>@<k>: integrator step = k x smallest spacing
        os.environ["PHOTON_STEP_SCALE"] = name.split("@")[1]
    else:
        os.environ.pop("PHOTON_STEP_SCALE", None)
    call = make(name.split("@")[0], d)
    s = ResidentScene(call)
    n = min(s.num_points, 10000)
    for k in range(2):
        s.trace_range(0, n)
    s.reset_counters(); s.zero_image()
    ms = []
    for k in range(3):
        s.trace_range(0, n); ms.append(s.last_trace_ms)
    rays = n * call.lightray_number_per_particle
    img = s.download_image()
    spr = s.ray_steps / max(s.rays, 1)
    print("%-16s %8.2f ms  %.3e rays/s  %.3e ray-steps/s  steps/ray %.1f  hits %.3f  image sum %.4g" % (
        name, np.mean(ms), rays / (np.mean(ms) * 1e-3), rays * spr / (np.mean(ms) * 1e-3), spr,
        s.sensor_hits / max(s.rays, 1), img.sum()), flush=True)
    s.close()
