#!/usr/bin/env python
"""bench.py -- throughput of the ray-tracing hot path on BASELINE.md config C3
(BOS, Gaussian density blob on a 512^3 NRRD grid, 1024x1024 sensor, RK4, thin lens,
diffraction splat, 1e6 source points x 1000 rays = 1e9 rays).

A *step* is one pass of the hot path over one batch of source_point_number = 10 000 source
points (1e7 rays, ~5e9 ray-steps): the reference's own batch granularity (prt.cu:3506-3559);
100 such steps make the full 1e9-ray job.  Consecutive steps walk consecutive batches.

  value : ray-steps/s, scene (sources, field, image) resident in HBM, CUDA-event timed
  e2e   : the same batches through the drop-in call start_ray_tracing with HOST buffers
          (source arrays H2D, image H2D + D2H inside the timed region; field cached)
  roofline : the trace kernel against the FP32 issue rate measured on this GPU
  cpu_baseline : the CPU oracle (or the reference's own source compiled for the host)
                 on a bounded sample of the same workload

Launch: python bench.py [--gpus N --steps K --warmup W]; for N>1 via torch.distributed.run.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CUBIC = os.environ.get("PHOTON_INTERPOLATION", "linear") == "cubic"
# RK4 + trilinear float4: SURVEY.md 8(d) / BASELINE.md 3.  RK4 + tricubic B-spline on scalar coefficients
# (PHOTON_INTERPOLATION=cubic): SURVEY 8(d)'s ~1330 FLOP, 3 x 64 taps x 4 B.
FLOP_PER_RAY_STEP = 1330.0 if CUBIC else 330.0
L1_BYTES_PER_RAY_STEP = 768.0 if CUBIC else 384.0    # linear: 3 samples x 8 texels x 16 B
# RK4 + tricubic B-spline on scalar coefficients ~1330 FLOP per ray-step (3 x ~400 + 135)
FLOP_PER_RAY_STEP_CUBIC = 1330.0
# generate ~60 + thick lens ~250 + sensor/bilinear ~20 FLOP per ray (SURVEY 8d "per ray (outside the loop)")
FLOP_PER_RAY_NO_FIELD = 330.0


def ncu_figures():
    """What `ncu --set full` reported for the timed kernel of THIS command, read from profiles/ncu_current.json
    (written by tools/ncu_summary.py from the capture kept under profiles/; it names the commit it was taken at).
    Not measured in this run: a capture under a profiler is never a bench value."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_current.json")))
    except Exception:
        return None


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--grid", type=int, default=512, help="field is grid^3 voxels (C3: 512)")
    ap.add_argument("--rays-per-point", type=int, default=1000)
    ap.add_argument("--points-per-step", type=int, default=10000)
    ap.add_argument("--sensor", type=int, default=1024)
    ap.add_argument("--cpu-sample-points", type=int, default=0, help="0 = auto (about 10-30 s of CPU work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity leg of the cpu_baseline section")
    ap.add_argument("--workloads", default="auto", help="comma list of c1_piv,c4_1024,c5,cubic,c3_euler,cubic_brick,c5_2048 | none | auto (all at N = 1, sharded C4/C5 at N = 8); N > 1: c5_full = the whole 1e10-ray C5 job")
    ap.add_argument("--no-in-call", action="store_true", help="N > 1: skip the PHOTON_GPUS leg")
    ap.add_argument("--workdir", default=None)
    return ap.parse_args()


def build_workload(args, workdir, rank):
    from photon_b200 import scenes
    nrrd = os.path.join(workdir, f"c3_blob_{args.grid}.nrrd")
    write = rank == 0 and not os.path.exists(nrrd)
    call = scenes.c3_bos(nrrd, n=args.grid, dots_per_side=100, points_per_dot=100,
                         rays_per_point=args.rays_per_point, nx=args.sensor, ny=args.sensor,
                         write_field=write, source_point_number=args.points_per_step)
    return call, nrrd


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        # index: one GPU ordinal, or "all" (rows then carry every GPU of the box, one line each)
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop.is_set():
            try:
                sel = [] if self.index == "all" else ["-i", str(self.index)]
                o = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"] + sel,
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                for ln in o.splitlines():
                    if ln.strip():
                        self.rows.append([t.strip() for t in ln.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_reference_lib():
    """(lib, kind): the reference's own source compiled for the host if present, else the oracle port."""
    from oracle.loader import load_oracle, load_ref_cpu
    # all host cores (torchrun exports OMP_NUM_THREADS=1 to its workers; the OpenMP runtime reads the
    # variable when the library is loaded)
    os.environ["OMP_NUM_THREADS"] = os.environ.get("PHOTON_BENCH_THREADS") or str(os.cpu_count() or 1)
    ref = load_ref_cpu()
    if ref is not None and not os.environ.get("PHOTON_BENCH_FORCE_PORT"):
        return ref, "reference"
    return load_oracle(), "port"


def time_cpu(call, npoints, lib, kind):
    """Trace the first npoints source points of the workload on the host cores.
    kind "port": the oracle restatement (counts committed steps itself).
    kind "reference": the reference's own sources compiled for the host (oracle/_ref); its kernel time
    is the time inside its parallel_ray_tracing launches and its ray-steps are tex3D fetches // 3
    (RK4 samples the field 3 times per committed step; the one discarded step per ray adds < 0.3 %)."""
    import ctypes
    sub = call.subset(slice(0, npoints))
    if kind == "reference":
        lib.photon_ref_last_trace_seconds.restype = ctypes.c_double
        lib.photon_ref_last_tex_fetches.restype = ctypes.c_longlong
        lib.photon_ref_reset_counters()
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    sys.stdout.flush()
    os.dup2(devnull, 1)                       # the reference prints progress lines on stdout
    try:
        t0 = time.perf_counter()
        sub.run(lib)
        wall = time.perf_counter() - t0
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    if kind == "port":
        secs = lib.photon_oracle_last_trace_seconds()
        steps = lib.photon_oracle_last_ray_steps()
        threads = lib.photon_oracle_last_threads()
    else:
        secs = lib.photon_ref_last_trace_seconds()
        steps = lib.photon_ref_last_tex_fetches() // 3
        threads = lib.photon_ref_last_threads()
    return {"seconds": secs, "wall": wall, "steps": int(steps), "rays": int(npoints * call.lightray_number_per_particle),
            "threads": int(threads)}


def parity_at_full_size(call, lib, cpu_lib, kind, workdir, npoints=96):
    """Parity AT THE BENCHMARK'S OWN CONFIGURATION (512^3 field, 16 GiB cell table): `npoints` source points spread over
    the whole dot grid, every ray of them, through the reference on the host cores (ray dumps on) and through the CUDA
    library -- its dump kernel for the per-ray figures and the non-dump kernel (the one timed above) for the image."""
    from tests import util
    stride = max(1, call.num_particles // npoints)
    idx = np.arange(0, call.num_particles, stride)[:npoints]
    sub = call.subset(idx)
    sub.source_point_number = len(idx)
    dirs = [os.path.join(workdir, "parity_" + t) for t in ("ref", "gpu")]
    for d in dirs:
        os.makedirs(d, exist_ok=True)
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)
    try:
        ref = util.run_with_dumps(cpu_lib, sub, dirs[0])
    finally:
        os.dup2(saved, 1); os.close(devnull); os.close(saved)
    got = util.run_with_dumps(lib, sub, dirs[1])
    v_dump = lib.photon_b200_last_variant().decode()
    plain = sub.run(lib)
    v_plain = lib.photon_b200_last_variant().decode()
    m = util.compare(ref, got, float(call.camera.pixel_pitch))
    i0 = ref[0].astype(np.float64)
    return {"rays": m["rays"], "dpos_px": m["dpos_px"], "dpos_px_p99": m["dpos_px_p99"], "ddir_rad": m["ddir_rad"],
            "image_rel_l2": m["image_rel_l2"],
            "image_rel_l2_nondump": float(np.linalg.norm(i0 - plain) / max(np.linalg.norm(i0), 1e-300)),
            "nan_mismatch": m["nan_mismatch"], "hit_fraction": m["hit_fraction"],
            "kernels": [v_dump, v_plain], "against": "reference (host build)" if kind == "reference" else "oracle port",
            "sample": f"{len(idx)} source points (every {stride}th of the workload) x {call.lightray_number_per_particle} rays",
            "bounds": {"dpos_px": 1e-3, "ddir_rad": 1e-5, "image_rel_l2": 1e-4}}


def run_reference_arm(args):
    """--impl reference: the CPU implementation of the path on the host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    workdir = args.workdir or tempfile.mkdtemp(prefix="photon_bench_")
    call, _ = build_workload(args, workdir, 0)
    lib, kind = cpu_reference_lib()
    pilot = time_cpu(call, 16, lib, kind)         # sizes the per-step sample: about 6 s of tracing per step
    npts = args.cpu_sample_points or int(max(16, min(2000, 6.0 * 16 / max(pilot["seconds"], 1e-3))))
    for _ in range(args.warmup):
        time_cpu(call, 8, lib, kind)
    secs, steps, rays, threads = 0.0, 0, 0, 1
    for k in range(args.steps):
        sub = call.subset(slice(k * npts, (k + 1) * npts))
        r = time_cpu(sub, npts, lib, kind)
        secs += r["seconds"]; steps += r["steps"]; rays += r["rays"]; threads = r["threads"]
    v = steps / secs
    line = {"impl": "reference", "metric": "ray_steps_per_s", "value": v, "unit": "ray-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "rays_per_s": rays / secs,
            "config": dict(workload_config(args, call), rays_per_step=npts * args.rays_per_point, points_per_step=npts,
                           note="a step of this arm is a bounded sample of the workload: %d source points x %d rays" % (npts, args.rays_per_point)),
            "cpu_baseline": {"value": v, "unit": "ray-steps/s", "cores": threads, "kind": kind,
                             "sample": f"{npts} source points x {args.rays_per_point} rays per step ({rays} rays total), "
                                       "time inside the tracing kernel launches; the per-call NRRD load + gradient precompute "
                                       "(about 10 s at 512^3, single-threaded in the reference) is excluded"},
            "e2e": {"value": v, "unit": "ray-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def workload_config(args, call):
    return {"workload": f"C3 BOS Gaussian blob {args.grid}^3 NRRD, {args.sensor}x{args.sensor} sensor, RK4 + {'tricubic B-spline (scalar coefficients)' if CUBIC else 'trilinear'}, "
                        f"thin lens f/11, diffraction 3 px, {call.num_particles} source points x {args.rays_per_point} rays",
            "rays_per_step": args.points_per_step * args.rays_per_point, "points_per_step": args.points_per_step,
            "field_bytes": 16 * args.grid ** 3, "coefficient_bytes": 4 * args.grid ** 3 if CUBIC else 0,
            # per-cell polynomial table of the coherent hot loop (8 float4 per cell; PHOTON_CELLS=0 reads the texels instead)
            "cell_table_bytes": 0 if (CUBIC or os.environ.get("PHOTON_CELLS") == "0") else 128 * args.grid ** 3,
            "interpolation": "cubic" if CUBIC else "linear", "rng": os.environ.get("PHOTON_RNG", "reference"),
            "sampler": os.environ.get("PHOTON_SAMPLER", "linear"),
            "l2_policy": "field (2 GiB of texels + 16 GiB cell table at 512^3) is larger than the 126 MB L2; consecutive steps trace different batches"}


# ---------------------------------------------------------------------------------------------------------------------
# The other BASELINE workloads, resident-scene throughput (CUDA-event time of the trace kernel), bounded to a few
# seconds each: C1-like PIV without a field, C4 (PIV through a turbulent 1024^3 field), C5 (2048 x 2048 x 1024 blob,
# three optical elements, step-size sweep), cubic mode on C3.  Fields are made on the GPU and handed over in memory.
# ---------------------------------------------------------------------------------------------------------------------
def _time_resident(scene, npoints, reps=3, warm=2):
    n = min(scene.num_points, npoints)
    for _ in range(warm):
        scene.trace_range(0, n)
    scene.reset_counters(); scene.zero_image()
    ms = []
    for _ in range(reps):
        scene.trace_range(0, n); ms.append(scene.last_trace_ms)
    return n, float(np.mean(ms)), scene.ray_steps / reps, scene.rays / reps


def _entry(lib, ms, steps, rays, fp32_peak, flop_per_step=330.0, **extra):
    sec = ms * 1e-3
    e = {"ms": ms, "rays_per_s": rays / sec, "kernel": lib.photon_b200_last_variant().decode()}
    if steps > 0:
        e.update(value=steps / sec, unit="ray-steps/s", steps_per_ray=steps / max(rays, 1),
                 roofline_frac=(steps * flop_per_step / sec / 1e12 / fp32_peak) if fp32_peak else None, flop_per_ray_step=flop_per_step)
    else:
        e.update(value=rays / sec, unit="rays/s",
                 roofline_frac=(rays * FLOP_PER_RAY_NO_FIELD / sec / 1e12 / fp32_peak) if fp32_peak else None, flop_per_ray=FLOP_PER_RAY_NO_FIELD)
    e.update(extra)
    return e


def _c5_field(lib, n=2048, nz=1024):
    """BASELINE C5's blob on an n x n x nz grid, made on the GPU and registered as "mem:c5" (16 GiB at 2048 x 2048 x 1024)."""
    import torch
    from photon_b200.nrrd import register_field
    extent = 511 * 176.0
    sx, sz = extent / (n - 1), extent / (nz - 1)
    dev = torch.device("cuda")
    cx = (torch.arange(n, dtype=torch.float64, device=dev) - (n - 1) / 2.0) * sx
    cz = (torch.arange(nz, dtype=torch.float64, device=dev) - (nz - 1) / 2.0) * sz
    gx = torch.exp(-cx ** 2 / (2.0 * 15e3 ** 2)).float(); gz = torch.exp(-cz ** 2 / (2.0 * 15e3 ** 2)).float()
    rho = gz[:, None, None] * gx[None, :, None]
    rho = rho * gx[None, None, :]
    rho.mul_(0.5).add_(1.225)
    torch.cuda.synchronize()
    register_field(lib, "c5", (rho.data_ptr(), (nz, n, n)), (sx, sx, sz), (-extent / 2, -extent / 2, 750e3 - extent / 2))
    return rho


def run_workloads(lib, args, c3_call, fp32_peak, which):
    import torch
    from photon_b200 import scenes
    from photon_b200.scene import ResidentScene
    out = {}

    def drop():
        lib.photon_b200_release_cache()
        torch.cuda.empty_cache()

    for name in which:
        t0 = time.perf_counter()
        try:
            if name == "c1_piv":
                c = scenes.c1_piv(particles=20000, rays_per_particle=10000)
                sc = ResidentScene(c)
                n, ms, steps, rays = _time_resident(sc, 10000)
                out[name] = _entry(lib, ms, 0, rays, fp32_peak, workload="PIV, no density field: 1e4 particles x 1e4 rays per launch, Mie, thick lens, bilinear splat, 1024^2",
                                   bound="instruction issue (per-ray generate / lens / splat arithmetic; the working set is the 4 MiB image, L2-resident by nature)")
                sc.close()
            elif name.startswith("c4_"):
                n3 = int(name.split("_")[1])
                c = scenes.c4_piv("mem:c4", n=n3, particles=100000, rays_per_particle=1000, device="cuda", lib=lib)
                torch.cuda.empty_cache()
                sc = ResidentScene(c)
                n, ms, steps, rays = _time_resident(sc, 10000)
                out[name] = _entry(lib, ms, steps, rays, fp32_peak, field_bytes=16 * n3 ** 3,
                                   workload=f"C4: PIV through a turbulent {n3}^3 field, Mie (bhmie table), thick lens, 2048^2 sensor, bilinear splat; 1e4 particles x 1000 rays per launch")
                sc.close(); lib.photon_b200_unregister_field(b"c4"); del c
            elif name == "c5":
                keep = _c5_field(lib)
                c = scenes.c5_stress("mem:c5", n=2048, nz=1024, write_field=False)
                for scale in ("1", "0.5", "0.25"):
                    os.environ["PHOTON_STEP_SCALE"] = scale
                    try:
                        sc = ResidentScene(c)
                        n, ms, steps, rays = _time_resident(sc, 10000, reps=2, warm=1)
                        out["c5_step_" + scale] = _entry(lib, ms, steps, rays, fp32_peak, field_bytes=16 * 2048 * 2048 * 1024,
                                                         workload="C5: BOS through a 2048 x 2048 x 1024 blob (the reference keeps 1024 planes), three optical elements, "
                                                                  f"step = {scale} x spacing; 1e4 points x 1000 rays per launch")
                        sc.close()
                    finally:
                        del os.environ["PHOTON_STEP_SCALE"]
                lib.photon_b200_unregister_field(b"c5"); del keep, c
            elif name == "c5_2048":
                # Optional (not in "auto"): C5 at its stated depth -- 2048^3 voxels, 32 GiB of density + 128 GiB of field
                # on one GPU, past the reference's 1024-plane cut (PHOTON_FULL_DEPTH) and past 32-bit voxel indices.
                os.environ["PHOTON_FULL_DEPTH"] = "1"
                try:
                    keep = _c5_field(lib, n=2048, nz=2048)
                    c = scenes.c5_stress("mem:c5", n=2048, nz=2048, write_field=False)
                    sc = ResidentScene(c)
                    del keep; torch.cuda.empty_cache()      # the field is built; the density is no longer read
                    n, ms, steps, rays = _time_resident(sc, 10000, reps=2, warm=1)
                    out[name] = _entry(lib, ms, steps, rays, fp32_peak, field_bytes=16 * 2048 ** 3,
                                       workload="C5 at full depth: BOS through a 2048^3 blob (PHOTON_FULL_DEPTH=1, 64-bit voxel index), "
                                                "three optical elements; 1e4 points x 1000 rays per launch")
                    sc.close()
                    lib.photon_b200_unregister_field(b"c5"); del c
                finally:
                    del os.environ["PHOTON_FULL_DEPTH"]
            elif name == "c3_euler":
                # C3's scene with the Euler integrator (algorithm 1: what the reference's shipped BOS sample runs); one sample
                # per step, SURVEY 8(d): ~100 FLOP and 128 B of texels per ray-step
                import copy
                ce = copy.copy(c3_call); ce.ray_tracing_algorithm = 1
                sc = ResidentScene(ce)
                n, ms, steps, rays = _time_resident(sc, args.points_per_step, reps=2, warm=1)
                out[name] = _entry(lib, ms, steps, rays, fp32_peak, flop_per_step=100.0,
                                   workload=f"C3 with the Euler integrator, {args.grid}^3 field; bound by the L1 data stage (8 LDG.128 per step), not by FP32")
                sc.close()
            elif name in ("cubic", "cubic_brick"):
                os.environ["PHOTON_INTERPOLATION"] = "cubic"
                if name == "cubic_brick":
                    os.environ["PHOTON_CUBIC_BRICK"] = "1"       # the same loop on TMA-staged bricks (A/B, not in "auto")
                try:
                    sc = ResidentScene(c3_call)
                    n, ms, steps, rays = _time_resident(sc, args.points_per_step, reps=2, warm=1)
                    out[name] = _entry(lib, ms, steps, rays, fp32_peak, flop_per_step=FLOP_PER_RAY_STEP_CUBIC,
                                       workload=f"C3 with the tricubic B-spline sampler (PHOTON_INTERPOLATION=cubic), {args.grid}^3 coefficients"
                                                + (", hot loop on TMA-staged bricks (PHOTON_CUBIC_BRICK=1)" if name == "cubic_brick" else ""))
                    if name == "cubic_brick":
                        out[name]["brick_timeouts"] = lib.photon_b200_brick_timeouts()
                    sc.close()
                finally:
                    del os.environ["PHOTON_INTERPOLATION"]
                    os.environ.pop("PHOTON_CUBIC_BRICK", None)
        except Exception as e:
            out[name] = {"error": repr(e)[:300]}
        drop()
        if name in out:
            out[name]["wall_s_including_setup"] = time.perf_counter() - t0
    return out


def run_workloads_sharded(lib, args, rank, world, local, fp32_peak, which="auto"):
    """N > 1: BASELINE configs 4 and 5 over all ranks.  C4 AT ITS STATED SIZE (1e6 particles x 1000 rays x 2 frames
    through the 1024^3 field, 4 MP): every rank makes the same field on its GPU and traces its shard of the particles;
    partial images are summed with one NCCL reduce per frame.  C5: a slice of the 1e7-point list (1e5 points x 1000
    rays per rank and step size) through the 2048 x 2048 x 1024 field.  Times are CUDA-event kernel times, max over ranks."""
    import torch
    import torch.distributed as dist
    from photon_b200 import scenes
    from photon_b200.scene import ResidentScene
    out = {}

    def reduce_max_sum(ms, steps, rays):
        t = torch.tensor([ms, steps, rays], dtype=torch.float64, device="cuda")
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        return float(mx[0]), float(sm[1]), float(sm[2])
    if "c5_full" in which.split(","):
        # Optional (not in "auto"): BASELINE config 5 AS STATED -- all 1e7 source points x 1000 rays = 1e10 rays through the
        # 2048 x 2048 x 1024 field (replicated per GPU), three optical elements, step = spacing; each rank traces 1e7 / N points
        # in ONE launch.  About 22 s of kernel time on 8 B200.
        try:
            keep = _c5_field(lib)
            c = scenes.c5_stress("mem:c5", n=2048, nz=1024, write_field=False, dots_per_side=316, points_per_dot=100)
            total = len(c.x)
            per_rank = (total + world - 1) // world
            sub = c.subset(slice(rank * per_rank, min(total, (rank + 1) * per_rank)))
            sc = ResidentScene(sub, device=local)
            del keep; torch.cuda.empty_cache()
            sc.reset_counters()
            torch.cuda.synchronize(); dist.barrier()
            t0 = time.perf_counter()
            sc.trace(); ms = sc.last_trace_ms
            img = torch.from_numpy(sc.download_image().reshape(-1)).cuda()
            dist.reduce(img, dst=0)
            torch.cuda.synchronize(); dist.barrier()
            wall = time.perf_counter() - t0
            m, st, ry = reduce_max_sum(ms, sc.ray_steps, sc.rays)
            out["c5_full_step_1"] = {"value": st / (m * 1e-3), "unit": "ray-steps/s", "rays": ry, "rays_per_s": ry / (m * 1e-3), "kernel_seconds": m * 1e-3,
                                     "wall_seconds_with_reduce": wall, "gpus": world, "kernel": lib.photon_b200_last_variant().decode(),
                                     "roofline_frac_per_gpu": st * 330.0 / (m * 1e-3) / 1e12 / fp32_peak / world if fp32_peak else None,
                                     "image_sum": float(img.double().sum()) if rank == 0 else None,
                                     "workload": f"BASELINE config 5 as stated: {total} source points x 1000 rays, 2048 x 2048 x 1024 field replicated per GPU, "
                                                 "three optical elements, step = spacing, one launch per GPU"}
            sc.close()
            lib.photon_b200_unregister_field(b"c5"); del c
        except Exception as e:
            out["c5_full_step_1"] = {"error": repr(e)[:300]}
        lib.photon_b200_release_cache(); torch.cuda.empty_cache()
        return out
    try:
        c = scenes.c4_piv("mem:c4", n=1024, particles=1000000, rays_per_particle=1000, device="cuda", lib=lib)
        torch.cuda.empty_cache()
        ny, nx = c.image_shape
        tot_ms, tot_steps, tot_rays = 0.0, 0.0, 0.0
        t0 = time.perf_counter()
        for frame in range(2):
            if frame == 1:
                c.x = (c.x + np.float32(17.0 / 0.176)).astype(np.float32)         # about one pixel on the sensor
            sc = ResidentScene(c, shard=(rank, world), device=local)
            sc.reset_counters()
            sc.trace(); ms = sc.last_trace_ms
            img = torch.from_numpy(sc.download_image().reshape(-1)).cuda()
            dist.reduce(img, dst=0)
            m, st, ry = reduce_max_sum(ms, sc.ray_steps, sc.rays)
            tot_ms += m; tot_steps += st; tot_rays += ry
            sc.close()
        torch.cuda.synchronize(); dist.barrier()
        wall = time.perf_counter() - t0
        out["c4_full"] = {"value": tot_steps / (tot_ms * 1e-3), "unit": "ray-steps/s", "rays": tot_rays, "rays_per_s": tot_rays / (tot_ms * 1e-3),
                          "kernel_seconds": tot_ms * 1e-3, "wall_seconds_with_upload_and_reduce": wall, "gpus": world,
                          "kernel": lib.photon_b200_last_variant().decode(),
                          "roofline_frac_per_gpu": tot_steps * 330.0 / (tot_ms * 1e-3) / 1e12 / fp32_peak / world if fp32_peak else None,
                          "workload": "BASELINE config 4 at its stated size: 1e6 Mie particles x 1000 rays x 2 frames, turbulent 1024^3 field replicated per GPU, 2048^2 sensor"}
        lib.photon_b200_unregister_field(b"c4"); del c
    except Exception as e:
        out["c4_full"] = {"error": repr(e)[:300]}
    lib.photon_b200_release_cache(); torch.cuda.empty_cache()
    try:
        keep = _c5_field(lib)
        c = scenes.c5_stress("mem:c5", n=2048, nz=1024, write_field=False, dots_per_side=316, points_per_dot=100)
        per_rank = 100000
        sub = c.subset(slice(rank * per_rank, (rank + 1) * per_rank))
        for scale in ("1", "0.5", "0.25"):
            os.environ["PHOTON_STEP_SCALE"] = scale
            try:
                sc = ResidentScene(sub, device=local)
                sc.reset_counters()
                sc.trace(); ms = sc.last_trace_ms
                m, st, ry = reduce_max_sum(ms, sc.ray_steps, sc.rays)
                out["c5_step_" + scale] = {"value": st / (m * 1e-3), "unit": "ray-steps/s", "rays": ry, "rays_per_s": ry / (m * 1e-3), "kernel_seconds": m * 1e-3,
                                           "gpus": world, "kernel": lib.photon_b200_last_variant().decode(),
                                           "full_job_seconds_extrapolated": 1e10 / (ry / (m * 1e-3)),
                                           "workload": f"BASELINE config 5, a slice: {per_rank} of the 1e7 source points x 1000 rays per GPU, 2048 x 2048 x 1024 field replicated per GPU, "
                                                       f"three optical elements, step = {scale} x spacing"}
                sc.close()
            finally:
                del os.environ["PHOTON_STEP_SCALE"]
        lib.photon_b200_unregister_field(b"c5"); del keep, c
    except Exception as e:
        out["c5"] = {"error": repr(e)[:300]}
    lib.photon_b200_release_cache(); torch.cuda.empty_cache()
    return out


def run_in_call(lib, call, world):
    """The library's OWN multi-GPU path: rank 0 alone makes ONE start_ray_tracing call over the whole C3 job
    with PHOTON_GPUS = N -- the library splits the source points over the N devices of the box, builds the field on each,
    and combines the partial images onto device 0 with its own kernel over NVLink peer memory (PHOTON_REDUCE=nccl: one ncclReduce).  Cold (nothing cached: field read + built on
    every device, communicator set-up) then warm (second image of a BOS pair), then the same call on one GPU."""
    rays = call.num_particles * call.lightray_number_per_particle
    lib.photon_b200_release_cache()
    res = {"gpus": world, "rays": rays}
    os.environ.pop("PHOTON_SHARD", None)
    os.environ.pop("PHOTON_DEVICE", None)
    os.environ["PHOTON_GPUS"] = str(world)
    os.environ["PHOTON_TIMING"] = "1"                 # the library's own per-stage laps go to stderr (kept by the driver's log)
    try:
        t0 = time.perf_counter(); img = call.run(lib); res["cold_s"] = time.perf_counter() - t0
        with ClockSampler("all") as ck:               # a multi-second kernel on every GPU at once: are the clocks held?
            t0 = time.perf_counter(); img = call.run(lib); res["warm_s"] = time.perf_counter() - t0
        res["clocks_warm"] = ck.summary()
        os.environ["PHOTON_GPUS"] = "1"
        t0 = time.perf_counter(); one = call.run(lib); res["one_gpu_warm_s"] = time.perf_counter() - t0
    finally:
        del os.environ["PHOTON_GPUS"]
        os.environ.pop("PHOTON_TIMING", None)
    res["speedup_vs_1"] = res["one_gpu_warm_s"] / res["warm_s"]
    res["rays_per_s_warm"] = rays / res["warm_s"]
    i1 = one.astype(np.float64)
    res["image_rel_l2_vs_sharded"] = float(np.linalg.norm(img - i1) / np.linalg.norm(i1))
    res["note"] = "one start_ray_tracing call with host buffers per figure; NRRD file in the page cache; cold = first multi-GPU call of the process (contexts, field read + builds on every device); images combined by the library's peer-memory kernel"
    return res


_JSON_FD = None


def claim_stdout():
    """The driver reads ONE JSON line from stdout.  Libraries write there too (NCCL prints its version line on
    stdout, the host-compiled reference prints progress): send file descriptor 1 to stderr for the whole run and
    keep the real stdout for emit()."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, data)


def main():
    args = parse()
    claim_stdout()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    from photon_b200.scene import ResidentScene
    from photon_b200.lib import load_library

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # a CPU-side group for the waits around the in-call leg: an NCCL barrier is a kernel that SPINS on the waiting
        # rank's GPU, and the library's in-call path runs its shards on those same GPUs from rank 0's process --
        # two processes on one GPU are time-sliced, which halved the in-call rate in the first run (profiles r4f)
        cpu_group = dist.new_group(backend="gloo")
    workdir = args.workdir or os.path.join(tempfile.gettempdir(), "photon_bench")
    os.makedirs(workdir, exist_ok=True)
    if rank == 0:
        call, nrrd = build_workload(args, workdir, 0)
    if world > 1:
        dist.barrier()
    if rank != 0:
        call, nrrd = build_workload(args, workdir, rank)

    lib = load_library(require_gpu=True)
    # weak scaling: every rank traces points_per_step points per step from its own shard of the point list
    scene = ResidentScene(call, shard=(rank, world), device=local)
    npts = scene.num_points
    pps = min(args.points_per_step, npts)
    nbatches = max(1, npts // pps)
    ny, nx = call.image_shape
    img_t = torch.zeros(ny * nx, dtype=torch.float32, device="cuda")

    class _Ptr:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}
    dev_image = torch.as_tensor(_Ptr(scene.image_device_ptr, ny * nx), device="cuda")

    def step(k):
        scene.trace_range((k % nbatches) * pps, pps, sync=False)

    def barrier():
        scene.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for k in range(args.warmup):
        step(k)
    barrier()
    scene.reset_counters()
    scene.zero_image()
    kernel_ms = []
    with ClockSampler(local) as clocks:
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            step(args.warmup + k)
            scene.sync()
            kernel_ms.append(scene.last_trace_ms)
        if world > 1:
            img_t.copy_(dev_image)
            dist.reduce(img_t, dst=0)       # partial sensor images -> rank 0 (one 4 MiB NCCL reduce)
        barrier()
        wall = time.perf_counter() - t0
    dev_ms = float(sum(kernel_ms))
    steps_done = scene.ray_steps
    rays_done = scene.rays
    launches = args.steps
    t = torch.tensor([dev_ms, wall * 1e3, float(steps_done), float(rays_done)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms_max, wall_ms_max = float(tmax[0]), float(tmax[1])
        steps_all, rays_all = float(tsum[2]), float(tsum[3])
    else:
        dev_ms_max, wall_ms_max, steps_all, rays_all = dev_ms, wall * 1e3, float(steps_done), float(rays_done)
    # whole-job time = max over ranks of the device time, plus the reduce (wall of the bracketed region)
    job_ms = max(dev_ms_max, 0.0) if world == 1 else wall_ms_max
    value = steps_all / (job_ms * 1e-3)

    # ---- e2e through the drop-in call with host buffers ---------------------------------
    e2e = None
    if not args.no_e2e:
        os.environ["PHOTON_SHARD"] = f"{rank}/{world}"
        os.environ["PHOTON_DEVICE"] = str(local)
        subs = [call.subset(slice(b * pps * world, (b + 1) * pps * world)) for b in range(min(nbatches, args.steps + 1))]
        img = np.zeros(ny * nx, np.float32)
        subs[0].run(lib, img)               # warm-up (field already cached by the resident scene)
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            out = subs[(k + 1) % len(subs)].run(lib, img)
        if world > 1:
            tt = torch.from_numpy(out.reshape(-1)).cuda(); dist.reduce(tt, dst=0)
        barrier()
        e2e_s = time.perf_counter() - t0
        te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        steps_per_step = steps_all / args.steps
        h2d = pps * 24 + ny * nx * 4 + call.lightray_number_per_particle * 8
        e2e = {"value": steps_per_step * args.steps / float(te[0]), "unit": "ray-steps/s",
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(ny * nx * 4),
               "note": "start_ray_tracing with host numpy buffers; NRRD field cached in HBM across calls"}
        del os.environ["PHOTON_SHARD"]

    # ---- the other workloads and the library's in-call multi-GPU path (after the headline legs) ------------------------
    fp32_peak = lib.photon_b200_measure_fp32_tflops(local)
    workloads, in_call = None, None
    wl = args.workloads
    if wl != "none":
        if world == 1:
            which = ["c1_piv", "c4_1024", "c5", "cubic", "c3_euler"] if wl == "auto" else [w for w in wl.split(",") if w]
            scene.close(); scene = None
            lib.photon_b200_release_cache()
            workloads = run_workloads(lib, args, call, fp32_peak, which)
        elif world == 8 or wl != "auto":
            scene.close(); scene = None
            lib.photon_b200_release_cache()
            workloads = run_workloads_sharded(lib, args, rank, world, local, fp32_peak, wl)
    if world > 1 and not args.no_in_call:
        if scene is not None:
            scene.close(); scene = None
        lib.photon_b200_release_cache()
        torch.cuda.empty_cache()
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)                   # every GPU idle, nobody spinning on one
        if rank == 0:
            try:
                in_call = run_in_call(lib, call, world)
            except Exception as e:
                in_call = {"error": repr(e)[:300]}
        dist.barrier(group=cpu_group)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        k_ms = float(np.mean(kernel_ms))
        ncu = ncu_figures()
        steps_per_launch = steps_done / max(args.steps, 1)
        achieved_tf = steps_per_launch * FLOP_PER_RAY_STEP / (k_ms * 1e-3) / 1e12
        line = {
            "metric": "ray_steps_per_s", "value": value, "unit": "ray-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": job_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "rays_per_s": rays_all / (job_ms * 1e-3), "steps_per_ray": steps_all / max(rays_all, 1.0),
            "config": workload_config(args, call),
            "gpu_launches": launches * world,
            "clocks": clocks.summary(),
            "e2e": e2e,
            "workloads": workloads, "in_call": in_call,
            "roofline": {"bound": "fp32_issue", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved_tf / fp32_peak if fp32_peak and fp32_peak > 0 else None,
                         "traffic": (ncu or {}).get("dram_bytes_per_launch") if not CUBIC else None,
                         "kernel": "trace_rays_kernel<BSplineField>" if CUBIC else "trace_rays_kernel<LinearFieldCells>", "kernel_ms": k_ms,
                         "flop_per_ray_step": FLOP_PER_RAY_STEP, "ray_steps_per_launch": steps_per_launch,
                         "peak_source": "FFMA loop measured live on this GPU (photon_b200_measure_fp32_tflops); "
                                        "MEASURED_PEAKS.json has no FP32 figure",
                         "l1_bytes_per_ray_step": L1_BYTES_PER_RAY_STEP,
                         "l1_gbs_achieved": steps_per_launch * L1_BYTES_PER_RAY_STEP / (k_ms * 1e-3) / 1e9,
                         "hbm_peak_gbs_measured": peaks.get("hbm_gbs"),
                         # what ncu --set full reported for this kernel and command (profiles/ncu_current.json; not measured in this run)
                         "ncu": ncu if not CUBIC else None},
        }
        if not args.no_cpu_baseline and world == 1:
            cl, kind = cpu_reference_lib()
            pilot = time_cpu(call, 16, cl, kind)    # sizes the sample: about 12 s of tracing on the host cores
            n = args.cpu_sample_points or int(max(16, min(4000, 12.0 * 16 / max(pilot["seconds"], 1e-3))))
            r = time_cpu(call, n, cl, kind)
            line["cpu_baseline"] = {"value": r["steps"] / r["seconds"], "unit": "ray-steps/s", "cores": r["threads"],
                                    "kind": kind, "rays_per_s": r["rays"] / r["seconds"],
                                    "sample": f"first {n} source points x {args.rays_per_point} rays of the same workload "
                                              f"({r['rays']} rays, {r['steps']} ray-steps, {r['seconds']:.1f} s); field load excluded"}
            if not args.no_parity:
                try:
                    line["parity"] = parity_at_full_size(call, lib, cl, kind, workdir)
                except Exception as e:            # the headline numbers stand on their own; say why parity is absent
                    line["parity"] = {"error": repr(e)[:300]}
        emit(line)
    if scene is not None:
        scene.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
