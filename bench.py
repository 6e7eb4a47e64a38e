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
# (PHOTON_INTERPOLATION=cubic): 3 x (9 coordinate + 40 weights + 192 FMA x 2) + 135 = 1434, 3 x 64 taps x 4 B.
FLOP_PER_RAY_STEP = 1434.0 if CUBIC else 330.0
L1_BYTES_PER_RAY_STEP = 768.0 if CUBIC else 384.0    # linear: 3 samples x 8 texels x 16 B
# dram__bytes_read.sum + dram__bytes_write.sum of one trace launch from the ncu --set full capture under
# profiles/ (r3k_trace_raw_selected.txt: 44.7 MB read + 1261.6 MB written, linear mode, cell-table kernel);
# None where no capture of this mode exists.  The writes are L2 evictions of the kernel's local-memory
# window (1184 B of stack per thread for the out-of-line generic step and the splat footprints, 135 MB
# resident, competing for the 126 MB L2 with the field streaming through it): 31 GB/s = 0.5 % of HBM.  The
# algorithmic HBM bytes per launch are ~4.4 MB (source points + one flush of the 4 MiB image) plus the cells
# the rays of the batch walk through once (10 000 columns x 511 cells x 128 B = 0.65 GB if nothing were shared;
# neighbouring points share cells and the L2 keeps them: 45 MB read).
NCU_DRAM_BYTES_PER_LAUNCH = None if CUBIC else 1306324232


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
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop.is_set():
            try:
                o = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                    str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([t.strip() for t in o.split(",")])
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
            "config": workload_config(args, call),
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
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
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

    if rank == 0:
        fp32_peak = lib.photon_b200_measure_fp32_tflops(local)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        k_ms = float(np.mean(kernel_ms))
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
            "roofline": {"bound": "fp32_issue", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved_tf / fp32_peak if fp32_peak and fp32_peak > 0 else None,
                         "traffic": NCU_DRAM_BYTES_PER_LAUNCH,
                         "kernel": "trace_rays_kernel<BSplineField>" if CUBIC else "trace_rays_kernel<LinearFieldCells>", "kernel_ms": k_ms,
                         "flop_per_ray_step": FLOP_PER_RAY_STEP, "ray_steps_per_launch": steps_per_launch,
                         "peak_source": "FFMA loop measured live on this GPU (photon_b200_measure_fp32_tflops); "
                                        "MEASURED_PEAKS.json has no FP32 figure",
                         "l1_bytes_per_ray_step": L1_BYTES_PER_RAY_STEP,
                         "l1_gbs_achieved": steps_per_launch * L1_BYTES_PER_RAY_STEP / (k_ms * 1e-3) / 1e9,
                         "hbm_peak_gbs_measured": peaks.get("hbm_gbs"),
                         # what ncu --set full reported for this kernel and command (not measured in this run)
                         "ncu": {"issue_slots_busy": 0.672, "fma_pipe_busy": 0.565, "dram_throughput": 0.005,
                                 "l1_hit_rate": 0.883, "warp_instructions_per_ray_step": 204,
                                 "source": "profiles/r3k_trace_details.txt, r3k_trace_raw_selected.txt"} if not CUBIC else None},
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
    scene.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
