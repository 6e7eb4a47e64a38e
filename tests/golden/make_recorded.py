#!/usr/bin/env python
"""Generates tests/golden/recorded_{bos,piv,bos_rotated}.npz: the EXACT start_ray_tracing inputs of BASELINE
configs C1 (sample_run_script.sh piv) and C2 (sample_run_script.sh bos), and of the shipped parameter file whose
rays cross the sample volume (sample-data/bos/sample-parameters.mat: camera turned 150 degrees about y, Euler,
SURVEY 3.2 / App. B-11), recorded from the reference's unchanged Python scripts, together with the images the
REFERENCE'S OWN SOURCES (oracle/_ref/libphoton_ref_cpu.so, host build of cuda_codes/*.cu) produced for them.

  python tests/golden/make_recorded.py [--work /tmp/photon_recorded] [--reuse]

Needs /root/reference (build container only).  What is stored per config:
  the call arguments of the first call (photon_b200.recording.save_npz) -- both calls of a run share
  everything except the fields listed below --
  call1_*          : the arguments that differ in the second call (BOS: density on, file, algorithm;
                     PIV: particle x/y/z of the second frame)
  ref_image0/1     : BOS: the reference's float32 images (1024x1024) of both calls
  ref_binned0/1    : PIV: 4x4-binned sums of the reference's images (256x256 float64 -> float32)
  ref_sum0/1, ref_nonzero0/1
  bos_rotated      : ref_image0/1 (512x512), and from the pos_0000.bin / dir_0000.bin files the reference wrote for
                     the second call (the parameter file has save_lightrays = 1): ray_index (every 180th ray id),
                     ref_pos / ref_dir at those rays, ref_hits (rays with a sensor position over all 722 500)
The sample NRRD is not stored: tests rebuild the path to tests/golden/sample-density.nrrd, a
byte-for-byte copy of sample-data/bos/sample-density.nrrd (input data, 1 MiB).
"""
import argparse
import os
import shutil
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from photon_b200 import recording  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def binned(img, b=4):
    ny, nx = img.shape
    return img.astype(np.float64).reshape(ny // b, b, nx // b, b).sum(axis=(1, 3)).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--work", default="/tmp/photon_recorded")
    ap.add_argument("--reuse", action="store_true", help="reuse recordings already under --work")
    a = ap.parse_args()
    ref = os.path.join(ROOT, "oracle", "_ref", "libphoton_ref_cpu.so")
    subprocess.run(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")], check=True)
    for kind in ("bos", "piv", "bos_rotated"):
        out = os.path.join(a.work, kind)
        rec = os.path.join(out, "recording_" + kind)
        if not (a.reuse and os.path.isdir(rec)):
            subprocess.run([sys.executable, os.path.join(ROOT, "tools", "stage_reference.py"), kind, "--out", out,
                            "--forward", ref], check=True)
        calls = recording.load_recording(rec)
        assert len(calls) == 2
        imgs = [np.fromfile(os.path.join(rec, "call_%02d" % i, "image_out.bin"), np.float32).reshape(calls[i].image_shape)
                for i in range(2)]
        extra = {}
        for i, im in enumerate(imgs):
            extra["ref_sum%d" % i] = np.float64(im.astype(np.float64).sum())
            extra["ref_nonzero%d" % i] = np.int64((im != 0).sum())
            if kind in ("bos", "bos_rotated"):
                extra["ref_image%d" % i] = im
            else:
                extra["ref_binned%d" % i] = binned(im)
        c0, c1 = calls
        if kind == "bos_rotated":
            assert np.array_equal(c0.x, c1.x) and c1.simulate_density_gradients and c1.save_lightrays
            extra["call1_simulate_density_gradients"] = np.asarray(c1.simulate_density_gradients)
            extra["call1_ray_tracing_algorithm"] = np.asarray(c1.ray_tracing_algorithm)
            extra["call1_density_grad_basename"] = np.asarray(os.path.basename(c1.density_grad_filename))
            stage_cwd = os.path.join(out, "stage", "python_codes")       # the dump paths are relative to the scripts' cwd
            pos = np.fromfile(os.path.join(stage_cwd, c1.lightray_position_save_path, "pos_0000.bin"), np.float32).reshape(-1, 3)
            dr = np.fromfile(os.path.join(stage_cwd, c1.lightray_direction_save_path, "dir_0000.bin"), np.float32).reshape(-1, 3)
            assert len(pos) == c1.num_lightrays_save == c1.num_particles * c1.lightray_number_per_particle
            idx = np.arange(0, len(pos), 180)
            extra["ray_index"], extra["ref_pos"], extra["ref_dir"] = idx, pos[idx], dr[idx]
            extra["ref_hits"] = np.int64(np.isfinite(pos[:, 0]).sum())
            c0.save_lightrays = False; c0.num_lightrays_save = 0
            c0.lightray_position_save_path = c0.lightray_direction_save_path = ""
        elif kind == "bos":
            assert np.array_equal(c0.x, c1.x) and np.array_equal(c0.radiance, c1.radiance)
            extra["call1_simulate_density_gradients"] = np.asarray(c1.simulate_density_gradients)
            extra["call1_ray_tracing_algorithm"] = np.asarray(c1.ray_tracing_algorithm)
            extra["call1_density_grad_basename"] = np.asarray(os.path.basename(c1.density_grad_filename))
            shutil.copyfile(c1.density_grad_filename, os.path.join(HERE, "sample-density.nrrd"))
        else:
            for k in ("x", "y", "z", "radiance", "diameter_index"):
                extra["call1_" + k] = getattr(c1, k)
        recording.save_npz(os.path.join(HERE, "recorded_%s.npz" % kind), c0, **extra)
        print(kind, "calls", len(calls), "points", c0.num_particles, "rays/point", c0.lightray_number_per_particle,
              "sums", [float(extra["ref_sum%d" % i]) for i in range(2)])


if __name__ == "__main__":
    main()
