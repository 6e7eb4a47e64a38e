"""Synthetic scene builders: produce a :class:`photon_b200.abi.RayTracingCall` (the 29
arguments of ``start_ray_tracing``) for the configurations in BASELINE.md section 4.

The optics follow what the reference's Python computes before it calls the library:
  * single biconvex lens, index solved from (f, R, thickness): ``run_simulation_02.py:259-363``
  * image distance, principal planes, lens z:               ``perform_ray_tracing_03.py:2016-2041``
  * element centre / plane parameters / system index:       ``perform_ray_tracing_03.py:18-345, 2058-2078``
  * camera rotation matrices:                               ``run_simulation_02.py:366-392, 1748-1756``
  * BOS dots filled with sunflower points at z_object:      ``run_simulation_02.py:999-1055, 1328-1551``
  * PIV particles in a Gaussian light sheet:                ``run_simulation_02.py:774-996``
All lengths are micrometres.  Everything here is deterministic (seeded numpy RNG only).
"""
import math

import numpy as np

from .abi import RayTracingCall, camera_design_struct, element_data_struct


def rotation_matrix(theta_x, theta_y, theta_z=0.0):
    """rs02.py:366-392 (R = Rx * Ry * Rz)."""
    cx, sx, cy, sy, cz, sz = (math.cos(theta_x), math.sin(theta_x), math.cos(theta_y),
                              math.sin(theta_y), math.cos(theta_z), math.sin(theta_z))
    rx = np.array([[1, 0, 0], [0, cx, sx], [0, -sx, cx]], float)
    ry = np.array([[cy, 0, -sy], [0, 1, 0], [sy, 0, cy]], float)
    rz = np.array([[cz, sz, 0], [-sz, cz, 0], [0, 0, 1]], float)
    return rx @ ry @ rz


def solve_lens(focal_length, f_number, radius, lens_model):
    """Thickness and refractive index of the single biconvex lens (rs02.py:259-363)."""
    pitch = focal_length / f_number
    if lens_model == "thin-lens":
        t = 0.0
    else:
        t = 2.0 * (radius - math.sqrt(radius ** 2 - (pitch / 2.0) ** 2))
    f, R = focal_length, radius
    disc = math.sqrt(-4.0 * t * f + (2.0 * f + R) ** 2)
    den = 2.0 * f * (t - 2.0 * R)
    n1 = (2.0 * t * f - 2.0 * f * R - R ** 2 - R * disc) / den
    n2 = (2.0 * t * f - 2.0 * f * R - R ** 2 + R * disc) / den
    cands = [n for n in (n1, n2) if n >= 1]
    return pitch, t, min(cands)


def camera_and_lens(call, *, focal_length, f_number, object_distance, lens_model="general",
                    lens_radius=1e5, nx=1024, ny=1024, pixel_pitch=17.0, diffraction_diameter=0.0,
                    implement_diffraction=False, x_camera_angle=0.0, y_camera_angle=0.0,
                    pixel_gain=25.0, pixel_bit_depth=10):
    """Fill the lens / camera arguments of ``call``; returns z_object (where sources sit)."""
    pitch, t, n = solve_lens(focal_length, f_number, lens_radius, lens_model)
    image_distance = (1.0 / focal_length - 1.0 / object_distance) ** (-1.0)
    h1 = -(focal_length * (n - 1.0) * t) / (-lens_radius * n)
    h2 = -(focal_length * (n - 1.0) * t) / (lens_radius * n)
    v2 = image_distance + h2
    v1 = v2 + t
    z_lens = (v1 + v2) / 2.0
    z_object = v1 - h1 + object_distance

    el = element_data_struct()
    el.axial_offset_distances[0] = 0.0
    el.axial_offset_distances[1] = 0.0
    el.element_number = 1.0
    el.element_type = {"thin-lens": b"t", "apparent": b"n"}.get(lens_model, b"l")
    el.elements_coplanar = 0.0
    el.z_inter_element_distance = 0.0
    g = el.element_geometry
    g.front_surface_radius, g.front_surface_spherical = lens_radius, True
    g.back_surface_radius, g.back_surface_spherical = -lens_radius, True
    g.pitch, g.vertex_distance = pitch, t
    p = el.element_properties
    p.abbe_number, p.absorbance_rate, p.refractive_index = float("nan"), 0.0, n
    p.thin_lens_focal_length, p.transmission_ratio = focal_length, 1.0

    call.elements = [el]
    call.element_center = np.array([[0.0, 0.0, z_lens]])
    call.element_plane_parameters = np.array([[0.0, 0.0, 1.0, -z_lens]])
    call.element_system_index = np.array([1], np.int32)
    call.lens_pitch = pitch
    call.image_distance = image_distance
    call.aperture_f_number = f_number
    call.object_distance = object_distance
    call.z_offset = z_object - object_distance

    cam = camera_design_struct()
    cam.pixel_bit_depth, cam.pixel_gain, cam.pixel_pitch = pixel_bit_depth, pixel_gain, pixel_pitch
    cam.x_camera_angle, cam.y_camera_angle = x_camera_angle, y_camera_angle
    cam.x_pixel_number, cam.y_pixel_number = nx, ny
    cam.z_sensor = 0.0
    cam.diffraction_diameter = diffraction_diameter
    cam.implement_diffraction = implement_diffraction
    R = rotation_matrix(x_camera_angle, y_camera_angle, 0.0)
    Rinv = R.T
    for i in range(3):
        for j in range(3):
            cam.rotation_matrix[i * 3 + j] = np.float32(R[i, j])
            cam.inverse_rotation_matrix[i * 3 + j] = np.float32(Rinv[i, j])
    call.camera = cam
    return z_object


def sunflower(diameter, count):
    """Deterministic restatement of rs02.py:999-1055 (random ring phase replaced by 0)."""
    area = math.pi * (diameter / 2.0) ** 2
    spacing = math.sqrt(area / count)
    nring = int(round((diameter / 2.0) / spacing))
    radii = np.linspace(spacing, diameter / 2.0, nring)
    xs, ys = [], []
    for r in radii:
        m = round((1.0 / spacing) * 2.0 * math.pi * r)
        th = (2.0 * math.pi / m) * (np.arange(1.0, m) - 1.0)
        xs.append(r * np.cos(th))
        ys.append(r * np.sin(th))
    xs.append(np.array([0.0]))
    ys.append(np.array([0.0]))
    return np.concatenate(xs), np.concatenate(ys)


def bos_dots(call, z_object, *, dots_per_side, extent, points_per_dot, dot_diameter, radiance=10.0):
    """Regular dots_per_side^2 grid over [-extent, +extent)^2, each dot filled with sunflower
    points (the 'regular' dot distribution of rs02.py:1417-1436).  Fully deterministic."""
    gx = np.linspace(-extent, extent, dots_per_side, endpoint=False)
    X, Y = np.meshgrid(gx, gx, indexing="xy")
    if points_per_dot > 1:
        sx, sy = sunflower(dot_diameter, points_per_dot)
    else:
        sx, sy = np.zeros(1), np.zeros(1)
    x = (X.reshape(-1, 1) + sx.reshape(1, -1)).reshape(-1)
    y = (Y.reshape(-1, 1) + sy.reshape(1, -1)).reshape(-1)
    call.x = x.astype(np.float32)
    call.y = y.astype(np.float32)
    call.z = np.full(x.shape, z_object, np.float32)
    call.radiance = np.full(x.shape, radiance, np.float64)
    call.diameter_index = np.ones(x.shape, np.int32)
    call.scattering_type = "diffuse"
    call.beam_wavelength = 0.0
    return sx.size


def gaussian_blob(n, spacing, sigma, rho0=1.225, amp=0.5):
    """rho = rho0 + amp * exp(-r^2 / 2 sigma^2) on an n^3 grid centred at the volume origin
    (BASELINE.md C3).  Returns float32 [z][y][x]."""
    c = (np.arange(n, dtype=np.float64) - (n - 1) / 2.0) * spacing
    g = np.exp(-c ** 2 / (2.0 * sigma ** 2))
    rho = rho0 + amp * (g[:, None, None] * g[None, :, None] * g[None, None, :])
    return rho.astype(np.float32)


def constant_gradient_slab(n, spacing, drho_dx, rho0=1.225):
    """rho = rho0 + drho_dx * x: the analytic KAT field of nrrd_functions.py:60-82."""
    c = (np.arange(n, dtype=np.float64) - (n - 1) / 2.0) * spacing
    rho = rho0 + drho_dx * c[None, None, :] + np.zeros((n, n, 1))
    return rho.astype(np.float32)


def c3_bos(nrrd_path, *, n=512, dots_per_side=100, points_per_dot=100, rays_per_point=1000,
           nx=1024, ny=1024, algorithm=2, write_field=True, source_point_number=10000, field_z_centre=0.0,
           blob_amplitude=0.5):
    """BOS through a Gaussian density blob on an n^3 NRRD grid.

    Geometry of sample-data/bos/sample-parameters.mat with y_camera_angle = 0: f = 105 mm,
    f/11, object distance 1e6 um, thin lens, pixel pitch 10 um, diffraction 3 px,
    ray_cone_pitch_ratio 1e-4, RK4.  Field extent 89 936 um regardless of n (spacing scales),
    sigma = 15 mm, centred at the volume-frame origin (NRRD z origin = 750e3 - extent/2), or at
    volume-frame z = field_z_centre (the dots sit at z = +250 mm; a box that encloses them makes rays
    start inside the field, which is the only way the reference's rk45 / adams-bashforth integrate).
    """
    from .nrrd import write_nrrd
    call = RayTracingCall()
    z_object = camera_and_lens(call, focal_length=105e3, f_number=11.0, object_distance=1e6,
                               lens_model="thin-lens", nx=nx, ny=ny, pixel_pitch=10.0,
                               diffraction_diameter=3.0, implement_diffraction=True)
    extent = 511 * 176.0
    spacing = extent / (n - 1)
    if write_field:
        rho = gaussian_blob(n, spacing, 15e3, amp=blob_amplitude)
        write_nrrd(nrrd_path, rho, (spacing,) * 3, (-extent / 2, -extent / 2, 750e3 + field_z_centre - extent / 2))
    bos_dots(call, z_object, dots_per_side=dots_per_side, extent=40e3,
             points_per_dot=points_per_dot, dot_diameter=300.0)
    call.lightray_number_per_particle = rays_per_point
    call.source_point_number = source_point_number
    call.simulate_density_gradients = True
    call.density_grad_filename = str(nrrd_path)
    call.ray_tracing_algorithm = algorithm
    call.ray_cone_pitch_ratio = 1e-4
    return call


def mie_table_synthetic(num_angles=255, num_diameters=27):
    """A smooth stand-in for the Bohren-Huffman table (bhmie.py) with the same shape the
    reference passes (255 angles 0..pi, 27 diameters): strong forward lobe, weaker back lobe."""
    ang = np.linspace(0.0, math.pi, num_angles)
    d = np.linspace(0.5, 3.0, num_diameters)
    irr = (1e6 * d[None, :] ** 2) * (np.exp(-(ang[:, None] * d[None, :]) ** 2 * 4.0)
                                    + 0.02 * (1 + np.cos(ang[:, None]) ** 2)
                                    + 0.05 * np.exp(-((math.pi - ang[:, None]) * d[None, :]) ** 2))
    return ang.astype(np.float32), irr.astype(np.float32)


def c1_piv(*, particles=50000, rays_per_particle=10000, nx=1024, ny=1024, seed=1105,
           mie=True, lens_model="general"):
    """BASELINE.md C1-like PIV case: particles uniform in +-75 mm x +-75 mm x +-7.5 mm around the
    object plane, Gaussian sheet radiance, Mie scattering, thick lens f = 105 mm f/8, object
    distance 700 mm, 17 um pixels, bilinear splat, ray_cone_pitch_ratio = 1 (rs02.py:774-996)."""
    call = RayTracingCall()
    z_object = camera_and_lens(call, focal_length=105e3, f_number=8.0, object_distance=700e3,
                               lens_model=lens_model, nx=nx, ny=ny, pixel_pitch=17.0,
                               implement_diffraction=False)
    rng = np.random.default_rng(seed)
    call.x = rng.uniform(-7.5e4, 7.5e4, particles).astype(np.float32)
    call.y = rng.uniform(-7.5e4, 7.5e4, particles).astype(np.float32)
    zl = rng.uniform(-7.5e3, 7.5e3, particles)
    call.z = (zl + z_object).astype(np.float32)
    sheet_sigma = 7.5e3 / 2.0 / 2.0
    call.radiance = 500.0 * np.exp(-zl ** 2 / (2 * sheet_sigma ** 2)) / (sheet_sigma * math.sqrt(2 * math.pi)) * 1e3
    call.diameter_index = np.ones(particles, np.int32)
    call.lightray_number_per_particle = rays_per_particle
    call.source_point_number = 10000
    call.beam_wavelength = 0.532
    call.ray_cone_pitch_ratio = 1.0
    if mie:
        call.scattering_type = "mie"
        call.mie_angles, call.mie_irradiance = mie_table_synthetic()
        call.mie_beam = np.array([0, 1, 0], np.float32)
        call.mie_inverse_rotation = np.eye(3, dtype=np.float32).reshape(-1)
    return call


def turbulent_field(n, *, seed=1, k_lo=4.0, k_hi=128.0, rms=0.05, rho0=1.225, device=None, keep_on_device=False):
    """Band-limited Gaussian random density field of BASELINE.md C4: white noise from
    ``default_rng(seed)`` shaped in Fourier space to E(k) ~ k^-5/3 for k_lo <= |k| <= k_hi (cycles
    per box), zero elsewhere, scaled to an rms of ``rms`` around ``rho0``.  float32 [z][y][x].
    The FFT runs with numpy, or with torch on ``device`` (the 1024^3 case takes minutes on a CPU);
    ``keep_on_device`` returns the torch tensor instead of copying the 4 GiB back to the host.
    Amplitude per mode = sqrt(E(k) / (4 pi k^2)) ~ k^(-11/6)."""
    rng = np.random.default_rng(seed)
    w = rng.standard_normal((n, n, n), dtype=np.float32)
    kz = np.fft.fftfreq(n, 1.0 / n).astype(np.float32)
    kx = np.fft.rfftfreq(n, 1.0 / n).astype(np.float32)
    if device is not None:
        import torch
        t = torch.from_numpy(w).to(device)
        F = torch.fft.rfftn(t)
        del t
        kzt, kxt = torch.from_numpy(kz).to(device), torch.from_numpy(kx).to(device)
        k = torch.sqrt(kzt[:, None, None] ** 2 + kzt[None, :, None] ** 2 + kxt[None, None, :] ** 2)
        amp = torch.where((k >= k_lo) & (k <= k_hi), k.clamp(min=1.0) ** (-11.0 / 6.0), torch.zeros_like(k))
        del k
        F *= amp
        del amp
        f = torch.fft.irfftn(F, s=(n, n, n))
        del F
        f *= rms / f.std()
        f += rho0
        f = f.to(torch.float32).contiguous()
        return f if keep_on_device else f.cpu().numpy()          # keep_on_device: hand it over with "mem:<name>"
    F = np.fft.rfftn(w)
    k = np.sqrt(kz[:, None, None] ** 2 + kz[None, :, None] ** 2 + kx[None, None, :] ** 2)
    amp = np.where((k >= k_lo) & (k <= k_hi), np.maximum(k, 1.0) ** (-11.0 / 6.0), 0.0)
    f = np.fft.irfftn(F * amp, s=(n, n, n), axes=(0, 1, 2))
    f *= rms / f.std()
    return (f + rho0).astype(np.float32)


def c4_piv(nrrd_path, *, n=1024, particles=1000000, rays_per_particle=1000, nx=2048, ny=2048,
           write_field=True, seed=2, device=None, frame_shift_um=0.0, source_point_number=10000, lib=None):
    """PIV with Mie-scattering particles seen through a turbulent density field.

    n^3 band-limited random field (:func:`turbulent_field`, k = 4..min(128, n/4)), extent 200 mm,
    NRRD origin (-100 mm, -100 mm, 450 mm): volume-frame z in [-300, -100] mm, between the light sheet
    (z = -50 mm for an object distance of 700 mm) and the lens.  Particles uniform in +-75 x +-75 x
    +-7.5 mm (``default_rng(seed)``), Gaussian sheet radiance, Mie table, thick lens f = 105 mm f/8,
    nx x ny sensor of 17 um pixels, bilinear splat, RK4.  ``frame_shift_um`` displaces the particles
    along x for the second frame of the image pair."""
    from .nrrd import write_nrrd
    call = c1_piv(particles=particles, rays_per_particle=rays_per_particle, nx=nx, ny=ny, seed=seed)
    call.x = (call.x + np.float32(frame_shift_um)).astype(np.float32)
    # the Bohren-Huffman table the reference builds for a PIV run (rs02.py:641-696): 27 diameters x 255 angles
    from .mie import calculate_mie_scattering_intensity
    ang, irr = calculate_mie_scattering_intensity(
        {"particle_field": {"medium_refractive_index": 1.0, "particle_refractive_index": 1.46,
                            "mie_scattering_angle_number": 128, "beam_wavelength": 0.532}}, np.linspace(0.1, 2.0, 27))
    call.mie_angles, call.mie_irradiance = ang.astype(np.float32), np.ascontiguousarray(irr, np.float32)
    call.diameter_index = np.random.default_rng(seed + 1).integers(0, 27, particles).astype(np.int32)
    extent = 200e3
    spacing = extent / (n - 1)
    if write_field and str(nrrd_path).startswith("mem:"):
        # the field is made on the GPU and handed over in memory (no 4 GiB round trip through a file); the call keeps it alive
        from .nrrd import register_field
        rho = turbulent_field(n, seed=1, k_hi=min(128.0, n / 4.0), device=device, keep_on_device=device is not None)
        if device is not None:
            call._field_keepalive = rho
            register_field(lib, str(nrrd_path)[4:], (rho.data_ptr(), tuple(rho.shape)), (spacing,) * 3, (-extent / 2, -extent / 2, 450e3))
        else:
            call._field_keepalive = register_field(lib, str(nrrd_path)[4:], rho, (spacing,) * 3, (-extent / 2, -extent / 2, 450e3))
    elif write_field:
        rho = turbulent_field(n, seed=1, k_hi=min(128.0, n / 4.0), device=device)
        write_nrrd(nrrd_path, rho, (spacing,) * 3, (-extent / 2, -extent / 2, 450e3))
    call.source_point_number = source_point_number
    call.simulate_density_gradients = True
    call.density_grad_filename = str(nrrd_path)
    call.ray_tracing_algorithm = 2
    return call


def three_element_system(call, *, focal_length=105e3, f_number=11.0, separation=8e3, lens_radius=1e5):
    """Replace the single lens of ``call`` (set by :func:`camera_and_lens`) by three sequential elements on
    the same axis: thick lens (f = 2 focal_length), aperture stop, thick lens (f = 2 focal_length), spaced
    ``separation`` / 2 apart around the original lens centre -- together about the original focal length
    (1/f = 1/f1 + 1/f2 - d/(f1 f2)).  System indices 3, 2, 1 in the order the rays meet them
    (prt.cu:1274-1381: stage s takes the element whose index is max - s)."""
    z_lens = float(call.element_center[0][2])
    pitch, t, n = solve_lens(2.0 * focal_length, 2.0 * f_number, lens_radius, "general")
    elems, centers, planes = [], [], []
    for kind, dz in ((b"l", +separation / 2.0), (b"a", 0.0), (b"l", -separation / 2.0)):
        el = element_data_struct()
        el.axial_offset_distances[0] = el.axial_offset_distances[1] = 0.0
        el.element_number = float(len(elems) + 1)
        el.element_type = kind
        el.elements_coplanar = 0.0
        el.z_inter_element_distance = 0.0
        g = el.element_geometry
        g.front_surface_radius, g.front_surface_spherical = lens_radius, True
        g.back_surface_radius, g.back_surface_spherical = -lens_radius, True
        g.pitch = pitch if kind == b"l" else 0.95 * pitch
        g.vertex_distance = t if kind == b"l" else 500.0
        p = el.element_properties
        p.abbe_number, p.absorbance_rate, p.refractive_index = float("nan"), 0.0, n
        p.thin_lens_focal_length, p.transmission_ratio = 2.0 * focal_length, 1.0
        elems.append(el)
        centers.append([0.0, 0.0, z_lens + dz])
        planes.append([0.0, 0.0, 1.0, -(z_lens + dz)])
    call.elements = elems
    call.element_center = np.array(centers)
    call.element_plane_parameters = np.array(planes)
    call.element_system_index = np.array([3, 2, 1], np.int32)
    return call


def c5_stress(field, *, n=2048, nz=1024, dots_per_side=100, points_per_dot=100, rays_per_point=1000,
              nx=1024, ny=1024, write_field=True, source_point_number=10000, lib=None):
    """BOS through a Gaussian blob on an n x n x nz grid (the reference keeps at most
    1024 z-planes, trace.h:2037-2038; 2048 x 2048 x 1024 = 2^32 voxels = 64 GiB of texels), three sequential
    optical elements (:func:`three_element_system`), diffraction splat, RK4.  ``field``: an NRRD path, or
    ``"mem:<name>"`` to hand the volume over in memory (``lib`` required; the returned call keeps it alive).
    The step-size sweep of C5 is ``PHOTON_STEP_SCALE`` = 1, 0.5, 0.25."""
    from .nrrd import register_field, write_nrrd
    call = RayTracingCall()
    z_object = camera_and_lens(call, focal_length=105e3, f_number=11.0, object_distance=1e6, lens_model="general",
                               nx=nx, ny=ny, pixel_pitch=10.0, diffraction_diameter=3.0, implement_diffraction=True)
    three_element_system(call, focal_length=105e3, f_number=11.0)
    extent = 511 * 176.0
    sx, sz = extent / (n - 1), extent / (nz - 1)
    if write_field:
        cx = (np.arange(n, dtype=np.float64) - (n - 1) / 2.0) * sx
        cz = (np.arange(nz, dtype=np.float64) - (nz - 1) / 2.0) * sz
        gx, gz = np.exp(-cx ** 2 / (2.0 * 15e3 ** 2)).astype(np.float32), np.exp(-cz ** 2 / (2.0 * 15e3 ** 2)).astype(np.float32)
        rho = np.float32(1.225) + np.float32(0.5) * (gz[:, None, None] * gx[None, :, None] * gx[None, None, :])
        origin = (-extent / 2, -extent / 2, 750e3 - extent / 2)
        if str(field).startswith("mem:"):
            call._field_keepalive = register_field(lib, str(field)[4:], rho, (sx, sx, sz), origin)
        else:
            write_nrrd(field, rho, (sx, sx, sz), origin)
    bos_dots(call, z_object, dots_per_side=dots_per_side, extent=40e3, points_per_dot=points_per_dot, dot_diameter=300.0)
    call.lightray_number_per_particle = rays_per_point
    call.source_point_number = source_point_number
    call.simulate_density_gradients = True
    call.density_grad_filename = str(field)
    call.ray_tracing_algorithm = 2
    call.ray_cone_pitch_ratio = 1e-4
    return call
