"""Source-point generation: host mirror of the generators in python_codes/run_simulation_02.py over the
C ABI (photon_b200_sunflower / _non_overlapping_dots / _expand_dots / _piv_particles).  Same function
names, arguments and return values as the reference; the one addition is ``rng``: a
``numpy.random.RandomState`` the draws come from (the reference uses numpy's global generator, re-seeded
from the clock).  With the same stream the results equal the reference's function by function."""
import ctypes as C

import numpy as np

from .lib import PhotonLibraryError, load_library

_D = C.c_double
_P = C.c_void_p
_LL = C.c_longlong


def _lib():
    lib = load_library(require_gpu=False)
    lib.photon_b200_sunflower.argtypes = [_D, _D, _P, _P, _P, _LL]
    lib.photon_b200_sunflower.restype = _LL
    lib.photon_b200_non_overlapping_dots.argtypes = [_D, _D, _D, _D, _LL, _D, _P, _LL, _P, _P]
    lib.photon_b200_non_overlapping_dots.restype = _LL
    lib.photon_b200_expand_dots.argtypes = [_P, _P, _LL, _P, _P, _LL, _D, _D, _P, _P, _P, _P, _P]
    lib.photon_b200_expand_dots.restype = C.c_int
    lib.photon_b200_piv_particles.argtypes = [_LL, _P, _P, _P, _P, _P, _D, _D, _P, _D, _P, _P, _P, _P, _P]
    lib.photon_b200_piv_particles.restype = C.c_int
    return lib


def _err(lib):
    return PhotonLibraryError(lib.photon_b200_last_error().decode())


def _rng(rng):
    return rng if rng is not None else np.random.RandomState()


def magnification(simulation_parameters):
    ld = simulation_parameters["lens_design"]
    return ld["focal_length"] / (ld["object_distance"] - ld["focal_length"])


def object_plane(simulation_parameters, optical_system):
    """z_object of rs02.py:816-835 / :1355-1371 from the first optical element."""
    ld = simulation_parameters["lens_design"]
    el = optical_system["design"]["optical_element"][0]["optical_element"][0]
    n = el["element_properties"]["refractive_index"]
    rf, rb = el["element_geometry"]["front_surface_radius"], el["element_geometry"]["back_surface_radius"]
    t = el["element_geometry"]["vertex_distance"]
    f, od = ld["focal_length"], ld["object_distance"]
    image_distance = (1.0 / f - 1.0 / od) ** -1
    h1 = -(f * (n - 1.0) * t) / (rb * n)
    h2 = -(f * (n - 1.0) * t) / (rf * n)
    v2 = image_distance + h2
    v1 = v2 + t
    return v1 - h1 + od


def calculate_sunflower_coordinates(grid_point_diameter, lightray_number_per_grid_point, rng=None):
    """rs02.py:999-1055 -> [x, y] offsets of the points that fill one dot (rings + the centre)."""
    lib = _lib()
    n = lib.photon_b200_sunflower(float(grid_point_diameter), float(lightray_number_per_grid_point), None, None, None, 0)
    if n < 0:
        raise _err(lib)
    area = np.pi * (grid_point_diameter / 2.0) ** 2.0
    nring = int(np.round((grid_point_diameter / 2.0) / np.sqrt(area / lightray_number_per_grid_point)))
    r = _rng(rng)
    phase = np.array([r.rand(1, 1)[0, 0] for _ in range(nring)], np.float64)
    x = np.zeros(n, np.float64)
    y = np.zeros(n, np.float64)
    lib.photon_b200_sunflower(float(grid_point_diameter), float(lightray_number_per_grid_point),
                              phase.ctypes.data if nring else None, x.ctypes.data, y.ctypes.data, n)
    return [x, y]


def create_non_overlapping_dot_coordinates(simulation_parameters, rng=None):
    """rs02.py:1251-1326 -> coordinates[n_generated, 2]."""
    bp, cam = simulation_parameters["bos_pattern"], simulation_parameters["camera_design"]
    num_dots = int(bp["grid_point_number"])
    d_diff = cam["diffraction_diameter"] if cam["implement_diffraction"] else 0.0
    d_diff_microns = d_diff * cam["pixel_pitch"] * 1 / magnification(simulation_parameters)
    dot_diameter = np.sqrt(bp["grid_point_diameter"] ** 2 + d_diff_microns ** 2)
    r = _rng(rng)
    max_iter = 50000
    state = r.get_state()
    u = r.rand(2 * max_iter)                       # drawn ahead; the stream is rewound to what the loop used
    xy = np.zeros((num_dots, 2), np.float64)
    lib = _lib()
    used = C.c_longlong(0)
    placed = lib.photon_b200_non_overlapping_dots(float(bp["X_Min"]), float(bp["X_Max"]), float(bp["Y_Min"]), float(bp["Y_Max"]),
                                                  num_dots, float(dot_diameter), u.ctypes.data, max_iter, xy.ctypes.data,
                                                  C.byref(used))
    if placed < 0:
        raise _err(lib)
    r.set_state(state)
    r.rand(2 * int(used.value))                    # leave the stream where the reference's loop leaves it
    return xy[:placed].copy()


def generate_bos_lightfield_data(simulation_parameters, optical_system, rng=None):
    """rs02.py:1328-1551 -> (lightfield_source, x_grid_point_coordinate_vector, y_grid_point_coordinate_vector)."""
    bp, cam, ld = simulation_parameters["bos_pattern"], simulation_parameters["camera_design"], simulation_parameters["lens_design"]
    grid_point_number = int(bp["grid_point_number"])
    per_dot = bp["particle_number_per_grid_point"]
    z_object = object_plane(simulation_parameters, optical_system)
    X_Min = bp["X_Min"] if bp["X_Min"] else -7.5e4
    X_Max = bp["X_Max"] if bp["X_Max"] else +7.5e4
    Y_Min = bp["Y_Min"] if bp["Y_Min"] else -7.5e4
    Y_Max = bp["Y_Max"] if bp["Y_Max"] else +7.5e4
    M = magnification(simulation_parameters)
    r = _rng(rng)
    if grid_point_number == 1:
        gx = np.array([X_Min + 1 * (X_Max - X_Min) / 2.0 + cam["pixel_pitch"] / M / 2.0])
        gy = np.array([Y_Min + 1 * (Y_Max - Y_Min) / 2.0 + cam["pixel_pitch"] / M / 2.0])
    elif bp["dot_overlap"]:
        rn = r.rand(int(grid_point_number) * 2)
        gx = X_Min + (X_Max - X_Min) * rn[:grid_point_number]
        gy = Y_Min + (Y_Max - Y_Min) * rn[grid_point_number:]
    else:
        if bp["dot_distribution"] == "regular":
            spacing = bp["dot_spacing"] * cam["pixel_pitch"] / M
            nxg = int((bp["X_Max"] - bp["X_Min"]) / spacing)
            nyg = int((bp["Y_Max"] - bp["Y_Min"]) / spacing)
            vx = np.linspace(bp["X_Min"], bp["X_Max"], nxg, endpoint=False)
            vy = np.linspace(bp["Y_Min"], bp["Y_Max"], nyg, endpoint=False)
            X, Y = np.meshgrid(vx, vy, indexing="xy")
            gx, gy = X.flatten("C"), Y.flatten("C")
        else:
            c = create_non_overlapping_dot_coordinates(simulation_parameters, r)
            gx, gy = c[:, 0].copy(), c[:, 1].copy()
        grid_point_number = gx.size
    if bp["grid_point_diameter"] > 0.0 and per_dot > 1.0:
        ox, oy = calculate_sunflower_coordinates(bp["grid_point_diameter"], per_dot, r)
    else:
        ox, oy = np.array([0.0]), np.array([0.0])
    if ox.size > per_dot:
        per_dot = ox.size
        bp["particle_number_per_grid_point"] = per_dot
    n = int(per_dot * grid_point_number)
    filled = ox.size * int(grid_point_number)
    x = np.zeros(max(n, filled), np.float32); y = np.zeros_like(x); z = np.zeros_like(x)
    rad = np.zeros(x.size, np.float64); diam = np.ones(x.size, np.int32)
    z0 = z_object + (ld["object_distance_buffer"] if "object_distance_buffer" in ld else 0.0)
    radiance = bp["lightray_radiance"] if "lightray_radiance" in bp else 10.0
    lib = _lib()
    gx = np.ascontiguousarray(gx, np.float64); gy = np.ascontiguousarray(gy, np.float64)
    ox = np.ascontiguousarray(ox, np.float64); oy = np.ascontiguousarray(oy, np.float64)
    if lib.photon_b200_expand_dots(gx.ctypes.data, gy.ctypes.data, gx.size, ox.ctypes.data, oy.ctypes.data, ox.size,
                                   float(z0), float(radiance), x.ctypes.data, y.ctypes.data, z.ctypes.data,
                                   rad.ctypes.data, diam.ctypes.data) != 0:
        raise _err(lib)
    if filled < n:                                   # rs02.py:1508-1510 keeps count+1 entries of the zero tail
        keep = filled + 1 if (filled + 1) < n else n
        z[filled:keep] = z0; rad[filled:keep] = radiance
        x, y, z, rad, diam = x[:keep], y[:keep], z[:keep], rad[:keep], diam[:keep]
    src = {"x": x, "y": y, "z": z, "radiance": rad, "diameter_index": diam, "source_point_number": 10000,
           "z_offset": z_object - ld["object_distance"], "object_distance": ld["object_distance"]}
    return src, gx, gy


def generate_piv_lightfield_data(simulation_parameters, optical_system, rotation_matrix, rng=None):
    """The generated-particles branch of load_lightfield_data (rs02.py:901-990): uniform box, Gaussian sheet,
    camera rotation (``rotation_matrix`` = calculate_rotation_matrix(x_angle, y_angle, 0), row-major 3x3)."""
    pf = simulation_parameters["particle_field"]
    n = int(pf["particle_number"])
    lo = np.array([pf.get("X_Min", -7.5e4), pf.get("Y_Min", -7.5e4), pf.get("Z_Min", -7.5e3)], np.float64)
    hi = np.array([pf.get("X_Max", 7.5e4), pf.get("Y_Max", 7.5e4), pf.get("Z_Max", 7.5e3)], np.float64)
    r = _rng(rng)
    ux, uy, uz = r.rand(n), r.rand(n), r.rand(n)
    const = 500.0 if pf["perform_mie_scattering"] else 1e4
    const = pf.get("lightray_radiance", const)
    z_object = object_plane(simulation_parameters, optical_system)
    R = np.ascontiguousarray(np.asarray(rotation_matrix, np.float64).reshape(9))
    x = np.zeros(n, np.float32); y = np.zeros(n, np.float32); z = np.zeros(n, np.float32)
    rad = np.zeros(n, np.float64); diam = np.zeros(n, np.int32)
    lib = _lib()
    if lib.photon_b200_piv_particles(n, ux.ctypes.data, uy.ctypes.data, uz.ctypes.data, lo.ctypes.data, hi.ctypes.data,
                                     float(pf["gaussian_beam_fwhm"]), float(const), R.ctypes.data, float(z_object),
                                     x.ctypes.data, y.ctypes.data, z.ctypes.data, rad.ctypes.data, diam.ctypes.data) != 0:
        raise _err(lib)
    ld = simulation_parameters["lens_design"]
    return {"x": x, "y": y, "z": z, "radiance": rad, "diameter_index": diam, "source_point_number": 10000,
            "lightray_number_per_particle": pf["lightray_number_per_particle"], "num_particles": n,
            "z_offset": z_object - ld["object_distance"], "object_distance": ld["object_distance"]}


def rotate_coordinates(X, Y, Z, Alpha, Beta, Gamma, XC, YC, ZC):
    """rs02.py:395-444: rotation about (XC, YC, ZC) by R = Rx * Ry * Rz."""
    from .scenes import rotation_matrix
    W = np.dot(rotation_matrix(Alpha, Beta, Gamma), np.array([X - XC, Y - YC, Z - ZC]))
    return [W[0] + XC, W[1] + YC, W[2] + ZC]


def generate_calibration_lightfield_data(simulation_parameters, optical_system, plane_index, rng=None):
    """rs02.py:1057-1249: one plane of the calibration target -- a regular grid of dots filled with sunflower
    points (x index outer, y index inner), two quarter-size origin markers at (-dx/2, 0) and (0, +dy/2), the plane
    at ``calibration_plane_spacing * (plane_index - (planes - 1) / 2)`` from the object plane, rotated by the
    camera angles about the world origin."""
    cg, cam, ld = simulation_parameters["calibration_grid"], simulation_parameters["camera_design"], simulation_parameters["lens_design"]
    planes = int(cg["calibration_plane_number"])
    nxg, nyg = int(cg["x_grid_point_number"]), int(cg["y_grid_point_number"])
    per_dot = int(cg["particle_number_per_grid_point"])
    z_object = object_plane(simulation_parameters, optical_system)
    plane_z = cg["calibration_plane_spacing"] * np.linspace(-(planes - 1.0) / 2.0, (planes - 1.0) / 2.0, planes, endpoint=True)
    vx = cg["x_grid_point_spacing"] * np.linspace(-(nxg - 1.0) / 2.0, (nxg - 1.0) / 2.0, nxg, endpoint=True)
    vy = cg["y_grid_point_spacing"] * np.linspace(-(nyg - 1.0) / 2.0, (nyg - 1.0) / 2.0, nyg, endpoint=True)
    r = _rng(rng)
    ox, oy = calculate_sunflower_coordinates(cg["grid_point_diameter"], per_dot, r)
    gx, gy = np.repeat(vx, nyg), np.tile(vy, nxg)
    x = (gx[:, None] + ox[None, :]).reshape(-1)
    y = (gy[:, None] + oy[None, :]).reshape(-1)
    mx, my = calculate_sunflower_coordinates(cg["grid_point_diameter"] / 4, per_dot / 16, r)
    x = np.concatenate([x, mx - cg["x_grid_point_spacing"] / 2.0, mx + 0])
    y = np.concatenate([y, my + 0, my + cg["y_grid_point_spacing"] / 2.0])
    z = plane_z[plane_index] * np.ones(x.shape)
    x, y, z = rotate_coordinates(x, y, z, cam["x_camera_angle"], cam["y_camera_angle"], 0.0, 0.0, 0.0, 0.0)
    return {"x": x, "y": y, "z": z + z_object, "radiance": np.ones(x.shape), "diameter_index": np.ones(x.shape),
            "source_point_number": 10000, "z_offset": z_object - ld["object_distance"], "object_distance": ld["object_distance"]}


def generate_bos_image_lightfield_data(simulation_parameters, optical_system, image=None):
    """rs02.py:1554-1697: a BOS target given as an image -- every pixel > 0 of channel 0 becomes one source point
    at its pixel centre (x mirrored), its grey value the radiance, the image spanning +-75 mm; then the camera
    rotation.  ``image`` is the array ``mpimg.imread`` would return ([rows][cols][channels] or [rows][cols]) or a
    file name (the reference reads a fixed ``./bos_pattern_black_dot_936.png`` that it does not ship).  The
    reference's dict lacks ``z_offset`` / ``object_distance``, which the ctypes layer needs; they are added."""
    cam, ld = simulation_parameters["camera_design"], simulation_parameters["lens_design"]
    if image is None or isinstance(image, str):
        import matplotlib.image as mpimg
        image = mpimg.imread(image or "./bos_pattern_black_dot_936.png")
    img = np.asarray(image)
    if img.ndim == 3:
        img = img[:, :, 0]
    z_object = object_plane(simulation_parameters, optical_system)
    X_Min, X_Max, Y_Max = -7.5e4, +7.5e4, +7.5e4
    height, width = img.shape
    pixel_width = (X_Max - X_Min) / width
    rc = np.argwhere(img > 0)
    x = X_Min + (width - rc[:, 1]) * pixel_width + pixel_width / 2.0
    y = Y_Max - (rc[:, 0] * pixel_width + pixel_width / 2.0)
    radiance = img[img > 0].astype(np.float64)
    x, y, z = rotate_coordinates(x, y, 0.0 * np.ones(x.shape), cam["x_camera_angle"], cam["y_camera_angle"], 0.0, 0.0, 0.0, 0.0)
    return {"x": x, "y": y, "z": z + z_object, "radiance": radiance, "diameter_index": np.ones(x.shape),
            "source_point_number": min(10000, x.size), "z_offset": z_object - ld["object_distance"],
            "object_distance": ld["object_distance"]}
