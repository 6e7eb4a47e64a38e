"""Parameter-file front end: a ``.mat`` file of ``simulation_parameters`` -> the ``start_ray_tracing``
calls the reference's scripts would make for it -> 16-bit and raw images.

Host-side mirror of the callers above the hot path, with the reference's function names:

  * ``loadmat``                                   run_simulation_02.py:22-31 (+ ``_check_keys``)
  * ``create_single_lens_optical_system``         run_simulation_02.py:33-257 (the fields the path reads)
  * ``create_camera_optical_system``              run_simulation_02.py:259-363
  * ``calculate_rotation_matrix``                 run_simulation_02.py:366-392
  * ``calculate_particle_diameter_distribution``  run_simulation_02.py:446-595
  * ``calculate_particle_diameter_indices``       run_simulation_02.py:597-638
  * ``create_mie_scattering_data``                run_simulation_02.py:699-772
  * ``prepare_call``                              perform_ray_tracing_03.py:1949-2187 and :1631-1947 --
    everything perform_ray_tracing_03 computes before the ctypes call, as a RayTracingCall
  * ``perform_ray_tracing_03``                    perform_ray_tracing_03.py:1949-2259 (call + image export)
  * ``run_simulation_02``                         run_simulation_02.py:1725-2110, 'piv', 'cal' and 'bos' branches

Source points come from :mod:`photon_b200.source_points`, the Mie table from :mod:`photon_b200.mie`,
the trace and the 16-bit export run in the CUDA library (no numpy fallback for either).  ``generate_bos_image_lightfield_data``
(a BOS target read from an image; never called by the reference's run loop either) lives in source_points too.
Random draws: the reference uses numpy's global generator; pass ``rng`` (a ``RandomState``) to pin them.
"""
import os

import numpy as np

from . import mie as _mie
from . import source_points as _sp
from .abi import RayTracingCall, camera_design_struct, element_data_struct
from .scenes import rotation_matrix as calculate_rotation_matrix


def _todict(obj):
    out = {}
    for name in obj._fieldnames:
        v = getattr(obj, name)
        out[name] = _todict(v) if hasattr(v, "_fieldnames") else v
    return out


def loadmat(filename):
    """scipy's loadmat with MATLAB structs turned into nested dicts (rs02.py:22-31)."""
    import scipy.io as sio
    data = sio.loadmat(filename, struct_as_record=False, squeeze_me=True)
    return {k: (_todict(v) if hasattr(v, "_fieldnames") else v) for k, v in data.items()}


def create_single_lens_optical_system():
    """The nested description of one biconvex lens inside one 'system' (rs02.py:33-257)."""
    nan = float("nan")
    lens = {
        "element_type": "lens", "element_number": nan, "elements_coplanar": nan, "z_inter_element_distance": 0.0,
        "axial_offset_distances": np.array([0.0, 0.0]), "rotation_angles": np.array([0.0, 0.0, 0.0]),
        "element_geometry": {"pitch": 100.0e3, "front_surface_spherical": True, "back_surface_spherical": True,
                             "front_surface_radius": 200.0e3, "back_surface_radius": -400.0e3, "vertex_distance": 10.0e3},
        "element_properties": {"refractive_index": 1.5, "abbe_number": nan, "thin_lens_focal_length": 85.0e3,
                               "transmission_ratio": 1.0, "absorbance_rate": 0.0},
    }
    inner = {"element_type": "system", "element_number": 1, "elements_coplanar": False, "z_inter_element_distance": 1.0e4,
             "axial_offset_distances": np.array([0.0, 0.0]), "rotation_angles": np.array([0.0, 0.0, 0.0]),
             "element_geometry": nan, "element_properties": nan, "optical_element": [lens]}
    design = {"element_type": "system", "element_number": 1, "elements_coplanar": False, "z_inter_element_distance": 0.0,
              "axial_offset_distances": np.array([0.0, 0.0]), "rotation_angles": np.array([0.0, 0.0, 0.0]),
              "element_geometry": {}, "element_properties": {}, "optical_element": [inner]}
    return {"distance_units": "micron", "angle_units": "degrees", "design": design}


def create_camera_optical_system(simulation_parameters):
    """Thickness from the lens pitch and radius, index from (f, R, t) -- rs02.py:259-363."""
    ld = simulation_parameters["lens_design"]
    f, R = float(ld["focal_length"]), float(ld["lens_radius_of_curvature"])
    pitch = f / float(ld["aperture_f_number"])
    t = 0.0 if ld["lens_model"] == "thin-lens" else 2.0 * (R - np.sqrt(R ** 2 - (pitch / 2.0) ** 2))
    root = np.sqrt(-4.0 * t * f + (2.0 * f + R) ** 2)
    den = 2.0 * f * (t - 2.0 * R)
    cands = [(2.0 * t * f - 2.0 * f * R - R ** 2 - R * root) / den, (2.0 * t * f - 2.0 * f * R - R ** 2 + R * root) / den]
    n = min([c if (np.isreal(c) and c >= 1) else np.inf for c in cands])
    system = create_single_lens_optical_system()
    lens = system["design"]["optical_element"][0]["optical_element"][0]
    lens["element_geometry"].update(pitch=pitch, front_surface_radius=+R, back_surface_radius=-R, vertex_distance=t)
    lens["element_properties"].update(refractive_index=float(n), thin_lens_focal_length=f)
    return system


def calculate_particle_diameter_distribution(simulation_parameters):
    """Log-normal diameters between the two points of equal density that leave ``threshhold`` of the mass
    outside (Newton iteration on the upper point) -- rs02.py:446-595.  -> [diameters, normalised pdf]."""
    from scipy.special import erf
    pf = simulation_parameters["particle_field"]
    mean, std = float(pf["particle_diameter_mean"]), float(pf["particle_diameter_std"])
    count, t = int(pf["particle_diameter_number"]), float(pf["particle_diameter_cdf_threshhold"])
    mu = np.log(mean) - 0.5 * np.log(1.0 + (std / mean) ** 2)
    sg = np.sqrt(np.log(1.0 + (std / mean) ** 2.0))

    def pdf(x):
        return np.exp(-(np.log(x) - mu) ** 2.0 / (2.0 * sg ** 2.0)) / (x * sg * np.sqrt(2.0 * np.pi))

    def cdf(x):
        return (1.0 + erf((np.log(x) - mu) / (sg * np.sqrt(2.0)))) / 2.0

    def both_roots(y):
        r = sg * np.sqrt(sg ** 2.0 - 2.0 * mu - 2.0 * np.log(y * sg * np.sqrt(2.0 * np.pi)))
        return np.exp(mu - sg ** 2 - r), np.exp(mu - sg ** 2 + r)

    hi = np.exp(mu + sg)
    while True:
        lo, hi = both_roots(pdf(hi))
        residual = 1.0 - (cdf(hi) - cdf(lo)) - t
        dlo_dhi = -np.exp(2.0 * mu - 2.0 * sg ** 2.0) / hi ** 2.0
        step = residual / (pdf(lo) * dlo_dhi - pdf(hi))
        if abs(step) < np.finfo(np.float64).eps * 1.0e2:
            break
        hi = hi - step
    spacing = (hi - lo) / count
    d = lo + spacing * (np.arange(0, count) + 0.5)
    p = pdf(d)
    return [d, p / np.sum(p)]


def calculate_particle_diameter_indices(simulation_parameters, particle_diameter_pdf, particle_diameter_vector, rng=None):
    """One diameter bin per particle drawn from the pdf (rs02.py:597-638; the last bin is never assigned
    there -- draws that fall in it keep index 0 -- reproduced)."""
    n = int(simulation_parameters["particle_field"]["particle_number"])
    edges = np.insert(np.cumsum(particle_diameter_pdf), 0, 0.0)
    u = (rng if rng is not None else np.random).rand(n)
    idx = np.zeros(n)
    for k in range(0, len(particle_diameter_vector) - 1):
        idx[np.logical_and(edges[k] <= u, edges[k + 1] > u)] = k
    return idx


def create_mie_scattering_data(simulation_parameters, rng=None):
    """rs02.py:699-772."""
    cam, pf = simulation_parameters["camera_design"], simulation_parameters["particle_field"]
    d, p = calculate_particle_diameter_distribution(simulation_parameters)
    idx = calculate_particle_diameter_indices(simulation_parameters, p, d, rng)
    angle, irradiance = _mie.calculate_mie_scattering_intensity(simulation_parameters, d)
    R = calculate_rotation_matrix(cam["x_camera_angle"], cam["y_camera_angle"], 0.0)
    beam = np.asarray(pf["beam_propogation_vector"], np.float64)
    return {"particle_diameter_vector": d, "particle_diameter_pdf": p, "particle_diameter_index_distribution": idx,
            "scattering_angle": angle, "scattering_irradiance": irradiance, "inverse_rotation_matrix": R.transpose(),
            "beam_propogation_vector": beam / np.linalg.norm(beam)}


def _first_lens(optical_system):
    return optical_system["design"]["optical_element"][0]["optical_element"][0]


def prepare_call(simulation_parameters, optical_system, pixel_gain, scattering_data, scattering_type, lightfield_source):
    """Everything perform_ray_tracing_03 works out before it calls the library (pr03.py:1949-2187) and the
    marshalling of prepare_data_for_cytpes_call (pr03.py:1631-1947), as a RayTracingCall."""
    sp = simulation_parameters
    ld, cd, out = sp["lens_design"], sp["camera_design"], sp["output_data"]
    dg = sp["density_gradients"]
    lens = _first_lens(optical_system)
    geo, prop = lens["element_geometry"], lens["element_properties"]
    f, od = float(ld["focal_length"]), float(ld["object_distance"])
    n, t = float(prop["refractive_index"]), float(geo["vertex_distance"])

    call = RayTracingCall()
    call.image_distance = (1.0 / f - 1.0 / od) ** (-1.0)                              # pr03.py:2016
    h2 = -(f * (n - 1.0) * t) / (float(geo["front_surface_radius"]) * n)               # pr03.py:2022
    v2 = call.image_distance + h2
    v1 = v2 + t
    z_sensor = 0.0
    if "perturbation" in ld:
        z_sensor += ld["perturbation"] * call.image_distance                           # pr03.py:2027-2033
    z_lens = (v1 + v2) / 2.0
    call.lens_pitch = f / float(ld["aperture_f_number"])
    call.aperture_f_number = float(ld["aperture_f_number"])
    call.beam_wavelength = float(sp["particle_field"]["beam_wavelength"]) if sp["simulation_type"] == "piv" else 0.0

    # one lens on the axis: centre (0, 0, 0) + z_lens, plane z = z_lens, first in the sequence (pr03.py:18-345, 2058-2078)
    call.element_center = np.array([[0.0, 0.0, z_lens]])
    call.element_plane_parameters = np.array([[0.0, 0.0, 1.0, -z_lens]])
    call.element_system_index = np.array([1], np.int32)
    lens_model = ld["lens_model"] if "lens_model" in ld else "general"
    el = element_data_struct()
    el.axial_offset_distances[0], el.axial_offset_distances[1] = [float(v) for v in lens["axial_offset_distances"]]
    el.element_number = lens["element_number"]
    el.element_type = {"thin-lens": b"t", "apparent": b"n"}.get(lens_model, b"l")      # pr03.py:1803-1808
    el.elements_coplanar = lens["elements_coplanar"]
    for j in range(3):
        el.rotation_angles[j] = float(lens["rotation_angles"][j])
    el.z_inter_element_distance = lens["z_inter_element_distance"]
    g = el.element_geometry
    g.front_surface_radius, g.front_surface_spherical = geo["front_surface_radius"], geo["front_surface_spherical"]
    g.back_surface_radius, g.back_surface_spherical = geo["back_surface_radius"], geo["back_surface_spherical"]
    g.pitch, g.vertex_distance = geo["pitch"], t
    p = el.element_properties
    p.abbe_number, p.absorbance_rate, p.refractive_index = prop["abbe_number"], prop["absorbance_rate"], n
    p.thin_lens_focal_length, p.transmission_ratio = prop["thin_lens_focal_length"], prop["transmission_ratio"]
    call.elements = [el]

    cam = camera_design_struct()
    cam.pixel_bit_depth, cam.pixel_gain, cam.pixel_pitch = int(cd["pixel_bit_depth"]), float(pixel_gain), float(cd["pixel_pitch"])
    cam.x_camera_angle, cam.y_camera_angle = float(cd["x_camera_angle"]), float(cd["y_camera_angle"])
    cam.x_pixel_number, cam.y_pixel_number = int(cd["x_pixel_number"]), int(cd["y_pixel_number"])
    cam.z_sensor = z_sensor
    cam.implement_diffraction = bool(cd.get("implement_diffraction", False))           # pr03.py:2113-2115
    cam.diffraction_diameter = float(cd.get("diffraction_diameter", 0.0)) if "implement_diffraction" in cd else 0.0
    R = calculate_rotation_matrix(cam.x_camera_angle, cam.y_camera_angle, 0.0)
    for i in range(3):
        for j in range(3):
            cam.rotation_matrix[i * 3 + j] = np.float32(R[i, j])
            cam.inverse_rotation_matrix[i * 3 + j] = np.float32(R[j, i])
    call.camera = cam

    call.scattering_type = scattering_type
    if scattering_type == "mie":                                                       # pr03.py:1660-1683
        call.mie_inverse_rotation = np.asarray(scattering_data["inverse_rotation_matrix"], np.float32).reshape(9)
        call.mie_beam = np.asarray(scattering_data["beam_propogation_vector"], np.float32)
        call.mie_angles = np.asarray(scattering_data["scattering_angle"], np.float32)
        irr = np.asarray(scattering_data["scattering_irradiance"], np.float32)
        call.mie_irradiance = irr.reshape(call.mie_angles.size, -1)

    src = lightfield_source
    call.x, call.y, call.z = (np.asarray(src[k], np.float32) for k in ("x", "y", "z"))
    call.radiance = np.asarray(src["radiance"], np.float64)
    call.diameter_index = np.asarray(src["diameter_index"]).astype(np.int32)
    call.lightray_number_per_particle = int(src["lightray_number_per_particle"])
    call.source_point_number = int(src["source_point_number"])
    call.z_offset, call.object_distance = float(src["z_offset"]), float(src["object_distance"])

    call.simulate_density_gradients = bool(dg["simulate_density_gradients"])           # pr03.py:2100-2109
    if call.simulate_density_gradients:
        call.density_grad_filename = os.path.abspath(str(dg["density_gradient_filename"]))
        call.ray_tracing_algorithm = int(dg["ray_tracing_algorithm"])
        if "add_pos_noise" in dg:
            call.add_pos_noise, call.pos_noise_std = bool(dg["add_pos_noise"]), float(dg["pos_noise_std"])
        if "add_ngrad_noise" in dg:
            call.add_ngrad_noise, call.ngrad_noise_std = bool(dg["add_ngrad_noise"]), float(dg["ngrad_noise_std"])
    call.ray_cone_pitch_ratio = float(ld["ray_cone_pitch_ratio"]) if "ray_cone_pitch_ratio" in ld else 1e-4
    call.save_lightrays = bool(out.get("save_lightrays", True))                        # pr03.py:2143-2162
    if call.save_lightrays:
        call.lightray_position_save_path = str(out["lightray_positions_filepath"])
        call.lightray_direction_save_path = str(out["lightray_directions_filepath"])
        call.num_lightrays_save = int(call.x.size) * call.lightray_number_per_particle
    call.save_intermediate_ray_data = bool(out.get("save_intermediate_ray_data", False))
    if call.save_intermediate_ray_data:
        call.num_intermediate_positions_save = int(out.get("num_intermediate_positions_save", 100))
    return call


def perform_ray_tracing_03(simulation_parameters, optical_system, pixel_gain, scattering_data, scattering_type,
                           lightfield_source, field_type=None, seed=0):
    """-> (I uint16, I_raw float32), pr03.py:1949-2259.  The trace is one ``start_ray_tracing`` call (so ray
    dumps land where the reference writes them); rescaling to the sensor's bit depth runs on the device
    (``photon_scene_export_u16``), image noise from a Philox stream keyed by ``seed``."""
    from .scene import ResidentScene, trace_image
    call = prepare_call(simulation_parameters, optical_system, pixel_gain, scattering_data, scattering_type, lightfield_source)
    I_raw = trace_image(call)
    cd, out = simulation_parameters["camera_design"], simulation_parameters["output_data"]
    points = call.subset(slice(0, 1))                       # the export needs the sensor, not the source points
    points.save_lightrays = False
    points.simulate_density_gradients = False
    scene = ResidentScene(points)
    try:
        scene.upload_image(I_raw)
        noise = float(cd.get("image_noise", 0.0))
        I = scene.export_u16(pixel_gain, int(cd["pixel_bit_depth"]), cd.get("intensity_rescaling", True), noise, seed)
        if noise > 0.0:
            I_raw = scene.download_image()                  # I_raw carries the noise afterwards (pr03.py:2203-2214)
    finally:
        scene.close()
    if out.get("crop_image", 0):                                                      # pr03.py:2245-2257
        nr_crop, nc_crop = int(out["r_crop"]), int(out["c_crop"])
        nr, nc = I.shape
        rows = slice(nr // 2 - nr_crop // 2, nr // 2 + nr_crop // 2 - 1)
        cols = slice(nc // 2 - nc_crop // 2, nc // 2 + nc_crop // 2 - 1)
        I, I_raw = I[rows, cols], I_raw[rows, cols]
    return I, I_raw


def _dump_dirs(simulation_parameters, image_directory, name):
    out = simulation_parameters["output_data"]
    out["lightray_positions_filepath"] = os.path.join(image_directory, "light-ray-positions", name) + "/"
    out["lightray_directions_filepath"] = os.path.join(image_directory, "light-ray-directions", name) + "/"
    for k in ("lightray_positions_filepath", "lightray_directions_filepath"):
        os.makedirs(out[k], exist_ok=True)


def run_simulation_02(simulation_parameters, rng=None, write_images=True):
    """The 'piv', 'cal' and 'bos' branches of rs02.py:1725-2110.  Returns a list of (name, I, I_raw); raw images
    are also written as ``<image_directory>/raw/<name>.bin`` like the reference (TIFF writing is left to the
    caller: the 16-bit arrays are what the reference hands to ``TIFF.imsave``)."""
    sp = simulation_parameters
    optical_system = create_camera_optical_system(sp)
    cd = sp["camera_design"]
    R = calculate_rotation_matrix(cd["x_camera_angle"], cd["y_camera_angle"], 0.0)
    cd["rotation_matrix"], cd["inverse_rotation_matrix"] = R, R.transpose()
    pixel_gain = cd["pixel_gain"]
    image_directory = sp["output_data"]["image_directory"]
    raw_dir = os.path.join(image_directory, "raw")
    if write_images:
        os.makedirs(raw_dir, exist_ok=True)
    results = []

    def emit(name, I, I_raw):
        if write_images:
            I_raw.tofile(os.path.join(raw_dir, name + ".bin"))
        results.append((name, I, I_raw))

    if sp["simulation_type"] == "piv":
        pf = sp["particle_field"]
        if not pf["generate_particle_field_images"]:
            return results
        if pf["perform_mie_scattering"]:
            scattering_data, scattering_type = create_mie_scattering_data(sp, rng), "mie"
        else:
            scattering_data, scattering_type = None, "diffuse"
        for frame_index in np.atleast_1d(pf["frame_vector"]):
            src = _sp.generate_piv_lightfield_data(sp, optical_system, R, rng)
            src["diameter_index"] = np.ones(src["x"].shape)                            # rs02.py:992
            _dump_dirs(sp, image_directory, "im%04d" % frame_index)
            I, I_raw = perform_ray_tracing_03(sp, optical_system, pixel_gain, scattering_data, scattering_type, src, "particle")
            emit("particle_image_frame_%04d" % frame_index, I, I_raw)
    elif sp["simulation_type"] == "bos":
        bp = sp["bos_pattern"]
        if not bp["generate_bos_pattern_images"]:
            return results
        src, gx, gy = _sp.generate_bos_lightfield_data(sp, optical_system, rng)
        bp["num_dots_generated"] = src["x"].size
        src["lightray_number_per_particle"] = bp["lightray_number_per_particle"]
        src["lightray_process_number"] = bp["lightray_process_number"]
        for number, with_gradients in ((1, False), (2, True)):                         # rs02.py:2013-2048
            sp["density_gradients"]["simulate_density_gradients"] = with_gradients
            _dump_dirs(sp, image_directory, "im%d" % number)
            I, I_raw = perform_ray_tracing_03(sp, optical_system, pixel_gain, float("nan"), "diffuse", src, "calibration")
            emit("bos_pattern_image_%d" % number, I, I_raw)
        sp["bos_pattern"]["x_grid_point_coordinate_vector"], sp["bos_pattern"]["y_grid_point_coordinate_vector"] = gx, gy
    elif sp["simulation_type"] == "cal":                                               # rs02.py:1877-1953
        cg = sp["calibration_grid"]
        if not cg["generate_calibration_grid_images"]:
            return results
        for plane_index in range(1, int(cg["calibration_plane_number"]) + 1):
            src = _sp.generate_calibration_lightfield_data(sp, optical_system, plane_index - 1, rng)
            src["lightray_number_per_particle"] = cg["lightray_number_per_particle"]
            src["lightray_process_number"] = cg["lightray_process_number"]
            if sp["output_data"].get("save_lightrays", True):
                _dump_dirs(sp, image_directory, "plane%02d" % plane_index)
            I, I_raw = perform_ray_tracing_03(sp, optical_system, pixel_gain, float("nan"), "diffuse", src, "calibration")
            emit("calibration_image_plane_%04d" % plane_index, I, I_raw)
    else:
        raise ValueError("simulation_type %r: 'piv', 'bos' and 'cal' are built" % (sp["simulation_type"],))
    return results
