"""ctypes mirror of photon's C ABI (host side of the drop-in boundary).

Field names, order and types follow the *live* classes the reference defines inline in
``python_codes/perform_ray_tracing_03.py:1651-1852`` (which match
``cuda_codes/parallel_ray_tracing.h:17-191``); ``python_codes/ctypes_classes.py`` is a
stale, unparseable copy of an older ABI and is not used.  ``include/photon_abi.h`` holds
the C side with static_asserts on the same sizes/offsets.
"""
import ctypes as C

import numpy as np


class scattering_data_struct(C.Structure):
    # pr03.py:1651-1659 (the reference spells it "beam_propogation_vector")
    _fields_ = [
        ("inverse_rotation_matrix", C.c_float * 9),
        ("beam_propogation_vector", C.c_float * 3),
        ("scattering_angle", C.c_void_p),
        ("scattering_irradiance", C.c_void_p),
        ("num_angles", C.c_int),
        ("num_diameters", C.c_int),
    ]


class lightfield_source_struct(C.Structure):
    # pr03.py:1705-1717
    _fields_ = [
        ("lightray_number_per_particle", C.c_int),
        ("source_point_number", C.c_int),
        ("diameter_index", C.c_void_p),
        ("radiance", C.c_void_p),
        ("x", C.c_void_p),
        ("y", C.c_void_p),
        ("z", C.c_void_p),
        ("num_particles", C.c_int),
        ("z_offset", C.c_float),
        ("object_distance", C.c_float),
    ]


class element_geometry_struct(C.Structure):
    # pr03.py:1756-1764
    _fields_ = [
        ("front_surface_radius", C.c_float),
        ("front_surface_spherical", C.c_bool),
        ("back_surface_radius", C.c_float),
        ("back_surface_spherical", C.c_bool),
        ("pitch", C.c_float),
        ("vertex_distance", C.c_double),
    ]


class element_properties_struct(C.Structure):
    # pr03.py:1766-1773
    _fields_ = [
        ("abbe_number", C.c_float),
        ("absorbance_rate", C.c_float),
        ("refractive_index", C.c_double),
        ("thin_lens_focal_length", C.c_float),
        ("transmission_ratio", C.c_float),
    ]


class element_data_struct(C.Structure):
    # pr03.py:1775-1785
    _fields_ = [
        ("axial_offset_distances", C.c_double * 2),
        ("element_geometry", element_geometry_struct),
        ("element_number", C.c_float),
        ("element_properties", element_properties_struct),
        ("element_type", C.c_char),
        ("elements_coplanar", C.c_float),
        ("rotation_angles", C.c_double * 3),
        ("z_inter_element_distance", C.c_float),
    ]


class camera_design_struct(C.Structure):
    # pr03.py:1838-1852
    _fields_ = [
        ("pixel_bit_depth", C.c_int),
        ("pixel_gain", C.c_float),
        ("pixel_pitch", C.c_float),
        ("x_camera_angle", C.c_float),
        ("y_camera_angle", C.c_float),
        ("x_pixel_number", C.c_int),
        ("y_pixel_number", C.c_int),
        ("z_sensor", C.c_float),
        ("diffraction_diameter", C.c_float),
        ("implement_diffraction", C.c_bool),
        ("rotation_matrix", C.c_float * 9),
        ("inverse_rotation_matrix", C.c_float * 9),
    ]


EXPECTED_SIZES = {
    scattering_data_struct: 72,
    lightfield_source_struct: 64,
    element_geometry_struct: 32,
    element_properties_struct: 24,
    element_data_struct: 120,
    camera_design_struct: 112,
}
for _cls, _sz in EXPECTED_SIZES.items():
    assert C.sizeof(_cls) == _sz, (_cls.__name__, C.sizeof(_cls), _sz)

# pr03.py:1914-1921
START_RAY_TRACING_ARGTYPES = [
    C.c_float, C.c_float, C.POINTER(scattering_data_struct), C.c_char_p,
    C.POINTER(lightfield_source_struct), C.c_int, C.c_float, C.c_float, C.c_int,
    C.c_void_p, C.POINTER(element_data_struct), C.c_void_p, C.c_void_p,
    C.POINTER(camera_design_struct), C.c_void_p,
    C.c_bool, C.c_char_p, C.c_bool, C.c_char_p, C.c_char_p,
    C.c_int, C.c_int, C.c_bool, C.c_float, C.c_bool, C.c_float, C.c_float, C.c_bool, C.c_int,
]


def bind_start_ray_tracing(lib):
    """Set argtypes/restype on ``lib.start_ray_tracing`` the way pr03.py:1914-1925 does."""
    fn = lib.start_ray_tracing
    fn.argtypes = START_RAY_TRACING_ARGTYPES
    fn.restype = None
    return fn


def _ptr(a):
    return a.ctypes.data if a is not None and a.size else None


class RayTracingCall:
    """All 29 arguments of one ``start_ray_tracing`` call, held as numpy arrays / scalars.

    This is the host-side object the tests, the bench and the recordings share: it can be
    built from a scene description, loaded from a recording of the unchanged reference
    scripts, and replayed into any library exporting the ABI.
    """

    SCALARS = (
        "lens_pitch", "image_distance", "scattering_type", "lightray_number_per_particle",
        "beam_wavelength", "aperture_f_number", "simulate_density_gradients",
        "density_grad_filename", "save_lightrays", "lightray_position_save_path",
        "lightray_direction_save_path", "num_lightrays_save", "ray_tracing_algorithm",
        "add_pos_noise", "pos_noise_std", "add_ngrad_noise", "ngrad_noise_std",
        "ray_cone_pitch_ratio", "save_intermediate_ray_data", "num_intermediate_positions_save",
        "source_point_number", "z_offset", "object_distance",
    )

    def __init__(self):
        self.lens_pitch = 0.0
        self.image_distance = 0.0
        self.scattering_type = "diffuse"
        self.mie_inverse_rotation = np.full(9, np.nan, np.float32)
        self.mie_beam = np.full(3, np.nan, np.float32)
        self.mie_angles = np.zeros(0, np.float32)
        self.mie_irradiance = np.zeros((0, 0), np.float32)  # [angle][diameter]
        self.lightray_number_per_particle = 1
        self.source_point_number = 10000
        self.x = np.zeros(0, np.float32)
        self.y = np.zeros(0, np.float32)
        self.z = np.zeros(0, np.float32)
        self.radiance = np.zeros(0, np.float64)
        self.diameter_index = np.zeros(0, np.int32)
        self.z_offset = 0.0
        self.object_distance = 0.0
        self.beam_wavelength = 0.0
        self.aperture_f_number = 1.0
        self.element_center = np.zeros((0, 3), np.float64)
        self.element_plane_parameters = np.zeros((0, 4), np.float64)
        self.element_system_index = np.zeros(0, np.int32)
        self.elements = []  # list of element_data_struct
        self.camera = camera_design_struct()
        self.simulate_density_gradients = False
        self.density_grad_filename = ""
        self.save_lightrays = False
        self.lightray_position_save_path = ""
        self.lightray_direction_save_path = ""
        self.num_lightrays_save = 0
        self.ray_tracing_algorithm = 0
        self.add_pos_noise = False
        self.pos_noise_std = 0.0
        self.add_ngrad_noise = False
        self.ngrad_noise_std = 0.0
        self.ray_cone_pitch_ratio = 1.0
        self.save_intermediate_ray_data = False
        self.num_intermediate_positions_save = 0

    # ------------------------------------------------------------------
    @property
    def num_particles(self):
        return int(self.diameter_index.size)

    @property
    def num_rays(self):
        return self.num_particles * int(self.lightray_number_per_particle)

    @property
    def image_shape(self):
        return int(self.camera.y_pixel_number), int(self.camera.x_pixel_number)

    def subset(self, idx):
        """A shallow copy keeping only source points ``idx`` (array of indices or slice)."""
        import copy
        o = copy.copy(self)
        for k in ("x", "y", "z", "radiance", "diameter_index"):
            setattr(o, k, np.ascontiguousarray(getattr(self, k)[idx]))
        return o

    def marshal(self, image):
        """Build the ctypes argument tuple.  Returns (args, keepalive)."""
        assert image.dtype == np.float32 and image.flags["C_CONTIGUOUS"]
        sc = scattering_data_struct()
        ang = np.ascontiguousarray(self.mie_angles, np.float32)
        irr = np.ascontiguousarray(self.mie_irradiance, np.float32)
        for i in range(9):
            sc.inverse_rotation_matrix[i] = float(self.mie_inverse_rotation[i])
        for i in range(3):
            sc.beam_propogation_vector[i] = float(self.mie_beam[i])
        sc.scattering_angle = _ptr(ang)
        sc.scattering_irradiance = _ptr(irr)
        sc.num_angles = int(irr.shape[0]) if irr.ndim == 2 else 0
        sc.num_diameters = int(irr.shape[1]) if irr.ndim == 2 else 0

        x = np.ascontiguousarray(self.x, np.float32)
        y = np.ascontiguousarray(self.y, np.float32)
        z = np.ascontiguousarray(self.z, np.float32)
        rad = np.ascontiguousarray(self.radiance, np.float64)
        di = np.ascontiguousarray(self.diameter_index, np.int32)
        src = lightfield_source_struct()
        src.lightray_number_per_particle = int(self.lightray_number_per_particle)
        src.source_point_number = int(self.source_point_number)
        src.diameter_index = _ptr(di)
        src.radiance = _ptr(rad)
        src.x, src.y, src.z = _ptr(x), _ptr(y), _ptr(z)
        src.num_particles = int(di.size)
        src.z_offset = float(self.z_offset)
        src.object_distance = float(self.object_distance)

        ne = len(self.elements)
        elems = (element_data_struct * max(ne, 1))(*self.elements)
        ec = np.ascontiguousarray(self.element_center, np.float64).reshape(ne, 3)
        ep = np.ascontiguousarray(self.element_plane_parameters, np.float64).reshape(ne, 4)
        esi = np.ascontiguousarray(self.element_system_index, np.int32)
        cam = camera_design_struct.from_buffer_copy(bytes(self.camera))
        args = (
            C.c_float(self.lens_pitch), C.c_float(self.image_distance), C.pointer(sc),
            self.scattering_type.encode("utf-8"), C.pointer(src),
            int(self.lightray_number_per_particle), C.c_float(self.beam_wavelength),
            C.c_float(self.aperture_f_number), ne, ec.ctypes.data, elems, ep.ctypes.data,
            esi.ctypes.data, C.pointer(cam), image.ctypes.data,
            bool(self.simulate_density_gradients), self.density_grad_filename.encode("utf-8"),
            bool(self.save_lightrays), self.lightray_position_save_path.encode("utf-8"),
            self.lightray_direction_save_path.encode("utf-8"), int(self.num_lightrays_save),
            int(self.ray_tracing_algorithm), bool(self.add_pos_noise), C.c_float(self.pos_noise_std),
            bool(self.add_ngrad_noise), C.c_float(self.ngrad_noise_std),
            C.c_float(self.ray_cone_pitch_ratio), bool(self.save_intermediate_ray_data),
            int(self.num_intermediate_positions_save),
        )
        keep = (sc, ang, irr, x, y, z, rad, di, src, elems, ec, ep, esi, cam, image)
        return args, keep

    def run(self, lib, image=None):
        """Call ``lib.start_ray_tracing`` and return the (ny, nx) float32 image."""
        ny, nx = self.image_shape
        if image is None:
            image = np.zeros(ny * nx, np.float32)
        else:
            image = np.ascontiguousarray(image, np.float32).reshape(-1).copy()
        fn = bind_start_ray_tracing(lib)
        args, keep = self.marshal(image)
        fn(*args)
        del keep
        return image.reshape(ny, nx)
