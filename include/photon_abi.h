/*
 * photon_abi.h -- the C ABI that photon's Python driver binds with ctypes.
 *
 * This is the drop-in boundary of libparallel_ray_tracing.so.  Every struct
 * below must stay byte-for-byte identical to the reference's
 *   cuda_codes/parallel_ray_tracing.h:17-191           (C side)
 *   python_codes/perform_ray_tracing_03.py:1651-1852   (live ctypes mirror)
 * The sizes and offsets are pinned with static_asserts (values observed with
 * ctypes.sizeof / ctypes offset on the pr03.py classes).
 *
 * Units: micrometres everywhere; angles in radians unless noted.
 * Ownership: the caller owns every buffer; the callee keeps no pointer after
 * it returns.
 */
#ifndef PHOTON_ABI_H_
#define PHOTON_ABI_H_

#include <stddef.h>
#ifndef __cplusplus
#include <stdbool.h>
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Mie scattering table (prt.h:17-33).  irradiance[angle * num_diameters + diameter]. */
typedef struct scattering_data_t {
    float  inverse_rotation_matrix[9];   /* row-major 3x3, camera -> world */
    float  beam_propagation_vector[3];
    float *scattering_angle;             /* [num_angles], radians, uniform spacing */
    float *scattering_irradiance;        /* [num_angles][num_diameters] */
    int    num_angles;
    int    num_diameters;
} scattering_data_t;

/* Source points: PIV particles or BOS dot sub-points (prt.h:36-60). */
typedef struct lightfield_source_t {
    int     lightray_number_per_particle;
    int     source_point_number;         /* points per batch (dump-file granularity) */
    int    *diameter_index;              /* [num_particles] column of the Mie table */
    double *radiance;                    /* [num_particles] */
    float  *x;
    float  *y;
    float  *z;
    int     num_particles;
    float   z_offset;
    float   object_distance;
} lightfield_source_t;

/* prt.h:115-131 */
typedef struct element_geometry_t {
    float  front_surface_radius;
    bool   front_surface_spherical;
    float  back_surface_radius;
    bool   back_surface_spherical;
    float  pitch;
    double vertex_distance;
} element_geometry_t;

/* prt.h:134-141 */
typedef struct element_properties_t {
    float  abbe_number;
    float  absorbance_rate;
    double refractive_index;
    float  thin_lens_focal_length;
    float  transmission_ratio;
} element_properties_t;

/* prt.h:144-162.  element_type: 't' thin lens, 'l' thick lens, 'n' apparent image,
 * anything else = aperture. */
typedef struct element_data_t {
    double               axial_offset_distances[2];
    element_geometry_t   element_geometry;
    float                element_number;
    element_properties_t element_properties;
    char                 element_type;
    float                elements_coplanar;
    double               rotation_angles[3];
    float                z_inter_element_distance;
} element_data_t;

/* prt.h:165-191 */
typedef struct camera_design_t {
    int   pixel_bit_depth;
    float pixel_gain;
    float pixel_pitch;
    float x_camera_angle;
    float y_camera_angle;
    int   x_pixel_number;
    int   y_pixel_number;
    float z_sensor;
    float diffraction_diameter;          /* pixels */
    bool  implement_diffraction;
    float rotation_matrix[9];            /* row-major, volume frame -> camera frame */
    float inverse_rotation_matrix[9];
} camera_design_t;

#ifdef __cplusplus
#define PHOTON_ABI_ASSERT(c, m) static_assert(c, m)
#else
#define PHOTON_ABI_ASSERT(c, m) _Static_assert(c, m)
#endif
PHOTON_ABI_ASSERT(sizeof(scattering_data_t) == 72, "scattering_data_t");
PHOTON_ABI_ASSERT(offsetof(scattering_data_t, beam_propagation_vector) == 36, "scattering.beam");
PHOTON_ABI_ASSERT(offsetof(scattering_data_t, scattering_angle) == 48, "scattering.angle");
PHOTON_ABI_ASSERT(offsetof(scattering_data_t, scattering_irradiance) == 56, "scattering.irr");
PHOTON_ABI_ASSERT(offsetof(scattering_data_t, num_angles) == 64, "scattering.na");
PHOTON_ABI_ASSERT(offsetof(scattering_data_t, num_diameters) == 68, "scattering.nd");
PHOTON_ABI_ASSERT(sizeof(lightfield_source_t) == 64, "lightfield_source_t");
PHOTON_ABI_ASSERT(offsetof(lightfield_source_t, diameter_index) == 8, "source.diam");
PHOTON_ABI_ASSERT(offsetof(lightfield_source_t, radiance) == 16, "source.radiance");
PHOTON_ABI_ASSERT(offsetof(lightfield_source_t, x) == 24, "source.x");
PHOTON_ABI_ASSERT(offsetof(lightfield_source_t, z) == 40, "source.z");
PHOTON_ABI_ASSERT(offsetof(lightfield_source_t, num_particles) == 48, "source.np");
PHOTON_ABI_ASSERT(offsetof(lightfield_source_t, z_offset) == 52, "source.zoff");
PHOTON_ABI_ASSERT(offsetof(lightfield_source_t, object_distance) == 56, "source.od");
PHOTON_ABI_ASSERT(sizeof(element_geometry_t) == 32, "element_geometry_t");
PHOTON_ABI_ASSERT(offsetof(element_geometry_t, back_surface_radius) == 8, "geom.rb");
PHOTON_ABI_ASSERT(offsetof(element_geometry_t, pitch) == 16, "geom.pitch");
PHOTON_ABI_ASSERT(offsetof(element_geometry_t, vertex_distance) == 24, "geom.vd");
PHOTON_ABI_ASSERT(sizeof(element_properties_t) == 24, "element_properties_t");
PHOTON_ABI_ASSERT(offsetof(element_properties_t, refractive_index) == 8, "prop.n");
PHOTON_ABI_ASSERT(offsetof(element_properties_t, thin_lens_focal_length) == 16, "prop.f");
PHOTON_ABI_ASSERT(sizeof(element_data_t) == 120, "element_data_t");
PHOTON_ABI_ASSERT(offsetof(element_data_t, element_geometry) == 16, "elem.geom");
PHOTON_ABI_ASSERT(offsetof(element_data_t, element_number) == 48, "elem.num");
PHOTON_ABI_ASSERT(offsetof(element_data_t, element_properties) == 56, "elem.prop");
PHOTON_ABI_ASSERT(offsetof(element_data_t, element_type) == 80, "elem.type");
PHOTON_ABI_ASSERT(offsetof(element_data_t, elements_coplanar) == 84, "elem.coplanar");
PHOTON_ABI_ASSERT(offsetof(element_data_t, rotation_angles) == 88, "elem.rot");
PHOTON_ABI_ASSERT(offsetof(element_data_t, z_inter_element_distance) == 112, "elem.zi");
PHOTON_ABI_ASSERT(sizeof(camera_design_t) == 112, "camera_design_t");
PHOTON_ABI_ASSERT(offsetof(camera_design_t, x_pixel_number) == 20, "cam.nx");
PHOTON_ABI_ASSERT(offsetof(camera_design_t, z_sensor) == 28, "cam.zs");
PHOTON_ABI_ASSERT(offsetof(camera_design_t, implement_diffraction) == 36, "cam.diff");
PHOTON_ABI_ASSERT(offsetof(camera_design_t, rotation_matrix) == 40, "cam.R");
PHOTON_ABI_ASSERT(offsetof(camera_design_t, inverse_rotation_matrix) == 76, "cam.Rinv");

/*
 * The one entry point photon's Python calls (replaces
 * cuda_codes/parallel_ray_tracing.cu:3078-3775; bound at
 * python_codes/perform_ray_tracing_03.py:1914-1938).
 *
 * Traces num_particles * lightray_number_per_particle rays and ACCUMULATES their
 * radiance onto image_array[y_pixel_number * x_pixel_number] (row-major).
 * No return value: failures are reported on stderr and leave image_array untouched.
 */
void start_ray_tracing(float lens_pitch, float image_distance,
                       scattering_data_t *scattering_data, char *scattering_type_str,
                       lightfield_source_t *lightfield_source, int lightray_number_per_particle,
                       float beam_wavelength, float aperture_f_number,
                       int num_elements, double (*element_center)[3], element_data_t *element_data,
                       double (*element_plane_parameters)[4], int *element_system_index,
                       camera_design_t *camera_design, float *image_array,
                       bool simulate_density_gradients, char *density_grad_filename,
                       bool save_lightrays, char *lightray_position_save_path,
                       char *lightray_direction_save_path,
                       int num_lightrays_save, int ray_tracing_algorithm,
                       bool add_pos_noise, float pos_noise_std,
                       bool add_ngrad_noise, float ngrad_noise_std,
                       float ray_cone_pitch_ratio,
                       bool save_intermediate_ray_data, int num_intermediate_positions_save);

#ifdef __cplusplus
}
#endif
#endif /* PHOTON_ABI_H_ */
