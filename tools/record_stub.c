/* record_stub.c -- a libparallel_ray_tracing.so stand-in that RECORDS the 29 arguments the unchanged
 * reference scripts pass to start_ray_tracing (pr03.py:1914-1938) and then, optionally, forwards the
 * call to a real implementation (PHOTON_RECORD_FORWARD=/path/to/lib.so).
 *
 * Output: $PHOTON_RECORD_DIR/call_%02d/{scalars.txt, *.bin}; photon_b200.recording.load_call() turns a
 * directory back into a RayTracingCall.  Test tooling only (tools/stage_reference.py).
 *
 * build: gcc -O1 -fPIC -shared -o libparallel_ray_tracing.so tools/record_stub.c -ldl
 */
#include <dlfcn.h>
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include "../include/photon_abi.h"

static int g_call = 0;

static void dump(const char *dir, const char *name, const void *p, size_t bytes) {
    char path[4096];
    snprintf(path, sizeof path, "%s/%s.bin", dir, name);
    FILE *f = fopen(path, "wb");
    if (!f) return;
    if (p && bytes) fwrite(p, 1, bytes, f);
    fclose(f);
}

typedef void (*start_fn)(float, float, scattering_data_t *, char *, lightfield_source_t *, int, float, float, int,
                         double (*)[3], element_data_t *, double (*)[4], int *, camera_design_t *, float *, bool,
                         char *, bool, char *, char *, int, int, bool, float, bool, float, float, bool, int);

void start_ray_tracing(float lens_pitch, float image_distance, scattering_data_t *scattering_data_p,
                       char *scattering_type_str, lightfield_source_t *lightfield_source_p,
                       int lightray_number_per_particle, float beam_wavelength, float aperture_f_number,
                       int num_elements, double (*element_center)[3], element_data_t *element_data_p,
                       double (*element_plane_parameters)[4], int *element_system_index,
                       camera_design_t *camera_design_p, float *image_array, bool simulate_density_gradients,
                       char *density_grad_filename, bool save_lightrays, char *lightray_position_save_path,
                       char *lightray_direction_save_path, int num_lightrays_save, int ray_tracing_algorithm,
                       bool add_pos_noise, float pos_noise_std, bool add_ngrad_noise, float ngrad_noise_std,
                       float ray_cone_pitch_ratio, bool save_intermediate_ray_data,
                       int num_intermediate_positions_save) {
    const char *root = getenv("PHOTON_RECORD_DIR");
    if (root) {
        char dir[4096], path[4200];
        mkdir(root, 0777);
        snprintf(dir, sizeof dir, "%s/call_%02d", root, g_call);
        mkdir(dir, 0777);
        snprintf(path, sizeof path, "%s/scalars.txt", dir);
        FILE *f = fopen(path, "w");
        const lightfield_source_t *s = lightfield_source_p;
        const scattering_data_t *m = scattering_data_p;
        if (f) {
            fprintf(f, "lens_pitch %.9g\nimage_distance %.9g\nscattering_type %s\n", lens_pitch, image_distance,
                    scattering_type_str);
            fprintf(f, "lightray_number_per_particle %d\nbeam_wavelength %.9g\naperture_f_number %.9g\n",
                    lightray_number_per_particle, beam_wavelength, aperture_f_number);
            fprintf(f, "num_elements %d\nsimulate_density_gradients %d\ndensity_grad_filename %s\n", num_elements,
                    (int)simulate_density_gradients, density_grad_filename);
            fprintf(f, "save_lightrays %d\nlightray_position_save_path %s\nlightray_direction_save_path %s\n",
                    (int)save_lightrays, lightray_position_save_path, lightray_direction_save_path);
            fprintf(f, "num_lightrays_save %d\nray_tracing_algorithm %d\nadd_pos_noise %d\npos_noise_std %.9g\n",
                    num_lightrays_save, ray_tracing_algorithm, (int)add_pos_noise, pos_noise_std);
            fprintf(f, "add_ngrad_noise %d\nngrad_noise_std %.9g\nray_cone_pitch_ratio %.9g\n", (int)add_ngrad_noise,
                    ngrad_noise_std, ray_cone_pitch_ratio);
            fprintf(f, "save_intermediate_ray_data %d\nnum_intermediate_positions_save %d\n",
                    (int)save_intermediate_ray_data, num_intermediate_positions_save);
            fprintf(f, "source_point_number %d\nnum_particles %d\nz_offset %.9g\nobject_distance %.9g\n",
                    s->source_point_number, s->num_particles, s->z_offset, s->object_distance);
            fprintf(f, "source_lightray_number_per_particle %d\n", s->lightray_number_per_particle);
            fprintf(f, "num_angles %d\nnum_diameters %d\n", m->num_angles, m->num_diameters);
            fprintf(f, "sizeof_scattering %zu\nsizeof_source %zu\nsizeof_element %zu\nsizeof_camera %zu\n",
                    sizeof(scattering_data_t), sizeof(lightfield_source_t), sizeof(element_data_t),
                    sizeof(camera_design_t));
            fclose(f);
        }
        size_t np = (size_t)s->num_particles;
        dump(dir, "x", s->x, np * 4); dump(dir, "y", s->y, np * 4); dump(dir, "z", s->z, np * 4);
        dump(dir, "radiance", s->radiance, np * 8); dump(dir, "diameter_index", s->diameter_index, np * 4);
        dump(dir, "mie_inverse_rotation", m->inverse_rotation_matrix, 36);
        dump(dir, "mie_beam", m->beam_propagation_vector, 12);
        size_t na = m->scattering_angle ? (size_t)m->num_angles : 0, nd = (size_t)m->num_diameters;
        dump(dir, "mie_angles", m->scattering_angle, na * 4);
        dump(dir, "mie_irradiance", m->scattering_irradiance, m->scattering_irradiance ? na * nd * 4 : 0);
        dump(dir, "element_center", element_center, (size_t)num_elements * 24);
        dump(dir, "element_plane_parameters", element_plane_parameters, (size_t)num_elements * 32);
        dump(dir, "element_system_index", element_system_index, (size_t)num_elements * 4);
        dump(dir, "elements", element_data_p, (size_t)num_elements * sizeof(element_data_t));
        dump(dir, "camera", camera_design_p, sizeof(camera_design_t));
        dump(dir, "image_in", image_array,
             (size_t)camera_design_p->x_pixel_number * camera_design_p->y_pixel_number * 4);
    }
    const char *fwd = getenv("PHOTON_RECORD_FORWARD");
    if (fwd && fwd[0]) {
        void *h = dlopen(fwd, RTLD_NOW | RTLD_LOCAL);
        start_fn fn = h ? (start_fn)dlsym(h, "start_ray_tracing") : NULL;
        if (!fn) { fprintf(stderr, "record_stub: cannot forward to %s: %s\n", fwd, dlerror()); }
        else fn(lens_pitch, image_distance, scattering_data_p, scattering_type_str, lightfield_source_p,
                lightray_number_per_particle, beam_wavelength, aperture_f_number, num_elements, element_center,
                element_data_p, element_plane_parameters, element_system_index, camera_design_p, image_array,
                simulate_density_gradients, density_grad_filename, save_lightrays, lightray_position_save_path,
                lightray_direction_save_path, num_lightrays_save, ray_tracing_algorithm, add_pos_noise,
                pos_noise_std, add_ngrad_noise, ngrad_noise_std, ray_cone_pitch_ratio,
                save_intermediate_ray_data, num_intermediate_positions_save);
    } else {
        image_array[0] = 1.0f;      /* the scripts divide by the image maximum (pr03.py:2236) */
    }
    if (root) {
        char dir[4096];
        snprintf(dir, sizeof dir, "%s/call_%02d", root, g_call);
        if (fwd && fwd[0])
            dump(dir, "image_out", image_array,
                 (size_t)camera_design_p->x_pixel_number * camera_design_p->y_pixel_number * 4);
    }
    g_call++;
}
