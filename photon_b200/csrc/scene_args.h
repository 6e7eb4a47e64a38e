// scene_args.h -- flattens the ABI structs into the kernel argument block (host only).
// Replaces the argument marshalling of start_ray_tracing, prt.cu:3146-3296: lens-sample
// draws, double -> float element arrays, the stage order of the optical train, camera.
// Device pointers (sources, Mie table, aim table, field, image) are set by the caller.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/photon_abi.h"
#include "trace_core.cuh"

namespace ph {

// Morton (Z-order) rank of the lens aim points of one source point: rays that are neighbours in the processing order
// aim at neighbouring points of the aperture, so the 32 lanes of a warp march through neighbouring cells of the
// density field and their cell loads fall on the same lines (north-star (b): "reordering rays by Morton key").
// 16 bits per axis over the bounding box of the aim points; stable counting sort is overkill for <= 1e4 rays.
inline void morton_ray_order(const std::vector<f2> &aim, std::vector<int> &order) {
    const size_t n = aim.size();
    order.resize(n);
    if (n == 0) return;
    float x0 = aim[0].x, x1 = x0, y0 = aim[0].y, y1 = y0;
    for (auto &p : aim) { x0 = std::min(x0, p.x); x1 = std::max(x1, p.x); y0 = std::min(y0, p.y); y1 = std::max(y1, p.y); }
    const float kx = x1 > x0 ? 65535.0f / (x1 - x0) : 0.0f, ky = y1 > y0 ? 65535.0f / (y1 - y0) : 0.0f;
    auto spread = [](uint32_t v) { v &= 0xffffu; v = (v | (v << 8)) & 0x00ff00ffu; v = (v | (v << 4)) & 0x0f0f0f0fu;
                                   v = (v | (v << 2)) & 0x33333333u; v = (v | (v << 1)) & 0x55555555u; return v; };
    std::vector<std::pair<uint32_t, int>> key(n);
    for (size_t i = 0; i < n; i++) {
        const uint32_t qx = (uint32_t)std::min(65535.0f, std::max(0.0f, (aim[i].x - x0) * kx));
        const uint32_t qy = (uint32_t)std::min(65535.0f, std::max(0.0f, (aim[i].y - y0) * ky));
        key[i] = {spread(qx) | (spread(qy) << 1), (int)i};
    }
    std::stable_sort(key.begin(), key.end(), [](const std::pair<uint32_t, int> &a, const std::pair<uint32_t, int> &b) { return a.first < b.first; });
    for (size_t i = 0; i < n; i++) order[i] = key[i].second;
}

inline bool fill_scene_args(TraceArgs &a, std::vector<f2> &aim, float lens_pitch, float image_distance,
                            const scattering_data_t *sc, const char *scattering_type_str,
                            const lightfield_source_t *src, int rpp, float beam_wavelength, float f_number,
                            int ne, const double (*element_center)[3], const element_data_t *element_data,
                            const double (*element_plane_parameters)[4], const int *element_system_index,
                            const camera_design_t *cam, int algorithm, bool add_pos_noise, float pos_noise_std,
                            bool add_ngrad_noise, float ngrad_noise_std, float cone_ratio, std::string &err) {
    if (ne < 1 || ne > PH_MAX_ELEMENTS) { err = "num_elements must be in [1, 8]"; return false; }
    if (rpp < 1) { err = "lightray_number_per_particle must be >= 1"; return false; }
    a.rpp = rpp;

    // Mie table scalars (pointers are set by the caller): prt.cu:3192-3210
    a.mie.enabled = strcmp(scattering_type_str, "mie") == 0;
    if (a.mie.enabled) {
        if (sc->num_angles < 2 || sc->num_diameters < 1) { err = "Mie table needs >= 2 angles and >= 1 diameter"; return false; }
        memcpy(a.mie.inv_rot, sc->inverse_rotation_matrix, sizeof a.mie.inv_rot);
        memcpy(a.mie.beam, sc->beam_propagation_vector, sizeof a.mie.beam);
        a.mie.num_angles = sc->num_angles; a.mie.num_diameters = sc->num_diameters;
    }

    // lens samples: prt.cu:3216-3243 (glibc rand after srand(10), r1/r2 interleaved; the same
    // rpp draws serve every source point, prt.cu:2006).  The aim point c*pitch*r1*(cos,sin)(2 pi r2)
    // depends only on the draw, so it is evaluated once here (in double, as prt.cu:123-124 does).
    const char *rng = getenv("PHOTON_RNG");
    a.philox = rng && strcmp(rng, "philox") == 0;
    a.seed = 10;
    if (const char *sd = getenv("PHOTON_SEED")) a.seed = (unsigned)strtoul(sd, nullptr, 10);
    a.cone_ratio = cone_ratio; a.lens_pitch = lens_pitch;
    aim.clear();
    if (!a.philox) {
        aim.resize((size_t)rpp);
        static std::mutex rand_mutex;
        std::lock_guard<std::mutex> lk(rand_mutex);
        srand(10);
        for (int k = 0; k < rpp; k++) {
            float r1 = (float)((double)rand() / (RAND_MAX));
            float r2 = (float)((double)rand() / (RAND_MAX));
            lens_aim(cone_ratio, lens_pitch, r1, r2, aim[k].x, aim[k].y);
        }
    }

    // optical train: prt.cu:3250-3284 (double -> float) and 1274-1381 (stage order)
    for (int k = 0; k < ne; k++) {
        Element &E = a.train.e[k];
        const element_data_t &d = element_data[k];
        E.centre = mk3((float)element_center[k][0], (float)element_center[k][1], (float)element_center[k][2]);
        E.plane.x = (float)element_plane_parameters[k][0]; E.plane.y = (float)element_plane_parameters[k][1];
        E.plane.z = (float)element_plane_parameters[k][2]; E.plane.w = (float)element_plane_parameters[k][3];
        E.pitch = d.element_geometry.pitch; E.vertex = d.element_geometry.vertex_distance;
        E.r_front = d.element_geometry.front_surface_radius; E.r_back = d.element_geometry.back_surface_radius;
        E.n_glass = d.element_properties.refractive_index; E.abbe = d.element_properties.abbe_number;
        E.transmission = d.element_properties.transmission_ratio; E.absorbance = d.element_properties.absorbance_rate;
        E.focal = d.element_properties.thin_lens_focal_length; E.type = d.element_type;
    }
    int seq = 0;
    for (int k = 0; k < ne; k++) if (seq <= element_system_index[k]) seq = element_system_index[k];
    if (seq > PH_MAX_ELEMENTS) { err = "more than 8 sequential optical stages"; return false; }
    // The reference always propagates through element 0 (prt.cu:1331-1333); identical for the
    // single-element scenes it ships.  Default here: the element that owns the stage.
    const bool literal_e0 = getenv("PHOTON_LITERAL_ELEMENT0") != nullptr;
    a.train.stages = seq;
    for (int st = 0; st < seq; st++) {
        int cnt = 0, which = -1;
        for (int k = 0; k < ne; k++) if (seq - element_system_index[k] == st) { cnt++; which = k; }
        a.train.order[st] = cnt == 1 ? (literal_e0 ? 0 : which) : -1;
    }

    // camera: prt.cu:3291-3296, pixel (0,0) centre prt.cu:1441-1442
    Camera &c = a.cam;
    c.pitch = cam->pixel_pitch; c.z_sensor = cam->z_sensor; c.diffraction_diameter = cam->diffraction_diameter;
    c.nx = cam->x_pixel_number; c.ny = cam->y_pixel_number;
    memcpy(c.rot, cam->rotation_matrix, sizeof c.rot);
    memcpy(c.inv_rot, cam->inverse_rotation_matrix, sizeof c.inv_rot);
    c.p1x = (float)((double)(-c.pitch * (float)(c.nx - 1)) / 2.0);
    c.p1y = (float)((double)(-c.pitch * (float)(c.ny - 1)) / 2.0);
    a.implement_diffraction = cam->implement_diffraction ? 1 : 0;
    if (c.nx < 1 || c.ny < 1) { err = "sensor has no pixels"; return false; }
    if ((long long)c.nx * (long long)c.ny >= (1ll << 31)) { err = "sensor has 2^31 pixels or more (pixel indices are 32-bit)"; return false; }
    // any diffraction diameter renders (the reference has no bound either, prt.cu:1466-1540): footprints of more than
    // 8 rows leave the register path for one pair of erf per pixel (ray_optics.cuh: erf_splat)

    a.image_distance = image_distance; a.wavelength = beam_wavelength;
    a.inv_fnum2 = 1 / (f_number * f_number);                             // prt.cu:233
    a.z_shift = (double)src->z_offset + 750e3;                           // prt.cu:2045
    a.apparent_z_object = src->object_distance + src->z_offset;          // prt.cu:2145
    a.apparent_z_offset = src->z_offset;
    a.add_pos_noise = add_pos_noise; a.pos_noise_std = pos_noise_std;
    a.add_ngrad_noise = add_ngrad_noise; a.ngrad_noise_std = ngrad_noise_std;
    a.algorithm = algorithm;
    return true;
}

}  // namespace ph
