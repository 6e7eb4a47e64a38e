// hostcheck.cpp -- TEST TOOL.  Compiles the product's per-ray stage code
// (photon_b200/csrc/trace_core.cuh, host/device inline functions) with g++ and drives it
// with a plain loop, so that the kernel's arithmetic can be compared with the oracle in the
// CPU container, before any GPU time is spent.  It is not part of the product library, is
// never loaded by it, and shares only headers with it.  The CUDA kernel's shared-memory
// tile / warp aggregation is NOT exercised here (that needs the GPU tests).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../photon_b200/csrc/nrrd_reader.h"
#include "../../photon_b200/csrc/scene_args.h"

using namespace ph;

namespace {
struct HostAcc {
    double *img; int nx;
    void add(int row, int col, float v) { img[(size_t)row * nx + col] += (double)v; }
    void add_linear(long idx, float v) { img[idx] += (double)v; }
    void add_slot(int, long idx, float v) { add_linear(idx, v); }
};
}

extern "C" long long g_hostcheck_steps = 0;
// the no-field kernel's queue rests on first_aperture_culls (ray_optics.cuh): a ray it culls must never reach the sensor.
// Counted here for every ray of a scene without a field: rays it culls, and culled rays that produced a sample anyway.
extern "C" long long g_hostcheck_culled = 0, g_hostcheck_cull_violations = 0, g_hostcheck_uncull_dead_in_train = 0;

extern "C" void start_ray_tracing(float lens_pitch, float image_distance, scattering_data_t *sc, char *scattering_type_str,
                                  lightfield_source_t *src, int rpp, float beam_wavelength, float f_number, int ne,
                                  double (*element_center)[3], element_data_t *element_data,
                                  double (*element_plane_parameters)[4], int *element_system_index, camera_design_t *cam,
                                  float *image_array, bool simulate_density_gradients, char *density_grad_filename,
                                  bool save_lightrays, char *pos_path, char *dir_path, int num_lightrays_save, int algorithm,
                                  bool add_pos_noise, float pos_noise_std, bool add_ngrad_noise, float ngrad_noise_std,
                                  float cone_ratio, bool save_intermediate, int num_intermediate) {
    TraceArgs a; memset(&a, 0, sizeof a);
    std::vector<f2> aim; std::string err;
    if (!fill_scene_args(a, aim, lens_pitch, image_distance, sc, scattering_type_str, src, rpp, beam_wavelength, f_number, ne,
                         element_center, element_data, element_plane_parameters, element_system_index, cam, algorithm,
                         add_pos_noise, pos_noise_std, add_ngrad_noise, ngrad_noise_std, cone_ratio, err)) {
        fprintf(stderr, "hostcheck: %s\n", err.c_str()); return;
    }
    a.sx = src->x; a.sy = src->y; a.sz = src->z; a.sradiance = src->radiance; a.sdiam = src->diameter_index;
    a.mie.angle = sc->scattering_angle; a.mie.irradiance = sc->scattering_irradiance;
    a.aim = aim.data();
    std::vector<f4> field;
    std::vector<float> coef;
    LinearField S{};
    BSplineField SC{};
    const char *ip = getenv("PHOTON_INTERPOLATION");
    const bool cubic = simulate_density_gradients && ip && strcmp(ip, "cubic") == 0;
    a.simulate_gradients = simulate_density_gradients;
    if (simulate_density_gradients) {
        pnrrd::Volume vol;
        if (!pnrrd::load(density_grad_filename, vol)) { fprintf(stderr, "hostcheck: %s\n", vol.error.c_str()); return; }
        int sx = vol.nx, sy = vol.ny, sz = vol.nz;
        double xmin = vol.origin[0], dx = vol.spacing[0], xmax = xmin + (sx - 1) * dx;
        double ymin = vol.origin[1], dy = vol.spacing[1], ymax = ymin + (sy - 1) * dy;
        double zmin = vol.origin[2] - 750e3, dz = vol.spacing[2], zmax = zmin + (sz - 1) * dz;
        if (sz > 1024) sz = 1024;
        FieldParams P{};
        P.mn = mk3((float)xmin, (float)ymin, (float)zmin); P.mx = mk3((float)xmax, (float)ymax, (float)zmax);
        P.scale = mk3(1.0f / (P.mx.x - P.mn.x), 1.0f / (P.mx.y - P.mn.y), 1.0f / (P.mx.z - P.mn.z));
        P.W = sx; P.H = sy; P.D = sz; P.Wf = (float)sx; P.Hf = (float)sy; P.Df = (float)sz;
        { double k[3] = {(double)P.scale.x * (sx - 2), (double)P.scale.y * (sy - 2), (double)P.scale.z * (sz - 2)};
          P.tk = mk3((float)k[0], (float)k[1], (float)k[2]);
          P.tc = mk3((float)(0.5 - (double)P.mn.x * k[0]), (float)(0.5 - (double)P.mn.y * k[1]), (float)(0.5 - (double)P.mn.z * k[2])); } P.wm2 = (float)(sx - 2); P.hm2 = (float)(sy - 2); P.dm2 = (float)(sz - 2);
        float step = (float)fmin(dx, dy); step = step < dz ? step : (float)dz; P.step = step;
        field.resize((size_t)sx * sy * sz);
        float dmin = INFINITY;
        const char *gi = getenv("PHOTON_GRADIENT_INDEX");
        const bool float_index = field.size() > ((size_t)1 << 24) && !(gi && strcmp(gi, "exact") == 0);
        for (int z = 0; z < sz; z++) for (int y = 0; y < sy; y++) for (int x = 0; x < sx; x++) {
            f4 o = float_index ? field_voxel<true>(vol.rho.data(), x, y, z, sx, sy, sz, (float)dx, (float)dy, (float)dz)
                               : field_voxel<false>(vol.rho.data(), x, y, z, sx, sy, sz, (float)dx, (float)dy, (float)dz);
            field[((size_t)z * sy + y) * sx + x] = o;
            if (o.w < dmin) dmin = o.w;
        }
        P.data_min = dmin; P.skip_step = P.step / (1 + P.data_min);
        S.g = field.data(); S.W = sx; S.H = sy; S.D = sz;
        if (cubic) {                                 // same ingest as driver.cu ensure_coefficients
            size_t W = sx, H = sy, D = sz;
            coef.resize(W * H * D);
            for (size_t i = 0; i < coef.size(); i++) coef[i] = field[i].w;
            for (size_t l = 0; l < H * D; l++) prefilter_line(&coef[l * W], W, 1);
            for (size_t l = 0; l < W * D; l++) prefilter_line(&coef[(l / W) * W * H + (l % W)], H, W);
            for (size_t l = 0; l < W * H; l++) prefilter_line(&coef[l], D, W * H);
            SC.c = coef.data(); SC.W = sx; SC.H = sy; SC.D = sz;
            SC.inv_sp = mk3(1.0f / (float)dx, 1.0f / (float)dy, 1.0f / (float)dz);
            P.data_min = P.data_min - 1e-4f * fabsf(P.data_min);
            P.skip_step = P.step / (1 + P.data_min);
        }
        a.field = P;
    }
    const int np = src->num_particles, nx = a.cam.nx, ny = a.cam.ny;
    std::vector<double> acc((size_t)nx * ny, 0.0);
    HostAcc hacc{acc.data(), nx};
    int spn = src->source_point_number; if (np < spn) spn = np;
    if (spn <= 0) return;
    int kmax = (np + spn - 1) / spn;
    const bool dumps = save_lightrays && num_lightrays_save > 0;
    std::vector<f3> fpos, fdir;
    if (dumps) { fpos.resize(num_lightrays_save); fdir.resize(num_lightrays_save); }
    const bool erf_model = a.implement_diffraction != 0 || a.train.e[0].type == 'n';
    const float erf_frac = a.train.e[0].type == 'n' ? 1.00f : 0.75f;
    unsigned long long steps = 0;
    for (int k = 0; k < kmax; k++) {
        int begin = k * spn, count = std::min(spn, np - begin);
        if (dumps) { for (auto &v : fpos) v = nan3(); for (auto &v : fdir) v = nan3(); }
        a.final_pos = dumps ? fpos.data() : nullptr; a.final_dir = dumps ? fdir.data() : nullptr;
        a.n_save = dumps ? num_lightrays_save : 0; a.dump_point_base = begin;
        for (int p = begin; p < begin + count; p++)
            for (int lr = 0; lr < rpp; lr++) {
                long long gid = (long long)lr + (long long)(p - begin) * rpp;
                bool save = dumps && gid < num_lightrays_save;
                SensorSample s = cubic ? trace_one<true>(a, SC, p, lr, gid, save, steps)
                                       : trace_one<true>(a, S, p, lr, gid, save, steps);
                if (!simulate_density_gradients && a.train.e[0].type != 'n') {
                    NormalStream ns0; Ray r0 = ray_prologue(a, p, lr, ns0);
                    const bool cull = first_aperture_culls(a.train, r0);
                    if (cull) { g_hostcheck_culled++; if (s.valid) g_hostcheck_cull_violations++; }
                    else { Ray r1 = r0; optical_train(a.train, r1); if (any_nan(r1.pos) || any_nan(r1.dir)) g_hostcheck_uncull_dead_in_train++; }
                }
                if (!s.valid) continue;
                if (erf_model) {
                    ErfFootprint fp; erf_footprint(fp, s.d_x, s.d_y, s.radiance, s.c4, a.cam.diffraction_diameter, erf_frac);
                    erf_splat(fp, a.cam, hacc);
                } else bilinear_splat(s.d_x, s.d_y, s.radiance, s.c4, a.cam, hacc);
            }
        if (dumps) {
            char name[32]; snprintf(name, sizeof name, "%04d.bin", k);
            std::string fp = std::string(pos_path) + "/pos_" + name, fd = std::string(dir_path) + "/dir_" + name;
            if (FILE *f = fopen(fp.c_str(), "wb")) { fwrite(fpos.data(), sizeof(f3), fpos.size(), f); fclose(f); }
            if (FILE *f = fopen(fd.c_str(), "wb")) { fwrite(fdir.data(), sizeof(f3), fdir.size(), f); fclose(f); }
        }
    }
    for (size_t i = 0; i < acc.size(); i++) image_array[i] = (float)((double)image_array[i] + acc[i]);
    g_hostcheck_steps = (long long)steps;
    (void)save_intermediate; (void)num_intermediate;
}
