// trace_core.cuh -- kernel argument block and the per-ray pipeline up to the sensor plane.
// Host/device: the CUDA kernel (trace_kernel.cuh) and the CPU host-check used by the tests
// (tests/hostcheck) compile this same code.
#pragma once
#include "ray_stages.cuh"

#if defined(__CUDA_ARCH__)
#define PH_LDG(p) __ldg(p)
#else
#define PH_LDG(p) (*(p))
#endif

namespace ph {

struct f2 { float x, y; };

struct TraceArgs {
    // scene
    float image_distance, wavelength, inv_fnum2, cone_ratio, lens_pitch;
    MieTable mie;
    OpticalTrain train;
    Camera cam;
    FieldParams field;
    int simulate_gradients, algorithm;
    double z_shift;                      // z_offset + 750e3 (double, prt.cu:2045)
    float apparent_z_object, apparent_z_offset;   // lens model 'n'
    int implement_diffraction;
    // sources (device arrays)
    const float *sx, *sy, *sz;
    const double *sradiance;
    const int *sdiam;
    int point_begin, point_count, rpp;
    const int *order;                    // processing order: CTA slot i traces point order[i] (null = i)
    int dump_point_base;                 // first point of the dump batch (batch-local ray ids)
    int global_point_offset;             // index of shard point 0 in the caller's arrays (RNG keys)
    // lens samples: table of aim points [rpp] (reference draws) or Philox per ray
    const f2 *aim;
    const int *ray_order;                // slot k of a source point traces ray ray_order[k] (Morton order of the aim points), or null
    int philox;
    unsigned int seed;
    // noise (reference: cuRAND XORWOW seeded with time(); here Philox streams)
    int add_pos_noise, add_ngrad_noise;
    float pos_noise_std, ngrad_noise_std;
    // outputs
    float *image;
    unsigned long long *image_fixed;     // deterministic mode: 64-bit fixed-point image, else null
    double fixed_scale;                  // 2^k: increments are rounded to multiples of 1/fixed_scale
    f3 *final_pos, *final_dir;           // [n_save] or null
    long long n_save;
    f3 *inter_pos, *inter_dir;           // [n_save][n_inter] or null
    int n_inter;
    unsigned long long *counters;        // [0] committed ray-steps, [1] rays that reached the sensor
    // dot statistics (the on-device form of light_ray_processing.py:211-275): per segment of rays_per_dot
    // consecutive rays, [0..2] sum of sensor positions, [3..5] sum of arccos(direction after the volume),
    // [6] rays with a sensor position, [7] rays with a direction.  null = off.
    double *dot_stats;
    long long rays_per_dot, n_dots;
};

// Result of tracing one ray up to the sensor plane
struct SensorSample {
    float d_x, d_y;
    double radiance, c4;
    bool valid;
    f3 sdir, spos;                       // STATS only: what the ray dumps would hold (prt.cu:2136-2141, 2236-2240)
    int sflags;                          // bit 0: sdir set, bit 1: spos set
};

// The per-ray pipeline in three pieces: generation (+ the move into the volume frame), the march, everything
// after it (back to the camera frame, lens system, sensor).
// aim point of ray lr of source point p on the lens, and the ray's global index (the key of its random streams)
PH_HD void ray_aim(const TraceArgs &a, int p, int lr, float &x_lens, float &y_lens, unsigned long long &ray_index) {
    ray_index = (unsigned long long)(p + a.global_point_offset) * (unsigned long long)a.rpp + (unsigned long long)lr;
    if (a.rpp == 1) { x_lens = 0.0f; y_lens = 0.0f; }
    else if (a.philox) {
        uint32_t c[4] = {(uint32_t)ray_index, (uint32_t)(ray_index >> 32), 0u, 0u};
        philox4x32_10(c, a.seed, 0x50484f54u);
        lens_aim(a.cone_ratio, a.lens_pitch, u01_24(c[0]), u01_24(c[1]), x_lens, y_lens);
    } else { x_lens = PH_LDG(&a.aim[lr].x); y_lens = PH_LDG(&a.aim[lr].y); }
}
PH_HD Ray ray_prologue(const TraceArgs &a, int p, int lr, NormalStream &ns) {
    float x_lens, y_lens;
    unsigned long long ray_index;
    ray_aim(a, p, lr, x_lens, y_lens, ray_index);
    ns.ray = ray_index; ns.ctr = 0; ns.seed = a.seed;
    Ray ray = generate_ray(a.image_distance, a.mie, PH_LDG(&a.sx[p]), PH_LDG(&a.sy[p]), PH_LDG(&a.sz[p]),
                           PH_LDG(&a.sradiance[p]), PH_LDG(&a.sdiam[p]), a.wavelength, a.inv_fnum2, x_lens, y_lens);
    if (a.simulate_gradients) {                                          // prt.cu:2033-2076
        f3 rp = ray.pos;
        rp.z = (float)((double)rp.z - a.z_shift);
        ray.pos = mat_mul(a.cam.inv_rot, rp);
        ray.dir = mat_mul(a.cam.inv_rot, ray.dir);
    }
    return ray;
}

template <bool DUMP, bool STATS>
PH_HD SensorSample ray_epilogue(const TraceArgs &a, Ray &ray, NormalStream &ns, long long gid, bool save, const f3 *entry_hit = nullptr) {
    SensorSample out; out.valid = false; out.d_x = out.d_y = 0; out.radiance = 0; out.c4 = 0; out.sflags = 0;
    if (a.simulate_gradients) {                                          // prt.cu:2100-2131
        f3 rp = mat_mul(a.cam.rot, ray.pos);
        ray.dir = normalize(mat_mul(a.cam.rot, ray.dir));
        rp.z = (float)((double)rp.z + a.z_shift);
        ray.pos = rp;
        if (any_nan(ray.dir) || any_nan(ray.pos)) return out;
    }
    if (DUMP && save) a.final_dir[gid] = ray.dir;                                // prt.cu:2136-2141
    if (STATS) { out.sdir = ray.dir; out.sflags |= 1; }
    bool flip = true;
    f3 hit;
    if (a.train.e[0].type == 'n') {                                      // prt.cu:1545-1733
        f3 dir = -ray.dir;
        f3 nrm = mk3(0.0f, 0.0f, -1.0f);
        float t = -(dot(nrm, ray.pos) + a.apparent_z_object) / dot(nrm, dir);
        hit = ray.pos + dir * t;
        float f = a.train.e[0].focal;
        float M = f / (a.apparent_z_object - a.apparent_z_offset - f);
        hit.x = -hit.x * M; hit.y = -hit.y * M;
        ray.dir = dir;
    } else {
        optical_train(a.train, ray, entry_hit);
        if (any_nan(ray.dir) || any_nan(ray.pos)) return out;            // prt.cu:2172-2176
        hit = sensor_plane_hit(ray, a.cam);
        flip = a.implement_diffraction != 0;
    }
    if (a.add_pos_noise) {
        float n1, n2; ns.next2(n1, n2);
        hit.x += n1 * a.pos_noise_std * a.cam.pitch; hit.y += n2 * a.pos_noise_std * a.cam.pitch;
    }
    bool on = pixel_coords(hit, a.cam, flip, out.d_x, out.d_y);
    if (!on || is_nan(hit.x) || is_nan(hit.y)) return out;               // dump stays NaN (prt.cu:1449-1455, 2196)
    if (DUMP && save) a.final_pos[gid] = hit;
    if (STATS) { out.spos = hit; out.sflags |= 2; }
    out.radiance = ray.radiance;
    out.c4 = cos4_alpha(ray.dir);
    out.valid = true;
    return out;
}

template <bool DUMP, class Sampler, bool STATS = false, int ALGS = 0>
PH_HD SensorSample trace_one(const TraceArgs &a, const Sampler &S, int p, int lr,
                                                  long long gid, bool save, unsigned long long &steps) {
    NormalStream ns;
    Ray ray = ray_prologue(a, p, lr, ns);
    if (a.simulate_gradients) {
        IntermediateDump im; im.n = 0; im.pos = nullptr; im.dir = nullptr;
        if (DUMP && a.inter_pos && gid < a.n_save) {
            im.pos = a.inter_pos + (size_t)gid * a.n_inter; im.dir = a.inter_dir + (size_t)gid * a.n_inter; im.n = a.n_inter;
        }
        steps += (unsigned long long)trace_volume<DUMP, Sampler, ALGS>(ray, a.field, S, a.algorithm, im, a.add_ngrad_noise != 0, a.ngrad_noise_std, ns);
    }
    return ray_epilogue<DUMP, STATS>(a, ray, ns, gid, save);
}

// One voxel of the (grad n, n-1) field: Gladstone-Dale n-1 = K rho (trace.h:1729-1748) and
// finite-difference gradients, one-sided 2nd order at the faces and central inside
// (setData, trace.h:1856-1995; literals such as 3.0/2 are double there, y uses 2.0*grid).
// FLOAT_INDEX: the reference addresses the neighbour samples as size_t(lookup.z*W*H + lookup.y*W + lookup.x) with a
// float3 lookup (trace.h:1873-1975) -- float products and sums, so above 2^24 voxels the low bits of the index are
// rounded away (at 512^3, x +- 1 lands on a multiple of 8 and the x gradient of most voxels is 0 or 4x too large).
// Reproduced by default for fields of more than 2^24 voxels (parity with the reference's images); an index rounded
// past the end (the reference reads out of bounds) is clamped.  PHOTON_GRADIENT_INDEX=exact selects the integer index.
// The index arithmetic lives in an Indexer so that the ingest kernel can hoist what is constant along a row:
//   VoxelIndex<false>  integer index;
//   VoxelIndex<true>   the reference's float index, every term evaluated per call (host check, oracle-style);
//   RowIndex           the same float index with fl(fl(z W) H) + fl(y W) precomputed for the five (y, z) pairs an
//                      interior voxel touches (trace_kernel.cuh) -- the same operations on the same operands, so the
//                      same bits.
template <bool FLOAT_INDEX>
struct VoxelIndex {
    int W, H; size_t n;
    PH_HD size_t at(int xx, int yy, int zz) const {
        if (!FLOAT_INDEX) return (size_t)zz * ((size_t)W * (size_t)H) + (size_t)yy * (size_t)W + (size_t)xx;
        const float f = add_rn(add_rn(mul_rn(mul_rn((float)zz, (float)W), (float)H), mul_rn((float)yy, (float)W)), (float)xx);
        const size_t i = (size_t)f;
        return i < n ? i : n - 1;
    }
};
template <class Indexer>
PH_HD f4 field_voxel_at(const Indexer &I, const float *rho, int x, int y, int z, int W, int H, int D, float gx, float gy, float gz) {
    const float K = 0.225e-3f;
    const size_t sW = (size_t)W, sWH = (size_t)W * (size_t)H;
#define PH_AT(xx, yy, zz) (K * (PH_LDG(&rho[I.at(xx, yy, zz)]) * 1.0f))
    float c = PH_AT(x, y, z);
    f4 o;
    if (x < 1) o.x = (float)((-3.0 / 2 * c + 2 * PH_AT(x + 1, y, z) - 1.0 / 2 * PH_AT(x + 2, y, z)) / gx);
    else if (x >= W - 1) o.x = (float)((3.0 / 2 * c - 2 * PH_AT(x - 1, y, z) + 1.0 / 2 * PH_AT(x - 2, y, z)) / gx);
    else o.x = (PH_AT(x + 1, y, z) - PH_AT(x - 1, y, z)) / (2 * gx);
    if (y < 1) o.y = (float)((-3.0 / 2 * c + 2 * PH_AT(x, y + 1, z) - 1.0 / 2 * PH_AT(x, y + 2, z)) / gy);
    else if (y >= H - 1) o.y = (float)((3.0 / 2 * c - 2 * PH_AT(x, y - 1, z) + 1.0 / 2 * PH_AT(x, y - 2, z)) / gy);
    else o.y = (float)((PH_AT(x, y + 1, z) - PH_AT(x, y - 1, z)) / (2.0 * gy));
    if (z < 1) o.z = (float)((-3.0 / 2 * c + 2 * PH_AT(x, y, z + 1) - 1.0 / 2 * PH_AT(x, y, z + 2)) / gz);
    else if (z >= D - 1) o.z = (float)((3.0 / 2 * c - 2 * PH_AT(x, y, z - 1) + 1.0 / 2 * PH_AT(x, y, z - 2)) / gz);
    else o.z = (PH_AT(x, y, z + 1) - PH_AT(x, y, z - 1)) / (2 * gz);
#undef PH_AT
    // the n-1 channel is addressed with integers in the reference too (trace.h:1990)
    o.w = K * (PH_LDG(&rho[(size_t)z * sWH + (size_t)y * sW + (size_t)x]) * 1.0f);
    return o;
}
template <bool FLOAT_INDEX>
PH_HD f4 field_voxel(const float *rho, int x, int y, int z, int W, int H, int D, float gx, float gy, float gz) {
    VoxelIndex<FLOAT_INDEX> I; I.W = W; I.H = H; I.n = (size_t)W * (size_t)H * (size_t)D;
    return field_voxel_at(I, rho, x, y, z, W, H, D, gx, gy, gz);
}

}  // namespace ph
