// ray_optics.cuh -- stage (c) optical train and stage (d) sensor projection + splat weights.
// Included from ray_stages.cuh.
#pragma once
#include "vecmath.cuh"

namespace ph {

// prt.cu:239-343.  Quadratic in float, discriminant and roots through double
// sub-expressions (the literals 4.0 and 2.0 promote).  Root choice: smaller root for
// R > 0, larger for R <= 0, for both surfaces (prt.cu:298-336).
PH_HD f3 ray_sphere(f3 c, float R, f3 d, f3 p) {
    float alpha = dot_rn(d, d);
    f3 pc = p - c;
    float beta = 2 * dot_rn(d, pc);
    float gamma = add_rn(dot_rn(pc, pc), -mul_rn(R, R));
    float disc = (float)((double)(beta * beta) - 4.0 * (double)alpha * (double)gamma);
    if (disc < 0.0f) return nan3();
    float sq = sqrtf(disc);
#if defined(__CUDA_ARCH__)
    // (float)((double)a / (2.0 * (double)alpha)) for floats a, alpha: 2 alpha is exact, and binary64 carries more than
    // 2 x 24 + 2 bits, so rounding its quotient to float gives the correctly rounded float quotient -- the same bits
    // from one float division instead of two double ones (74 of the 1840 warp instructions of a PIV ray, ncu r4o)
    const float den = __fmul_rn(2.0f, alpha);
    float t1 = __fdiv_rn(__fadd_rn(-beta, sq), den);
    float t2 = __fdiv_rn(__fadd_rn(-beta, -sq), den);
#else
    double den = 2.0 * (double)alpha;
    float t1 = (float)((double)(-beta + sq) / den);
    float t2 = (float)((double)(-beta - sq) / den);
#endif
    float t = (R > 0) ? (t1 <= t2 ? t1 : t2) : (t1 >= t2 ? t1 : t2);
    return p + d * t;
}

// prt.cu:345-380
PH_HD float axis_distance(f3 p, f3 c, f4 pl) {
    f3 n = mk3(pl.x, pl.y, pl.z);
    float tmin = dot(n, p - c) / (pl.x * pl.x + pl.y * pl.y + pl.z * pl.z);
    f3 q = c + n * tmin;
    f3 e = p - q;
    return sqrtf(dot(e, e));
}

// Cauchy dispersion with the Abbe number (prt.cu:622-636 front, 774-784 back).
PH_HD double dispersed_index(double n_glass, float abbe, float wavelength, bool front) {
    const float lD = 589.3f, lF = 486.1f, lC = 656.3f;
    double inv_l2 = 1.0 / (double)(wavelength * wavelength);
    double inv_d2 = front ? (double)(1 / (lD * lD)) : 1.0 / (double)(lD * lD);
    return n_glass + (inv_l2 - inv_d2) * ((n_glass - 1) / (double)(abbe * (1 / (lF * lF) - 1 / (lC * lC))));
}

// Snell at a surface with unit normal N facing the incoming ray; eta = n_in/n_out
// (prt.cu:655-690).  Returns false on total internal reflection.
PH_HD bool refract(f3 &dir, f3 N, float eta) {
    float cosi = -dot(dir, N);
    float rad = (float)(1.0 - (double)(eta * eta) * (1.0 - (double)(cosi * cosi)));
    if (rad < 0.0f) return false;
    dir = dir * eta + (eta * cosi - sqrtf(rad)) * N;
    dir = normalize(dir);
    return true;
}

// prt.cu:383-1011
// entry_hit: where the ray meets the element's first surface, when the caller already has it (first_aperture_culls of the
// no-field kernel computed it, and that the ray passes the aperture there); null = compute and test it here.
PH_HD void refract_element(const Element &el, Ray &ray, const f3 *entry_hit = nullptr) {
    f3 dir = ray.dir, pos = ray.pos;
    f4 pl = el.plane;
    f3 n = mk3(pl.x, pl.y, pl.z);
    const double half_pitch = (double)el.pitch / 2.0;
    if (el.type == 't') {
        f3 hit;
        if (entry_hit) hit = *entry_hit;
        else {
            float t = -(dot(n, pos) + pl.w) / dot(n, dir);
            hit = pos + dir * t;
            if ((double)axis_distance(hit, el.centre, pl) > half_pitch) { kill(ray); return; }
        }
        pos = hit;
        dir = normalize(-(pos - el.centre) / el.focal + dir);            // prt.cu:494-496
    } else if (el.type == 'l') {
        float nmag = sqrtf(pl.x * pl.x + pl.y * pl.y + pl.z * pl.z);
        float ds = (float)(el.vertex / 2.0 - (double)el.r_front);        // prt.cu:551
        f3 cf = el.centre + n * ds / nmag;
        f3 hit;
        if (entry_hit) hit = *entry_hit;
        else {
            hit = ray_sphere(cf, el.r_front, dir, pos);
            if ((double)axis_distance(hit, el.centre, pl) > half_pitch) { kill(ray); return; }
        }
        f3 N = normalize(hit - cf);
        float eta = is_nan(el.abbe) ? (float)(1.0 / el.n_glass)
                                    : (float)(1.0 / dispersed_index(el.n_glass, el.abbe, ray.wavelength, true));
        if (!refract(dir, N, eta)) { kill(ray); return; }
        pos = hit;
        ds = (float)(-el.vertex / 2 - (double)el.r_back);                // prt.cu:714
        f3 cb = el.centre + n * ds / nmag;
        hit = ray_sphere(cb, el.r_back, dir, pos);
        if ((double)axis_distance(hit, el.centre, pl) > half_pitch) { kill(ray); return; }
        N = normalize(-(hit - cb));
        eta = is_nan(el.abbe) ? (float)el.n_glass : (float)dispersed_index(el.n_glass, el.abbe, ray.wavelength, false);
        if (!refract(dir, N, eta)) { kill(ray); return; }
        if (el.absorbance != 0) {                                        // prt.cu:838-853
            f3 e = hit - pos;
            float dist = sqrtf(dot(e, e));
            ray.radiance = (1.0 - (double)el.absorbance) * ray.radiance * (double)dist;
        } else ray.radiance = (double)el.transmission * ray.radiance;
        pos = hit;
    } else {                                                             // aperture, prt.cu:868-992
        float nmag = sqrtf(pl.x * pl.x + pl.y * pl.y + pl.z * pl.z);
        float ds = (float)(-el.vertex / 2.0);
        float dtmp = pl.w - ds * nmag;
        f3 hit;
        if (!entry_hit) {
            float t = -(dot(n, pos) + dtmp) / dot(n, dir);
            hit = pos + dir * t;
            if ((double)axis_distance(hit, el.centre, pl) > half_pitch) { kill(ray); return; }
        }
        ds = (float)(el.vertex / 2);
        dtmp = pl.w - ds * nmag;
        float t = -(dot(n, pos) + dtmp) / dot(n, dir);
        hit = pos + dir * t;
        if ((double)axis_distance(hit, el.centre, pl) > half_pitch) { kill(ray); return; }
        pos = hit;
    }
    ray.dir = dir; ray.pos = pos;
}

// The FIRST aperture test a ray meets in the optical train, on its own: true when refract_element would cull the ray
// at the entry of the first stage's element (prt.cu:478-489 thin lens, 560-575 front surface of a thick lens, 880-895
// first plane of an aperture) -- the same expressions on the same operands, so the same decision.  About half of all
// rays end here (the lens samples cover a disc of radius lens_pitch, twice the aperture's: prt.cu:123-124 with
// ray_cone_pitch_ratio = 1), which is what the no-field kernel's compaction uses (trace_kernel.cuh).  A NaN hit is
// not a cull here either: it propagates and the ray is dropped by the NaN test after the train.
PH_HD bool first_aperture_culls(const OpticalTrain &ot, const Ray &ray, f3 *hit_out = nullptr) {
    int e = -1;
    for (int s = 0; s < ot.stages && e < 0; s++) e = ot.order[s];
    if (e < 0) return false;
    const Element &el = ot.e[e];
    const f3 dir = ray.dir, pos = ray.pos;
    const f4 pl = el.plane;
    const f3 n = mk3(pl.x, pl.y, pl.z);
    const double half_pitch = (double)el.pitch / 2.0;
    f3 hit;
    if (el.type == 't') {
        float t = -(dot(n, pos) + pl.w) / dot(n, dir);
        hit = pos + dir * t;
    } else if (el.type == 'l') {
        float nmag = sqrtf(pl.x * pl.x + pl.y * pl.y + pl.z * pl.z);
        float ds = (float)(el.vertex / 2.0 - (double)el.r_front);
        f3 cf = el.centre + n * ds / nmag;
        hit = ray_sphere(cf, el.r_front, dir, pos);
    } else {
        float nmag = sqrtf(pl.x * pl.x + pl.y * pl.y + pl.z * pl.z);
        float ds = (float)(-el.vertex / 2.0);
        float dtmp = pl.w - ds * nmag;
        float t = -(dot(n, pos) + dtmp) / dot(n, dir);
        hit = pos + dir * t;
    }
    if (hit_out) *hit_out = hit;
    return (double)axis_distance(hit, el.centre, pl) > half_pitch;
}

// prt.cu:1274-1381 with the intended element-per-stage order (deviation B-3).
PH_HD void optical_train(const OpticalTrain &ot, Ray &ray, const f3 *entry_hit = nullptr) {
    for (int s = 0; s < ot.stages; s++) {
        int e = ot.order[s];
        if (e < 0) continue;
        refract_element(ot.e[e], ray, entry_hit);
        entry_hit = nullptr;                                             // only the first element the ray meets
        if (any_nan(ray.pos) || any_nan(ray.dir)) return;
    }
}

// ----------------------------------------------------------------------------------
// (d) sensor
// ----------------------------------------------------------------------------------
PH_HD f3 sensor_plane_hit(const Ray &ray, const Camera &cam) {          // prt.cu:1403-1415
    f3 nrm = mk3(0.0f, 0.0f, 1.0f);
    float t = -(dot(nrm, ray.pos) + (-cam.z_sensor)) / dot(nrm, ray.dir);
    return ray.pos + ray.dir * t;
}
PH_HD double cos4_alpha(f3 d) {                                         // prt.cu:1461-1463
    float qx = d.x / d.z, qy = d.y / d.z;
    float alpha = atanf(sqrtf(qx * qx + qy * qy));
    float c = cosf(alpha);
    return (double)(c * c * c * c);
}

// Separable erf-Gaussian footprint of one ray (prt.cu:1466-1540).  The reference evaluates 4 erf per rendered
// pixel; the column factor depends only on the column and the row factor only on the row, so evaluating them once
// per column / row gives bit-identical increments with ~4x fewer erf calls.  The factors are never stored in an
// indexable array: the row factors of a footprint of up to N rows live in registers (loops over them are fully
// unrolled), the column factor is computed when its column is walked, and a footprint with more rows than the
// largest N falls back to one pair of erf per pixel -- any diameter renders, nothing sits in local memory.
struct ErfFootprint {
    int c0, r0, nc, nr;        // first column/row and counts (unclipped)
    float X, Y, radius, diam;
    double scale;              // peak*pi/32
    bool valid;
};

PH_HD void erf_footprint(ErfFootprint &fp, float d_x, float d_y, double radiance, double c4, float diam, float frac) {
    const double pi = 3.141592653589793;
    fp.X = (float)((double)d_x - 0.5); fp.Y = (float)((double)d_y - 0.5);
    float peak = (float)(radiance * c4 * 8.0 / pi);
    fp.radius = frac * diam; fp.diam = diam;
    fp.c0 = (int)floorf(fp.X - fp.radius); int c1 = (int)ceilf(fp.X + fp.radius);
    fp.r0 = (int)floorf(fp.Y - fp.radius); int r1 = (int)ceilf(fp.Y + fp.radius);
    fp.nc = c1 - fp.c0 + 1; fp.nr = r1 - fp.r0 + 1;
    fp.valid = fp.nc >= 1 && fp.nr >= 1;
    fp.scale = (double)peak * pi / 32.0;
}
// erf(sqrt8 (u - 1/2) / D) - erf(sqrt8 (u + 1/2) / D) at pixel index idx of the axis whose centre is c
// (prt.cu:1516-1523; the literal 0.5 promotes to double there)
PH_HD double erf_factor(const ErfFootprint &fp, int idx, float c) {
    const float sqrt8 = sqrtf(8.0f);
    const double u = (double)((float)idx - c);
    return erf((double)sqrt8 * (u - 0.5) / (double)fp.diam) - erf((double)sqrt8 * (u + 0.5) / (double)fp.diam);
}
PH_HD bool erf_pixel_inside(const ErfFootprint &fp, int col, int row) {
    const float dx = (float)col - fp.X, dy = (float)row - fp.Y;
    return sqrtf(dx * dx + dy * dy) <= fp.radius;
}

// One ray's footprint into an accumulator; row factors of up to N rows in registers.
template <int N, class Acc>
PH_HD void erf_splat_rows(const ErfFootprint &fp, const Camera &cam, Acc &acc) {
    double ey[N];
#pragma unroll
    for (int q = 0; q < N; q++) ey[q] = 0.0;
#pragma unroll 1                                             // ONE inlined copy of the erf pair; the value is steered
    for (int j = 0; j < fp.nr; j++) {                        // into its register by an unrolled select
        const int row = fp.r0 + j;
        if (row < 0 || row > cam.ny - 1) continue;
        const double v = erf_factor(fp, row, fp.Y);
#pragma unroll
        for (int q = 0; q < N; q++) if (q == j) ey[q] = v;
    }
#pragma unroll 1
    for (int i = 0; i < fp.nc; i++) {
        const int col = fp.c0 + i;
        if (col < 0 || col > cam.nx - 1) continue;
        const double ex = erf_factor(fp, col, fp.X);
#pragma unroll
        for (int j = 0; j < N; j++) {
            const int row = fp.r0 + j;
            if (j >= fp.nr || row < 0 || row > cam.ny - 1) continue;
            if (!erf_pixel_inside(fp, col, row)) continue;
            acc.add(row, col, (float)(fp.scale * ex * ey[j]));
        }
    }
}
// Any footprint: a pair of erf per column and per pixel (large diffraction diameters; rare).
// Arguments by value: a reference into the caller's frame would pin its footprint record in local memory.
template <class Acc>
PH_COLD void erf_splat_any(const ErfFootprint fp, const Camera &cam, Acc acc) {
    for (int i = 0; i < fp.nc; i++) {
        const int col = fp.c0 + i;
        if (col < 0 || col > cam.nx - 1) continue;
        const double ex = erf_factor(fp, col, fp.X);
        for (int j = 0; j < fp.nr; j++) {
            const int row = fp.r0 + j;
            if (row < 0 || row > cam.ny - 1) continue;
            if (!erf_pixel_inside(fp, col, row)) continue;
            acc.add(row, col, (float)(fp.scale * ex * erf_factor(fp, row, fp.Y)));
        }
    }
}
template <class Acc>
PH_HD void erf_splat(const ErfFootprint &fp, const Camera &cam, Acc &acc) {
    if (!fp.valid) return;
    if (fp.nr <= 8) erf_splat_rows<8>(fp, cam, acc);
    else erf_splat_any(fp, cam, acc);
}

// pixel coordinates of a sensor hit; flip_x for the diffraction / apparent models
// (prt.cu:1446-1447) and not for the bilinear model (prt.cu:1807-1808).
PH_HD bool pixel_coords(f3 hit, const Camera &cam, bool flip_x, float &d_x, float &d_y) {
    float u = (hit.x - cam.p1x) / cam.pitch;
    d_x = flip_x ? (float)(cam.nx - 1) - u : u;
    d_y = (hit.y - cam.p1y) / cam.pitch;
    return !(d_x >= cam.nx || d_y >= cam.ny || d_x < 0 || d_y < 0);
}

// Bilinear 4-pixel splat (prt.cu:1810-1850 weights, 2219-2234 accumulation).  The
// reference subtracts 1 from both indices after its bounds test; a resulting negative
// linear index (its out-of-bounds write) is dropped.  Acc::add_slot(k, idx, v): k = corner 0..3.
template <class Acc>
PH_HD void bilinear_splat(float d_x, float d_y, double radiance, double c4, const Camera &cam, Acc &acc) {
    float yl = (float)((double)d_y - 0.5), xl = (float)((double)d_x - 0.5);
    float cy = ceilf(yl), cx = ceilf(xl);
    double dii = (double)(cy - yl), djj = (double)(cx - xl);
    int ii0 = (int)(cy - 1), jj0 = (int)(cx - 1);
    double rc = radiance;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        int ii = ii0 + (k >> 1), jj = jj0 + (k & 1);
        double w = ((k >> 1) ? (1 - dii) : dii) * ((k & 1) ? (1 - djj) : djj);
        if (ii < 0 || ii >= cam.ny || jj < 0 || jj >= cam.nx) continue;
        long idx = (long)(ii - 1) * cam.nx + jj - 1;
        if (idx < 0) continue;
        float inc = (float)(w * rc * c4);
        acc.add_slot(k, idx, inc);
    }
}

}  // namespace ph
