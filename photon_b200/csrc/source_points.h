// source_points.h -- source-point generation, the step before the path (host, double precision).
// Replaces the Python loops of run_simulation_02.py: calculate_sunflower_coordinates (:999-1055),
// create_non_overlapping_dot_coordinates (:1251-1326, an O(n^2) rejection loop), the per-dot expansion
// of generate_bos_lightfield_data (:1485-1506) and the particle field of load_lightfield_data (:901-975).
// The reference draws its random numbers from numpy's global generator seeded from the clock; these
// functions take the draws as arrays, so a caller can hand over numpy's own stream (what the tests do to
// compare with the reference function by function) or any counter-based stream.
#pragma once
#include <cmath>
#include <cstdint>
#include <unordered_map>
#include <vector>

namespace ph {

// rs02.py:999-1055.  ring_phase_u: one uniform draw per ring (np.random.rand(1,1) inside the loop), or null = 0.
// Writes up to capacity points; returns the number of points the pattern has (rings + the centre point).
inline long long sunflower_points(double diameter, double count, const double *ring_phase_u, double *x, double *y,
                                  long long capacity) {
    const double pi = 3.141592653589793;
    const double area = pi * std::pow(diameter / 2.0, 2.0);
    const double spacing = std::sqrt(area / count);
    const long long nring = (long long)std::nearbyint((diameter / 2.0) / spacing);       // np.round: half to even
    const double rho = 1.0 / spacing;
    long long n = 0;
    for (long long r = 0; r < nring; r++) {
        // np.linspace(spacing, diameter/2, nring)[r]
        const double radius = nring == 1 ? spacing : spacing + (double)r * ((diameter / 2.0 - spacing) / (double)(nring - 1));
        const double m = std::nearbyint(rho * (2.0 * pi * radius));
        const double phase = ring_phase_u ? 2.0 * pi * ring_phase_u[r] : 0.0;
        // theta = (2 pi / m) * (arange(1, m) - 1) + phase: m - 1 points (the reference drops the last one)
        for (double k = 1.0; k < m; k += 1.0) {
            const double th = (2.0 * pi / m) * (k - 1.0) + phase;
            if (n < capacity) { x[n] = radius * std::cos(th); y[n] = radius * std::sin(th); }
            n++;
        }
    }
    if (n < capacity) { x[n] = 0.0; y[n] = 0.0; }
    return n + 1;
}

// rs02.py:1251-1326 with a uniform hash grid instead of the distance to every stored dot.  uniforms: 2 per
// iteration (np.random.rand(2)).  Quirk kept: the reference measures the distance to ALL num_dots slots of
// its zero-initialised arrays, so from the second iteration on a candidate closer than the threshold to
// (0, 0) is rejected as long as a slot is still empty -- i.e. always.  Returns the number of dots placed.
inline long long non_overlapping_dots(double xmin, double xmax, double ymin, double ymax, long long num_dots,
                                      double dot_diameter, const double *uniforms, long long n_pairs,
                                      long long max_iter, double *xy, long long *iterations_used) {
    const double thr = dot_diameter * 1.5;
    const double cell = thr > 0.0 ? thr : 1.0;
    std::unordered_map<uint64_t, std::vector<long long>> grid;
    auto key = [&](double px, double py) {
        const int64_t cx = (int64_t)std::floor(px / cell), cy = (int64_t)std::floor(py / cell);
        return ((uint64_t)(uint32_t)cx << 32) | (uint64_t)(uint32_t)cy;
    };
    long long placed = 0, it = 0;
    while (placed < num_dots && it < max_iter && it < n_pairs) {
        const double u0 = uniforms[2 * it], u1 = uniforms[2 * it + 1];
        it++;
        const double cx = xmin + dot_diameter / 2 + (xmax - xmin - dot_diameter) * u0;
        const double cy = ymin + dot_diameter / 2 + (ymax - ymin - dot_diameter) * u1;
        bool ok = true;
        if (it > 1) {
            if (std::sqrt(cx * cx + cy * cy) <= thr) ok = false;                       // the empty slots at (0, 0)
            const int64_t gx = (int64_t)std::floor(cx / cell), gy = (int64_t)std::floor(cy / cell);
            for (int64_t dy = -1; dy <= 1 && ok; dy++)
                for (int64_t dx = -1; dx <= 1 && ok; dx++) {
                    auto f = grid.find(((uint64_t)(uint32_t)(gx + dx) << 32) | (uint64_t)(uint32_t)(gy + dy));
                    if (f == grid.end()) continue;
                    for (long long j : f->second) {
                        const double ddx = xy[2 * j] - cx, ddy = xy[2 * j + 1] - cy;
                        if (std::sqrt(ddx * ddx + ddy * ddy) <= thr) { ok = false; break; }
                    }
                }
        }
        if (ok) {
            xy[2 * placed] = cx; xy[2 * placed + 1] = cy;
            grid[key(cx, cy)].push_back(placed);
            placed++;
        }
    }
    if (iterations_used) *iterations_used = it;
    return placed;
}

// rs02.py:1485-1530: every dot centre gets the same pattern of offsets; dot-major order.
inline void expand_dots(const double *dot_x, const double *dot_y, long long ndots, const double *off_x, const double *off_y,
                        long long noff, double z, double radiance, float *x, float *y, float *zz, double *rad, int *diam) {
    for (long long d = 0; d < ndots; d++)
        for (long long k = 0; k < noff; k++) {
            const long long i = d * noff + k;
            x[i] = (float)(off_x[k] + dot_x[d]); y[i] = (float)(off_y[k] + dot_y[d]); zz[i] = (float)z;
            rad[i] = radiance; diam[i] = 1;
        }
}

// rs02.py:941-975: particles uniform in a box (3 draws each, X then Y then Z as three np.random.rand(n) calls),
// Gaussian light-sheet radiance from the unrotated Z, rotation about the origin by the camera angles (row-major
// 3x3 R, rs02.py:395-443), shift to the object plane.
inline void piv_particles(long long n, const double *ux, const double *uy, const double *uz, const double lo[3],
                          const double hi[3], double fwhm, double irradiance_constant, const double R[9], double z_object,
                          float *x, float *y, float *z, double *rad, int *diam) {
    const double sigma = fwhm / (2.0 * std::sqrt(2.0 * std::log(2.0)));
    const double pi = 3.141592653589793;
    for (long long i = 0; i < n; i++) {
        const double X = (hi[0] - lo[0]) * ux[i] + lo[0], Y = (hi[1] - lo[1]) * uy[i] + lo[1], Z = (hi[2] - lo[2]) * uz[i] + lo[2];
        rad[i] = irradiance_constant * (1.0 / (sigma * std::sqrt(2.0 * pi))) * std::exp(-1.0 * (std::pow(Z, 2.0) / (2.0 * std::pow(sigma, 2.0))));
        x[i] = (float)(R[0] * X + R[1] * Y + R[2] * Z);
        y[i] = (float)(R[3] * X + R[4] * Y + R[5] * Z);
        z[i] = (float)((R[6] * X + R[7] * Y + R[8] * Z) + z_object);
        diam[i] = 1;
    }
}

}  // namespace ph
