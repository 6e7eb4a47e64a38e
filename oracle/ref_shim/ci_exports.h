/*
 * ci_exports.h -- TEST INFRASTRUCTURE.  C entry points over CubicInterpolationCUDA's plain-C++ CPU reference
 * (examples/referenceCubicTexture3D/cubicPrefilter3D.hpp, cubicFilter3D_kernel.hpp and code/internal/
 * bspline_kernel.cu, cubicPrefilter_kernel.cu), compiled from where they lie by oracle/build_ref_ci.sh.  They are
 * the known answers for the product's cubic mode: the prefilter, the 64-tap value, the B-spline weights and their
 * first derivative.
 */
extern "C" {
/* CubicBSplinePrefilter3DTimer (cubicPrefilter3D.hpp:96-150): in place, x then y then z */
void photon_ci_prefilter3d(float *volume, unsigned w, unsigned h, unsigned d) {
    ci_prefilter::CubicBSplinePrefilter3DTimer(volume, w, h, d);
}
/* interpolate_tricubic_simple (cubicFilter3D_kernel.hpp:66-93) at n coordinates (texel centres at +0.5) */
void photon_ci_interpolate(float *coeffs, unsigned w, unsigned h, unsigned d, int n, const float *xyz, float *out) {
    uint3 ext = make_uint3(w, h, d);
    for (int i = 0; i < n; i++)
        out[i] = ci_filter::interpolate_tricubic_simple(coeffs, make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), ext);
}
/* bspline_weights / bspline_weights_1st_derivative (bspline_kernel.cu:83-104) */
void photon_ci_weights(float t, float *w, float *dw) {
    ci_bspline::bspline_weights<float>(t, w[0], w[1], w[2], w[3]);
    ci_bspline::bspline_weights_1st_derivative<float>(t, dw[0], dw[1], dw[2], dw[3]);
}
/* bspline / bspline_1st_derivative (bspline_kernel.cu:50-70): the piecewise definitions */
float photon_ci_bspline(float t) { return ci_bspline::bspline(t); }
float photon_ci_bspline_1st_derivative(float t) { return ci_bspline::bspline_1st_derivative(t); }
}
