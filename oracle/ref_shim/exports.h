/*
 * exports.h -- TEST INFRASTRUCTURE.  Definitions of the shim's globals and the counters that
 * bench.py / the tests read from oracle/_ref/libphoton_ref_cpu.so.  Appended after the
 * reference translation unit by oracle/build_ref.sh.
 */
thread_local uint3 threadIdx, blockIdx;
thread_local dim3 blockDim, gridDim;
namespace shim {
int g_tex_fixed8 = 0;
thread_local long long t_tex_fetches = 0;
thread_local std::string t_biff;
double g_kernel_seconds = 0.0;
long long g_tex_fetches = 0, g_threads_launched = 0;
int g_host_threads = 1;
}
extern "C" {
void photon_ref_reset_counters() {
    shim::g_kernel_seconds = 0.0; shim::g_tex_fetches = 0; shim::g_threads_launched = 0;
    const char *f = getenv("PHOTON_REF_TEXFILTER");
    shim::g_tex_fixed8 = f && strcmp(f, "fixed8") == 0;
}
double photon_ref_last_trace_seconds() { return shim::g_kernel_seconds; }
long long photon_ref_last_tex_fetches() { return shim::g_tex_fetches; }
long long photon_ref_last_threads_launched() { return shim::g_threads_launched; }
int photon_ref_last_threads() { return shim::g_host_threads; }
}
