// driver.cu -- host side of libparallel_ray_tracing.so (B200 build).
//
// Replaces cuda_codes/parallel_ray_tracing.cu:3078-3775 (start_ray_tracing) and the field
// setup of cuda_codes/trace_rays_through_density_gradients.h:1612-2105.  Differences in
// structure (not in results): one stream, asynchronous uploads from the caller's buffers,
// the density field is built on the GPU and cached across calls, all source points go
// through one persistent launch unless ray dumps force the reference's per-batch files,
// every CUDA call is checked, and nothing ever calls exit() inside the host process.
#include <cuda.h>              // CUtensorMap types only; cuTensorMapEncodeTiled is fetched through the runtime
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <chrono>
#include <limits>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/photon_b200.h"
#include "nrrd_reader.h"
#include "scene_args.h"
#include "trace_kernel.cuh"
#include "mie_table.h"
#include "source_points.h"
#if defined(__has_include)
#if __has_include(<nccl.h>)
#include <nccl.h>
#define PH_HAVE_NCCL 1
#endif
#endif

using namespace ph;

namespace {

thread_local std::string g_error;
thread_local std::string g_variant;       // trace kernel variant of the last launch (photon_b200_last_variant)
bool verbose() { static int v = getenv("PHOTON_VERBOSE") ? atoi(getenv("PHOTON_VERBOSE")) : 0; return v != 0; }
#define VLOG(...) do { if (verbose()) { printf(__VA_ARGS__); fflush(stdout); } } while (0)

bool cuda_ok(cudaError_t e, const char *what, const char *file, int line) {
    if (e == cudaSuccess) return true;
    char buf[512];
    snprintf(buf, sizeof buf, "%s failed: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
    g_error = buf;
    fprintf(stderr, "libparallel_ray_tracing(b200): %s\n", buf);
    return false;
}
#define CK(x) do { if (!cuda_ok((x), #x, __FILE__, __LINE__)) return false; } while (0)

// ---------------------------------------------------------------------------------
// Density field resident on one device; cached across calls (the BOS script calls the
// library twice per run and the reference reloads + recomputes the field each time).
// ---------------------------------------------------------------------------------
struct DeviceField {
    std::mutex mu;                       // held while the entry is built or one of its on-demand tables is added
    int device = -1;                     // the device the buffers live on (never changes once set)
    int refs = 0;                        // scenes / ingest calls holding the entry (guarded by g_field_mutex)
    bool failed = false;                 // the build failed: waiters give up, the entry is dropped
    std::string path; long long mtime = 0, fsize = 0;
    FieldParams P{};
    f3 spacing{};
    f4 *lin = nullptr;                   // [D][H][W] float4 (grad n, n-1)
    float *coef = nullptr;               // cubic mode: [D][H][W] B-spline coefficients of n-1 (built on demand)
    float4 *coefq = nullptr;             // ... and their quad rows (c[x..x+3] per voxel) for the cubic hot loop
    void *coefq_tmap = nullptr;          // ... and a CUtensorMap of the quad rows in device memory (PHOTON_CUBIC_BRICK=1)
    float4 *cells = nullptr;             // coherent hot loop: [D][H][W][8] per-cell polynomial table (built on demand)
    bool cells_refused = false;          // it did not fit; do not try again for this field
    cudaArray_t arr = nullptr;           // optional 3-D array for the texture path
    cudaTextureObject_t tex = 0;
    size_t bytes() const {
        const size_t n = (size_t)P.W * P.H * P.D;
        return (lin ? n * 16 : 0) + (coef ? n * 4 : 0) + (coefq ? n * 16 : 0) + (cells ? n * 128 : 0) + (arr ? n * 16 : 0);
    }
    void release() {
        if (device >= 0) cudaSetDevice(device);
        if (tex) cudaDestroyTextureObject(tex);
        if (arr) cudaFreeArray(arr);
        if (lin) cudaFree(lin);
        if (coef) cudaFree(coef);
        if (coefq) cudaFree(coefq);
        if (coefq_tmap) cudaFree(coefq_tmap);
        if (cells) cudaFree(cells);
        tex = 0; arr = nullptr; lin = nullptr; coef = nullptr; coefq = nullptr; coefq_tmap = nullptr; cells = nullptr; cells_refused = false;
    }
};
// The field cache.  An entry is keyed by (device, path, mtime, size) and REFERENCE COUNTED: every scene (and every
// ingest call) that uses a field holds a reference until it is destroyed, so loading another file, rewriting the
// NRRD or a concurrent call from another host thread can never free buffers a live scene still reads.  Entries
// nobody holds stay cached (the BOS script calls the library twice per run with the same file) until another
// field is loaded on that device or photon_b200_release_cache() is called.
std::mutex g_field_mutex;
std::vector<DeviceField *> g_fields;

// Fields registered from memory (photon_b200_register_field): density_grad_filename "mem:<name>" resolves
// here instead of the file system -- for synthetic fields that would otherwise make a round trip through a
// multi-GiB NRRD file.  The caller keeps the array alive until every device that uses it has built its field.
struct MemField { const float *rho; int nx, ny, nz; double spacing[3], origin[3]; long long version; };
std::mutex g_mem_mutex;
std::vector<std::pair<std::string, MemField>> g_mem_fields;
long long g_mem_version = 0;
bool find_mem_field(const char *path, MemField &out) {
    if (strncmp(path, "mem:", 4) != 0) return false;
    std::lock_guard<std::mutex> lk(g_mem_mutex);
    for (auto &e : g_mem_fields) if (e.first == path + 4) { out = e.second; return true; }
    return false;
}

// The NRRD file is read ONCE per (path, mtime, size) however many devices build a field from it at the same time (the
// device threads of a PHOTON_GPUS call all arrive here together; eight reads + eight 512 MB vectors cost the cold call
// of round 1 seconds).  The first thread reads, the others wait on the entry's mutex and share the host copy; the copy
// lives as long as somebody holds it (weak cache).
struct SharedVolume { std::mutex mu; bool loaded = false, ok = false; pnrrd::Volume vol; };
std::mutex g_vol_mutex;
std::vector<std::pair<std::string, std::weak_ptr<SharedVolume>>> g_volumes;
std::shared_ptr<SharedVolume> shared_volume(const char *path) {
    struct stat sb;
    std::string key = path;
    if (stat(path, &sb) == 0) key += "|" + std::to_string((long long)sb.st_mtim.tv_sec) + "." + std::to_string((long long)sb.st_mtim.tv_nsec) + "|" + std::to_string((long long)sb.st_size);
    std::shared_ptr<SharedVolume> v;
    {
        std::lock_guard<std::mutex> lk(g_vol_mutex);
        for (size_t i = 0; i < g_volumes.size();) {
            if (g_volumes[i].second.expired()) { g_volumes.erase(g_volumes.begin() + i); continue; }
            if (g_volumes[i].first == key) v = g_volumes[i].second.lock();
            i++;
        }
        if (!v) { v = std::make_shared<SharedVolume>(); g_volumes.emplace_back(key, v); }
    }
    std::lock_guard<std::mutex> lk(v->mu);
    if (!v->loaded) { v->ok = pnrrd::load(path, v->vol); v->loaded = true; }
    return v;
}

bool build_field(DeviceField &F, const char *path, int device, cudaStream_t st) {
    pnrrd::Volume vol;                   // header fields only (the voxels stay in the shared copy)
    std::shared_ptr<SharedVolume> shared;
    MemField mem{};
    const bool from_memory = strncmp(path, "mem:", 4) == 0;
    const float *src = nullptr;
    if (from_memory) {
        if (!find_mem_field(path, mem)) {
            g_error = std::string("no field registered as \"") + path + "\"";
            fprintf(stderr, "libparallel_ray_tracing(b200): %s\n", g_error.c_str());
            return false;
        }
        vol.nx = mem.nx; vol.ny = mem.ny; vol.nz = mem.nz;
        for (int i = 0; i < 3; i++) { vol.spacing[i] = mem.spacing[i]; vol.origin[i] = mem.origin[i]; }
        src = mem.rho;
    } else {
        shared = shared_volume(path);
        if (!shared->ok) {
            g_error = std::string("failed to load \"") + path + "\": " + shared->vol.error;
            fprintf(stderr, "libparallel_ray_tracing(b200): %s\n", g_error.c_str());
            return false;
        }
        vol.nx = shared->vol.nx; vol.ny = shared->vol.ny; vol.nz = shared->vol.nz;
        for (int i = 0; i < 3; i++) { vol.spacing[i] = shared->vol.spacing[i]; vol.origin[i] = shared->vol.origin[i]; }
        src = shared->vol.rho.data();
    }
    int sx = vol.nx, sy = vol.ny, sz = vol.nz;
    // bounds from the header (trace.h:1698-1706); NRRD z origin is shifted by -750e3
    double xmin = vol.origin[0], dx = vol.spacing[0], xmax = xmin + (sx - 1) * dx;
    double ymin = vol.origin[1], dy = vol.spacing[1], ymax = ymin + (sy - 1) * dy;
    double zmin = vol.origin[2] - 750e3, dz = vol.spacing[2], zmax = zmin + (sz - 1) * dz;
    // the reference keeps at most 1024 planes (data_max = 1024, trace.h:2037-2038, 1713-1716)
    // while the bounds keep the full extent; PHOTON_FULL_DEPTH=1 lifts the cut.
    if (sz > 1024 && !getenv("PHOTON_FULL_DEPTH")) sz = 1024;
    FieldParams P{};
    P.mn = mk3((float)xmin, (float)ymin, (float)zmin);
    P.mx = mk3((float)xmax, (float)ymax, (float)zmax);
    P.scale = mk3(1.0f / (P.mx.x - P.mn.x), 1.0f / (P.mx.y - P.mn.y), 1.0f / (P.mx.z - P.mn.z));
    P.W = sx; P.H = sy; P.D = sz;
    P.Wf = (float)sx; P.Hf = (float)sy; P.Df = (float)sz;
        { double k[3] = {(double)P.scale.x * (sx - 2), (double)P.scale.y * (sy - 2), (double)P.scale.z * (sz - 2)};
          P.tk = mk3((float)k[0], (float)k[1], (float)k[2]);
          P.tc = mk3((float)(0.5 - (double)P.mn.x * k[0]), (float)(0.5 - (double)P.mn.y * k[1]), (float)(0.5 - (double)P.mn.z * k[2])); } P.wm2 = (float)(sx - 2); P.hm2 = (float)(sy - 2); P.dm2 = (float)(sz - 2);
    float step = (float)fmin(dx, dy);
    step = step < dz ? step : (float)dz;
    P.step = step;
    size_t n = (size_t)sx * sy * sz;
    // Up to 2^32 voxels the hot loops index voxels with 32 bits; beyond (2048^3 with PHOTON_FULL_DEPTH: 128 GiB of
    // texels) the scene runs the 64-bit-index instantiation (LinearFieldWide, ray_stages.cuh).
    if (n > ((size_t)1 << 34)) {
        g_error = "density field has more than 2^34 voxels; not supported by this build";
        fprintf(stderr, "libparallel_ray_tracing(b200): %s\n", g_error.c_str());
        return false;
    }
    float *d_rho = nullptr; unsigned int *d_min = nullptr;
    // a registered array that already lives on this device is read in place; anything else (host memory, another
    // device: unified addressing) is copied next to the texels first
    bool in_place = false;
    if (from_memory) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeDevice && at.device == device) in_place = true;
        cudaGetLastError();
    }
    CK(cudaMalloc(&F.lin, n * sizeof(f4)));
    CK(cudaMalloc(&d_min, sizeof(unsigned int)));
    if (in_place) d_rho = const_cast<float *>(src);
    else {
        CK(cudaMalloc(&d_rho, n * sizeof(float)));
        CK(cudaMemcpyAsync(d_rho, src, n * sizeof(float), cudaMemcpyDefault, st));
    }
    CK(cudaMemsetAsync(d_min, 0xff, sizeof(unsigned int), st));
    // rows (y, z) in a grid-stride loop: exactly the CTAs that are resident at once (2368 = 2.67 waves of the 888 that fit
    // left the last third of the kernel to a partial wave, ncu r4j)
    const char *gi = getenv("PHOTON_GRADIENT_INDEX");
    const bool float_index = n > ((size_t)1 << 24) && !(gi && strcmp(gi, "exact") == 0);
    int per_sm = 0, sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if ((float_index ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, build_field_kernel<true>, 256, 0)
                     : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, build_field_kernel<false>, 256, 0)) != cudaSuccess || per_sm < 1) {
        cudaGetLastError(); per_sm = 4;
    }
    int blocks = (int)std::min<size_t>((size_t)sy * sz, (size_t)sms * per_sm);
    // (float_index: the reference's float-arithmetic voxel index (trace_core.cuh: field_voxel) differs from the integer one
    // only above 2^24 voxels)
    if (float_index) build_field_kernel<true><<<blocks, 256, 0, st>>>(d_rho, F.lin, sx, sy, sz, (float)dx, (float)dy, (float)dz, d_min);
    else build_field_kernel<false><<<blocks, 256, 0, st>>>(d_rho, F.lin, sx, sy, sz, (float)dx, (float)dy, (float)dz, d_min);
    CK(cudaGetLastError());
    unsigned int key = 0;
    CK(cudaMemcpyAsync(&key, d_min, sizeof key, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (!in_place) cudaFree(d_rho);
    cudaFree(d_min);
    P.data_min = float_from_order_key(key);
    P.skip_step = P.step / (1 + P.data_min);
    F.P = P; F.device = device; F.path = path; F.spacing = mk3((float)dx, (float)dy, (float)dz);
    VLOG("field %s: %d x %d x %d, step %.4f um, min(n-1) %.6g\n", path, sx, sy, sz, P.step, P.data_min);
    return true;
}

bool ensure_texture(DeviceField &F, cudaStream_t st) {
    if (F.tex) return true;
    cudaChannelFormatDesc desc = cudaCreateChannelDesc<float4>();
    cudaExtent ext = make_cudaExtent(F.P.W, F.P.H, F.P.D);
    CK(cudaMalloc3DArray(&F.arr, &desc, ext));
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr((void *)F.lin, (size_t)F.P.W * sizeof(float4), F.P.W, F.P.H);
    cp.dstArray = F.arr; cp.extent = ext; cp.kind = cudaMemcpyDeviceToDevice;
    CK(cudaMemcpy3DAsync(&cp, st));
    cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = F.arr;
    cudaTextureDesc td = {};                                             // trace.h:1628-1631
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
    CK(cudaCreateTextureObject(&F.tex, &rd, &td, nullptr));
    CK(cudaStreamSynchronize(st));
    return true;
}

bool ensure_coefficients(DeviceField &F, cudaStream_t st) {
    if (F.coef) return true;
    size_t n = (size_t)F.P.W * F.P.H * F.P.D;
    CK(cudaMalloc(&F.coef, n * sizeof(float)));
    int blocks = (int)std::min<size_t>((n + 255) / 256, (size_t)148 * 16);
    extract_nm1_kernel<<<blocks, 256, 0, st>>>(F.lin, F.coef, n);
    for (int axis = 0; axis < 3; axis++) {
        size_t lines = axis == 0 ? (size_t)F.P.H * F.P.D : axis == 1 ? (size_t)F.P.W * F.P.D : (size_t)F.P.W * F.P.H;
        int b = (int)std::min<size_t>((lines + 127) / 128, (size_t)148 * 16);
        prefilter_kernel<<<b, 128, 0, st>>>(F.coef, F.P.W, F.P.H, F.P.D, axis);
    }
    CK(cudaGetLastError());
    // quad rows for the hot loop (16 B per voxel); without them (out of memory) every step takes the generic path
    if (cudaMalloc(&F.coefq, n * sizeof(float4)) == cudaSuccess) {
        const size_t rows = (size_t)F.P.H * F.P.D;
        quad_rows_kernel<<<(unsigned)std::min<size_t>(rows, (size_t)148 * 16), 256, 0, st>>>(F.coef, F.coefq, F.P.W, rows);
        CK(cudaGetLastError());
    } else { cudaGetLastError(); F.coefq = nullptr; }
    CK(cudaStreamSynchronize(st));
    return true;
}

// Tensor map of the quad-row table for the TMA-staged cubic hot loop (CubicGatherBrick): the table seen as
// float[D][H][4 W], box = BXQ quad rows x BY x BZ, no swizzle, out-of-range rows filled with zeros (never read).
// The descriptor lives in device memory (128 bytes) and is handed to the kernel as a pointer.
bool ensure_brick_map(DeviceField &F, cudaStream_t st) {
    if (F.coefq_tmap) return true;
    if (!F.coefq) { g_error = "PHOTON_CUBIC_BRICK needs the quad-row table (it did not fit)"; return false; }
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess || !fn) {
        cudaGetLastError(); g_error = "cuTensorMapEncodeTiled is not available from this driver"; return false;
    }
    alignas(64) CUtensorMap tm;
    const cuuint64_t dims[3] = {(cuuint64_t)F.P.W * 4u, (cuuint64_t)F.P.H, (cuuint64_t)F.P.D};
    const cuuint64_t strides[2] = {(cuuint64_t)F.P.W * 16u, (cuuint64_t)F.P.W * (cuuint64_t)F.P.H * 16u};   // bytes, dims 1 and 2
    const cuuint32_t box[3] = {CubicBrickShape::BXQ * 4u, (cuuint32_t)CubicBrickShape::BY, (cuuint32_t)CubicBrickShape::BZ};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    const CUresult r = reinterpret_cast<encode_fn>(fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, F.coefq, dims, strides, box, estr,
                                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { g_error = "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")"; return false; }
    CK(cudaMalloc(&F.coefq_tmap, sizeof tm));
    CK(cudaMemcpyAsync(F.coefq_tmap, &tm, sizeof tm, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    return true;
}

// The per-cell polynomial table costs 128 B per voxel (8x the texels: 16 GiB at 512^3).  It is built when it fits
// in PHOTON_CELLS_MAX_GIB (default 40) and leaves a quarter of the free memory alone; otherwise the coherent loop
// reads the texels (LinearFieldCoherent).  Returns true when F.cells is there.
bool ensure_cells(DeviceField &F, cudaStream_t st) {
    if (F.cells) return true;
    if (F.cells_refused) return false;
    const size_t n = (size_t)F.P.W * F.P.H * F.P.D, bytes = n * 128;
    double cap_gib = 40.0;
    if (const char *e = getenv("PHOTON_CELLS_MAX_GIB")) cap_gib = atof(e);
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || (double)bytes > cap_gib * 1073741824.0 || bytes > free_b - free_b / 4) {
        cudaGetLastError(); F.cells_refused = true; return false;
    }
    if (cudaMalloc(&F.cells, bytes) != cudaSuccess) { cudaGetLastError(); F.cells = nullptr; F.cells_refused = true; return false; }
    int blocks = (int)std::min<size_t>((size_t)F.P.H * F.P.D, (size_t)148 * 16);
    build_cells_kernel<<<blocks, 256, 0, st>>>(F.lin, F.cells, F.P.W, F.P.H, F.P.D);
    if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
        cudaGetLastError(); cudaFree(F.cells); F.cells = nullptr; F.cells_refused = true; return false;
    }
    VLOG("field %s: per-cell polynomial table, %.1f GiB\n", F.path.c_str(), (double)bytes / 1073741824.0);
    return true;
}

void put_field(DeviceField *F) {
    if (!F) return;
    std::lock_guard<std::mutex> lk(g_field_mutex);
    if (F->refs > 0) F->refs--;
}
// Takes a reference (release it with put_field).  Builds the entry on a miss; concurrent callers of the same key
// wait for the builder on the entry's mutex.  Entries of this device that nobody holds are dropped first, so at
// most one idle field per device stays resident.
DeviceField *get_field(const char *path, int device, cudaStream_t st) {
    struct stat sb;
    long long mt = 0, fs = 0;
    if (stat(path, &sb) == 0) { mt = (long long)sb.st_mtim.tv_sec * 1000000000LL + sb.st_mtim.tv_nsec; fs = (long long)sb.st_size; }
    { MemField m{}; if (find_mem_field(path, m)) { mt = m.version; fs = (long long)m.nx * m.ny * m.nz; } }
    const bool no_cache = getenv("PHOTON_NO_FIELD_CACHE") != nullptr;
    DeviceField *F = nullptr;
    bool builder = false;
    std::vector<DeviceField *> drop;
    {
        std::lock_guard<std::mutex> lk(g_field_mutex);
        if (!no_cache)
            for (auto *f : g_fields)
                if (f->device == device && !f->failed && f->path == path && f->mtime == mt && f->fsize == fs) { F = f; break; }
        if (!F) {
            for (size_t i = 0; i < g_fields.size();) {
                DeviceField *f = g_fields[i];
                if (f->device == device && f->refs == 0) { drop.push_back(f); g_fields.erase(g_fields.begin() + i); } else i++;
            }
            F = new DeviceField();
            F->device = device; F->path = path; F->mtime = mt; F->fsize = fs;
            F->mu.lock();                                    // released below, after the build
            g_fields.push_back(F);
            builder = true;
        }
        F->refs++;
    }
    for (auto *f : drop) { { std::lock_guard<std::mutex> w(f->mu); f->release(); } delete f; }
    if (builder) {
        const bool ok = build_field(*F, path, device, st);
        if (!ok) { F->release(); F->failed = true; }
        F->mu.unlock();
        if (!ok) {
            std::lock_guard<std::mutex> lk(g_field_mutex);
            F->refs--;
            if (F->refs == 0) { for (size_t i = 0; i < g_fields.size(); i++) if (g_fields[i] == F) { g_fields.erase(g_fields.begin() + i); break; } delete F; }
            return nullptr;
        }
        return F;
    }
    bool failed;
    { std::lock_guard<std::mutex> w(F->mu); failed = F->failed; }       // wait for the builder
    if (failed) {
        std::lock_guard<std::mutex> lk(g_field_mutex);
        F->refs--;
        if (F->refs == 0) { for (size_t i = 0; i < g_fields.size(); i++) if (g_fields[i] == F) { g_fields.erase(g_fields.begin() + i); break; } delete F; }
        return nullptr;
    }
    return F;
}

// Per-call buffers come from a stream-ordered memory pool of the LIBRARY'S OWN (one per device, release threshold =
// max: after the first call a scene costs no cudaMalloc/cudaFree).  The device's default pool, which PyTorch / CuPy
// cudaMallocAsync users of the same process share, is left alone.  photon_b200_release_cache() trims it.
std::mutex g_pool_mutex;
std::vector<std::pair<int, cudaMemPool_t>> g_pools;
cudaMemPool_t device_pool(int device) {
    std::lock_guard<std::mutex> lk(g_pool_mutex);
    for (auto &e : g_pools) if (e.first == device) return e.second;
    cudaMemPool_t pool = nullptr;
    cudaMemPoolProps props; memset(&props, 0, sizeof props);
    props.allocType = cudaMemAllocationTypePinned; props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice; props.location.id = device;
    if (cudaMemPoolCreate(&pool, &props) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    } else { cudaGetLastError(); pool = nullptr; }          // fall back to plain cudaMallocAsync
    g_pools.emplace_back(device, pool);
    return pool;
}
template <class T>
bool pool_alloc(T *&dst, size_t bytes, cudaStream_t st, cudaMemPool_t pool) {
    dst = nullptr;
    if (pool) CK(cudaMallocFromPoolAsync((void **)&dst, bytes ? bytes : 1, pool, st));
    else CK(cudaMallocAsync((void **)&dst, bytes ? bytes : 1, st));
    return true;
}
template <class T>
void pool_free(T *p, cudaStream_t st) { if (p) cudaFreeAsync((void *)p, st); }

// Pinned staging for the image copies (one buffer per host thread, grown on demand).
struct PinnedStage {
    void *p = nullptr; size_t cap = 0;
    void *get(size_t bytes) {
        if (bytes > cap) {
            if (p) cudaFreeHost(p);
            p = nullptr; cap = 0;
            if (cudaMallocHost(&p, bytes) != cudaSuccess) { cudaGetLastError(); p = nullptr; return nullptr; }
            cap = bytes;
        }
        return p;
    }
    ~PinnedStage() { if (p) cudaFreeHost(p); }
};
thread_local PinnedStage g_stage;

template <class T>
bool upload(T *&dst, const T *src, size_t n, cudaStream_t st, cudaMemPool_t pool) {
    dst = nullptr;
    if (n == 0) return true;
    if (!pool_alloc(dst, n * sizeof(T), st, pool)) return false;
    CK(cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyHostToDevice, st));
    return true;
}

}  // namespace

// ---------------------------------------------------------------------------------
// A scene resident on one device
// ---------------------------------------------------------------------------------
struct photon_scene {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaMemPool_t pool = nullptr;        // the library's own stream-ordered pool of this device
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    TraceArgs args{};
    DeviceField *field = nullptr;
    bool use_texture = false;
    bool use_cubic = false;              // PHOTON_INTERPOLATION=cubic
    bool use_brick = false;              // ... with the hot loop on TMA-staged bricks (PHOTON_CUBIC_BRICK=1, coherent RK4 scenes)
    int npoints = 0, point_offset = 0;   // shard of the caller's source points
    int total_points = 0, source_point_number = 0;
    float *d_sx = nullptr, *d_sy = nullptr, *d_sz = nullptr; double *d_srad = nullptr; int *d_sdiam = nullptr;
    float *d_angle = nullptr, *d_irr = nullptr;
    f2 *d_aim = nullptr;
    int *d_ray_order = nullptr;          // Morton order of the aim points within a source point, or null
    float *d_image = nullptr;
    unsigned long long *d_fixed = nullptr;   // deterministic mode (PHOTON_DETERMINISTIC=1)
    double fixed_scale = 0.0;
    unsigned long long *d_counters = nullptr;
    unsigned long long h_counters[2] = {0, 0};
    long long rays_traced = 0;
    float last_ms = 0.0f;
    int last_launches = 0, total_launches = 0;
    int grid_blocks = 0;
    int occupancy = 3;                   // resident CTAs per SM the trace kernel is compiled for (PHOTON_OCC)
    bool occupancy_set = false;          // PHOTON_OCC was given (otherwise each gatherer runs its measured best)
    double *d_dots = nullptr;            // dot statistics [n_dots][8], see TraceArgs::dot_stats
    int *d_order = nullptr;              // processing order of the source points (Morton order of x, y), or null
    bool wide = false;                   // field of more than 2^32 voxels (or PHOTON_INDEX64=1): 64-bit voxel index in the hot loop
    bool coherent = false;               // rays of a source point stay within ~a cell of each other: v5 hot loop
    bool use_cells = false;              // ... and the field has its per-cell polynomial table (PHOTON_CELLS, default on)
    long long n_dots = 0, rays_per_dot = 0;
};

namespace {

// Morton (Z-order) permutation of points by (x, y) quantised to 16 bits per axis over their bounding box:
// 4-pass LSD radix sort of the 32-bit keys (stable, O(n)).
inline uint32_t spread16(uint32_t v) {
    v &= 0xffffu; v = (v | (v << 8)) & 0x00ff00ffu; v = (v | (v << 4)) & 0x0f0f0f0fu;
    v = (v | (v << 2)) & 0x33333333u; v = (v | (v << 1)) & 0x55555555u;
    return v;
}
std::vector<int> morton_order(const float *x, const float *y, size_t n) {
    float x0 = x[0], x1 = x[0], y0 = y[0], y1 = y[0];
    for (size_t i = 1; i < n; i++) { x0 = std::min(x0, x[i]); x1 = std::max(x1, x[i]); y0 = std::min(y0, y[i]); y1 = std::max(y1, y[i]); }
    const float kx = x1 > x0 ? 65535.0f / (x1 - x0) : 0.0f, ky = y1 > y0 ? 65535.0f / (y1 - y0) : 0.0f;
    std::vector<uint32_t> key(n), key2(n);
    std::vector<int> idx(n), idx2(n);
    for (size_t i = 0; i < n; i++) {
        const uint32_t qx = (uint32_t)std::min(65535.0f, std::max(0.0f, (x[i] - x0) * kx));
        const uint32_t qy = (uint32_t)std::min(65535.0f, std::max(0.0f, (y[i] - y0) * ky));
        key[i] = spread16(qx) | (spread16(qy) << 1); idx[i] = (int)i;
    }
    for (int pass = 0; pass < 4; pass++) {
        size_t cnt[257] = {0};
        const int sh = pass * 8;
        for (size_t i = 0; i < n; i++) cnt[((key[i] >> sh) & 0xffu) + 1]++;
        for (int b = 0; b < 256; b++) cnt[b + 1] += cnt[b];
        for (size_t i = 0; i < n; i++) { const size_t d = cnt[(key[i] >> sh) & 0xffu]++; key2[d] = key[i]; idx2[d] = idx[i]; }
        key.swap(key2); idx.swap(idx2);
    }
    return idx;
}

int pick_device(int requested) {
    if (requested >= 0) return requested;
    if (const char *e = getenv("PHOTON_DEVICE")) return atoi(e);
    int cur = 0;
    if (cudaGetDevice(&cur) != cudaSuccess) cur = 0;
    return cur;
}

bool launch_trace(photon_scene *s, int point_begin, int point_count, int dump_base,
                  f3 *fpos, f3 *fdir, long long n_save, f3 *ipos, f3 *idir, int n_inter) {
    if (point_count <= 0) return true;
    TraceArgs a = s->args;
    a.point_begin = point_begin; a.point_count = point_count; a.dump_point_base = dump_base;
    a.final_pos = fpos; a.final_dir = fdir; a.n_save = n_save;
    a.inter_pos = ipos; a.inter_dir = idir; a.n_inter = n_inter;
    const bool dump = fpos != nullptr || ipos != nullptr;
    int grid = std::min(point_count, s->grid_blocks);

    a.dot_stats = s->d_dots; a.rays_per_dot = s->rays_per_dot; a.n_dots = s->n_dots;
    a.order = dump ? nullptr : s->d_order;
    const bool stats = s->d_dots != nullptr;
    if (stats && (dump || s->d_fixed || s->use_texture)) {
        g_error = "dot statistics are not available together with ray dumps, PHOTON_DETERMINISTIC or PHOTON_SAMPLER=texture"; return false; }
    if (s->use_texture && s->d_fixed) { g_error = "PHOTON_DETERMINISTIC is not available with PHOTON_SAMPLER=texture"; return false; }
    if (s->use_cubic && s->d_fixed) { g_error = "PHOTON_DETERMINISTIC is not available with PHOTON_INTERPOLATION=cubic"; return false; }
    // Dump launches (the per-ray parity channel) run the SAME gatherer as the non-dump launch of the scene: cell table,
    // coherent corner texels or fan-out corner texels.  g_variant names what ran (photon_b200_last_variant).
    const char *mode = dump ? "dump" : stats ? "stats" : "image";
    const char *alg = a.algorithm == 1 ? "euler" : a.algorithm == 2 ? "rk4" : a.algorithm == 3 ? "rk45" : a.algorithm == 4 ? "ab4" : "?";
    const char *gat = "none";
    const bool slow_alg = a.simulate_gradients && (a.algorithm == 3 || a.algorithm == 4);
    if (slow_alg && stats) { g_error = "dot statistics are not available with the rk45 / adams-bashforth integrators"; return false; }
    if (slow_alg) {
        // rk45 / adams-bashforth (trace.h:304-718, 1293-1453): literal restatements on the generic samplers, in kernel
        // instantiations of their own so that their stack frames stay out of the RK4 / Euler kernels
        gat = s->use_cubic ? "cubic" : s->use_texture ? "texture" : "generic";
        LinearField S; S.g = s->field ? s->field->lin : nullptr; S.W = a.field.W; S.H = a.field.H; S.D = a.field.D;
        if (s->use_cubic) {
            BSplineField B; B.c = s->field->coef; B.quad = reinterpret_cast<const f4 *>(s->field->coefq); B.W = a.field.W; B.H = a.field.H; B.D = a.field.D;
            B.inv_sp = mk3(1.0f / s->field->spacing.x, 1.0f / s->field->spacing.y, 1.0f / s->field->spacing.z);
            if (dump) trace_rays_kernel<true, 2, BSplineField, false, PH_BLOCK, false, 2><<<grid, PH_BLOCK, 0, s->stream>>>(a, B);
            else trace_rays_kernel<false, 2, BSplineField, false, PH_BLOCK, false, 2><<<grid, PH_BLOCK, 0, s->stream>>>(a, B);
        } else if (s->use_texture) {
            TextureField T; T.tex = s->field->tex;
            if (dump) trace_rays_kernel<true, 2, TextureField, false, PH_BLOCK, false, 2><<<grid, PH_BLOCK, 0, s->stream>>>(a, T);
            else trace_rays_kernel<false, 2, TextureField, false, PH_BLOCK, false, 2><<<grid, PH_BLOCK, 0, s->stream>>>(a, T);
        } else if (s->d_fixed) {
            if (dump) trace_rays_kernel<true, 2, LinearField, true, PH_BLOCK, false, 2><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
            else trace_rays_kernel<false, 2, LinearField, true, PH_BLOCK, false, 2><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        } else {
            if (dump) trace_rays_kernel<true, 2, LinearField, false, PH_BLOCK, false, 2><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
            else trace_rays_kernel<false, 2, LinearField, false, PH_BLOCK, false, 2><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        }
    } else if (s->use_cubic) {
        gat = "cubic";
        BSplineField S; S.c = s->field->coef; S.quad = reinterpret_cast<const f4 *>(s->field->coefq); S.W = a.field.W; S.H = a.field.H; S.D = a.field.D;
        S.inv_sp = mk3(1.0f / s->field->spacing.x, 1.0f / s->field->spacing.y, 1.0f / s->field->spacing.z);
        BSplineFieldBrick SB; static_cast<BSplineField &>(SB) = S; SB.tmap = s->field->coefq_tmap;
        const bool brick = s->use_brick && SB.tmap && !stats && a.algorithm == 2;
        if (brick) gat = "cubic-brick";
        if (brick && dump) trace_rays_kernel<true, 2, BSplineFieldBrick><<<grid, PH_BLOCK, 0, s->stream>>>(a, SB);
        else if (brick) trace_rays_kernel<false, 2, BSplineFieldBrick><<<grid, PH_BLOCK, 0, s->stream>>>(a, SB);
        else if (stats) trace_rays_kernel<false, 2, BSplineField, false, PH_BLOCK, true><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        else if (dump) trace_rays_kernel<true, 2, BSplineField><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        else trace_rays_kernel<false, 2, BSplineField><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
    } else if (s->use_texture) {
        gat = "texture";
        TextureField S; S.tex = s->field->tex;
        if (dump) trace_rays_kernel<true, 2, TextureField><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        else trace_rays_kernel<false, 2, TextureField><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
    } else {
        LinearField S; S.g = s->field ? s->field->lin : nullptr;
        S.W = a.field.W; S.H = a.field.H; S.D = a.field.D;
        LinearFieldCells SC; static_cast<LinearField &>(SC) = S; SC.cells = s->use_cells ? s->field->cells : nullptr;
        LinearFieldCoherent SH; static_cast<LinearField &>(SH) = S;
        LinearFieldWide SW; static_cast<LinearField &>(SW) = S;
        const bool fixed = s->d_fixed != nullptr;
        const bool wide = s->wide && a.simulate_gradients;
        if (wide && (stats || fixed)) { g_error = "dot statistics / PHOTON_DETERMINISTIC are not available for fields of more than 2^32 voxels"; return false; }
        // the gatherer of the scene (what the non-dump launch runs; PHOTON_OCC variants exist for some only)
        enum { G_FANOUT, G_COHERENT, G_CELLS };
        const int occ = s->occupancy;
        int g = (stats || fixed || !a.simulate_gradients) ? G_FANOUT : s->use_cells ? G_CELLS : s->coherent ? G_COHERENT : G_FANOUT;
        if (g == G_CELLS && occ != 3 && occ != 7) g = G_COHERENT;
        if (g == G_COHERENT && occ != 3 && occ != 5) g = G_FANOUT;
        gat = !a.simulate_gradients ? "none" : g == G_CELLS ? "cells" : g == G_COHERENT ? "coherent" : "fanout";
        if (wide) gat = "wide";
        // scenes without a field, plain image launch: the two-phase kernel (rays that pass the first aperture are queued
        // in shared memory and walk the lens system in full warps); PHOTON_QUEUE=0 keeps the fused kernel
        const char *qe = getenv("PHOTON_QUEUE");
        const bool queue_off = qe && atoi(qe) == 0;
        const bool queued = !a.simulate_gradients && !dump && !stats && !fixed && !queue_off && (occ == 3 || occ == 6);
        // 6 CTAs per SM at 40 registers: the kernel is issue-bound on libm and IEEE division / square-root sequences, and
        // 48 warps per SM hide more of their latency than the spills cost (C1 3.98 ms at 3 CTAs / 80 registers, 3.81 at 4,
        // 3.67 at 5, 3.65 at 6, 3.69 at 8 / 32 registers: profiles r6e-r6g).  PHOTON_OCC=3 selects the 80-register build.
        if (queued && occ == 3 && s->occupancy_set) { gat = "queue"; trace_nofield_kernel<3><<<grid, PH_BLOCK, 0, s->stream>>>(a); }
        else if (queued) { gat = "queue"; trace_nofield_kernel<6><<<grid, PH_BLOCK, 0, s->stream>>>(a); }
        else if (wide && dump) trace_rays_kernel<true, 2, LinearFieldWide><<<grid, PH_BLOCK, 0, s->stream>>>(a, SW);
        else if (wide) trace_rays_kernel<false, 3, LinearFieldWide><<<grid, PH_BLOCK, 0, s->stream>>>(a, SW);
        else if (stats) trace_rays_kernel<false, 3, LinearField, false, PH_BLOCK, true><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        else if (fixed && dump) trace_rays_kernel<true, 2, LinearField, true><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        else if (fixed) trace_rays_kernel<false, 2, LinearField, true><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        else if (dump && g == G_CELLS) trace_rays_kernel<true, 2, LinearFieldCells><<<grid, PH_BLOCK, 0, s->stream>>>(a, SC);
        else if (dump && g == G_COHERENT) trace_rays_kernel<true, 2, LinearFieldCoherent><<<grid, PH_BLOCK, 0, s->stream>>>(a, SH);
        else if (dump) trace_rays_kernel<true, 2, LinearField><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        // the cell-table loop alone gains from more, smaller CTAs: 7 x 128 threads at 72 registers = 28 warps per SM against
        // 3 x 256 at 80 = 24 (C3 1.162e11 -> 1.198e11; 6 x 128 at 80 registers 1.171e11, 8 x 128 at 64 registers 1.112e11:
        // spills).  The corner-texel loops lose with it (C4 -23 %, C5 -8 %, Euler -11 %, profiles r5k-r5m) and keep 3 x 256.
        else if (g == G_CELLS && (occ == 7 || !s->occupancy_set)) trace_rays_kernel<false, 7, LinearFieldCells, false, 128><<<grid, 128, 0, s->stream>>>(a, SC);
        else if (g == G_CELLS) trace_rays_kernel<false, 3, LinearFieldCells><<<grid, PH_BLOCK, 0, s->stream>>>(a, SC);
        else if (g == G_COHERENT && occ == 5) trace_rays_kernel<false, 5, LinearFieldCoherent, false, 128><<<grid, 128, 0, s->stream>>>(a, SH);
        else if (g == G_COHERENT) trace_rays_kernel<false, 3, LinearFieldCoherent><<<grid, PH_BLOCK, 0, s->stream>>>(a, SH);
        else if (occ == 5) trace_rays_kernel<false, 5, LinearField, false, 128><<<grid, 128, 0, s->stream>>>(a, S);   // 96 regs, 20 warps/SM
        else if (occ == 6) trace_rays_kernel<false, 3, LinearField, false, 192><<<grid, 192, 0, s->stream>>>(a, S);   // 112 regs, 18 warps/SM
        else if (occ == 4) trace_rays_kernel<false, 4, LinearField><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        else if (occ == 3) trace_rays_kernel<false, 3, LinearField><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
        else trace_rays_kernel<false, 2, LinearField><<<grid, PH_BLOCK, 0, s->stream>>>(a, S);
    }
    g_variant = std::string(a.simulate_gradients ? alg : "nofield") + "/" + gat + "/" + mode + (s->d_fixed ? "/fixed" : "");
    CK(cudaGetLastError());
    s->last_launches++; s->total_launches++;
    s->rays_traced += (long long)point_count * a.rpp;
    return true;
}

}  // namespace

// ---------------------------------------------------------------------------------
// Extended C ABI (include/photon_b200.h)
// ---------------------------------------------------------------------------------
extern "C" {

int photon_b200_abi_version(void) { return 1; }

int photon_b200_register_field(const char *name, const float *rho, int nx, int ny, int nz,
                               const double spacing[3], const double origin[3]) {
    g_error.clear();
    if (!name || !*name || !rho || nx < 3 || ny < 3 || nz < 3 || !spacing || !origin) { g_error = "register_field: bad arguments"; return 1; }
    MemField m{rho, nx, ny, nz, {spacing[0], spacing[1], spacing[2]}, {origin[0], origin[1], origin[2]}, 0};
    std::lock_guard<std::mutex> lk(g_mem_mutex);
    m.version = ++g_mem_version;
    for (auto &e : g_mem_fields) if (e.first == name) { e.second = m; return 0; }
    g_mem_fields.emplace_back(std::string(name), m);
    return 0;
}
int photon_b200_unregister_field(const char *name) {
    std::lock_guard<std::mutex> lk(g_mem_mutex);
    for (size_t i = 0; i < g_mem_fields.size(); i++) if (name && g_mem_fields[i].first == name) { g_mem_fields.erase(g_mem_fields.begin() + i); return 0; }
    return 1;
}

// ---- source-point generation (host; replaces the Python loops of run_simulation_02.py) ----
long long photon_b200_sunflower(double diameter, double count, const double *ring_phase_u, double *x, double *y, long long capacity) {
    g_error.clear();
    if (!(diameter > 0.0) || !(count >= 1.0)) { g_error = "sunflower: diameter and count must be positive"; return -1; }
    return ph::sunflower_points(diameter, count, ring_phase_u, x, y, (x && y) ? capacity : 0);
}
long long photon_b200_non_overlapping_dots(double xmin, double xmax, double ymin, double ymax, long long num_dots,
                                           double dot_diameter, const double *uniforms, long long n_pairs, double *xy,
                                           long long *iterations_used) {
    g_error.clear();
    if (!uniforms || !xy || num_dots < 0) { g_error = "non_overlapping_dots: bad arguments"; return -1; }
    return ph::non_overlapping_dots(xmin, xmax, ymin, ymax, num_dots, dot_diameter, uniforms, n_pairs, 50000, xy, iterations_used);
}
int photon_b200_expand_dots(const double *dot_x, const double *dot_y, long long ndots, const double *off_x, const double *off_y,
                            long long noff, double z, double radiance, float *x, float *y, float *zz, double *rad, int *diam) {
    g_error.clear();
    if (!dot_x || !dot_y || !off_x || !off_y || !x || !y || !zz || !rad || !diam) { g_error = "expand_dots: null argument"; return 1; }
    ph::expand_dots(dot_x, dot_y, ndots, off_x, off_y, noff, z, radiance, x, y, zz, rad, diam);
    return 0;
}
int photon_b200_piv_particles(long long n, const double *ux, const double *uy, const double *uz, const double lo[3],
                              const double hi[3], double gaussian_beam_fwhm, double irradiance_constant, const double rotation[9],
                              double z_object, float *x, float *y, float *z, double *rad, int *diam) {
    g_error.clear();
    if (!ux || !uy || !uz || !lo || !hi || !rotation || !x || !y || !z || !rad || !diam) { g_error = "piv_particles: null argument"; return 1; }
    ph::piv_particles(n, ux, uy, uz, lo, hi, gaussian_beam_fwhm, irradiance_constant, rotation, z_object, x, y, z, rad, diam);
    return 0;
}

// ---- Mie table (host; replaces bhmie.py and run_simulation_02.py:641-696) ----
int photon_b200_bhmie(double x, double refrel_re, double refrel_im, int nang, double *s1, double *s2, double *q4) {
    g_error.clear();
    if (nang < 2 || !s1 || !s2) { g_error = "bhmie: nang must be >= 2 and the outputs non-null"; return 1; }
    if (!(x > 0.0)) { g_error = "bhmie: size parameter must be positive"; return 1; }
    if (ph::bhmie_host(x, refrel_re, refrel_im, nang, s1, s2, q4) != 0) { g_error = "bhmie: nang > 1000 or nmx > 150000"; return 1; }
    return 0;
}
int photon_b200_mie_irradiance_table(const double *particle_sizes, int num_sizes, double medium_refractive_index,
                                     double particle_refractive_index, double beam_wavelength, int nang,
                                     double *scattering_angle, double *scattering_irradiance) {
    g_error.clear();
    if (nang < 2 || num_sizes < 1 || !particle_sizes || !scattering_angle || !scattering_irradiance) {
        g_error = "mie table: bad arguments"; return 1; }
    const int rows = 2 * nang - 1;
    std::vector<double> s1(2 * rows), s2(2 * rows);
    const double pi = 3.141592653589793;
    for (int c = 0; c < num_sizes; c++) {
        const double x = 2 * pi * particle_sizes[c] * medium_refractive_index / beam_wavelength;      // rs02.py:676
        if (!(x > 0.0) || ph::bhmie_host(x, particle_refractive_index / medium_refractive_index, 0.0, nang, s1.data(), s2.data(), nullptr) != 0) {
            g_error = "mie table: bhmie failed for size index " + std::to_string(c); return 1; }
        for (int r = 0; r < rows; r++)                                                                // rs02.py:690-692
            scattering_irradiance[(size_t)r * num_sizes + c] =
                0.5 * ((s1[2 * r] * s1[2 * r] + s1[2 * r + 1] * s1[2 * r + 1]) + (s2[2 * r] * s2[2 * r] + s2[2 * r + 1] * s2[2 * r + 1]));
    }
    const double dang = 0.5 * pi / (nang - 1);
    for (int r = 0; r < rows; r++) scattering_angle[r] = ((double)(r + 1) - 1.0) * dang;            // rs02.py:688
    return 0;
}

int photon_b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char *photon_b200_last_error(void) { return g_error.c_str(); }

static bool scene_fill(photon_scene *s, float lens_pitch, float image_distance,
                       const scattering_data_t *sc, const char *scattering_type_str,
                       const lightfield_source_t *src, int rpp, float beam_wavelength, float f_number,
                       int ne, const double (*element_center)[3], const element_data_t *element_data,
                       const double (*element_plane_parameters)[4], const int *element_system_index,
                       const camera_design_t *cam, bool simulate_density_gradients, const char *density_grad_filename,
                       int algorithm, bool add_pos_noise, float pos_noise_std, bool add_ngrad_noise,
                       float ngrad_noise_std, float cone_ratio, int shard_index, int shard_count) {
    CK(cudaSetDevice(s->device));
    s->pool = device_pool(s->device);
    CK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&s->ev0));
    CK(cudaEventCreate(&s->ev1));
    TraceArgs &a = s->args;
    memset(&a, 0, sizeof a);
    if (ne < 1 || ne > PH_MAX_ELEMENTS) { g_error = "num_elements must be in [1, 8]"; return false; }
    if (rpp < 1) { g_error = "lightray_number_per_particle must be >= 1"; return false; }

    // ---- source points (shard) : prt.cu:3162-3180
    int np = src->num_particles;
    if (shard_count < 1) { shard_count = 1; shard_index = 0; }
    long long b = (long long)np * shard_index / shard_count, e = (long long)np * (shard_index + 1) / shard_count;
    s->total_points = np; s->point_offset = (int)b; s->npoints = (int)(e - b);
    s->source_point_number = src->source_point_number;
    size_t n = (size_t)s->npoints;
    if (!upload(s->d_sx, src->x + b, n, s->stream, s->pool) || !upload(s->d_sy, src->y + b, n, s->stream, s->pool) ||
        !upload(s->d_sz, src->z + b, n, s->stream, s->pool) || !upload(s->d_srad, src->radiance + b, n, s->stream, s->pool) ||
        !upload(s->d_sdiam, src->diameter_index + b, n, s->stream, s->pool)) return false;
    a.sx = s->d_sx; a.sy = s->d_sy; a.sz = s->d_sz; a.sradiance = s->d_srad; a.sdiam = s->d_sdiam;
    a.rpp = rpp; a.global_point_offset = s->point_offset;
    // Processing order (north-star (b): Morton-ordered rays): consecutive CTAs take source points that are
    // neighbours in (x, y), so the field columns they march through share texels in L1/L2 whatever order the
    // caller generated the points in (random PIV particles: +7 % on C4).  Ray ids, dumps and dot statistics keep
    // the caller's order -- only the point a CTA picks changes -- and dump launches use the identity order.
    std::vector<int> order;              // lives until the stream is synchronised at the end of this function
    {
        const char *mo = getenv("PHOTON_MORTON");
        const bool want = mo ? atoi(mo) != 0 : simulate_density_gradients;
        if (want && n > 1) {
            order = morton_order(src->x + b, src->y + b, n);
            if (!upload(s->d_order, order.data(), n, s->stream, s->pool)) return false;
        }
    }

    // ---- Mie table : prt.cu:3192-3210
    a.mie.enabled = strcmp(scattering_type_str, "mie") == 0;
    if (a.mie.enabled) {
        if (!upload(s->d_angle, sc->scattering_angle, (size_t)sc->num_angles, s->stream, s->pool) ||
            !upload(s->d_irr, sc->scattering_irradiance, (size_t)sc->num_angles * sc->num_diameters, s->stream, s->pool)) return false;
        memcpy(a.mie.inv_rot, sc->inverse_rotation_matrix, sizeof a.mie.inv_rot);
        memcpy(a.mie.beam, sc->beam_propagation_vector, sizeof a.mie.beam);
        a.mie.angle = s->d_angle; a.mie.irradiance = s->d_irr;
        a.mie.num_angles = sc->num_angles; a.mie.num_diameters = sc->num_diameters;
    }

    // ---- scalar scene parameters, lens samples, optical train, camera (scene_args.h)
    std::vector<f2> aim;
    if (!fill_scene_args(a, aim, lens_pitch, image_distance, sc, scattering_type_str, src, rpp, beam_wavelength, f_number,
                         ne, element_center, element_data, element_plane_parameters, element_system_index, cam,
                         algorithm, add_pos_noise, pos_noise_std, add_ngrad_noise, ngrad_noise_std, cone_ratio, g_error))
        return false;
    std::vector<int> ray_order;
    if (!a.philox) {
        if (!upload(s->d_aim, aim.data(), aim.size(), s->stream, s->pool)) return false;
        // rays of a source point in Morton order of their aim points (PHOTON_RAY_ORDER, default on with a density field
        // and at least two warps of rays per point)
        const char *ro = getenv("PHOTON_RAY_ORDER");
        const bool want = ro ? atoi(ro) != 0 : (simulate_density_gradients && rpp >= 64);
        if (want) {
            morton_ray_order(aim, ray_order);
            if (!upload(s->d_ray_order, ray_order.data(), ray_order.size(), s->stream, s->pool)) return false;
            a.ray_order = s->d_ray_order;
        }
        CK(cudaStreamSynchronize(s->stream));    // aim and ray_order are local buffers
        a.aim = s->d_aim;
    }
    Camera &c = a.cam;

    // ---- density field : prt.cu:3323-3357
    a.simulate_gradients = simulate_density_gradients ? 1 : 0;
    if (simulate_density_gradients) {
        s->field = get_field(density_grad_filename, s->device, s->stream);
        if (!s->field) return false;
        a.field = s->field->P;
        // extension (BASELINE C5's step-size sweep; the reference always steps by the smallest spacing, trace.h:2086-2098)
        if (const char *e = getenv("PHOTON_STEP_SCALE")) {
            const double k = atof(e);
            if (k > 0.0) { a.field.step = (float)((double)a.field.step * k); a.field.skip_step = a.field.step / (1 + a.field.data_min); }
        }
        // coherent rays: the cone of a source point at the lens (cone_ratio * lens_pitch) is narrower than a cell
        s->coherent = rpp == 1 || cone_ratio * lens_pitch < a.field.step;
        if (const char *h = getenv("PHOTON_HOT")) s->coherent = atoi(h) >= 5;
        std::lock_guard<std::mutex> field_lock(s->field->mu);            // the on-demand tables below are added under it
        { const char *w = getenv("PHOTON_INDEX64");
          s->wide = (size_t)a.field.W * a.field.H * a.field.D > ((size_t)1 << 32) || (w && atoi(w) != 0); }
        { const char *pc = getenv("PHOTON_CELLS");
          s->use_cells = !s->wide && s->coherent && algorithm == 2 && !(pc && atoi(pc) == 0) && ensure_cells(*s->field, s->stream); }   // only rk4_hot reads it
        const char *smp = getenv("PHOTON_SAMPLER");
        s->use_texture = smp && strcmp(smp, "texture") == 0;
        if (s->use_texture && !ensure_texture(*s->field, s->stream)) return false;
        const char *ip = getenv("PHOTON_INTERPOLATION");
        s->use_cubic = ip && strcmp(ip, "cubic") == 0;
        if (s->use_cubic) {
            if (s->use_texture) { g_error = "PHOTON_INTERPOLATION=cubic is not available with PHOTON_SAMPLER=texture"; return false; }
            if (!ensure_coefficients(*s->field, s->stream)) return false;
            // experiment kept as an option: the hot loop of coherent RK4 scenes on TMA-staged bricks of the quad-row table
            { const char *bk = getenv("PHOTON_CUBIC_BRICK");
              s->use_brick = bk && atoi(bk) != 0 && s->coherent && algorithm == 2;
              if (s->use_brick && !ensure_brick_map(*s->field, s->stream)) return false; }
            // the spline of a constant region reproduces it to rounding only: keep the "below the global
            // minimum" fallback (trace.h:1056) from firing on the last bit
            a.field.data_min = a.field.data_min - 1e-4f * fabsf(a.field.data_min);
            a.field.skip_step = a.field.step / (1 + a.field.data_min);
        }
    }

    // ---- image + counters
    if (!pool_alloc(s->d_image, (size_t)c.nx * c.ny * sizeof(float), s->stream, s->pool)) return false;
    CK(cudaMemsetAsync(s->d_image, 0, (size_t)c.nx * c.ny * sizeof(float), s->stream));
    if (!pool_alloc(s->d_counters, 2 * sizeof(unsigned long long), s->stream, s->pool)) return false;
    if (getenv("PHOTON_DETERMINISTIC") && atoi(getenv("PHOTON_DETERMINISTIC")) != 0) {
        // Upper bound of the radiance one ray can deposit in one pixel, times ALL rays of the call
        // (not of the shard: every shard must use the same scale), must stay below 2^62.
        double rmax = 0.0;
        for (int i = 0; i < np; i++) { double r = fabs(src->radiance[i]); if (r > rmax) rmax = r; }
        if (a.mie.enabled) {
            double imax = 0.0;
            size_t nt = (size_t)sc->num_angles * sc->num_diameters;
            for (size_t i = 0; i < nt; i++) { double v = fabs((double)sc->scattering_irradiance[i]); if (v > imax) imax = v; }
            rmax *= imax;
        }
        double tr = 1.0;
        for (int k = 0; k < ne; k++) tr *= fmax(1.0, fabs((double)element_data[k].element_properties.transmission_ratio));
        double bound = rmax * (double)a.inv_fnum2 * tr * (double)np * (double)rpp;
        int e = 0;
        if (bound > 0 && std::isfinite(bound)) frexp(bound, &e);
        s->fixed_scale = ldexp(1.0, 62 - e - 1);
        size_t nb = (size_t)c.nx * c.ny * sizeof(unsigned long long);
        if (!pool_alloc(s->d_fixed, nb, s->stream, s->pool)) return false;
        CK(cudaMemsetAsync(s->d_fixed, 0, nb, s->stream));
    }
    a.image_fixed = s->d_fixed; a.fixed_scale = s->fixed_scale;
    CK(cudaMemsetAsync(s->d_counters, 0, 2 * sizeof(unsigned long long), s->stream));
    a.image = s->d_image; a.counters = s->d_counters;

    // persistent grid: a multiple of the SM count (148 on B200), 2 CTAs resident per SM
    int sm_count = 148;
    CK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, s->device));
    if (const char *o = getenv("PHOTON_OCC")) { s->occupancy = std::max(2, std::min(7, atoi(o))); s->occupancy_set = true; }
    // One source point per CTA up to 64 CTAs per SM-slot: the block scheduler balances the tail of a launch at the
    // granularity of one point (r1w: 12 CTAs per SM striding over the points lost 2.3 % to the last wave).
    int mult = 64;
    if (const char *g = getenv("PHOTON_GRID_MULT")) mult = std::max(1, atoi(g));
    s->grid_blocks = sm_count * mult;
    CK(cudaStreamSynchronize(s->stream));
    return true;
}

photon_scene *photon_scene_create(float lens_pitch, float image_distance,
                                  const scattering_data_t *sc, const char *scattering_type_str,
                                  const lightfield_source_t *src, int rpp, float beam_wavelength, float f_number,
                                  int ne, const double (*element_center)[3], const element_data_t *element_data,
                                  const double (*element_plane_parameters)[4], const int *element_system_index,
                                  const camera_design_t *cam, bool simulate_density_gradients,
                                  const char *density_grad_filename, int algorithm, bool add_pos_noise,
                                  float pos_noise_std, bool add_ngrad_noise, float ngrad_noise_std, float cone_ratio,
                                  int shard_index, int shard_count, int device) {
    g_error.clear();
    if (photon_b200_device_count() < 1) {
        g_error = "no CUDA device available (this library has no CPU path)";
        fprintf(stderr, "libparallel_ray_tracing(b200): %s\n", g_error.c_str());
        return nullptr;
    }
    photon_scene *s = new photon_scene();
    s->device = pick_device(device);
    if (!scene_fill(s, lens_pitch, image_distance, sc, scattering_type_str, src, rpp, beam_wavelength, f_number, ne,
                    element_center, element_data, element_plane_parameters, element_system_index, cam,
                    simulate_density_gradients, density_grad_filename, algorithm, add_pos_noise, pos_noise_std,
                    add_ngrad_noise, ngrad_noise_std, cone_ratio, shard_index, shard_count)) {
        if (!g_error.empty()) fprintf(stderr, "libparallel_ray_tracing(b200): %s\n", g_error.c_str());
        photon_scene_destroy(s);
        return nullptr;
    }
    return s;
}

void photon_scene_destroy(photon_scene *s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    cudaStream_t st = s->stream;
    pool_free(s->d_sx, st); pool_free(s->d_sy, st); pool_free(s->d_sz, st); pool_free(s->d_srad, st); pool_free(s->d_sdiam, st);
    pool_free(s->d_angle, st); pool_free(s->d_irr, st); pool_free(s->d_aim, st); pool_free(s->d_image, st); pool_free(s->d_counters, st);
    pool_free(s->d_fixed, st); pool_free(s->d_dots, st); pool_free(s->d_order, st); pool_free(s->d_ray_order, st);
    if (st) cudaStreamSynchronize(st);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->stream) cudaStreamDestroy(s->stream);
    put_field(s->field);   // the field stays cached until another one is loaded on this device or the cache is released
    delete s;
}

static bool scene_ok(photon_scene *s) { if (!s) { g_error = "null scene"; return false; } return cudaSetDevice(s->device) == cudaSuccess; }

int photon_scene_zero_image(photon_scene *s) {
    if (!scene_ok(s)) return 1;
    if (s->d_fixed) cudaMemsetAsync(s->d_fixed, 0, (size_t)s->args.cam.nx * s->args.cam.ny * sizeof(unsigned long long), s->stream);
    return cuda_ok(cudaMemsetAsync(s->d_image, 0, (size_t)s->args.cam.nx * s->args.cam.ny * sizeof(float), s->stream),
                   "zero image", __FILE__, __LINE__) ? 0 : 1;
}
int photon_scene_upload_image(photon_scene *s, const float *h) {
    if (!scene_ok(s)) return 1;
    const size_t bytes = (size_t)s->args.cam.nx * s->args.cam.ny * sizeof(float);
    // Through the thread's pinned staging buffer (DMA at full PCIe rate instead of the driver's pageable path).  The
    // buffer is shared by every scene this thread touches, so the copy is waited for before the call returns: no
    // other scene can overwrite the buffer while this copy is in flight.
    if (void *stage = g_stage.get(bytes)) {
        memcpy(stage, h, bytes);
        if (!cuda_ok(cudaMemcpyAsync(s->d_image, stage, bytes, cudaMemcpyHostToDevice, s->stream), "image H2D", __FILE__, __LINE__)) return 1;
        return cuda_ok(cudaStreamSynchronize(s->stream), "image H2D", __FILE__, __LINE__) ? 0 : 1;
    }
    return cuda_ok(cudaMemcpyAsync(s->d_image, h, bytes, cudaMemcpyHostToDevice, s->stream), "image H2D", __FILE__, __LINE__) ? 0 : 1;
}
static bool fold_fixed(photon_scene *s) {
    if (!s->d_fixed) return true;
    size_t n = (size_t)s->args.cam.nx * s->args.cam.ny;
    fold_fixed_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 8), 256, 0, s->stream>>>(s->d_image, s->d_fixed, 1.0 / s->fixed_scale, n);
    return cuda_ok(cudaGetLastError(), "fold_fixed_kernel", __FILE__, __LINE__);
}
int photon_scene_deterministic(photon_scene *s) { return s && s->d_fixed ? 1 : 0; }
int photon_scene_download_image_fixed(photon_scene *s, long long *h, double *scale) {
    if (!scene_ok(s)) return 1;
    if (!s->d_fixed) { g_error = "scene was not created with PHOTON_DETERMINISTIC=1"; return 1; }
    if (scale) *scale = s->fixed_scale;
    size_t bytes = (size_t)s->args.cam.nx * s->args.cam.ny * sizeof(long long);
    if (!cuda_ok(cudaMemcpyAsync(h, s->d_fixed, bytes, cudaMemcpyDeviceToHost, s->stream), "fixed D2H", __FILE__, __LINE__)) return 1;
    return cuda_ok(cudaStreamSynchronize(s->stream), "sync", __FILE__, __LINE__) ? 0 : 1;
}
int photon_scene_download_image(photon_scene *s, float *h) {
    if (!scene_ok(s)) return 1;
    if (!fold_fixed(s)) return 1;
    const size_t bytes = (size_t)s->args.cam.nx * s->args.cam.ny * sizeof(float);
    void *stage = g_stage.get(bytes);
    if (!cuda_ok(cudaMemcpyAsync(stage ? stage : (void *)h, s->d_image, bytes, cudaMemcpyDeviceToHost, s->stream), "image D2H", __FILE__, __LINE__)) return 1;
    if (!cuda_ok(cudaStreamSynchronize(s->stream), "sync", __FILE__, __LINE__)) return 1;
    if (stage) memcpy(h, stage, bytes);
    return 0;
}
int photon_scene_export_u16(photon_scene *s, float pixel_gain, int pixel_bit_depth, int intensity_rescaling,
                            float image_noise, unsigned int seed, unsigned short *host_out) {
    if (!scene_ok(s)) return 1;
    if (!host_out || pixel_bit_depth < 1 || pixel_bit_depth > 16) { g_error = "export_u16: bad arguments"; return 1; }
    if (!fold_fixed(s)) return 1;
    const size_t n = (size_t)s->args.cam.nx * s->args.cam.ny;
    float *tmp = nullptr; unsigned short *d16 = nullptr; unsigned int *dmax = nullptr;
    bool ok = pool_alloc(tmp, n * sizeof(float), s->stream, s->pool) && pool_alloc(d16, n * sizeof(unsigned short), s->stream, s->pool) &&
              pool_alloc(dmax, sizeof(unsigned int), s->stream, s->pool);
    if (ok) {
        const float gain = (float)std::pow(10.0, (double)pixel_gain / 20.0);
        const float levels = (float)((1 << pixel_bit_depth) - 1);
        const float k = (float)((65536.0 - 1.0) / ((double)(1 << pixel_bit_depth) - 1.0));
        const unsigned blocks = (unsigned)std::min<size_t>((n + 255) / 256, 148 * 8);
        ok = cuda_ok(cudaMemsetAsync(dmax, 0, sizeof(unsigned int), s->stream), "export", __FILE__, __LINE__);
        if (ok) {
            export_prepare_kernel<<<blocks, 256, 0, s->stream>>>(s->d_image, tmp, n, gain, image_noise * 100.0f, seed,
                                                                 intensity_rescaling, dmax);
            export_u16_kernel<<<blocks, 256, 0, s->stream>>>(tmp, d16, n, levels, k, intensity_rescaling, dmax);
            ok = cuda_ok(cudaGetLastError(), "export kernels", __FILE__, __LINE__) &&
                 cuda_ok(cudaMemcpyAsync(host_out, d16, n * sizeof(unsigned short), cudaMemcpyDeviceToHost, s->stream), "export D2H", __FILE__, __LINE__) &&
                 cuda_ok(cudaStreamSynchronize(s->stream), "sync", __FILE__, __LINE__);
        }
    }
    pool_free(tmp, s->stream); pool_free(d16, s->stream); pool_free(dmax, s->stream);
    return ok ? 0 : 1;
}
void *photon_scene_image_device_ptr(photon_scene *s) { return s ? (void *)s->d_image : nullptr; }

int photon_scene_trace(photon_scene *s) {
    if (!scene_ok(s)) return 1;
    s->last_launches = 0;
    if (!cuda_ok(cudaEventRecord(s->ev0, s->stream), "event", __FILE__, __LINE__)) return 1;
    if (!launch_trace(s, 0, s->npoints, 0, nullptr, nullptr, 0, nullptr, nullptr, 0)) return 1;
    if (!cuda_ok(cudaEventRecord(s->ev1, s->stream), "event", __FILE__, __LINE__)) return 1;
    return 0;
}

int photon_scene_trace_range(photon_scene *s, int point_begin, int point_count) {
    if (!scene_ok(s)) return 1;
    if (point_begin < 0 || point_count < 0 || point_begin + point_count > s->npoints) { g_error = "point range outside the shard"; return 1; }
    s->last_launches = 0;
    if (!cuda_ok(cudaEventRecord(s->ev0, s->stream), "event", __FILE__, __LINE__)) return 1;
    if (!launch_trace(s, point_begin, point_count, point_begin, nullptr, nullptr, 0, nullptr, nullptr, 0)) return 1;
    if (!cuda_ok(cudaEventRecord(s->ev1, s->stream), "event", __FILE__, __LINE__)) return 1;
    return 0;
}

int photon_scene_sync(photon_scene *s) {
    if (!scene_ok(s)) return 1;
    if (!cuda_ok(cudaStreamSynchronize(s->stream), "trace kernel", __FILE__, __LINE__)) return 1;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, s->ev0, s->ev1) == cudaSuccess) s->last_ms = ms; else cudaGetLastError();
    return 0;
}

int photon_scene_trace_batch(photon_scene *s, int batch, long long n_save, float *final_pos, float *final_dir,
                             int n_inter, float *inter_pos, float *inter_dir) {
    if (!scene_ok(s)) return 1;
    int spn = s->source_point_number;
    if (s->npoints < spn) spn = s->npoints;                              // prt.cu:3366-3368
    if (spn <= 0) return 0;
    int begin = batch * spn, count = std::min(spn, s->npoints - begin);
    if (count <= 0 || n_save < 0) return 0;
    f3 *dpos = nullptr, *ddir = nullptr, *dipos = nullptr, *didir = nullptr;
    bool ok = true;
    size_t nb = (size_t)n_save * sizeof(f3), nib = nb * (size_t)std::max(n_inter, 0);
    const bool inter = s->args.simulate_gradients && n_inter > 0 && inter_pos && inter_dir && n_save > 0;
    auto fail = [&]() { cudaFree(dpos); cudaFree(ddir); cudaFree(dipos); cudaFree(didir); return 1; };
    if (n_save > 0) {
        ok = cuda_ok(cudaMalloc(&dpos, nb), "dump alloc", __FILE__, __LINE__) && cuda_ok(cudaMalloc(&ddir, nb), "dump alloc", __FILE__, __LINE__);
        if (!ok) return fail();
        cudaMemsetAsync(dpos, 0xff, nb, s->stream); cudaMemsetAsync(ddir, 0xff, nb, s->stream);   // all-ones = NaN (prt.cu:3527-3533)
    }
    if (inter) {
        ok = cuda_ok(cudaMalloc(&dipos, nib), "dump alloc", __FILE__, __LINE__) && cuda_ok(cudaMalloc(&didir, nib), "dump alloc", __FILE__, __LINE__);
        if (!ok) return fail();
        cudaMemsetAsync(dipos, 0xff, nib, s->stream); cudaMemsetAsync(didir, 0xff, nib, s->stream);
    }
    s->last_launches = 0;
    cudaEventRecord(s->ev0, s->stream);
    if (!launch_trace(s, begin, count, begin, dpos, ddir, n_save, dipos, didir, inter ? n_inter : 0)) return fail();
    cudaEventRecord(s->ev1, s->stream);
    if (n_save > 0 && final_pos) cudaMemcpyAsync(final_pos, dpos, nb, cudaMemcpyDeviceToHost, s->stream);
    if (n_save > 0 && final_dir) cudaMemcpyAsync(final_dir, ddir, nb, cudaMemcpyDeviceToHost, s->stream);
    if (inter) {
        cudaMemcpyAsync(inter_pos, dipos, nib, cudaMemcpyDeviceToHost, s->stream);
        cudaMemcpyAsync(inter_dir, didir, nib, cudaMemcpyDeviceToHost, s->stream);
    }
    int rc = photon_scene_sync(s);
    cudaFree(dpos); cudaFree(ddir); cudaFree(dipos); cudaFree(didir);
    return rc;
}

static bool fetch_counters(photon_scene *s) {
    if (!scene_ok(s)) return false;
    CK(cudaMemcpyAsync(s->h_counters, s->d_counters, sizeof s->h_counters, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    return true;
}
int photon_scene_reset_counters(photon_scene *s) {
    if (!scene_ok(s)) return 1;
    s->rays_traced = 0; s->total_launches = 0;
    return cuda_ok(cudaMemsetAsync(s->d_counters, 0, 2 * sizeof(unsigned long long), s->stream), "reset", __FILE__, __LINE__) ? 0 : 1;
}
// ---- dot statistics (light_ray_processing.py:211-275 on the device; include/photon_b200.h) ----
int photon_scene_enable_dot_stats(photon_scene *s, long long rays_per_dot) {
    if (!scene_ok(s)) return 1;
    if (rays_per_dot < 1) { g_error = "rays_per_dot must be >= 1"; return 1; }
    const long long rays = (long long)s->npoints * s->args.rpp;
    const long long n = (rays + rays_per_dot - 1) / rays_per_dot;
    if (s->d_dots) { pool_free(s->d_dots, s->stream); s->d_dots = nullptr; }
    if (!pool_alloc(s->d_dots, (size_t)n * 8 * sizeof(double), s->stream, s->pool)) return 1;
    s->n_dots = n; s->rays_per_dot = rays_per_dot;
    return cuda_ok(cudaMemsetAsync(s->d_dots, 0, (size_t)n * 8 * sizeof(double), s->stream), "dot stats", __FILE__, __LINE__) ? 0 : 1;
}
long long photon_scene_num_dots(photon_scene *s) { return s ? s->n_dots : -1; }
int photon_scene_reset_dot_stats(photon_scene *s) {
    if (!scene_ok(s)) return 1;
    if (!s->d_dots) { g_error = "dot statistics are not enabled"; return 1; }
    return cuda_ok(cudaMemsetAsync(s->d_dots, 0, (size_t)s->n_dots * 8 * sizeof(double), s->stream), "dot stats", __FILE__, __LINE__) ? 0 : 1;
}
int photon_scene_download_dot_stats(photon_scene *s, double *host, long long capacity_dots) {
    if (!scene_ok(s)) return 1;
    if (!s->d_dots) { g_error = "dot statistics are not enabled"; return 1; }
    if (capacity_dots < s->n_dots) { g_error = "dot statistics buffer too small"; return 1; }
    if (!cuda_ok(cudaMemcpyAsync(host, s->d_dots, (size_t)s->n_dots * 8 * sizeof(double), cudaMemcpyDeviceToHost, s->stream), "dot stats", __FILE__, __LINE__)) return 1;
    return cuda_ok(cudaStreamSynchronize(s->stream), "dot stats", __FILE__, __LINE__) ? 0 : 1;
}
long long photon_scene_ray_steps(photon_scene *s) { return fetch_counters(s) ? (long long)s->h_counters[0] : -1; }
long long photon_scene_sensor_hits(photon_scene *s) { return fetch_counters(s) ? (long long)s->h_counters[1] : -1; }
long long photon_scene_rays(photon_scene *s) { return s ? s->rays_traced : -1; }
float photon_scene_last_trace_ms(photon_scene *s) { return s ? s->last_ms : -1.0f; }
int photon_scene_kernel_launches(photon_scene *s) { return s ? s->last_launches : -1; }
int photon_scene_num_points(photon_scene *s) { return s ? s->npoints : -1; }

int photon_b200_load_field(const char *path, float meta16[16], float *out, long long capacity_voxels) {
    g_error.clear();
    if (photon_b200_device_count() < 1) { g_error = "no CUDA device available"; return 1; }
    int dev = pick_device(-1);
    if (cudaSetDevice(dev) != cudaSuccess) return 1;
    cudaStream_t st;
    if (cudaStreamCreate(&st) != cudaSuccess) return 1;
    DeviceField *F = get_field(path, dev, st);
    int rc = 1;
    if (F) {
        const FieldParams &P = F->P;
        float m[16] = {P.mn.x, P.mn.y, P.mn.z, P.mx.x, P.mx.y, P.mx.z, 0, 0, 0, P.step, P.data_min,
                       (float)P.W, (float)P.H, (float)P.D, 0, 0};
        m[6] = F->spacing.x; m[7] = F->spacing.y; m[8] = F->spacing.z;
        memcpy(meta16, m, sizeof m);
        size_t n = (size_t)P.W * P.H * P.D;
        rc = 0;
        if (out && capacity_voxels >= (long long)n)
            rc = cudaMemcpy(out, F->lin, n * sizeof(f4), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : 1;
    }
    put_field(F);
    cudaStreamDestroy(st);
    return rc;
}

int photon_b200_load_coefficients(const char *path, float *out, long long capacity_voxels) {
    g_error.clear();
    if (photon_b200_device_count() < 1) { g_error = "no CUDA device available"; return 1; }
    int dev = pick_device(-1);
    if (cudaSetDevice(dev) != cudaSuccess) return 1;
    cudaStream_t st;
    if (cudaStreamCreate(&st) != cudaSuccess) return 1;
    DeviceField *F = get_field(path, dev, st);
    int rc = 1;
    bool have = false;
    if (F) { std::lock_guard<std::mutex> lk(F->mu); have = ensure_coefficients(*F, st); }
    if (have) {
        size_t n = (size_t)F->P.W * F->P.H * F->P.D;
        rc = 0;
        if (out && capacity_voxels >= (long long)n)
            rc = cudaMemcpy(out, F->coef, n * sizeof(float), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : 1;
    }
    put_field(F);
    cudaStreamDestroy(st);
    return rc;
}

int photon_b200_sample_field(const char *path, int cubic, int n, const float *xyz_texel, float *out4) {
    g_error.clear();
    if (photon_b200_device_count() < 1) { g_error = "no CUDA device available"; return 1; }
    if (n <= 0) return 0;
    int dev = pick_device(-1);
    if (cudaSetDevice(dev) != cudaSuccess) return 1;
    cudaStream_t st;
    if (cudaStreamCreate(&st) != cudaSuccess) return 1;
    DeviceField *F = get_field(path, dev, st);
    int rc = 1;
    float *d_in = nullptr, *d_out = nullptr;
    bool have = F != nullptr;
    if (have && cubic) { std::lock_guard<std::mutex> lk(F->mu); have = cubic == 2 ? ensure_texture(*F, st) : ensure_coefficients(*F, st); }
    if (have &&
        cudaMalloc(&d_in, (size_t)n * 3 * sizeof(float)) == cudaSuccess &&
        cudaMalloc(&d_out, (size_t)n * 4 * sizeof(float)) == cudaSuccess &&
        cudaMemcpyAsync(d_in, xyz_texel, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, st) == cudaSuccess) {
        int blocks = (n + 127) / 128;
        if (cubic == 2) {
            TextureField S; S.tex = F->tex;
            sample_field_kernel<<<blocks, 128, 0, st>>>(S, n, d_in, d_out);
        } else if (cubic) {
            BSplineField S; S.c = F->coef; S.quad = reinterpret_cast<const f4 *>(F->coefq); S.W = F->P.W; S.H = F->P.H; S.D = F->P.D;
            S.inv_sp = mk3(1.0f / F->spacing.x, 1.0f / F->spacing.y, 1.0f / F->spacing.z);
            sample_field_kernel<<<blocks, 128, 0, st>>>(S, n, d_in, d_out);
        } else {
            LinearField S; S.g = F->lin; S.W = F->P.W; S.H = F->P.H; S.D = F->P.D;
            sample_field_kernel<<<blocks, 128, 0, st>>>(S, n, d_in, d_out);
        }
        if (cudaGetLastError() == cudaSuccess &&
            cudaMemcpyAsync(out4, d_out, (size_t)n * 4 * sizeof(float), cudaMemcpyDeviceToHost, st) == cudaSuccess &&
            cudaStreamSynchronize(st) == cudaSuccess) rc = 0;
    }
    if (rc) { cudaError_t e = cudaGetLastError(); if (e != cudaSuccess && g_error.empty()) g_error = cudaGetErrorString(e); }
    cudaFree(d_in); cudaFree(d_out);
    put_field(F);
    cudaStreamDestroy(st);
    return rc;
}

// Drop what the library keeps between calls: every cached density field nobody is using (texels, cell table,
// spline coefficients, texture array -- 18 GiB per device for a 512^3 field with its cell table) and the memory
// its stream-ordered pools hold.  Fields of live scenes stay.  Returns the number of bytes of field data released.
long long photon_b200_release_cache(void) {
    std::vector<DeviceField *> drop;
    {
        std::lock_guard<std::mutex> lk(g_field_mutex);
        for (size_t i = 0; i < g_fields.size();) {
            if (g_fields[i]->refs == 0) { drop.push_back(g_fields[i]); g_fields.erase(g_fields.begin() + i); } else i++;
        }
    }
    long long freed = 0;
    for (auto *f : drop) { { std::lock_guard<std::mutex> w(f->mu); freed += (long long)f->bytes(); f->release(); } delete f; }
    {
        std::lock_guard<std::mutex> lk(g_pool_mutex);
        for (auto &e : g_pools) if (e.second) { cudaSetDevice(e.first); cudaDeviceSynchronize(); cudaMemPoolTrimTo(e.second, 0); }
    }
    cudaGetLastError();
    return freed;
}
// Bytes of field data currently resident in the cache, over all devices (held or idle).
long long photon_b200_cache_bytes(void) {
    std::lock_guard<std::mutex> lk(g_field_mutex);
    long long b = 0;
    for (auto *f : g_fields) b += (long long)f->bytes();
    return b;
}
// Which trace kernel the last launch of this host thread ran, e.g. "rk4/cells/dump" (tests assert on it).
const char *photon_b200_last_variant(void) { return g_variant.c_str(); }
// brick refills of the TMA-staged cubic loop whose mbarrier wait gave up, on the current device (0 unless something is broken)
int photon_b200_brick_timeouts(void) {
    unsigned int v = 0;
    if (cudaMemcpyFromSymbol(&v, ph_brick_timeouts, sizeof v) != cudaSuccess) { cudaGetLastError(); return -1; }
    return (int)v;
}

}  // extern "C"

// ---------------------------------------------------------------------------------
// The entry point photon's Python binds (include/photon_abi.h)
// ---------------------------------------------------------------------------------
namespace {

bool write_f3_file(const std::string &dir, const char *prefix, int k, const float *data, size_t count_f3) {
    char name[64];
    snprintf(name, sizeof name, "/%s%04d.bin", prefix, k);               // prt.cu:3566-3571
    std::string path = dir + name;
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { fprintf(stderr, "libparallel_ray_tracing(b200): cannot write %s\n", path.c_str()); return false; }
    if (count_f3) fwrite(data, 3 * sizeof(float), count_f3, f);
    fclose(f);
    return true;
}

struct ShardResult { bool ok = false; photon_scene *scene = nullptr; std::vector<float> image; std::vector<long long> fixed; double scale = 0; };

// The library's OWN image reduce over NVLink peer memory -- the default route of a PHOTON_GPUS call.  The partial
// images live in the per-device pools of this library; device 0 is granted access to the other pools
// (cudaMemPoolSetAccess, once per pool) and ONE kernel on device 0 reads every peer image through NVLink / NVSwitch and
// adds them in device order onto its own (float sum in a fixed order: the same image for the same shards, run after
// run; deterministic mode: exact 64-bit integer sum).  8 x 4 MiB cross the switch in tens of microseconds, and nothing
// has to be set up first: the NCCL communicators this replaces cost the first call of a process seconds
// (ncclCommInitAll + the first collective, contending with context creation on the other devices: the cold 8-GPU call
// of r4k waited 1 s for them after its last device had finished tracing).  PHOTON_REDUCE=nccl keeps that route,
// PHOTON_REDUCE=host sums through the host.
#define PH_MAX_PEERS 16
struct PeerImages { const void *p[PH_MAX_PEERS]; int n; };
template <class T>
__global__ void __launch_bounds__(256) peer_sum_kernel(T *__restrict__ dst, const PeerImages src, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        T acc = dst[i];
        for (int g = 0; g < src.n; g++) acc += static_cast<const T *>(src.p[g])[i];
        dst[i] = acc;
    }
}
std::mutex g_peer_mutex;
std::vector<int> g_peer_granted;         // devices whose pool device 0 may read
// 0 = reduced onto device 0, 1 = peer access not available (nothing was enqueued)
// Let device `reader` read the pool of scene s's device (once per pool).  Called by each device thread of a PHOTON_GPUS
// call as soon as its scene exists -- the grants (tens of ms each in a fresh process: the driver sets up the peer
// mapping) then run side by side behind the field builds instead of one after the other in front of the reduce (the cold
// 8-GPU call of r6b spent 195 ms there) -- and again, as a check, by the reduce itself.
bool grant_peer_access(photon_scene *s, int reader) {
    if (s->device == reader) return true;
    {
        std::lock_guard<std::mutex> lk(g_peer_mutex);
        for (int d : g_peer_granted) if (d == s->device) return true;
    }
    int can = 0;
    if (!s->pool || cudaDeviceCanAccessPeer(&can, reader, s->device) != cudaSuccess || !can) { cudaGetLastError(); return false; }
    cudaMemAccessDesc d; memset(&d, 0, sizeof d);
    d.location.type = cudaMemLocationTypeDevice; d.location.id = reader; d.flags = cudaMemAccessFlagsProtReadWrite;
    if (cudaMemPoolSetAccess(s->pool, &d, 1) != cudaSuccess) { cudaGetLastError(); return false; }
    std::lock_guard<std::mutex> lk(g_peer_mutex);
    g_peer_granted.push_back(s->device);
    return true;
}
int peer_reduce_images(std::vector<ShardResult> &res, int n, size_t npix) {
    if (n - 1 > PH_MAX_PEERS) return 1;
    photon_scene *s0 = res[0].scene;
    const bool fixed = s0->d_fixed != nullptr;
    for (int g = 1; g < n; g++) if (!grant_peer_access(res[g].scene, s0->device)) return 1;
    PeerImages src; src.n = n - 1;
    for (int g = 1; g < n; g++) src.p[g - 1] = fixed ? (const void *)res[g].scene->d_fixed : (const void *)res[g].scene->d_image;
    if (cudaSetDevice(s0->device) != cudaSuccess) { cudaGetLastError(); return 1; }
    const int blocks = (int)std::min<size_t>((npix + 255) / 256, (size_t)148 * 8);
    if (fixed) peer_sum_kernel<unsigned long long><<<blocks, 256, 0, s0->stream>>>(s0->d_fixed, src, npix);
    else peer_sum_kernel<float><<<blocks, 256, 0, s0->stream>>>(s0->d_image, src, npix);
    if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(s0->stream) != cudaSuccess) {
        cudaGetLastError();
        fprintf(stderr, "libparallel_ray_tracing(b200): peer reduce failed\n");
        return 2;                                            // device 0's image may be partly summed: the caller gives up
    }
    return 0;
}

// NCCL for the one exchange step the path has: the partial sensor images of a PHOTON_GPUS call are combined
// with ONE ncclReduce over NVLink onto device 0.  The library is bound at run time (dlopen of
// libnccl.so.2: no link-time dependency; without it the images are summed on the host, which is only a
// different route for the same 4 MiB, not a compute fallback).  Communicators are created once per device
// count with ncclCommInitAll and kept for the life of the process (no teardown at exit: the CUDA context may be
// gone by then).
#if PH_HAVE_NCCL
struct NcclApi {
    void *h = nullptr; bool tried = false;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::vector<ncclComm_t> comms; int ncomms = 0;
    std::mutex mu;
    bool load() {
        if (tried) return h != nullptr;
        tried = true;
        if (const char *e = getenv("PHOTON_NCCL")) if (atoi(e) == 0) return false;
        // PHOTON_NCCL_LIB names the file (photon_b200/lib.py points it at the NCCL that PyTorch bundles, when there is one:
        // two different builds of libnccl.so.2 cannot live in one process, and whichever is loaded first would break
        // the other's user -- found on an 8-GPU box, where `import torch` failed after this library had loaded the
        // system's NCCL)
        if (const char *f = getenv("PHOTON_NCCL_LIB")) if (*f) h = dlopen(f, RTLD_NOW | RTLD_LOCAL);
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) { if (h) break; h = dlopen(name, RTLD_NOW | RTLD_LOCAL); }
        if (!h) return false;
        CommInitAll = (decltype(CommInitAll))dlsym(h, "ncclCommInitAll");
        GroupStart = (decltype(GroupStart))dlsym(h, "ncclGroupStart");
        GroupEnd = (decltype(GroupEnd))dlsym(h, "ncclGroupEnd");
        Reduce = (decltype(Reduce))dlsym(h, "ncclReduce");
        GetErrorString = (decltype(GetErrorString))dlsym(h, "ncclGetErrorString");
        if (!CommInitAll || !GroupStart || !GroupEnd || !Reduce) { dlclose(h); h = nullptr; }
        return h != nullptr;
    }
    bool ensure(int n) {
        if (!load()) return false;
        if (ncomms == n) return true;
        if (ncomms != 0) return false;                 // a different device count than the cached communicators
        std::vector<int> devs(n); for (int i = 0; i < n; i++) devs[i] = i;
        comms.resize(n);
        ncclResult_t r = CommInitAll(comms.data(), n, devs.data());
        if (r != ncclSuccess) { fprintf(stderr, "libparallel_ray_tracing(b200): ncclCommInitAll: %s; summing on the host\n", GetErrorString ? GetErrorString(r) : "?"); comms.clear(); return false; }
        ncomms = n;
        return true;
    }
};
NcclApi g_nccl;
// Sum the shard images onto scene 0 of device 0.  float mode: ncclFloat sum of the images; deterministic mode:
// ncclUint64 sum of the fixed-point images (exact, order-free), folded into the float image once on device 0.
// Returns 0 = reduced, 1 = NCCL not available (nothing was enqueued: the caller may sum through the host),
// 2 = failed after the reduce was enqueued (device 0's image may be partly reduced: the caller must give up).
int nccl_reduce_images(std::vector<ShardResult> &res, int n, size_t npix) {
    std::lock_guard<std::mutex> lk(g_nccl.mu);
    if (!g_nccl.ensure(n)) return 1;
    const bool fixed = res[0].scene->d_fixed != nullptr;
    ncclResult_t r = g_nccl.GroupStart();
    for (int g = 0; g < n && r == ncclSuccess; g++) {
        photon_scene *s = res[g].scene;
        cudaSetDevice(s->device);
        if (fixed) r = g_nccl.Reduce(s->d_fixed, s->d_fixed, npix, ncclUint64, ncclSum, 0, g_nccl.comms[g], s->stream);
        else r = g_nccl.Reduce(s->d_image, s->d_image, npix, ncclFloat, ncclSum, 0, g_nccl.comms[g], s->stream);
    }
    ncclResult_t e = g_nccl.GroupEnd();
    if (r == ncclSuccess) r = e;
    if (r != ncclSuccess) { fprintf(stderr, "libparallel_ray_tracing(b200): ncclReduce: %s\n", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?"); return 2; }
    for (int g = 0; g < n; g++) { cudaSetDevice(res[g].scene->device); if (cudaStreamSynchronize(res[g].scene->stream) != cudaSuccess) { cudaGetLastError(); return 2; } }
    return 0;
}
// Communicator set-up AND the first collective (which loads NCCL's kernels and connects its channels: another
// ~0.5 s on top of ncclCommInitAll, profiles r4g) on a throw-away buffer, so that the reduce of the first call costs
// what it costs in every later call.  Runs on its own thread beside the field builds.
void nccl_prepare(int n) {
    std::lock_guard<std::mutex> lk(g_nccl.mu);
    const bool first = g_nccl.ncomms == 0;
    if (!g_nccl.ensure(n) || !first) return;
    std::vector<float *> buf((size_t)n, nullptr);
    std::vector<cudaStream_t> st((size_t)n, nullptr);
    bool ok = true;
    for (int g = 0; g < n && ok; g++)
        ok = cudaSetDevice(g) == cudaSuccess && cudaStreamCreateWithFlags(&st[g], cudaStreamNonBlocking) == cudaSuccess &&
             cudaMalloc(&buf[g], 256 * sizeof(float)) == cudaSuccess && cudaMemsetAsync(buf[g], 0, 256 * sizeof(float), st[g]) == cudaSuccess;
    if (ok) {
        g_nccl.GroupStart();
        for (int g = 0; g < n; g++) { cudaSetDevice(g); g_nccl.Reduce(buf[g], buf[g], 256, ncclFloat, ncclSum, 0, g_nccl.comms[g], st[g]); }
        g_nccl.GroupEnd();
    }
    for (int g = 0; g < n; g++) {
        cudaSetDevice(g);
        if (st[g]) { cudaStreamSynchronize(st[g]); cudaStreamDestroy(st[g]); }
        if (buf[g]) cudaFree(buf[g]);
    }
    cudaGetLastError();
}
#else
int nccl_reduce_images(std::vector<ShardResult> &, int, size_t) { return 1; }
void nccl_prepare(int) {}
#endif

}  // namespace

extern "C" void start_ray_tracing(float lens_pitch, float image_distance,
                                  scattering_data_t *scattering_data_p, char *scattering_type_str,
                                  lightfield_source_t *lightfield_source_p, int lightray_number_per_particle,
                                  float beam_wavelength, float aperture_f_number,
                                  int num_elements, double (*element_center)[3], element_data_t *element_data_p,
                                  double (*element_plane_parameters)[4], int *element_system_index,
                                  camera_design_t *camera_design_p, float *image_array,
                                  bool simulate_density_gradients, char *density_grad_filename,
                                  bool save_lightrays, char *lightray_position_save_path,
                                  char *lightray_direction_save_path,
                                  int num_lightrays_save, int ray_tracing_algorithm,
                                  bool add_pos_noise, float pos_noise_std,
                                  bool add_ngrad_noise, float ngrad_noise_std,
                                  float ray_cone_pitch_ratio,
                                  bool save_intermediate_ray_data, int num_intermediate_positions_save) {
    g_error.clear();
    const int np = lightfield_source_p->num_particles;
    VLOG("number of particles: %d\n", np);
    int shard_index = 0, shard_count = 1;
    if (const char *sh = getenv("PHOTON_SHARD")) {
        if (sscanf(sh, "%d/%d", &shard_index, &shard_count) != 2 || shard_count < 1 || shard_index < 0 || shard_index >= shard_count) {
            fprintf(stderr, "libparallel_ray_tracing(b200): bad PHOTON_SHARD \"%s\" (want r/n)\n", sh);
            return;
        }
    }
    int ngpus = 1;
    if (const char *g = getenv("PHOTON_GPUS")) ngpus = std::max(1, std::min(atoi(g), photon_b200_device_count()));
    const bool dumps = save_lightrays && num_lightrays_save > 0;
    const bool empty_dump_files = save_lightrays && num_lightrays_save <= 0;     // the reference writes empty files then
    const bool inter = simulate_density_gradients && save_intermediate_ray_data && num_intermediate_positions_save > 0 && num_lightrays_save > 0;
    if (dumps || inter) ngpus = 1;       // dump files are per batch of the whole point list
    const size_t npix = (size_t)camera_design_p->x_pixel_number * camera_design_p->y_pixel_number;

    auto make_scene = [&](int si, int sc, int dev) {
        return photon_scene_create(lens_pitch, image_distance, scattering_data_p, scattering_type_str, lightfield_source_p,
                                   lightray_number_per_particle, beam_wavelength, aperture_f_number, num_elements,
                                   element_center, element_data_p, element_plane_parameters, element_system_index,
                                   camera_design_p, simulate_density_gradients, density_grad_filename, ray_tracing_algorithm,
                                   add_pos_noise, pos_noise_std, add_ngrad_noise, ngrad_noise_std, ray_cone_pitch_ratio,
                                   si, sc, dev);
    };

    const bool timing = getenv("PHOTON_TIMING") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[photon timing] %-16s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    if (ngpus == 1) {
        photon_scene *s = make_scene(shard_index, shard_count, -1);
        if (!s) return;                                                  // image untouched
        lap("scene create");
        bool ok = photon_scene_upload_image(s, image_array) == 0;        // accumulate onto the caller's image (prt.cu:3309)
        lap("image H2D");
        // batches of the points THIS scene holds (all of them, or the PHOTON_SHARD range): prt.cu:3506-3510
        const int held = photon_scene_num_points(s);
        const int spn = std::min(lightfield_source_p->source_point_number, held);
        const int kmax = spn > 0 ? (held + spn - 1) / spn : 0;
        if (ok && (dumps || inter)) {
            size_t ns = (size_t)std::max(num_lightrays_save, 0);
            const float qnan = std::numeric_limits<float>::quiet_NaN();
            std::vector<float> fpos(3 * ns, qnan), fdir(3 * ns, qnan), ipos, idir;   // prt.cu:3527-3533
            if (inter) { ipos.assign(3 * ns * num_intermediate_positions_save, qnan); idir.assign(ipos.size(), qnan); }
            for (int k = 0; k < kmax && ok; k++) {
                VLOG("%d out of %d\n", k + 1, kmax);
                ok = photon_scene_trace_batch(s, k, (long long)ns, fpos.data(), fdir.data(),
                                              inter ? num_intermediate_positions_save : 0,
                                              inter ? ipos.data() : nullptr, inter ? idir.data() : nullptr) == 0;
                if (ok && dumps) {
                    write_f3_file(lightray_position_save_path, "pos_", k, fpos.data(), ns);
                    write_f3_file(lightray_direction_save_path, "dir_", k, fdir.data(), ns);
                }
                if (ok && inter) {                                       // prt.cu:3613-3670
                    write_f3_file(lightray_position_save_path, "intermediate_pos_", k, ipos.data(), ns * num_intermediate_positions_save);
                    write_f3_file(lightray_direction_save_path, "intermediate_dir_", k, idir.data(), ns * num_intermediate_positions_save);
                }
            }
        } else if (ok) {
            ok = photon_scene_trace(s) == 0 && photon_scene_sync(s) == 0;
            for (int k = 0; k < kmax && ok && empty_dump_files; k++) {    // save_lightrays with nothing to save: empty files
                write_f3_file(lightray_position_save_path, "pos_", k, nullptr, 0);
                write_f3_file(lightray_direction_save_path, "dir_", k, nullptr, 0);
            }
        }
        lap("trace");
        if (ok) {
            photon_scene_download_image(s, image_array);         // through pinned staging: all or nothing
            lap("image D2H");
            VLOG("traced %lld rays, %lld ray-steps, kernel %.3f ms\n", photon_scene_rays(s), photon_scene_ray_steps(s), photon_scene_last_trace_ms(s));
        }
        photon_scene_destroy(s);
        lap("scene destroy");
        return;
    }

    // ---- several GPUs inside one call: contiguous source-point shards, field replicated per
    // GPU, private images, one sum at the end.  One host thread per device.
    std::vector<ShardResult> res((size_t)ngpus);
    std::vector<std::thread> th;
    // how the partial images are combined: peer (default: this library's own kernel over NVLink peer memory), nccl, host
    const char *rm = getenv("PHOTON_REDUCE");
    const bool use_nccl = rm && strcmp(rm, "nccl") == 0, use_host = rm && strcmp(rm, "host") == 0;
    // NCCL route only: communicator set-up (seconds on the first call of a process) runs beside the field builds
    std::thread nccl_init([ngpus, use_nccl]() { if (use_nccl) nccl_prepare(ngpus); });
    for (int g = 0; g < ngpus; g++) {
        th.emplace_back([&, g]() {
            auto t0 = std::chrono::steady_clock::now();
            photon_scene *s = make_scene(shard_index * ngpus + g, shard_count * ngpus, g);
            if (!s) return;
            auto t1 = std::chrono::steady_clock::now();
            res[g].scene = s;
            if (!use_nccl && !use_host) grant_peer_access(s, 0);              // early and best effort; the reduce checks again
            res[g].ok = (g != 0 || photon_scene_upload_image(s, image_array) == 0)     // accumulate onto the caller's image
                        && photon_scene_trace(s) == 0 && photon_scene_sync(s) == 0;
            if (timing) fprintf(stderr, "[photon timing] gpu %d: scene + field %8.3f ms, trace %8.3f ms (kernel %8.3f ms)\n", g,
                                std::chrono::duration<double, std::milli>(t1 - t0).count(),
                                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count(),
                                (double)photon_scene_last_trace_ms(s));
        });
    }
    for (auto &t : th) t.join();
    nccl_init.join();
    bool all_ok = true;
    for (int g = 0; g < ngpus; g++) if (!res[g].ok) { fprintf(stderr, "libparallel_ray_tracing(b200): GPU %d failed; image untouched\n", g); all_ok = false; }
    auto destroy_all = [&]() { for (int g = 0; g < ngpus; g++) if (res[g].scene) photon_scene_destroy(res[g].scene); };
    if (!all_ok) { destroy_all(); return; }
    lap("trace (all GPUs)");
    int red = 1;                                                      // one reduce over NVLink, then one D2H
    if (use_nccl) red = nccl_reduce_images(res, ngpus, npix);
    else if (!use_host) red = peer_reduce_images(res, ngpus, npix);
    if (red == 0) {
        lap(use_nccl ? "nccl reduce" : "peer reduce");
        photon_scene_download_image(res[0].scene, image_array);      // folds the fixed-point sum first, if any
        lap("image D2H");
        destroy_all();
        return;
    }
    if (red == 2) {                                                   // device 0 may hold a partly reduced image: no second sum
        fprintf(stderr, "libparallel_ray_tracing(b200): image reduce failed; image untouched\n");
        destroy_all();
        return;
    }
    // no peer access / no NCCL on this machine (or PHOTON_REDUCE=host): the same sum through the host
    const bool fixed = photon_scene_deterministic(res[0].scene) != 0;
    for (int g = 0; g < ngpus && all_ok; g++) {
        if (fixed) { res[g].fixed.resize(npix); all_ok = photon_scene_download_image_fixed(res[g].scene, res[g].fixed.data(), &res[g].scale) == 0; }
        else if (g != 0) { res[g].image.resize(npix); all_ok = photon_scene_download_image(res[g].scene, res[g].image.data()) == 0; }
    }
    if (!all_ok) { fprintf(stderr, "libparallel_ray_tracing(b200): image download failed; image untouched\n"); destroy_all(); return; }
    if (fixed) {                          // deterministic: exact integer sum, one rounding (as on one GPU)
        const double inv = 1.0 / res[0].scale;
        for (size_t i = 0; i < npix; i++) {
            long long q = 0;
            for (int g = 0; g < ngpus; g++) q += res[g].fixed[i];
            if (q != 0) image_array[i] += (float)((double)q * inv);
        }
    } else {
        std::vector<float> img0(npix);
        if (photon_scene_download_image(res[0].scene, img0.data()) == 0) {     // holds the caller's image + shard 0
            for (size_t i = 0; i < npix; i++) {
                float acc = img0[i];
                for (int g = 1; g < ngpus; g++) acc += res[g].image[i];
                image_array[i] = acc;
            }
        }
    }
    destroy_all();
}

// ---------------------------------------------------------------------------------
// Roofline denominator measured on the box: FP32 FMA issue rate (the tracer is bound by
// FP32 issue / L1 latency, not by HBM; MEASURED_PEAKS.json only has HBM and bf16 figures).
// ---------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) fp32_peak_kernel(float *out, int iters, float a, float b) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == 123.456f) out[0] = s;
}
// the same with Blackwell's packed FFMA2 (fma.rn.f32x2): two FP32 FMAs per issued instruction
__global__ void __launch_bounds__(256) fp32x2_peak_kernel(float *out, int iters, float a, float b) {
    float t = threadIdx.x * 1e-3f;
    float2 x0 = make_float2(t, t + 1), x1 = make_float2(t + 2, t + 3), x2 = make_float2(t + 4, t + 5), x3 = make_float2(t + 6, t + 7);
    float2 x4 = make_float2(t + 8, t + 9), x5 = make_float2(t + 10, t + 11), x6 = make_float2(t + 12, t + 13), x7 = make_float2(t + 14, t + 15);
    const float2 aa = make_float2(a, a), bb = make_float2(b, b);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            x0 = __ffma2_rn(x0, aa, bb); x1 = __ffma2_rn(x1, aa, bb); x2 = __ffma2_rn(x2, aa, bb); x3 = __ffma2_rn(x3, aa, bb);
            x4 = __ffma2_rn(x4, aa, bb); x5 = __ffma2_rn(x5, aa, bb); x6 = __ffma2_rn(x6, aa, bb); x7 = __ffma2_rn(x7, aa, bb);
        }
    }
    float s = x0.x + x1.x + x2.x + x3.x + x4.x + x5.x + x6.x + x7.x + x0.y + x1.y + x2.y + x3.y + x4.y + x5.y + x6.y + x7.y;
    if (s == 123.456f) out[0] = s;
}
}  // namespace

static double measure_fp32(int device, bool packed);
extern "C" double photon_b200_measure_fp32_tflops(int device) { return measure_fp32(device, false); }
extern "C" double photon_b200_measure_fp32x2_tflops(int device) { return measure_fp32(device, true); }
static double measure_fp32(int device, bool packed) {
    if (photon_b200_device_count() < 1) return -1.0;
    int dev = pick_device(device);
    if (cudaSetDevice(dev) != cudaSuccess) return -1.0;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return -1.0;
    float *d = nullptr;
    if (cudaMalloc(&d, 4) != cudaSuccess) return -1.0;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096, blocks = prop.multiProcessorCount * 8, threads = 256;
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        if (packed) fp32x2_peak_kernel<<<blocks, threads>>>(d, iters, 0.999f, 0.001f);
        else fp32_peak_kernel<<<blocks, threads>>>(d, iters, 0.999f, 0.001f);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        double flops = (packed ? 2.0 : 1.0) * 2.0 * 8 * 16 * (double)iters * threads * blocks;
        double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    return best;
}

// ---------------------------------------------------------------------------------
// photon_b200_export_u16_static: image export that allocates nothing (include/photon_b200.h).
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// photon_scene_export_u16 takes its two scratch buffers -- the float image after gain (the input of pass 2) and the
// word the block maxima are folded into -- from the library's memory pool on every call.  Here they live in one
// file-scope device array of fixed size: [0, 256) the maximum (cleared by every call), [256, 256 + 4 n) the
// intermediate image (overwritten by every call, never cleared).  Every GPU that loads the library holds the
// 16 MiB + 256 B until the process exits.  In host code `g_export_scratch` is a host-side stand-in for the symbol, not
// a device address: the address the kernels get comes from cudaGetSymbolAddress, asked once per device with that
// device current.  All calls on one device share the array: calls that overlap on the GPU (two streams, two
// threads) race on it.
// ---------------------------------------------------------------------------------
#define PH_EXPORT_STATIC_MAX_PIXELS ((size_t)PHOTON_B200_EXPORT_STATIC_MAX_PIXELS)
static __device__ __align__(256) unsigned char g_export_scratch[256 + PH_EXPORT_STATIC_MAX_PIXELS * sizeof(float)];

namespace {
constexpr int kMaxScratchDevices = 64;
std::mutex g_export_scratch_mutex;
unsigned char *g_export_scratch_ptr[kMaxScratchDevices] = {nullptr};

// Device address of g_export_scratch on `device`, which must be the current device.
unsigned char *export_scratch_on(int device) {
    if (device < 0 || device >= kMaxScratchDevices) { g_error = "export_u16_static: device ordinal out of range"; return nullptr; }
    std::lock_guard<std::mutex> lk(g_export_scratch_mutex);
    if (!g_export_scratch_ptr[device]) {
        unsigned char *p = nullptr;
        if (!cuda_ok(cudaGetSymbolAddress((void **)&p, g_export_scratch), "cudaGetSymbolAddress(g_export_scratch)", __FILE__, __LINE__)) return nullptr;
        g_export_scratch_ptr[device] = p;
    }
    return g_export_scratch_ptr[device];
}
// The device a pointer lives on, or -1 if it is not device (or managed) memory.
int device_of(const void *p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return -1; }
    if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged) return -1;
    return at.device;
}
}  // namespace

extern "C" int photon_b200_export_u16_static(float *image_dev, long long n_pixels, float pixel_gain, int pixel_bit_depth,
                                             int intensity_rescaling, float image_noise, unsigned int seed,
                                             unsigned short *out_dev, void *stream) {
    g_error.clear();
    if (!image_dev || !out_dev || n_pixels < 0 || pixel_bit_depth < 1 || pixel_bit_depth > 16) { g_error = "export_u16_static: bad arguments"; return 1; }
    if ((unsigned long long)n_pixels > (unsigned long long)PH_EXPORT_STATIC_MAX_PIXELS) {
        char buf[256];
        snprintf(buf, sizeof buf, "export_u16_static: %lld pixels exceed the static scratch of %zu pixels (PHOTON_B200_EXPORT_STATIC_MAX_PIXELS); "
                                  "use photon_scene_export_u16", n_pixels, PH_EXPORT_STATIC_MAX_PIXELS);
        g_error = buf;
        return 1;
    }
    if (n_pixels == 0) return 0;
    const int dev = device_of(image_dev);
    if (dev < 0 || device_of(out_dev) != dev) { g_error = "export_u16_static: image_dev and out_dev must be device memory of one GPU"; return 1; }
    int prev = -1;
    if (!cuda_ok(cudaGetDevice(&prev), "cudaGetDevice", __FILE__, __LINE__)) return 1;
    if (prev != dev && !cuda_ok(cudaSetDevice(dev), "cudaSetDevice", __FILE__, __LINE__)) return 1;
    bool ok = false;
    if (unsigned char *scratch = export_scratch_on(dev)) {
        cudaStream_t st = (cudaStream_t)stream;
        unsigned int *dmax = reinterpret_cast<unsigned int *>(scratch);
        float *tmp = reinterpret_cast<float *>(scratch + 256);
        const size_t n = (size_t)n_pixels;
        const float gain = (float)std::pow(10.0, (double)pixel_gain / 20.0);
        const float levels = (float)((1 << pixel_bit_depth) - 1);
        const float k = (float)((65536.0 - 1.0) / ((double)(1 << pixel_bit_depth) - 1.0));
        const unsigned blocks = (unsigned)std::min<size_t>((n + 255) / 256, 148 * 8);
        // the array keeps the last call's contents: the maximum must start at zero; tmp[0, n) is fully rewritten
        ok = cuda_ok(cudaMemsetAsync(dmax, 0, sizeof(unsigned int), st), "export_u16_static: clear", __FILE__, __LINE__);
        if (ok) {
            export_prepare_kernel<<<blocks, 256, 0, st>>>(image_dev, tmp, n, gain, image_noise * 100.0f, seed, intensity_rescaling, dmax);
            export_u16_kernel<<<blocks, 256, 0, st>>>(tmp, out_dev, n, levels, k, intensity_rescaling, dmax);
            ok = cuda_ok(cudaGetLastError(), "export_u16_static: kernels", __FILE__, __LINE__);
        }
    }
    if (prev != dev) cudaSetDevice(prev);
    return ok ? 0 : 1;
}
