/*
 * photon_b200.h -- extended C ABI of libparallel_ray_tracing.so (B200 build).
 *
 * photon's Python binds exactly one symbol, start_ray_tracing (include/photon_abi.h).
 * That call does everything per image: host->device copies, NRRD load + gradient
 * precompute, the batch loop, device->host copy (cuda_codes/parallel_ray_tracing.cu:
 * 3162-3309 uploads, 3323-3357 field, 3506-3559 launch loop, 3675 image copy-back).
 * The entry points below expose the same stages separately so that a caller can keep
 * a scene resident in HBM and re-trace it (what bench.py times as `value`), read
 * counters, and shard source points over several GPUs / processes.  Plain pointers and
 * sizes only; no C++ or torch types.
 *
 * Environment variables read by the library (the positional ABI cannot grow):
 *   PHOTON_VERBOSE=1        progress lines like the reference's printf output
 *   PHOTON_DEVICE=<n>       CUDA device ordinal (default: current device / 0)
 *   PHOTON_RNG=reference    lens samples from srand(10)/rand(), the same rpp draws for
 *                           every source point (prt.cu:3228-3236, 2006) [default]
 *   PHOTON_RNG=philox       counter-based Philox4x32-10 keyed by the global ray index
 *   PHOTON_SEED=<n>         seed for Philox lens samples and for the noise streams
 *   PHOTON_SAMPLER=linear   software trilinear from a linear float4 buffer [default]
 *   PHOTON_SAMPLER=texture  hardware tex3D<float4> (the reference's 9-bit filter weights)
 *   PHOTON_SHARD=r/n        trace only the r-th of n contiguous source-point shards
 *   PHOTON_GPUS=<n>         split the source points over n GPUs inside one call
 *   PHOTON_DETERMINISTIC=1  accumulate the image in 64-bit fixed point with integer atomics: the
 *                           result is bit-identical from run to run and for any sharding
 *   PHOTON_OCC=<2..6>       resident CTAs per SM the trace kernel variant is compiled for
 *   PHOTON_TIMING=1         per-stage wall times of start_ray_tracing on stderr
 *   PHOTON_INTERPOLATION=cubic  tricubic B-spline sampler on a prefiltered scalar volume (default: linear)
 *   PHOTON_MORTON=<0|1>     CTAs take the source points in Morton order of (x, y) [default 1 with a density field]
 *   PHOTON_RAY_ORDER=<0|1>  the rays of a source point are walked in Morton order of their aim points on the lens, so a
 *                           warp's lanes cross the field side by side [default 1 with a density field]; ray ids, dumps,
 *                           noise streams and dot statistics keep the caller's order
 *   PHOTON_GRADIENT_INDEX=exact  integer voxel index in the gradient stencil instead of the reference's float-arithmetic
 *                           index (which loses its low bits above 2^24 voxels, trace.h:1873-1975) [default: reference]
 *   PHOTON_HOT=<4|5>        RK4 hot loop: 5 = z-neighbour short path (coherent rays), 4 = without; default by the
 *                           width of a source point's cone against the grid spacing
 *   PHOTON_NCCL=0           PHOTON_GPUS: sum the partial images through the host instead of one ncclReduce
 *   PHOTON_STEP_SCALE=<k>   integrator step = k x smallest grid spacing (the reference has k = 1 only)
 *   PHOTON_FULL_DEPTH=1     keep more than the 1024 z-planes the reference keeps (trace.h:2037-2038)
 *   PHOTON_INDEX64=1        run the 64-bit-voxel-index hot loop (automatic above 2^32 voxels, e.g. 2048^3 with PHOTON_FULL_DEPTH)
 *   PHOTON_CELLS=<0|1>      coherent RK4 loop reads the per-cell polynomial table (128 B per voxel) [default 1 when it fits]
 *   PHOTON_CELLS_MAX_GIB=<g>  largest table that is built [default 40]; beyond it the loop reads the corner texels
 *   PHOTON_NO_FIELD_CACHE=1 rebuild the density field on every call
 */
#ifndef PHOTON_B200_H_
#define PHOTON_B200_H_

#include "photon_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct photon_scene photon_scene;   /* opaque: one scene resident on one GPU */

/* Library / device queries. */
int         photon_b200_abi_version(void);          /* bumps when this header changes */
int         photon_b200_device_count(void);         /* 0 if no CUDA device is usable */
const char *photon_b200_last_error(void);           /* "" if none; thread-local */

/*
 * Upload a scene: same arguments as start_ray_tracing minus the image and the dump
 * paths (replaces prt.cu:3162-3357).  shard_index/shard_count select a contiguous range
 * of source points (0/1 = all).  Returns NULL on failure (see photon_b200_last_error).
 */
photon_scene *photon_scene_create(float lens_pitch, float image_distance,
                                  const scattering_data_t *scattering_data, const char *scattering_type_str,
                                  const lightfield_source_t *lightfield_source, int lightray_number_per_particle,
                                  float beam_wavelength, float aperture_f_number,
                                  int num_elements, const double (*element_center)[3], const element_data_t *element_data,
                                  const double (*element_plane_parameters)[4], const int *element_system_index,
                                  const camera_design_t *camera_design,
                                  bool simulate_density_gradients, const char *density_grad_filename,
                                  int ray_tracing_algorithm,
                                  bool add_pos_noise, float pos_noise_std,
                                  bool add_ngrad_noise, float ngrad_noise_std,
                                  float ray_cone_pitch_ratio,
                                  int shard_index, int shard_count, int device);
void photon_scene_destroy(photon_scene *s);

/* Device image (float32 [ny*nx], accumulated onto). */
int   photon_scene_zero_image(photon_scene *s);
int   photon_scene_upload_image(photon_scene *s, const float *host_image);      /* H2D */
int   photon_scene_download_image(photon_scene *s, float *host_image);          /* D2H, overwrites */
void *photon_scene_image_device_ptr(photon_scene *s);
/* The step after the path (perform_ray_tracing_03.py:2189-2243) on the device: optional N(0, image_noise*100)
 * added to the raw image (Philox keyed by pixel and seed; the reference seeds numpy from the clock), negatives
 * and non-finite values to 0, x 10^(pixel_gain/20), rescale by the maximum to 2^bit_depth-1 levels, round
 * half-even, stretch to the 16-bit range, narrow to uint16 -- in float32 like numpy on the float32 image.
 * host_out: ny*nx uint16 (half the D2H bytes of the raw image). */
int   photon_scene_export_u16(photon_scene *s, float pixel_gain, int pixel_bit_depth, int intensity_rescaling,
                              float image_noise, unsigned int seed, unsigned short *host_out);                           /* for NCCL reduce by the caller */
/* Deterministic mode: 1 if the scene accumulates in fixed point; the raw int64 image (increments in
 * units of 1/scale, the same scale for every shard of a call) for an exact cross-GPU sum. */
int   photon_scene_deterministic(photon_scene *s);
int   photon_scene_download_image_fixed(photon_scene *s, long long *host_fixed, double *scale);

/*
 * Trace all source points of the scene's shard onto the device image (replaces the
 * launch loop prt.cu:3506-3559).  Asynchronous on the scene's stream; returns 0 on a
 * successful launch.  photon_scene_sync waits and returns the CUDA status.
 */
int photon_scene_trace(photon_scene *s);
int photon_scene_trace_range(photon_scene *s, int point_begin, int point_count);   /* shard-local points */
int photon_scene_sync(photon_scene *s);

/*
 * Trace one dump batch (source points [batch*source_point_number, ...)) and copy the
 * per-ray sensor positions and post-volume directions to host arrays of n_save float3
 * (NaN where a ray was culled), as written to pos_%04d.bin / dir_%04d.bin by
 * prt.cu:3561-3611.  inter_* may be NULL; otherwise n_save*n_inter float3 each.
 */
int photon_scene_trace_batch(photon_scene *s, int batch, long long n_save, float *final_pos, float *final_dir,
                             int n_inter, float *inter_pos, float *inter_dir);

/* Counters of the work done since the last photon_scene_reset_counters. */
int       photon_scene_reset_counters(photon_scene *s);
long long photon_scene_ray_steps(photon_scene *s);      /* committed integrator steps */
long long photon_scene_rays(photon_scene *s);           /* rays generated */
long long photon_scene_sensor_hits(photon_scene *s);    /* rays that reached the sensor */
float     photon_scene_last_trace_ms(photon_scene *s);  /* CUDA-event time of the last trace */
int       photon_scene_kernel_launches(photon_scene *s);/* kernels launched by the last trace */
int       photon_scene_num_points(photon_scene *s);     /* source points in this shard */

/*
 * Dot statistics: the on-device form of the reference's ray-dump consumer
 * (python_codes/light_ray_processing.py:74-141 load_light_ray_positions/_directions,
 * :211-237 calculate_lightray_deflections, :239-275 calculate_dot_average).  The reference writes
 * 24 bytes per ray to pos_%04d.bin / dir_%04d.bin (prt.cu:3561-3611), reads them back with numpy,
 * takes arccos of the direction cosines and sums segments of num_rays_per_dot consecutive rays.
 * Here the same sums are accumulated during the trace, per segment of rays_per_dot consecutive rays
 * of this scene (ray id = point * rays_per_point + ray, the dump order):
 *   stats[dot*8 + 0..2]  sum of sensor positions x, y, z (um)
 *   stats[dot*8 + 3..5]  sum of arccos(direction cosine) x, y, z after the volume (rad)
 *   stats[dot*8 + 6]     rays that have a sensor position (the dump holds NaN for the others)
 *   stats[dot*8 + 7]     rays that have a direction
 * Enabled per scene; every later photon_scene_trace* adds to the sums until they are reset.
 * Not available together with ray dumps, PHOTON_DETERMINISTIC or PHOTON_SAMPLER=texture.
 */
int       photon_scene_enable_dot_stats(photon_scene *s, long long rays_per_dot);
long long photon_scene_num_dots(photon_scene *s);
int       photon_scene_reset_dot_stats(photon_scene *s);
int       photon_scene_download_dot_stats(photon_scene *s, double *host_stats, long long capacity_dots);

/*
 * Source-point generation, the step before the path (host code).  Replaces the Python loops of
 * python_codes/run_simulation_02.py.  Random draws are INPUTS (the reference uses numpy's clock-seeded global
 * generator): hand over numpy's own stream to reproduce a reference run, or any other stream.
 *   photon_b200_sunflower            calculate_sunflower_coordinates (:999-1055); ring_phase_u = one draw per ring
 *                                    or NULL for phase 0; x, y may be NULL to query the count (return value)
 *   photon_b200_non_overlapping_dots create_non_overlapping_dot_coordinates (:1251-1326), hash grid instead of the
 *                                    O(n^2) distance scan; uniforms = 2 per iteration; returns dots placed into xy[2*i]
 *   photon_b200_expand_dots          the per-dot loop of generate_bos_lightfield_data (:1485-1530) into the
 *                                    lightfield_source_t arrays (x, y, z float32; radiance float64; diameter_index 1)
 *   photon_b200_piv_particles        load_lightfield_data (:941-975): uniform box, Gaussian sheet radiance, camera
 *                                    rotation (row-major 3x3), shift to the object plane
 */
long long photon_b200_sunflower(double diameter, double count, const double *ring_phase_u, double *x, double *y, long long capacity);
long long photon_b200_non_overlapping_dots(double xmin, double xmax, double ymin, double ymax, long long num_dots,
                                           double dot_diameter, const double *uniforms, long long n_pairs, double *xy,
                                           long long *iterations_used /* may be NULL */);
int photon_b200_expand_dots(const double *dot_x, const double *dot_y, long long ndots, const double *off_x, const double *off_y,
                            long long noff, double z, double radiance, float *x, float *y, float *zz, double *rad, int *diam);
int photon_b200_piv_particles(long long n, const double *ux, const double *uy, const double *uz, const double lo[3],
                              const double hi[3], double gaussian_beam_fwhm, double irradiance_constant, const double rotation[9],
                              double z_object, float *x, float *y, float *z, double *rad, int *diam);

/*
 * Mie scattering table (host code, double precision).
 * photon_b200_bhmie replaces python_codes/bhmie.py:3-173 (Bohren & Huffman): size parameter x, relative
 * refractive index (re, im), nang angles in [0, 90] deg -> s1, s2 = 2*nang-1 complex amplitudes over
 * [0, 180] deg (interleaved re, im) and q4 = {qext, qsca, qback, gsca} (q4 may be NULL).
 * photon_b200_mie_irradiance_table replaces calculate_mie_scattering_intensity
 * (run_simulation_02.py:641-696): for each particle size, x = 2 pi size n_medium / wavelength and
 * irradiance[angle*num_sizes + size] = (|s1|^2 + |s2|^2) / 2 -- the row-major table scattering_data_t
 * carries (prt.cu:196-198) -- with scattering_angle[2*nang-1] in radians.
 */
int photon_b200_bhmie(double x, double refrel_re, double refrel_im, int nang, double *s1, double *s2, double *q4);
int photon_b200_mie_irradiance_table(const double *particle_sizes, int num_sizes, double medium_refractive_index,
                                     double particle_refractive_index, double beam_wavelength, int nang,
                                     double *scattering_angle, double *scattering_irradiance);

/* Field ingest alone (replaces trace.h:1663-2105): loads the NRRD, builds (grad n, n-1) on
 * the GPU and copies it back.  meta16 = {min xyz, max xyz, spacing xyz, step, data_min, W, H, D, 0, 0}.
 * out may be NULL to query the size. */
int photon_b200_load_field(const char *path, float meta16[16], float *out_float4, long long capacity_voxels);

/* A density volume handed over in memory instead of an NRRD file: rho[nz][ny][nx] float32 (x fastest) in host
 * memory or on any device of the box, spacing / origin (x, y, z, um) as the NRRD header would give them
 * (`spacings`, `space origin`: loadNRRD, trace.h:1663-1817).  Afterwards density_grad_filename "mem:<name>" selects it.
 * The array must stay valid until every device that traces with it has built its field (first call per device);
 * registering the name again invalidates the cached fields. */
int photon_b200_register_field(const char *name, const float *rho, int nx, int ny, int nz,
                               const double spacing[3], const double origin[3]);
int photon_b200_unregister_field(const char *name);

/* Cubic mode (PHOTON_INTERPOLATION=cubic): the B-spline coefficient volume of n-1 (float[D][H][W]),
 * prefiltered on the GPU (replaces CubicBSplinePrefilter3DTimer, trace.h:1655).  out may be NULL. */
int photon_b200_load_coefficients(const char *path, float *out, long long capacity_voxels);
/* One field sample per point at texel coordinates xyz_texel[3n] (the convention of tex3D in
 * trace.h:1052: texel centres at +0.5, clamp addressing) -> out4[4n] = (dn/dx, dn/dy, dn/dz, n-1).
 * cubic = 0: trilinear blend of the (grad n, n-1) texels with exact float weights; 1: tricubic B-spline with
 * analytic gradient; 2: the hardware unit, tex3D<float4> on the field's texture object (the reference's sampler). */
int photon_b200_sample_field(const char *path, int cubic, int n, const float *xyz_texel, float *out4);

/* What the library keeps between calls, and how to drop it.  A density field stays resident on its device after the
 * call that loaded it (the BOS script calls twice per run with the same file): 16 B per voxel of texels, plus 128 B
 * per voxel when the coherent RK4 loop built its cell table, plus 4 B (cubic mode) or 16 B (texture mode) -- 18 GiB
 * for a 512^3 field with its table.  Entries are reference counted: a field is never freed under a live scene, and a
 * field nobody holds is dropped when another one is loaded on that device or by photon_b200_release_cache(), which
 * also trims the library's own stream-ordered memory pools.  Returns the bytes of field data released. */
long long photon_b200_release_cache(void);
long long photon_b200_cache_bytes(void);           /* field bytes resident now, all devices */
/* Which trace kernel the calling thread's last launch ran: "<integrator>/<gatherer>/<dump|image|stats>[/fixed]",
 * e.g. "rk4/cells/image" (tests use it to prove which instantiation they compared). */
const char *photon_b200_last_variant(void);
/* PHOTON_CUBIC_BRICK=1 (with PHOTON_INTERPOLATION=cubic, coherent RK4 scenes) runs the cubic hot loop on bricks of the
 * quad-row table staged per warp in shared memory by TMA (variant "rk4/cubic-brick/..."); same results bit for bit.
 * This counter is the number of brick refills on the current device whose wait gave up and fell back to the table:
 * 0 unless something is broken (tests assert it).  No counterpart in the reference. */
int photon_b200_brick_timeouts(void);

/* FP32 FMA throughput of the device in TFLOP/s, measured with a register-resident FFMA loop
 * (the roofline denominator for the trace kernel, which is FP32-issue bound, not HBM bound). */
double photon_b200_measure_fp32_tflops(int device);
/* The same loop with Blackwell's packed FFMA2 (two FP32 FMAs per issued instruction). */
double photon_b200_measure_fp32x2_tflops(int device);

/* photon_scene_export_u16 for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  This entry point has no counterpart in the reference.
 * The caller owns input and output -- image_dev: float32[n_pixels] on a GPU (rewritten in
 * place with the noisy image when image_noise > 0, exactly as photon_scene_export_u16 does to the scene's image);
 * out_dev: uint16[n_pixels] on the same GPU -- and the two scratch buffers of the export kernels (the image after
 * gain, the running maximum) live in one static device array of
 *     256 + 4 * PHOTON_B200_EXPORT_STATIC_MAX_PIXELS bytes (16 MiB + 256 B, a 2048 x 2048 sensor),
 * which every GPU that loads the library holds until the process exits.  The call allocates and frees nothing, on
 * the device or through any allocating runtime routine; it enqueues one 4-byte memset and the two kernels on
 * `stream` (a cudaStream_t; NULL = the default stream) and returns without waiting: out_dev is valid once the
 * stream has been synchronised.  n_pixels beyond the maximum is rejected (returns 1, photon_b200_last_error() says
 * so).  All calls on one GPU share the array: calls that overlap on the GPU (different streams or threads without an
 * ordering between them) race on it.  The first call on each device looks the array's address up
 * (cudaGetSymbolAddress); make that call outside a stream capture.  Returns 0 on success. */
#define PHOTON_B200_EXPORT_STATIC_MAX_PIXELS (2048 * 2048)
int photon_b200_export_u16_static(float *image_dev, long long n_pixels, float pixel_gain, int pixel_bit_depth,
                                  int intensity_rescaling, float image_noise, unsigned int seed,
                                  unsigned short *out_dev, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PHOTON_B200_H_ */
