/* sunsky_b200.h — C ABI of the B200-native sunsky emitter hot path.
 *
 * Drop-in boundary for the `sunsky` emitter plugin of matttsss/mitsuba3-sunsky
 * (reference: src/emitters/sunsky.cpp, class SunskyEmitter, lines 153-1034). The reference exposes
 * the path only as a C++ class behind Mitsuba's plugin loader (plugin_name()/plugin_descr(),
 * include/mitsuba/core/class.h:220-225); a Mitsuba adapter plugin or any other host binds the
 * entry points below instead. Each entry point names the reference member it replaces.
 *
 * Conventions
 *   - all arithmetic is fp32 and follows the reference's llvm_ad_rgb / llvm_ad_spectral op order
 *   - arrays are planar (SoA): one pointer per component, `n` floats each, caller-owned
 *   - `*_device` pointers are CUDA device pointers on the context's device; kernels are enqueued
 *     on the caller's stream (a cudaStream_t passed as void*; NULL = legacy default stream) and
 *     the call returns without synchronising
 *   - `*_host` entry points take host pointers, stage through pinned memory, run the same kernels
 *     and return after the results are in the caller's buffers
 *   - every function returns SSK_OK (0) or a negative error code; ssk_last_error() describes the
 *     last failure on the calling thread. There is no CPU fallback: without a CUDA device the
 *     compute entry points fail with SSK_ERR_CUDA.
 *   - a context is immutable between ssk_update() calls, so launches may be issued concurrently
 *   - ordering of updates against launches: ssk_update / ssk_update_batch / ssk_set_scene never touch device memory
 *     that launches already queued still read. Each context keeps a ring of 3 device blocks; an update writes the
 *     next one (after waiting, on the host, for the launches that last used it -- those issued two updates ago) and
 *     only launches issued AFTER the update call returns see the new state. So the Dr.Jit / torch pattern
 *     "launch, parameters_changed, launch" on one stream needs no synchronisation by the caller. The wait is on
 *     the event recorded after the most recent launch per block: a caller that issues launches on one context from
 *     SEVERAL streams and updates more than twice while the oldest of them are still queued must synchronise
 *     those streams itself.
 *   - every call restores the calling thread's current CUDA device before it returns
 */
#ifndef SUNSKY_B200_H
#define SUNSKY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSK_OK                 0
#define SSK_ERR_INVALID       -1   /* bad argument / property out of range (reference: Log(Error,..) throws) */
#define SSK_ERR_IO            -2   /* dataset file missing or malformed */
#define SSK_ERR_CUDA          -3   /* CUDA runtime failure, or no device */
#define SSK_ERR_UNSUPPORTED   -4

#define SSK_MODE_RGB           0   /* llvm_ad_rgb:      3 channels */
#define SSK_MODE_SPECTRAL      1   /* llvm_ad_spectral: 4 wavelengths per ray, 11 dataset bins */

/* ssk_params.flags */
#define SSK_FLAG_SUN_DIR_NORMALIZED 1u  /* sun_direction is already unit length: use it verbatim */
#define SSK_FLAG_TO_LOCAL_GIVEN     2u  /* to_local holds the inverse the caller's Transform4f carries
                                           (include/mitsuba/core/transform.h:44-45, 59-62, 77-79: a Transform stores
                                           matrix AND inverse_transpose, built analytically by rotate()/translate()/...);
                                           without the flag the inverse is computed as Properties does for a to_world
                                           property: fp64 cofactor inverse, rounded to fp32 */

/* ssk_params.arith */
#define SSK_ARITH_EXACT        0   /* replay the reference's op sequence (IEEE div/sqrt, CEPHES polynomials):
                                      bit-identical to the llvm_ad_* variants */
#define SSK_ARITH_FAST         1   /* same formulas; geometry, cancellation-prone terms and sampled directions keep
                                      the reference sequence, per-channel / per-Gaussian exp, pow and divisions
                                      use the MUFU unit (ex2 / rsqrt / rcp). Radiance, pdf and weight stay within
                                      the 1e-5 relative parity tolerance, directions are bit-identical */

/* The plugin's property surface (reference: src/emitters/sunsky.cpp:19-104, init_from_props :889-948). */
typedef struct ssk_params {
    float    turbidity;          /* [1, 10], default 3 */
    int32_t  n_albedo;           /* 1 = one value for all channels, 3 (RGB) or 11 (spectral bins 320..720 nm) */
    float    albedo[11];         /* each in [0, 1], default 0.3 */
    int32_t  use_record;         /* 0: sun_direction, 1: location + date/time record */
    float    sun_direction[3];   /* world space, pointing towards the sun */
    float    latitude, longitude, timezone;     /* defaults 35.6894, 139.6917, 9 */
    int32_t  year, month, day;                  /* defaults 2010, 7, 10 */
    float    hour, minute, second;              /* defaults 15, 0, 0 */
    float    sun_scale, sky_scale;              /* >= 0, default 1 */
    float    sun_aperture_deg;                  /* (0, 180), default 0.5358 */
    float    to_world[16];                      /* row-major 4x4 affine, default identity */
    uint32_t flags;
    int32_t  arith;
    float    to_local[16];                      /* row-major inverse of to_world; read only with SSK_FLAG_TO_LOCAL_GIVEN */
} ssk_params;

/* Interpolated per-emitter state, for parity checks against the reference's private members
 * (m_sky_params, m_sky_radiance, m_sun_radiance, m_gaussians, m_gaussian_distr, m_sky_sampling_w,
 * m_sun_angles, m_local_sun_frame; src/emitters/sunsky.cpp:995-1032). */
typedef struct ssk_state {
    int32_t n_channels;              /* 3 or 11 */
    float   sky_params[99];          /* [channel][9] */
    float   sky_radiance[11];
    int32_t n_sun_radiance;          /* 3240 (RGB) or 1980 (spectral) */
    float   sun_radiance[3240];
    float   sun_ld[66];              /* [wavelength][6], spectral only */
    float   gaussians[100];          /* [20][mu_phi, mu_theta, sigma_phi, sigma_theta, weight] */
    float   pmf[20], cdf[20];
    float   cdf_sum, cdf_normalization;
    float   sky_sampling_w;
    float   sun_angles[2];           /* (phi, theta) of the sun in emitter-local coordinates */
    float   frame_s[3], frame_t[3], frame_n[3];
    float   sun_half_aperture;       /* radians */
    float   sun_dir_world[3];
    float   bsphere_center[3], bsphere_radius;
    float   spectral_pdf[10], spectral_cdf[10];   /* wavelength distribution over 360..720 nm */
    float   spectral_integral, spectral_normalization;
    float   area_ratio;
    int32_t sun_below_horizon;       /* the reference logs a warning in this case */
    float   to_local[9];             /* row-major 3x3 linear part of m_to_world.inverse() as used per lane */
} ssk_state;

typedef struct ssk_ctx ssk_ctx;

const char *ssk_last_error(void);
const char *ssk_version(void);

/* Fill `p` with the plugin's defaults (src/emitters/sunsky.cpp:889-948). */
void ssk_params_default(ssk_params *p);

/* SunskyEmitter(const Properties&) (src/emitters/sunsky.cpp:162-218). `dataset_dir` holds the
 * reference's resources/sunsky/datasets/ *.bin files (format: sunsky.h:515-561). `device` is the CUDA
 * ordinal; -1 builds a host-only context (state queries work, compute entry points fail). */
int ssk_create(const ssk_params *p, const char *dataset_dir, int device, int mode, ssk_ctx **out);

/* parameters_changed() after mi.traverse(emitter) updates (src/emitters/sunsky.cpp:220-285). */
int ssk_update(ssk_ctx *ctx, const ssk_params *p);
/* parameters_changed for n emitters of one device and mode at once (time-of-day sweeps, animations): the same
 * result as n ssk_update calls, but the setup quadratures of all emitters share one chain-sum launch and one
 * synchronisation (0.14 ms per emitter one by one; see DESIGN.md for the batched figure). On failure no context is
 * modified. */
int ssk_update_batch(ssk_ctx *const *ctxs, const ssk_params *params, int n);
/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). Same result as
 * ssk_update_batch -- every context ends in exactly the state n ssk_update calls produce -- but the call makes no
 * cudaMalloc / cudaMallocHost / cudaFree and creates no stream: the scratch of the setup quadrature (node values of
 * the two 200 x 200 integrands, chain partials, Gauss-Legendre table) lives in one file-scope `static __device__`
 * array of SSK_NOALLOC_MAX_EMITTERS x 4.125 MiB (+ 2 KiB) that every GPU which loads the library holds until the
 * process exits; kernels and copies run on the default stream and the call returns when the blocks are on the
 * device. n > SSK_NOALLOC_MAX_EMITTERS is rejected with SSK_ERR_INVALID (split the batch), so are contexts without
 * a device. All calls share the one array per GPU and it keeps the previous call's contents (the partial sums are
 * cleared at the start of every call, the node values are overwritten in full): calls on one device serialise on a
 * mutex inside the library, and nothing else may use the array -- two processes' worth of calls cannot overlap on
 * the GPU because each process has its own copy. Contexts must exist already (ssk_create allocates). On failure no
 * context is modified. */
#define SSK_NOALLOC_MAX_EMITTERS 2
int ssk_update_batch_noalloc(ssk_ctx *const *ctxs, const ssk_params *params, int n);

/* set_scene(): bounding sphere from the scene bbox; NULL pointers = invalid bbox (:287-301). */
int ssk_set_scene(ssk_ctx *ctx, const float *bbox_min, const float *bbox_max);

void ssk_destroy(ssk_ctx *ctx);
int ssk_get_state(const ssk_ctx *ctx, ssk_state *out);
int ssk_channels(const ssk_ctx *ctx);           /* 3 (RGB) or 4 (spectral: wavelengths per ray) */
int ssk_device(const ssk_ctx *ctx);

/* to_string() (src/emitters/sunsky.cpp:502-519) */
int ssk_to_string(const ssk_ctx *ctx, char *buf, size_t len);

/* ---- device entry points ------------------------------------------------------------------- */

/* eval(si) (src/emitters/sunsky.cpp:303-352): radiance arriving along -wi.
 * wi[3]: si.wi, world space. wavelengths[4]: si.wavelengths in nm (spectral mode; NULL in RGB mode).
 * out[C]: C = 3 or 4 planes. */
int ssk_eval_device(const ssk_ctx *ctx, const float *const wi[3], const float *const wavelengths[4],
                    size_t n, float *const out[4], void *stream);

/* eval_direction(it, ds) (:453-461): same as eval with wi = -ds.d. */
int ssk_eval_direction_device(const ssk_ctx *ctx, const float *const d[3], const float *const wavelengths[4],
                              size_t n, float *const out[4], void *stream);

/* pdf_direction(it, ds) (:443-451): solid-angle density of sample_direction at world direction ds.d. */
int ssk_pdf_direction_device(const ssk_ctx *ctx, const float *const d[3], size_t n, float *pdf, void *stream);

/* sample_direction(it, u) (:399-441), fused: technique selection, TGMM / sun-cone sample, pdf, eval,
 * weight = eval / pdf (non-finite -> 0).
 *   u[2]: the Point2f sample. it_p[3]: it.p (NULL = origin). wavelengths[4]: it.wavelengths (spectral).
 *   d[3]: ds.d. pdf: ds.pdf. weight[C]: returned spectrum.
 *   p[3] (nullable): ds.p = it.p + d * dist. dist (nullable): ds.dist = 2 * max(r_bsphere, |it.p - c|).
 *   ds.n = -d, ds.uv = u, ds.delta = false, ds.time = it.time are trivially reconstructed by the caller. */
int ssk_sample_direction_device(const ssk_ctx *ctx, const float *const u[2], const float *const it_p[3],
                                const float *const wavelengths[4], size_t n, float *const d[3], float *pdf,
                                float *const weight[4], float *const p[3], float *dist, void *stream);

/* Forward derivative of eval() along a tangent of the Differentiable parameters `turbidity`, `albedo` and
 * `sun_direction` (traverse(), src/emitters/sunsky.cpp:222-237; sunsky-testing/traversal_test.py:94-144 drives it
 * with dr.forward_from):
 *   out = d eval/d turbidity * d_turbidity + sum_c d eval/d albedo[c] * d_albedo[c] + sum_i d eval/d sun_direction[i] * d_sun_direction[i]
 * d_albedo has 0, 1 (same tangent for every channel) or n_channels (3 RGB / 11 spectral bins) entries; d_sun_direction
 * is NULL or 3 floats (world space; emitters built from a sun direction only). Planes as ssk_eval_device. The
 * derivative is the one forward-mode AD of the reference's code produces: the interpolated state is piecewise linear
 * in turbidity and albedo and a quintic Bezier in the sun elevation; gamma = unit_angle(sun, wo) and cos(psi) are
 * differentiated as written; masks, bin and segment indices, the sampling weight and the TGMM (not part of eval())
 * carry no tangent. The tangent block travels as a kernel parameter: calls on one context may be issued
 * concurrently. Independent oracle: oracle/jvp_oracle.py (fp64 dual numbers), tests/test_parity_gpu.py. */
int ssk_eval_jvp_device(const ssk_ctx *ctx, float d_turbidity, const float *d_albedo, int n_d_albedo, const float *d_sun_direction,
                        const float *const wi[3], const float *const wav[4], size_t n, float *const out[4], void *stream);

/* sample_wavelengths(si, sample) (src/emitters/sunsky.cpp:463-480).
 *   spectral: wavelengths[k] = ContinuousDistribution::sample_pdf(sample_shifted<4>(sample)[k]) over the emitter's
 *   mean sun+sky spectrum on [360, 720] nm (include/mitsuba/core/distr_1d.h:468-502, core/math.h:408-431), and
 *   weight = eval(si with those wavelengths) / pdf. sample: one float per ray. wavelengths[4], weight[4]: outputs.
 *   RGB: the reference returns Wavelength(0) and eval(si); sample and wavelengths are ignored (may be NULL),
 *   weight[3] = eval. */
int ssk_sample_wavelengths_device(const ssk_ctx *ctx, const float *const wi[3], const float *sample, size_t n,
                                  float *const wavelengths[4], float *const weight[4], void *stream);

/* sample_ray(time, wavelength_sample, sample2, sample3) (src/emitters/sunsky.cpp:354-397), fused: origin offset
 * on the bounding-sphere disc (sample2), technique selection + TGMM / sun-cone direction (sample3), pdf including
 * the disc, wavelengths + eval, weight = eval / (pdf_wavelength * pdf) (non-finite -> 0).
 *   wl_sample: wavelength sample per ray (spectral; NULL for RGB). s2[2], s3[2]: the two Point2f samples.
 *   o[3], d[3]: ray origin and direction (d points from the sky into the scene). wavelengths[4]: ray.wavelengths
 *   (spectral; NULL for RGB). weight[C]: returned spectrum. ray.time = time is reconstructed by the caller. */
int ssk_sample_ray_device(const ssk_ctx *ctx, const float *wl_sample, const float *const s2[2], const float *const s3[2],
                          size_t n, float *const o[3], float *const d[3], float *const wavelengths[4],
                          float *const weight[4], void *stream);

/* Equirectangular environment maps (BASELINE configs 4 and 5): one map per context, direction
 * generated from the pixel centre, no direction buffer is read. Pixel (x, y) looks along
 *   phi = 2*pi*(x + 0.5)/width, theta = pi*(y + 0.5)/height, d = (sin(theta)cos(phi), sin(theta)sin(phi), cos(theta))
 * and stores eval(wi = -d). out: n_maps * C planes of width*height floats (map-major, then channel).
 * wavelengths4: the 4 wavelengths used for every pixel of a spectral map (n_maps * 4 floats, host memory). */
int ssk_envmap_device(ssk_ctx *const *ctxs, int n_maps, int width, int height, const float *wavelengths4,
                      float *out, void *stream);

/* ---- host entry points (same semantics; H2D + kernel + D2H inside the call) ------------------- */
int ssk_eval_host(const ssk_ctx *ctx, const float *const wi[3], const float *const wavelengths[4],
                  size_t n, float *const out[4]);
int ssk_eval_direction_host(const ssk_ctx *ctx, const float *const d[3], const float *const wavelengths[4],
                            size_t n, float *const out[4]);
int ssk_pdf_direction_host(const ssk_ctx *ctx, const float *const d[3], size_t n, float *pdf);
int ssk_sample_direction_host(const ssk_ctx *ctx, const float *const u[2], const float *const it_p[3],
                              const float *const wavelengths[4], size_t n, float *const d[3], float *pdf,
                              float *const weight[4], float *const p[3], float *dist);
int ssk_sample_wavelengths_host(const ssk_ctx *ctx, const float *const wi[3], const float *sample, size_t n,
                                float *const wavelengths[4], float *const weight[4]);
int ssk_sample_ray_host(const ssk_ctx *ctx, const float *wl_sample, const float *const s2[2], const float *const s3[2],
                        size_t n, float *const o[3], float *const d[3], float *const wavelengths[4],
                        float *const weight[4]);

/* ---- multi-GPU (one 8 x B200 box) -----------------------------------------------------------------------
 * The reference has no multi-device layer (no NCCL / MPI under src/ or include/). Rays, samples and pixels are
 * independent given the 17 KB emitter block, which every GPU builds for itself: query batches shard as contiguous
 * ray ranges, env-map sweeps as whole frames, with NO data-path collective. The one optional exchange is collecting
 * finished frames on one GPU. These entry points serve both deployment shapes: one process per GPU (bench.py under
 * torchrun; handles and ids travel through any host channel) and one process driving all GPUs (ssk_group_*). */
#define SSK_SHARD_BLOCKED      0   /* consecutive frames per GPU (configs[3]: 256 maps, 32 per GPU) */
#define SSK_SHARD_ROUND_ROBIN  1   /* frame f on GPU f % world (configs[4]: animations of any length stay balanced) */
#define SSK_PEER_HANDLE_BYTES  64
#define SSK_COMM_ID_BYTES      128

const char *ssk_multi_last_error(void);

/* [begin, end) of an n-ray batch for `rank`: boundaries are multiples of `align` rays (planes stay 16-byte aligned
 * with align = 4), the last rank takes the ragged tail, shards differ by at most `align` rays. */
int ssk_shard_ray_range(size_t n, int rank, int world, size_t align, size_t *begin, size_t *end);
/* Frame indices rendered by `rank` (frames may be NULL to query the count). */
int ssk_shard_frames(int n_frames, int rank, int world, int mode, int *frames, int *count);

/* Frame buffers other GPUs can store into. ssk_peer_alloc: device memory on `device` plus (handle != NULL) an
 * inter-process handle for it; ssk_peer_open maps such a buffer into another process on `device` and enables peer
 * access, so that kernels there -- ssk_envmap_device / ssk_sweep_render with `out` pointing into the mapping --
 * write their pixels straight into the owner's HBM over NVLink: the gather happens inside the kernel that computes
 * the frame, nothing is staged or copied afterwards. */
int ssk_peer_alloc(int device, size_t bytes, void **ptr, unsigned char handle[SSK_PEER_HANDLE_BYTES]);
int ssk_peer_open(int device, const unsigned char handle[SSK_PEER_HANDLE_BYTES], void **ptr);
int ssk_peer_close(int device, void *ptr);
int ssk_peer_free(int device, void *ptr);

/* BASELINE configs[3] / [4]: n_frames emitters (full property sets, same mode and arith), one equirectangular map
 * each (pixel convention of ssk_envmap_device) into out[f] (device memory, local or peer-mapped). Emitter contexts
 * are pooled per (device, mode) and re-parametrised `chunk` at a time (ssk_update_batch); the setup of one chunk
 * overlaps the kernels of the previous one. wavelengths4: 4 per frame (spectral). Enqueued on `stream`. */
int ssk_sweep_render(const ssk_params *frames, int n_frames, const char *dataset_dir, int device, int mode,
                     int width, int height, const float *wavelengths4, float *const *out, int chunk, void *stream);
void ssk_sweep_release(void);      /* destroys the pooled contexts */

/* The optional NCCL gather of the north star. libnccl.so.2 is loaded at run time (the copy already in the process
 * if there is one). The 128-byte id comes from ssk_comm_unique_id on one rank and reaches the others through any
 * host channel. ssk_comm_gather_frames: frames first .. first + count - 1 of every rank's list (ssk_shard_frames
 * order; `local` holds exactly those) to their final place in `full` on `root`, one grouped ncclSend / ncclRecv
 * launch. ssk_sweep_render_gather_nccl: the sweep with that gather running on a second stream, double-buffered
 * against the next chunk's kernels (staging = 2 * chunk frames). */
typedef struct ssk_comm ssk_comm;
int  ssk_comm_unique_id(unsigned char id[SSK_COMM_ID_BYTES]);
int  ssk_comm_create(const unsigned char id[SSK_COMM_ID_BYTES], int rank, int world, int device, ssk_comm **out);
void ssk_comm_destroy(ssk_comm *c);
int  ssk_comm_version(void);
int  ssk_comm_gather_frames(ssk_comm *c, const float *local, size_t frame_floats, int n_frames, int shard_mode, int first,
                            int count, float *full, int root, void *stream);
int  ssk_sweep_render_gather_nccl(ssk_comm *c, const ssk_params *frames, int n_frames_total, int shard_mode,
                                  const char *dataset_dir, int mode, int width, int height, const float *wavelengths4,
                                  float *staging, float *full, int root, int chunk, void *stream);

/* One process, N GPUs: a context per device built from the same properties, one host thread per device inside each
 * call. Query batches are split with ssk_shard_ray_range (align 4); sweeps with ssk_shard_frames. For a gathered
 * sweep every out[f] lives on one GPU and the others store into it through peer access (enabled at creation). */
typedef struct ssk_group ssk_group;
int  ssk_group_create(const int *devices, int n_devices, const ssk_params *p, const char *dataset_dir, int mode, ssk_group **out);
void ssk_group_destroy(ssk_group *g);
int  ssk_group_size(const ssk_group *g);
ssk_ctx *ssk_group_context(const ssk_group *g, int i);
int  ssk_group_update(ssk_group *g, const ssk_params *p);
int  ssk_group_eval_host(const ssk_group *g, const float *const wi[3], const float *const wavelengths[4], size_t n,
                         float *const out[4]);
int  ssk_group_sample_direction_host(const ssk_group *g, const float *const u[2], const float *const wavelengths[4], size_t n,
                                     float *const d[3], float *pdf, float *const weight[4]);
int  ssk_group_sweep(const ssk_group *g, const ssk_params *frames, int n_frames, int shard_mode, int width, int height,
                     const float *wavelengths4, float *const *out, int chunk);

/* Test hook: out[i] = a[i] / b[i] computed with the reciprocal-based division the kernels use for
 * per-emitter constant divisors (device pointers); the tests compare it with IEEE division. */
int ssk_debug_fdiv_by_device(const float *a, const float *b, float *out, size_t n, void *stream);
/* Test hook for EXACT mode's guarded shortcuts, each beside the sequence it replaces (device pointers, n even):
 * op 0 / 1: normal-range division a / b by reciprocal refinement / IEEE division; op 2 / 3: exp(a) without / with the
 * range selects; op 4 / 5: pow(a, 1.5) without / with the special cases; op 6 / 7: square root without / with the
 * range test; op 8: scalar form of op 0. b may be NULL for ops 2-7. */
int ssk_debug_math_device(int op, const float *a, const float *b, float *out, size_t n, void *stream);
/* Test hook: overrides the per-emitter guard of EXACT mode's select-free sky term (rays with 1 / (cos(theta) + 0.01) >
 * lim_s or gamma > lim_gamma take the generic sequence behind a call). With the shipped datasets no ray ever fails the
 * guard the library derives, so the tests shrink it to drive rays through the fallback and compare the results, which
 * must not change by a bit. The next ssk_update restores the derived limits. */
int ssk_debug_set_exact_limits(ssk_ctx *ctx, float lim_s, float lim_gamma);

/* Measurement aid: the time the bare transfers of a *_host call take with nothing else in the way -- h2d_bytes from
 * pinned host memory on one stream while d2h_bytes go back on another, in pieces of piece_bytes (0 = one copy per
 * direction): best and (nullable) mean of `reps` passes after a warm-up. With several GPUs sharing the host's memory
 * and PCIe root the mean of passes that every rank runs at the same time is the contended ceiling. bench.py reports
 * the *_host entry points against it. */
int ssk_host_copy_ceiling(int device, size_t h2d_bytes, size_t d2h_bytes, size_t piece_bytes, int reps, double *best_seconds,
                          double *mean_seconds);

/* Number of kernel launches issued through this library since load (bench bookkeeping). */
uint64_t ssk_launch_count(void);

/* Host-side helpers that mirror src/render/python/sunsky_v.cpp and sunsky.h utilities */
/* compute_sun_coordinates (sunsky.h:283-374): date/time/location -> local sun direction */
int ssk_sun_coordinates(int year, int month, int day, float hour, float minute, float second,
                        float latitude, float longitude, float timezone, float out_dir[3]);
/* mi.hosek_sun_rad (src/render/python/sunsky_v.cpp:19 = arhosekskymodel_solar_radiance_internal2,
 * include/mitsuba/render/sunsky/ArHosekSkyModel.c:686-786): direct solar radiance of the original fp64 Hosek-Wilkie
 * model at one wavelength [nm], solar elevation [rad] and angle gamma from the disc centre [rad], evaluated on the
 * shipped fp64 tables (sun_spec_rad.bin, sun_spec_ld.bin). Host, fp64; 0 outside 320..720 nm or turbidity 1..10. */
int ssk_hosek_sun_rad(const char *dataset_dir, double turbidity, double wavelength, double elevation, double gamma, double *out);
/* dr::normalize as the constructor applies it to the sun_direction property (sunsky.cpp:922), same rounding.
 * parameters_changed() does NOT normalise (the reference works on m_sun_dir as stored), so a caller that feeds a raw
 * property to ssk_update and wants constructor semantics normalises it with this first; ssk_sweep_render does. */
int ssk_normalize_direction(const float in[3], float out[3]);
/* array_from_file (sunsky.h:515-561): reads a dataset as fp32. *count receives the element count;
 * shape (nullable) receives up to 8 dims, *ndim their number. `out` may be NULL to query sizes. */
int ssk_array_from_file(const char *path, int file_is_f64, float *out, size_t capacity, size_t *count,
                        size_t *shape, size_t *ndim);
/* array_to_file (sunsky.h:572-597): writes fp32 or fp64 payload with the "SKY" header. */
int ssk_array_to_file(const char *path, const float *data, const size_t *shape, size_t ndim, int as_f64);

#ifdef __cplusplus
}
#endif
#endif /* SUNSKY_B200_H */
