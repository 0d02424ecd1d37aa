// C ABI of the B200 sunsky hot path (include/sunsky_b200.h): context management, launch wrappers for
// the device entry points, and the pinned-staging pipeline behind the host entry points.
#include "ssk_host.cuh"
#include "ssk_kernels.cuh"
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <mutex>
#include <string>
#include <vector>
#include <algorithm>

using namespace ssk;

namespace {

thread_local std::string g_err;
std::atomic<uint64_t> g_launches{ 0 };

int fail(int code, const std::string &msg) { g_err = msg; return code; }
#define SSK_CUDA(call)                                                                             \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(SSK_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));         \
    } while (0)

struct Staging {            // per-context pinned + device scratch for the *_host entry points
    static constexpr int kStreams = 6;      // slots allocated; `active` of them are used (SSK_HOST_STREAMS, default 3)
    int active = 3;
    size_t chunk = 0;       // elements per chunk
    int planes = 0;
    float *pinned[kStreams] = {};
    float *dev[kStreams] = {};
    cudaStream_t stream[kStreams] = {};
    std::mutex mu;
};

} // namespace

/// Device copies of the emitter block. Kernels take the hot part by value at launch but read the sun table and the
/// shared-memory staging from global memory when they EXECUTE, so an update must not overwrite a block that queued
/// launches still read. Every update therefore moves to the next block of a small ring and, before writing it,
/// waits for the event recorded after the last launch that used it (launches issued two updates ago).
constexpr int kBlockRing = 3;

struct ssk_ctx {
    HostEmitter E;
    int device = -1;
    int mode = SSK_MODE_RGB;
    EmitterBlock *d_block = nullptr;                 // = ring[cur]: what new launches read
    EmitterBlock *ring[kBlockRing] = {};
    cudaEvent_t used[kBlockRing] = {};               // recorded on the launch stream after every launch on ring[i]
    int cur = 0;
    int sm_count = 0;
    mutable Staging st;
};

namespace {

constexpr int kSampleRays = 1;
/// eval / pdf: one ray per thread measured 3-9 % faster than four (fewer staging registers and selects, and
/// these kernels sit at 25 % of HBM peak, so access width is not what limits them); SSK_RAYS_PER_THREAD=4
/// selects the 128-bit variant when the planes are 16-byte aligned
bool force_scalar() { static const bool v = !(getenv("SSK_RAYS_PER_THREAD") && getenv("SSK_RAYS_PER_THREAD")[0] == '4'); return v; }
bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
bool aligned8(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 7u) == 0; }

/// Persistent grid: enough CTAs to fill every SM at the kernel's occupancy, never more than the work needs.
template <typename K> int occupancy_grid(const ssk_ctx *ctx, K kernel, size_t n_items, int *grid_out, int rays_per_thread = 4) {
    int per_sm = 0;
    SSK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, 0));
    if (per_sm < 1) per_sm = 1;
    size_t want = (n_items + (size_t) kThreads * rays_per_thread - 1) / ((size_t) kThreads * rays_per_thread);
    // Grid = SSK_GRID_WAVES (default 4) waves of resident CTAs, each CTA still a grid-stride loop. One persistent wave
    // (rounds 1-2) left the SMs finishing at different times with nothing to pick up: ncu showed 19.6 of 24 resident
    // warps on average. Measured on B200, 2^24 rays, FAST, 1 / 2 / 4 / 8 / 16 waves: spectral eval 66.3 / 68.3 / 69.4 /
    // 69.2 / 67.2 Gevals/s, RGB eval 137.5 / 140.2 / 142.6 / 142.3 / 138.4, pdf 131.7 / 138.8 / 142.6 / 142.4 / 137.9.
    static const int waves = getenv("SSK_GRID_WAVES") ? std::max(1, atoi(getenv("SSK_GRID_WAVES"))) : 4;
    size_t cap = (size_t) ctx->sm_count * per_sm * waves;
    *grid_out = (int) std::max<size_t>(1, std::min(want, cap));
    return SSK_OK;
}

/// Makes the context's device current for the scope of one C-ABI call and restores the caller's device on exit
/// (a host application that drives several GPUs from one thread must not find its current device changed).
struct DeviceGuard {
    int prev = -1; bool active = false;
    int enter(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
        if (prev != device) { SSK_CUDA(cudaSetDevice(device)); active = prev >= 0; }
        return SSK_OK;
    }
    ~DeviceGuard() { if (active) cudaSetDevice(prev); }
};

int alloc_blocks(ssk_ctx *ctx) {
    for (int i = 0; i < kBlockRing; ++i) {
        SSK_CUDA(cudaMalloc(&ctx->ring[i], sizeof(EmitterBlock)));
        SSK_CUDA(cudaEventCreateWithFlags(&ctx->used[i], cudaEventDisableTiming));
    }
    ctx->cur = 0; ctx->d_block = ctx->ring[0];
    return SSK_OK;
}

/// Next block of the ring becomes the one updates write and new launches read; waits (host side) until the launches
/// that last read it have finished. Launches queued on the current block keep reading the current block.
int rotate_block(ssk_ctx *ctx) {
    if (ctx->device < 0) return SSK_OK;
    const int next = (ctx->cur + 1) % kBlockRing;
    SSK_CUDA(cudaEventSynchronize(ctx->used[next]));
    ctx->cur = next; ctx->d_block = ctx->ring[next];
    return SSK_OK;
}

/// Bookkeeping after a launch on `s` that reads the context's current device block
int launched(const ssk_ctx *ctx, cudaStream_t s, int n = 1) {
    g_launches += n;
    SSK_CUDA(cudaGetLastError());
    SSK_CUDA(cudaEventRecord(ctx->used[ctx->cur], s));
    return SSK_OK;
}

int upload_block(ssk_ctx *ctx) {
    if (ctx->device < 0) return SSK_OK;
    DeviceGuard dg; if (int rc = dg.enter(ctx->device)) return rc;
    // synchronous: when the update call returns the block is on the device, and nothing queued reads ring[cur]
    // (rotate_block) -- except after ssk_set_scene, which patches the current block: see there
    SSK_CUDA(cudaMemcpy(ctx->d_block, &ctx->E.block, sizeof(EmitterBlock), cudaMemcpyHostToDevice));
    return SSK_OK;
}

// ---- device execution of the setup quadrature ---------------------------------------------------------------
struct SetupScratch {       // one per device, shared by every context on it (setup calls serialise on `mu`)
    float *d_gl = nullptr, *d_buf = nullptr, *d_partial = nullptr, *h_partial = nullptr;
    int capacity = 0;       // emitters d_buf / the partial buffers hold (1, grown to kSetupBatch by the first batch)
    cudaStream_t stream = nullptr;
    std::mutex mu;
};
constexpr int kMaxDevices = 64;
SetupScratch g_setup[kMaxDevices];
constexpr int kMaxChains = 2 * kMaxChannels * kQuadChunks;
constexpr int kSetupBatch = 16;                       // emitters per batched quadrature (69 MB of scratch when used)
constexpr size_t kEmitterStride = (size_t) kMaxChains * kQuadChunk;

int ensure_setup_scratch(SetupScratch &sc, int emitters) {
    if (!sc.d_gl) {
        // highest priority: when query / env-map kernels of other streams occupy the device, the few setup CTAs are
        // scheduled as soon as any CTA retires instead of after everything already queued
        int prio_lo = 0, prio_hi = 0;
        SSK_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        SSK_CUDA(cudaStreamCreateWithPriority(&sc.stream, cudaStreamNonBlocking, prio_hi));
        SSK_CUDA(cudaMalloc(&sc.d_gl, 2 * kQuadN * sizeof(float)));
        SSK_CUDA(cudaMemcpy(sc.d_gl, gauss_legendre_200().data(), 2 * kQuadN * sizeof(float), cudaMemcpyHostToDevice));
    }
    if (sc.capacity >= emitters) return SSK_OK;
    if (sc.d_buf) { cudaFree(sc.d_buf); cudaFree(sc.d_partial); cudaFreeHost(sc.h_partial); sc.d_buf = nullptr; sc.capacity = 0; }
    SSK_CUDA(cudaMalloc(&sc.d_buf, (size_t) emitters * kEmitterStride * sizeof(float)));
    SSK_CUDA(cudaMalloc(&sc.d_partial, (size_t) emitters * kMaxChains * sizeof(float)));
    SSK_CUDA(cudaMallocHost(&sc.h_partial, (size_t) emitters * kMaxChains * sizeof(float)));
    sc.capacity = emitters;
    return SSK_OK;
}

/// Quadratures of `n` emitters of one device and mode in groups of kSetupBatch: per emitter one block upload and one
/// node-evaluation launch, per group ONE chain-sum launch, one copy back and one synchronisation.
int quadrature_device_batch(ssk_ctx *const *ctxs, int n, float (*sky_rad)[11], float (*sun_rad)[11]) {
    const int device = ctxs[0]->device;
    if (device >= kMaxDevices) return fail(SSK_ERR_UNSUPPORTED, "device ordinal too large");
    SetupScratch &sc = g_setup[device];
    std::lock_guard<std::mutex> lock(sc.mu);
    DeviceGuard dg; if (int rc = dg.enter(device)) return rc;
    if (int rc = ensure_setup_scratch(sc, n > 1 ? kSetupBatch : 1)) return rc;
    const int C = ctxs[0]->E.n_channels, nchains = 2 * C * kQuadChunks;
    for (int g0 = 0; g0 < n; g0 += sc.capacity) {
        const int m = std::min(sc.capacity, n - g0);
        for (int e = 0; e < m; ++e) {
            ssk_ctx *ctx = ctxs[g0 + e];
            SSK_CUDA(cudaMemcpyAsync(ctx->d_block, &ctx->E.block, sizeof(EmitterBlock), cudaMemcpyHostToDevice, sc.stream));
            setup_quadrature_kernel<<<(2 * kQuadNodes + kThreads - 1) / kThreads, kThreads, 0, sc.stream>>>(
                ctx->E.block, ctx->d_block, sc.d_gl, sc.d_buf + (size_t) e * kEmitterStride, C);
        }
        setup_sum_kernel<<<m * nchains, 32, 0, sc.stream>>>(sc.d_buf, sc.d_partial, nchains, kEmitterStride);
        g_launches += m + 1;
        SSK_CUDA(cudaGetLastError());
        SSK_CUDA(cudaMemcpyAsync(sc.h_partial, sc.d_partial, (size_t) m * nchains * sizeof(float), cudaMemcpyDeviceToHost, sc.stream));
        SSK_CUDA(cudaStreamSynchronize(sc.stream));
        for (int e = 0; e < m; ++e)
            for (int which = 0; which < 2; ++which) {
                const float J = which == 0 ? quad_sky_jacobian() : quad_sun_jacobian(ctxs[g0 + e]->E.block);
                for (int c = 0; c < C; ++c) {
                    float total = 0.f;                   // llvm_block_sum: partials added in chunk order
                    for (int k = 0; k < kQuadChunks; ++k) total = total + sc.h_partial[e * nchains + (which * C + c) * kQuadChunks + k];
                    (which == 0 ? sky_rad : sun_rad)[g0 + e][c] = total * J;
                }
            }
    }
    return SSK_OK;
}

/// QuadratureFn of a context with a device: a batch of one
int quadrature_device(ssk_ctx *ctx, const HostEmitter &, float sky_rad[11], float sun_rad[11]) {
    float sky[1][11], sun[1][11];
    if (int rc = quadrature_device_batch(&ctx, 1, sky, sun)) return rc;
    memcpy(sky_rad, sky[0], sizeof(sky[0])); memcpy(sun_rad, sun[0], sizeof(sun[0]));
    return SSK_OK;
}

/// build_emitter with the quadrature on the context's device (host execution for device-less contexts)
void build_on(ssk_ctx *ctx, const ssk_params &p, bool from_constructor) {
    if (ctx->device < 0 || getenv("SSK_HOST_SETUP")) { build_emitter(ctx->E, p, from_constructor); return; }
    auto t0 = std::chrono::steady_clock::now();
    struct Trace { decltype(t0) t; ~Trace() { if (getenv("SSK_SETUP_TRACE")) fprintf(stderr, "ssk setup: build_emitter total %.1f us\n",
        std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t).count()); } } tr{ t0 };
    build_emitter(ctx->E, p, from_constructor, [ctx](const HostEmitter &E, float *sky, float *sun) {
        if (quadrature_device(ctx, E, sky, sun)) throw std::runtime_error("setup quadrature: " + g_err);
    });
}

int need_device(const ssk_ctx *ctx, DeviceGuard &dg) {
    if (!ctx) return fail(SSK_ERR_INVALID, "null context");
    if (ctx->device < 0 || !ctx->d_block)
        return fail(SSK_ERR_CUDA, "context has no CUDA device: the sunsky kernels have no CPU fallback");
    return dg.enter(ctx->device);
}

const HotBlock &hot(const ssk_ctx *ctx) { return ctx->E.block; }
bool fast(const ssk_ctx *ctx) { return ctx->E.block.arith == kFast; }

/// spectral eval with the inputs staged by the TMA (every input plane 16-byte aligned). Opt-in (SSK_EVAL_TMA=1):
/// measured on B200 it runs at 53.4-54.0 Gevals/s against 54.7 for the register-prefetch kernel -- the ring itself
/// sustains 6.5 TB/s on this 7-in / 4-out plane pattern (profiles/ubench/stream_probe.cu), but the kernel is bound by
/// its dependency chains at ~62-66 % issue utilisation, not by its loads (profiles/r2_summary.md)
template <int A>
int launch_eval_spec_tma(const ssk_ctx *ctx, const Planes3 &i3, const Planes4 &w4, float *const out[4], size_t n,
                         float dir_sign, cudaStream_t s) {
    int per_sm = 0;
    SSK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, eval_spec_tma_kernel<A>, kThreads, 0));
    if (per_sm < 1) per_sm = 1;
    const size_t tiles = (n + kThreads - 1) / kThreads;
    const int grid = (int) std::max<size_t>(1, std::min(tiles, (size_t) ctx->sm_count * per_sm));
    OutPlanes4 o4 = { out[0], out[1], out[2], out[3] };
    eval_spec_tma_kernel<A><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, i3, w4, o4, n, dir_sign);
    return SSK_OK;
}

template <int A, int R>
int launch_eval_t(const ssk_ctx *ctx, bool spectral, const Planes3 &i3, const Planes4 &w4, float *const out[4], size_t n,
                  float dir_sign, cudaStream_t s) {
    int grid = 1;
    if (spectral) {
        static const bool use_tma = getenv("SSK_EVAL_TMA") != nullptr;
        if (R == 1 && use_tma && aligned16(i3.x) && aligned16(i3.y) && aligned16(i3.z) && aligned16(w4.a) && aligned16(w4.b) &&
            aligned16(w4.c) && aligned16(w4.d))
            return launch_eval_spec_tma<A>(ctx, i3, w4, out, n, dir_sign, s);
        OutPlanes4 o4 = { out[0], out[1], out[2], out[3] };
        if (int rc = occupancy_grid(ctx, eval_spec_kernel<A, R>, n, &grid, R)) return rc;
        eval_spec_kernel<A, R><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, i3, w4, o4, n, dir_sign);
    } else {
        OutPlanes3 o3 = { out[0], out[1], out[2] };
        if (int rc = occupancy_grid(ctx, eval_rgb_kernel<A, R>, n, &grid, R)) return rc;
        eval_rgb_kernel<A, R><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, i3, o3, n, dir_sign);
    }
    return SSK_OK;
}

int launch_eval(const ssk_ctx *ctx, const float *const in[3], const float *const wav[4], size_t n,
                float *const out[4], float dir_sign, cudaStream_t s) {
    DeviceGuard dg; if (int rc = need_device(ctx, dg)) return rc;
    if (!in || !out || !in[0] || !in[1] || !in[2]) return fail(SSK_ERR_INVALID, "null direction / output planes");
    if (n == 0) return SSK_OK;
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    const int C = spectral ? 4 : 3;
    bool vec = !force_scalar() && aligned16(in[0]) && aligned16(in[1]) && aligned16(in[2]);
    for (int c = 0; c < C; ++c) {
        if (!out[c]) return fail(SSK_ERR_INVALID, "null output plane");
        vec = vec && aligned16(out[c]);
    }
    Planes3 i3 = { in[0], in[1], in[2] };
    Planes4 w4 = { nullptr, nullptr, nullptr, nullptr };
    if (spectral) {
        if (!wav || !wav[0] || !wav[1] || !wav[2] || !wav[3])
            return fail(SSK_ERR_INVALID, "spectral mode needs 4 wavelength planes");
        for (int c = 0; c < 4; ++c) vec = vec && aligned16(wav[c]);
        w4 = { wav[0], wav[1], wav[2], wav[3] };
    }
    int rc;
    // RGB: two rays per thread on packed f32x2 arithmetic (planes 8-byte aligned); SSK_RAYS_PER_THREAD=1 or 4
    // selects the older one-ray / 128-bit kernels
    bool pair = !spectral && !getenv("SSK_RAYS_PER_THREAD") && aligned8(in[0]) && aligned8(in[1]) && aligned8(in[2]) &&
                aligned8(out[0]) && aligned8(out[1]) && aligned8(out[2]);
    if (pair) {
        int grid = 1;
        OutPlanes3 o3 = { out[0], out[1], out[2] };
        if (fast(ctx)) {
            if (int rc2 = occupancy_grid(ctx, eval_rgb_pair_kernel<kFast>, n, &grid, 2)) return rc2;
            eval_rgb_pair_kernel<kFast><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, i3, o3, n, dir_sign);
        } else {
            if (int rc2 = occupancy_grid(ctx, eval_rgb_pair_kernel<kExact>, n, &grid, 2)) return rc2;
            eval_rgb_pair_kernel<kExact><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, i3, o3, n, dir_sign);
        }
        rc = SSK_OK;
    } else
    if (fast(ctx)) rc = vec ? launch_eval_t<kFast, 4>(ctx, spectral, i3, w4, out, n, dir_sign, s)
                            : launch_eval_t<kFast, 1>(ctx, spectral, i3, w4, out, n, dir_sign, s);
    else rc = vec ? launch_eval_t<kExact, 4>(ctx, spectral, i3, w4, out, n, dir_sign, s)
                  : launch_eval_t<kExact, 1>(ctx, spectral, i3, w4, out, n, dir_sign, s);
    if (rc) return rc;
    return launched(ctx, s);
}

template <int A, int R>
int launch_pdf_t(const ssk_ctx *ctx, const Planes3 &i3, float *pdf, size_t n, cudaStream_t s) {
    int grid = 1;
    if (int rc = occupancy_grid(ctx, pdf_kernel<A, R>, n, &grid, R)) return rc;
    pdf_kernel<A, R><<<grid, kThreads, 0, s>>>(hot(ctx), i3, pdf, n);
    return SSK_OK;
}

int launch_pdf(const ssk_ctx *ctx, const float *const d[3], size_t n, float *pdf, cudaStream_t s) {
    DeviceGuard dg; if (int rc = need_device(ctx, dg)) return rc;
    if (!d || !d[0] || !d[1] || !d[2] || !pdf) return fail(SSK_ERR_INVALID, "null direction / pdf planes");
    if (n == 0) return SSK_OK;
    bool vec = !force_scalar() && aligned16(d[0]) && aligned16(d[1]) && aligned16(d[2]) && aligned16(pdf);
    Planes3 i3 = { d[0], d[1], d[2] };
    int rc;
    if (!getenv("SSK_RAYS_PER_THREAD") && aligned8(d[0]) && aligned8(d[1]) && aligned8(d[2]) && aligned8(pdf)) {
        int grid = 1;
        if (fast(ctx)) {
            if (int rc2 = occupancy_grid(ctx, pdf_pair_kernel<kFast>, n, &grid, 2)) return rc2;
            pdf_pair_kernel<kFast><<<grid, kThreads, 0, s>>>(hot(ctx), i3, pdf, n);
        } else {
            if (int rc2 = occupancy_grid(ctx, pdf_pair_kernel<kExact>, n, &grid, 2)) return rc2;
            pdf_pair_kernel<kExact><<<grid, kThreads, 0, s>>>(hot(ctx), i3, pdf, n);
        }
        rc = SSK_OK;
    } else
    if (fast(ctx)) rc = vec ? launch_pdf_t<kFast, 4>(ctx, i3, pdf, n, s) : launch_pdf_t<kFast, 1>(ctx, i3, pdf, n, s);
    else rc = vec ? launch_pdf_t<kExact, 4>(ctx, i3, pdf, n, s) : launch_pdf_t<kExact, 1>(ctx, i3, pdf, n, s);
    if (rc) return rc;
    return launched(ctx, s);
}

template <int A, int R, bool Spectral>
int launch_sample_t(const ssk_ctx *ctx, const SampleArgs &a, size_t n, cudaStream_t s) {
    int grid = 1;
    if (int rc = occupancy_grid(ctx, sample_kernel<A, R, Spectral>, n, &grid, R)) return rc;
    sample_kernel<A, R, Spectral><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
    return SSK_OK;
}
/// sample_direction moves 9 to 21 planes: one ray per thread (32-bit accesses, full 128-byte lines per warp)
/// keeps the staging registers low enough for 2-3 resident CTAs per SM
template <int A, bool Spectral>
int launch_sample_pair_t(const ssk_ctx *ctx, const SampleArgs &a, size_t n, cudaStream_t s) {
    int grid = 1;
    if (int rc = occupancy_grid(ctx, sample_pair_kernel<A, Spectral>, n, &grid, 2)) return rc;
    sample_pair_kernel<A, Spectral><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
    return SSK_OK;
}
template <int A, bool Spectral>
int launch_sample_sorted_t(const ssk_ctx *ctx, const SampleArgs &a, size_t n, cudaStream_t s) {
    int grid = 1;
    if (int rc = occupancy_grid(ctx, sample_sorted_kernel<A, Spectral>, n, &grid, 2)) return rc;
    sample_sorted_kernel<A, Spectral><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
    return SSK_OK;
}
/// every non-null plane of the argument block 8-byte aligned: the two-samples-per-thread kernel can be used
bool sample_planes_aligned8(const SampleArgs &a) {
    const void *ps[] = { a.u0, a.u1, a.px, a.py, a.pz, a.w0, a.w1, a.w2, a.w3, a.dx, a.dy, a.dz, a.pdf,
                         a.q0, a.q1, a.q2, a.q3, a.ox, a.oy, a.oz, a.dist };
    uintptr_t acc = 0;
    for (const void *q : ps) acc |= reinterpret_cast<uintptr_t>(q);
    return (acc & 7u) == 0;
}
/// spectral sample_direction as two kernels (see weight_spec_kernel): sampler pass (technique-sorted when both
/// techniques are in play), then eval / pdf
template <int A>
int launch_sample_two_pass(const ssk_ctx *ctx, const SampleArgs &a, size_t n, bool sorted, cudaStream_t s) {
    int grid = 1;
    if (sorted) {
        if (int rc = occupancy_grid(ctx, sample_sorted_kernel<A, false, false>, n, &grid, 2)) return rc;
        sample_sorted_kernel<A, false, false><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
    } else {
        if (int rc = occupancy_grid(ctx, sample_pair_kernel<A, false, false>, n, &grid, 2)) return rc;
        sample_pair_kernel<A, false, false><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
    }
    if (sorted) {                                  // same tiles, same slot order as the first pass
        if (int rc = occupancy_grid(ctx, weight_spec_sorted_kernel<A>, n, &grid, 2)) return rc;
        weight_spec_sorted_kernel<A><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
    } else {
        if (int rc = occupancy_grid(ctx, weight_spec_kernel<A>, n, &grid, 1)) return rc;
        weight_spec_kernel<A><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
    }
    g_launches++;
    return SSK_OK;
}
template <int A>
int launch_sample_a(const ssk_ctx *ctx, bool spectral, bool vec, const SampleArgs &a, size_t n, cudaStream_t s) {
    (void) vec;
    static const bool fused = getenv("SSK_SAMPLE_FUSED") != nullptr, two_pass_nosort = getenv("SSK_SAMPLE_2P_NOSORT") != nullptr;
    if (spectral && !fused && !getenv("SSK_RAYS_PER_THREAD") && sample_planes_aligned8(a)) {
        const float w2 = ctx->E.block.sky_w;
        return launch_sample_two_pass<A>(ctx, a, n, w2 > 0.f && w2 < 1.f && !two_pass_nosort, s);
    }
    // both techniques in play (0 < w < 1): technique-sorted tiles; one technique only: plain pairs
    static const bool no_sort = getenv("SSK_SAMPLE_NOSORT") != nullptr;
    const float w = ctx->E.block.sky_w;
    // Measured on B200 (2^24 samples, t=6 a=0.5 eta=50.2), fused kernels sorted vs unsorted: with one round per tile
    // FAST RGB 18.1 -> 20.9 Gsamples/s, EXACT RGB 13.8 -> 13.0, FAST spectral 14.8 -> 13.6, EXACT spectral 9.4 -> 6.6
    // (73-112 KB bodies that already stall on instruction fetch); with four rounds per tile (ssk_kernels.cuh,
    // SSK_TILE_ROUNDS) FAST RGB 24.2 and EXACT RGB 14.3 -> 15.3. Both RGB modes sort; spectral runs as two passes (above).
    static const bool no_sort_exact = getenv("SSK_SAMPLE_NOSORT_EXACT") != nullptr;
    if ((A == kFast || !no_sort_exact) && !spectral && !getenv("SSK_RAYS_PER_THREAD") && sample_planes_aligned8(a) && w > 0.f && w < 1.f && !no_sort)
        return launch_sample_sorted_t<A, false>(ctx, a, n, s);
    if (!getenv("SSK_RAYS_PER_THREAD") && sample_planes_aligned8(a))
        return spectral ? launch_sample_pair_t<A, true>(ctx, a, n, s) : launch_sample_pair_t<A, false>(ctx, a, n, s);
    if (spectral) return launch_sample_t<A, kSampleRays, true>(ctx, a, n, s);
    return launch_sample_t<A, kSampleRays, false>(ctx, a, n, s);
}

int launch_sample(const ssk_ctx *ctx, const float *const u[2], const float *const it_p[3],
                  const float *const wav[4], size_t n, float *const d[3], float *pdf, float *const weight[4],
                  float *const p[3], float *dist, cudaStream_t s) {
    DeviceGuard dg; if (int rc = need_device(ctx, dg)) return rc;
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    const int C = spectral ? 4 : 3;
    if (!u || !u[0] || !u[1] || !d || !d[0] || !d[1] || !d[2] || !pdf || !weight)
        return fail(SSK_ERR_INVALID, "null sample / output planes");
    if (n == 0) return SSK_OK;
    SampleArgs a = {};
    a.u0 = u[0]; a.u1 = u[1];
    bool vec = aligned16(u[0]) && aligned16(u[1]) && aligned16(d[0]) && aligned16(d[1]) && aligned16(d[2]) && aligned16(pdf);
    if (it_p && it_p[0]) {
        if (!it_p[1] || !it_p[2]) return fail(SSK_ERR_INVALID, "it_p needs 3 planes");
        a.px = it_p[0]; a.py = it_p[1]; a.pz = it_p[2];
        vec = vec && aligned16(a.px) && aligned16(a.py) && aligned16(a.pz);
    }
    if (spectral) {
        if (!wav || !wav[0] || !wav[1] || !wav[2] || !wav[3])
            return fail(SSK_ERR_INVALID, "spectral mode needs 4 wavelength planes");
        a.w0 = wav[0]; a.w1 = wav[1]; a.w2 = wav[2]; a.w3 = wav[3];
        for (int c = 0; c < 4; ++c) vec = vec && aligned16(wav[c]);
    }
    a.dx = d[0]; a.dy = d[1]; a.dz = d[2]; a.pdf = pdf;
    float *q[4] = { nullptr, nullptr, nullptr, nullptr };
    for (int c = 0; c < C; ++c) {
        if (!weight[c]) return fail(SSK_ERR_INVALID, "null weight plane");
        q[c] = weight[c]; vec = vec && aligned16(q[c]);
    }
    a.q0 = q[0]; a.q1 = q[1]; a.q2 = q[2]; a.q3 = q[3];
    if (p && p[0]) {
        if (!p[1] || !p[2]) return fail(SSK_ERR_INVALID, "p needs 3 planes");
        a.ox = p[0]; a.oy = p[1]; a.oz = p[2];
        vec = vec && aligned16(a.ox) && aligned16(a.oy) && aligned16(a.oz);
    }
    if (dist) { a.dist = dist; vec = vec && aligned16(dist); }
    int rc = fast(ctx) ? launch_sample_a<kFast>(ctx, spectral, vec, a, n, s) : launch_sample_a<kExact>(ctx, spectral, vec, a, n, s);
    if (rc) return rc;
    return launched(ctx, s);
}

/// sample_wavelengths: spectral -> wavelength importance sampling + eval / pdf; RGB -> eval (sunsky.cpp:463-480)
int launch_sample_wavelengths(const ssk_ctx *ctx, const float *const wi[3], const float *sample, size_t n,
                              float *const wav[4], float *const weight[4], cudaStream_t s) {
    DeviceGuard dg; if (int rc = need_device(ctx, dg)) return rc;
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    if (!spectral) return launch_eval(ctx, wi, nullptr, n, weight, 1.f, s);      // Wavelength(0), eval(si)
    if (!wi || !wi[0] || !wi[1] || !wi[2] || !sample) return fail(SSK_ERR_INVALID, "null direction / sample planes");
    if (!wav || !weight) return fail(SSK_ERR_INVALID, "null output planes");
    for (int c = 0; c < 4; ++c) if (!wav[c] || !weight[c]) return fail(SSK_ERR_INVALID, "null output plane");
    if (n == 0) return SSK_OK;
    Planes3 i3 = { wi[0], wi[1], wi[2] };
    OutPlanes4 w4 = { wav[0], wav[1], wav[2], wav[3] }, q4 = { weight[0], weight[1], weight[2], weight[3] };
    int grid = 1;
    if (fast(ctx)) {
        if (int rc = occupancy_grid(ctx, sample_wavelengths_kernel<kFast>, n, &grid, 1)) return rc;
        sample_wavelengths_kernel<kFast><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, i3, sample, w4, q4, n);
    } else {
        if (int rc = occupancy_grid(ctx, sample_wavelengths_kernel<kExact>, n, &grid, 1)) return rc;
        sample_wavelengths_kernel<kExact><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, i3, sample, w4, q4, n);
    }
    return launched(ctx, s);
}

template <int A, bool Spectral>
int launch_sample_ray_t(const ssk_ctx *ctx, const RayArgs &a, size_t n, cudaStream_t s) {
    int grid = 1;
    static const bool fused = getenv("SSK_SAMPLE_FUSED") != nullptr;
    static const bool scalar_rays = getenv("SSK_RAY_SCALAR") != nullptr;
    (void) scalar_rays;
#if !SSK_SORT_PER_WARP
    if (Spectral && !fused && !scalar_rays) {
        // two technique-sorted passes, as spectral sample_direction: geometry (direction, origin, pdf), then
        // wavelengths + eval + weight in the sorted order
        const void *planes[] = { a.s2x, a.s2y, a.s3x, a.s3y, a.ox, a.oy, a.oz, a.dx, a.dy, a.dz, a.q0, a.q1, a.q2, a.q3 };
        uintptr_t acc = 0;
        for (const void *q : planes) acc |= reinterpret_cast<uintptr_t>(q);
        if ((acc & 7u) == 0) {
            SampleArgs sa = {};
            sa.u0 = a.s3x; sa.u1 = a.s3y; sa.s2x = a.s2x; sa.s2y = a.s2y;
            sa.dx = a.dx; sa.dy = a.dy; sa.dz = a.dz; sa.ox = a.ox; sa.oy = a.oy; sa.oz = a.oz;
            sa.q0 = a.q0; sa.q1 = a.q1; sa.q2 = a.q2; sa.q3 = a.q3;
            sa.wl_sample = a.wl_sample; sa.ow0 = a.w0; sa.ow1 = a.w1; sa.ow2 = a.w2; sa.ow3 = a.w3;
            if (int rc = occupancy_grid(ctx, sample_sorted_kernel<A, false, false, true>, n, &grid, 2)) return rc;
            sample_sorted_kernel<A, false, false, true><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, sa, n);
            if (int rc = occupancy_grid(ctx, weight_spec_sorted_kernel<A, true>, n, &grid, 2)) return rc;
            weight_spec_sorted_kernel<A, true><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, sa, n);
            g_launches++;
            return SSK_OK;
        }
    }
#endif
    if (Spectral && !fused) {                      // two kernels, see weight_spec_kernel
        if (int rc = occupancy_grid(ctx, sample_ray_kernel<A, false, false>, n, &grid, 1)) return rc;
        sample_ray_kernel<A, false, false><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
        if (int rc = occupancy_grid(ctx, ray_weight_spec_kernel<A>, n, &grid, 1)) return rc;
        ray_weight_spec_kernel<A><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
        g_launches++;
        return SSK_OK;
    }
#if !SSK_SORT_PER_WARP
    // RGB: the technique-sorted ray-pair kernel of sample_direction in ray mode (origin, scaled pdf, -d). Measured on
    // 2^24 rays with four rounds per tile: FAST 21.0 vs 17.9 Grays/s for the scalar kernel below, EXACT 14.0 vs 13.2
    const void *planes[] = { a.s2x, a.s2y, a.s3x, a.s3y, a.ox, a.oy, a.oz, a.dx, a.dy, a.dz, a.q0, a.q1, a.q2 };
    uintptr_t acc = 0;
    for (const void *q : planes) acc |= reinterpret_cast<uintptr_t>(q);
    if (!Spectral && !scalar_rays && (acc & 7u) == 0) {
        SampleArgs sa = {};
        sa.u0 = a.s3x; sa.u1 = a.s3y; sa.s2x = a.s2x; sa.s2y = a.s2y;
        sa.dx = a.dx; sa.dy = a.dy; sa.dz = a.dz; sa.ox = a.ox; sa.oy = a.oy; sa.oz = a.oz;
        sa.q0 = a.q0; sa.q1 = a.q1; sa.q2 = a.q2;
        if (int rc = occupancy_grid(ctx, sample_sorted_kernel<A, false, true, true>, n, &grid, 2)) return rc;
        sample_sorted_kernel<A, false, true, true><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, sa, n);
        return SSK_OK;
    }
#endif
    if (int rc = occupancy_grid(ctx, sample_ray_kernel<A, Spectral>, n, &grid, 1)) return rc;
    sample_ray_kernel<A, Spectral><<<grid, kThreads, 0, s>>>(hot(ctx), ctx->d_block, a, n);
    return SSK_OK;
}
/// sample_ray (sunsky.cpp:354-397); `time` passes through the reference untouched and is not an argument
int launch_sample_ray(const ssk_ctx *ctx, const float *wl_sample, const float *const s2[2], const float *const s3[2],
                      size_t n, float *const o[3], float *const d[3], float *const wav[4], float *const weight[4],
                      cudaStream_t s) {
    DeviceGuard dg; if (int rc = need_device(ctx, dg)) return rc;
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    const int C = spectral ? 4 : 3;
    if (!s2 || !s2[0] || !s2[1] || !s3 || !s3[0] || !s3[1]) return fail(SSK_ERR_INVALID, "null sample planes");
    if (!o || !o[0] || !o[1] || !o[2] || !d || !d[0] || !d[1] || !d[2] || !weight)
        return fail(SSK_ERR_INVALID, "null output planes");
    RayArgs a = {};
    a.s2x = s2[0]; a.s2y = s2[1]; a.s3x = s3[0]; a.s3y = s3[1];
    a.ox = o[0]; a.oy = o[1]; a.oz = o[2]; a.dx = d[0]; a.dy = d[1]; a.dz = d[2];
    if (spectral) {
        if (!wl_sample) return fail(SSK_ERR_INVALID, "spectral mode needs the wavelength sample plane");
        if (!wav || !wav[0] || !wav[1] || !wav[2] || !wav[3])
            return fail(SSK_ERR_INVALID, "spectral mode needs 4 wavelength output planes");
        a.wl_sample = wl_sample; a.w0 = wav[0]; a.w1 = wav[1]; a.w2 = wav[2]; a.w3 = wav[3];
    }
    float *q[4] = { nullptr, nullptr, nullptr, nullptr };
    for (int c = 0; c < C; ++c) {
        if (!weight[c]) return fail(SSK_ERR_INVALID, "null weight plane");
        q[c] = weight[c];
    }
    a.q0 = q[0]; a.q1 = q[1]; a.q2 = q[2]; a.q3 = q[3];
    if (n == 0) return SSK_OK;
    int rc;
    if (fast(ctx)) rc = spectral ? launch_sample_ray_t<kFast, true>(ctx, a, n, s) : launch_sample_ray_t<kFast, false>(ctx, a, n, s);
    else rc = spectral ? launch_sample_ray_t<kExact, true>(ctx, a, n, s) : launch_sample_ray_t<kExact, false>(ctx, a, n, s);
    if (rc) return rc;
    return launched(ctx, s);
}

// ---- host entry points: chunked H2D -> kernel -> D2H over 3 streams with pinned staging ---------
int ensure_staging(const ssk_ctx *ctx, int planes) {
    Staging &st = ctx->st;
    // elements per chunk and plane and streams in rotation (SSK_HOST_CHUNK_LOG2 / SSK_HOST_STREAMS). Measured on
    // B200, spectral eval, 2^24 rays, pinned buffers, round-2 pipeline (one strided copy per chunk and direction, no
    // host synchronisation between chunks): 2^18 1.80, 2^19 1.84, 2^20 1.82, 2^21 1.76 Gevals/s at one GPU (the copy
    // ceiling of the same bytes is 1.85), 2^19 3.48 vs 2^21 3.39 at eight; 3 or 4 streams alike. Small chunks shorten
    // the pipeline's fill and drain (first H2D, last D2H), which nothing overlaps. Round 1's pipeline (11 plane copies
    // and a host synchronisation per chunk) had its optimum at 2^21.
    static const int chunk_log2 = getenv("SSK_HOST_CHUNK_LOG2") ? std::min(24, std::max(12, atoi(getenv("SSK_HOST_CHUNK_LOG2")))) : 19;
    static const int n_streams = getenv("SSK_HOST_STREAMS") ? std::min(Staging::kStreams, std::max(1, atoi(getenv("SSK_HOST_STREAMS")))) : 3;
    const size_t chunk = (size_t) 1 << chunk_log2;
    if (st.planes >= planes && st.chunk == chunk) return SSK_OK;
    st.active = n_streams;
    for (int i = 0; i < st.active; ++i) {
        if (st.pinned[i]) cudaFreeHost(st.pinned[i]);
        if (st.dev[i]) cudaFree(st.dev[i]);
        st.pinned[i] = nullptr; st.dev[i] = nullptr;
        if (!st.stream[i]) SSK_CUDA(cudaStreamCreateWithFlags(&st.stream[i], cudaStreamNonBlocking));
        SSK_CUDA(cudaMallocHost(&st.pinned[i], (size_t) planes * chunk * sizeof(float)));
        SSK_CUDA(cudaMalloc(&st.dev[i], (size_t) planes * chunk * sizeof(float)));
    }
    st.planes = planes; st.chunk = chunk;
    return SSK_OK;
}

bool is_pinned(const void *p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}

/// Planes k = 0..n-1 start at equal byte distances (views of one (n, N) array): one strided copy moves all of them
template <typename P> bool equally_spaced(P *const *planes, int n, ptrdiff_t *stride_bytes) {
    if (n < 2) { *stride_bytes = 0; return n == 1; }
    const ptrdiff_t d = reinterpret_cast<const char *>(planes[1]) - reinterpret_cast<const char *>(planes[0]);
    if (d <= 0) return false;
    for (int k = 2; k < n; ++k)
        if (reinterpret_cast<const char *>(planes[k]) - reinterpret_cast<const char *>(planes[k - 1]) != d) return false;
    *stride_bytes = d;
    return true;
}

/// Generic pipeline: `n_in` input planes, `n_out` output planes, `launch(dev_in[], dev_out[], m, stream)`.
/// The batch moves in chunks over `active` slots, each slot a stream with its own device staging: H2D, kernel and
/// D2H of one chunk are ordered on the slot's stream, chunks of different slots overlap (the two copy engines run
/// H2D of one chunk while another chunk's results go back). Pinned caller buffers (the drop-in case: Dr.Jit and
/// torch hand over pinned or device memory) need no host-side synchronisation until the end; planes that are views
/// of one array travel as ONE strided copy per chunk and direction (cudaMemcpy2DAsync, rows = planes) instead of one
/// copy per plane. Pageable buffers go through the slot's pinned staging with a host memcpy on either side.
template <typename Launch>
int host_pipeline(const ssk_ctx *ctx, const float *const *in, int n_in, float *const *out, int n_out, size_t n,
                  Launch &&launch) {
    DeviceGuard dg; if (int rc = need_device(ctx, dg)) return rc;
    if (n == 0) return SSK_OK;
    std::lock_guard<std::mutex> lock(ctx->st.mu);
    if (int rc = ensure_staging(ctx, n_in + n_out)) return rc;
    Staging &st = ctx->st;
    const size_t chunk = st.chunk;
    bool pin_in = true, pin_out = true;
    for (int k = 0; k < n_in; ++k) pin_in = pin_in && is_pinned(in[k]);
    for (int k = 0; k < n_out; ++k) pin_out = pin_out && is_pinned(out[k]);
    static const bool no_2d = getenv("SSK_HOST_NO_2D") != nullptr;
    ptrdiff_t in_stride = 0, out_stride = 0;
    const bool in_2d = !no_2d && pin_in && n_in > 1 && equally_spaced(in, n_in, &in_stride);
    const bool out_2d = !no_2d && pin_out && n_out > 1 && equally_spaced(out, n_out, &out_stride);
    const size_t n_chunks = (n + chunk - 1) / chunk;
    // pending[i]: chunk index whose results still sit in staging slot i (pageable outputs only)
    std::vector<long long> pending(Staging::kStreams, -1);
    const int n_slots = st.active;
    auto drain = [&](int slot) -> int {
        if (pending[slot] < 0) return SSK_OK;
        SSK_CUDA(cudaStreamSynchronize(st.stream[slot]));
        if (!pin_out) {
            size_t off = (size_t) pending[slot] * chunk, m = std::min(chunk, n - off);
            for (int k = 0; k < n_out; ++k)
                memcpy(out[k] + off, st.pinned[slot] + (size_t) (n_in + k) * chunk, m * sizeof(float));
        }
        pending[slot] = -1;
        return SSK_OK;
    };
    const bool host_sync = !(pin_in && pin_out);          // pinned on both sides: stream order is all the ordering needed
    for (size_t ci = 0; ci < n_chunks; ++ci) {
        int slot = (int) (ci % n_slots);
        if (host_sync) if (int rc = drain(slot)) return rc;
        size_t off = ci * chunk, m = std::min(chunk, n - off);
        cudaStream_t s = st.stream[slot];
        const float *din[16]; float *dout[16];
        for (int k = 0; k < n_in; ++k) din[k] = st.dev[slot] + (size_t) k * chunk;
        for (int k = 0; k < n_out; ++k) dout[k] = st.dev[slot] + (size_t) (n_in + k) * chunk;
        if (in_2d) {
            SSK_CUDA(cudaMemcpy2DAsync(st.dev[slot], chunk * sizeof(float), in[0] + off, (size_t) in_stride,
                                       m * sizeof(float), n_in, cudaMemcpyHostToDevice, s));
        } else {
            for (int k = 0; k < n_in; ++k) {
                const float *src = in[k] + off;
                if (!pin_in) {
                    float *stage = st.pinned[slot] + (size_t) k * chunk;
                    memcpy(stage, src, m * sizeof(float));
                    src = stage;
                }
                SSK_CUDA(cudaMemcpyAsync(const_cast<float *>(din[k]), src, m * sizeof(float), cudaMemcpyHostToDevice, s));
            }
        }
        if (int rc = launch(din, dout, m, s)) return rc;
        if (out_2d) {
            SSK_CUDA(cudaMemcpy2DAsync(out[0] + off, (size_t) out_stride, dout[0], chunk * sizeof(float),
                                       m * sizeof(float), n_out, cudaMemcpyDeviceToHost, s));
        } else {
            for (int k = 0; k < n_out; ++k) {
                float *dst = pin_out ? out[k] + off : st.pinned[slot] + (size_t) (n_in + k) * chunk;
                SSK_CUDA(cudaMemcpyAsync(dst, dout[k], m * sizeof(float), cudaMemcpyDeviceToHost, s));
            }
        }
        pending[slot] = (long long) ci;
    }
    for (int i = 0; i < n_slots; ++i)
        if (int rc = drain(i)) return rc;
    return SSK_OK;
}

template <int A, bool Spectral>
int launch_envmap_t(const ssk_ctx *c, float4 wl, float *out, int width, int height, cudaStream_t s) {
    int grid = 1;
    if (int rc = occupancy_grid(c, envmap_kernel<A, Spectral>, (size_t) width * height, &grid, 4)) return rc;
    envmap_kernel<A, Spectral><<<grid, kThreads, 0, s>>>(hot(c), c->d_block, wl, out, width, height);
    return SSK_OK;
}

template <typename F> int guarded(F &&f) {
    try { return f(); }
    catch (const std::invalid_argument &e) { return fail(SSK_ERR_INVALID, e.what()); }
    catch (const std::runtime_error &e) {
        std::string m = e.what();
        bool io = m.find("file") != std::string::npos || m.find("dataset") != std::string::npos ||
                  m.find("truncated") != std::string::npos || m.find("header") != std::string::npos;
        return fail(io ? SSK_ERR_IO : SSK_ERR_INVALID, m);
    }
    catch (const std::exception &e) { return fail(SSK_ERR_INVALID, e.what()); }
}

} // namespace

extern "C" {

const char *ssk_last_error(void) { return g_err.c_str(); }
const char *ssk_version(void) { return "sunsky_b200 0.1 (sm_100a)"; }
uint64_t ssk_launch_count(void) { return g_launches.load(); }

void ssk_params_default(ssk_params *p) {
    memset(p, 0, sizeof(*p));
    p->turbidity = 3.f;
    p->n_albedo = 1; p->albedo[0] = 0.3f;
    p->use_record = 1;
    p->sun_direction[2] = 1.f;
    p->latitude = 35.6894f; p->longitude = 139.6917f; p->timezone = 9.f;
    p->year = 2010; p->month = 7; p->day = 10;
    p->hour = 15.f; p->minute = 0.f; p->second = 0.f;
    p->sun_scale = 1.f; p->sky_scale = 1.f;
    p->sun_aperture_deg = (float) 0.5358;
    for (int i = 0; i < 4; ++i) p->to_world[i * 5] = p->to_local[i * 5] = 1.f;
    p->flags = 0; p->arith = SSK_ARITH_EXACT;
}

int ssk_create(const ssk_params *p, const char *dataset_dir, int device, int mode, ssk_ctx **out) {
    if (!p || !dataset_dir || !out) return fail(SSK_ERR_INVALID, "null argument");
    if (mode != SSK_MODE_RGB && mode != SSK_MODE_SPECTRAL)
        return fail(SSK_ERR_UNSUPPORTED, "Unsupported spectrum type, can only render in Spectral or RGB modes!");
    *out = nullptr;
    ssk_ctx *ctx = new ssk_ctx();
    int rc = guarded([&]() -> int {
        ctx->mode = mode;
        ctx->device = device;
        ctx->E.spectral = mode == SSK_MODE_SPECTRAL;
        ctx->E.n_channels = ctx->E.spectral ? 11 : 3;
        ctx->E.data = datasets_for(dataset_dir, ctx->E.spectral);
        if (device >= 0) {
            int count = 0;
            cudaError_t e = cudaGetDeviceCount(&count);
            if (e != cudaSuccess || device >= count) {
                validate_params(*p, ctx->E.spectral);          // property errors win over the missing device
                return fail(SSK_ERR_CUDA, std::string("CUDA device ") + std::to_string(device) + " is not available: " +
                                              (e != cudaSuccess ? cudaGetErrorString(e) : "ordinal out of range"));
            }
            DeviceGuard dg; if (int rc2 = dg.enter(device)) return rc2;
            SSK_CUDA(cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device));
            if (int rc2 = alloc_blocks(ctx)) return rc2;
        }
        build_on(ctx, *p, true);
        return upload_block(ctx);
    });
    if (rc) { ssk_destroy(ctx); return rc; }
    *out = ctx;
    return SSK_OK;
}

int ssk_update(ssk_ctx *ctx, const ssk_params *p) {
    if (!ctx || !p) return fail(SSK_ERR_INVALID, "null argument");
    return guarded([&]() -> int {
        HostEmitter backup = ctx->E;
        const int cur0 = ctx->cur;
        // the new state goes to the next block of the ring: launches already queued keep reading the old one
        if (int rc = rotate_block(ctx)) return rc;
        try { build_on(ctx, *p, false); }
        catch (...) { ctx->E = backup; ctx->cur = cur0; ctx->d_block = ctx->ring[cur0]; throw; }
        return upload_block(ctx);
    });
}

int ssk_update_batch(ssk_ctx *const *ctxs, const ssk_params *params, int n) {
    if (n < 0 || (n > 0 && (!ctxs || !params))) return fail(SSK_ERR_INVALID, "null argument");
    if (n == 0) return SSK_OK;
    for (int i = 0; i < n; ++i) {
        if (!ctxs[i]) return fail(SSK_ERR_INVALID, "null context");
        if (ctxs[i]->device != ctxs[0]->device || ctxs[i]->mode != ctxs[0]->mode)
            return fail(SSK_ERR_INVALID, "ssk_update_batch: contexts must share one device and one mode");
    }
    if (ctxs[0]->device < 0 || getenv("SSK_HOST_SETUP")) {
        std::vector<HostEmitter> backup(n);
        for (int i = 0; i < n; ++i) backup[i] = ctxs[i]->E;
        for (int i = 0; i < n; ++i)
            if (int rc = ssk_update(ctxs[i], &params[i])) {        // all or nothing: undo the ones already updated
                std::string msg = g_err;
                for (int j = 0; j < i; ++j) { ctxs[j]->E = backup[j]; if (ctxs[j]->device >= 0) { rotate_block(ctxs[j]); upload_block(ctxs[j]); } }
                return fail(rc, msg);
            }
        return SSK_OK;
    }
    return guarded([&]() -> int {
        std::vector<HostEmitter> backup(n);
        std::vector<int> cur0(n);
        for (int i = 0; i < n; ++i) { backup[i] = ctxs[i]->E; cur0[i] = ctxs[i]->cur; }
        // on failure every context goes back to the block it was on (untouched: updates write the next one)
        auto rollback = [&] { for (int i = 0; i < n; ++i) { ctxs[i]->E = backup[i]; ctxs[i]->cur = cur0[i]; ctxs[i]->d_block = ctxs[i]->ring[cur0[i]]; } };
        for (int i = 0; i < n; ++i) if (int rc = rotate_block(ctxs[i])) { rollback(); return rc; }
        static const bool trace = getenv("SSK_SETUP_TRACE") != nullptr;
        auto now = [] { return std::chrono::steady_clock::now(); };
        auto us = [](auto a, auto b) { return std::chrono::duration<double, std::micro>(b - a).count(); };
        auto t0 = now(), t1 = t0, t2 = t0, t3 = t0;
        try {
            for (int i = 0; i < n; ++i) build_emitter_pre(ctxs[i]->E, params[i], false);
            t1 = now();
            std::vector<float> sky((size_t) n * 11), sun((size_t) n * 11);
            if (int rc = quadrature_device_batch(ctxs, n, reinterpret_cast<float (*)[11]>(sky.data()),
                                                 reinterpret_cast<float (*)[11]>(sun.data()))) { rollback(); return rc; }
            t2 = now();
            for (int i = 0; i < n; ++i)
                build_emitter_post(ctxs[i]->E, &sky[(size_t) i * 11], &sun[(size_t) i * 11]);
            t3 = now();
        } catch (...) { rollback(); throw; }
        {   // final blocks: enqueued together on the setup stream, one synchronisation
            SetupScratch &sc = g_setup[ctxs[0]->device];
            std::lock_guard<std::mutex> lock(sc.mu);
            DeviceGuard dg; if (int rc = dg.enter(ctxs[0]->device)) return rc;
            for (int i = 0; i < n; ++i)
                SSK_CUDA(cudaMemcpyAsync(ctxs[i]->d_block, &ctxs[i]->E.block, sizeof(EmitterBlock), cudaMemcpyHostToDevice, sc.stream));
            SSK_CUDA(cudaStreamSynchronize(sc.stream));
        }
        if (trace)
            fprintf(stderr, "ssk update_batch n=%d: backup+pre %.1f us, quadrature %.1f us, post %.1f us, upload %.1f us\n", n,
                    us(t0, t1), us(t1, t2), us(t2, t3), us(t3, now()));
        return SSK_OK;
    });
}

// ---- parameters_changed without any CUDA allocation (no counterpart in the reference) ----------
// The setup quadrature needs scratch: per emitter the node values of both integrands (chain-major, kEmitterStride
// floats), the chain partials, and the Gauss-Legendre table. ssk_update / ssk_update_batch keep these in lazily
// cudaMalloc'ed buffers; this variant keeps them in ONE file-scope device array of fixed size, resident on every GPU
// that loads the library until the process exits:
//   [ kNoAllocMaxEmitters x kEmitterStride | kNoAllocMaxEmitters x kMaxChains partials (padded) | 2 x kQuadN table ]
constexpr int kNoAllocMaxEmitters = SSK_NOALLOC_MAX_EMITTERS;
constexpr size_t kNoAllocPartialOff = (size_t) kNoAllocMaxEmitters * kEmitterStride;
constexpr size_t kNoAllocPartialLen = ((size_t) kNoAllocMaxEmitters * kMaxChains + 63) / 64 * 64;      // keeps the table 256-byte aligned
constexpr size_t kNoAllocTableOff = kNoAllocPartialOff + kNoAllocPartialLen;
constexpr size_t kNoAllocFloats = kNoAllocTableOff + 2 * kQuadN;
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocFloats];

struct NoAllocDevice {      // one per device: the array's address ON THAT DEVICE (the symbol itself is a host-side stand-in)
    float *ptr = nullptr;
    std::mutex mu;          // calls on one device serialise here; calls that overlap on the GPU would race on the array
};
NoAllocDevice g_noalloc[kMaxDevices];

/// Address of the scratch array on `device` (current device on entry); the first call per device also puts the
/// Gauss-Legendre table into its region. No allocation: cudaGetSymbolAddress only resolves the module's symbol.
int noalloc_scratch(int device, float **out) {
    NoAllocDevice &d = g_noalloc[device];
    if (!d.ptr) {
        float *p = nullptr;
        SSK_CUDA(cudaGetSymbolAddress((void **) &p, g_noalloc_scratch));
        if (!p) return fail(SSK_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the no-alloc scratch array");
        SSK_CUDA(cudaMemcpy(p + kNoAllocTableOff, gauss_legendre_200().data(), 2 * kQuadN * sizeof(float), cudaMemcpyHostToDevice));
        d.ptr = p;
    }
    *out = d.ptr;
    return SSK_OK;
}

/// quadrature_device_batch on the static scratch, default stream, synchronous copies into stack buffers
int quadrature_device_noalloc(ssk_ctx *const *ctxs, int n, float (*sky_rad)[11], float (*sun_rad)[11]) {
    const int device = ctxs[0]->device;
    if (device >= kMaxDevices) return fail(SSK_ERR_UNSUPPORTED, "device ordinal too large");
    std::lock_guard<std::mutex> lock(g_noalloc[device].mu);
    DeviceGuard dg; if (int rc = dg.enter(device)) return rc;
    float *base = nullptr;
    if (int rc = noalloc_scratch(device, &base)) return rc;
    float *d_buf = base, *d_partial = base + kNoAllocPartialOff, *d_gl = base + kNoAllocTableOff;
    const int C = ctxs[0]->E.n_channels, nchains = 2 * C * kQuadChunks;
    // the array keeps the previous call's contents: the partials start from zero so that a launch that failed to run
    // cannot hand back the last call's sums (the node values are overwritten in full for every chain that is read)
    SSK_CUDA(cudaMemsetAsync(d_partial, 0, kNoAllocPartialLen * sizeof(float), 0));
    for (int e = 0; e < n; ++e) {
        ssk_ctx *ctx = ctxs[e];
        SSK_CUDA(cudaMemcpyAsync(ctx->d_block, &ctx->E.block, sizeof(EmitterBlock), cudaMemcpyHostToDevice, 0));
        setup_quadrature_kernel<<<(2 * kQuadNodes + kThreads - 1) / kThreads, kThreads>>>(
            ctx->E.block, ctx->d_block, d_gl, d_buf + (size_t) e * kEmitterStride, C);
    }
    setup_sum_kernel<<<n * nchains, 32>>>(d_buf, d_partial, nchains, kEmitterStride);
    g_launches += n + 1;
    SSK_CUDA(cudaGetLastError());
    float h_partial[kNoAllocMaxEmitters * kMaxChains];
    SSK_CUDA(cudaMemcpy(h_partial, d_partial, (size_t) n * nchains * sizeof(float), cudaMemcpyDeviceToHost));
    for (int e = 0; e < n; ++e)
        for (int which = 0; which < 2; ++which) {
            const float J = which == 0 ? quad_sky_jacobian() : quad_sun_jacobian(ctxs[e]->E.block);
            for (int c = 0; c < C; ++c) {
                float total = 0.f;                   // llvm_block_sum: partials added in chunk order
                for (int k = 0; k < kQuadChunks; ++k) total = total + h_partial[e * nchains + (which * C + c) * kQuadChunks + k];
                (which == 0 ? sky_rad : sun_rad)[e][c] = total * J;
            }
        }
    return SSK_OK;
}

int ssk_update_batch_noalloc(ssk_ctx *const *ctxs, const ssk_params *params, int n) {
    if (n < 0 || (n > 0 && (!ctxs || !params))) return fail(SSK_ERR_INVALID, "null argument");
    if (n == 0) return SSK_OK;
    if (n > kNoAllocMaxEmitters)
        return fail(SSK_ERR_INVALID, "ssk_update_batch_noalloc: " + std::to_string(n) + " emitters exceed the fixed scratch of SSK_NOALLOC_MAX_EMITTERS = " +
                                     std::to_string(kNoAllocMaxEmitters) + " per call; split the batch or use ssk_update_batch");
    for (int i = 0; i < n; ++i) {
        if (!ctxs[i]) return fail(SSK_ERR_INVALID, "null context");
        if (ctxs[i]->device != ctxs[0]->device || ctxs[i]->mode != ctxs[0]->mode)
            return fail(SSK_ERR_INVALID, "ssk_update_batch_noalloc: contexts must share one device and one mode");
        for (int j = 0; j < i; ++j)
            if (ctxs[j] == ctxs[i]) return fail(SSK_ERR_INVALID, "ssk_update_batch_noalloc: a context appears twice");
    }
    if (ctxs[0]->device < 0 || !ctxs[0]->d_block)
        return fail(SSK_ERR_CUDA, "ssk_update_batch_noalloc: context has no CUDA device (the setup kernels have no CPU fallback here)");
    return guarded([&]() -> int {
        HostEmitter backup[kNoAllocMaxEmitters];
        int cur0[kNoAllocMaxEmitters];
        for (int i = 0; i < n; ++i) { backup[i] = ctxs[i]->E; cur0[i] = ctxs[i]->cur; }
        auto rollback = [&] { for (int i = 0; i < n; ++i) { ctxs[i]->E = backup[i]; ctxs[i]->cur = cur0[i]; ctxs[i]->d_block = ctxs[i]->ring[cur0[i]]; } };
        for (int i = 0; i < n; ++i) if (int rc = rotate_block(ctxs[i])) { rollback(); return rc; }
        try {
            for (int i = 0; i < n; ++i) build_emitter_pre(ctxs[i]->E, params[i], false);
            float sky[kNoAllocMaxEmitters][11], sun[kNoAllocMaxEmitters][11];
            if (int rc = quadrature_device_noalloc(ctxs, n, sky, sun)) { rollback(); return rc; }
            for (int i = 0; i < n; ++i) build_emitter_post(ctxs[i]->E, sky[i], sun[i]);
        } catch (...) { rollback(); throw; }
        for (int i = 0; i < n; ++i)
            if (int rc = upload_block(ctxs[i])) return rc;
        return SSK_OK;
    });
}

int ssk_set_scene(ssk_ctx *ctx, const float *bmin, const float *bmax) {
    if (!ctx) return fail(SSK_ERR_INVALID, "null context");
    // set_scene (sunsky.cpp:287-301); math::RayEpsilon<float> = 89.40697e-6 (Epsilon * 1500)
    const float ray_eps = SSK_EPSILON * 1500.f;
    HostEmitter &E = ctx->E;
    bool valid = bmin && bmax && bmin[0] <= bmax[0] && bmin[1] <= bmax[1] && bmin[2] <= bmax[2];
    if (valid) {
        // ScalarBoundingBox3f::bounding_sphere(): center = (min + max) * 0.5, radius = |max - center|
        float c[3], r2 = 0.f;
        for (int i = 0; i < 3; ++i) { c[i] = (bmax[i] + bmin[i]) * 0.5f; }
        for (int i = 0; i < 3; ++i) { float d = bmax[i] - c[i]; r2 += d * d; }
        float radius = sqrtf(r2);
        for (int i = 0; i < 3; ++i) E.bs_center[i] = c[i];
        E.bs_radius = fmaxf(ray_eps, radius * (1.f + ray_eps));
    } else {
        E.bs_center[0] = E.bs_center[1] = E.bs_center[2] = 0.f;
        E.bs_radius = ray_eps;
    }
    for (int i = 0; i < 3; ++i) E.block.bs_center[i] = E.bs_center[i];
    E.block.bs_radius = E.bs_radius;
    set_ray_pdf_scale(E.block);
    if (int rc = rotate_block(ctx)) return rc;
    return upload_block(ctx);
}

void ssk_destroy(ssk_ctx *ctx) {
    if (!ctx) return;
    if (ctx->device >= 0) {
        DeviceGuard dg; dg.enter(ctx->device);
        for (int i = 0; i < kBlockRing; ++i) {
            if (ctx->used[i]) { cudaEventSynchronize(ctx->used[i]); cudaEventDestroy(ctx->used[i]); }
            if (ctx->ring[i]) cudaFree(ctx->ring[i]);
        }
        for (int i = 0; i < Staging::kStreams; ++i) {
            if (ctx->st.stream[i]) { cudaStreamSynchronize(ctx->st.stream[i]); cudaStreamDestroy(ctx->st.stream[i]); }
            if (ctx->st.pinned[i]) cudaFreeHost(ctx->st.pinned[i]);
            if (ctx->st.dev[i]) cudaFree(ctx->st.dev[i]);
        }
    }
    delete ctx;
}

int ssk_channels(const ssk_ctx *ctx) { return ctx ? (ctx->mode == SSK_MODE_SPECTRAL ? 4 : 3) : 0; }
int ssk_device(const ssk_ctx *ctx) { return ctx ? ctx->device : -1; }

int ssk_get_state(const ssk_ctx *ctx, ssk_state *s) {
    if (!ctx || !s) return fail(SSK_ERR_INVALID, "null argument");
    const HostEmitter &E = ctx->E;
    memset(s, 0, sizeof(*s));
    s->n_channels = E.n_channels;
    memcpy(s->sky_params, E.sky_params, sizeof(float) * E.n_channels * 9);
    memcpy(s->sky_radiance, E.sky_radiance, sizeof(float) * E.n_channels);
    s->n_sun_radiance = (int32_t) E.sun_radiance.size();
    memcpy(s->sun_radiance, E.sun_radiance.data(), sizeof(float) * E.sun_radiance.size());
    if (E.spectral) memcpy(s->sun_ld, E.data->sun_ld.data(), sizeof(float) * 66);
    memcpy(s->gaussians, E.gaussians, sizeof(E.gaussians));
    memcpy(s->pmf, E.pmf, sizeof(E.pmf)); memcpy(s->cdf, E.cdf, sizeof(E.cdf));
    s->cdf_sum = E.cdf_sum; s->cdf_normalization = E.cdf_norm; s->sky_sampling_w = E.sky_w;
    s->sun_angles[0] = E.sun_angles[0]; s->sun_angles[1] = E.sun_angles[1];
    for (int i = 0; i < 3; ++i) {
        s->frame_s[i] = E.block.sun_s[i]; s->frame_t[i] = E.block.sun_t[i]; s->frame_n[i] = E.block.sun_n[i];
        s->bsphere_center[i] = E.bs_center[i];
    }
    s->sun_dir_world[0] = E.sun_dir_world.x; s->sun_dir_world[1] = E.sun_dir_world.y; s->sun_dir_world[2] = E.sun_dir_world.z;
    s->sun_half_aperture = E.half_aperture;
    s->bsphere_radius = E.bs_radius;
    memcpy(s->spectral_pdf, E.spectral_pdf, sizeof(E.spectral_pdf));
    memcpy(s->spectral_cdf, E.spectral_cdf, sizeof(E.spectral_cdf));
    s->spectral_integral = E.spectral_integral; s->spectral_normalization = E.spectral_norm;
    s->area_ratio = E.area_ratio;
    s->sun_below_horizon = E.sun_below_horizon ? 1 : 0;
    memcpy(s->to_local, E.block.to_local, sizeof(s->to_local));
    return SSK_OK;
}

int ssk_to_string(const ssk_ctx *ctx, char *buf, size_t len) {
    if (!ctx || !buf) return fail(SSK_ERR_INVALID, "null argument");
    const HostEmitter &E = ctx->E; const ssk_params &p = E.params;
    char tmp[1024];
    int k = snprintf(tmp, sizeof(tmp),
                     "SunskyEmitter[\n  bsphere = [center = [%g, %g, %g], radius = %g]\n  turbidity = %g\n"
                     "  sky_scale = %g\n  sun_scale = %g\n  albedo = %g%s\n  sun aperture (deg) = %g\n",
                     E.bs_center[0], E.bs_center[1], E.bs_center[2], E.bs_radius, p.turbidity, p.sky_scale,
                     p.sun_scale, p.albedo[0], p.n_albedo > 1 ? ", ..." : "", p.sun_aperture_deg);
    if (E.active_record)
        snprintf(tmp + k, sizeof(tmp) - k,
                 "  location = LocationRecord[latitude = %g, longitude = %g, timezone = %g]\n"
                 "  date_time = DateTimeRecord[year = %d, month= %d, day = %d, hour = %g, minute = %g, second = %g]\n]",
                 p.latitude, p.longitude, p.timezone, p.year, p.month, p.day, p.hour, p.minute, p.second);
    else
        snprintf(tmp + k, sizeof(tmp) - k, "  sun_dir = [%g, %g, %g]\n]", E.sun_dir_world.x, E.sun_dir_world.y, E.sun_dir_world.z);
    snprintf(buf, len, "%s", tmp);
    return SSK_OK;
}

int ssk_eval_device(const ssk_ctx *ctx, const float *const wi[3], const float *const wav[4], size_t n,
                    float *const out[4], void *stream) {
    return launch_eval(ctx, wi, wav, n, out, 1.f, (cudaStream_t) stream);
}
int ssk_eval_direction_device(const ssk_ctx *ctx, const float *const d[3], const float *const wav[4], size_t n,
                              float *const out[4], void *stream) {
    return launch_eval(ctx, d, wav, n, out, -1.f, (cudaStream_t) stream);
}
int ssk_pdf_direction_device(const ssk_ctx *ctx, const float *const d[3], size_t n, float *pdf, void *stream) {
    return launch_pdf(ctx, d, n, pdf, (cudaStream_t) stream);
}
int ssk_sample_direction_device(const ssk_ctx *ctx, const float *const u[2], const float *const it_p[3],
                                const float *const wav[4], size_t n, float *const d[3], float *pdf,
                                float *const weight[4], float *const p[3], float *dist, void *stream) {
    return launch_sample(ctx, u, it_p, wav, n, d, pdf, weight, p, dist, (cudaStream_t) stream);
}

int ssk_envmap_device(ssk_ctx *const *ctxs, int n_maps, int width, int height, const float *wavelengths4,
                      float *out, void *stream) {
    if (!ctxs || n_maps <= 0 || width <= 0 || height <= 0 || !out) return fail(SSK_ERR_INVALID, "bad envmap arguments");
    const ssk_ctx *c0 = ctxs[0];
    DeviceGuard dg; if (int rc = need_device(c0, dg)) return rc;
    const bool spectral = c0->mode == SSK_MODE_SPECTRAL;
    if (spectral && !wavelengths4) return fail(SSK_ERR_INVALID, "spectral env maps need 4 wavelengths per map");
    for (int i = 0; i < n_maps; ++i)
        if (!ctxs[i] || ctxs[i]->device != c0->device || ctxs[i]->mode != c0->mode || !ctxs[i]->d_block)
            return fail(SSK_ERR_INVALID, "env-map contexts must share device and mode");
    cudaStream_t s = (cudaStream_t) stream;
    const size_t map_floats = (size_t) (spectral ? 4 : 3) * width * height;
    // one launch per map: each carries its emitter block as the kernel parameter, launches queue back to back
    for (int i = 0; i < n_maps; ++i) {
        const ssk_ctx *c = ctxs[i];
        float4 wl = make_float4(0.f, 0.f, 0.f, 0.f);
        if (spectral) wl = make_float4(wavelengths4[4 * i], wavelengths4[4 * i + 1], wavelengths4[4 * i + 2], wavelengths4[4 * i + 3]);
        float *o = out + (size_t) i * map_floats;
        int rc;
        if (fast(c)) rc = spectral ? launch_envmap_t<kFast, true>(c, wl, o, width, height, s) : launch_envmap_t<kFast, false>(c, wl, o, width, height, s);
        else rc = spectral ? launch_envmap_t<kExact, true>(c, wl, o, width, height, s) : launch_envmap_t<kExact, false>(c, wl, o, width, height, s);
        if (rc) return rc;
        if (int rc2 = launched(c, s)) return rc2;
    }
    return SSK_OK;
}

int ssk_debug_fdiv_by_device(const float *a, const float *b, float *out, size_t n, void *stream) {
    if (!a || !b || !out) return fail(SSK_ERR_INVALID, "null argument");
    if (n == 0) return SSK_OK;
    fdiv_by_kernel<<<1184, 256, 0, (cudaStream_t) stream>>>(a, b, out, n);
    g_launches++;
    SSK_CUDA(cudaGetLastError());
    return SSK_OK;
}

int ssk_debug_set_exact_limits(ssk_ctx *ctx, float lim_s, float lim_gamma) {
    if (!ctx) return fail(SSK_ERR_INVALID, "null context");
    ctx->E.block.exact_lim_s = lim_s; ctx->E.block.exact_lim_gamma = lim_gamma;
    if (int rc = rotate_block(ctx)) return rc;
    return upload_block(ctx);
}

int ssk_debug_math_device(int op, const float *a, const float *b, float *out, size_t n, void *stream) {
    if (!a || !out || op < 0 || op > 8 || ((op < 2 || op == 8) && !b) || (n & 1)) return fail(SSK_ERR_INVALID, "bad argument");
    if (n == 0) return SSK_OK;
    math_probe_kernel<<<1184, 256, 0, (cudaStream_t) stream>>>(op, a, b, out, n);
    g_launches++;
    SSK_CUDA(cudaGetLastError());
    return SSK_OK;
}

int ssk_host_copy_ceiling(int device, size_t h2d_bytes, size_t d2h_bytes, size_t piece_bytes, int reps, double *best_seconds,
                          double *mean_seconds) {
    if (!best_seconds || reps < 1) return fail(SSK_ERR_INVALID, "bad argument");
    DeviceGuard dg; if (int rc = dg.enter(device)) return rc;
    char *h_in = nullptr, *h_out = nullptr, *d_in = nullptr, *d_out = nullptr;
    cudaStream_t s0 = nullptr, s1 = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr;
    auto cleanup = [&] {
        if (h_in) cudaFreeHost(h_in); if (h_out) cudaFreeHost(h_out); if (d_in) cudaFree(d_in); if (d_out) cudaFree(d_out);
        if (s0) cudaStreamDestroy(s0); if (s1) cudaStreamDestroy(s1);
        if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); if (e2) cudaEventDestroy(e2);
    };
    struct Final { decltype(cleanup) &f; ~Final() { f(); } } fin{ cleanup };
    if (h2d_bytes) { SSK_CUDA(cudaMallocHost(&h_in, h2d_bytes)); SSK_CUDA(cudaMalloc(&d_in, h2d_bytes)); memset(h_in, 1, h2d_bytes); }
    if (d2h_bytes) { SSK_CUDA(cudaMallocHost(&h_out, d2h_bytes)); SSK_CUDA(cudaMalloc(&d_out, d2h_bytes)); SSK_CUDA(cudaMemset(d_out, 2, d2h_bytes)); }
    SSK_CUDA(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
    SSK_CUDA(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
    SSK_CUDA(cudaEventCreate(&e0)); SSK_CUDA(cudaEventCreate(&e1)); SSK_CUDA(cudaEventCreate(&e2));
    if (!piece_bytes) piece_bytes = std::max(h2d_bytes, d2h_bytes);
    double best = 1e30, total = 0.0;
    for (int r = 0; r < reps + 1; ++r) {                       // first pass is a warm-up
        SSK_CUDA(cudaDeviceSynchronize());
        auto t0 = std::chrono::steady_clock::now();
        for (size_t o = 0; o < h2d_bytes; o += piece_bytes)
            SSK_CUDA(cudaMemcpyAsync(d_in + o, h_in + o, std::min(piece_bytes, h2d_bytes - o), cudaMemcpyHostToDevice, s0));
        for (size_t o = 0; o < d2h_bytes; o += piece_bytes)
            SSK_CUDA(cudaMemcpyAsync(h_out + o, d_out + o, std::min(piece_bytes, d2h_bytes - o), cudaMemcpyDeviceToHost, s1));
        SSK_CUDA(cudaStreamSynchronize(s0)); SSK_CUDA(cudaStreamSynchronize(s1));
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (r > 0) { best = std::min(best, dt); total += dt; }
    }
    *best_seconds = best;
    if (mean_seconds) *mean_seconds = total / reps;
    return SSK_OK;
}

int ssk_envmap_directions_device(const ssk_ctx *ctx, int width, int height, float *const d[3], void *stream) {
    DeviceGuard dg; if (int rc = need_device(ctx, dg)) return rc;
    if (!d || !d[0] || !d[1] || !d[2]) return fail(SSK_ERR_INVALID, "null planes");
    envmap_dirs_kernel<<<ctx->sm_count * 4, 256, 0, (cudaStream_t) stream>>>(d[0], d[1], d[2], width, height);
    g_launches++;
    SSK_CUDA(cudaGetLastError());
    return SSK_OK;
}

int ssk_eval_host(const ssk_ctx *ctx, const float *const wi[3], const float *const wav[4], size_t n, float *const out[4]) {
    if (!ctx || !wi || !out) return fail(SSK_ERR_INVALID, "null argument");
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    if (spectral && !wav) return fail(SSK_ERR_INVALID, "spectral mode needs 4 wavelength planes");
    const float *in[7] = { wi[0], wi[1], wi[2], spectral ? wav[0] : nullptr, spectral ? wav[1] : nullptr,
                           spectral ? wav[2] : nullptr, spectral ? wav[3] : nullptr };
    const int n_in = spectral ? 7 : 3, n_out = spectral ? 4 : 3;
    for (int k = 0; k < n_in; ++k) if (!in[k]) return fail(SSK_ERR_INVALID, "null input plane");
    for (int k = 0; k < n_out; ++k) if (!out[k]) return fail(SSK_ERR_INVALID, "null output plane");
    return host_pipeline(ctx, in, n_in, out, n_out, n, [&](const float **din, float **dout, size_t m, cudaStream_t s) {
        return launch_eval(ctx, din, din + 3, m, dout, 1.f, s);
    });
}

int ssk_eval_direction_host(const ssk_ctx *ctx, const float *const d[3], const float *const wav[4], size_t n,
                            float *const out[4]) {
    if (!ctx || !d || !out) return fail(SSK_ERR_INVALID, "null argument");
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    if (spectral && !wav) return fail(SSK_ERR_INVALID, "spectral mode needs 4 wavelength planes");
    const float *in[7] = { d[0], d[1], d[2], spectral ? wav[0] : nullptr, spectral ? wav[1] : nullptr,
                           spectral ? wav[2] : nullptr, spectral ? wav[3] : nullptr };
    const int n_in = spectral ? 7 : 3, n_out = spectral ? 4 : 3;
    for (int k = 0; k < n_in; ++k) if (!in[k]) return fail(SSK_ERR_INVALID, "null input plane");
    for (int k = 0; k < n_out; ++k) if (!out[k]) return fail(SSK_ERR_INVALID, "null output plane");
    return host_pipeline(ctx, in, n_in, out, n_out, n, [&](const float **din, float **dout, size_t m, cudaStream_t s) {
        return launch_eval(ctx, din, din + 3, m, dout, -1.f, s);
    });
}

int ssk_pdf_direction_host(const ssk_ctx *ctx, const float *const d[3], size_t n, float *pdf) {
    if (!ctx || !d || !pdf || !d[0] || !d[1] || !d[2]) return fail(SSK_ERR_INVALID, "null argument");
    const float *in[3] = { d[0], d[1], d[2] };
    float *out[1] = { pdf };
    return host_pipeline(ctx, in, 3, out, 1, n, [&](const float **din, float **dout, size_t m, cudaStream_t s) {
        return launch_pdf(ctx, din, m, dout[0], s);
    });
}

int ssk_sample_direction_host(const ssk_ctx *ctx, const float *const u[2], const float *const it_p[3],
                              const float *const wav[4], size_t n, float *const d[3], float *pdf,
                              float *const weight[4], float *const p[3], float *dist) {
    if (!ctx || !u || !d || !pdf || !weight) return fail(SSK_ERR_INVALID, "null argument");
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    const int C = spectral ? 4 : 3;
    const bool has_p = it_p && it_p[0], want_p = p && p[0], want_dist = dist != nullptr;
    if (spectral && !wav) return fail(SSK_ERR_INVALID, "spectral mode needs 4 wavelength planes");
    const float *in[9]; int n_in = 0;
    in[n_in++] = u[0]; in[n_in++] = u[1];
    if (has_p) { in[n_in++] = it_p[0]; in[n_in++] = it_p[1]; in[n_in++] = it_p[2]; }
    if (spectral) for (int c = 0; c < 4; ++c) in[n_in++] = wav[c];
    float *out[12]; int n_out = 0;
    out[n_out++] = d[0]; out[n_out++] = d[1]; out[n_out++] = d[2]; out[n_out++] = pdf;
    for (int c = 0; c < C; ++c) out[n_out++] = weight[c];
    if (want_p) { out[n_out++] = p[0]; out[n_out++] = p[1]; out[n_out++] = p[2]; }
    if (want_dist) out[n_out++] = dist;
    for (int k = 0; k < n_in; ++k) if (!in[k]) return fail(SSK_ERR_INVALID, "null input plane");
    for (int k = 0; k < n_out; ++k) if (!out[k]) return fail(SSK_ERR_INVALID, "null output plane");
    return host_pipeline(ctx, in, n_in, out, n_out, n, [&](const float **din, float **dout, size_t m, cudaStream_t s) {
        int k = 2;
        const float *dp[3] = { nullptr, nullptr, nullptr };
        if (has_p) { dp[0] = din[k]; dp[1] = din[k + 1]; dp[2] = din[k + 2]; k += 3; }
        const float *dw[4] = { nullptr, nullptr, nullptr, nullptr };
        if (spectral) for (int c = 0; c < 4; ++c) dw[c] = din[k++];
        int o = 4;
        float *wq[4] = { nullptr, nullptr, nullptr, nullptr };
        for (int c = 0; c < C; ++c) wq[c] = dout[o++];
        float *op[3] = { nullptr, nullptr, nullptr };
        if (want_p) { op[0] = dout[o]; op[1] = dout[o + 1]; op[2] = dout[o + 2]; o += 3; }
        float *od = want_dist ? dout[o++] : nullptr;
        return launch_sample(ctx, din, has_p ? dp : nullptr, spectral ? dw : nullptr, m, dout, dout[3], wq,
                             want_p ? op : nullptr, od, s);
    });
}

int ssk_eval_jvp_device(const ssk_ctx *ctx, float d_turbidity, const float *d_albedo, int n_d_albedo, const float *d_sun_direction,
                        const float *const wi[3], const float *const wav[4], size_t n, float *const out[4], void *stream) {
    DeviceGuard dg; if (int rc = need_device(ctx, dg)) return rc;
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    const int C = spectral ? 4 : 3;
    if (n_d_albedo != 0 && n_d_albedo != 1 && n_d_albedo != ctx->E.n_channels)
        return fail(SSK_ERR_INVALID, "d_albedo needs 0, 1 or n_channels (3 RGB / 11 spectral) entries");
    if (n_d_albedo && !d_albedo) return fail(SSK_ERR_INVALID, "null d_albedo");
    if (d_sun_direction && ctx->E.active_record)
        return fail(SSK_ERR_INVALID, "sun_direction is not a parameter of an emitter built from a date / location record");
    if (n == 0) return SSK_OK;                     // empty batches carry null planes
    if (!wi || !out || !wi[0] || !wi[1] || !wi[2]) return fail(SSK_ERR_INVALID, "null direction / output planes");
    for (int c = 0; c < C; ++c) if (!out[c]) return fail(SSK_ERR_INVALID, "null output plane");
    if (spectral && (!wav || !wav[0] || !wav[1] || !wav[2] || !wav[3]))
        return fail(SSK_ERR_INVALID, "spectral mode needs 4 wavelength planes");
    return guarded([&]() -> int {
        cudaStream_t s = (cudaStream_t) stream;
        // the tangent block is a kernel parameter: built here, copied into the launch, nothing shared between calls
        static thread_local TangentBlock T;
        state_tangent(ctx->E, d_turbidity, d_albedo, n_d_albedo, d_sun_direction, T);
        Planes3 i3 = { wi[0], wi[1], wi[2] };
        Planes4 w4 = { nullptr, nullptr, nullptr, nullptr };
        if (spectral) w4 = { wav[0], wav[1], wav[2], wav[3] };
        OutPlanes4 o4 = { out[0], out[1], out[2], spectral ? out[3] : nullptr };
        int grid = 1;
        auto launch = [&](auto kernel) -> int {
            if (int rc = occupancy_grid(ctx, kernel, n, &grid, 1)) return rc;
            kernel<<<grid, kThreads, 0, s>>>(hot(ctx), T, ctx->d_block, i3, w4, o4, n);
            return SSK_OK;
        };
        int rc = fast(ctx) ? (spectral ? launch(eval_jvp_kernel<kFast, true>) : launch(eval_jvp_kernel<kFast, false>))
                           : (spectral ? launch(eval_jvp_kernel<kExact, true>) : launch(eval_jvp_kernel<kExact, false>));
        if (rc) return rc;
        return launched(ctx, s);
    });
}

int ssk_sample_wavelengths_device(const ssk_ctx *ctx, const float *const wi[3], const float *sample, size_t n,
                                  float *const wav[4], float *const weight[4], void *stream) {
    return launch_sample_wavelengths(ctx, wi, sample, n, wav, weight, (cudaStream_t) stream);
}
int ssk_sample_ray_device(const ssk_ctx *ctx, const float *wl_sample, const float *const s2[2], const float *const s3[2],
                          size_t n, float *const o[3], float *const d[3], float *const wav[4], float *const weight[4],
                          void *stream) {
    return launch_sample_ray(ctx, wl_sample, s2, s3, n, o, d, wav, weight, (cudaStream_t) stream);
}
int ssk_sample_wavelengths_host(const ssk_ctx *ctx, const float *const wi[3], const float *sample, size_t n,
                                float *const wav[4], float *const weight[4]) {
    if (!ctx || !wi || !weight) return fail(SSK_ERR_INVALID, "null argument");
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    if (!spectral) return ssk_eval_host(ctx, wi, nullptr, n, weight);
    if (!sample || !wav) return fail(SSK_ERR_INVALID, "spectral mode needs the sample plane and 4 wavelength output planes");
    const float *in[4] = { wi[0], wi[1], wi[2], sample };
    float *out[8] = { wav[0], wav[1], wav[2], wav[3], weight[0], weight[1], weight[2], weight[3] };
    for (int k = 0; k < 4; ++k) if (!in[k]) return fail(SSK_ERR_INVALID, "null input plane");
    for (int k = 0; k < 8; ++k) if (!out[k]) return fail(SSK_ERR_INVALID, "null output plane");
    return host_pipeline(ctx, in, 4, out, 8, n, [&](const float **din, float **dout, size_t m, cudaStream_t s) {
        return launch_sample_wavelengths(ctx, din, din[3], m, dout, dout + 4, s);
    });
}
int ssk_sample_ray_host(const ssk_ctx *ctx, const float *wl_sample, const float *const s2[2], const float *const s3[2],
                        size_t n, float *const o[3], float *const d[3], float *const wav[4], float *const weight[4]) {
    if (!ctx || !s2 || !s3 || !o || !d || !weight) return fail(SSK_ERR_INVALID, "null argument");
    const bool spectral = ctx->mode == SSK_MODE_SPECTRAL;
    const int C = spectral ? 4 : 3;
    if (spectral && (!wl_sample || !wav)) return fail(SSK_ERR_INVALID, "spectral mode needs the wavelength sample and output planes");
    const float *in[5]; int n_in = 0;
    in[n_in++] = s2[0]; in[n_in++] = s2[1]; in[n_in++] = s3[0]; in[n_in++] = s3[1];
    if (spectral) in[n_in++] = wl_sample;
    float *out[14]; int n_out = 0;
    for (int k = 0; k < 3; ++k) out[n_out++] = o[k];
    for (int k = 0; k < 3; ++k) out[n_out++] = d[k];
    if (spectral) for (int k = 0; k < 4; ++k) out[n_out++] = wav[k];
    for (int k = 0; k < C; ++k) out[n_out++] = weight[k];
    for (int k = 0; k < n_in; ++k) if (!in[k]) return fail(SSK_ERR_INVALID, "null input plane");
    for (int k = 0; k < n_out; ++k) if (!out[k]) return fail(SSK_ERR_INVALID, "null output plane");
    return host_pipeline(ctx, in, n_in, out, n_out, n, [&](const float **din, float **dout, size_t m, cudaStream_t s) {
        return launch_sample_ray(ctx, spectral ? din[4] : nullptr, din, din + 2, m, dout, dout + 3,
                                 spectral ? dout + 6 : nullptr, dout + (spectral ? 10 : 6), s);
    });
}

int ssk_sun_coordinates(int year, int month, int day, float hour, float minute, float second, float latitude,
                        float longitude, float timezone, float out_dir[3]) {
    if (!out_dir) return fail(SSK_ERR_INVALID, "null argument");
    DateTime dt = { year, month, day, hour, minute, second };
    Location loc = { latitude, longitude, timezone };
    Vec3 v = sun_coordinates(dt, loc);
    out_dir[0] = v.x; out_dir[1] = v.y; out_dir[2] = v.z;
    return SSK_OK;
}

/// fp64 payload of a dataset file (the files are fp64 on disk; the plugin itself keeps them in fp32)
static std::vector<double> read_dataset_f64(const std::string &path, size_t expect) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("\"" + path + "\": file does not exist!");
    char tag[4] = { 0, 0, 0, 0 }; uint32_t version; uint64_t ndim = 0, count = 1;
    bool ok = fread(tag, 1, 3, f) == 3 && fread(&version, 4, 1, f) == 1 && fread(&ndim, 8, 1, f) == 1 && ndim > 0 && ndim <= 8;
    for (uint64_t i = 0; ok && i < ndim; ++i) { uint64_t d; ok = fread(&d, 8, 1, f) == 1; count *= d; }
    std::vector<double> v(ok && count == expect ? count : 0);
    ok = ok && count == expect && fread(v.data(), 8, count, f) == count;
    fclose(f);
    if (!ok) throw std::runtime_error(path + ": truncated or unexpected dataset file");
    return v;
}

int ssk_hosek_sun_rad(const char *dataset_dir, double turbidity, double wavelength, double elevation, double gamma, double *out) {
    if (!dataset_dir || !out) return fail(SSK_ERR_INVALID, "null argument");
    return guarded([&]() -> int {
        // arhosekskymodel_solar_radiance_internal2 (ArHosekSkyModel.c:686-786) on the shipped tables:
        // sun_spec_rad.bin = solarDatasets as [turbidity][segment][lambda][power of x], sun_spec_ld.bin = limbDarkeningDatasets
        static std::mutex mu;
        static std::string dir;
        static std::vector<double> rad, ld;
        std::lock_guard<std::mutex> lock(mu);
        if (dir != dataset_dir || rad.empty()) {
            std::string d = dataset_dir; if (!d.empty() && d.back() != '/') d += '/';
            rad = read_dataset_f64(d + "sun_spec_rad.bin", 10 * 45 * 11 * 4);
            ld = read_dataset_f64(d + "sun_spec_ld.bin", 66);
            dir = dataset_dir;
        }
        *out = 0.0;
        if (!(wavelength >= 320.0 && wavelength <= 720.0 && turbidity >= 1.0 && turbidity <= 10.0)) return SSK_OK;
        int turb_low = (int) turbidity - 1;
        double turb_frac = turbidity - (double) (turb_low + 1);
        if (turb_low == 9) { turb_low = 8; turb_frac = 1.0; }
        int wl_low = (int) ((wavelength - 320.0) / 40.0);
        double wl_frac = fmod(wavelength, 40.0) / 40.0;
        if (wl_low == 10) { wl_low = 9; wl_frac = 1.0; }
        auto sr = [&](int t, int wl) {                       // arhosekskymodel_sr_internal (:656-684)
            const double pi = 3.14159265358979323846;
            int pos = (int) (pow(2.0 * elevation / pi, 1.0 / 3.0) * 45);
            if (pos > 44) pos = 44;
            const double break_x = pow((double) pos / 45.0, 3.0) * (pi * 0.5), x = elevation - break_x;
            const double *c = &rad[(((size_t) t * 45 + pos) * 11 + wl) * 4];
            double res = 0.0, x_exp = 1.0;
            for (int i = 0; i < 4; ++i) { res += x_exp * c[i]; x_exp *= x; }
            return res;
        };
        double direct = (1.0 - turb_frac) * ((1.0 - wl_frac) * sr(turb_low, wl_low) + wl_frac * sr(turb_low, wl_low + 1)) +
                        turb_frac * ((1.0 - wl_frac) * sr(turb_low + 1, wl_low) + wl_frac * sr(turb_low + 1, wl_low + 1));
        double c[6];
        for (int i = 0; i < 6; ++i) c[i] = (1.0 - wl_frac) * ld[wl_low * 6 + i] + wl_frac * ld[(wl_low + 1) * 6 + i];
        const double sol = sin(0.5358 / 2.0 * (3.14159265358979323846 / 180.0)), ar2 = 1 / (sol * sol), sg = sin(gamma);
        double sc2 = 1.0 - ar2 * sg * sg;
        if (sc2 < 0.0) sc2 = 0.0;
        const double sc = sqrt(sc2);
        *out = direct * (c[0] + c[1] * sc + c[2] * pow(sc, 2.0) + c[3] * pow(sc, 3.0) + c[4] * pow(sc, 4.0) + c[5] * pow(sc, 5.0));
        return SSK_OK;
    });
}

int ssk_normalize_direction(const float in[3], float out[3]) {
    if (!in || !out) return fail(SSK_ERR_INVALID, "null argument");
    Vec3 v = normalize_sun_direction({ in[0], in[1], in[2] });
    out[0] = v.x; out[1] = v.y; out[2] = v.z;
    return SSK_OK;
}

int ssk_array_from_file(const char *path, int file_is_f64, float *out, size_t capacity, size_t *count,
                        size_t *shape, size_t *ndim) {
    if (!path) return fail(SSK_ERR_INVALID, "null path");
    return guarded([&]() -> int {
        std::vector<size_t> sh;
        std::vector<float> v = read_dataset(path, file_is_f64 != 0, &sh);
        if (count) *count = v.size();
        if (ndim) *ndim = sh.size();
        if (shape) for (size_t i = 0; i < sh.size() && i < 8; ++i) shape[i] = sh[i];
        if (out) {
            if (capacity < v.size()) return fail(SSK_ERR_INVALID, "output buffer too small");
            memcpy(out, v.data(), v.size() * sizeof(float));
        }
        return SSK_OK;
    });
}

int ssk_array_to_file(const char *path, const float *data, const size_t *shape, size_t ndim, int as_f64) {
    if (!path || !data || !shape || ndim == 0) return fail(SSK_ERR_INVALID, "null argument");
    FILE *f = fopen(path, "wb");
    if (!f) return fail(SSK_ERR_IO, std::string("cannot open ") + path);
    size_t count = 1;
    uint32_t version = 0; uint64_t nd = ndim;
    fwrite("SKY", 1, 3, f); fwrite(&version, 4, 1, f); fwrite(&nd, 8, 1, f);
    for (size_t i = 0; i < ndim; ++i) {
        if (!shape[i]) { fclose(f); return fail(SSK_ERR_INVALID, "Got dimension with 0 elements"); }
        uint64_t d = shape[i]; fwrite(&d, 8, 1, f); count *= shape[i];
    }
    if (as_f64) { for (size_t i = 0; i < count; ++i) { double v = data[i]; fwrite(&v, 8, 1, f); } }
    else fwrite(data, 4, count, f);
    fclose(f);
    return SSK_OK;
}

} // extern "C"
