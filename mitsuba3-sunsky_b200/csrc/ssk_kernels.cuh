// sm_100a kernels of the sunsky hot path. One kernel per Emitter entry point, each a single fused
// pass: planar fp32 in, planar fp32 out, 128-bit loads/stores (4 consecutive rays per thread; the
// many-plane sampling kernel moves one ray per thread to stay within 2 resident CTAs' registers), a
// persistent grid sized to a multiple of the SM count.
//
// The per-emitter coefficient block (HotBlock, 3.4 KB) is a __grid_constant__ kernel parameter: every
// coefficient read with a compile-time index (all of RGB eval and of the 20-Gaussian density) becomes a
// constant-bank operand of the FFMA/FMUL that uses it, with no load instruction at all. Only reads with a
// lane-dependent index (spectral bins, the sampled mixture component) go through a shared-memory copy.
// The work is FP32/MUFU-issue bound (hundreds of flops per 16-44 bytes), so there is no TMA /
// tensor-core path here by design.
#pragma once
#include "ssk_model.cuh"
#include <cuda_runtime.h>

namespace ssk {

#ifndef SSK_THREADS
#define SSK_THREADS 256
#endif
constexpr int kThreads = SSK_THREADS;

// Resident CTAs per SM each kernel family is compiled for (the register cap of __launch_bounds__); overridable
// for tuning builds (-DSSK_..._MINB=n).
#ifndef SSK_SAMPLE_MINB
#define SSK_SAMPLE_MINB 4      /* resident CTAs per SM the sampling kernels are compiled for (register cap); */
#endif                         /* measured on B200, sky-only RGB: 2 -> 30.5, 3 -> 33.5, 4 -> 35.3 Gsamples/s */
#ifndef SSK_SAMPLE_MINB_SPEC
#define SSK_SAMPLE_MINB_SPEC 3
#endif
#ifndef SSK_RGBP_MINB
#define SSK_RGBP_MINB 3
#endif
#ifndef SSK_SPECF_MINB
#define SSK_SPECF_MINB 3
#endif
#ifndef SSK_ENVMAP_SPEC_MINB
#define SSK_ENVMAP_SPEC_MINB 2     /* measured, spectral 4096x2048 map: FAST 1 -> 53.7, 2 -> 66.9, 3 -> 64.4 Gpixels/s; EXACT 30.7 / 38.4 / 39.2 */
#endif
#ifndef SSK_ENVMAP_MINB
#define SSK_ENVMAP_MINB 3
#endif
#ifndef SSK_PAIR_PREFETCH
#define SSK_PAIR_PREFETCH 1
#endif
#ifndef SSK_EVAL_PREFETCH
#define SSK_EVAL_PREFETCH 1      /* FAST spectral eval: measured 54.5 (pipelined loads, 3 CTAs/SM, 85 regs) vs 50.5 Gevals/s; */
                                 /* with 4 CTAs/SM (64 regs) the prefetched values spill and the gain is lost; a non-binding       */
                                 /* prefetch.global.L1 of the next iteration instead measured no gain (profiles/r1_summary.md)  */
#endif
#ifndef SSK_SPECX_MINB
#define SSK_SPECX_MINB 1
#endif
// EXACT spectral eval, one ray per thread. Measured on B200 (2^24 rays): no prefetch, 3 CTAs / SM (80 registers) 28.8
// Gevals/s with 19 % of the stall samples on the first use of the loaded inputs; software-pipelined loads at 3 CTAs
// (80 registers, prefetch registers spilled) 28.4; pipelined loads at 2 CTAs (96 registers, nothing spilled) 30.3;
// inputs through the TMA ring (SSK_EVAL_TMA=1) 27.6. Without the pipelined loads 2 CTAs are 5 % slower than 3.
#ifndef SSK_EVAL_PREFETCH_EXACT
#define SSK_EVAL_PREFETCH_EXACT 1
#endif
#ifndef SSK_SPECX1_MINB
#define SSK_SPECX1_MINB 2
#endif
#ifndef SSK_SPECT_MINB
#define SSK_SPECT_MINB 3           /* TMA-staged FAST spectral eval: resident CTAs per SM (4 = 64 registers, 8 bytes of spill) */
#endif

/// Copy the emitter block from global to shared memory (16-byte accesses) and, for FAST spectral
/// kernels, build the bin-pair table of eval_spec (build_pair_table's layout). Ends with a barrier.
__device__ __forceinline__ void stage_block(HotBlock *dst, float2 *pairs, const EmitterBlock *src, bool exact_pairs = false) {
    const uint4 *s = reinterpret_cast<const uint4 *>(src);
    uint4 *d = reinterpret_cast<uint4 *>(dst);
    for (int i = threadIdx.x; i < (int) (sizeof(HotBlock) / 16); i += blockDim.x) d[i] = __ldg(s + i);
    if (pairs && exact_pairs) {                    // EXACT spectral kernels: build_exact_pair_table's layout
        for (int i = threadIdx.x; i < kExactPairTable; i += blockDim.x) {
            int c = i % kChanPad, k = i / kChanPad;
            const float *row = k < kSkyParams ? &src->sky_params[k][0] : &src->sky_rad[0];
            pairs[i] = make_float2(__ldg(row + c), c + 1 < kChanPad ? __ldg(row + c + 1) : 0.f);
        }
    } else if (pairs)
        for (int i = threadIdx.x; i < kPairTable; i += blockDim.x) {
            int c = i % kChanPad, k = i / kChanPad;
            const float *row = &src->sky_fast[k][0];
            pairs[pair_index(k, c)] = make_float2(__ldg(row + c), c + 1 < kChanPad ? __ldg(row + c + 1) : 0.f);
        }
    __syncthreads();
}

/// R rays per thread: R = 4 moves 4 consecutive rays with one 128-bit access per plane (planes must be
/// 16-byte aligned), R = 1 moves one ray with a 32-bit access (a warp still covers one full 128-byte line).
/// Which one a kernel uses is a register-pressure trade: every plane costs R staging registers per thread.
template <int R> __device__ __forceinline__ void loadr(const float *p, size_t i, size_t n, float v[4]) {
    if (R == 4) {
        if (i + 3 < n) {
            float4 t = __ldcs(reinterpret_cast<const float4 *>(p + i));      // streamed once: evict first
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = (i + k < n) ? __ldg(p + i + k) : 0.f;
        }
    } else {
        v[0] = __ldcs(p + i);
    }
}
template <int R> __device__ __forceinline__ void storer(float *p, size_t i, size_t n, const float v[4]) {
    if (R == 4) {
        if (i + 3 < n) {
            __stcs(reinterpret_cast<float4 *>(p + i), make_float4(v[0], v[1], v[2], v[3]));
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) if (i + k < n) p[i + k] = v[k];
        }
    } else {
        __stcs(p + i, v[0]);
    }
}

/// Element k of a 4-vector held in registers, k not known at compile time (3 selects, no local memory).
/// The per-ray body is deliberately NOT unrolled over the rays of a thread: unrolled, the kernels grow to
/// 14-30 k instructions and stall on instruction fetch (ncu: `no_instructions` was the top stall reason).
template <int R> __device__ __forceinline__ float getk(const float v[4], int k) {
    if (R == 1) return v[0];
    float a = k & 1 ? v[1] : v[0], b = k & 1 ? v[3] : v[2];
    return k & 2 ? b : a;
}
template <int R> __device__ __forceinline__ void setk(float v[4], int k, float x) {
    if (R == 1) { v[0] = x; return; }
    v[0] = k == 0 ? x : v[0]; v[1] = k == 1 ? x : v[1]; v[2] = k == 2 ? x : v[2]; v[3] = k == 3 ? x : v[3];
}

struct Planes3 { const float *x, *y, *z; };
struct Planes4 { const float *a, *b, *c, *d; };
struct OutPlanes3 { float *x, *y, *z; };
struct OutPlanes4 { float *a, *b, *c, *d; };

/// eval / eval_direction, RGB. dir_sign = +1: input is si.wi (the model negates it);
/// dir_sign = -1: input is ds.d (eval_direction passes wi = -d).
template <int A, int R>
__global__ void __launch_bounds__(kThreads, R == 1 ? 4 : 3)
eval_rgb_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, Planes3 in, OutPlanes3 out,
                size_t n, float dir_sign) {
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x * R;
    for (size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * R; i < n; i += stride) {
        float x[4], y[4], z[4], r[4], g[4], b[4];
        loadr<R>(in.x, i, n, x); loadr<R>(in.y, i, n, y); loadr<R>(in.z, i, n, z);
#pragma unroll 1
        for (int k = 0; k < R; ++k) {
            Vec3 wi = { dir_sign * getk<R>(x, k), dir_sign * getk<R>(y, k), dir_sign * getk<R>(z, k) };
            float o[3];
            eval_rgb<A>(U, sun_rad, wi, true, o);
            setk<R>(r, k, o[0]); setk<R>(g, k, o[1]); setk<R>(b, k, o[2]);
        }
        storer<R>(out.x, i, n, r); storer<R>(out.y, i, n, g); storer<R>(out.z, i, n, b);
    }
}

// ---- two rays per thread: 64-bit plane accesses (planes 8-byte aligned) ------------------------------------
__device__ __forceinline__ float2 ld2(const float *p, size_t i, bool two, float pad) {
    return two ? __ldcs(reinterpret_cast<const float2 *>(p + i)) : make_float2(__ldcs(p + i), pad);
}
__device__ __forceinline__ void st2(float *p, size_t i, bool two, float2 v) {
    if (two) __stcs(reinterpret_cast<float2 *>(p + i), v); else __stcs(p + i, v.x);
}
/// Three direction planes of ray pair i (pads 0, 0, 1 for the missing second ray of an odd tail)
__device__ __forceinline__ Vec3p ld_dir2(const Planes3 &in, size_t i, size_t n) {
    const bool two = i + 1 < n;
    return { ld2(in.x, i, two, 0.f), ld2(in.y, i, two, 0.f), ld2(in.z, i, two, 1.f) };
}

/// eval / eval_direction, RGB, two consecutive rays per thread as float2 pairs (64-bit plane accesses, planes
/// 8-byte aligned): every FP32 instruction of the model works on both rays (FFMA2 / FMUL2 / FADD2).
template <int A>
__global__ void __launch_bounds__(kThreads, SSK_RGBP_MINB)
eval_rgb_pair_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, Planes3 in, OutPlanes3 out,
                     size_t n, float dir_sign) {
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x * 2;
    size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * 2;
    // software-pipelined: the next grid-stride iteration's directions are requested before this pair's arithmetic
    Vec3p next = { f2(0.f, 0.f), f2(0.f, 0.f), f2(1.f, 1.f) };
    if (i < n) next = ld_dir2(in, i, n);
    for (; i < n; i += stride) {
        const bool two = i + 1 < n;
        Vec3p wi = { mul2s(next.x, dir_sign), mul2s(next.y, dir_sign), mul2s(next.z, dir_sign) };
        if (SSK_PAIR_PREFETCH && i + stride < n) next = ld_dir2(in, i + stride, n);
        float2 o[3];
        eval_rgb_p<A>(U, sun_rad, wi, true, two, o);
        st2(out.x, i, two, o[0]); st2(out.y, i, two, o[1]); st2(out.z, i, two, o[2]);
        if (!SSK_PAIR_PREFETCH && i + stride < n) next = ld_dir2(in, i + stride, n);
    }
}

/// Ray loop of eval_spec_kernel, one ray per thread, software-pipelined: the 7 input planes of the next grid-stride
/// iteration are requested before this ray's arithmetic starts (ncu: long_scoreboard was the top stall without it).
/// The prefetch index is clamped to the last ray instead of guarded by a branch (the tail re-reads ray n - 1).
template <int A, typename I>
__device__ __forceinline__ void eval_spec_pipelined(const HotBlock &U, const HotBlock &S, const float2 *pairs, const float *sun_rad,
                                                    const Planes3 &in, const Planes4 &wav, const OutPlanes4 &out, I n, float dir_sign) {
    const I stride = (I) gridDim.x * (I) blockDim.x, last = n - 1;
    I i = (I) blockIdx.x * (I) blockDim.x + (I) threadIdx.x;
    if (i >= n) return;
    float nx = __ldcs(in.x + i), ny = __ldcs(in.y + i), nz = __ldcs(in.z + i);
    float na = __ldcs(wav.a + i), nb = __ldcs(wav.b + i), nc = __ldcs(wav.c + i), nd = __ldcs(wav.d + i);
    for (;;) {
        Vec3 wi = { dir_sign * nx, dir_sign * ny, dir_sign * nz };
        float wl[4] = { na, nb, nc, nd }, o[4];
        const I j = i + stride, jc = j < last ? j : last;
        nx = __ldcs(in.x + jc); ny = __ldcs(in.y + jc); nz = __ldcs(in.z + jc);
        na = __ldcs(wav.a + jc); nb = __ldcs(wav.b + jc); nc = __ldcs(wav.c + jc); nd = __ldcs(wav.d + jc);
        eval_spec<A, A == kExact>(U, S, pairs, sun_rad, wi, wl, true, o);
        __stcs(out.a + i, o[0]); __stcs(out.b + i, o[1]); __stcs(out.c + i, o[2]); __stcs(out.d + i, o[3]);
        if (j >= n) break;
        i = j;
    }
}

/// eval / eval_direction, spectral (4 wavelengths per ray)
template <int A, int R>
__global__ void __launch_bounds__(kThreads, (A == kFast && R == 1) ? SSK_SPECF_MINB : (R == 1 ? SSK_SPECX1_MINB : SSK_SPECX_MINB))
eval_spec_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, Planes3 in, Planes4 wav,
                 OutPlanes4 out, size_t n, float dir_sign) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[A == kFast ? kPairTable : kExactPairTable];
    stage_block(&S, pairs, blk, A == kExact);
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x * R;
    if (R == 1 && (A == kFast || SSK_EVAL_PREFETCH_EXACT == 1) && SSK_EVAL_PREFETCH == 1) {
        // one ray per thread, software-pipelined (eval_spec_pipelined); 32-bit ray indices whenever they fit: one
        // IMAD.WIDE per plane address instead of a 64-bit add pair
        if (n <= 0x7fffffffu) eval_spec_pipelined<A, uint32_t>(U, S, pairs, sun_rad, in, wav, out, (uint32_t) n, dir_sign);
        else eval_spec_pipelined<A, size_t>(U, S, pairs, sun_rad, in, wav, out, n, dir_sign);
        return;
    }
    for (size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * R; i < n; i += stride) {
        float x[4], y[4], z[4], w0[4], w1[4], w2[4], w3[4], o0[4], o1[4], o2[4], o3[4];
        loadr<R>(in.x, i, n, x); loadr<R>(in.y, i, n, y); loadr<R>(in.z, i, n, z);
        loadr<R>(wav.a, i, n, w0); loadr<R>(wav.b, i, n, w1); loadr<R>(wav.c, i, n, w2); loadr<R>(wav.d, i, n, w3);
#pragma unroll 1
        for (int k = 0; k < R; ++k) {
            Vec3 wi = { dir_sign * getk<R>(x, k), dir_sign * getk<R>(y, k), dir_sign * getk<R>(z, k) };
            float wl[4] = { getk<R>(w0, k), getk<R>(w1, k), getk<R>(w2, k), getk<R>(w3, k) }, o[4];
            eval_spec<A, A == kExact>(U, S, pairs, sun_rad, wi, wl, true, o);
            setk<R>(o0, k, o[0]); setk<R>(o1, k, o[1]); setk<R>(o2, k, o[2]); setk<R>(o3, k, o[3]);
        }
        storer<R>(out.a, i, n, o0); storer<R>(out.b, i, n, o1); storer<R>(out.c, i, n, o2); storer<R>(out.d, i, n, o3);
    }
}

// ---- TMA-staged input tiles ------------------------------------------------------------------------------------
// Round-2 profile of eval_spec_kernel (profiles/r2_summary.md): the register-held software prefetch of the 7 input
// planes spilled (80 registers at 3 CTAs / SM) and the spill store of a just-requested value waits for the load --
// 34 % of all stall samples sat on one `STL` with long_scoreboard, and 28 IADD3 per ray went into 64-bit plane
// addressing. Here the planes of a 256-ray tile are brought in by the TMA instead: one thread issues one 1 KB
// cp.async.bulk per plane into a 3-stage shared-memory ring, completion is signalled on an mbarrier, the consumers
// read their element with a fixed-offset LDS and release the stage with an mbarrier arrive. No registers are tied
// up by data in flight, nothing stalls on a load, and the per-thread global address arithmetic of the inputs is gone.
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "SSK_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra SSK_DONE;\n"
                 "bra SSK_WAIT;\n"
                 "SSK_DONE:\n"
                 "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
/// 1-D bulk copy global -> shared (16-byte aligned on both sides, size a multiple of 16), completion on `bar`
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

constexpr int kTmaStages = 3;
constexpr int kWarps = kThreads / 32;
/// Ring of input tiles: P planes x kThreads floats per stage, one `full` mbarrier per stage. A stage is refilled by
/// whichever warp releases it LAST (a shared-memory counter, one atomic per warp and tile): nobody ever waits for a
/// stage to become free, and the ~100 instructions of issuing P bulk copies are paid by one warp in eight per tile.
/// (Measured alternatives, profiles/r2_summary.md: thread 0 as fixed producer behind an `empty` mbarrier made every
/// consumer spin whenever warp 0 ran late; one pipeline per warp cost +48 % instructions in copy issue.)
template <int P> struct TileRing {
    float data[kTmaStages][P][kThreads];
    uint64_t full[kTmaStages];
    int released[kTmaStages];
};
/// Issued by ONE thread: planes src[p] + tile * kThreads -> stage `s`
template <int P> __device__ __forceinline__ void tile_issue(TileRing<P> &ring, int s, const float *const (&src)[P], size_t tile) {
    mbar_expect_tx(&ring.full[s], P * kThreads * (uint32_t) sizeof(float));
#pragma unroll
    for (int p = 0; p < P; ++p)
        tma_load_1d(&ring.data[s][p][0], src[p] + tile * kThreads, kThreads * (uint32_t) sizeof(float), &ring.full[s]);
}
template <int P> __device__ __forceinline__ void ring_init(TileRing<P> &ring) {
    if (threadIdx.x == 0) {
        for (int s = 0; s < kTmaStages; ++s) { mbar_init(&ring.full[s], 1); ring.released[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
}
/// Called by every thread after it has read its element of stage `s`: the last warp to get here refills the stage
/// with tile `next` (if there is one)
template <int P> __device__ __forceinline__ void ring_release(TileRing<P> &ring, int s, const float *const (&src)[P], size_t next, size_t n_tiles) {
    __syncwarp();
    if ((threadIdx.x & 31) == 0) {
        __threadfence_block();
        if (atomicAdd(&ring.released[s], 1) == kWarps - 1) {
            ring.released[s] = 0;
            __threadfence_block();
            if (next < n_tiles) tile_issue(ring, s, src, next);
        }
    }
}

/// eval / eval_direction, spectral, one ray per thread, inputs through the TMA ring. Planes must be 16-byte
/// aligned (the launch wrapper checks); the rays past the last full 256-ray tile are read with plain loads.
template <int A>
__global__ void __launch_bounds__(kThreads, A == kFast ? SSK_SPECT_MINB : 2)
eval_spec_tma_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, Planes3 in, Planes4 wav,
                     OutPlanes4 out, size_t n, float dir_sign) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[A == kFast ? kPairTable : 1];
    __shared__ __align__(128) TileRing<7> ring;
    const float *const src[7] = { in.x, in.y, in.z, wav.a, wav.b, wav.c, wav.d };
    const size_t n_tiles = n / kThreads;
    ring_init(ring);
    stage_block(&S, A == kFast ? pairs : nullptr, blk);          // ends with __syncthreads(): barriers are initialised
    const float *sun_rad = blk->sun_rad;
    if (threadIdx.x == 0)
        for (int j = 0; j < kTmaStages; ++j) {
            size_t t = (size_t) blockIdx.x + (size_t) j * gridDim.x;
            if (t < n_tiles) tile_issue(ring, j, src, t);
        }
    uint32_t k = 0;
    for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
        const int s = (int) (k % kTmaStages);
        mbar_wait(&ring.full[s], (k / kTmaStages) & 1u);
        const float (*d)[kThreads] = ring.data[s];
        Vec3 wi = { dir_sign * d[0][threadIdx.x], dir_sign * d[1][threadIdx.x], dir_sign * d[2][threadIdx.x] };
        float wl[4] = { d[3][threadIdx.x], d[4][threadIdx.x], d[5][threadIdx.x], d[6][threadIdx.x] }, o[4];
        ring_release(ring, s, src, t + (size_t) kTmaStages * gridDim.x, n_tiles);
        eval_spec<A>(U, S, pairs, sun_rad, wi, wl, true, o);
        const size_t i = t * kThreads + threadIdx.x;
        __stcs(out.a + i, o[0]); __stcs(out.b + i, o[1]); __stcs(out.c + i, o[2]); __stcs(out.d + i, o[3]);
    }
    // ragged tail (n % 256 rays): the CTA whose turn the next tile would be
    const size_t i = n_tiles * kThreads + threadIdx.x;
    if (blockIdx.x == n_tiles % gridDim.x && i < n) {
        Vec3 wi = { dir_sign * __ldcs(in.x + i), dir_sign * __ldcs(in.y + i), dir_sign * __ldcs(in.z + i) };
        float wl[4] = { __ldcs(wav.a + i), __ldcs(wav.b + i), __ldcs(wav.c + i), __ldcs(wav.d + i) }, o[4];
        eval_spec<A>(U, S, pairs, sun_rad, wi, wl, true, o);
        __stcs(out.a + i, o[0]); __stcs(out.b + i, o[1]); __stcs(out.c + i, o[2]); __stcs(out.d + i, o[3]);
    }
}

/// pdf_direction
template <int A, int R>
__global__ void __launch_bounds__(kThreads)
pdf_kernel(const __grid_constant__ HotBlock U, Planes3 in, float *__restrict__ pdf, size_t n) {
    const size_t stride = (size_t) gridDim.x * blockDim.x * R;
    for (size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * R; i < n; i += stride) {
        float x[4], y[4], z[4], o[4];
        loadr<R>(in.x, i, n, x); loadr<R>(in.y, i, n, y); loadr<R>(in.z, i, n, z);
#pragma unroll 1
        for (int k = 0; k < R; ++k) {
            Vec3 d = { getk<R>(x, k), getk<R>(y, k), getk<R>(z, k) };
            setk<R>(o, k, pdf_direction<A>(U, d));
        }
        storer<R>(pdf, i, n, o);
    }
}

struct SampleArgs {
    const float *u0, *u1;
    const float *px, *py, *pz;          // it.p, nullable
    const float *w0, *w1, *w2, *w3;     // it.wavelengths (spectral)
    float *dx, *dy, *dz, *pdf;
    float *q0, *q1, *q2, *q3;           // weight planes
    float *ox, *oy, *oz, *dist;         // ds.p / ds.dist, nullable (sample_ray through sample_sorted_kernel<.., Ray>: ray origin)
    const float *s2x, *s2y;             // sample_ray only: the origin-disc sample
    const float *wl_sample;             // spectral sample_ray only: the wavelength sample and the sampled wavelengths
    float *ow0, *ow1, *ow2, *ow3;
};

/// sample_direction, fused: select technique -> sample -> pdf -> eval -> weight (sunsky.cpp:399-441)
template <int A, int R, bool Spectral>
__global__ void __launch_bounds__(kThreads)
sample_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, SampleArgs a, size_t n) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[(A == kFast && Spectral) ? kPairTable : 1];
    stage_block(&S, (A == kFast && Spectral) ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x * R;
    for (size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * R; i < n; i += stride) {
        float u0[4], u1[4], wl[4][4], dx[4], dy[4], dz[4], pdf[4], q[4][4];
        float px[4] = { 0, 0, 0, 0 }, py[4] = { 0, 0, 0, 0 }, pz[4] = { 0, 0, 0, 0 };
        float ox[4] = { 0, 0, 0, 0 }, oy[4] = { 0, 0, 0, 0 }, oz[4] = { 0, 0, 0, 0 }, dist[4] = { 0, 0, 0, 0 };
        loadr<R>(a.u0, i, n, u0); loadr<R>(a.u1, i, n, u1);
        if (Spectral) {
            loadr<R>(a.w0, i, n, wl[0]); loadr<R>(a.w1, i, n, wl[1]);
            loadr<R>(a.w2, i, n, wl[2]); loadr<R>(a.w3, i, n, wl[3]);
        }
        if (a.px) { loadr<R>(a.px, i, n, px); loadr<R>(a.py, i, n, py); loadr<R>(a.pz, i, n, pz); }
#pragma unroll 1
        for (int k = 0; k < R; ++k) {
            DirSample s = sample_dir<A>(U, S, getk<R>(u0, k), getk<R>(u1, k));
            setk<R>(dx, k, s.d_world.x); setk<R>(dy, k, s.d_world.y); setk<R>(dz, k, s.d_world.z); setk<R>(pdf, k, s.pdf);
            // eval(si) with si.wi = -d (sunsky.cpp:430-438)
            Vec3 wi = { -s.d_world.x, -s.d_world.y, -s.d_world.z };
            float e[4];
            if (Spectral) {
                float w4[4] = { getk<R>(wl[0], k), getk<R>(wl[1], k), getk<R>(wl[2], k), getk<R>(wl[3], k) };
                eval_spec<A>(U, S, pairs, sun_rad, wi, w4, s.active, e);
            } else {
                eval_rgb<A>(U, sun_rad, wi, s.active, e);
                e[3] = 0.f;
            }
            // Spectrum / Float -> eval * rcp(pdf) (drjit/array_router.h:225-228)
            const float inv_pdf = A == kFast ? hw_rcp(s.pdf) : fdiv(1.f, s.pdf);
#pragma unroll
            for (int c = 0; c < (Spectral ? 4 : 3); ++c) {
                float w = e[c] * inv_pdf;
                setk<R>(q[c], k, (fabsf(w) < INFINITY) ? w : 0.f);         // res &= isfinite(res)
            }
            // bounding-sphere bookkeeping (sunsky.cpp:415-421)
            if (a.ox || a.dist) {
                float pxk = getk<R>(px, k), pyk = getk<R>(py, k), pzk = getk<R>(pz, k);
                Vec3 rel = { pxk - U.bs_center[0], pyk - U.bs_center[1], pzk - U.bs_center[2] };
                float radius = fmaxf(U.bs_radius, sqrtf(sqnorm3(rel)));
                float dd = 2.f * radius;
                setk<R>(dist, k, dd);
                setk<R>(ox, k, fmaf(s.d_world.x, dd, pxk)); setk<R>(oy, k, fmaf(s.d_world.y, dd, pyk)); setk<R>(oz, k, fmaf(s.d_world.z, dd, pzk));
            }
        }
        storer<R>(a.dx, i, n, dx); storer<R>(a.dy, i, n, dy); storer<R>(a.dz, i, n, dz);
        storer<R>(a.pdf, i, n, pdf);
        storer<R>(a.q0, i, n, q[0]); storer<R>(a.q1, i, n, q[1]); storer<R>(a.q2, i, n, q[2]);
        if (Spectral) storer<R>(a.q3, i, n, q[3]);
        if (a.ox) { storer<R>(a.ox, i, n, ox); storer<R>(a.oy, i, n, oy); storer<R>(a.oz, i, n, oz); }
        if (a.dist) storer<R>(a.dist, i, n, dist);
    }
}

/// pdf_direction, two directions per thread
template <int A>
__global__ void __launch_bounds__(kThreads, 4)
pdf_pair_kernel(const __grid_constant__ HotBlock U, Planes3 in, float *__restrict__ pdf, size_t n) {
    const size_t stride = (size_t) gridDim.x * blockDim.x * 2;
    size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * 2;
    Vec3p next = { f2(0.f, 0.f), f2(0.f, 0.f), f2(1.f, 1.f) };
    if (i < n) next = ld_dir2(in, i, n);
    for (; i < n; i += stride) {
        const bool two = i + 1 < n;
        const Vec3p d = next;
        if (SSK_PAIR_PREFETCH && i + stride < n) next = ld_dir2(in, i + stride, n);       // see eval_rgb_pair_kernel
        st2(pdf, i, two, pdf_direction_p<A>(U, d, true, two));
        if (!SSK_PAIR_PREFETCH && i + stride < n) next = ld_dir2(in, i + stride, n);
    }
}

/// sample_direction, two samples per thread (see sample_kernel)
/// Weight = false: first pass of the two-pass spectral path (direction, pdf and the `active` flag in q0; the
/// weight comes from weight_spec_kernel)
template <int A, bool Spectral, bool Weight = true>
__global__ void __launch_bounds__(kThreads, Spectral ? SSK_SAMPLE_MINB_SPEC : SSK_SAMPLE_MINB)
sample_pair_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, SampleArgs a, size_t n) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[(A == kFast && Spectral) ? kPairTable : 1];
    stage_block(&S, (A == kFast && Spectral) ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x * 2;
    for (size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * 2; i < n; i += stride) {
        const bool two = i + 1 < n;
        DirSampleP s = sample_dir_p<A>(U, S, ld2(a.u0, i, two, 0.5f), ld2(a.u1, i, two, 0.5f));
        st2(a.dx, i, two, s.d_world.x); st2(a.dy, i, two, s.d_world.y); st2(a.dz, i, two, s.d_world.z);
        st2(a.pdf, i, two, s.pdf);
        if (!Weight) st2(a.q0, i, two, f2(s.active0 ? 1.f : 0.f, s.active1 ? 1.f : 0.f));
        // eval(si) with si.wi = -d (sunsky.cpp:430-438)
        float2 e[4];
        if (Weight) {
        if (Spectral) {
            float2 w0 = ld2(a.w0, i, two, 0.f), w1 = ld2(a.w1, i, two, 0.f), w2 = ld2(a.w2, i, two, 0.f), w3 = ld2(a.w3, i, two, 0.f);
            float ea[4], eb[4];
            const float wa[4] = { w0.x, w1.x, w2.x, w3.x }, wb[4] = { w0.y, w1.y, w2.y, w3.y };
            eval_spec<A>(U, S, pairs, sun_rad, Vec3{ -s.d_world.x.x, -s.d_world.y.x, -s.d_world.z.x }, wa, s.active0, ea);
            eval_spec<A>(U, S, pairs, sun_rad, Vec3{ -s.d_world.x.y, -s.d_world.y.y, -s.d_world.z.y }, wb, s.active1 && two, eb);
#pragma unroll
            for (int c = 0; c < 4; ++c) e[c] = f2(ea[c], eb[c]);
        } else {
            Vec3p wi = { neg2(s.d_world.x), neg2(s.d_world.y), neg2(s.d_world.z) };
            eval_rgb_p<A>(U, sun_rad, wi, s.active0, s.active1 && two, e);
        }
        // Spectrum / Float -> eval * rcp(pdf) (drjit/array_router.h:225-228); res &= isfinite(res)
        const float2 inv_pdf = A == kFast ? f2(hw_rcp(s.pdf.x), hw_rcp(s.pdf.y)) : f2(fdiv(1.f, s.pdf.x), fdiv(1.f, s.pdf.y));
#pragma unroll
        for (int c = 0; c < (Spectral ? 4 : 3); ++c) {
            float2 w = mul2(e[c], inv_pdf);
            w = f2((fabsf(w.x) < INFINITY) ? w.x : 0.f, (fabsf(w.y) < INFINITY) ? w.y : 0.f);
            st2(c == 0 ? a.q0 : c == 1 ? a.q1 : c == 2 ? a.q2 : a.q3, i, two, w);
        }
        }
        // bounding-sphere bookkeeping (sunsky.cpp:415-421)
        if (a.ox || a.dist) {
            float2 px = f2(0.f, 0.f), py = px, pz = px;
            if (a.px) { px = ld2(a.px, i, two, 0.f); py = ld2(a.py, i, two, 0.f); pz = ld2(a.pz, i, two, 0.f); }
            Vec3p rel = { add2s(px, -U.bs_center[0]), add2s(py, -U.bs_center[1]), add2s(pz, -U.bs_center[2]) };
            float2 nr = sqrt2(sqnorm3_p(rel));
            float2 dd = mul2s(f2(fmaxf(U.bs_radius, nr.x), fmaxf(U.bs_radius, nr.y)), 2.f);
            if (a.dist) st2(a.dist, i, two, dd);
            if (a.ox) {
                st2(a.ox, i, two, fma2(s.d_world.x, dd, px)); st2(a.oy, i, two, fma2(s.d_world.y, dd, py));
                st2(a.oz, i, two, fma2(s.d_world.z, dd, pz));
            }
        }
    }
}

// ---- spectral sample_direction, second pass -------------------------------------------------------------
// The fused spectral kernels carry the sampler AND the 4-wavelength evaluation in one 73-112 KB body and stall on
// instruction fetch. Two kernels that each fit the instruction caches do the same arithmetic per sample: the first
// pass writes d, pdf and the sampler's `active` flag (in weight plane 0), this one evaluates eval(wi = -d) / pdf.
template <int A>
__global__ void __launch_bounds__(kThreads, A == kFast ? SSK_SPECF_MINB : SSK_SPECX_MINB)
weight_spec_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, SampleArgs a, size_t n) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[A == kFast ? kPairTable : 1];
    stage_block(&S, A == kFast ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    float nx = 0.f, ny = 0.f, nz = 0.f, np = 1.f, nf = 0.f, na = 0.f, nb = 0.f, nc = 0.f, nd = 0.f;
    if (i < n) {
        nx = __ldcs(a.dx + i); ny = __ldcs(a.dy + i); nz = __ldcs(a.dz + i); np = __ldcs(a.pdf + i); nf = __ldcs(a.q0 + i);
        na = __ldcs(a.w0 + i); nb = __ldcs(a.w1 + i); nc = __ldcs(a.w2 + i); nd = __ldcs(a.w3 + i);
    }
    for (; i < n; i += stride) {
        const Vec3 wi = { -nx, -ny, -nz };
        const float wl[4] = { na, nb, nc, nd }, pdf = np;
        const bool active = nf != 0.f;
        float o[4];
        const size_t j = i + stride;
        if (j < n) {
            nx = __ldcs(a.dx + j); ny = __ldcs(a.dy + j); nz = __ldcs(a.dz + j); np = __ldcs(a.pdf + j); nf = __ldcs(a.q0 + j);
            na = __ldcs(a.w0 + j); nb = __ldcs(a.w1 + j); nc = __ldcs(a.w2 + j); nd = __ldcs(a.w3 + j);
        }
        eval_spec<A>(U, S, pairs, sun_rad, wi, wl, active, o);
        const float inv_pdf = A == kFast ? hw_rcp(pdf) : fdiv(1.f, pdf);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float w = o[c] * inv_pdf;
            w = (fabsf(w) < INFINITY) ? w : 0.f;
            __stcs((c == 0 ? a.q0 : c == 1 ? a.q1 : c == 2 ? a.q2 : a.q3) + i, w);
        }
    }
}

// ---- sample_direction with both techniques active: technique-sorted tiles -------------------------------
// With 0 < m_sky_sampling_w < 1 about half of the samples take the sun-cone technique and half the TGMM, at
// random: in sample_pair_kernel every warp then executes both samplers and the sun-disc evaluation for a few
// lanes each (measured: 18 vs 36 Gsamples/s sky-only). Here a CTA takes a tile of 512 consecutive samples,
// ranks them by technique with warp ballots (sky samples first, sun samples after) and hands each thread two
// samples of the SAME technique, so all warps but the one on the boundary are homogeneous. Inputs of the tile are
// staged in shared memory, outputs are scattered back to their own slots (32-bit stores inside the tile's lines,
// merged in L2). Per-sample arithmetic is the pair code unchanged: results are bit-identical to any other pairing.
// A tile is kTileRounds x 512 samples: every thread sorts 2 * kTileRounds samples into the tile and then works through
// kTileRounds pairs, round r taking sorted slots 512 r + 2 t, + 1. With one round (round 1 of the project) the first
// warps of a CTA got TGMM samples and the last ones sun-cone samples, which cost 1.7 x as much: 24 % of the stall
// samples sat at the tile barrier. With several rounds every warp gets its share of each technique, and all warps of
// the CTA still run the same technique's code at the same time (instruction cache). Measured on B200, FAST, 2^24
// samples (sun + sky RGB / spectral / sample_ray RGB, Gsamples/s): 1 round 22.1 / 18.9 / 19.4, 2 rounds 23.4 / 20.0 /
// 20.2, 4 rounds 24.2 / 20.8 / 21.0, 8 rounds 23.0 / 20.4 / 19.5 (40 KB of shared memory per CTA).
#ifndef SSK_TILE_ROUNDS
#define SSK_TILE_ROUNDS 4
#endif
#ifndef SSK_SORT_PER_WARP
#define SSK_SORT_PER_WARP 0
#endif
constexpr int kTileRounds = SSK_SORT_PER_WARP ? 1 : SSK_TILE_ROUNDS;
constexpr int kTile = 2 * kThreads * kTileRounds;
__device__ __forceinline__ float2 gld2(const float *p, size_t ia, size_t ib, bool two, float pad) {
    return make_float2(__ldcs(p + ia), two ? __ldcs(p + ib) : pad);
}
__device__ __forceinline__ void gst2(float *p, size_t ia, size_t ib, bool two, float2 v) {
    __stcs(p + ia, v.x);
    if (two) __stcs(p + ib, v.y);
}
// Two implementations, bit-identical results. SSK_SORT_PER_WARP = 0 (default): the CTA sorts a 512-sample tile
// (three __syncthreads per tile). = 1: every warp sorts its own 512-sample tile (ballots and __syncwarp only). ncu of
// the CTA version showed 23 % of stall samples at the barriers, but without them the warps of an SM drift apart and
// the two techniques' code paths (65 KB of SASS) thrash the instruction cache: `no_inst` becomes the top stall (29 %)
// and the kernel is 4 % SLOWER (20.9 vs 21.7 Gsamples/s, profiles/r2_summary.md). The barriers were acting as a
// phase lock that keeps the instruction working set small.
#if SSK_SORT_PER_WARP
template <int A, bool Spectral, bool Weight = true>
__global__ void __launch_bounds__(kThreads, Spectral ? SSK_SAMPLE_MINB_SPEC : SSK_SAMPLE_MINB)
sample_sorted_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, SampleArgs a, size_t n) {
    // Round 2: every WARP sorts and processes its own 512-sample tile (ballots + __syncwarp only). The CTA-wide
    // version spent 23 % of its stall samples at __syncthreads: with w ~ 0.5 the first four warps of a CTA always
    // got TGMM samples and the last four sun-cone samples, and the tile took as long as the slower technique.
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[(A == kFast && Spectral) ? kPairTable : 1];
    __shared__ __align__(8) float tu0_all[kThreads / 32][kTile], tu1_all[kThreads / 32][kTile];
    __shared__ unsigned short perm_all[kThreads / 32][kTile];
    stage_block(&S, (A == kFast && Spectral) ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *tu0 = tu0_all[warp], *tu1 = tu1_all[warp];
    unsigned short *perm = perm_all[warp];
    const unsigned lt = (1u << lane) - 1u;
    const size_t n_tiles = (n + kTile - 1) / kTile;
    const size_t warps_total = (size_t) gridDim.x * (kThreads / 32);
    for (size_t tile = (size_t) blockIdx.x * (kThreads / 32) + warp; tile < n_tiles; tile += warps_total) {
        const size_t base = tile * kTile;
        const int valid = (int) (n - base < (size_t) kTile ? n - base : (size_t) kTile);
        // 1. + 2. stage the tile's samples (coalesced 64-bit loads), classify, and give every sample its slot: sky
        //    samples fill the tile from the front, sun samples from the back (of the valid part)
        int n_sky = 0, n_sun = 0;
#pragma unroll 2
        for (int k = 0; k < kTile / 64; ++k) {
            const int l0 = 64 * k + 2 * lane, l1 = l0 + 1;
            float2 v0 = f2(0.f, 0.f), v1 = f2(0.f, 0.f);
            if (l1 < valid) { v0 = __ldcs(reinterpret_cast<const float2 *>(a.u0 + base + l0)); v1 = __ldcs(reinterpret_cast<const float2 *>(a.u1 + base + l0)); }
            else if (l0 < valid) { v0.x = __ldcs(a.u0 + base + l0); v1.x = __ldcs(a.u1 + base + l0); }
            *reinterpret_cast<float2 *>(tu0 + l0) = v0; *reinterpret_cast<float2 *>(tu1 + l0) = v1;
            const bool ok0 = l0 < valid, ok1 = l1 < valid;
            const bool sky0 = ok0 && v0.x < U.sky_w, sky1 = ok1 && v0.y < U.sky_w;
            const unsigned bs0 = __ballot_sync(0xffffffffu, sky0), bs1 = __ballot_sync(0xffffffffu, sky1);
            const unsigned bn0 = __ballot_sync(0xffffffffu, ok0 && !sky0), bn1 = __ballot_sync(0xffffffffu, ok1 && !sky1);
            if (sky0) perm[n_sky + __popc(bs0 & lt)] = (unsigned short) l0;
            else if (ok0) perm[valid - 1 - (n_sun + __popc(bn0 & lt))] = (unsigned short) l0;
            if (sky1) perm[n_sky + __popc(bs0) + __popc(bs1 & lt)] = (unsigned short) l1;
            else if (ok1) perm[valid - 1 - (n_sun + __popc(bn0) + __popc(bn1 & lt))] = (unsigned short) l1;
            n_sky += __popc(bs0) + __popc(bs1); n_sun += __popc(bn0) + __popc(bn1);
        }
        __syncwarp();
#pragma unroll 1
        for (int k = 0; k < kTile / 64; ++k) {
        const int l0 = 64 * k + 2 * lane, l1 = l0 + 1;
        // 3. two samples of one technique per thread
        if (l0 < valid) {
            const bool two = l1 < valid;
            const int la = perm[l0], lb = two ? perm[l1] : la;
            const size_t ia = base + la, ib = base + lb;
            DirSampleP s = sample_dir_p<A>(U, S, f2(tu0[la], two ? tu0[lb] : 0.5f), f2(tu1[la], two ? tu1[lb] : 0.5f));
            gst2(a.dx, ia, ib, two, s.d_world.x); gst2(a.dy, ia, ib, two, s.d_world.y); gst2(a.dz, ia, ib, two, s.d_world.z);
            gst2(a.pdf, ia, ib, two, s.pdf);
            // first pass: weight plane 0 keeps, per SLOT of the sorted tile, the sample it holds and its `active`
            // flag, so that weight_spec_sorted_kernel walks the samples in the same technique order
            if (!Weight) {
                __stcs(a.q0 + base + l0, __int_as_float(la | (s.active0 ? kTile : 0)));
                if (two) __stcs(a.q0 + base + l1, __int_as_float(lb | (s.active1 ? kTile : 0)));
            }
            float2 e[4];
            if (Weight) {
            if (Spectral) {
                float2 w0 = gld2(a.w0, ia, ib, two, 0.f), w1 = gld2(a.w1, ia, ib, two, 0.f), w2 = gld2(a.w2, ia, ib, two, 0.f),
                       w3 = gld2(a.w3, ia, ib, two, 0.f);
                float ea[4], eb[4];
                const float wa[4] = { w0.x, w1.x, w2.x, w3.x }, wb[4] = { w0.y, w1.y, w2.y, w3.y };
                eval_spec<A>(U, S, pairs, sun_rad, Vec3{ -s.d_world.x.x, -s.d_world.y.x, -s.d_world.z.x }, wa, s.active0, ea);
                eval_spec<A>(U, S, pairs, sun_rad, Vec3{ -s.d_world.x.y, -s.d_world.y.y, -s.d_world.z.y }, wb, s.active1 && two, eb);
#pragma unroll
                for (int c = 0; c < 4; ++c) e[c] = f2(ea[c], eb[c]);
            } else {
                Vec3p wi = { neg2(s.d_world.x), neg2(s.d_world.y), neg2(s.d_world.z) };
                eval_rgb_p<A>(U, sun_rad, wi, s.active0, s.active1 && two, e);
            }
            const float2 inv_pdf = A == kFast ? f2(hw_rcp(s.pdf.x), hw_rcp(s.pdf.y)) : f2(fdiv(1.f, s.pdf.x), fdiv(1.f, s.pdf.y));
#pragma unroll
            for (int c = 0; c < (Spectral ? 4 : 3); ++c) {
                float2 w = mul2(e[c], inv_pdf);
                w = f2((fabsf(w.x) < INFINITY) ? w.x : 0.f, (fabsf(w.y) < INFINITY) ? w.y : 0.f);
                gst2(c == 0 ? a.q0 : c == 1 ? a.q1 : c == 2 ? a.q2 : a.q3, ia, ib, two, w);
            }
            }
            if (a.ox || a.dist) {          // bounding-sphere bookkeeping (sunsky.cpp:415-421), as sample_pair_kernel
                float2 px = f2(0.f, 0.f), py = px, pz = px;
                if (a.px) { px = gld2(a.px, ia, ib, two, 0.f); py = gld2(a.py, ia, ib, two, 0.f); pz = gld2(a.pz, ia, ib, two, 0.f); }
                Vec3p rel = { add2s(px, -U.bs_center[0]), add2s(py, -U.bs_center[1]), add2s(pz, -U.bs_center[2]) };
                float2 nr = sqrt2(sqnorm3_p(rel));
                float2 dd = mul2s(f2(fmaxf(U.bs_radius, nr.x), fmaxf(U.bs_radius, nr.y)), 2.f);
                if (a.dist) gst2(a.dist, ia, ib, two, dd);
                if (a.ox) {
                    gst2(a.ox, ia, ib, two, fma2(s.d_world.x, dd, px)); gst2(a.oy, ia, ib, two, fma2(s.d_world.y, dd, py));
                    gst2(a.oz, ia, ib, two, fma2(s.d_world.z, dd, pz));
                }
            }
        }
        }
        __syncwarp();                  // tu0 / tu1 / perm are rewritten by the next tile
    }
}

#else
/// Ray = true (RGB sample_ray, sunsky.cpp:354-397): u0 / u1 are sample3, the pdf carries the origin disc's
/// 1 / (pi r^2), the stored direction is the RAY's (-d), and the origin is computed from a.s2x / a.s2y into a.ox..oz
template <int A, bool Spectral, bool Weight = true, bool Ray = false>
__global__ void __launch_bounds__(kThreads, Spectral ? SSK_SAMPLE_MINB_SPEC : SSK_SAMPLE_MINB)
sample_sorted_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, SampleArgs a, size_t n) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[(A == kFast && Spectral) ? kPairTable : 1];
    __shared__ __align__(8) float tu0[kTile], tu1[kTile];
    __shared__ unsigned short perm[kTile];
    __shared__ int wcount[kTileRounds][kThreads / 32][2];
    stage_block(&S, (A == kFast && Spectral) ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const size_t n_tiles = (n + kTile - 1) / kTile;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t base = tile * kTile;
        const int valid = (int) (n - base < (size_t) kTile ? n - base : (size_t) kTile);
        // 1. stage the tile's samples (coalesced 64-bit loads) and classify; sample 512 h + 2 t (+ 1) in part h
        unsigned bs0[kTileRounds], bs1[kTileRounds], bn0[kTileRounds], bn1[kTileRounds];
#pragma unroll
        for (int h = 0; h < kTileRounds; ++h) {
            const int l0 = 2 * kThreads * h + 2 * t, l1 = l0 + 1;
            float2 v0 = f2(0.f, 0.f), v1 = f2(0.f, 0.f);
            if (l1 < valid) { v0 = __ldcs(reinterpret_cast<const float2 *>(a.u0 + base + l0)); v1 = __ldcs(reinterpret_cast<const float2 *>(a.u1 + base + l0)); }
            else if (l0 < valid) { v0.x = __ldcs(a.u0 + base + l0); v1.x = __ldcs(a.u1 + base + l0); }
            tu0[l0] = v0.x; tu0[l1] = v0.y; tu1[l0] = v1.x; tu1[l1] = v1.y;
            const bool ok0 = l0 < valid, ok1 = l1 < valid;
            const bool sky0 = ok0 && v0.x < U.sky_w, sky1 = ok1 && v0.y < U.sky_w;
            bs0[h] = __ballot_sync(0xffffffffu, sky0); bs1[h] = __ballot_sync(0xffffffffu, sky1);
            bn0[h] = __ballot_sync(0xffffffffu, ok0 && !sky0); bn1[h] = __ballot_sync(0xffffffffu, ok1 && !sky1);
            if (lane == 0) { wcount[h][warp][0] = __popc(bs0[h]) + __popc(bs1[h]); wcount[h][warp][1] = __popc(bn0[h]) + __popc(bn1[h]); }
        }
        __syncthreads();
        // 2. slot of each sample: sky samples first (part-major, then warp-major), then sun samples
        int sky_total = 0;
#pragma unroll
        for (int h = 0; h < kTileRounds; ++h)
#pragma unroll
            for (int w = 0; w < kThreads / 32; ++w) sky_total += wcount[h][w][0];
        int sky_before = 0, sun_before = 0;
#pragma unroll
        for (int h = 0; h < kTileRounds; ++h) {
            int sky_here = sky_before, sun_here = sun_before;
#pragma unroll
            for (int w = 0; w < kThreads / 32; ++w) {
                const int cs = wcount[h][w][0], cn = wcount[h][w][1];
                if (w < warp) { sky_here += cs; sun_here += cn; }
                sky_before += cs; sun_before += cn;
            }
            const int l0 = 2 * kThreads * h + 2 * t, l1 = l0 + 1;
            const bool in0 = (bs0[h] >> lane) & 1u, in1 = (bs1[h] >> lane) & 1u, sn0 = (bn0[h] >> lane) & 1u, sn1 = (bn1[h] >> lane) & 1u;
            if (in0) perm[sky_here + __popc(bs0[h] & lt)] = (unsigned short) l0;
            else if (sn0) perm[sky_total + sun_here + __popc(bn0[h] & lt)] = (unsigned short) l0;
            if (in1) perm[sky_here + __popc(bs0[h]) + __popc(bs1[h] & lt)] = (unsigned short) l1;
            else if (sn1) perm[sky_total + sun_here + __popc(bn0[h]) + __popc(bn1[h] & lt)] = (unsigned short) l1;
        }
        __syncthreads();
#pragma unroll 1
        for (int round = 0; round < kTileRounds; ++round) {
        const int l0 = 2 * kThreads * round + 2 * t, l1 = l0 + 1;
        // 3. two samples of one technique per thread
        if (l0 < valid) {
            const bool two = l1 < valid;
            const int la = perm[l0], lb = two ? perm[l1] : la;
            const size_t ia = base + la, ib = base + lb;
            DirSampleP s = sample_dir_p<A>(U, S, f2(tu0[la], two ? tu0[lb] : 0.5f), f2(tu1[la], two ? tu1[lb] : 0.5f));
            if (Ray) {
                static_assert(!Ray || !Spectral, "ray mode: RGB with fused weights, or the first pass of the spectral path");
                s.pdf = mul2s(s.pdf, U.ray_pdf_scale);
                s.active0 = s.active0 && s.pdf.x > 0.f; s.active1 = s.active1 && s.pdf.y > 0.f;
                const float2 rx = neg2(s.d_world.x), ry = neg2(s.d_world.y), rz = neg2(s.d_world.z);
                gst2(a.dx, ia, ib, two, rx); gst2(a.dy, ia, ib, two, ry); gst2(a.dz, ia, ib, two, rz);
                const float2 s2x = gld2(a.s2x, ia, ib, two, 0.5f), s2y = gld2(a.s2y, ia, ib, two, 0.5f);
                float dxa, dya, dxb, dyb;
                disk_concentric(s2x.x, s2y.x, dxa, dya); disk_concentric(s2x.y, s2y.y, dxb, dyb);
                const Vec3 oa = ray_origin(U, Vec3{ rx.x, ry.x, rz.x }, dxa, dya), ob = ray_origin(U, Vec3{ rx.y, ry.y, rz.y }, dxb, dyb);
                gst2(a.ox, ia, ib, two, f2(oa.x, ob.x)); gst2(a.oy, ia, ib, two, f2(oa.y, ob.y)); gst2(a.oz, ia, ib, two, f2(oa.z, ob.z));
                if (!Weight) gst2(a.q1, ia, ib, two, s.pdf);      // no pdf output in sample_ray: weight plane 1 carries it to the second pass
            } else {
                gst2(a.dx, ia, ib, two, s.d_world.x); gst2(a.dy, ia, ib, two, s.d_world.y); gst2(a.dz, ia, ib, two, s.d_world.z);
                gst2(a.pdf, ia, ib, two, s.pdf);
            }
            // first pass: weight plane 0 keeps, per SLOT of the sorted tile, the sample it holds and its `active`
            // flag, so that weight_spec_sorted_kernel walks the samples in the same technique order
            if (!Weight) {
                __stcs(a.q0 + base + l0, __int_as_float(la | (s.active0 ? kTile : 0)));
                if (two) __stcs(a.q0 + base + l1, __int_as_float(lb | (s.active1 ? kTile : 0)));
            }
            float2 e[4];
            if (Weight) {
            if (Spectral) {
                float2 w0 = gld2(a.w0, ia, ib, two, 0.f), w1 = gld2(a.w1, ia, ib, two, 0.f), w2 = gld2(a.w2, ia, ib, two, 0.f),
                       w3 = gld2(a.w3, ia, ib, two, 0.f);
                float ea[4], eb[4];
                const float wa[4] = { w0.x, w1.x, w2.x, w3.x }, wb[4] = { w0.y, w1.y, w2.y, w3.y };
                eval_spec<A>(U, S, pairs, sun_rad, Vec3{ -s.d_world.x.x, -s.d_world.y.x, -s.d_world.z.x }, wa, s.active0, ea);
                eval_spec<A>(U, S, pairs, sun_rad, Vec3{ -s.d_world.x.y, -s.d_world.y.y, -s.d_world.z.y }, wb, s.active1 && two, eb);
#pragma unroll
                for (int c = 0; c < 4; ++c) e[c] = f2(ea[c], eb[c]);
            } else {
                Vec3p wi = { neg2(s.d_world.x), neg2(s.d_world.y), neg2(s.d_world.z) };
                eval_rgb_p<A>(U, sun_rad, wi, s.active0, s.active1 && two, e);
            }
            const float2 inv_pdf = A == kFast ? f2(hw_rcp(s.pdf.x), hw_rcp(s.pdf.y)) : f2(fdiv(1.f, s.pdf.x), fdiv(1.f, s.pdf.y));
#pragma unroll
            for (int c = 0; c < (Spectral ? 4 : 3); ++c) {
                float2 w = mul2(e[c], inv_pdf);
                w = f2((fabsf(w.x) < INFINITY) ? w.x : 0.f, (fabsf(w.y) < INFINITY) ? w.y : 0.f);
                gst2(c == 0 ? a.q0 : c == 1 ? a.q1 : c == 2 ? a.q2 : a.q3, ia, ib, two, w);
            }
            }
            if (!Ray && (a.ox || a.dist)) { // bounding-sphere bookkeeping (sunsky.cpp:415-421), as sample_pair_kernel
                float2 px = f2(0.f, 0.f), py = px, pz = px;
                if (a.px) { px = gld2(a.px, ia, ib, two, 0.f); py = gld2(a.py, ia, ib, two, 0.f); pz = gld2(a.pz, ia, ib, two, 0.f); }
                Vec3p rel = { add2s(px, -U.bs_center[0]), add2s(py, -U.bs_center[1]), add2s(pz, -U.bs_center[2]) };
                float2 nr = sqrt2(sqnorm3_p(rel));
                float2 dd = mul2s(f2(fmaxf(U.bs_radius, nr.x), fmaxf(U.bs_radius, nr.y)), 2.f);
                if (a.dist) gst2(a.dist, ia, ib, two, dd);
                if (a.ox) {
                    gst2(a.ox, ia, ib, two, fma2(s.d_world.x, dd, px)); gst2(a.oy, ia, ib, two, fma2(s.d_world.y, dd, py));
                    gst2(a.oz, ia, ib, two, fma2(s.d_world.z, dd, pz));
                }
            }
        }
        }
        __syncthreads();               // tu0 / tu1 / perm are rewritten by the next tile
    }
}

#endif

/// Second pass after sample_sorted_kernel<A,false,false>: the tile's samples in technique order (sky samples in the
/// low slots, sun-cone samples -- all inside the sun disc -- in the high ones), thread t takes slots t and
/// t + kThreads, so only the warp on the boundary runs the sun-disc evaluation for some lanes and not others
/// (unsorted, every warp did: 0.63 vs 0.33 ms sky-only for 2^24 samples).
/// Ray = true: second pass of spectral sample_ray (sunsky.cpp:384-396): the stored direction is the ray's, the pdf
/// comes from weight plane 1, the wavelengths are importance-sampled here from a.wl_sample and written to a.ow0..3
template <int A, bool Ray = false>
__global__ void __launch_bounds__(kThreads, A == kFast ? SSK_SPECF_MINB : SSK_SPECX_MINB)
weight_spec_sorted_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, SampleArgs a, size_t n) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[A == kFast ? kPairTable : 1];
    stage_block(&S, A == kFast ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    const size_t n_tiles = (n + kTile - 1) / kTile;
    // slot codes are read one tile ahead: the dependent chain code -> sample index -> planes loses one memory round
    // trip per tile, and the barrier below no longer waits for the slowest warp's code load. Thread t takes the sorted
    // slots t, t + 256, ...: one from every stretch of the tile, sky samples first and sun-cone samples last
    constexpr int kSlots = kTile / kThreads;
    __shared__ int codes[2][kTile];                 // [tile parity][slot]; every thread reads back only what it wrote
    int buf = 0;
    if ((size_t) blockIdx.x < n_tiles) {
        const size_t base = (size_t) blockIdx.x * kTile;
        const int valid = (int) (n - base < (size_t) kTile ? n - base : (size_t) kTile);
#pragma unroll
        for (int k = 0; k < kSlots; ++k) {
            const int sk = (int) threadIdx.x + kThreads * k;
            codes[0][sk] = sk < valid ? __float_as_int(__ldcs(a.q0 + base + sk)) : -1;
        }
    }
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, buf ^= 1) {
        const size_t base = tile * kTile;
        const size_t next = tile + gridDim.x;
        if (next < n_tiles) {
            const size_t nbase = next * kTile;
            const int nvalid = (int) (n - nbase < (size_t) kTile ? n - nbase : (size_t) kTile);
#pragma unroll
            for (int k = 0; k < kSlots; ++k) {
                const int sk = (int) threadIdx.x + kThreads * k;
                codes[buf ^ 1][sk] = sk < nvalid ? __float_as_int(__ldcs(a.q0 + nbase + sk)) : -1;
            }
        }
        // every slot code of THIS tile was read (an iteration ago, before the previous barrier, or in the prologue) by
        // all threads before weight plane 0 of the tile is overwritten below; the next tile's codes are read before
        // this barrier and its plane is written after the next one
        __syncthreads();
        // (requesting the next slot's planes before this slot's arithmetic measured slower: 18.4 vs 18.7
        // Gsamples/s -- the FAST variant sits at its 80-register cap)
#pragma unroll 1
        for (int k = 0; k < kSlots; ++k) {
            const int cd = codes[buf][(int) threadIdx.x + kThreads * k];
            if (cd < 0) continue;
            const size_t i = base + (size_t) (cd & (kTile - 1));
            float o[4], pdf;
            if (Ray) {
                const Vec3 dr = { __ldcs(a.dx + i), __ldcs(a.dy + i), __ldcs(a.dz + i) };
                pdf = __ldcs(a.q1 + i);
                float w[4];
                sample_wavelengths_spec<A>(U, S, pairs, sun_rad, dr, __ldcs(a.wl_sample + i), (cd & kTile) != 0, w, o);
                __stcs(a.ow0 + i, w[0]); __stcs(a.ow1 + i, w[1]); __stcs(a.ow2 + i, w[2]); __stcs(a.ow3 + i, w[3]);
            } else {
                const Vec3 wi = { -__ldcs(a.dx + i), -__ldcs(a.dy + i), -__ldcs(a.dz + i) };
                const float wl[4] = { __ldcs(a.w0 + i), __ldcs(a.w1 + i), __ldcs(a.w2 + i), __ldcs(a.w3 + i) };
                pdf = __ldcs(a.pdf + i);
                eval_spec<A>(U, S, pairs, sun_rad, wi, wl, (cd & kTile) != 0, o);
            }
            const float inv_pdf = A == kFast ? hw_rcp(pdf) : fdiv(1.f, pdf);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float w = o[c] * inv_pdf;
                w = (fabsf(w) < INFINITY) ? w : 0.f;
                __stcs((c == 0 ? a.q0 : c == 1 ? a.q1 : c == 2 ? a.q2 : a.q3) + i, w);
            }
        }
    }
}

// ---- emitter setup on the device: the two 200 x 200 quadratures of estimate_sky_sun_ratio ------------------
// The reference integrates on the host / in a traced kernel and reduces with the LLVM backend's block sum:
// sequential fp32 adds inside 16384-wide chunks, then over the chunk partials (ssk_host.cuh llvm_block_sum).
// Here: one thread per node evaluates all channels (80 000 independent threads) and stores them chain-major
// (chain = integrand x channel x chunk, at most 66); one warp per chain then adds its values in index order.
constexpr int kQuadNodes = kQuadN * kQuadN;
constexpr int kQuadChunk = 16384;
constexpr int kQuadChunks = (kQuadNodes + kQuadChunk - 1) / kQuadChunk;
static_assert(kQuadNodes > 2 * kQuadChunk, "below 2 x 16384 elements the LLVM block sum is a single chain");

__global__ void __launch_bounds__(kThreads)
setup_quadrature_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, const float *__restrict__ gl,
                        float *__restrict__ buf, int C) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 2 * kQuadNodes) return;
    const int which = idx / kQuadNodes, j = idx - which * kQuadNodes, iy = j / kQuadN, ix = j - iy * kQuadN;
    const float gx_x = __ldg(gl + ix), gx_y = __ldg(gl + iy), gw_x = __ldg(gl + kQuadN + ix), gw_y = __ldg(gl + kQuadN + iy);
    float v[kMaxChannels];
    if (which == 0) quad_sky_node(U, C, gx_x, gx_y, gw_x, gw_y, v);
    else quad_sun_node(U, blk->sun_rad, C, gx_x, gx_y, gw_x, gw_y, v);
    const int chunk = j / kQuadChunk, pos = j - chunk * kQuadChunk;
    for (int c = 0; c < C; ++c) buf[(size_t) ((which * C + c) * kQuadChunks + chunk) * kQuadChunk + pos] = v[c];
}

/// One warp per chain. Lanes fetch 32 consecutive values with one coalesced load (several tiles in flight), the
/// values are then broadcast lane by lane and added in index order: the only serial dependency left is the
/// 16384-long chain of fp32 adds itself (every lane carries the same accumulator).
__global__ void __launch_bounds__(32)
setup_sum_kernel(const float *__restrict__ buf, float *__restrict__ partial, int nchains, size_t emitter_stride) {
    // blockIdx.x = emitter * nchains + chain (batched setup: several emitters' chains in one launch)
    const int emitter = blockIdx.x / nchains, chain = blockIdx.x - emitter * nchains, lane = threadIdx.x;
    const int chunk = chain % kQuadChunks;
    const int len = min(kQuadChunk, kQuadNodes - chunk * kQuadChunk);
    const float *p = buf + (size_t) emitter * emitter_stride + (size_t) chain * kQuadChunk;
    constexpr int kTiles = 8;
    float acc = 0.f;
    // the next group of tiles is requested before the current group is added: the chain of adds never waits for memory
    float nv[kTiles];
#pragma unroll
    for (int u = 0; u < kTiles; ++u) {
        const int j = u * 32 + lane;
        nv[u] = j < len ? __ldcg(p + j) : 0.f;
    }
    for (int j0 = 0; j0 < len; j0 += 32 * kTiles) {
        float v[kTiles];
#pragma unroll
        for (int u = 0; u < kTiles; ++u) {
            v[u] = nv[u];
            const int j = j0 + 32 * kTiles + u * 32 + lane;
            nv[u] = j < len ? __ldcg(p + j) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < kTiles; ++u) {
            const int left = len - (j0 + u * 32);                  // warp-uniform
            if (left >= 32) {
#pragma unroll
                for (int k = 0; k < 32; ++k) acc = acc + __shfl_sync(0xffffffffu, v[u], k);
            } else {
                for (int k = 0; k < left; ++k) acc = acc + __shfl_sync(0xffffffffu, v[u], k);
            }
        }
    }
    if (lane == 0) partial[blockIdx.x] = acc;
}

// ---- forward derivative of eval (turbidity / albedo / sun-direction tangents) -----------------------------------
/// The 14 KB tangent block travels as a kernel parameter (sm_70+ takes up to 32 KB): nothing to allocate, copy or
/// race on between calls. Spectral lanes index the sky coefficients and their tangent with their own wavelength bin;
/// from the parameter bank that serialises per distinct channel, so both are copied to shared memory first.
template <int A, bool Spectral>
__global__ void __launch_bounds__(kThreads)
eval_jvp_kernel(const __grid_constant__ HotBlock U, const __grid_constant__ TangentBlock T, const EmitterBlock *__restrict__ blk,
                Planes3 in, Planes4 wav, OutPlanes4 out, size_t n) {
    __shared__ SkyCoefs SP, ST;
    if (Spectral) {
        for (int i = threadIdx.x; i < (int) (sizeof(SkyCoefs) / 4); i += blockDim.x) {
            const int row = i / kChanPad, c = i % kChanPad;
            if (row < kSkyParams) { SP.sky_params[row][c] = U.sky_params[row][c]; ST.sky_params[row][c] = T.sky_params[row][c]; }
            else                  { SP.sky_rad[c] = U.sky_rad[c];                 ST.sky_rad[c] = T.sky_rad[c]; }
        }
        __syncthreads();
    }
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Vec3 wi = { __ldcs(in.x + i), __ldcs(in.y + i), __ldcs(in.z + i) };
        float o[4] = { 0.f, 0.f, 0.f, 0.f };
        if (Spectral) {
            const float wl[4] = { __ldcs(wav.a + i), __ldcs(wav.b + i), __ldcs(wav.c + i), __ldcs(wav.d + i) };
            eval_spec_jvp<A>(U, SP, ST, T, sun_rad, wi, wl, o);
            __stcs(out.d + i, o[3]);
        } else {
            eval_rgb_jvp<A>(U, T, sun_rad, wi, o);
        }
        __stcs(out.a + i, o[0]); __stcs(out.b + i, o[1]); __stcs(out.c + i, o[2]);
    }
}

// ---- sample_wavelengths / sample_ray -------------------------------------------------------------------
/// sample_wavelengths, spectral (sunsky.cpp:463-473): one ray per thread; 4 planes in (wi, sample), 8 planes out
template <int A>
__global__ void __launch_bounds__(kThreads, 3)
sample_wavelengths_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, Planes3 in,
                          const float *__restrict__ sample, OutPlanes4 wav, OutPlanes4 weight, size_t n) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[A == kFast ? kPairTable : 1];
    stage_block(&S, A == kFast ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    // software-pipelined like eval_spec_pipelined: the next grid-stride iteration's four planes are requested before
    // this ray's arithmetic (prefetch index clamped to the last ray)
    const size_t stride = (size_t) gridDim.x * blockDim.x, last = n - 1;
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float nx = __ldcs(in.x + i), ny = __ldcs(in.y + i), nz = __ldcs(in.z + i), ns = __ldcs(sample + i);
    for (;;) {
        Vec3 wi = { nx, ny, nz };
        const float smp = ns;
        const size_t j = i + stride, jc = j < last ? j : last;
        nx = __ldcs(in.x + jc); ny = __ldcs(in.y + jc); nz = __ldcs(in.z + jc); ns = __ldcs(sample + jc);
        float w[4], q[4];
        sample_wavelengths_spec<A>(U, S, pairs, sun_rad, wi, smp, true, w, q);
        __stcs(wav.a + i, w[0]); __stcs(wav.b + i, w[1]); __stcs(wav.c + i, w[2]); __stcs(wav.d + i, w[3]);
        __stcs(weight.a + i, q[0]); __stcs(weight.b + i, q[1]); __stcs(weight.c + i, q[2]); __stcs(weight.d + i, q[3]);
        if (j >= n) break;
        i = j;
    }
}

struct RayArgs {
    const float *wl_sample;             // wavelength sample (spectral only)
    const float *s2x, *s2y, *s3x, *s3y; // sample2 (origin disc), sample3 (direction)
    float *ox, *oy, *oz, *dx, *dy, *dz; // ray origin and direction
    float *w0, *w1, *w2, *w3;           // sampled wavelengths (spectral only)
    float *q0, *q1, *q2, *q3;           // weight planes
};

/// sample_ray, fused (sunsky.cpp:354-397): origin disc -> technique -> direction -> pdf -> wavelengths -> eval -> weight
/// Weight = false: first pass of the two-pass spectral path (origin, direction, `active` in q0, pdf in q1)
template <int A, bool Spectral, bool Weight = true>
__global__ void __launch_bounds__(kThreads, 3)
sample_ray_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, RayArgs a, size_t n) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[(A == kFast && Spectral) ? kPairTable : 1];
    stage_block(&S, (A == kFast && Spectral) ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        RaySample r = sample_ray_geom<A>(U, S, __ldcs(a.s2x + i), __ldcs(a.s2y + i), __ldcs(a.s3x + i), __ldcs(a.s3y + i));
        __stcs(a.ox + i, r.origin.x); __stcs(a.oy + i, r.origin.y); __stcs(a.oz + i, r.origin.z);
        __stcs(a.dx + i, r.d_world.x); __stcs(a.dy + i, r.d_world.y); __stcs(a.dz + i, r.d_world.z);
        if (!Weight) { __stcs(a.q0 + i, r.active ? 1.f : 0.f); __stcs(a.q1 + i, r.pdf); continue; }
        float e[4];
        if (Spectral) {
            float w[4];
            sample_wavelengths_spec<A>(U, S, pairs, sun_rad, r.d_world, __ldcs(a.wl_sample + i), r.active, w, e);
            __stcs(a.w0 + i, w[0]); __stcs(a.w1 + i, w[1]); __stcs(a.w2 + i, w[2]); __stcs(a.w3 + i, w[3]);
        } else {
            eval_rgb<A>(U, sun_rad, r.d_world, r.active, e);
        }
        // weight /= pdf (Spectrum / Float -> weight * rcp(pdf)); weight &= isfinite(weight)
        const float inv_pdf = A == kFast ? hw_rcp(r.pdf) : fdiv(1.f, r.pdf);
#pragma unroll
        for (int c = 0; c < (Spectral ? 4 : 3); ++c) {
            float w = e[c] * inv_pdf;
            __stcs((c == 0 ? a.q0 : c == 1 ? a.q1 : c == 2 ? a.q2 : a.q3) + i, (fabsf(w) < INFINITY) ? w : 0.f);
        }
    }
}

/// spectral sample_ray, second pass (see weight_spec_kernel): wavelengths, eval and weight = eval / pdf for the rays
/// the first pass generated
template <int A>
__global__ void __launch_bounds__(kThreads, 3)
ray_weight_spec_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, RayArgs a, size_t n) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[A == kFast ? kPairTable : 1];
    stage_block(&S, A == kFast ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const Vec3 d = { __ldcs(a.dx + i), __ldcs(a.dy + i), __ldcs(a.dz + i) };
        const bool active = __ldcs(a.q0 + i) != 0.f;
        const float pdf = __ldcs(a.q1 + i);
        float w[4], e[4];
        sample_wavelengths_spec<A>(U, S, pairs, sun_rad, d, __ldcs(a.wl_sample + i), active, w, e);
        __stcs(a.w0 + i, w[0]); __stcs(a.w1 + i, w[1]); __stcs(a.w2 + i, w[2]); __stcs(a.w3 + i, w[3]);
        const float inv_pdf = A == kFast ? hw_rcp(pdf) : fdiv(1.f, pdf);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float q = e[c] * inv_pdf;
            __stcs((c == 0 ? a.q0 : c == 1 ? a.q1 : c == 2 ? a.q2 : a.q3) + i, (fabsf(q) < INFINITY) ? q : 0.f);
        }
    }
}

/// Direction of equirectangular pixel (x, y): phi = 2*pi*(x+.5)/W, theta = pi*(y+.5)/H
SSK_HD Vec3 envmap_dir(int x, int y, int width, int height) {
    float phi = fdiv(SSK_TWO_PI * ((float) x + 0.5f), (float) width),
          theta = fdiv(SSK_PI * ((float) y + 0.5f), (float) height);
    return sphdir(theta, phi);
}

/// Equirect env map of one emitter. Write-only: no direction buffer. A thread handles 4 consecutive
/// pixels of one row (width % 4 == 0 on the vector path), so sin/cos(theta) is computed once per thread.
template <int A, bool Spectral>
__global__ void __launch_bounds__(kThreads, Spectral ? (A == kFast ? SSK_ENVMAP_SPEC_MINB : 3) : SSK_ENVMAP_MINB)
envmap_kernel(const __grid_constant__ HotBlock U, const EmitterBlock *__restrict__ blk, float4 wav4,
              float *__restrict__ out, int width, int height) {
    __shared__ __align__(16) HotBlock S;
    __shared__ __align__(16) float2 pairs[(A == kFast && Spectral) ? kPairTable : 1];
    if (Spectral) stage_block(&S, A == kFast ? pairs : nullptr, blk);
    const float *sun_rad = blk->sun_rad;
    constexpr int C = Spectral ? 4 : 3;
    const size_t npix = (size_t) width * height;
    const float wl[4] = { wav4.x, wav4.y, wav4.z, wav4.w };
    const bool vec = (width % 4) == 0 && (reinterpret_cast<uintptr_t>(out) & 15u) == 0;
    constexpr int R = 4;
    const size_t stride = (size_t) gridDim.x * blockDim.x * R;
    for (size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * R; i < npix; i += stride) {
        float o[4][4];
        if (!Spectral && vec) {
            // RGB, 4 pixels of one row: row / column from one division, theta once, the reciprocal of the width once
            // (fdiv_by_safe returns the IEEE quotient of envmap_dir bit for bit), phi pairs on the packed path
            const int y = (int) (i / (size_t) width), x = (int) (i - (size_t) y * width);
            const float fw = (float) width, rw = fdiv(1.f, fw);
            const float theta = fdiv(SSK_PI * ((float) y + 0.5f), (float) height);
#pragma unroll 1
            for (int k = 0; k < 4; k += 2) {
                float2 phi = f2(fdiv_by_safe(SSK_TWO_PI * ((float) (x + k) + 0.5f), fw, rw),
                                fdiv_by_safe(SSK_TWO_PI * ((float) (x + k + 1) + 0.5f), fw, rw));
                Vec3p d = sphdir_p(f2(theta, theta), phi);
                Vec3p wi = { neg2(d.x), neg2(d.y), neg2(d.z) };
                float2 e[3];
                eval_rgb_p<A>(U, sun_rad, wi, true, true, e);
#pragma unroll
                for (int c = 0; c < 3; ++c) { setk<R>(o[c], k, e[c].x); setk<R>(o[c], k + 1, e[c].y); }
            }
        } else if (!Spectral) {
            // RGB: two pixels at a time on the packed two-ray path (the same functions as eval_rgb_pair_kernel)
#pragma unroll 1
            for (int k = 0; k < 4; k += 2) {
                size_t p0 = i + k < npix ? i + k : npix - 1, p1 = i + k + 1 < npix ? i + k + 1 : npix - 1;
                int y0 = (int) (p0 / (size_t) width), x0 = (int) (p0 - (size_t) y0 * width);
                int y1 = (int) (p1 / (size_t) width), x1 = (int) (p1 - (size_t) y1 * width);
                Vec3 d0 = envmap_dir(x0, y0, width, height), d1 = envmap_dir(x1, y1, width, height);
                Vec3p wi = { f2(-d0.x, -d1.x), f2(-d0.y, -d1.y), f2(-d0.z, -d1.z) };
                float2 e[3];
                eval_rgb_p<A>(U, sun_rad, wi, true, true, e);
#pragma unroll
                for (int c = 0; c < 3; ++c) { setk<R>(o[c], k, e[c].x); setk<R>(o[c], k + 1, e[c].y); }
            }
        } else if (Spectral && vec) {
            // spectral, 4 pixels of one row: as the RGB branch (sin / cos of theta shared by the 4 pixels)
            const int y = (int) (i / (size_t) width), x = (int) (i - (size_t) y * width);
            const float fw = (float) width, rw = fdiv(1.f, fw);
            float st, ct;
            dsincos(fdiv(SSK_PI * ((float) y + 0.5f), (float) height), st, ct);
#pragma unroll 1
            for (int k = 0; k < 4; ++k) {
                float sp, cp;
                dsincos(fdiv_by_safe(SSK_TWO_PI * ((float) (x + k) + 0.5f), fw, rw), sp, cp);
                Vec3 wi = { -(cp * st), -(sp * st), -ct };
                float e[4] = { 0, 0, 0, 0 };
                eval_spec<A>(U, S, pairs, sun_rad, wi, wl, true, e);
#pragma unroll
                for (int c = 0; c < C; ++c) setk<R>(o[c], k, e[c]);
            }
        } else
#pragma unroll 1
        for (int k = 0; k < 4; ++k) {
            size_t pix = i + k < npix ? i + k : npix - 1;
            int py = (int) (pix / (size_t) width), px = (int) (pix - (size_t) py * width);
            Vec3 d = envmap_dir(px, py, width, height);
            Vec3 wi = { -d.x, -d.y, -d.z };
            float e[4] = { 0, 0, 0, 0 };
            if (Spectral) eval_spec<A>(U, S, pairs, sun_rad, wi, wl, true, e);
            else eval_rgb<A>(U, sun_rad, wi, true, e);
#pragma unroll
            for (int c = 0; c < C; ++c) setk<R>(o[c], k, e[c]);
        }
#pragma unroll
        for (int c = 0; c < C; ++c) {
            if (vec) storer<4>(out + (size_t) c * npix, i, npix, o[c]);
            else for (int k = 0; k < 4; ++k) if (i + k < npix) out[(size_t) c * npix + i + k] = o[c][k];
        }
    }
}

/// Directions of an env map (for parity checks of the pixel -> direction convention)
__global__ void envmap_dirs_kernel(float *dx, float *dy, float *dz, int width, int height) {
    size_t npix = (size_t) width * height;
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t) gridDim.x * blockDim.x) {
        int py = (int) (i / (size_t) width), px = (int) (i - (size_t) py * width);
        Vec3 d = envmap_dir(px, py, width, height);
        dx[i] = d.x; dy[i] = d.y; dz[i] = d.z;
    }
}

/// Test hook: out[i] = fdiv_by(a[i], b[i], RN(1 / b[i])), compared with IEEE division by the tests
__global__ void fdiv_by_kernel(const float *a, const float *b, float *out, size_t n) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
        out[i] = fdiv_by(a[i], b[i], fdiv(1.f, b[i]));
}

/// Test hook for the guarded shortcuts of EXACT mode, each next to the sequence it replaces (bit-compared by the tests):
///   op 0 / 1: fdiv_normal_2(a, b) / IEEE a / b      op 2 / 3: dexp_2<true>(a) / dexp(a)
///   op 4 / 5: dpow15_2_safe(a) / dpow15(a)          op 6 / 7: sqrt_normal(a) / sqrtf(a)      op 8: fdiv_normal(a, b)
__global__ void math_probe_kernel(int op, const float *a, const float *b, float *out, size_t n) {
    for (size_t i = ((size_t) blockIdx.x * blockDim.x + threadIdx.x) * 2; i + 1 < n; i += (size_t) gridDim.x * blockDim.x * 2) {
        float2 x = f2(a[i], a[i + 1]), y = b ? f2(b[i], b[i + 1]) : splat2(1.f), r;
        switch (op) {
            case 0: r = fdiv_normal_2(x, y); break;
            case 1: r = f2(fdiv(x.x, y.x), fdiv(x.y, y.y)); break;
            case 2: r = dexp_2<true>(x); break;
            case 3: r = f2(dexp(x.x), dexp(x.y)); break;
            case 4: r = dpow15_2_safe(x); break;
            case 5: r = f2(dpow15(x.x), dpow15(x.y)); break;
            case 6: r = f2(sqrt_normal(x.x), sqrt_normal(x.y)); break;
            case 7: r = f2(sqrtf(x.x), sqrtf(x.y)); break;
            default: r = f2(fdiv_normal(x.x, y.x), fdiv_normal(x.y, y.y)); break;
        }
        out[i] = r.x; out[i + 1] = r.y;
    }
}

} // namespace ssk
