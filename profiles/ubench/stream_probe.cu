// Memory ceiling of the spectral-eval access pattern: 7 input planes + 4 output planes of n floats, no arithmetic
// to speak of. Variants: 4 B / thread LDG (256-ray tiles), 16 B / thread LDG (1024-ray tiles), TMA bulk tiles of
// 1 KB and 4 KB per plane. Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o stream_probe stream_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
struct P7 { const float *p[7]; };
struct P4 { float *p[4]; };

__global__ void __launch_bounds__(256) k_ldg4(P7 in, P4 out, size_t n) {
    const size_t stride = (size_t) gridDim.x * 256;
    for (size_t i = (size_t) blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
        float s = 0.f;
#pragma unroll
        for (int p = 0; p < 7; ++p) s += __ldcs(in.p[p] + i);
#pragma unroll
        for (int q = 0; q < 4; ++q) __stcs(out.p[q] + i, s + q);
    }
}
__global__ void __launch_bounds__(256) k_ldg16(P7 in, P4 out, size_t n) {
    const size_t stride = (size_t) gridDim.x * 1024;
    for (size_t i = ((size_t) blockIdx.x * 256 + threadIdx.x) * 4; i < n; i += stride) {
        float4 s = make_float4(0, 0, 0, 0);
#pragma unroll
        for (int p = 0; p < 7; ++p) { float4 v = __ldcs(reinterpret_cast<const float4 *>(in.p[p] + i)); s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
#pragma unroll
        for (int q = 0; q < 4; ++q) __stcs(reinterpret_cast<float4 *>(out.p[q] + i), s);
    }
}
__device__ __forceinline__ uint32_t su32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
template <int TILE, int STAGES>
__global__ void __launch_bounds__(256) k_tma(P7 in, P4 out, size_t n) {
    extern __shared__ __align__(128) unsigned char raw[];
    float (*data)[7][TILE] = reinterpret_cast<float (*)[7][TILE]>(raw);
    uint64_t *full = reinterpret_cast<uint64_t *>(raw + sizeof(float) * STAGES * 7 * TILE);
    int *released = reinterpret_cast<int *>(full + STAGES);
    const size_t n_tiles = n / TILE;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(su32(&full[s]))); released[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int s, size_t t) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(su32(&full[s])), "r"(7u * TILE * 4u) : "memory");
        for (int p = 0; p < 7; ++p)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(su32(&data[s][p][0])), "l"(in.p[p] + t * TILE), "r"((uint32_t) (TILE * 4)), "r"(su32(&full[s])) : "memory");
    };
    if (threadIdx.x == 0) for (int j = 0; j < STAGES; ++j) { size_t t = blockIdx.x + (size_t) j * gridDim.x; if (t < n_tiles) issue(j, t); }
    uint32_t k = 0;
    for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++k) {
        const int s = k % STAGES; const uint32_t par = (k / STAGES) & 1;
        asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(su32(&full[s])), "r"(par) : "memory");
        float acc[TILE / 256];
#pragma unroll
        for (int j = 0; j < TILE / 256; ++j) { float v = 0.f; for (int p = 0; p < 7; ++p) v += data[s][p][j * 256 + threadIdx.x]; acc[j] = v; }
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
            __threadfence_block();
            if (atomicAdd(&released[s], 1) == 7) { released[s] = 0; __threadfence_block(); size_t tn = t + (size_t) STAGES * gridDim.x; if (tn < n_tiles) issue(s, tn); }
        }
#pragma unroll
        for (int j = 0; j < TILE / 256; ++j)
            for (int q = 0; q < 4; ++q) __stcs(out.p[q] + t * TILE + j * 256 + threadIdx.x, acc[j] + q);
    }
}
int main() {
    const size_t n = 1 << 24;
    P7 in; P4 out;
    for (int p = 0; p < 7; ++p) { float *d; CK(cudaMalloc(&d, n * 4)); CK(cudaMemset(d, 0, n * 4)); in.p[p] = d; }
    for (int q = 0; q < 4; ++q) CK(cudaMalloc(&out.p[q], n * 4));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto timeit = [&](const char *name, auto launch) {
        for (int i = 0; i < 3; ++i) launch();
        cudaEventRecord(e0);
        for (int i = 0; i < 20; ++i) launch();
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 20;
        printf("%-28s %.4f ms  %.0f GB/s  (%s)\n", name, ms, 44.0 * n / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
    };
    for (int per_sm : { 3, 4, 6, 8 }) {
        int grid = 148 * per_sm; char nm[64];
        snprintf(nm, 64, "ldg 4B/thread, %d CTA/SM", per_sm); timeit(nm, [&] { k_ldg4<<<grid, 256>>>(in, out, n); });
        snprintf(nm, 64, "ldg 16B/thread, %d CTA/SM", per_sm); timeit(nm, [&] { k_ldg16<<<grid, 256>>>(in, out, n); });
    }
    auto tma = [&](auto kern, int tile, int stages, int per_sm, const char *nm) {
        size_t sm = (size_t) stages * 7 * tile * 4 + stages * 16 + 64;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sm);
        timeit(nm, [&] { kern<<<148 * per_sm, 256, sm>>>(in, out, n); });
    };
    tma(k_tma<256, 3>, 256, 3, 3, "tma 1KB x3 stages, 3 CTA/SM");
    tma(k_tma<256, 3>, 256, 3, 4, "tma 1KB x3 stages, 4 CTA/SM");
    tma(k_tma<256, 6>, 256, 6, 4, "tma 1KB x6 stages, 4 CTA/SM");
    tma(k_tma<1024, 2>, 1024, 2, 3, "tma 4KB x2 stages, 3 CTA/SM");
    tma(k_tma<1024, 3>, 1024, 3, 2, "tma 4KB x3 stages, 2 CTA/SM");
    tma(k_tma<2048, 2>, 2048, 2, 1, "tma 8KB x2 stages, 1 CTA/SM");
    return 0;
}
