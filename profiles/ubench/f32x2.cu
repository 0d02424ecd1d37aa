// Microbenchmark: issue throughput of scalar FP32 (FFMA/FMUL/FADD), packed FP32 (FFMA2/FMUL2/FADD2 =
// fma/mul/add.rn.f32x2, sm_100+) and MUFU.EX2 on B200, per SM per clock. Decides whether packing the two
// wavelength bins / two rays of the sunsky kernels into f32x2 instructions buys issue slots.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2 f32x2.cu && ./f32x2
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 2048, CH = 8;

template <int MODE> __global__ void __launch_bounds__(256) k(float *out, float a, float b) {
    float2 v[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) v[i] = make_float2(threadIdx.x * 1e-3f + i, blockIdx.x * 1e-4f - i);
    float2 A = make_float2(a, a * 1.0001f), B = make_float2(b, b * 0.999f);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (MODE == 0) { v[i].x = __fmaf_rn(v[i].x, A.x, B.x); v[i].y = __fmaf_rn(v[i].y, A.y, B.y); }   // 2 FFMA
            if (MODE == 1) v[i] = __ffma2_rn(v[i], A, B);                                                    // 1 FFMA2
            if (MODE == 2) { v[i].x = __fmul_rn(v[i].x, A.x); v[i].y = __fmul_rn(v[i].y, A.y); }
            if (MODE == 3) v[i] = __fmul2_rn(v[i], A);
            if (MODE == 4) { v[i].x = __fadd_rn(v[i].x, B.x); v[i].y = __fadd_rn(v[i].y, B.y); }
            if (MODE == 5) v[i] = __fadd2_rn(v[i], B);
            if (MODE == 6) { v[i].x = exp2f(v[i].x); v[i].y = exp2f(v[i].y); }   // -use_fast_math: MUFU.EX2
            if (MODE == 7) {   // mix: 1 FFMA2 + 1 scalar ALU op (FSEL-like integer min) per pair
                v[i] = __ffma2_rn(v[i], A, B);
                v[i].x = fminf(v[i].x, 1e30f);
            }
            if (MODE == 8) {   // mix: 2 FFMA + 1 FMNMX
                v[i].x = __fmaf_rn(v[i].x, A.x, B.x); v[i].y = __fmaf_rn(v[i].y, A.y, B.y);
                v[i].x = fminf(v[i].x, 1e30f);
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s += v[i].x + v[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE> void run(const char *name, float *out, int sms, double mhz, int lanes_per_iter) {
    int grid = sms * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 3; ++w) k<MODE><<<grid, 256>>>(out, 0.999f, 1e-3f);
    cudaEventRecord(e0);
    const int reps = 10;
    for (int r = 0; r < reps; ++r) k<MODE><<<grid, 256>>>(out, 0.999f, 1e-3f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= reps;
    double lane_ops = (double) grid * 256 * ITERS * CH * lanes_per_iter;
    double per_clk_sm = lane_ops / (ms * 1e-3) / (mhz * 1e6) / sms;
    printf("%-28s %8.3f ms  %7.1f lane-ops/clk/SM (at %.0f MHz)\n", name, ms, per_clk_sm, mhz);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount; int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double mhz = khz / 1e3;
    printf("%s, %d SMs, %.0f MHz\n", p.name, sms, mhz);
    float *out; cudaMalloc(&out, (size_t) sms * 8 * 256 * 4);
    run<0>("2 x FFMA", out, sms, mhz, 2);
    run<1>("1 x FFMA2", out, sms, mhz, 2);
    run<2>("2 x FMUL", out, sms, mhz, 2);
    run<3>("1 x FMUL2", out, sms, mhz, 2);
    run<4>("2 x FADD", out, sms, mhz, 2);
    run<5>("1 x FADD2", out, sms, mhz, 2);
    run<6>("2 x MUFU.EX2", out, sms, mhz, 2);
    run<7>("FFMA2 + FMNMX (3 lane-ops)", out, sms, mhz, 3);
    run<8>("2 FFMA + FMNMX (3 lane-ops)", out, sms, mhz, 3);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
