// GPU debug aid: packed-pair math functions against their scalar twins, bit for bit, on the device.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false ...
#include "../../mitsuba3-sunsky_b200/csrc/ssk_math.cuh"
#include <cstdio>
#include <vector>
#include <random>
using namespace ssk;
__global__ void k(const float *a, const float *b, float *o, int n, int op) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float2 x = f2(a[i], b[i]), r; float s0, s1;
    switch (op) {
        case 0: r = dexp_2(x); s0 = dexp(x.x); s1 = dexp(x.y); break;
        case 1: r = dexp2_2(x); s0 = dexp2(x.x); s1 = dexp2(x.y); break;
        case 2: r = dlog2_2(x); s0 = dlog2(x.x); s1 = dlog2(x.y); break;
        case 3: r = dpow15_2(x); s0 = dpow15(x.x); s1 = dpow15(x.y); break;
        default: r = x; s0 = x.x; s1 = x.y;
    }
    o[4 * i] = r.x; o[4 * i + 1] = r.y; o[4 * i + 2] = s0; o[4 * i + 3] = s1;
}
int main() {
    const int n = 1 << 20;
    std::mt19937 g(1);
    const char *names[] = { "exp", "exp2", "log2", "pow15" };
    for (int op = 0; op < 4; ++op) {
        std::vector<float> a(n), b(n), o(4 * n);
        std::uniform_real_distribution<float> d(op >= 2 ? 1e-6f : -20.f, op >= 2 ? 8.f : 20.f);
        for (int i = 0; i < n; ++i) { a[i] = d(g); b[i] = d(g); }
        float *da, *db, *dout;
        cudaMalloc(&da, n * 4); cudaMalloc(&db, n * 4); cudaMalloc(&dout, 16 * n);
        cudaMemcpy(da, a.data(), n * 4, cudaMemcpyHostToDevice); cudaMemcpy(db, b.data(), n * 4, cudaMemcpyHostToDevice);
        k<<<n / 256, 256>>>(da, db, dout, n, op);
        cudaMemcpy(o.data(), dout, 16 * n, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int i = 0; i < n; ++i)
            for (int e = 0; e < 2; ++e)
                if (f2u(o[4 * i + e]) != f2u(o[4 * i + 2 + e])) {
                    if (bad < 4) printf("  %s(%a) packed %a scalar %a (element %d)\n", names[op], e ? b[i] : a[i], o[4 * i + e], o[4 * i + 2 + e], e);
                    ++bad;
                }
        printf("%s: %d mismatches of %d (%s)\n", names[op], bad, 2 * n, cudaGetErrorString(cudaDeviceSynchronize()));
    }
}
