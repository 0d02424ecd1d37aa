// DEBUG AID for development in the GPU-less container only (never shipped, never used by the product,
// the GPU tests, smoke() or bench.py): runs the __host__ __device__ per-lane model of
// mitsuba3-sunsky_b200/csrc/ssk_model.cuh on the host so restatement bugs show up before a GPU run.
#include "../../mitsuba3-sunsky_b200/csrc/ssk_host.cuh"
namespace ssk { long long g_fast_calls = 0, g_fast_fallbacks = 0; }
using namespace ssk;
extern "C" void hostdbg_fallback_stats(long long *calls, long long *fb) { *calls = g_fast_calls; *fb = g_fast_fallbacks; g_fast_calls = g_fast_fallbacks = 0; }
static const float2 *E_pairs(const HostEmitter &E) {
    static thread_local float2 pairs[kPairTable];
    build_pair_table(E.block, pairs);
    return pairs;
}
static const float2 *E_xpairs(const HostEmitter &E) {
    static thread_local float2 xp[kExactPairTable];
    build_exact_pair_table(E.block, xp);
    return xp;
}
extern "C" {
void *hostdbg_create(const ssk_params *p, const char *dir, int spectral) {
    HostEmitter *E = new HostEmitter();
    try {
        E->spectral = spectral; E->n_channels = spectral ? 11 : 3;
        E->data = datasets_for(dir, spectral);
        build_emitter(*E, *p, true);
    } catch (const std::exception &e) { fprintf(stderr, "hostdbg: %s\n", e.what()); delete E; return nullptr; }
    return E;
}
void hostdbg_eval(void *h, size_t n, const float *wi, const float *wav, float *out) {
    HostEmitter &E = *(HostEmitter *) h;
    for (size_t i = 0; i < n; ++i) {
        Vec3 v = { wi[i], wi[n + i], wi[2 * n + i] };
        if (E.spectral) {
            float w[4] = { wav[i], wav[n + i], wav[2 * n + i], wav[3 * n + i] }, o[4];
            if (E.block.arith) eval_spec<1>(E.block, E.block, E_pairs(E), E.block.sun_rad, v, w, true, o); else if (getenv("HOSTDBG_NOPAIRS")) eval_spec<0>(E.block, E.block, nullptr, E.block.sun_rad, v, w, true, o); else eval_spec<0, true>(E.block, E.block, E_xpairs(E), E.block.sun_rad, v, w, true, o);
            for (int c = 0; c < 4; ++c) out[c * n + i] = o[c];
        } else if (getenv("HOSTDBG_PAIR")) {       // the two-rays-per-thread path of eval_rgb_pair_kernel
            if (i & 1) continue;
            bool two = i + 1 < n;
            Vec3p w = { f2(wi[i], two ? wi[i + 1] : 0.f), f2(wi[n + i], two ? wi[n + i + 1] : 0.f), f2(wi[2 * n + i], two ? wi[2 * n + i + 1] : 1.f) };
            float2 o[3];
            if (E.block.arith) eval_rgb_p<1>(E.block, E.block.sun_rad, w, true, two, o); else eval_rgb_p<0>(E.block, E.block.sun_rad, w, true, two, o);
            for (int c = 0; c < 3; ++c) { out[c * n + i] = o[c].x; if (two) out[c * n + i + 1] = o[c].y; }
        } else {
            float o[3];
            if (E.block.arith) eval_rgb<1>(E.block, E.block.sun_rad, v, true, o); else eval_rgb<0>(E.block, E.block.sun_rad, v, true, o);
            for (int c = 0; c < 3; ++c) out[c * n + i] = o[c];
        }
    }
}
void hostdbg_eval_jvp(void *h, float dt, const float *da, int nda, const float *ds, size_t n, const float *wi, const float *wav, float *out) {
    HostEmitter &E = *(HostEmitter *) h;
    static TangentBlock T;
    state_tangent(E, dt, da, nda, ds, T);
    for (size_t i = 0; i < n; ++i) {
        Vec3 v = { wi[i], wi[n + i], wi[2 * n + i] };
        float o[4];
        if (E.spectral) { float w[4] = { wav[i], wav[n + i], wav[2 * n + i], wav[3 * n + i] }; if (E.block.arith) eval_spec_jvp<1>(E.block, E.block, T, T, E.block.sun_rad, v, w, o); else eval_spec_jvp<0>(E.block, E.block, T, T, E.block.sun_rad, v, w, o); }
        else if (E.block.arith) eval_rgb_jvp<1>(E.block, T, E.block.sun_rad, v, o); else eval_rgb_jvp<0>(E.block, T, E.block.sun_rad, v, o);
        for (int c = 0; c < (E.spectral ? 4 : 3); ++c) out[c * n + i] = o[c];
    }
}
void hostdbg_pdf(void *h, size_t n, const float *d, float *pdf) {
    HostEmitter &E = *(HostEmitter *) h;
    if (getenv("HOSTDBG_PAIR")) {
        for (size_t i = 0; i < n; i += 2) {
            bool two = i + 1 < n;
            Vec3p v = { f2(d[i], two ? d[i + 1] : 0.f), f2(d[n + i], two ? d[n + i + 1] : 0.f), f2(d[2 * n + i], two ? d[2 * n + i + 1] : 1.f) };
            float2 r = E.block.arith ? pdf_direction_p<1>(E.block, v, true, two) : pdf_direction_p<0>(E.block, v, true, two);
            pdf[i] = r.x; if (two) pdf[i + 1] = r.y;
        }
        return;
    }
    for (size_t i = 0; i < n; ++i) { Vec3 v{ d[i], d[n + i], d[2 * n + i] }; pdf[i] = E.block.arith ? pdf_direction<1>(E.block, v) : pdf_direction<0>(E.block, v); }
}
void hostdbg_sample(void *h, size_t n, const float *u, const float *wav, float *d, float *pdf, float *weight) {
    HostEmitter &E = *(HostEmitter *) h;
    int C = E.spectral ? 4 : 3;
    if (getenv("HOSTDBG_PAIR")) {
        for (size_t i = 0; i < n; i += 2) {
            bool two = i + 1 < n;
            float2 a = f2(u[i], two ? u[i + 1] : .5f), b = f2(u[n + i], two ? u[n + i + 1] : .5f);
            DirSampleP s = E.block.arith ? sample_dir_p<1>(E.block, E.block, a, b) : sample_dir_p<0>(E.block, E.block, a, b);
            float2 e[4];
            if (E.spectral) {
                float ea[4], eb[4] = { 0, 0, 0, 0 };
                size_t j = two ? i + 1 : i;
                float wa[4] = { wav[i], wav[n + i], wav[2 * n + i], wav[3 * n + i] }, wb[4] = { wav[j], wav[n + j], wav[2 * n + j], wav[3 * n + j] };
                Vec3 va = { -s.d_world.x.x, -s.d_world.y.x, -s.d_world.z.x }, vb = { -s.d_world.x.y, -s.d_world.y.y, -s.d_world.z.y };
                if (E.block.arith) { eval_spec<1>(E.block, E.block, E_pairs(E), E.block.sun_rad, va, wa, s.active0, ea); eval_spec<1>(E.block, E.block, E_pairs(E), E.block.sun_rad, vb, wb, s.active1, eb); }
                else { eval_spec<0>(E.block, E.block, nullptr, E.block.sun_rad, va, wa, s.active0, ea); eval_spec<0>(E.block, E.block, nullptr, E.block.sun_rad, vb, wb, s.active1, eb); }
                for (int c = 0; c < 4; ++c) e[c] = f2(ea[c], eb[c]);
            } else {
                Vec3p wi = { neg2(s.d_world.x), neg2(s.d_world.y), neg2(s.d_world.z) };
                if (E.block.arith) eval_rgb_p<1>(E.block, E.block.sun_rad, wi, s.active0, s.active1, e); else eval_rgb_p<0>(E.block, E.block.sun_rad, wi, s.active0, s.active1, e);
            }
            for (int k = 0; k < (two ? 2 : 1); ++k) {
                size_t o = i + k;
                d[o] = k ? s.d_world.x.y : s.d_world.x.x; d[n + o] = k ? s.d_world.y.y : s.d_world.y.x; d[2 * n + o] = k ? s.d_world.z.y : s.d_world.z.x;
                float pv = k ? s.pdf.y : s.pdf.x; pdf[o] = pv;
                for (int c = 0; c < C; ++c) { float w = (k ? e[c].y : e[c].x) * (1.f / pv); weight[c * n + o] = (fabsf(w) < INFINITY) ? w : 0.f; }
            }
        }
        return;
    }
    for (size_t i = 0; i < n; ++i) {
        DirSample s = E.block.arith ? sample_dir<1>(E.block, E.block, u[i], u[n + i]) : sample_dir<0>(E.block, E.block, u[i], u[n + i]);
        d[i] = s.d_world.x; d[n + i] = s.d_world.y; d[2 * n + i] = s.d_world.z; pdf[i] = s.pdf;
        Vec3 wi = { -s.d_world.x, -s.d_world.y, -s.d_world.z };
        float e[4];
        if (E.spectral) { float w[4] = { wav[i], wav[n + i], wav[2 * n + i], wav[3 * n + i] }; if (E.block.arith) eval_spec<1>(E.block, E.block, E_pairs(E), E.block.sun_rad, wi, w, s.active, e); else eval_spec<0>(E.block, E.block, nullptr, E.block.sun_rad, wi, w, s.active, e); }
        else if (E.block.arith) eval_rgb<1>(E.block, E.block.sun_rad, wi, s.active, e); else eval_rgb<0>(E.block, E.block.sun_rad, wi, s.active, e);
        for (int c = 0; c < C; ++c) { float w = e[c] * (1.f / s.pdf); weight[c * n + i] = (fabsf(w) < INFINITY) ? w : 0.f; }
    }
}
void hostdbg_set_bsphere(void *h, const float *c4) {
    HostEmitter &E = *(HostEmitter *) h;
    for (int i = 0; i < 3; ++i) E.bs_center[i] = E.block.bs_center[i] = c4[i];
    E.bs_radius = E.block.bs_radius = c4[3];
    set_ray_pdf_scale(E.block);
}
void hostdbg_sample_wavelengths(void *h, size_t n, const float *wi, const float *sample, float *wav, float *weight) {
    HostEmitter &E = *(HostEmitter *) h;
    for (size_t i = 0; i < n; ++i) {
        Vec3 v = { wi[i], wi[n + i], wi[2 * n + i] };
        float w[4], q[4];
        if (E.block.arith) sample_wavelengths_spec<1>(E.block, E.block, E_pairs(E), E.block.sun_rad, v, sample[i], true, w, q);
        else sample_wavelengths_spec<0>(E.block, E.block, nullptr, E.block.sun_rad, v, sample[i], true, w, q);
        for (int c = 0; c < 4; ++c) { wav[c * n + i] = w[c]; weight[c * n + i] = q[c]; }
    }
}
void hostdbg_sample_ray(void *h, size_t n, const float *wl, const float *s2, const float *s3, float *o, float *d, float *wav, float *weight) {
    HostEmitter &E = *(HostEmitter *) h;
    int C = E.spectral ? 4 : 3;
    for (size_t i = 0; i < n; ++i) {
        RaySample r = E.block.arith ? sample_ray_geom<1>(E.block, E.block, s2[i], s2[n + i], s3[i], s3[n + i])
                                    : sample_ray_geom<0>(E.block, E.block, s2[i], s2[n + i], s3[i], s3[n + i]);
        o[i] = r.origin.x; o[n + i] = r.origin.y; o[2 * n + i] = r.origin.z;
        d[i] = r.d_world.x; d[n + i] = r.d_world.y; d[2 * n + i] = r.d_world.z;
        float e[4] = { 0, 0, 0, 0 }, w[4] = { 0, 0, 0, 0 };
        if (E.spectral) {
            if (E.block.arith) sample_wavelengths_spec<1>(E.block, E.block, E_pairs(E), E.block.sun_rad, r.d_world, wl[i], r.active, w, e);
            else sample_wavelengths_spec<0>(E.block, E.block, nullptr, E.block.sun_rad, r.d_world, wl[i], r.active, w, e);
            for (int c = 0; c < 4; ++c) wav[c * n + i] = w[c];
        } else {
            if (E.block.arith) eval_rgb<1>(E.block, E.block.sun_rad, r.d_world, r.active, e); else eval_rgb<0>(E.block, E.block.sun_rad, r.d_world, r.active, e);
        }
        for (int c = 0; c < C; ++c) { float q = e[c] * (1.f / r.pdf); weight[c * n + i] = (fabsf(q) < INFINITY) ? q : 0.f; }
    }
}
void hostdbg_math(int op, size_t n, const float *a, const float *b, float *o) {
    for (size_t i = 0; i < n; ++i) {
        float x = a[i];
        switch (op) {
            case 0: o[i] = dsin(x); break; case 1: o[i] = dcos(x); break; case 2: o[i] = dasin(x); break;
            case 3: o[i] = dacos(x); break; case 4: o[i] = datan2(x, b[i]); break; case 5: o[i] = dexp(x); break;
            case 6: o[i] = dexp2(x); break; case 7: o[i] = dlog(x); break; case 8: o[i] = dlog2(x); break;
            case 9: o[i] = dcbrt(x); break; case 10: o[i] = derf(x); break; case 11: o[i] = derfinv(x); break;
            case 12: o[i] = dpow15(x); break; case 13: o[i] = dtan(x); break; case 14: o[i] = sqrtf(x); break;
        }
    }
}
}
