// Per-lane sunsky model: Hosek-Wilkie sky, sun disc with limb darkening, TGMM sampler and pdf.
// Every function states the reference lines it reproduces (src/emitters/sunsky.cpp and
// include/mitsuba/render/sunsky/sunsky.h of matttsss/mitsuba3-sunsky). The op order follows the
// llvm_ad_* variants so results can be compared bit for bit with the oracle.
#pragma once
#include "ssk_math.cuh"

namespace ssk {

constexpr int kGauss = 20;          // 4 mixtures x NB_GAUSSIAN (sunsky.h:49, sunsky.cpp:746)
constexpr int kSunSegments = 45;    // NB_SUN_SEGMENTS (sunsky.h:42)
constexpr int kSkyParams = 9;       // NB_SKY_PARAMS (sunsky.h:37)
constexpr int kLd = 6;              // NB_SUN_LD_PARAMS (sunsky.h:44)
constexpr int kMaxChannels = 11;    // NB_WAVELENGTHS (sunsky.h:20)

/// Per-emitter parameter block (host-built, replicated to the device; ~2.9 KB without the sun table).
/// Layout is SoA per coefficient so lane-divergent channel / component indices hit distinct shared
/// memory banks.
struct EmitterBlock {
    // --- hot part, staged into shared memory by every CTA ---
    float sky_params[kSkyParams][kMaxChannels + 1];  // [param][channel]; channel 11 is the zero pad that
                                                     // stands in for the reference's out-of-bounds
                                                     // gather at 720 nm (sunsky.cpp:329-338)
    float sky_rad[kMaxChannels + 1];
    float sun_ld[kLd][kMaxChannels + 1];             // spectral only
    float g_mu_phi[kGauss], g_mu_theta[kGauss], g_sigma_phi[kGauss], g_sigma_theta[kGauss];
    float g_weight[kGauss], g_volume[kGauss];
    float g_cdf_a_phi[kGauss], g_cdf_b_phi[kGauss], g_cdf_a_theta[kGauss], g_cdf_b_theta[kGauss];
    float g_cdf[kGauss];                             // unnormalised inclusive CDF
    float g_pmf_n[kGauss];                           // pmf[i] * normalization
    float g_cdf_prev_n[kGauss];                      // cdf[i-1] * normalization, 0 for i == 0
    float cdf_sum;
    float to_world[9], to_local[9];                  // row-major 3x3 (linear part of to_world / inverse)
    float sun_n[3], sun_s[3], sun_t[3];              // Frame3f(local sun direction)
    float sun_phi_shift;                             // m_sun_angles.x() - 0.5f * Pi
    float cos_cutoff;                                // dr::cos(m_sun_half_aperture)
    float one_minus_cos_cutoff;
    float ar2;                                       // 1 / sin(half_aperture)^2 (sunsky.h:386-387)
    float area_ratio;                                // get_area_ratio (sunsky.h:98-101)
    float sun_pdf_const;                             // InvTwoPi / (1 - cos_cutoff) (warp.h:567-577)
    float sky_scale, sun_scale;
    float sky_w;                                     // m_sky_sampling_w
    float bs_center[3], bs_radius;
    int32_t spectral;
    int32_t hot_end_marker;
    // --- cold part: only read by lanes inside the sun disc, straight from global / L2 ---
    float sun_rad[kSunSegments * 3 * 4 * kLd];       // RGB: [seg][ch][k][j]; spectral: [seg][11][k]
};
constexpr int kHotFloats = (int) (offsetof(EmitterBlock, sun_rad) / 4);

// ---------------------------------------------------------------------------------------------
SSK_HD Vec3 xform(const float *m, const Vec3 &v) {
    // Transform4f::operator*(Vector) (include/mitsuba/core/transform.h:147-157): column 0 times x,
    // then one fma per remaining column
    return { fmaf(m[2], v.z, fmaf(m[1], v.y, m[0] * v.x)),
             fmaf(m[5], v.z, fmaf(m[4], v.y, m[3] * v.x)),
             fmaf(m[8], v.z, fmaf(m[7], v.y, m[6] * v.x)) };
}

/// Quantities shared by every channel of one ray (sunsky.cpp:547-552)
struct RayTerms {
    float cos_theta, gamma, cos_gamma, cos_gamma_sqr, inv_denom_c1, one_plus_cg2, sqrt_cos_theta;
};
SSK_HD RayTerms ray_terms(float cos_theta, float gamma) {
    RayTerms r;
    r.cos_theta = cos_theta; r.gamma = gamma;
    r.cos_gamma = dcos(gamma);
    r.cos_gamma_sqr = r.cos_gamma * r.cos_gamma;
    // `coefs[1] / (cos_theta + 0.01f)` divides a Spectrum by a Float: Dr.Jit routes an array divided by
    // a shallower operand to a1 * rcp(a2) (drjit/array_router.h:218-232), rcp = 1/x on the LLVM backend
    r.inv_denom_c1 = fdiv(1.f, cos_theta + 0.01f);
    r.one_plus_cg2 = 1.f + r.cos_gamma_sqr;
    r.sqrt_cos_theta = safe_sqrt(cos_theta);
    return r;
}

/// render_sky for one channel (sunsky.cpp:538-555)
SSK_HD float render_sky(const EmitterBlock &P, int ch, const RayTerms &r) {
    float A0 = P.sky_params[0][ch], A1 = P.sky_params[1][ch], A2 = P.sky_params[2][ch],
          A3 = P.sky_params[3][ch], A4 = P.sky_params[4][ch], A5 = P.sky_params[5][ch],
          A6 = P.sky_params[6][ch], A7 = P.sky_params[7][ch], A8 = P.sky_params[8][ch];
    float c1 = 1.f + A0 * dexp(A1 * r.inv_denom_c1);
    float chi = fdiv(r.one_plus_cg2, dpow15(1.f + A8 * A8 - 2.f * A8 * r.cos_gamma));
    float c2 = A2 + A3 * dexp(A4 * r.gamma) + A5 * r.cos_gamma_sqr + A6 * chi + A7 * r.sqrt_cos_theta;
    return c1 * c2 * P.sky_rad[ch];
}

/// compute_cos_psi (sunsky.h:384-392); `1.0 - ...` is evaluated in fp64 by the reference, which
/// rounds to the same fp32 value as the fp32 subtraction
SSK_HD float cos_psi(const EmitterBlock &P, float gamma) {
    float sg = dsin(gamma);
    float sc2 = 1.f - P.ar2 * sg * sg;
    return safe_sqrt(sc2);
}

/// Sun segment lookup shared by the channels of a ray (sunsky.cpp:580-587)
struct SunSeg { uint32_t pos; float x; };
SSK_HD SunSeg sun_segment(float cos_theta) {
    float elevation = SSK_HALF_PI - dacos(cos_theta);
    float f = floorf(dcbrt(2.f * elevation * SSK_INV_PI) * (float) kSunSegments);
    uint32_t pos = (uint32_t) (int32_t) f;
    pos = pos < (uint32_t) (kSunSegments - 1) ? pos : (uint32_t) (kSunSegments - 1);
    float q = (float) pos * (1.f / (float) kSunSegments);   // Float / int -> mul by rcp (array_router.h:225-228)
    float break_x = SSK_HALF_PI * (q * (q * q));   // dr::pow(q, 3.f) = powi: q * (q*q)
    return { pos, elevation - break_x };
}
SSK_HD void powers4(float x, float *p) { p[0] = 1.f; p[1] = x; p[2] = x * x; p[3] = x * p[2]; }
SSK_HD void powers6(float c, float *p) {
    // detail::powi repeated squaring (drjit/math.h:990-1005)
    float c2 = c * c, c4 = c2 * c2;
    p[0] = 1.f; p[1] = c; p[2] = c2; p[3] = c * c2; p[4] = c4; p[5] = c * c4;
}

/// render_sun, RGB branch (sunsky.cpp:597-611)
SSK_HD float render_sun_rgb(const float *sun_rad, int ch, const SunSeg &s, const float *xk, const float *cj) {
    const float *S = sun_rad + s.pos * (3 * 4 * kLd) + ch * (4 * kLd);
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int j = 0; j < kLd; ++j)
            acc += xk[k] * cj[j] * S[k * kLd + j];
    return acc;
}
/// render_sun, spectral branch (sunsky.cpp:590-596)
SSK_HD float render_sun_spec(const float *sun_rad, int ch, const SunSeg &s, const float *xk) {
    const float *S = sun_rad + s.pos * (kMaxChannels * 4) + ch * 4;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        acc += xk[k] * S[k];
    return acc;
}
/// compute_sun_ld (sunsky.cpp:631-650)
SSK_HD float sun_ld(const EmitterBlock &P, int lo, int hi, float f, const float *cj) {
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < kLd; ++j)
        acc += cj[j] * lerp(P.sun_ld[j][lo], P.sun_ld[j][hi], f);
    return acc;
}

/// Geometry of eval(): world -> local, cos(theta), gamma, masks (sunsky.cpp:309-314)
struct EvalGeom { RayTerms r; bool active, hit_sun; };
SSK_HD EvalGeom eval_geom(const EmitterBlock &P, const Vec3 &wi, bool active_in) {
    Vec3 neg = { -wi.x, -wi.y, -wi.z };
    Vec3 wo = xform(P.to_local, neg);
    Vec3 n = { P.sun_n[0], P.sun_n[1], P.sun_n[2] };
    float cos_theta = wo.z, gamma = unit_angle(n, wo);
    EvalGeom g;
    g.active = active_in && cos_theta >= 0.f;
    g.hit_sun = g.active && dot3(n, wo) >= P.cos_cutoff;
    if (g.active) g.r = ray_terms(cos_theta, gamma);
    else { g.r.cos_theta = cos_theta; g.r.gamma = gamma; }
    return g;
}

/// SunskyEmitter::eval, RGB (sunsky.cpp:303-323, 351). The sun term and the CIE-Y normalisation are
/// carried out in fp64 by the reference (`Array<float> * double` promotes, sunsky.cpp:321-323).
SSK_HD void eval_rgb(const EmitterBlock &P, const float *sun_rad, const Vec3 &wi, bool active_in, float out[3]) {
    EvalGeom g = eval_geom(P, wi, active_in);
    if (!g.active) { out[0] = out[1] = out[2] = 0.f; return; }
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; ++c)
        res[c] = P.sky_scale * render_sky(P, c, g.r);
    if (g.hit_sun) {
        SunSeg s = sun_segment(g.r.cos_theta);
        float xk[4], cj[6];
        powers4(s.x, xk);
        powers6(cos_psi(P, g.r.gamma), cj);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float sun = P.sun_scale * render_sun_rgb(sun_rad, c, s, xk, cj) * P.area_ratio;
            res[c] = (float) ((double) res[c] + (double) sun * 467.069280386);
        }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c)
        out[c] = (float) ((double) res[c] * (1.0 / 106.7502593994140625));
}

/// SunskyEmitter::eval, spectral, for the 4 wavelengths of one ray (sunsky.cpp:325-351)
SSK_HD void eval_spec(const EmitterBlock &P, const float *sun_rad, const Vec3 &wi, const float wav[4],
                      bool active_in, float out[4]) {
    EvalGeom g = eval_geom(P, wi, active_in);
    if (!g.active) { out[0] = out[1] = out[2] = out[3] = 0.f; return; }
    SunSeg s; float xk[4], cj[6];
    if (g.hit_sun) {
        s = sun_segment(g.r.cos_theta);
        powers4(s.x, xk);
        powers6(cos_psi(P, g.r.gamma), cj);
    }
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        float nw = (wav[w] - 320.f) * (1.f / 40.f);               // Spectrum / size_t -> mul by rcp(40.f)
        bool valid = 0.f <= nw && nw <= 10.f;
        if (!valid) { out[w] = 0.f; continue; }
        int lo = (int) floorf(nw), hi = lo + 1;
        float f = nw - (float) lo;
        float res = P.sky_scale * lerp(render_sky(P, lo, g.r), render_sky(P, hi, g.r), f);
        if (g.hit_sun) {
            float sun_lo = render_sun_spec(sun_rad, lo, s, xk),
                  sun_hi = hi < kMaxChannels ? render_sun_spec(sun_rad, hi, s, xk) : 0.f,
                  sun = lerp(sun_lo, sun_hi, f);
            float ld = sun_ld(P, lo, hi, f, cj);
            res += P.sun_scale * sun * ld * P.area_ratio;
        }
        out[w] = res;
    }
}

// ---------------------------------------------------------------------------------------------
/// tgmm_pdf (sunsky.cpp:732-763). mu/sigma/weight and `volume` are per-emitter constants that the
/// JIT folds on the host (m_gaussians[i] is a scalar read); `volume` is precomputed with the same ops.
SSK_HD float tgmm_pdf(const EmitterBlock &P, float phi, float theta, bool active) {
    phi -= P.sun_phi_shift;
    phi = phi < 0.f ? phi + SSK_TWO_PI : phi;
    phi = phi > SSK_TWO_PI ? phi - SSK_TWO_PI : phi;
    active = active && theta >= 0.f && theta <= SSK_HALF_PI;
    if (!active) return 0.f;
    float pdf = 0.f;
#pragma unroll 4
    for (int i = 0; i < kGauss; ++i) {
        float sx = fdiv(phi - P.g_mu_phi[i], P.g_sigma_phi[i]),
              sy = fdiv(theta - P.g_mu_theta[i], P.g_sigma_theta[i]);
        float g = SSK_INV_TWO_PI * dexp(-.5f * fmaf(sy, sy, sx * sx));   // warp.h:189-191
        pdf += fdiv(P.g_weight[i] * g, P.g_volume[i]);
    }
    return pdf;
}

/// compute_pdfs (sunsky.cpp:711-723) followed by the lerp of pdf_direction / sample_direction
SSK_HD float combined_pdf(const EmitterBlock &P, const Vec3 &d, bool check_sun, bool active) {
    float sin_theta = safe_sqrt(fmaf(d.x, d.x, d.y * d.y));        // Frame3f::sin_theta (frame.h:59-65)
    active = active && d.z >= 0.f && sin_theta != 0.f;
    sin_theta = fmaxf(sin_theta, SSK_EPSILON);
    float phi = datan2(d.y, d.x), theta = unit_angle_z(d);          // from_spherical (sunsky.h:83-89)
    float sky_pdf = fdiv(tgmm_pdf(P, phi, theta, active), sin_theta);
    Vec3 n = { P.sun_n[0], P.sun_n[1], P.sun_n[2] };
    float sun_pdf = (!check_sun || dot3(n, d) >= P.cos_cutoff) ? P.sun_pdf_const : 0.f;
    return lerp(sun_pdf, sky_pdf, P.sky_w);
}

/// pdf_direction (sunsky.cpp:443-451)
SSK_HD float pdf_direction(const EmitterBlock &P, const Vec3 &d_world) {
    return combined_pdf(P, xform(P.to_local, d_world), true, true);
}

/// DiscreteDistribution::sample_reuse, JIT flavour (distr_1d.h:116-134,172-182; drjit/util.h:136-181)
SSK_HD int tgmm_pick(const EmitterBlock &P, float u, float &reuse) {
    float s = u * P.cdf_sum;
    uint32_t start = 0, end = kGauss - 1;
#pragma unroll
    for (int it = 0; it < 5; ++it) {                 // log2i(19) + 1 iterations
        uint32_t middle = (start + end) >> 1;
        float v = P.g_cdf[middle];
        bool cond = ((v < s) || v == 0.f) && v != P.cdf_sum;
        uint32_t nstart = middle + 1 < end ? middle + 1 : end;
        start = cond ? nstart : start;
        end = cond ? end : middle;
    }
    reuse = fdiv(u - P.g_cdf_prev_n[start], P.g_pmf_n[start]);
    return (int) start;
}

/// sample_sky (sunsky.cpp:661-689). cdf_a / cdf_b depend only on the component and are precomputed.
SSK_HD Vec3 sample_sky(const EmitterBlock &P, float u0, float u1) {
    float t;
    int i = tgmm_pick(P, u0, t);
    float sx = lerp(P.g_cdf_a_phi[i], P.g_cdf_b_phi[i], t),
          sy = lerp(P.g_cdf_a_theta[i], P.g_cdf_b_theta[i], u1);
    sx = fmaxf(fminf(sx, SSK_ONE_MINUS_EPSILON), SSK_EPSILON);
    sy = fmaxf(fminf(sy, SSK_ONE_MINUS_EPSILON), SSK_EPSILON);
    float phi = SSK_SQRT_TWO * derfinv(2.f * sx - 1.f) * P.g_sigma_phi[i] + P.g_mu_phi[i],
          theta = SSK_SQRT_TWO * derfinv(2.f * sy - 1.f) * P.g_sigma_theta[i] + P.g_mu_theta[i];
    phi += P.sun_phi_shift;
    theta = fminf(theta, SSK_HALF_PI - SSK_EPSILON);
    return sphdir(theta, phi);
}

/// sample_sun (sunsky.cpp:697-701): warp::square_to_uniform_cone via the concentric disk mapping
/// (warp.h:54-90, 533-551), rotated into the sun frame (frame.h:39-41)
SSK_HD Vec3 sample_sun(const EmitterBlock &P, float u0, float u1) {
    float x = fmaf(2.f, u0, -1.f), y = fmaf(2.f, u1, -1.f);
    bool is_zero = x == 0.f && y == 0.f, q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = fdiv(0.25f * SSK_PI * rp, r);
    if (q13) phi = SSK_HALF_PI - phi;
    if (is_zero) phi = 0.f;
    float s, c;
    dsincos(phi, s, c);
    float px = r * c, py = r * s;
    float pn = fmaf(py, py, px * px);
    float z = P.cos_cutoff + P.one_minus_cos_cutoff * (1.f - pn);
    float k = safe_sqrt(P.one_minus_cos_cutoff * (2.f - P.one_minus_cos_cutoff * pn));
    px *= k; py *= k;
    return { fmaf(P.sun_n[0], z, fmaf(P.sun_t[0], py, P.sun_s[0] * px)),
             fmaf(P.sun_n[1], z, fmaf(P.sun_t[1], py, P.sun_s[1] * px)),
             fmaf(P.sun_n[2], z, fmaf(P.sun_t[2], py, P.sun_s[2] * px)) };
}

/// Direction part of sample_direction (sunsky.cpp:404-419, 434-436)
struct DirSample { Vec3 d_world; float pdf; bool active; };
SSK_HD DirSample sample_dir(const EmitterBlock &P, float u0, float u1) {
    bool pick_sky = u0 < P.sky_w;
    Vec3 dir = pick_sky ? sample_sky(P, fdiv(u0, P.sky_w), u1)
                        : sample_sun(P, fdiv(u0 - P.sky_w, 1.f - P.sky_w), u1);
    DirSample r;
    r.active = dir.z >= 0.f;
    r.d_world = xform(P.to_world, dir);
    r.pdf = combined_pdf(P, dir, pick_sky, r.active);
    return r;
}

} // namespace ssk
