// Per-lane sunsky model: Hosek-Wilkie sky, sun disc with limb darkening, TGMM sampler and pdf.
// Every function states the reference lines it reproduces (src/emitters/sunsky.cpp and
// include/mitsuba/render/sunsky/sunsky.h of matttsss/mitsuba3-sunsky).
//
// Two arithmetic modes (ssk_params.arith):
//   EXACT  the op order of the llvm_ad_* variants, so results compare bit for bit with the oracle.
//   FAST   same formulas; everything a result is ill-conditioned in (the geometry: cos(theta), gamma,
//          cos(gamma), phi, theta, the 1 + A8^2 - 2 A8 cos(gamma) cancellation, the sampled direction) is
//          still computed with the exact sequence, while the well-conditioned per-channel / per-Gaussian
//          transcendentals (exp, pow 1.5, divisions) go to the MUFU unit. Measured deviation from the
//          oracle: see tests/test_parity_gpu.py::test_fast_mode_* and DESIGN.md.
//
// Functions take the emitter block twice where it matters: `U` is read with compile-time indices
// only (in the kernels it is the __grid_constant__ parameter, so coefficients become constant-bank
// operands of the FFMA/FMUL that use them), `S` is read with lane-dependent indices (in the kernels: a
// shared-memory copy). On the host both are the same object.
#pragma once
#include "ssk_math.cuh"
#include <cstddef>

namespace ssk {

constexpr int kGauss = 20;          // 4 mixtures x NB_GAUSSIAN (sunsky.h:49, sunsky.cpp:746)
constexpr int kSunSegments = 45;    // NB_SUN_SEGMENTS (sunsky.h:42)
constexpr int kSkyParams = 9;       // NB_SKY_PARAMS (sunsky.h:37)
constexpr int kLd = 6;              // NB_SUN_LD_PARAMS (sunsky.h:44)
constexpr int kMaxChannels = 11;    // NB_WAVELENGTHS (sunsky.h:20)
constexpr int kChanPad = kMaxChannels + 1;
// FAST-mode coefficient table, per channel: c1 coefficients (T0, T1, T11: see c1_fast), A2, A3, A4 log2e, A5, A6, A7,
// 1 + A8^2, 2 A8, radiance x scale
constexpr int kFastCoefs = 12;

// How FAST mode evaluates c1 = 1 + A0 exp(A1 / (cos(theta) + 0.01)), which cancels to ~0.1 towards the zenith:
//   0  the reference's exp polynomial with a cheaper range reduction (fexp): 24 instructions per bin pair
//   1  c1 = fma(A0, MUFU ex2 of a pre-scaled exponent, 1), guarded: |c1| < 1/16 (error amplification > 16, A0 e ~ -1)
//      -> reference path
//   2  u = A1 s + ln|A0|; c1 = -expm1(u) by a degree-6 polynomial where A0 < 0 and |u| < 1/4 (no cancellation at
//      all: the small quantity is computed directly), 1 + sgn(A0) 2^(u log2 e) on the MUFU elsewhere (c1 >= 0.22
//      there, amplification < 5). Select, not branch. No fallback needed.
#ifndef SSK_C1_MODE
#define SSK_C1_MODE 1
#endif
// 1: FAST spectral coefficient table in shared memory as 128-bit rows {T_2q(bin), T_2q(bin+1), T_2q+1(bin), T_2q+1(bin+1)},
// 6 LDS.128 per wavelength instead of 12 LDS.64. Measured SLOWER (spectral eval 54.9 vs 65.0 Gevals/s with c1 mode 1):
// lanes of a quarter-warp whose bins differ by 8 collide, and the kernel is latency-bound on exactly these loads.
#ifndef SSK_QUAD_TABLE
#define SSK_QUAD_TABLE 0
#endif

enum { kExact = 0, kFast = 1 };

/// Per-emitter parameter block without the sun table (host-built, ~3.4 KB). Passed to every kernel by
/// value as a __grid_constant__ parameter and mirrored in global memory for the shared-memory copies.
struct HotBlock {
    // --- sky: [param][channel]; channel 11 is the zero pad that stands in for the reference's
    //     out-of-bounds gather at 720 nm (sunsky.cpp:329-338)
    float sky_params[kSkyParams][kChanPad];
    float sky_rad[kChanPad];
    float sun_ld[kLd][kChanPad];                     // spectral only
    float sky_fast[kFastCoefs][kChanPad];            // FAST mode coefficients (see sky_channel_fast)
    // --- TGMM, reference component order: the sampler (lane-dependent component index)
    float g_mu_phi[kGauss], g_mu_theta[kGauss], g_sigma_phi[kGauss], g_sigma_theta[kGauss];
    float g_cdf_a_phi[kGauss], g_cdf_b_phi[kGauss], g_cdf_a_theta[kGauss], g_cdf_b_theta[kGauss];
    float g_cdf[kGauss];                             // unnormalised inclusive CDF
    float g_pmf_n[kGauss], g_rpmf_n[kGauss];         // pmf[i] * normalization and RN(1 / that)
    float g_cdf_prev_n[kGauss];                      // cdf[i-1] * normalization, 0 for i == 0
    // --- TGMM pdf: components with weight 0 dropped (they add +0 to the density), order kept
    int32_t n_pdf;
    int32_t pdf_generic;                             // 1: a reciprocal is not finite, keep the generic divisions
    float p_mu_phi[kGauss], p_mu_theta[kGauss];
    float p_sigma_phi[kGauss], p_sigma_theta[kGauss], p_rsigma_phi[kGauss], p_rsigma_theta[kGauss];
    float p_weight[kGauss], p_volume[kGauss], p_rvolume[kGauss];
    float p_k_phi[kGauss], p_k_theta[kGauss];        // FAST: sqrt(log2(e) / 2) / sigma
    float p_wv[kGauss];                              // FAST: weight / (2 pi volume)
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
    float sky_w, r_sky_w;                            // m_sky_sampling_w and RN(1 / w)
    float one_minus_sky_w, r_one_minus_sky_w;
    float bs_center[3], bs_radius;
    int32_t spectral;
    int32_t arith;
    // EXACT mode's per-ray guard (eval_spec / render_sky_2<true>): a ray with 0 <= 1 / (cos(theta) + 0.01) <= exact_lim_s
    // and gamma <= exact_lim_gamma keeps every exponential argument of every channel below 86 and, for the gamma
    // term, above -86 (no overflow / underflow select can apply, results normal; the c1 exponent is clamped from
    // below, see render_sky_2), and the host has checked |A0| <= 1e20 and that 1 + A8^2 - 2 A8 cos(gamma) stays in
    // [2^-39, 2^39] for every channel (no special case of pow, quotient by it normal). Both -1 when that fails.
    float exact_lim_s, exact_lim_gamma;
    int32_t pad_[2];
    // --- wavelength importance sampling: ContinuousDistribution over [360, 720] nm, 10 nodes
    //     (sunsky.cpp:869-881, distr_1d.h:286-593); RGB emitters leave it zeroed
    float wl_pdf[10], wl_cdf[10];                    // unnormalised node densities, cumulative trapezoid sums (9 used)
    float wl_integral, wl_norm;                      // cdf[valid.y], 1 / that
    float wl_interval, wl_inv_interval;              // 40 nm, 1 / 40
    int32_t wl_valid_lo, wl_valid_hi, wl_iters;      // m_valid and the binary search trip count
    float ray_pdf_scale;                             // InvPi * rcp(radius^2) of sample_ray (sunsky.cpp:381)
    int32_t pad2_[4];
    // per interval i of the wavelength density: RN(1 / (pdf[i] - pdf[i+1])) and RN(1 / pdf[i]), the two IEEE quotients
    // ContinuousDistribution::sample_pdf forms per sample (distr_1d.h:489-493), taken once on the host (same operands,
    // same correctly rounded result, infinities included)
    float wl_rdiff[10], wl_ry0[10];
};

/// HotBlock followed by the sun radiance table, which only lanes inside the sun disc read (from L2).
struct EmitterBlock : HotBlock {
    float sun_rad[kSunSegments * 3 * 4 * kLd];       // RGB: [seg][ch][k][j]; spectral: [seg][11][k]
};
static_assert(sizeof(HotBlock) % 16 == 0, "HotBlock is copied with 16-byte accesses; sun_rad must start 16-byte aligned");
static_assert(sizeof(HotBlock) <= 3800, "HotBlock must fit the 4 KB kernel parameter space with the plane pointers");

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
template <int A> SSK_HD RayTerms ray_terms(float cos_theta, float gamma) {
    RayTerms r;
    r.cos_theta = cos_theta; r.gamma = gamma;
    r.cos_gamma = dcos(gamma);
    r.cos_gamma_sqr = r.cos_gamma * r.cos_gamma;
    // `coefs[1] / (cos_theta + 0.01f)` divides a Spectrum by a Float: Dr.Jit routes an array divided by
    // a shallower operand to a1 * rcp(a2) (drjit/array_router.h:218-232), rcp = 1/x on the LLVM backend
    // (only evaluated for active rays, cos_theta >= 0: the divisor lies in [0.01, 1.01] and fdiv_normal is the IEEE
    // quotient without the range test)
    r.inv_denom_c1 = fdiv_normal(1.f, cos_theta + 0.01f);
    r.one_plus_cg2 = 1.f + r.cos_gamma_sqr;
    r.sqrt_cos_theta = A == kFast ? hw_sqrt(fmaxf(cos_theta, 0.f)) : safe_sqrt(cos_theta);
    return r;
}

/// render_sky behind a call: the cancellation fallback of FAST mode and the sun-disc lanes
SSK_COLD float render_sky_cold(const HotBlock &P, int ch, RayTerms r);

/// FAST ray terms -> the reference's (only the square root differs)
SSK_HD RayTerms exact_terms(RayTerms r) { r.sqrt_cos_theta = safe_sqrt(r.cos_theta); return r; }

/// render_sky for one channel (sunsky.cpp:538-555), reference op order
SSK_HD float render_sky(const HotBlock &P, int ch, const RayTerms &r) {
    float A0 = P.sky_params[0][ch], A1 = P.sky_params[1][ch], A2 = P.sky_params[2][ch],
          A3 = P.sky_params[3][ch], A4 = P.sky_params[4][ch], A5 = P.sky_params[5][ch],
          A6 = P.sky_params[6][ch], A7 = P.sky_params[7][ch], A8 = P.sky_params[8][ch];
    float c1 = 1.f + A0 * dexp(A1 * r.inv_denom_c1);
    float chi = fdiv(r.one_plus_cg2, dpow15(1.f + A8 * A8 - 2.f * A8 * r.cos_gamma));
    float c2 = A2 + A3 * dexp(A4 * r.gamma) + A5 * r.cos_gamma_sqr + A6 * chi + A7 * r.sqrt_cos_theta;
    return c1 * c2 * P.sky_rad[ch];
}

SSK_COLD float render_sky_cold(const HotBlock &P, int ch, RayTerms r) { return render_sky(P, ch, r); }
/// render_sky for channels ch and ch + 1 at once (the two bins a wavelength falls between): the reference
/// op order on float2, each element bit-identical to render_sky(). The one IEEE division stays per element.
template <bool Safe = false> SSK_HD float2 render_sky_2(const HotBlock &P, int ch, const RayTerms &r);
template <bool Safe> SSK_HD float2 render_sky_2(const HotBlock &P, int ch, const RayTerms &r) {
    const float (*A)[kChanPad] = P.sky_params;
    float2 A0 = f2(A[0][ch], A[0][ch + 1]), A1 = f2(A[1][ch], A[1][ch + 1]), A2 = f2(A[2][ch], A[2][ch + 1]),
           A3 = f2(A[3][ch], A[3][ch + 1]), A4 = f2(A[4][ch], A[4][ch + 1]), A5 = f2(A[5][ch], A[5][ch + 1]),
           A6 = f2(A[6][ch], A[6][ch + 1]), A7 = f2(A[7][ch], A[7][ch + 1]), A8 = f2(A[8][ch], A[8][ch + 1]);
    // every product that feeds a sum goes through add2x (see there: ptxas would contract it otherwise)
    // Safe: see the pair-table overload below (same guard, same clamp)
    float2 x1 = mul2s(A1, r.inv_denom_c1);
    if (Safe) x1 = f2(fmaxf(x1.x, -80.f), fmaxf(x1.y, -80.f));
    float2 c1 = add2x(mul2(A0, dexp_2<Safe>(x1)), splat2(1.f));
    // 1 + A8 * A8 - 2 * A8 * cos_gamma: ((1 + A8 A8) - ((2 A8) cos_gamma))
    float2 arg = add2x(neg2(mul2s(mul2s(A8, 2.f), r.cos_gamma)), add2x(mul2(A8, A8), splat2(1.f)));
    float2 pw = Safe ? dpow15_2_safe(arg) : dpow15_2(arg);
    float2 chi = Safe ? fdiv_normal_2(splat2(r.one_plus_cg2), pw) : f2(fdiv(r.one_plus_cg2, pw.x), fdiv(r.one_plus_cg2, pw.y));
    // A2 + A3 exp(A4 gamma) + A5 cos^2 + A6 chi + A7 sqrt(cos_theta), summed left to right
    float2 c2 = add2x(mul2(A3, dexp_2<Safe>(mul2s(A4, r.gamma))), A2);
    c2 = add2x(mul2s(A5, r.cos_gamma_sqr), c2);
    c2 = add2x(mul2(A6, chi), c2);
    c2 = add2x(mul2s(A7, r.sqrt_cos_theta), c2);
    return mul2(mul2(c1, c2), f2(P.sky_rad[ch], P.sky_rad[ch + 1]));
}


/// c1 of FAST mode on a pair (SSK_C1_MODE above). T0, T1, T11 per mode:
///   0: A0, A1, -            1: A0, A1 log2(e), -            2: sgn(A0), A1, ln|A0| (-1e30 for A0 == 0)
/// `s` = 1 / (cos(theta) + 0.01) on both elements (bin pair) or per element (ray pair).
constexpr float kC1Guard = 0.0625f;
SSK_HD float2 c1_fast_2(float2 T0, float2 T1, float2 T11, float2 s, bool &ok) {
#if SSK_C1_MODE == 0
    (void) T11;
    return add2x(mul2(T0, fexp_2(mul2(T1, s), ok)), splat2(1.f));
#elif SSK_C1_MODE == 1
    (void) T11;
    float2 a = mul2(T1, s);
    float2 c1 = fma2(T0, f2(hw_ex2(a.x), hw_ex2(a.y)), splat2(1.f));
    ok = fabsf(c1.x) >= kC1Guard && fabsf(c1.y) >= kC1Guard;
    return c1;
#else
    ok = true;
    float2 u = fma2(T1, s, T11);
    float2 q = fma2(u, splat2(1.f / 720.f), splat2(1.f / 120.f));
    q = fma2(q, u, splat2(1.f / 24.f));
    q = fma2(q, u, splat2(1.f / 6.f));
    q = fma2(q, u, splat2(0.5f));
    q = fma2(q, u, splat2(1.f));
    float2 small = mul2(neg2(u), q);                              // -expm1(u)
    float2 a = mul2s(u, SSK_INV_LOG_TWO);
    float2 big = fma2(T0, f2(hw_ex2(a.x), hw_ex2(a.y)), splat2(1.f));
    return f2((T0.x < 0.f && fabsf(u.x) < 0.25f) ? small.x : big.x, (T0.y < 0.f && fabsf(u.y) < 0.25f) ? small.y : big.y);
#endif
}
SSK_HD float c1_fast(float T0, float T1, float T11, float s, bool &ok) {
#if SSK_C1_MODE == 0
    (void) T11;
    return T0 * fexp(T1 * s, ok) + 1.f;
#elif SSK_C1_MODE == 1
    (void) T11;
    float c1 = fmaf(T0, hw_ex2(T1 * s), 1.f);
    ok = fabsf(c1) >= kC1Guard;
    return c1;
#else
    ok = true;
    float u = fmaf(T1, s, T11);
    float q = fmaf(u, 1.f / 720.f, 1.f / 120.f);
    q = fmaf(q, u, 1.f / 24.f);
    q = fmaf(q, u, 1.f / 6.f);
    q = fmaf(q, u, 0.5f);
    q = fmaf(q, u, 1.f);
    float small = -u * q;
    float big = fmaf(T0, hw_ex2(u * SSK_INV_LOG_TWO), 1.f);
    return (T0 < 0.f && fabsf(u) < 0.25f) ? small : big;
#endif
}

/// render_sky_2 with the coefficient pairs {A_k(ch), A_k(ch + 1)} read as ONE 64-bit value each from a pair table
/// (EXACT spectral kernels, exact_pair_index): 10 loads per wavelength instead of 20 scalar ones plus their address
/// arithmetic. Same operations in the same order as render_sky_2(P, ch, r).
constexpr int kExactPairRows = kSkyParams + 1;           // 9 coefficients + mean radiance
constexpr int kExactPairTable = kExactPairRows * kChanPad;
/// Safe = true (HotBlock::exact_lim_*: the ray passed the guard): the exponentials skip their range selects, the power
/// its special cases, and the two quotients come from fdiv_normal_2. Every value is the one the generic sequence gives.
template <bool Safe = false> SSK_HD float2 render_sky_2(const float2 *xp, int ch, const RayTerms &r) {
    float2 A0 = xp[0 * kChanPad + ch], A1 = xp[1 * kChanPad + ch], A2 = xp[2 * kChanPad + ch], A3 = xp[3 * kChanPad + ch],
           A4 = xp[4 * kChanPad + ch], A5 = xp[5 * kChanPad + ch], A6 = xp[6 * kChanPad + ch], A7 = xp[7 * kChanPad + ch],
           A8 = xp[8 * kChanPad + ch], L = xp[9 * kChanPad + ch];
    // Safe: the exponent is clamped at -80 from below. Below that the exponential is < 2e-35 either way and, with
    // |A0| <= 1e20 (checked on the host), c1 = 1 + A0 e rounds to exactly 1 in both sequences; without the clamp the
    // ~1 % of rays next to the horizon would fail the guard and, scattered over the warps, drag a fifth of them
    // through the generic sequence
    float2 x1 = mul2s(A1, r.inv_denom_c1);
    if (Safe) x1 = f2(fmaxf(x1.x, -80.f), fmaxf(x1.y, -80.f));
    float2 c1 = add2x(mul2(A0, dexp_2<Safe>(x1)), splat2(1.f));
    float2 arg = add2x(neg2(mul2s(mul2s(A8, 2.f), r.cos_gamma)), add2x(mul2(A8, A8), splat2(1.f)));
    float2 pw = Safe ? dpow15_2_safe(arg) : dpow15_2(arg);
    float2 chi = Safe ? fdiv_normal_2(splat2(r.one_plus_cg2), pw) : f2(fdiv(r.one_plus_cg2, pw.x), fdiv(r.one_plus_cg2, pw.y));
    float2 c2 = add2x(mul2(A3, dexp_2<Safe>(mul2s(A4, r.gamma))), A2);
    c2 = add2x(mul2s(A5, r.cos_gamma_sqr), c2);
    c2 = add2x(mul2(A6, chi), c2);
    c2 = add2x(mul2s(A7, r.sqrt_cos_theta), c2);
    return mul2(mul2(c1, c2), L);
}
SSK_HD void build_exact_pair_table(const HotBlock &B, float2 *xp) {
    for (int k = 0; k < kExactPairRows; ++k)
        for (int c = 0; c < kChanPad; ++c) {
            const float *row = k < kSkyParams ? B.sky_params[k] : B.sky_rad;
            xp[k * kChanPad + c] = f2(row[c], c + 1 < kChanPad ? row[c + 1] : 0.f);
        }
}

/// render_sky for one channel, FAST. `T` is one column of HotBlock::sky_fast (already gathered for
/// lane-dependent channels); the radiance scale (mean radiance x sky_scale and, in RGB mode, the CIE-Y
/// normalisation) is folded into the five c2 coefficients T2, T3, T5, T6, T7 on the host, T[10] keeps it for
/// reference. What is approximated and what is not follows the conditioning of the formula:
///  * c1 = 1 + A0 exp(A1 / (cos(theta) + 0.01)) has A0 ~ -1.1: it cancels to ~0.1 over most of the sky, a
///    10-20x amplification of any error in the exponential. SSK_C1_MODE selects how it is evaluated (default: MUFU
///    ex2 and one FMA, with a fallback to the reference sequence where |c1| < 1/16).
///  * the argument of the -3/2 power keeps the reference's two roundings (it cancels to ~(1 - A8)^2 towards
///    the sun); the power itself is rsqrt^3.
///  * c2 is a sum of terms of either sign that is well conditioned except for a low sun near the horizon,
///    where the radiance passes through zero and no approximation keeps a relative tolerance. The sum
///    carries its absolute-value twin (every variable is >= 0, so |T x| = |T| x) and `ok` is cleared when
///    |c2| falls below kCancel of it; the caller then evaluates this channel with the reference sequence.
constexpr float kCancel = 0.125f;
/// All three variants (scalar, bin pair, ray pair) run ONE explicit op sequence, so they agree bit for bit:
///   c1 = fma(T0, 2^(T1 / (cos(theta) + .01)), 1)   one rounding (SSK_C1_MODE 1; it cancels to ~0.1, guarded by |c1| >= 1/16)
///   t  = T8 - T9 cos(gamma)                        two roundings, as the reference (it cancels towards the sun)
///   c2 = fma chain T2 + T3 2^(T4 gamma) + T5 cos^2(gamma) + T6 chi + T7 sqrt(cos(theta)), s2 its |.| twin
SSK_HD float sky_channel_fast(const float T[kFastCoefs], const RayTerms &r, bool &ok) {
    bool ok_e;
    float c1 = c1_fast(T[0], T[1], T[11], r.inv_denom_c1, ok_e);
    float t = T[8] - T[9] * r.cos_gamma;
    float rs = hw_rsqrt(t);
    float chi = (rs * rs * rs) * r.one_plus_cg2;
    float ex = hw_ex2(T[4] * r.gamma);
    float c2 = fmaf(r.sqrt_cos_theta, T[7], fmaf(chi, T[6], fmaf(r.cos_gamma_sqr, T[5], fmaf(ex, T[3], T[2]))));
    float s2 = fmaf(r.sqrt_cos_theta, fabsf(T[7]), fmaf(chi, fabsf(T[6]), fmaf(r.cos_gamma_sqr, fabsf(T[5]), fmaf(ex, fabsf(T[3]), fabsf(T[2])))));
    ok = ok_e && fabsf(c2) >= kCancel * s2;
#if defined(SSK_COUNT_FALLBACK) && !defined(__CUDA_ARCH__)
    extern long long g_fast_calls, g_fast_fallbacks;
    g_fast_calls++; g_fast_fallbacks += !ok;
#endif
    return c1 * c2;
}

/// sky_channel_fast on a pair of channels (the two neighbouring bins of a wavelength, two RGB channels):
/// the same op sequence on float2, so each element equals the scalar function's result bit for bit.
/// T[k] = { coefficient k of channel a, of channel b }. ok = neither element cancelled.
SSK_HD float2 sky_channel_fast_2(const float2 T[kFastCoefs], const RayTerms &r, bool &ok) {
    bool ok_e;
    float2 c1 = c1_fast_2(T[0], T[1], T[11], splat2(r.inv_denom_c1), ok_e);
    float2 t = subp2(T[8], mul2s(T[9], r.cos_gamma));
    float2 rs = f2(hw_rsqrt(t.x), hw_rsqrt(t.y));
    float2 chi = mul2s(mul2(mul2(rs, rs), rs), r.one_plus_cg2);
    float2 a = mul2s(T[4], r.gamma);
    float2 ex = f2(hw_ex2(a.x), hw_ex2(a.y));
    float2 c2 = fma2(splat2(r.sqrt_cos_theta), T[7], fma2(chi, T[6], fma2(splat2(r.cos_gamma_sqr), T[5], fma2(ex, T[3], T[2]))));
    float2 s2 = fma2(splat2(r.sqrt_cos_theta), abs2(T[7]), fma2(chi, abs2(T[6]), fma2(splat2(r.cos_gamma_sqr), abs2(T[5]), fma2(ex, abs2(T[3]), abs2(T[2])))));
    ok = ok_e && fabsf(c2.x) >= kCancel * s2.x && fabsf(c2.y) >= kCancel * s2.y;
    return mul2(c1, c2);
}

/// sky_channel_fast_2 for the spectral eval loop: instead of a flag per bin pair it keeps two running margins for the
/// whole ray -- g1 = min |c1| (c1 = 1 + p cancels exactly when |c1| is small: the ray is fine if g1 >= kC1Guard) and
/// g2 = min (|c2| - kCancel s2) (fine if g2 >= 0) -- so that the per-pair compares and their scale multiplies become one
/// min each and the two compares happen once per ray. A NaN margin is dropped by the min: such a ray does not fall back,
/// and its fast result is the NaN the reference sequence would propagate as well.
SSK_HD float2 sky_channel_fast_2g(const float2 T[kFastCoefs], const RayTerms &r, float &g1, float &g2) {
#if SSK_C1_MODE == 1
    float2 a1 = mul2(T[1], splat2(r.inv_denom_c1));
    float2 c1 = fma2(T[0], f2(hw_ex2(a1.x), hw_ex2(a1.y)), splat2(1.f));
    g1 = fminf(g1, fminf(fabsf(c1.x), fabsf(c1.y)));
#else
    bool ok_e;
    float2 c1 = c1_fast_2(T[0], T[1], T[11], splat2(r.inv_denom_c1), ok_e);
    if (!ok_e) g1 = 0.f;
#endif
    float2 t = subp2(T[8], mul2s(T[9], r.cos_gamma));
    float2 rs = f2(hw_rsqrt(t.x), hw_rsqrt(t.y));
    float2 chi = mul2s(mul2(mul2(rs, rs), rs), r.one_plus_cg2);
    float2 a = mul2s(T[4], r.gamma);
    float2 ex = f2(hw_ex2(a.x), hw_ex2(a.y));
    float2 c2 = fma2(splat2(r.sqrt_cos_theta), T[7], fma2(chi, T[6], fma2(splat2(r.cos_gamma_sqr), T[5], fma2(ex, T[3], T[2]))));
    float2 s2 = fma2(splat2(r.sqrt_cos_theta), abs2(T[7]), fma2(chi, abs2(T[6]), fma2(splat2(r.cos_gamma_sqr), abs2(T[5]), fma2(ex, abs2(T[3]), abs2(T[2])))));
    float2 d = fma2(s2, splat2(-kCancel), abs2(c2));
    g2 = fminf(g2, fminf(d.x, d.y));
    return mul2(c1, c2);
}

/// compute_cos_psi (sunsky.h:384-392); `1.0 - ...` is evaluated in fp64 by the reference, which
/// rounds to the same fp32 value as the fp32 subtraction
SSK_HD float cos_psi(const HotBlock &P, float gamma) {
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

/// render_sun, RGB branch (sunsky.cpp:597-611). The 24 coefficients of a (segment, channel) pair are
/// contiguous and 16-byte aligned: six 128-bit loads.
/// Global = false: the table is not in global memory (the tangent block of eval_jvp is a kernel parameter)
template <bool Global = true>
SSK_HD float render_sun_rgb(const float *sun_rad, int ch, const SunSeg &s, const float *xk, const float *cj) {
    const float *S = sun_rad + s.pos * (3 * 4 * kLd) + ch * (4 * kLd);
    float c[4 * kLd];
#if defined(__CUDA_ARCH__)
    if (Global) {
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            float4 v = __ldg(reinterpret_cast<const float4 *>(S) + q);
            c[4 * q] = v.x; c[4 * q + 1] = v.y; c[4 * q + 2] = v.z; c[4 * q + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int q = 0; q < 4 * kLd; ++q) c[q] = S[q];
    }
#else
    for (int q = 0; q < 4 * kLd; ++q) c[q] = S[q];
#endif
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int j = 0; j < kLd; ++j)
            acc += xk[k] * cj[j] * c[k * kLd + j];
    return acc;
}
/// render_sun, spectral branch (sunsky.cpp:590-596): one 128-bit load per (segment, wavelength bin)
template <bool Global = true>
SSK_HD float render_sun_spec(const float *sun_rad, int ch, const SunSeg &s, const float *xk) {
    const float *S = sun_rad + s.pos * (kMaxChannels * 4) + ch * 4;
#if defined(__CUDA_ARCH__)
    float4 v = Global ? __ldg(reinterpret_cast<const float4 *>(S)) : make_float4(S[0], S[1], S[2], S[3]);
    const float c[4] = { v.x, v.y, v.z, v.w };
#else
    const float c[4] = { S[0], S[1], S[2], S[3] };
#endif
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        acc += xk[k] * c[k];
    return acc;
}
/// compute_sun_ld (sunsky.cpp:631-650)
SSK_HD float sun_ld(const HotBlock &P, int lo, int hi, float f, const float *cj) {
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < kLd; ++j)
        acc += cj[j] * lerp(P.sun_ld[j][lo], P.sun_ld[j][hi], f);
    return acc;
}

/// Geometry of eval(): world -> local, cos(theta), gamma, masks (sunsky.cpp:309-314). Same sequence in
/// both modes: the sky formula is ill-conditioned in cos(gamma) close to the sun.
struct EvalGeom { RayTerms r; bool active, hit_sun; };
template <int A> SSK_HD EvalGeom eval_geom(const HotBlock &U, const Vec3 &wi, bool active_in) {
    Vec3 neg = { -wi.x, -wi.y, -wi.z };
    Vec3 wo = xform(U.to_local, neg);
    Vec3 n = { U.sun_n[0], U.sun_n[1], U.sun_n[2] };
    float cos_theta = wo.z, gamma = unit_angle<A == kFast>(n, wo);
    EvalGeom g;
    g.active = active_in && cos_theta >= 0.f;
    g.hit_sun = g.active && dot3(n, wo) >= U.cos_cutoff;
    if (g.active) g.r = ray_terms<A>(cos_theta, gamma);
    else { g.r.cos_theta = cos_theta; g.r.gamma = gamma; }
    return g;
}

/// The sun term of SunskyEmitter::eval, RGB, added to the scaled sky radiance `res` in fp64 as the
/// reference does (`Array<float> * double` promotes, sunsky.cpp:321-323). Reference op order in both modes.
SSK_COLD void add_sun_rgb(const HotBlock &U, const float *sun_rad, const RayTerms &r, float res[3]) {
    SunSeg s = sun_segment(r.cos_theta);
    float xk[4], cj[6];
    powers4(s.x, xk);
    powers6(cos_psi(U, r.gamma), cj);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float sun = U.sun_scale * render_sun_rgb(sun_rad, c, s, xk, cj) * U.area_ratio;
        res[c] = (float) ((double) res[c] + (double) sun * 467.069280386);
    }
}

/// Sun-disc lane of SunskyEmitter::eval, RGB, reference sequence from the ray terms on (sunsky.cpp:317-323)
SSK_COLD void eval_rgb_sun_cold(const HotBlock &U, const float *sun_rad, RayTerms r, float out[3]) {
    float res[3];
    for (int c = 0; c < 3; ++c) res[c] = U.sky_scale * render_sky(U, c, r);
    add_sun_rgb(U, sun_rad, r, res);
    for (int c = 0; c < 3; ++c) out[c] = (float) ((double) res[c] * (1.0 / 106.7502593994140625));
}

/// FAST mode, sun-disc lane: normalised sun term on top of the fast sky radiance `sky` (by value in and
/// out: a pointer argument would push the caller's registers to local memory)
SSK_COLD Vec3 add_sun_rgb_fast_cold(const HotBlock &U, const float *sun_rad, RayTerms r, Vec3 sky) {
    float sun[3] = { 0.f, 0.f, 0.f }, out[3] = { sky.x, sky.y, sky.z };
    add_sun_rgb(U, sun_rad, r, sun);
    bool ok = true;
    for (int c = 0; c < 3; ++c) {
        float sn = (float) ((double) sun[c] * (1.0 / 106.7502593994140625)), sum = out[c] + sn;
        ok = ok && fabsf(sum) >= kCancel * (fabsf(out[c]) + fabsf(sn));
        out[c] = sum;
    }
    if (!ok) eval_rgb_sun_cold(U, sun_rad, exact_terms(r), out);
    return { out[0], out[1], out[2] };
}

/// SunskyEmitter::eval, RGB (sunsky.cpp:303-323, 351)
template <int A>
SSK_HD void eval_rgb(const HotBlock &U, const float *sun_rad, const Vec3 &wi, bool active_in, float out[3]) {
    EvalGeom g = eval_geom<A>(U, wi, active_in);
    if (!g.active) { out[0] = out[1] = out[2] = 0.f; return; }
    if (A == kFast) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float T[kFastCoefs];
#pragma unroll
            for (int k = 0; k < kFastCoefs; ++k) T[k] = U.sky_fast[k][c];
            bool ok;
            out[c] = sky_channel_fast(T, g.r, ok);
            if (!ok)   // cancellation: this channel with the reference sequence
                out[c] = (float) ((double) (U.sky_scale * render_sky_cold(U, c, exact_terms(g.r))) *
                                  (1.0 / 106.7502593994140625));
        }
        // sun disc lanes (half of the lanes of sample_direction when the sun is on): the reference's sun
        // term, normalised, on top of the fast sky. It usually outweighs the sky by 4 orders of magnitude,
        // but the fitted sun radiance turns negative for a low sun in a turbid atmosphere and can cancel the
        // sky: such lanes take the reference sequence for all three channels.
        if (g.hit_sun) {
            Vec3 v = add_sun_rgb_fast_cold(U, sun_rad, g.r, Vec3{ out[0], out[1], out[2] });
            out[0] = v.x; out[1] = v.y; out[2] = v.z;
        }
        return;
    }
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; ++c)
        res[c] = U.sky_scale * render_sky(U, c, g.r);
    if (g.hit_sun) add_sun_rgb(U, sun_rad, g.r, res);
#pragma unroll
    for (int c = 0; c < 3; ++c)
        out[c] = (float) ((double) res[c] * (1.0 / 106.7502593994140625));
}

// ---- two rays per thread (RGB): the same sequences on float2, one element per ray ----------------------
struct RayTermsP { float2 cos_theta, gamma, cos_gamma, cos_gamma_sqr, inv_denom_c1, one_plus_cg2, sqrt_cos_theta; };
SSK_HD RayTerms lane_terms(const RayTermsP &r, int e) {
    return e == 0 ? RayTerms{ r.cos_theta.x, r.gamma.x, r.cos_gamma.x, r.cos_gamma_sqr.x, r.inv_denom_c1.x, r.one_plus_cg2.x, r.sqrt_cos_theta.x }
                  : RayTerms{ r.cos_theta.y, r.gamma.y, r.cos_gamma.y, r.cos_gamma_sqr.y, r.inv_denom_c1.y, r.one_plus_cg2.y, r.sqrt_cos_theta.y };
}
template <int A> SSK_HD RayTermsP ray_terms_p(float2 cos_theta, float2 gamma) {
    RayTermsP r;
    r.cos_theta = cos_theta; r.gamma = gamma;
    r.cos_gamma = dcos_2(gamma);
    r.cos_gamma_sqr = mul2(r.cos_gamma, r.cos_gamma);
    float2 den = add2s(cos_theta, 0.01f);
    r.inv_denom_c1 = fdiv_normal_2(splat2(1.f), den);        // IEEE quotient for active rays (den in [0.01, 1.01]); others are discarded
    r.one_plus_cg2 = add2x(r.cos_gamma_sqr, splat2(1.f));
    r.sqrt_cos_theta = A == kFast ? f2(hw_sqrt(fmaxf(cos_theta.x, 0.f)), hw_sqrt(fmaxf(cos_theta.y, 0.f))) : safe_sqrt2(cos_theta);
    return r;
}
SSK_HD Vec3p xform_p(const float *m, const Vec3p &v) {
    return { fma2(splat2(m[2]), v.z, fma2(splat2(m[1]), v.y, mul2s(v.x, m[0]))),
             fma2(splat2(m[5]), v.z, fma2(splat2(m[4]), v.y, mul2s(v.x, m[3]))),
             fma2(splat2(m[8]), v.z, fma2(splat2(m[7]), v.y, mul2s(v.x, m[6]))) };
}
/// render_sky for one channel and two rays, reference op order (see render_sky). Safe = true: both rays passed the
/// guard of HotBlock::exact_lim_* (or are inactive and discarded): select-free exp / pow, quotients by fdiv_normal_2,
/// c1 exponent clamped from below (render_sky_2 explains why that keeps every bit)
template <bool Safe = false> SSK_HD float2 render_sky_p(const HotBlock &P, int ch, const RayTermsP &r) {
    float A0 = P.sky_params[0][ch], A1 = P.sky_params[1][ch], A2 = P.sky_params[2][ch],
          A3 = P.sky_params[3][ch], A4 = P.sky_params[4][ch], A5 = P.sky_params[5][ch],
          A6 = P.sky_params[6][ch], A7 = P.sky_params[7][ch], A8 = P.sky_params[8][ch];
    float2 x1 = mul2s(r.inv_denom_c1, A1);
    if (Safe) x1 = f2(fmaxf(x1.x, -80.f), fmaxf(x1.y, -80.f));
    float2 c1 = add2x(mul2s(dexp_2<Safe>(x1), A0), splat2(1.f));
    const float k1 = 1.f + A8 * A8, k2 = 2.f * A8;          // the same for every lane: uniform scalars
    float2 arg = subp2(splat2(k1), mul2s(r.cos_gamma, k2));
    float2 chi = Safe ? fdiv_normal_2(r.one_plus_cg2, dpow15_2_safe(arg)) : fdiv2(r.one_plus_cg2, dpow15_2(arg));
    float2 c2 = add2x(mul2s(dexp_2<Safe>(mul2s(r.gamma, A4)), A3), splat2(A2));
    c2 = add2x(mul2s(r.cos_gamma_sqr, A5), c2);
    c2 = add2x(mul2s(chi, A6), c2);
    c2 = add2x(mul2s(r.sqrt_cos_theta, A7), c2);
    return mul2s(mul2(c1, c2), P.sky_rad[ch]);
}
/// sky_channel_fast for one channel and two rays (see sky_channel_fast_2 for what is approximated)
SSK_HD float2 sky_channel_fast_p(const float T[kFastCoefs], const RayTermsP &r, bool &ok0, bool &ok1) {
    bool ok_e;
    float2 c1 = c1_fast_2(splat2(T[0]), splat2(T[1]), splat2(T[11]), r.inv_denom_c1, ok_e);
    float2 t = subp2(splat2(T[8]), mul2s(r.cos_gamma, T[9]));
    float2 rs = f2(hw_rsqrt(t.x), hw_rsqrt(t.y));
    float2 chi = mul2(mul2(mul2(rs, rs), rs), r.one_plus_cg2);
    float2 a = mul2s(r.gamma, T[4]);
    float2 ex = f2(hw_ex2(a.x), hw_ex2(a.y));
    float2 c2 = fma2(r.sqrt_cos_theta, splat2(T[7]), fma2(chi, splat2(T[6]), fma2(r.cos_gamma_sqr, splat2(T[5]), fma2(ex, splat2(T[3]), splat2(T[2])))));
    float2 s2 = fma2(r.sqrt_cos_theta, splat2(fabsf(T[7])), fma2(chi, splat2(fabsf(T[6])), fma2(r.cos_gamma_sqr, splat2(fabsf(T[5])), fma2(ex, splat2(fabsf(T[3])), splat2(fabsf(T[2]))))));
    ok0 = ok_e && fabsf(c2.x) >= kCancel * s2.x;
    ok1 = ok_e && fabsf(c2.y) >= kCancel * s2.y;
    return mul2(c1, c2);
}
/// sky_channel_fast_p with running cancellation margins per ray (see sky_channel_fast_2g): g1 = min |c1|, g2 = min
/// (|c2| - kCancel s2) over the channels, element x / y for ray 0 / 1
SSK_HD float2 min2(float2 a, float2 b) { return f2(fminf(a.x, b.x), fminf(a.y, b.y)); }
SSK_HD float2 sky_channel_fast_pg(const float T[kFastCoefs], const RayTermsP &r, float2 &g1, float2 &g2) {
#if SSK_C1_MODE == 1
    float2 a1 = mul2s(r.inv_denom_c1, T[1]);
    float2 c1 = fma2(splat2(T[0]), f2(hw_ex2(a1.x), hw_ex2(a1.y)), splat2(1.f));
    g1 = min2(g1, abs2(c1));
#else
    bool ok_e;
    float2 c1 = c1_fast_2(splat2(T[0]), splat2(T[1]), splat2(T[11]), r.inv_denom_c1, ok_e);
    if (!ok_e) g1 = f2(0.f, 0.f);
#endif
    float2 t = subp2(splat2(T[8]), mul2s(r.cos_gamma, T[9]));
    float2 rs = f2(hw_rsqrt(t.x), hw_rsqrt(t.y));
    float2 chi = mul2(mul2(mul2(rs, rs), rs), r.one_plus_cg2);
    float2 a = mul2s(r.gamma, T[4]);
    float2 ex = f2(hw_ex2(a.x), hw_ex2(a.y));
    float2 c2 = fma2(r.sqrt_cos_theta, splat2(T[7]), fma2(chi, splat2(T[6]), fma2(r.cos_gamma_sqr, splat2(T[5]), fma2(ex, splat2(T[3]), splat2(T[2])))));
    float2 s2 = fma2(r.sqrt_cos_theta, splat2(fabsf(T[7])), fma2(chi, splat2(fabsf(T[6])), fma2(r.cos_gamma_sqr, splat2(fabsf(T[5])), fma2(ex, splat2(fabsf(T[3])), splat2(fabsf(T[2]))))));
    g2 = min2(g2, fma2(s2, splat2(-kCancel), abs2(c2)));
    return mul2(c1, c2);
}
struct EvalGeomP { RayTermsP r; bool active0, active1, hit0, hit1; };
template <int A> SSK_HD EvalGeomP eval_geom_p(const HotBlock &U, const Vec3p &wi, bool act0, bool act1) {
    Vec3p neg = { neg2(wi.x), neg2(wi.y), neg2(wi.z) };
    Vec3p wo = xform_p(U.to_local, neg);
    Vec3 n = { U.sun_n[0], U.sun_n[1], U.sun_n[2] };
    float2 cos_theta = wo.z, gamma = unit_angle_p<A == kFast>(n, wo), dn = dot3_p(n, wo);
    EvalGeomP g;
    g.active0 = act0 && cos_theta.x >= 0.f; g.active1 = act1 && cos_theta.y >= 0.f;
    g.hit0 = g.active0 && dn.x >= U.cos_cutoff; g.hit1 = g.active1 && dn.y >= U.cos_cutoff;
    g.r = ray_terms_p<A>(cos_theta, gamma);
    return g;
}
/// One element of the EXACT pair path, after the sky channels: sun term and fp64 normalisation (sunsky.cpp:317-323)
SSK_COLD Vec3 finish_rgb_sun_cold(const HotBlock &U, const float *sun_rad, RayTerms r, Vec3 res) {
    float v[3] = { res.x, res.y, res.z };
    add_sun_rgb(U, sun_rad, r, v);
    return { v[0], v[1], v[2] };
}
SSK_COLD float sky_rgb_exact_cold(const HotBlock &U, int c, RayTerms r) {
    return (float) ((double) (U.sky_scale * render_sky(U, c, exact_terms(r))) * (1.0 / 106.7502593994140625));
}
SSK_COLD float2 render_sky_p_generic_cold(const HotBlock &U, int c, RayTermsP r) { return render_sky_p<false>(U, c, r); }
/// SunskyEmitter::eval, RGB, two rays (see eval_rgb)
template <int A>
SSK_HD void eval_rgb_p(const HotBlock &U, const float *sun_rad, const Vec3p &wi, bool act0, bool act1, float2 out[3]) {
    EvalGeomP g = eval_geom_p<A>(U, wi, act0, act1);
    if (!g.active0 && !g.active1) { out[0] = out[1] = out[2] = f2(0.f, 0.f); return; }
    if (A == kFast) {
        float2 g1 = f2(1e30f, 1e30f), g2 = g1;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float T[kFastCoefs];
#pragma unroll
            for (int k = 0; k < kFastCoefs; ++k) T[k] = U.sky_fast[k][c];
            out[c] = sky_channel_fast_pg(T, g.r, g1, g2);
        }
        // a ray on which some channel cancels redoes its three channels with the reference sequence
        if (g.active0 && !(g1.x >= kC1Guard && g2.x >= 0.f)) {
#pragma unroll 1
            for (int c = 0; c < 3; ++c) out[c].x = sky_rgb_exact_cold(U, c, lane_terms(g.r, 0));
        }
        if (g.active1 && !(g1.y >= kC1Guard && g2.y >= 0.f)) {
#pragma unroll 1
            for (int c = 0; c < 3; ++c) out[c].y = sky_rgb_exact_cold(U, c, lane_terms(g.r, 1));
        }
        if (g.hit0) {
            Vec3 v = add_sun_rgb_fast_cold(U, sun_rad, lane_terms(g.r, 0), Vec3{ out[0].x, out[1].x, out[2].x });
            out[0].x = v.x; out[1].x = v.y; out[2].x = v.z;
        }
        if (g.hit1) {
            Vec3 v = add_sun_rgb_fast_cold(U, sun_rad, lane_terms(g.r, 1), Vec3{ out[0].y, out[1].y, out[2].y });
            out[0].y = v.x; out[1].y = v.y; out[2].y = v.z;
        }
    } else {
        float2 res[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) res[c] = mul2s(render_sky_p<true>(U, c, g.r), U.sky_scale);
        {   // rays outside the guard (none with the dataset's coefficients): generic sequence for the pair
            const bool in0 = g.r.inv_denom_c1.x >= 0.f && g.r.inv_denom_c1.x <= U.exact_lim_s && fabsf(g.r.gamma.x) <= U.exact_lim_gamma,
                       in1 = g.r.inv_denom_c1.y >= 0.f && g.r.inv_denom_c1.y <= U.exact_lim_s && fabsf(g.r.gamma.y) <= U.exact_lim_gamma;
            if (!((in0 || !g.active0) && (in1 || !g.active1))) {
#pragma unroll 1
                for (int c = 0; c < 3; ++c) res[c] = mul2s(render_sky_p_generic_cold(U, c, g.r), U.sky_scale);
            }
        }
        if (g.hit0) {
            Vec3 v = finish_rgb_sun_cold(U, sun_rad, lane_terms(g.r, 0), Vec3{ res[0].x, res[1].x, res[2].x });
            res[0].x = v.x; res[1].x = v.y; res[2].x = v.z;
        }
        if (g.hit1) {
            Vec3 v = finish_rgb_sun_cold(U, sun_rad, lane_terms(g.r, 1), Vec3{ res[0].y, res[1].y, res[2].y });
            res[0].y = v.x; res[1].y = v.y; res[2].y = v.z;
        }
#pragma unroll
        for (int c = 0; c < 3; ++c)
            out[c] = f2((float) ((double) res[c].x * (1.0 / 106.7502593994140625)), (float) ((double) res[c].y * (1.0 / 106.7502593994140625)));
    }
    if (!g.active0) { out[0].x = 0.f; out[1].x = 0.f; out[2].x = 0.f; }
    if (!g.active1) { out[0].y = 0.f; out[1].y = 0.f; out[2].y = 0.f; }
}

/// Spectral bin lookup of one wavelength (sunsky.cpp:326-338)
struct Bin { int lo; float f; bool valid; };
SSK_HD Bin wavelength_bin(float wav) {
    float nw = (wav - 320.f) * (1.f / 40.f);                  // Spectrum / size_t -> mul by rcp(40.f)
    Bin b;
    b.valid = 0.f <= nw && nw <= 10.f;
    float fl = floorf(nw);
    b.lo = b.valid ? (int) fl : 0;
    b.f = nw - fl;
    return b;
}

/// Sun term of the spectral eval for one wavelength (sunsky.cpp:339-347), reference op order
SSK_HD float sun_spec_term(const HotBlock &U, const HotBlock &S, const float *sun_rad, const SunSeg &s,
                           const float *xk, const float *cj, int lo, float f) {
    int hi = lo + 1;
    float sun_lo = render_sun_spec(sun_rad, lo, s, xk),
          sun_hi = hi < kMaxChannels ? render_sun_spec(sun_rad, hi, s, xk) : 0.f,
          sun = lerp(sun_lo, sun_hi, f);
    float ld = sun_ld(S, lo, hi, f, cj);
    return U.sun_scale * sun * ld * U.area_ratio;
}

/// Layout of the FAST spectral coefficient table in shared memory: pairs[k * kChanPad + bin] =
/// { sky_fast[k][bin], sky_fast[k][bin + 1] }, the coefficient k of the two dataset bins a wavelength falls
/// between, in one 64-bit load. For a given k the 11 bins span 88 bytes, so whatever bins the lanes of a warp
/// ask for, the load is a single conflict-free wavefront (a row-per-bin layout read with 128-bit loads was
/// measured at 2.2x the ideal wavefront count and became the bottleneck: profiles/r1_summary.md).
/// With SSK_QUAD_TABLE the same pairs are interleaved two coefficients at a time, so that one 128-bit load returns
/// the pairs of coefficients 2q and 2q + 1: pairs[((k / 2) * kChanPad + bin) * 2 + (k & 1)]. For a given q the 11 bins
/// span 176 bytes: two lanes of a quarter-warp collide only when their bins differ by exactly 8.
constexpr int kPairTable = kFastCoefs * kChanPad;
SSK_HD int pair_index(int k, int bin) {
#if SSK_QUAD_TABLE
    return ((k >> 1) * kChanPad + bin) * 2 + (k & 1);
#else
    return k * kChanPad + bin;
#endif
}
SSK_HD void build_pair_table(const HotBlock &B, float2 *pairs) {
    for (int k = 0; k < kFastCoefs; ++k)
        for (int c = 0; c < kChanPad; ++c)
            pairs[pair_index(k, c)] = f2(B.sky_fast[k][c], c + 1 < kChanPad ? B.sky_fast[k][c + 1] : 0.f);
}

/// Sky radiance of the 4 wavelengths with the reference sequence (FAST mode's fallback for cancelling lanes)
SSK_COLD float4 sky_spec_exact_cold(const HotBlock &U, const HotBlock &S, RayTerms r, float4 wav) {
    const float wl[4] = { wav.x, wav.y, wav.z, wav.w };
    float o[4];
    RayTerms re = exact_terms(r);
    for (int w = 0; w < 4; ++w) {
        Bin b = wavelength_bin(wl[w]);
        float2 sk = render_sky_2(S, b.lo, re);
        o[w] = U.sky_scale * lerp(sk.x, sk.y, b.f);
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

/// EXACT spectral sky of the 4 wavelengths from the pair table, generic sequence (rays that fail the exact_lim guard)
SSK_COLD float4 sky_spec_pairs_generic_cold(const HotBlock &U, const float2 *pairs, RayTerms r, float4 wav) {
    const float wl[4] = { wav.x, wav.y, wav.z, wav.w };
    float o[4];
    for (int w = 0; w < 4; ++w) {
        Bin b = wavelength_bin(wl[w]);
        float2 sk = render_sky_2<false>(pairs, b.lo, r);
        o[w] = U.sky_scale * lerp(sk.x, sk.y, b.f);
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

/// Sun-disc lanes of the spectral eval: the sun term of every wavelength on top of `sky` (sunsky.cpp:339-347),
/// reference op order. `fast`: guard against the sun cancelling the sky (the fitted sun radiance can be
/// negative for a low sun in a turbid atmosphere); such lanes redo the sky with the reference sequence.
SSK_COLD float4 add_sun_spec_cold(const HotBlock &U, const HotBlock &S, const float *sun_rad, RayTerms r, float4 wav,
                                  float4 sky, bool fast) {
    const float wl[4] = { wav.x, wav.y, wav.z, wav.w };
    float res[4] = { sky.x, sky.y, sky.z, sky.w }, sun[4];
    SunSeg s = sun_segment(r.cos_theta);
    float xk[4], cj[6];
    powers4(s.x, xk);
    powers6(cos_psi(U, r.gamma), cj);
    bool ok = true;
    for (int w = 0; w < 4; ++w) {
        Bin b = wavelength_bin(wl[w]);
        sun[w] = sun_spec_term(U, S, sun_rad, s, xk, cj, b.lo, b.f);
        float sum = res[w] + sun[w];
        ok = ok && fabsf(sum) >= kCancel * (fabsf(res[w]) + fabsf(sun[w]));
        res[w] = sum;
    }
    if (fast && !ok) {
        float4 e = sky_spec_exact_cold(U, S, r, wav);
        res[0] = e.x + sun[0]; res[1] = e.y + sun[1]; res[2] = e.z + sun[2]; res[3] = e.w + sun[3];
    }
    return make_float4(res[0], res[1], res[2], res[3]);
}

/// SunskyEmitter::eval, spectral, for the 4 wavelengths of one ray (sunsky.cpp:325-351). The two bins of a
/// wavelength are evaluated as one float2 pair. FAST mode reads `pairs` (build_pair_table); lanes where a
/// channel cancels redo all four wavelengths with the reference sequence, sun-disc lanes add the sun term,
/// both behind one call per ray.
template <int A, bool ExactPairs = false>
SSK_HD void eval_spec(const HotBlock &U, const HotBlock &S, const float2 *pairs, const float *sun_rad,
                      const Vec3 &wi, const float wav[4], bool active_in, float out[4]) {
    EvalGeom g = eval_geom<A>(U, wi, active_in);
    if (!g.active) { out[0] = out[1] = out[2] = out[3] = 0.f; return; }
    bool ok = true, valid[4];
    float g1 = 1e30f, g2 = 1e30f;                   // FAST: cancellation margins of the ray (sky_channel_fast_2g)
    // EXACT: rays inside the guard take the select-free sequence, the others redo the four wavelengths with the
    // generic one behind a call (same values; see HotBlock::exact_lim_s)
    const bool safe = A == kExact && g.r.inv_denom_c1 >= 0.f && g.r.inv_denom_c1 <= U.exact_lim_s &&
                      fabsf(g.r.gamma) <= U.exact_lim_gamma;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
        Bin b = wavelength_bin(wav[w]);
        valid[w] = b.valid;
        if (A == kFast) {
            float2 T[kFastCoefs];
#if SSK_QUAD_TABLE
            const float4 *quads = reinterpret_cast<const float4 *>(pairs);
#pragma unroll
            for (int q = 0; q < kFastCoefs / 2; ++q) {
                float4 v = quads[q * kChanPad + b.lo];
                T[2 * q] = f2(v.x, v.y); T[2 * q + 1] = f2(v.z, v.w);
            }
#else
#pragma unroll
            for (int k = 0; k < kFastCoefs; ++k) T[k] = pairs[k * kChanPad + b.lo];
#endif
            float2 sk = sky_channel_fast_2g(T, g.r, g1, g2);
            out[w] = lerp(sk.x, sk.y, b.f);
        } else {
            // EXACT: with ExactPairs the kernel staged build_exact_pair_table's table in `pairs`
            if (ExactPairs) {
                // guard failed: the value computed here is discarded below (no branch inside the unrolled loop)
                float2 sk = render_sky_2<true>(pairs, b.lo, g.r);
                out[w] = U.sky_scale * lerp(sk.x, sk.y, b.f);
            } else {
                float2 sk = render_sky_2<true>(S, b.lo, g.r);
                out[w] = U.sky_scale * lerp(sk.x, sk.y, b.f);
            }
        }
    }
    const float4 wav4 = make_float4(wav[0], wav[1], wav[2], wav[3]);
    if (A == kFast) ok = g1 >= kC1Guard && g2 >= 0.f;
    if (A == kExact && !safe) {
        float4 e = ExactPairs ? sky_spec_pairs_generic_cold(U, pairs, g.r, wav4) : sky_spec_exact_cold(U, S, g.r, wav4);
        out[0] = e.x; out[1] = e.y; out[2] = e.z; out[3] = e.w;
    }
    if (A == kFast && !ok) {
        float4 e = sky_spec_exact_cold(U, S, g.r, wav4);
        out[0] = e.x; out[1] = e.y; out[2] = e.z; out[3] = e.w;
    }
    if (g.hit_sun) {
        float4 e = add_sun_spec_cold(U, S, sun_rad, A == kFast ? exact_terms(g.r) : g.r, wav4,
                                     make_float4(out[0], out[1], out[2], out[3]), A == kFast);
        out[0] = e.x; out[1] = e.y; out[2] = e.z; out[3] = e.w;
    }
#pragma unroll
    for (int w = 0; w < 4; ++w) out[w] = valid[w] ? out[w] : 0.f;
}

// ---------------------------------------------------------------------------------------------
/// tgmm_pdf (sunsky.cpp:732-763), reference op order, generic IEEE divisions. mu / sigma / weight and
/// `volume` are per-emitter constants that the JIT folds on the host (m_gaussians[i] is a scalar read).
SSK_COLD float tgmm_sum_generic(const HotBlock &U, float phi, float theta) {
    float pdf = 0.f;
#pragma unroll 1
    for (int i = 0; i < U.n_pdf; ++i) {
        float sx = fdiv(phi - U.p_mu_phi[i], U.p_sigma_phi[i]),
              sy = fdiv(theta - U.p_mu_theta[i], U.p_sigma_theta[i]);
        float g = SSK_INV_TWO_PI * dexp(-.5f * fmaf(sy, sy, sx * sx));   // warp.h:189-191
        pdf += fdiv(U.p_weight[i] * g, U.p_volume[i]);
    }
    return pdf;
}

/// exp(x) of a Gaussian exponent x <= 0 for the EXACT density sums below: the select-free dexp_2<true> on the
/// exponent clamped at -86. A clamped element comes out as e^-86 = 4.5e-38 instead of something smaller (or the 0 the
/// reference's underflow select gives): its numerator weight * g / 2 pi is far below 2^-70 (weights <= 1e10, checked on
/// the host), so the sums flag it `tiny` like any other negligible term -- a density >= 2^-30 absorbs it without a
/// trace, anything smaller is recomputed with the generic sequence. NaN / infinite exponents clamp the same way and
/// end in the generic sum, which propagates them as the reference does.
SSK_HD float2 gauss_exp_2(float2 x) { return dexp_2<true>(f2(fmaxf(x.x, -86.f), fmaxf(x.y, -86.f))); }

/// The same sum with the three divisions per component done by fdiv_by() on reciprocals rounded at
/// setup (identical quotients, 3 instructions each). A numerator below 2^-70 could make the residual
/// inexact; such a term only matters when the whole density is tiny, and then the generic sum is redone.
SSK_HD float tgmm_sum_exact(const HotBlock &U, float phi, float theta) {
    float pdf = 0.f;
    bool tiny = false;
    // two components per pass on float2 (same per-element op sequence), accumulated in component order
#pragma unroll
    for (int i = 0; i < kGauss; i += 2) {
        if (i + 1 < U.n_pdf) {
            float2 sx = fdiv_by_2(sub2(splat2(phi), f2(U.p_mu_phi[i], U.p_mu_phi[i + 1])),
                                  f2(U.p_sigma_phi[i], U.p_sigma_phi[i + 1]), f2(U.p_rsigma_phi[i], U.p_rsigma_phi[i + 1])),
                   sy = fdiv_by_2(sub2(splat2(theta), f2(U.p_mu_theta[i], U.p_mu_theta[i + 1])),
                                  f2(U.p_sigma_theta[i], U.p_sigma_theta[i + 1]), f2(U.p_rsigma_theta[i], U.p_rsigma_theta[i + 1]));
            float2 g = mul2s(gauss_exp_2(mul2s(fma2(sy, sy, mul2(sx, sx)), -.5f)), SSK_INV_TWO_PI);
            float2 num = mul2(f2(U.p_weight[i], U.p_weight[i + 1]), g);
            tiny = tiny || fdiv_by_tiny(num.x) || fdiv_by_tiny(num.y);
            float2 q = fdiv_by_2(num, f2(U.p_volume[i], U.p_volume[i + 1]), f2(U.p_rvolume[i], U.p_rvolume[i + 1]));
            pdf += q.x;
            pdf += q.y;
        } else if (i < U.n_pdf) {
            float sx = fdiv_by(phi - U.p_mu_phi[i], U.p_sigma_phi[i], U.p_rsigma_phi[i]),
                  sy = fdiv_by(theta - U.p_mu_theta[i], U.p_sigma_theta[i], U.p_rsigma_theta[i]);
            float g = SSK_INV_TWO_PI * dexp(-.5f * fmaf(sy, sy, sx * sx));
            float num = U.p_weight[i] * g;
            tiny = tiny || fdiv_by_tiny(num);
            pdf += fdiv_by(num, U.p_volume[i], U.p_rvolume[i]);
        }
    }
    if (tiny && !(pdf >= 9.313225746154785e-10f /* 2^-30 */)) pdf = tgmm_sum_generic(U, phi, theta);
    return pdf;
}

/// FAST: 4.5 instructions per component, one MUFU. phi - mu is the reference's own rounding (the Gaussian
/// argument cancels there); k = sqrt(log2(e) / 2) / sigma, wv = weight / (2 pi volume). Two components per
/// pass on float2, even and odd components in separate accumulators.
SSK_HD float tgmm_sum_fast(const HotBlock &U, float phi, float theta) {
    float2 acc = f2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < kGauss; i += 2) {
        if (i + 1 < U.n_pdf) {
            float2 sx = mul2(sub2(splat2(phi), f2(U.p_mu_phi[i], U.p_mu_phi[i + 1])), f2(U.p_k_phi[i], U.p_k_phi[i + 1])),
                   sy = mul2(sub2(splat2(theta), f2(U.p_mu_theta[i], U.p_mu_theta[i + 1])), f2(U.p_k_theta[i], U.p_k_theta[i + 1]));
            float2 e = fma2(sy, sy, mul2(sx, sx));
            acc = fma2(f2(U.p_wv[i], U.p_wv[i + 1]), f2(hw_ex2(-e.x), hw_ex2(-e.y)), acc);
        } else if (i < U.n_pdf) {
            float sx = (phi - U.p_mu_phi[i]) * U.p_k_phi[i], sy = (theta - U.p_mu_theta[i]) * U.p_k_theta[i];
            acc.x = fmaf(U.p_wv[i], hw_ex2(-fmaf(sy, sy, sx * sx)), acc.x);
        }
    }
    return acc.x + acc.y;
}

template <int A> SSK_HD float tgmm_pdf(const HotBlock &U, float phi, float theta, bool active) {
    phi -= U.sun_phi_shift;
    phi = phi < 0.f ? phi + SSK_TWO_PI : phi;
    phi = phi > SSK_TWO_PI ? phi - SSK_TWO_PI : phi;
    active = active && theta >= 0.f && theta <= SSK_HALF_PI;
    if (!active) return 0.f;
    if (A == kFast) return tgmm_sum_fast(U, phi, theta);
    return U.pdf_generic ? tgmm_sum_generic(U, phi, theta) : tgmm_sum_exact(U, phi, theta);
}

/// compute_pdfs (sunsky.cpp:711-723) followed by the lerp of pdf_direction / sample_direction
template <int A> SSK_HD float combined_pdf(const HotBlock &U, const Vec3 &d, bool check_sun, bool active) {
    float sin_theta = safe_sqrt(fmaf(d.x, d.x, d.y * d.y));        // Frame3f::sin_theta (frame.h:59-65)
    active = active && d.z >= 0.f && sin_theta != 0.f;
    sin_theta = fmaxf(sin_theta, SSK_EPSILON);
    float phi = datan2(d.y, d.x), theta = unit_angle_z(d);          // from_spherical (sunsky.h:83-89)
    float t = tgmm_pdf<A>(U, phi, theta, active);
    float sky_pdf = A == kFast ? t * hw_rcp(sin_theta) : fdiv(t, sin_theta);
    Vec3 n = { U.sun_n[0], U.sun_n[1], U.sun_n[2] };
    float sun_pdf = (!check_sun || dot3(n, d) >= U.cos_cutoff) ? U.sun_pdf_const : 0.f;
    return lerp(sun_pdf, sky_pdf, U.sky_w);
}

/// pdf_direction (sunsky.cpp:443-451)
template <int A> SSK_HD float pdf_direction(const HotBlock &U, const Vec3 &d_world) {
    return combined_pdf<A>(U, xform(U.to_local, d_world), true, true);
}

// ---- two rays per thread: pdf --------------------------------------------------------------------------
/// tgmm_sum_exact for two rays: per element the same ops and the same accumulation order
SSK_HD float2 tgmm_sum_exact_p(const HotBlock &U, float2 phi, float2 theta) {
    float2 pdf = f2(0.f, 0.f);
    bool tiny0 = false, tiny1 = false;
#pragma unroll
    for (int i = 0; i < kGauss; ++i) {
        if (i < U.n_pdf) {
            float2 sx = fdiv_by_2(add2s(phi, -U.p_mu_phi[i]), splat2(U.p_sigma_phi[i]), splat2(U.p_rsigma_phi[i])),
                   sy = fdiv_by_2(add2s(theta, -U.p_mu_theta[i]), splat2(U.p_sigma_theta[i]), splat2(U.p_rsigma_theta[i]));
            float2 g = mul2s(gauss_exp_2(mul2s(fma2(sy, sy, mul2(sx, sx)), -.5f)), SSK_INV_TWO_PI);
            float2 num = mul2s(g, U.p_weight[i]);
            tiny0 = tiny0 || fdiv_by_tiny(num.x); tiny1 = tiny1 || fdiv_by_tiny(num.y);
            pdf = add2(pdf, fdiv_by_2(num, splat2(U.p_volume[i]), splat2(U.p_rvolume[i])));
        }
    }
    if (tiny0 && !(pdf.x >= 9.313225746154785e-10f)) pdf.x = tgmm_sum_generic(U, phi.x, theta.x);
    if (tiny1 && !(pdf.y >= 9.313225746154785e-10f)) pdf.y = tgmm_sum_generic(U, phi.y, theta.y);
    return pdf;
}
/// tgmm_sum_fast for two rays: even and odd components in separate accumulators, as the scalar version
SSK_HD float2 tgmm_sum_fast_p(const HotBlock &U, float2 phi, float2 theta) {
    float2 acc_e = f2(0.f, 0.f), acc_o = f2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < kGauss; ++i) {
        if (i < U.n_pdf) {
            float2 sx = mul2s(add2s(phi, -U.p_mu_phi[i]), U.p_k_phi[i]), sy = mul2s(add2s(theta, -U.p_mu_theta[i]), U.p_k_theta[i]);
            float2 e = fma2(sy, sy, mul2(sx, sx));
            float2 ex = f2(hw_ex2(-e.x), hw_ex2(-e.y));
            if (i & 1) acc_o = fma2(splat2(U.p_wv[i]), ex, acc_o);
            else acc_e = fma2(splat2(U.p_wv[i]), ex, acc_e);
        }
    }
    return add2(acc_e, acc_o);
}
template <int A> SSK_HD float2 tgmm_pdf_p(const HotBlock &U, float2 phi, float2 theta, bool act0, bool act1) {
    phi = add2s(phi, -U.sun_phi_shift);
    float2 up = add2s(phi, SSK_TWO_PI);
    phi = sel2(phi.x < 0.f, phi.y < 0.f, up, phi);
    float2 dn = add2s(phi, -SSK_TWO_PI);
    phi = sel2(phi.x > SSK_TWO_PI, phi.y > SSK_TWO_PI, dn, phi);
    act0 = act0 && theta.x >= 0.f && theta.x <= SSK_HALF_PI;
    act1 = act1 && theta.y >= 0.f && theta.y <= SSK_HALF_PI;
    if (!act0 && !act1) return f2(0.f, 0.f);
    float2 r;
    if (A == kFast) r = tgmm_sum_fast_p(U, phi, theta);
    else if (U.pdf_generic) r = f2(tgmm_sum_generic(U, phi.x, theta.x), tgmm_sum_generic(U, phi.y, theta.y));
    else r = tgmm_sum_exact_p(U, phi, theta);
    return f2(act0 ? r.x : 0.f, act1 ? r.y : 0.f);
}
/// combined_pdf for two rays
template <int A> SSK_HD float2 combined_pdf_p(const HotBlock &U, const Vec3p &d, bool cs0, bool cs1, bool act0, bool act1) {
    float2 sin_theta = safe_sqrt2(fma2(d.x, d.x, mul2(d.y, d.y)));
    act0 = act0 && d.z.x >= 0.f && sin_theta.x != 0.f;
    act1 = act1 && d.z.y >= 0.f && sin_theta.y != 0.f;
    sin_theta = f2(fmaxf(sin_theta.x, SSK_EPSILON), fmaxf(sin_theta.y, SSK_EPSILON));
    float2 phi = datan2_2(d.y, d.x), theta = unit_angle_z_p(d);
    float2 t = tgmm_pdf_p<A>(U, phi, theta, act0, act1);
    float2 sky_pdf = A == kFast ? mul2(t, f2(hw_rcp(sin_theta.x), hw_rcp(sin_theta.y))) : fdiv2(t, sin_theta);
    Vec3 n = { U.sun_n[0], U.sun_n[1], U.sun_n[2] };
    float2 dn = dot3_p(n, d);
    float2 sun_pdf = f2((!cs0 || dn.x >= U.cos_cutoff) ? U.sun_pdf_const : 0.f, (!cs1 || dn.y >= U.cos_cutoff) ? U.sun_pdf_const : 0.f);
    return lerp2(sun_pdf, sky_pdf, splat2(U.sky_w));
}
template <int A> SSK_HD float2 pdf_direction_p(const HotBlock &U, const Vec3p &d_world, bool act0, bool act1) {
    return combined_pdf_p<A>(U, xform_p(U.to_local, d_world), true, true, act0, act1);
}

/// DiscreteDistribution::sample_reuse, JIT flavour (distr_1d.h:116-134,172-182; drjit/util.h:136-181)
SSK_HD int tgmm_pick(const HotBlock &U, const HotBlock &S, float u, float &reuse) {
    float s = u * U.cdf_sum;
    uint32_t start = 0, end = kGauss - 1;
#pragma unroll
    for (int it = 0; it < 5; ++it) {                 // log2i(19) + 1 iterations
        uint32_t middle = (start + end) >> 1;
        float v = S.g_cdf[middle];
        bool cond = ((v < s) || v == 0.f) && v != U.cdf_sum;
        uint32_t nstart = middle + 1 < end ? middle + 1 : end;
        start = cond ? nstart : start;
        end = cond ? end : middle;
    }
    reuse = fdiv_by_safe(u - S.g_cdf_prev_n[start], S.g_pmf_n[start], S.g_rpmf_n[start]);
    return (int) start;
}

/// sample_sky (sunsky.cpp:661-689). cdf_a / cdf_b depend only on the component and are precomputed.
/// Reference sequence in both modes: the sampled direction is held to 1e-6 absolute.
SSK_HD Vec3 sample_sky(const HotBlock &U, const HotBlock &S, float u0, float u1) {
    float t;
    int i = tgmm_pick(U, S, u0, t);
    float sx = lerp(S.g_cdf_a_phi[i], S.g_cdf_b_phi[i], t),
          sy = lerp(S.g_cdf_a_theta[i], S.g_cdf_b_theta[i], u1);
    sx = fmaxf(fminf(sx, SSK_ONE_MINUS_EPSILON), SSK_EPSILON);
    sy = fmaxf(fminf(sy, SSK_ONE_MINUS_EPSILON), SSK_EPSILON);
    float phi = SSK_SQRT_TWO * derfinv(2.f * sx - 1.f) * S.g_sigma_phi[i] + S.g_mu_phi[i],
          theta = SSK_SQRT_TWO * derfinv(2.f * sy - 1.f) * S.g_sigma_theta[i] + S.g_mu_theta[i];
    phi += U.sun_phi_shift;
    theta = fminf(theta, SSK_HALF_PI - SSK_EPSILON);
    return sphdir(theta, phi);
}

/// sample_sun (sunsky.cpp:697-701): warp::square_to_uniform_cone via the concentric disk mapping
/// (warp.h:54-90, 533-551), rotated into the sun frame (frame.h:39-41)
SSK_HD Vec3 sample_sun(const HotBlock &P, float u0, float u1) {
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
template <int A> SSK_HD DirSample sample_dir(const HotBlock &U, const HotBlock &S, float u0, float u1) {
    bool pick_sky = u0 < U.sky_w;
    Vec3 dir = pick_sky ? sample_sky(U, S, fdiv_by_safe(u0, U.sky_w, U.r_sky_w), u1)
                        : sample_sun(U, fdiv_by_safe(u0 - U.sky_w, U.one_minus_sky_w, U.r_one_minus_sky_w), u1);
    DirSample r;
    r.active = dir.z >= 0.f;
    r.d_world = xform(U.to_world, dir);
    r.pdf = combined_pdf<A>(U, dir, pick_sky, r.active);
    return r;
}

// ---- forward derivative of eval() w.r.t. the Differentiable parameters ------------------------------------
// traverse() flags turbidity, albedo and sun_direction as Differentiable (sunsky.cpp:222-237). They enter eval()
//   * through the interpolated state: sky coefficients and mean radiance (piecewise linear in turbidity and albedo,
//     quintic Bezier in x = cbrt(2 eta / pi) of the sun elevation eta, sunsky.h:157-231) and the sun table (linear in
//     turbidity, :403-419): the host builds that tangent (state_tangent, ssk_host.cu);
//   * sun_direction also directly, through gamma = unit_angle(sun, wo) in render_sky and through cos(psi) in the
//     sun-disc term: the kernel differentiates those formulas as written (what forward-mode AD does; masks, bin and
//     segment indices are constants).
// The independent oracle is oracle/jvp_oracle.py (fp64 dual numbers).
struct TangentBlock {
    float sky_params[kSkyParams][kChanPad];
    float sky_rad[kChanPad];
    float d_sun_n[4];                                 // tangent of the local sun direction (m_local_sun_frame.n)
    float sun_rad[kSunSegments * 3 * 4 * kLd];
};
/// The per-channel sky part of a HotBlock or a TangentBlock (same member layout as the head of both): what the
/// spectral derivative indexes with a per-lane channel, so the kernel keeps a shared-memory copy of each
struct SkyCoefs {
    float sky_params[kSkyParams][kChanPad];
    float sky_rad[kChanPad];
};
/// d gamma for the tangent dn of the sun direction: unit_angle (drjit/sphere.h:53-59) differentiated as written
SSK_HD float unit_angle_jvp(const Vec3 &n, const Vec3 &wo, const float dn[3]) {
    const float dot = dot3(n, wo), sg = dot >= 0.f ? 1.f : -1.f;
    const float dx = wo.x - sg * n.x, dy = wo.y - sg * n.y, dz = wo.z - sg * n.z;
    const float D2 = dx * dx + dy * dy + dz * dz;
    if (!(D2 > 0.f)) return 0.f;
    const float D = sqrtf(D2);
    const float dD = -sg * (dx * dn[0] + dy * dn[1] + dz * dn[2]) / D;
    return sg * dD / sqrtf(fmaxf(1.f - 0.25f * D2, 1e-30f));
}
/// d render_sky(ch) for the state tangent T and the angle tangent dgamma (product rule over sunsky.cpp:538-555).
/// A = kFast: MUFU ex2 / rsq for the transcendentals (the derivative has no bit-level reference to match)
template <int A, class PB, class TB>
SSK_HD float render_sky_jvp(const PB &P, const TB &T, int ch, const RayTerms &r, float dgamma, float sin_gamma) {
    float A0 = P.sky_params[0][ch], A1 = P.sky_params[1][ch], A2 = P.sky_params[2][ch], A3 = P.sky_params[3][ch],
          A4 = P.sky_params[4][ch], A5 = P.sky_params[5][ch], A6 = P.sky_params[6][ch], A7 = P.sky_params[7][ch],
          A8 = P.sky_params[8][ch], L = P.sky_rad[ch];
    float d0 = T.sky_params[0][ch], d1 = T.sky_params[1][ch], d2 = T.sky_params[2][ch], d3 = T.sky_params[3][ch],
          d4 = T.sky_params[4][ch], d5 = T.sky_params[5][ch], d6 = T.sky_params[6][ch], d7 = T.sky_params[7][ch],
          d8 = T.sky_params[8][ch], dL = T.sky_rad[ch];
    constexpr float kLog2e = 1.44269504088896340736f;
    float e1 = A == kFast ? hw_ex2(A1 * r.inv_denom_c1 * kLog2e) : dexp(A1 * r.inv_denom_c1);
    float c1 = 1.f + A0 * e1, dc1 = d0 * e1 + A0 * e1 * (d1 * r.inv_denom_c1);
    float arg = 1.f + A8 * A8 - 2.f * A8 * r.cos_gamma;
    float chi, chi_over_arg, inv_pow;
    if (A == kFast) { float rs = hw_rsqrt(arg); inv_pow = rs * rs * rs; chi = r.one_plus_cg2 * inv_pow; chi_over_arg = chi * (rs * rs); }
    else { float pw = dpow15(arg); inv_pow = fdiv(1.f, pw); chi = fdiv(r.one_plus_cg2, pw); chi_over_arg = fdiv(chi, arg); }
    const float dcg = -sin_gamma * dgamma, dcg2 = 2.f * r.cos_gamma * dcg;
    float darg = (2.f * A8 - 2.f * r.cos_gamma) * d8 - 2.f * A8 * dcg;
    float dchi = dcg2 * inv_pow - 1.5f * chi_over_arg * darg;
    float e4 = A == kFast ? hw_ex2(A4 * r.gamma * kLog2e) : dexp(A4 * r.gamma);
    float c2 = A2 + A3 * e4 + A5 * r.cos_gamma_sqr + A6 * chi + A7 * r.sqrt_cos_theta;
    float dc2 = d2 + d3 * e4 + A3 * e4 * (r.gamma * d4 + A4 * dgamma) + d5 * r.cos_gamma_sqr + A5 * dcg2 + d6 * chi + A6 * dchi +
                d7 * r.sqrt_cos_theta;
    return (dc1 * c2 + c1 * dc2) * L + c1 * c2 * dL;
}
/// Powers of cos(psi) and their tangents for the sun-disc term: cj[j] = c^j, dcj[j] = j c^(j-1) dc with
/// c = sqrt(1 - sin^2(gamma) / sin^2(half aperture)) (sunsky.h:384-392)
SSK_HD void cos_psi_powers_jvp(const HotBlock &U, float gamma, float sin_gamma, float cos_gamma, float dgamma, float cj[6], float dcj[6]) {
    const float c = cos_psi(U, gamma);
    const float dc = c > 0.f ? -U.ar2 * sin_gamma * cos_gamma * dgamma / c : 0.f;
    powers6(c, cj);
    dcj[0] = 0.f;
    for (int j = 1; j < 6; ++j) dcj[j] = (float) j * cj[j - 1] * dc;
}
/// d eval, RGB (mirrors eval_rgb). sun_rad: the primal sun table (for the cos(psi) tangent); T.sun_rad: its tangent
template <int A> SSK_HD void eval_rgb_jvp(const HotBlock &U, const TangentBlock &T, const float *sun_rad, const Vec3 &wi, float out[3]) {
    EvalGeom g = eval_geom<kExact>(U, wi, true);
    out[0] = out[1] = out[2] = 0.f;
    if (!g.active) return;
    Vec3 n = { U.sun_n[0], U.sun_n[1], U.sun_n[2] };
    Vec3 wo = xform(U.to_local, Vec3{ -wi.x, -wi.y, -wi.z });
    const float dgamma = unit_angle_jvp(n, wo, T.d_sun_n), sg = dsin(g.r.gamma);
    float xk[4], cj[6], dcj[6];
    SunSeg s = { 0, 0.f };
    if (g.hit_sun) { s = sun_segment(g.r.cos_theta); powers4(s.x, xk); cos_psi_powers_jvp(U, g.r.gamma, sg, g.r.cos_gamma, dgamma, cj, dcj); }
    for (int c = 0; c < 3; ++c) {
        double res = (double) (U.sky_scale * render_sky_jvp<A>(U, T, c, g.r, dgamma, sg));
        if (g.hit_sun) {
            float dsun = render_sun_rgb<false>(T.sun_rad, c, s, xk, cj) + render_sun_rgb(sun_rad, c, s, xk, dcj);
            res += (double) (U.sun_scale * dsun * U.area_ratio) * 467.069280386;
        }
        out[c] = (float) (res * (1.0 / 106.7502593994140625));
    }
}
/// d eval, spectral (mirrors eval_spec). P / TS: the sky coefficients and their tangent, indexed by a per-lane channel
/// (the kernel passes shared-memory copies; U and T themselves work too)
template <int A, class PB, class TB>
SSK_HD void eval_spec_jvp(const HotBlock &U, const PB &P, const TB &TS, const TangentBlock &T, const float *sun_rad,
                          const Vec3 &wi, const float wav[4], float out[4]) {
    EvalGeom g = eval_geom<kExact>(U, wi, true);
    out[0] = out[1] = out[2] = out[3] = 0.f;
    if (!g.active) return;
    Vec3 n = { U.sun_n[0], U.sun_n[1], U.sun_n[2] };
    Vec3 wo = xform(U.to_local, Vec3{ -wi.x, -wi.y, -wi.z });
    const float dgamma = unit_angle_jvp(n, wo, T.d_sun_n), sg = dsin(g.r.gamma);
    float xk[4], cj[6], dcj[6];
    SunSeg s = { 0, 0.f };
    if (g.hit_sun) { s = sun_segment(g.r.cos_theta); powers4(s.x, xk); cos_psi_powers_jvp(U, g.r.gamma, sg, g.r.cos_gamma, dgamma, cj, dcj); }
    for (int w = 0; w < 4; ++w) {
        Bin b = wavelength_bin(wav[w]);
        if (!b.valid) continue;
        const int hi = b.lo + 1;                   // channel 11 is the zero pad (see HotBlock)
        float d = U.sky_scale * lerp(render_sky_jvp<A>(P, TS, b.lo, g.r, dgamma, sg), render_sky_jvp<A>(P, TS, hi, g.r, dgamma, sg), b.f);
        if (g.hit_sun) {
            float t_lo = render_sun_spec<false>(T.sun_rad, b.lo, s, xk), t_hi = hi < kMaxChannels ? render_sun_spec<false>(T.sun_rad, hi, s, xk) : 0.f;
            float p_lo = render_sun_spec(sun_rad, b.lo, s, xk), p_hi = hi < kMaxChannels ? render_sun_spec(sun_rad, hi, s, xk) : 0.f;
            d += U.sun_scale * (lerp(t_lo, t_hi, b.f) * sun_ld(U, b.lo, hi, b.f, cj) + lerp(p_lo, p_hi, b.f) * sun_ld(U, b.lo, hi, b.f, dcj)) * U.area_ratio;
        }
        out[w] = d;
    }
}

// ---- setup quadrature (estimate_sky_sun_ratio, sunsky.cpp:772-886): one node of the 200 x 200 grids --------
constexpr int kQuadN = 200;
/// Sky over the hemisphere, [-1,1]^2 -> [0,2pi] x [0,1] (sunsky.cpp:789-815): per-channel integrand x weights
SSK_HD void quad_sky_node(const HotBlock &B, int C, float gx_x, float gx_y, float gw_x, float gw_y, float *out) {
    float cos_theta = 0.5f * (gx_y + 1.f);
    float sin_theta = safe_sqrt(1.f - cos_theta * cos_theta);
    float phi = SSK_PI * (gx_x + 1.f), sp, cp;
    dsincos(phi, sp, cp);
    Vec3 n = { B.sun_n[0], B.sun_n[1], B.sun_n[2] };
    Vec3 wo = { sin_theta * cp, sin_theta * sp, cos_theta };
    RayTerms r = ray_terms<kExact>(cos_theta, unit_angle(n, wo));
    for (int c = 0; c < C; ++c) out[c] = render_sky(B, c, r) * gw_x * gw_y;
}
SSK_HD float quad_sky_jacobian() { return 0.5f * SSK_PI; }
/// Sun over its cone, [-1,1]^2 -> [0,2pi] x [cos(half aperture), 1] (sunsky.cpp:818-849)
SSK_HD void quad_sun_node(const HotBlock &B, const float *sun_rad, int C, float gx_x, float gx_y, float gw_x, float gw_y,
                          float *out) {
    const float cutoff = B.cos_cutoff;
    float cos_gamma = 0.5f * ((1.f - cutoff) * gx_y + (1.f + cutoff));
    float sin_gamma = safe_sqrt(1.f - cos_gamma * cos_gamma);
    float phi = SSK_PI * (gx_x + 1.f), sp, cp;
    dsincos(phi, sp, cp);
    Vec3 v = { sin_gamma * cp, sin_gamma * sp, cos_gamma };
    float gamma = unit_angle_z(v);
    Vec3 wo = { fmaf(B.sun_n[0], v.z, fmaf(B.sun_t[0], v.y, B.sun_s[0] * v.x)),
                fmaf(B.sun_n[1], v.z, fmaf(B.sun_t[1], v.y, B.sun_s[1] * v.x)),
                fmaf(B.sun_n[2], v.z, fmaf(B.sun_t[2], v.y, B.sun_s[2] * v.x)) };
    float cos_theta = wo.z;
    bool active = cos_theta >= 0.f;
    float xk[4], cj[6];
    SunSeg s = { 0, 0.f };
    if (active) {
        s = sun_segment(cos_theta);
        powers4(s.x, xk);
        powers6(cos_psi(B, gamma), cj);
    }
    for (int c = 0; c < C; ++c) {
        float v0 = 0.f;
        if (active) {
            if (B.spectral) {
                v0 = render_sun_spec(sun_rad, c, s, xk) * gw_x * gw_y;
                v0 *= sun_ld(B, c, c, 0.f, cj);
            } else {
                v0 = render_sun_rgb(sun_rad, c, s, xk, cj) * gw_x * gw_y;
            }
        }
        out[c] = v0;
    }
}
SSK_HD float quad_sun_jacobian(const HotBlock &B) { return 0.5f * SSK_PI * (1.f - B.cos_cutoff); }

// ---- sample_wavelengths / sample_ray (sunsky.cpp:463-480, 354-397) ---------------------------------------
/// ContinuousDistribution::sample_pdf (distr_1d.h:468-502) over the emitter's 10-node wavelength density.
/// U: constant-bank copy (uniform scalars), S: copy with lane-indexed tables (shared memory on the device).
SSK_HD void wavelength_sample_pdf(const HotBlock &U, const HotBlock &S, float sample, float &wav, float &pdf) {
    sample *= U.wl_integral;
    uint32_t start = (uint32_t) U.wl_valid_lo, end = (uint32_t) U.wl_valid_hi;
#pragma unroll
    for (int it = 0; it < 4; ++it) {                 // at most log2i(8) + 1 iterations (drjit/util.h:136-181)
        if (it < U.wl_iters) {
            uint32_t middle = (start + end) >> 1;
            bool cond = S.wl_cdf[middle] < sample;
            uint32_t nstart = middle + 1 < end ? middle + 1 : end;
            start = cond ? nstart : start;
            end = cond ? end : middle;
        }
    }
    const uint32_t index = start;
    float y0 = S.wl_pdf[index], y1 = S.wl_pdf[index + 1], c0 = index > 0 ? S.wl_cdf[index - 1] : 0.f;
    sample = (sample - c0) * U.wl_inv_interval;
    float t_linear = (y0 - safe_sqrt(fmaf(y0, y0, 2.f * sample * (y1 - y0)))) * S.wl_rdiff[index],
          t_const = sample * S.wl_ry0[index],
          t = y0 == y1 ? t_const : t_linear;
    wav = fmaf((float) index + t, U.wl_interval, 360.f);
    pdf = fmaf(t, y1 - y0, y0) * U.wl_norm;
}

/// sample_wavelengths, spectral (sunsky.cpp:466-473): math::sample_shifted<Wavelength> (core/math.h:408-415),
/// the density above per wavelength, weight = eval / pdf (Spectrum / Wavelength: element-wise division)
template <int A>
SSK_HD void sample_wavelengths_spec(const HotBlock &U, const HotBlock &S, const float2 *pairs, const float *sun_rad,
                                    const Vec3 &wi, float sample, bool active, float wav[4], float weight[4]) {
    float pdf[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float v = sample + 0.25f * (float) k;
        if (v > 1.f) v -= 1.f;
        wavelength_sample_pdf(U, S, v, wav[k], pdf[k]);
    }
    float e[4];
    eval_spec<A>(U, S, pairs, sun_rad, wi, wav, active, e);
#pragma unroll
    for (int k = 0; k < 4; ++k) weight[k] = A == kFast ? e[k] * hw_rcp(pdf[k]) : fdiv(e[k], pdf[k]);
}

/// warp::square_to_uniform_disk_concentric (warp.h:54-90)
SSK_HD void disk_concentric(float u0, float u1, float &px, float &py) {
    float x = fmaf(2.f, u0, -1.f), y = fmaf(2.f, u1, -1.f);
    bool is_zero = x == 0.f && y == 0.f, q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = fdiv(0.25f * SSK_PI * rp, r);
    if (q13) phi = SSK_HALF_PI - phi;
    if (is_zero) phi = 0.f;
    float s, c;
    dsincos(phi, s, c);
    px = r * c; py = r * s;
}

/// coordinate_system (core/vector.h:116-136), the s and t of Frame3f(n)
SSK_HD void coordinate_system_dev(const Vec3 &n, Vec3 &s, Vec3 &t) {
    float sign = copysignf(1.f, n.z), a = -fdiv(1.f, sign + n.z), b = n.x * n.y * a;   // dr::sign = copysign(1, v)
    s = { mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z) };
    t = { b, fmaf(n.y, n.y * a, sign), -n.y };
}

/// Ray origin of sample_ray (sunsky.cpp:390-392) from the ray direction and the concentric-disc point (ox, oy):
/// Frame3f(d_world).to_world((ox, oy, 0)) (frame.h:39-41), then center + (offset - d_world) * radius
SSK_HD Vec3 ray_origin(const HotBlock &U, const Vec3 &d_world, float ox, float oy) {
    Vec3 fs, ft;
    coordinate_system_dev(d_world, fs, ft);
    Vec3 perp = { fmaf(d_world.x, 0.f, fmaf(ft.x, oy, fs.x * ox)), fmaf(d_world.y, 0.f, fmaf(ft.y, oy, fs.y * ox)),
                  fmaf(d_world.z, 0.f, fmaf(ft.z, oy, fs.z * ox)) };
    return { U.bs_center[0] + (perp.x - d_world.x) * U.bs_radius, U.bs_center[1] + (perp.y - d_world.y) * U.bs_radius,
             U.bs_center[2] + (perp.z - d_world.z) * U.bs_radius };
}
/// Geometry part of sample_ray (sunsky.cpp:360-383, 390-392): direction, pdf (incl. the origin disc), origin
struct RaySample { Vec3 d_world, origin; float pdf; bool active; };
template <int A> SSK_HD RaySample sample_ray_geom(const HotBlock &U, const HotBlock &S, float s2x, float s2y, float s3x, float s3y) {
    float ox, oy;
    disk_concentric(s2x, s2y, ox, oy);
    bool pick_sky = s3x < U.sky_w;
    Vec3 d = pick_sky ? sample_sky(U, S, fdiv_by_safe(s3x, U.sky_w, U.r_sky_w), s3y)
                      : sample_sun(U, fdiv_by_safe(s3x - U.sky_w, U.one_minus_sky_w, U.r_one_minus_sky_w), s3y);
    RaySample r;
    r.d_world = xform(U.to_world, Vec3{ -d.x, -d.y, -d.z });
    r.active = d.z >= 0.f;
    r.pdf = combined_pdf<A>(U, d, pick_sky, r.active) * U.ray_pdf_scale;
    r.active = r.active && r.pdf > 0.f;
    r.origin = ray_origin(U, r.d_world, ox, oy);
    return r;
}

// ---- two samples per thread: sample_direction ------------------------------------------------------------
SSK_HD Vec3p sphdir_p(float2 theta, float2 phi) {
    float2 st, ct, sp, cp;
    dsincos_2(theta, st, ct);
    dsincos_2(phi, sp, cp);
    return { mul2(cp, st), mul2(sp, st), ct };
}
/// sample_sky for two samples (see sample_sky)
SSK_HD Vec3p sample_sky_p(const HotBlock &U, const HotBlock &S, float2 u0, float2 u1) {
    float t0, t1;
    int i0 = tgmm_pick(U, S, u0.x, t0), i1 = tgmm_pick(U, S, u0.y, t1);
    float2 sx = lerp2(f2(S.g_cdf_a_phi[i0], S.g_cdf_a_phi[i1]), f2(S.g_cdf_b_phi[i0], S.g_cdf_b_phi[i1]), f2(t0, t1)),
           sy = lerp2(f2(S.g_cdf_a_theta[i0], S.g_cdf_a_theta[i1]), f2(S.g_cdf_b_theta[i0], S.g_cdf_b_theta[i1]), u1);
    sx = f2(fmaxf(fminf(sx.x, SSK_ONE_MINUS_EPSILON), SSK_EPSILON), fmaxf(fminf(sx.y, SSK_ONE_MINUS_EPSILON), SSK_EPSILON));
    sy = f2(fmaxf(fminf(sy.x, SSK_ONE_MINUS_EPSILON), SSK_EPSILON), fmaxf(fminf(sy.y, SSK_ONE_MINUS_EPSILON), SSK_EPSILON));
    float2 ex = mul2s(derfinv_2(add2x(mul2s(sx, 2.f), splat2(-1.f))), SSK_SQRT_TWO),
           ey = mul2s(derfinv_2(add2x(mul2s(sy, 2.f), splat2(-1.f))), SSK_SQRT_TWO);
    float2 phi = add2x(mul2(ex, f2(S.g_sigma_phi[i0], S.g_sigma_phi[i1])), f2(S.g_mu_phi[i0], S.g_mu_phi[i1])),
           theta = add2x(mul2(ey, f2(S.g_sigma_theta[i0], S.g_sigma_theta[i1])), f2(S.g_mu_theta[i0], S.g_mu_theta[i1]));
    phi = add2s(phi, U.sun_phi_shift);
    theta = f2(fminf(theta.x, SSK_HALF_PI - SSK_EPSILON), fminf(theta.y, SSK_HALF_PI - SSK_EPSILON));
    return sphdir_p(theta, phi);
}
/// sample_sun for two samples (see sample_sun)
SSK_HD Vec3p sample_sun_p(const HotBlock &P, float2 u0, float2 u1) {
    float2 x = fma2(splat2(2.f), u0, splat2(-1.f)), y = fma2(splat2(2.f), u1, splat2(-1.f));
    bool z0 = x.x == 0.f && y.x == 0.f, z1 = x.y == 0.f && y.y == 0.f;
    bool q0 = fabsf(x.x) < fabsf(y.x), q1 = fabsf(x.y) < fabsf(y.y);
    float2 r = sel2(q0, q1, y, x), rp = sel2(q0, q1, x, y);
    float2 phi = fdiv2(mul2(splat2(0.25f * SSK_PI), rp), r);
    float2 alt = add2(splat2(SSK_HALF_PI), neg2(phi));
    phi = sel2(q0, q1, alt, phi);
    phi = sel2(z0, z1, f2(0.f, 0.f), phi);
    float2 s, c;
    dsincos_2(phi, s, c);
    float2 px = mul2(r, c), py = mul2(r, s);
    float2 pn = fma2(py, py, mul2(px, px));
    float2 z = add2x(mul2s(add2(splat2(1.f), neg2(pn)), P.one_minus_cos_cutoff), splat2(P.cos_cutoff));
    float2 k = safe_sqrt2(mul2s(subp2(splat2(2.f), mul2s(pn, P.one_minus_cos_cutoff)), P.one_minus_cos_cutoff));
    px = mul2(px, k); py = mul2(py, k);
    return { fma2(splat2(P.sun_n[0]), z, fma2(splat2(P.sun_t[0]), py, mul2s(px, P.sun_s[0]))),
             fma2(splat2(P.sun_n[1]), z, fma2(splat2(P.sun_t[1]), py, mul2s(px, P.sun_s[1]))),
             fma2(splat2(P.sun_n[2]), z, fma2(splat2(P.sun_t[2]), py, mul2s(px, P.sun_s[2]))) };
}
/// Direction part of sample_direction for two samples (see sample_dir)
struct DirSampleP { Vec3p d_world; float2 pdf; bool active0, active1; };
template <int A> SSK_HD DirSampleP sample_dir_p(const HotBlock &U, const HotBlock &S, float2 u0, float2 u1) {
    bool sky0 = u0.x < U.sky_w, sky1 = u0.y < U.sky_w;
    Vec3p dir = { f2(0.f, 0.f), f2(0.f, 0.f), f2(0.f, 0.f) };
    if (sky0 || sky1) {
        float2 v = f2(sky0 ? fdiv_by_safe(u0.x, U.sky_w, U.r_sky_w) : 0.5f, sky1 ? fdiv_by_safe(u0.y, U.sky_w, U.r_sky_w) : 0.5f);
        dir = sample_sky_p(U, S, v, u1);
    }
    if (!sky0 || !sky1) {
        float2 v = f2(sky0 ? 0.5f : fdiv_by_safe(u0.x - U.sky_w, U.one_minus_sky_w, U.r_one_minus_sky_w),
                      sky1 ? 0.5f : fdiv_by_safe(u0.y - U.sky_w, U.one_minus_sky_w, U.r_one_minus_sky_w));
        Vec3p ds = sample_sun_p(U, v, u1);
        dir = { sel2(sky0, sky1, dir.x, ds.x), sel2(sky0, sky1, dir.y, ds.y), sel2(sky0, sky1, dir.z, ds.z) };
    }
    DirSampleP r;
    r.active0 = dir.z.x >= 0.f; r.active1 = dir.z.y >= 0.f;
    r.d_world = xform_p(U.to_world, dir);
    r.pdf = combined_pdf_p<A>(U, dir, sky0, sky1, r.active0, r.active1);
    return r;
}

} // namespace ssk
