// Host setup of a sunsky emitter: see ssk_host.cuh. Compiled by nvcc with -fmad=false and the host
// compiler's contraction off, so the fp32 op order below is the one that executes.
#include "ssk_host.cuh"
#include <map>
#include <mutex>
#include <cstring>

namespace ssk {

static std::string fmt(const char *f, double v) { char b[256]; snprintf(b, sizeof(b), f, v); return b; }

/// Property validation of init_from_props / parameters_changed (sunsky.cpp:243-252, 889-948, 974-975)
void validate_params(const ssk_params &p, bool spectral) {
    if (!(p.sun_scale >= 0.f)) throw std::invalid_argument(fmt("Invalid sun scale: %f, must be positive!", p.sun_scale));
    if (!(p.sky_scale >= 0.f)) throw std::invalid_argument(fmt("Invalid sky scale: %f, must be positive!", p.sky_scale));
    if (!(p.turbidity >= 1.f && p.turbidity <= 10.f))
        throw std::invalid_argument(fmt("Turbidity value %f is out of range [1, 10]", p.turbidity));
    float half = 0.5f * p.sun_aperture_deg * (float) (3.14159265358979323846 / 180.0);
    if (!(half > 0.f && half < SSK_HALF_PI))
        throw std::invalid_argument(fmt("Invalid sun aperture angle: %f, must be in ]0, 90[ degrees!", p.sun_aperture_deg));
    int want = spectral ? 11 : 3;
    if (p.n_albedo != 1 && p.n_albedo != want)
        throw std::invalid_argument("albedo must hold 1 value or one value per channel (3 RGB / 11 spectral)");
    for (int i = 0; i < p.n_albedo; ++i)
        if (!(p.albedo[i] >= 0.f && p.albedo[i] <= 1.f))
            throw std::invalid_argument(fmt("Albedo values must be in [0, 1], got: %f", p.albedo[i]));
    if (!p.use_record) {
        float n2 = p.sun_direction[0] * p.sun_direction[0] + p.sun_direction[1] * p.sun_direction[1] +
                   p.sun_direction[2] * p.sun_direction[2];
        if (!(n2 > 0.f) || !std::isfinite(n2)) throw std::invalid_argument("sun_direction must be a non-zero vector");
    }
}

/// dr::normalize on a ScalarVector3f (sunsky.cpp:922): v * rsqrt(sum(v * v)); the 3-wide horizontal sum adds
/// (x^2 + z^2) + y^2 (drjit/packet_sse42.h:1659-1666) and a scalar rsqrt is 1 / sqrt (drjit/array_utils.h:213-219)
Vec3 normalize_sun_direction(const Vec3 &v) {
    float inv = 1.f / sqrtf((v.x * v.x + v.z * v.z) + v.y * v.y);
    return { v.x * inv, v.y * inv, v.z * inv };
}

static void mat3_from_4x4(const float *m16, float *m9) {
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) m9[r * 3 + c] = m16[r * 4 + c];
}
/// The inverse a `to_world` property carries in the reference. Properties stores transforms as
/// Transform<Point<double, 4>> (include/mitsuba/core/properties.h:70), whose constructor computes
/// dr::inverse_transpose in fp64 (transform.h:55-57); reading it back as ScalarTransform4f casts both matrices to
/// fp32 (transform.h:64-67). This is drjit/matrix.h:440-500 (4 x 4 cofactor expansion on 4-wide packets) lane by
/// lane in the same order, so the rounded result is the one the reference ends up with. `inv` = row-major inverse.
namespace {
struct V4 { double v[4]; };
template <int A, int B, int C, int D> V4 shuf(const V4 &a) { return { { a.v[A], a.v[B], a.v[C], a.v[D] } }; }
V4 vmul(const V4 &a, const V4 &b) { V4 r; for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] * b.v[i]; return r; }
V4 vfmadd(const V4 &a, const V4 &b, const V4 &c) { V4 r; for (int i = 0; i < 4; ++i) r.v[i] = fma(a.v[i], b.v[i], c.v[i]); return r; }
V4 vfmsub(const V4 &a, const V4 &b, const V4 &c) { V4 r; for (int i = 0; i < 4; ++i) r.v[i] = fma(a.v[i], b.v[i], -c.v[i]); return r; }
V4 vfnmadd(const V4 &a, const V4 &b, const V4 &c) { V4 r; for (int i = 0; i < 4; ++i) r.v[i] = fma(-a.v[i], b.v[i], c.v[i]); return r; }
}
static bool invert4_f64(const float *m16, float *inv16) {
    V4 row0, row1, row2, row3;
    for (int c = 0; c < 4; ++c) { row0.v[c] = m16[c]; row1.v[c] = m16[4 + c]; row2.v[c] = m16[8 + c]; row3.v[c] = m16[12 + c]; }
    row1 = shuf<2, 3, 0, 1>(row1);
    row3 = shuf<2, 3, 0, 1>(row3);
    V4 temp, col0, col1, col2, col3;
    temp = shuf<1, 0, 3, 2>(vmul(row2, row3));
    col0 = vmul(row1, temp);
    col1 = vmul(row0, temp);
    temp = shuf<2, 3, 0, 1>(temp);
    col0 = vfmsub(row1, temp, col0);
    col1 = shuf<2, 3, 0, 1>(vfmsub(row0, temp, col1));

    temp = shuf<1, 0, 3, 2>(vmul(row1, row2));
    col0 = vfmadd(row3, temp, col0);
    col3 = vmul(row0, temp);
    temp = shuf<2, 3, 0, 1>(temp);
    col0 = vfnmadd(row3, temp, col0);
    col3 = shuf<2, 3, 0, 1>(vfmsub(row0, temp, col3));

    temp = shuf<1, 0, 3, 2>(vmul(shuf<2, 3, 0, 1>(row1), row3));
    row2 = shuf<2, 3, 0, 1>(row2);
    col0 = vfmadd(row2, temp, col0);
    col2 = vmul(row0, temp);
    temp = shuf<2, 3, 0, 1>(temp);
    col0 = vfnmadd(row2, temp, col0);
    col2 = shuf<2, 3, 0, 1>(vfmsub(row0, temp, col2));

    temp = shuf<1, 0, 3, 2>(vmul(row0, row1));
    col2 = vfmadd(row3, temp, col2);
    col3 = vfmsub(row2, temp, col3);
    temp = shuf<2, 3, 0, 1>(temp);
    col2 = vfmsub(row3, temp, col2);
    col3 = vfnmadd(row2, temp, col3);

    temp = shuf<1, 0, 3, 2>(vmul(row0, row3));
    col1 = vfnmadd(row2, temp, col1);
    col2 = vfmadd(row1, temp, col2);
    temp = shuf<2, 3, 0, 1>(temp);
    col1 = vfmadd(row2, temp, col1);
    col2 = vfnmadd(row1, temp, col2);

    temp = shuf<1, 0, 3, 2>(vmul(row0, row2));
    col1 = vfmadd(row3, temp, col1);
    col3 = vfnmadd(row1, temp, col3);
    temp = shuf<2, 3, 0, 1>(temp);
    col1 = vfnmadd(row3, temp, col1);
    col3 = vfmadd(row1, temp, col3);

    // dot() of two 4-wide double packets: sum(a * b) = (p0 + p2) + (p1 + p3) (drjit/packet_avx.h:738)
    V4 pr = vmul(row0, col0);
    double det = (pr.v[0] + pr.v[2]) + (pr.v[1] + pr.v[3]);
    if (det == 0.0 || !std::isfinite(det)) return false;
    double inv_det = 1.0 / det;
    const V4 *cols[4] = { &col0, &col1, &col2, &col3 };
    // cols[r] is row r of the inverse TRANSPOSE; the inverse itself is its transpose
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c)
            inv16[c * 4 + r] = (float) (cols[r]->v[c] * inv_det);
    return true;
}

/// The constructor (from_constructor = true, sunsky.cpp:162-218) or parameters_changed()
/// (false, sunsky.cpp:242-285) applied to `p`.
void build_emitter(HostEmitter &E, const ssk_params &p, bool from_constructor, const QuadratureFn &quad) {
    build_emitter_pre(E, p, from_constructor);
    float sky_rad[11], sun_rad[11];
    if (quad) quad(E, sky_rad, sun_rad); else quadrature_host(E, sky_rad, sun_rad);
    build_emitter_post(E, sky_rad, sun_rad);
}

/// Everything up to the quadrature: after this the block is complete except for the sampling weight and the
/// wavelength density (which need the two integrals)
void build_emitter_pre(HostEmitter &E, const ssk_params &p, bool from_constructor) {
    validate_params(p, E.spectral);
    if (!from_constructor && (p.use_record != 0) != E.active_record)
        throw std::invalid_argument("an emitter keeps the sun parametrisation (direction or date/location) it was created with");
    E.params = p;
    E.active_record = p.use_record != 0;
    EmitterBlock &B = E.block;
    const int C = E.n_channels;
    memset(static_cast<HotBlock *>(&B), 0, sizeof(HotBlock));
    B.spectral = E.spectral ? 1 : 0;
    B.arith = p.arith == SSK_ARITH_FAST ? kFast : kExact;

    // --- transforms
    mat3_from_4x4(p.to_world, B.to_world);
    if (p.flags & SSK_FLAG_TO_LOCAL_GIVEN) {
        // the caller's Transform4f carries its own inverse (transform.h:59-62, 77-79): use it verbatim
        mat3_from_4x4(p.to_local, B.to_local);
        for (int i = 0; i < 9; ++i)
            if (!std::isfinite(B.to_local[i])) throw std::invalid_argument("to_local is not finite");
    } else {
        float inv16[16];
        if (!invert4_f64(p.to_world, inv16)) throw std::invalid_argument("to_world is not invertible");
        mat3_from_4x4(inv16, B.to_local);
    }

    // --- albedo per channel (extract_albedo, sunsky.cpp:956-978)
    float albedo[11];
    for (int c = 0; c < C; ++c) albedo[c] = p.n_albedo == 1 ? p.albedo[0] : p.albedo[c];

    // --- sun direction (init_from_props :912-946, ctor :174-179, parameters_changed :257-267)
    E.half_aperture = 0.5f * p.sun_aperture_deg * (float) (3.14159265358979323846 / 180.0);
    Vec3 local_sun;
    if (E.active_record) {
        DateTime dt = { p.year, p.month, p.day, p.hour, p.minute, p.second };
        Location loc = { p.latitude, p.longitude, p.timezone };
        Vec3 coords = sun_coordinates(dt, loc);
        E.sun_below_horizon = coords.z < 0.f;
        E.sun_dir_world = xform(B.to_world, coords);
        local_sun = from_constructor ? xform(B.to_local, E.sun_dir_world) : coords;
    } else {
        Vec3 v = { p.sun_direction[0], p.sun_direction[1], p.sun_direction[2] };
        if (!(p.flags & SSK_FLAG_SUN_DIR_NORMALIZED) && from_constructor) {
            v = normalize_sun_direction(v);
        }
        E.sun_dir_world = v;
        local_sun = xform(B.to_local, v);
        E.sun_below_horizon = false;
    }
    E.sun_angles[0] = datan2(local_sun.y, local_sun.x);          // from_spherical (sunsky.h:83-89)
    E.sun_angles[1] = unit_angle_z(local_sun);
    Vec3 fs, ft;
    coordinate_system(local_sun, fs, ft);
    B.sun_n[0] = local_sun.x; B.sun_n[1] = local_sun.y; B.sun_n[2] = local_sun.z;
    B.sun_s[0] = fs.x; B.sun_s[1] = fs.y; B.sun_s[2] = fs.z;
    B.sun_t[0] = ft.x; B.sun_t[1] = ft.y; B.sun_t[2] = ft.z;
    B.sun_phi_shift = E.sun_angles[0] - SSK_HALF_PI;
    const float eta = SSK_HALF_PI - E.sun_angles[1];
    // A to_world that is not rigid leaves the local sun direction far from unit length and unit_angle_z returns NaN;
    // the reference then dies in the TGMM weights (NaN fails `pmf >= 0`, distr_1d.h:221-222). Same failure, said earlier.
    if (std::isnan(eta))
        throw std::invalid_argument("local sun direction is not a unit vector (to_world must be a rigid transform): "
                                    "DiscreteDistribution: entries must be non-negative!");

    // --- aperture-dependent constants
    B.cos_cutoff = dcos(E.half_aperture);
    B.one_minus_cos_cutoff = 1.f - B.cos_cutoff;
    float s = dsin(E.half_aperture);
    B.ar2 = 1.f / (s * s);
    // get_area_ratio (sunsky.h:98-101): fp64 numerator 1 - dr::cos(deg_to_rad(0.5358 / 2.0)) (value of
    // Dr.Jit's fp64 cosine polynomial, a compile-time constant of the reference), fp32 denominator
    const double area_num = 0x1.6eca9ffb60000p-17;
    E.area_ratio = B.area_ratio = (float) (area_num / (double) (1.f - dcos(E.half_aperture)));
    B.sun_pdf_const = fdiv(SSK_INV_TWO_PI, 1.f - B.cos_cutoff);
    B.sky_scale = p.sky_scale; B.sun_scale = p.sun_scale;
    B.bs_center[0] = E.bs_center[0]; B.bs_center[1] = E.bs_center[1]; B.bs_center[2] = E.bs_center[2];
    B.bs_radius = E.bs_radius;

    // --- sky coefficients and mean radiance (sunsky.cpp:185-186 / 270-271)
    radiance_params(E.data->sky_params, albedo, C, p.turbidity, eta, E.sky_params);
    radiance_params(E.data->sky_rad, albedo, C, p.turbidity, eta, E.sky_radiance);
    for (int c = 0; c < C; ++c) {
        for (int k = 0; k < kSkyParams; ++k) B.sky_params[k][c] = E.sky_params[c * 9 + k];
        B.sky_rad[c] = E.sky_radiance[c];
        // FAST mode coefficients (sky_channel_fast): exponents pre-multiplied by log2(e); 1 + A8^2 and 2 A8
        // rounded exactly as the reference rounds them; radiance folded with the scales
        const float *A = &E.sky_params[c * 9];
        const double log2e = 1.4426950408889634074;
        const double norm = E.spectral ? 1.0 : 1.0 / 106.7502593994140625;
        // c1 coefficients per SSK_C1_MODE (c1_fast in ssk_model.cuh)
#if SSK_C1_MODE == 0
        const float t0 = A[0], t1 = A[1], t11 = 0.f;
#elif SSK_C1_MODE == 1
        const float t0 = A[0], t1 = (float) ((double) A[1] * log2e), t11 = 0.f;
#else
        const float t0 = A[0] < 0.f ? -1.f : (A[0] > 0.f ? 1.f : 0.f), t1 = A[1],
                    t11 = A[0] != 0.f ? (float) log(fabs((double) A[0])) : -1e30f;
#endif
        // radiance scale folded into the c2 coefficients (one multiply and one table load less per evaluation)
        const double L = (double) E.sky_radiance[c] * (double) p.sky_scale * norm;
        const float fast[kFastCoefs] = { t0, t1, (float) (A[2] * L), (float) (A[3] * L), (float) ((double) A[4] * log2e),
                                         (float) (A[5] * L), (float) (A[6] * L), (float) (A[7] * L), 1.f + A[8] * A[8], 2.f * A[8],
                                         (float) L, t11 };
        for (int k = 0; k < kFastCoefs; ++k) B.sky_fast[k][c] = fast[k];
    }
    for (int c = C; c < kChanPad; ++c) B.sky_fast[8][c] = 1.f;      // pad channels: keep the -3/2 power finite
    // EXACT mode's per-ray guard (HotBlock::exact_lim_s): written so that a NaN anywhere disables it
    {
        float m1 = 0.f, m4 = 0.f;
        bool pow_ok = true;
        for (int c = 0; c < C; ++c) {
            const float *A = &E.sky_params[c * 9];
            m1 = fmaxf(m1, A[1]); m4 = fmaxf(m4, fabsf(A[4]));        // c1: only positive exponents are bounded by the guard
            const double a8 = fabs((double) A[8]), lo = (1.0 - a8) * (1.0 - a8), hi = (1.0 + a8) * (1.0 + a8);
            pow_ok = pow_ok && lo >= 0x1p-39 && hi <= 0x1p39 && fabsf(A[1]) <= 1e20f && A[4] == A[4] && fabsf(A[0]) <= 1e20f;
        }
        B.exact_lim_s = pow_ok ? (m1 > 0.f ? 86.f / m1 : INFINITY) : -1.f;
        B.exact_lim_gamma = pow_ok ? (m4 > 0.f ? 86.f / m4 : INFINITY) : -1.f;
    }
    // --- sun radiance table and limb darkening (:191-195 / 274)
    E.sun_radiance.resize(E.data->sun_rad.size() / 10);
    sun_params(E.data->sun_rad, p.turbidity, E.sun_radiance.data());
    memset(B.sun_rad, 0, sizeof(B.sun_rad));
    memcpy(B.sun_rad, E.sun_radiance.data(), E.sun_radiance.size() * sizeof(float));
    if (E.spectral)
        for (int c = 0; c < 11; ++c)
            for (int j = 0; j < kLd; ++j) B.sun_ld[j][c] = E.data->sun_ld[c * kLd + j];

    // --- TGMM mixture, CDF (JIT DiscreteDistribution: fp32 inclusive prefix sum, 1/sum), and the
    //     per-component constants the JIT folds at trace time (sunsky.cpp:201-204, 668-675, 742-755)
    tgmm_distribution(E.data->tgmm, p.turbidity, eta, E.gaussians);
    float acc = 0.f;
    bool any_pos = false;
    for (int i = 0; i < kGauss; ++i) {
        E.pmf[i] = E.gaussians[i * 5 + 4];
        // distr_1d.h:221-224 (a NaN weight -- local sun direction far from unit length under a sheared
        // to_world -- fails the same test in the reference)
        if (!(E.pmf[i] >= 0.f)) throw std::invalid_argument("DiscreteDistribution: entries must be non-negative!");
        any_pos |= E.pmf[i] > 0.f;
        acc = acc + E.pmf[i];
        E.cdf[i] = acc;
    }
    if (!any_pos) throw std::runtime_error("DiscreteDistribution: no probability mass found!");
    E.cdf_sum = E.cdf[kGauss - 1];
    E.cdf_norm = 1.f / E.cdf_sum;
    B.cdf_sum = E.cdf_sum;
    B.n_pdf = 0; B.pdf_generic = 0;
    auto finite_nz = [](float v) { return std::isfinite(v) && v != 0.f && fabsf(v) > 1e-30f && fabsf(v) < 1e30f; };
    for (int i = 0; i < kGauss; ++i) {
        float mu_p = E.gaussians[i * 5 + 0], mu_t = E.gaussians[i * 5 + 1],
              sg_p = E.gaussians[i * 5 + 2], sg_t = E.gaussians[i * 5 + 3];
        B.g_mu_phi[i] = mu_p; B.g_mu_theta[i] = mu_t; B.g_sigma_phi[i] = sg_p; B.g_sigma_theta[i] = sg_t;
        B.g_cdf_a_phi[i] = gaussian_cdf(mu_p, sg_p, 0.f);
        B.g_cdf_a_theta[i] = gaussian_cdf(mu_t, sg_t, 0.f);
        B.g_cdf_b_phi[i] = gaussian_cdf(mu_p, sg_p, SSK_TWO_PI);
        B.g_cdf_b_theta[i] = gaussian_cdf(mu_t, sg_t, SSK_HALF_PI);
        float volume = (B.g_cdf_b_phi[i] - B.g_cdf_a_phi[i]) * (B.g_cdf_b_theta[i] - B.g_cdf_a_theta[i]) * (sg_p * sg_t);
        B.g_cdf[i] = E.cdf[i];
        B.g_pmf_n[i] = E.pmf[i] * E.cdf_norm;
        B.g_rpmf_n[i] = 1.f / B.g_pmf_n[i];
        B.g_cdf_prev_n[i] = i > 0 ? E.cdf[i - 1] * E.cdf_norm : 0.f;
        // density table: a component of weight exactly 0 (integer turbidity / on-grid elevation) adds +0 to
        // the sum as long as its sigma and volume are ordinary numbers, so it is dropped
        const bool ordinary = finite_nz(sg_p) && finite_nz(sg_t) && finite_nz(volume);
        if (E.pmf[i] == 0.f && ordinary) continue;
        const int k = B.n_pdf++;
        B.p_mu_phi[k] = mu_p; B.p_mu_theta[k] = mu_t; B.p_sigma_phi[k] = sg_p; B.p_sigma_theta[k] = sg_t;
        B.p_rsigma_phi[k] = 1.f / sg_p; B.p_rsigma_theta[k] = 1.f / sg_t;
        B.p_weight[k] = E.pmf[i]; B.p_volume[k] = volume; B.p_rvolume[k] = 1.f / volume;
        if (!ordinary || !(E.pmf[i] <= 1e10f)) B.pdf_generic = 1;     // (weight bound: gauss_exp_2 in ssk_model.cuh)
        const double kk = sqrt(0.5 * 1.4426950408889634074);
        B.p_k_phi[k] = (float) (kk / (double) sg_p); B.p_k_theta[k] = (float) (kk / (double) sg_t);
        B.p_wv[k] = (float) ((double) E.pmf[i] * 0.15915494309189533577 / (double) volume);
    }
}

/// Sky / sun sampling weight and wavelength distribution (:207-209 / 282-284) from the two integrals
void build_emitter_post(HostEmitter &E, const float sky_rad[11], const float sun_rad[11]) {
    EmitterBlock &B = E.block;
    estimate_sky_sun_ratio(E, sky_rad, sun_rad);
    B.sky_w = E.sky_w; B.r_sky_w = 1.f / E.sky_w;
    B.one_minus_sky_w = 1.f - E.sky_w; B.r_one_minus_sky_w = 1.f / B.one_minus_sky_w;
    for (int i = 0; i < 10; ++i) { B.wl_pdf[i] = E.spectral_pdf[i]; B.wl_cdf[i] = E.spectral_cdf[i]; }
    for (int i = 0; i < 10; ++i) {
        const float y0 = B.wl_pdf[i], y1 = i + 1 < 10 ? B.wl_pdf[i + 1] : 0.f;
        B.wl_rdiff[i] = fdiv(1.f, y0 - y1); B.wl_ry0[i] = fdiv(1.f, y0);
    }
    B.wl_integral = E.spectral_integral; B.wl_norm = E.spectral_norm;
    B.wl_interval = (float) ((720.0 - 360.0) / 9.0); B.wl_inv_interval = 1.f / B.wl_interval;
    B.wl_valid_lo = E.spectral_valid[0]; B.wl_valid_hi = E.spectral_valid[1];
    // dr::binary_search trip count: log2i(end - start) + 1 when start < end (drjit/util.h:138-139)
    B.wl_iters = 0;
    for (uint32_t w = (uint32_t) (B.wl_valid_hi - B.wl_valid_lo); w; w >>= 1) ++B.wl_iters;
    set_ray_pdf_scale(B);
}

/// pdf of the origin offset in sample_ray (sunsky.cpp:381): InvPi * rcp(square(radius)), fp32
void set_ray_pdf_scale(HotBlock &B) { B.ray_pdf_scale = SSK_INV_PI * (1.f / (B.bs_radius * B.bs_radius)); }

std::shared_ptr<const Datasets> datasets_for(const std::string &dir, bool spectral) {
    static std::mutex mu;
    static std::map<std::pair<std::string, bool>, std::shared_ptr<const Datasets>> cache;
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_pair(dir, spectral);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    auto d = std::make_shared<const Datasets>(load_datasets(dir, spectral));
    cache[key] = d;
    return d;
}

const std::vector<float> &gauss_legendre_200() {
    static const std::vector<float> g = [] {
        std::vector<float> x, w, r;
        gauss_legendre(kQuadN, x, w);
        r = x; r.insert(r.end(), w.begin(), w.end());
        return r;
    }();
    return g;
}

/// Host execution of the two 200 x 200 quadratures (contexts without a device; the device path is
/// setup_quadrature_kernel + setup_sum_kernel in ssk_kernels.cuh, same per-node functions, same sum order)
void quadrature_host(const HostEmitter &E, float sky_rad[11], float sun_rad[11]) {
    const EmitterBlock &B = E.block;
    const int C = E.n_channels, N = kQuadN;
    const std::vector<float> &g = gauss_legendre_200();
    const float *gx = g.data(), *gw = g.data() + N;
    std::vector<float> buf((size_t) C * N * N);
    float v[kMaxChannels];
    for (int which = 0; which < 2; ++which) {
        for (int iy = 0; iy < N; ++iy)
            for (int ix = 0; ix < N; ++ix) {
                if (which == 0) quad_sky_node(B, C, gx[ix], gx[iy], gw[ix], gw[iy], v);
                else quad_sun_node(B, B.sun_rad, C, gx[ix], gx[iy], gw[ix], gw[iy], v);
                for (int c = 0; c < C; ++c) buf[(size_t) c * N * N + iy * N + ix] = v[c];
            }
        const float J = which == 0 ? quad_sky_jacobian() : quad_sun_jacobian(B);
        for (int c = 0; c < C; ++c)
            (which == 0 ? sky_rad : sun_rad)[c] = llvm_block_sum(&buf[(size_t) c * N * N], N * N) * J;
    }
}

/// estimate_sky_sun_ratio, array branch (sunsky.cpp:772-886): 200x200 Gauss-Legendre quadrature of the
/// sky over the hemisphere and of the sun over its cone (`sky_rad`, `sun_rad`), luminance ratio, and (spectral)
/// the per-bin mean radiance that drives wavelength sampling.
void estimate_sky_sun_ratio(HostEmitter &E, const float sky_rad[11], const float sun_rad[11]) {
    const EmitterBlock &B = E.block;

    float sky_lum = B.sky_scale, sun_lum = B.sun_scale;
    if (!E.spectral) {
        // luminance(Color3f) (include/mitsuba/core/spectrum.h:431-434)
        float ls = sky_rad[0] * 0.212671f + sky_rad[1] * 0.715160f + sky_rad[2] * 0.072169f;
        float lu = sun_rad[0] * 0.212671f + sun_rad[1] * 0.715160f + sun_rad[2] * 0.072169f;
        sky_lum = sky_lum * ls;
        sun_lum = (float) ((double) sun_lum * ((double) (lu * B.area_ratio) * 467.069280386));
    } else {
        // luminance(spectrum, wavelengths) = mean(cie1931_y(lambda) * value) (spectrum.h:415-429, 271-295).
        // The 11 bins sit on the 5 nm CIE grid, so the table lookup is exact: CIE 1931 2-degree y-bar at
        // 360, 400, ..., 720 nm; 320 nm lies outside [360, 830] and contributes 0.
        const float cie_y[11] = { 0.f, 0.000003917f, 0.000396f, 0.023f, 0.13902f, 0.71f, 0.995f, 0.631f,
                                  0.175f, 0.017f, 0.001047f };
        auto lum = [&](const float *v) {
            float acc = cie_y[0] * v[0];
            for (int c = 1; c < 11; ++c) acc += cie_y[c] * v[c];
            return acc * (1.f / 11.f);
        };
        sky_lum = sky_lum * lum(sky_rad);
        sun_lum = sun_lum * (lum(sun_rad) * B.area_ratio);
    }
    float res = fdiv(sky_lum, sky_lum + sun_lum);
    E.sky_w = std::isnan(res) ? 0.f : res;

    if (E.spectral) {
        // ContinuousDistribution(range = [360, 720], 10 bins) (distr_1d.h:283-...): scalar cdf build
        bool all_zero = true;
        for (int i = 0; i < 10; ++i) {
            E.spectral_pdf[i] = sun_rad[i + 1] + sky_rad[i + 1];
            all_zero &= E.spectral_pdf[i] == 0.f;
        }
        if (all_zero) for (int i = 0; i < 10; ++i) E.spectral_pdf[i] += 1.f;
        // compute_cdf_scalar (distr_1d.h:540-590): fp64 running trapezoid sum, cast to fp32 per entry;
        // m_valid = first / last interval with positive mass; m_integral = cdf[valid.y]
        double interval = (720.0 - 360.0) / 9.0, integral = 0.0;
        int lo = -1, hi = -1;
        for (int i = 0; i < 9; ++i) {
            double y0 = E.spectral_pdf[i], y1 = E.spectral_pdf[i + 1];
            double value = 0.5 * interval * (y0 + y1);
            integral += value;
            E.spectral_cdf[i] = (float) integral;
            if (value > 0.0) { if (lo < 0) lo = i; hi = i; }
        }
        E.spectral_cdf[9] = 0.f;
        if (lo < 0) { lo = 0; hi = 8; }               // unreachable: the reference throws "no probability mass"
        E.spectral_valid[0] = lo; E.spectral_valid[1] = hi;
        E.spectral_integral = E.spectral_cdf[hi];
        E.spectral_norm = 1.f / E.spectral_integral;  // dr::rcp(m_integral) in fp32
    } else {
        memset(E.spectral_pdf, 0, sizeof(E.spectral_pdf));
        memset(E.spectral_cdf, 0, sizeof(E.spectral_cdf));
        E.spectral_valid[0] = E.spectral_valid[1] = 0;
        E.spectral_integral = E.spectral_norm = 0.f;
    }
}

/// d compute_radiance_params / d eta (sunsky.h:157-231): only the Bezier weights C_i x^i (1 - x)^(5 - i) depend on the
/// sun elevation (x = cbrt(2 eta / pi)); the turbidity and albedo lerps that follow are linear. fp64.
static void radiance_params_deta(const std::vector<float> &dataset, const float *albedo, int n_channels, float turbidity,
                                 float eta, std::vector<double> &out) {
    const uint32_t dataset_size = (uint32_t) dataset.size();
    const uint32_t t_block = dataset_size / 10, a_block = t_block / 2, result_size = a_block / 6,
                   nb_params = result_size / (uint32_t) n_channels;
    out.assign(result_size, 0.0);
    if (!(eta > 0.f && eta <= SSK_HALF_PI)) return;
    const double x = cbrt(2.0 / 3.14159265358979323846 * (double) eta);
    if (x < 1e-6) return;
    const double dx = (2.0 / 3.14159265358979323846) / (3.0 * x * x);
    const double coefs[6] = { 1, 5, 10, 10, 5, 1 };
    double dw[6];
    for (int i = 0; i < 6; ++i) {
        const double a = i ? i * pow(x, i - 1) * pow(1.0 - x, 5 - i) : 0.0;
        const double b = i < 5 ? (5 - i) * pow(x, i) * pow(1.0 - x, 4 - i) : 0.0;
        dw[i] = coefs[i] * (a - b) * dx;
    }
    uint32_t t_high = (uint32_t) (int32_t) floorf(turbidity), t_low = t_high - 1;
    const double t_rem = (double) turbidity - (double) t_high;
    auto bez = [&](uint32_t t_idx, uint32_t a_idx, uint32_t i) {
        if (t_idx >= 10) return 0.0;
        double acc = 0.0;
        for (int c = 0; c < 6; ++c) acc += dw[c] * (double) dataset[t_idx * t_block + a_idx * a_block + c * result_size + i];
        return acc;
    };
    for (uint32_t i = 0; i < result_size; ++i) {
        const double lo = bez(t_low, 0, i) + (bez(t_high, 0, i) - bez(t_low, 0, i)) * t_rem,
                     hi = bez(t_low, 1, i) + (bez(t_high, 1, i) - bez(t_low, 1, i)) * t_rem;
        out[i] = lo + (hi - lo) * (double) albedo[i / nb_params];
    }
}

/// State tangent for the direction (d_turbidity, d_albedo[]): the interpolated state is linear in turbidity inside a
/// slab [t, t+1] and linear in each channel's albedo (sunsky.h:157-231 for the sky, :403-419 for the sun), so the
/// exact derivative is a difference of states rebuilt at the slab ends / at albedo 0 and 1.
void state_tangent(const HostEmitter &E, float d_turbidity, const float *d_albedo, int n_d_albedo, const float *d_sun_direction,
                   TangentBlock &T) {
    memset(&T, 0, sizeof(T));
    const int C = E.n_channels;
    if (d_sun_direction && (d_sun_direction[0] != 0.f || d_sun_direction[1] != 0.f || d_sun_direction[2] != 0.f)) {
        // local sun direction v = to_world^-1 sun_dir (linear), theta = unit_angle_z(v) differentiated as written
        // (drjit/sphere.h:73-81), eta = pi/2 - theta, then the Bezier weights of both sky tables
        const EmitterBlock &B = E.block;
        Vec3 dv = xform(B.to_local, Vec3{ d_sun_direction[0], d_sun_direction[1], d_sun_direction[2] });
        T.d_sun_n[0] = dv.x; T.d_sun_n[1] = dv.y; T.d_sun_n[2] = dv.z;
        const double vx = B.sun_n[0], vy = B.sun_n[1], vz = B.sun_n[2], sg = vz >= 0.0 ? 1.0 : -1.0;
        const double D2 = vx * vx + vy * vy + (vz - sg) * (vz - sg);
        double d_eta = 0.0;
        if (D2 > 0.0) {
            const double D = sqrt(D2), dD = (vx * dv.x + vy * dv.y + (vz - sg) * dv.z) / D;
            d_eta = -sg * dD / sqrt(std::max(1.0 - 0.25 * D2, 1e-300));
        }
        if (d_eta != 0.0) {
            float albedo[11];
            for (int c = 0; c < C; ++c) albedo[c] = E.params.n_albedo == 1 ? E.params.albedo[0] : E.params.albedo[c];
            const float eta = SSK_HALF_PI - E.sun_angles[1];
            std::vector<double> dp, dr;
            radiance_params_deta(E.data->sky_params, albedo, C, E.params.turbidity, eta, dp);
            radiance_params_deta(E.data->sky_rad, albedo, C, E.params.turbidity, eta, dr);
            for (int c = 0; c < C; ++c) {
                for (int k = 0; k < kSkyParams; ++k) T.sky_params[k][c] += (float) (d_eta * dp[c * 9 + k]);
                T.sky_rad[c] += (float) (d_eta * dr[c]);
            }
        }
    }
    auto rebuilt = [&](float turbidity, float albedo) {
        HostEmitter H = E;
        ssk_params p = E.params;
        if (turbidity > 0.f) p.turbidity = turbidity;
        if (albedo >= 0.f) { p.n_albedo = 1; p.albedo[0] = albedo; }
        build_emitter_pre(H, p, false);
        return H;
    };
    auto axpy = [&](float w, const HostEmitter &hi, const HostEmitter &lo, int c0, int c1) {
        for (int c = c0; c < c1; ++c) {
            for (int k = 0; k < kSkyParams; ++k) T.sky_params[k][c] += w * (hi.block.sky_params[k][c] - lo.block.sky_params[k][c]);
            T.sky_rad[c] += w * (hi.block.sky_rad[c] - lo.block.sky_rad[c]);
        }
    };
    if (d_turbidity != 0.f) {
        // slab [floor(t), floor(t) + 1]; reads above turbidity 10 are masked to zero (sunsky.h:176-186, :403-419)
        float t = fminf(fmaxf(E.params.turbidity, 1.f), 10.f), t_lo = floorf(t);
        HostEmitter lo = rebuilt(t_lo, -1.f), hi = lo;
        if (t_lo < 10.f) hi = rebuilt(t_lo + 1.f, -1.f);
        else {                                     // both tables read 0 above the last slab
            memset(hi.block.sky_params, 0, sizeof(hi.block.sky_params));
            memset(hi.block.sky_rad, 0, sizeof(hi.block.sky_rad));
            memset(hi.block.sun_rad, 0, sizeof(hi.block.sun_rad));
        }
        axpy(d_turbidity, hi, lo, 0, C);
        for (size_t i = 0; i < sizeof(T.sun_rad) / sizeof(float); ++i)
            T.sun_rad[i] += d_turbidity * (hi.block.sun_rad[i] - lo.block.sun_rad[i]);
    }
    bool any = false;
    for (int c = 0; c < n_d_albedo; ++c) any = any || d_albedo[c] != 0.f;
    if (any) {
        HostEmitter a0 = rebuilt(-1.f, 0.f), a1 = rebuilt(-1.f, 1.f);
        for (int c = 0; c < C; ++c) axpy(n_d_albedo == 1 ? d_albedo[0] : d_albedo[c], a1, a0, c, c + 1);
    }
}

} // namespace ssk
