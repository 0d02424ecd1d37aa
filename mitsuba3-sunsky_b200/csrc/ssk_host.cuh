// Host-side setup of one sunsky emitter: dataset loading, Bezier / turbidity / albedo interpolation,
// sun position, TGMM mixture + CDF, and the sky/sun sampling weight. Mirrors the constructor and
// parameters_changed() of the reference plugin (src/emitters/sunsky.cpp:162-285) and the helpers in
// include/mitsuba/render/sunsky/sunsky.h. All of it is a few thousand flops except the 200x200
// Gauss-Legendre quadrature, which runs the same per-lane functions as the kernels.
#pragma once
#include "ssk_model.cuh"
#include "../../include/sunsky_b200.h"
#include <string>
#include <functional>
#include <memory>
#include <vector>
#include <stdexcept>
#include <cstdio>
#include <cmath>

namespace ssk {

struct Datasets {
    std::vector<float> sky_params, sky_rad, sun_rad, sun_ld, tgmm;
    bool spectral = false;
};

/// array_from_file (sunsky.h:515-561): "SKY"/"SUN" tag, u32 version, u64 ndim, u64 shape[], payload.
/// The element type is chosen by the caller, not the file, exactly as in the reference.
inline std::vector<float> read_dataset(const std::string &path, bool f64, std::vector<size_t> *shape_out = nullptr) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("\"" + path + "\": file does not exist!");
    auto fail = [&](const char *m) { fclose(f); throw std::runtime_error(path + ": " + m); };
    char tag[4] = { 0, 0, 0, 0 };
    if (fread(tag, 1, 3, f) != 3) fail("truncated header");
    if (strcmp(tag, "SKY") != 0 && strcmp(tag, "SUN") != 0) fail("not a sunsky dataset (bad tag)");
    uint32_t version; uint64_t ndim;
    if (fread(&version, 4, 1, f) != 1 || fread(&ndim, 8, 1, f) != 1 || ndim == 0 || ndim > 8) fail("bad header");
    size_t count = 1;
    std::vector<size_t> shape(ndim);
    for (uint64_t i = 0; i < ndim; ++i) {
        uint64_t d;
        if (fread(&d, 8, 1, f) != 1) fail("truncated shape");
        if (!d) fail("Got dimension with 0 elements");
        shape[i] = (size_t) d; count *= (size_t) d;
    }
    std::vector<float> out(count);
    if (f64) {
        std::vector<double> tmp(count);
        if (fread(tmp.data(), 8, count, f) != count) fail("truncated payload");
        for (size_t i = 0; i < count; ++i) out[i] = (float) tmp[i];
    } else {
        if (fread(out.data(), 4, count, f) != count) fail("truncated payload");
    }
    fclose(f);
    if (shape_out) *shape_out = shape;
    return out;
}

inline Datasets load_datasets(const std::string &dir, bool spectral) {
    // path_to_dataset (sunsky.h:124-141); element types as in sunsky.cpp:182-199
    std::string d = dir;
    if (!d.empty() && d.back() != '/') d += '/';
    const char *t = spectral ? "_spec_" : "_rgb_";
    Datasets ds;
    ds.spectral = spectral;
    ds.sky_params = read_dataset(d + "sky" + t + "params.bin", true);
    ds.sky_rad = read_dataset(d + "sky" + t + "rad.bin", true);
    ds.sun_rad = read_dataset(d + "sun" + t + "rad.bin", true);
    if (spectral) ds.sun_ld = read_dataset(d + "sun_spec_ld.bin", true);
    ds.tgmm = read_dataset(d + "tgmm_tables.bin", false);
    const size_t C = spectral ? 11 : 3;
    if (ds.sky_params.size() != 10 * 2 * 6 * C * 9 || ds.sky_rad.size() != 10 * 2 * 6 * C ||
        ds.sun_rad.size() != 10 * 45 * C * 4 * (spectral ? 1 : 6) || ds.tgmm.size() != 9 * 30 * 5 * 5 ||
        (spectral && ds.sun_ld.size() != 66))
        throw std::runtime_error("sunsky dataset has an unexpected shape");
    return ds;
}

/// detail::powi (drjit/math.h:990-1005)
inline float powi(float x, int n) {
    float r = 1.f;
    while (n) { if (n & 1) r *= x; x *= x; n >>= 1; }
    return r;
}

/// compute_radiance_params + bezier_interpolate (sunsky.h:157-231)
inline void radiance_params(const std::vector<float> &dataset, const float *albedo, int n_channels,
                            float turbidity, float eta, float *out /* result_size */) {
    const uint32_t dataset_size = (uint32_t) dataset.size();
    float x = dcbrt(2.f * SSK_INV_PI * eta);
    uint32_t t_high = (uint32_t) (int32_t) floorf(turbidity), t_low = t_high - 1;
    float t_rem = turbidity - (float) t_high;
    const uint32_t t_block = dataset_size / 10, a_block = t_block / 2, result_size = a_block / 6,
                   nb_params = result_size / (uint32_t) n_channels;
    const float coefs[6] = { 1, 5, 10, 10, 5, 1 };
    auto bezier = [&](uint32_t offset, bool active, std::vector<float> &res) {
        res.assign(result_size, 0.f);
        for (int c = 0; c < 6; ++c) {
            float wgt = coefs[c] * powi(x, c) * powi(1.f - x, 5 - c);
            for (uint32_t i = 0; i < result_size; ++i) {
                float data = active ? dataset[offset + i + c * result_size] : 0.f;
                res[i] += wgt * data;
            }
        }
    };
    std::vector<float> ll, hl, lh, hh;
    bezier(t_low * t_block + 0 * a_block, t_low < 10, ll);
    bezier(t_high * t_block + 0 * a_block, t_high < 10, hl);
    bezier(t_low * t_block + 1 * a_block, t_low < 10, lh);
    bezier(t_high * t_block + 1 * a_block, t_high < 10, hh);
    bool in_range = 0.f <= eta && eta <= SSK_HALF_PI;
    for (uint32_t i = 0; i < result_size; ++i) {
        float a_low = lerp(ll[i], hl[i], t_rem), a_high = lerp(lh[i], hh[i], t_rem);
        float r = lerp(a_low, a_high, albedo[i / nb_params]);
        out[i] = in_range ? r : 0.f;
    }
}

/// compute_sun_params (sunsky.h:403-419)
inline void sun_params(const std::vector<float> &dataset, float turbidity, float *out) {
    uint32_t t_high = (uint32_t) (int32_t) floorf(turbidity), t_low = t_high - 1;
    float t_rem = turbidity - (float) t_high;
    const uint32_t t_block = (uint32_t) dataset.size() / 10;
    for (uint32_t i = 0; i < t_block; ++i) {
        float lo = t_low < 10 ? dataset[t_low * t_block + i] : 0.f,
              hi = t_high < 10 ? dataset[t_high * t_block + i] : 0.f;
        out[i] = lerp(lo, hi, t_rem);
    }
}

/// compute_tgmm_distribution (sunsky.h:437-501): 4 neighbouring 5-Gaussian mixtures, weights scaled
/// by the bilinear (turbidity, elevation) factors. out: 20 x {mu_phi, mu_theta, sigma_phi, sigma_theta, w}
inline void tgmm_distribution(const std::vector<float> &tables, float turbidity, float eta, float *out100) {
    float eta_deg = eta * (float) (180.0 / 3.14159265358979323846);       // dr::rad_to_deg
    float eta_idx_f = fmaxf(fminf((eta_deg - 2.f) * (1.f / 3.f), 29.f), 0.f)   /* Float / int -> mul by rcp */,
          t_idx_f = fmaxf(fminf(turbidity - 2.f, 8.f), 0.f);
    uint32_t eta_lo = (uint32_t) (int32_t) floorf(eta_idx_f), t_lo = (uint32_t) (int32_t) floorf(t_idx_f);
    uint32_t eta_hi = eta_lo + 1 < 29u ? eta_lo + 1 : 29u, t_hi = t_lo + 1 < 8u ? t_lo + 1 : 8u;
    float eta_rem = eta_idx_f - (float) eta_lo, t_rem = t_idx_f - (float) t_lo;
    const uint32_t t_block = (uint32_t) tables.size() / 9, result_size = t_block / 30;   // 750, 25
    const uint32_t idx[4] = { t_lo * t_block + eta_lo * result_size, t_lo * t_block + eta_hi * result_size,
                              t_hi * t_block + eta_lo * result_size, t_hi * t_block + eta_hi * result_size };
    const float fac[4] = { (1.f - t_rem) * (1.f - eta_rem), (1.f - t_rem) * eta_rem,
                           t_rem * (1.f - eta_rem), t_rem * eta_rem };
    for (int m = 0; m < 4; ++m)
        for (uint32_t i = 0; i < result_size; ++i) {
            float v = tables[idx[m] + i];
            if (i % 5 == 4) v = v * fac[m];
            out100[m * result_size + i] = v;
        }
}

/// gaussian_cdf (sunsky.h:112-115), array erf branch
inline float gaussian_cdf(float mu, float sigma, float x) {
    return 0.5f * (1.f + derf(fdiv(SSK_INV_SQRT_TWO * (x - mu), sigma)));
}

struct DateTime { int32_t year, month, day; float hour, minute, second; };
struct Location { float latitude, longitude, timezone; };

/// compute_sun_coordinates (sunsky.h:283-374), Blanco-Muriel et al. in fp32 as the reference runs it
inline Vec3 sun_coordinates(const DateTime &dt, const Location &loc) {
    const float deg2rad = (float) (3.14159265358979323846 / 180.0);
    // Float / scalar divisions are a1 * rcp(a2) in Dr.Jit (drjit/array_router.h:225-228)
    float dec_hours = dt.hour - loc.timezone + (dt.minute + dt.second * (1.f / 60.f)) * (1.f / 60.f);
    int32_t aux1 = (dt.month - 14) / 12;
    int32_t aux2 = (1461 * (dt.year + 4800 + aux1)) / 4 + (367 * (dt.month - 2 - 12 * aux1)) / 12 -
                   (3 * ((dt.year + 4900 + aux1) / 100)) / 4 + dt.day - 32075;
    float julian = (float) aux2 - 0.5f + dec_hours * (1.f / 24.f);
    float elapsed = julian - 2451545.f;

    float omega = 2.1429f - 0.0010394594f * elapsed;
    float mean_longitude = 4.8950630f + 0.017202791698f * elapsed;
    float anomaly = 6.2400600f + 0.0172019699f * elapsed;
    float ecl_long = mean_longitude + 0.03341607f * dsin(anomaly) + 0.00034894f * dsin(2.f * anomaly) -
                     0.0001134f - 0.0000203f * dsin(omega);
    float ecl_obl = 0.4090928f - 6.2140e-9f * elapsed + 0.0000396f * dcos(omega);

    float sin_ecl = dsin(ecl_long);
    float dy = dcos(ecl_obl) * sin_ecl, dx = dcos(ecl_long);
    float ra = datan2(dy, dx);
    ra += ra < 0.f ? SSK_TWO_PI : 0.f;
    float decl = dasin(dsin(ecl_obl) * sin_ecl);

    float gmst = 6.6974243242f + 0.0657098283f * elapsed + dec_hours;
    float lmst = (gmst * 15.f + loc.longitude) * deg2rad;
    float lat = loc.latitude * deg2rad;
    float cos_lat = dcos(lat), sin_lat = dsin(lat);
    float hour_angle = lmst - ra, cos_ha = dcos(hour_angle);
    float elevation = dacos(cos_lat * cos_ha * dcos(decl) + dsin(decl) * sin_lat);
    dy = -dsin(hour_angle);
    dx = dtan(decl) * cos_lat - sin_lat * cos_ha;
    float azimuth = datan2(dy, dx);
    azimuth += azimuth < 0.f ? SSK_TWO_PI : 0.f;
    // parallax correction: the ratio is a double constant, so the update happens in fp64
    elevation = (float) ((double) elevation + (6371.01 / 149597890.0) * (double) dsin(elevation));
    return sphdir(elevation, azimuth - SSK_PI);
}

/// coordinate_system (include/mitsuba/core/vector.h:116-136), as used by Frame3f(v)
inline void coordinate_system(const Vec3 &n, Vec3 &s, Vec3 &t) {
    float sign = copysignf(1.f, n.z), a = -(1.f / (sign + n.z)), b = n.x * n.y * a;
    s = { mulsign(n.x * n.x * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z) };
    t = { b, fmaf(n.y, n.y * a, sign), -n.y };
}

/// quad::gauss_legendre(n) (include/mitsuba/core/quad.h:27-86): Newton iteration in fp64 on the
/// Legendre recurrence (core/math.h:91-120), rounded to fp32 at the end
inline void gauss_legendre(int n_, std::vector<float> &nodes, std::vector<float> &weights) {
    int n = n_;
    nodes.assign(n, 0.f); weights.assign(n, 0.f);
    auto legendre_pd = [](int l, double x, double &p, double &d) {
        double l_cur = 0, d_cur = 0;
        if (l > 1) {
            double l_pp = 1, l_p = x, d_pp = 0, d_p = 1, k0 = 3, k1 = 2, k2 = 1;
            for (int ki = 2; ki <= l; ++ki) {
                l_cur = (k0 * x * l_p - k2 * l_pp) / k1;
                d_cur = d_pp + k0 * l_p;
                l_pp = l_p; l_p = l_cur; d_pp = d_p; d_p = d_cur;
                k2 = k1; k0 += 2; k1 += 1;
            }
        } else if (l == 0) { l_cur = 1; d_cur = 0; } else { l_cur = x; d_cur = 1; }
        p = l_cur; d = d_cur;
    };
    n--;
    if (n == 0) { nodes[0] = 0.f; weights[0] = 2.f; return; }
    int m = (n + 1) / 2;
    for (int i = 0; i < m; ++i) {
        double x = -cos((double) (2 * i + 1) / (double) (2 * n + 2) * 3.14159265358979323846);
        for (int it = 0;; ) {
            if (++it > 20) throw std::runtime_error("gauss_legendre(): did not converge");
            double p, d; legendre_pd(n + 1, x, p, d);
            double step = p / d;
            x -= step;
            if (fabs(step) <= 4 * fabs(x) * 1.1102230246251565e-16) break;
        }
        double p, d; legendre_pd(n + 1, x, p, d);
        weights[i] = weights[n - i] = (float) (2 / ((1 - x * x) * (d * d)));
        nodes[i] = (float) x; nodes[n - i] = (float) -x;
    }
    if ((n % 2) == 0) {
        double p, d; legendre_pd(n + 1, 0.0, p, d);
        weights[n / 2] = (float) (2.0 / (d * d));
        nodes[n / 2] = 0.f;
    }
}

/// fp32 sum in the order the LLVM backend's block reduction uses (drjit-core src/llvm_ts.cpp:232-316,
/// llvm_red.h:78-116): sequential within 16384-wide chunks when size > 32768, then over the partials
inline float llvm_block_sum(const float *a, uint32_t size) {
    const uint32_t ideal = 16384;
    uint32_t chunk = size > 2 * ideal ? ideal : size;
    float total = 0.f;
    for (uint32_t start = 0; start < size; start += chunk) {
        uint32_t end = start + chunk < size ? start + chunk : size;
        float acc = 0.f;
        for (uint32_t j = start; j < end; ++j) acc = acc + a[j];
        if (chunk == size) return acc;
        total = total + acc;
    }
    return total;
}

/// Everything the C ABI reports back plus the device block
struct HostEmitter {
    ssk_params params;
    bool spectral = false;
    int n_channels = 3;
    std::shared_ptr<const Datasets> data;      // shared by every emitter created from the same directory
    EmitterBlock block;
    // reference-layout copies for ssk_get_state
    float sky_params[99], sky_radiance[11], gaussians[100], pmf[20], cdf[20];
    std::vector<float> sun_radiance;
    float cdf_sum = 0, cdf_norm = 0, sky_w = 0;
    float sun_angles[2] = { 0, 0 };
    Vec3 sun_dir_world = { 0, 0, 1 };
    float half_aperture = 0;
    float spectral_pdf[10], spectral_cdf[10], spectral_integral = 0, spectral_norm = 0;
    int spectral_valid[2] = { 0, 0 };
    float area_ratio = 1;
    bool sun_below_horizon = false;
    bool active_record = false;
    float bs_center[3] = { 0, 0, 0 }, bs_radius = 1.f;
};

void validate_params(const ssk_params &p, bool spectral);
Vec3 normalize_sun_direction(const Vec3 &v);
/// Computes the hemisphere integral of the sky and the cone integral of the sun per channel (x Jacobian) for an
/// emitter whose block is complete except for the sampling weight; empty = quadrature_host
using QuadratureFn = std::function<void(const HostEmitter &E, float sky_rad[11], float sun_rad[11])>;
void build_emitter(HostEmitter &E, const ssk_params &p, bool from_constructor, const QuadratureFn &quad = {});
void quadrature_host(const HostEmitter &E, float sky_rad[11], float sun_rad[11]);
void build_emitter_pre(HostEmitter &E, const ssk_params &p, bool from_constructor);
void build_emitter_post(HostEmitter &E, const float sky_rad[11], const float sun_rad[11]);
void estimate_sky_sun_ratio(HostEmitter &E, const float sky_rad[11], const float sun_rad[11]);
std::shared_ptr<const Datasets> datasets_for(const std::string &dir, bool spectral);   // process-wide cache
const std::vector<float> &gauss_legendre_200();     // 200 nodes followed by 200 weights
void set_ray_pdf_scale(HotBlock &B);
void state_tangent(const HostEmitter &E, float d_turbidity, const float *d_albedo, int n_d_albedo, const float *d_sun_direction,
                   TangentBlock &T);

} // namespace ssk
