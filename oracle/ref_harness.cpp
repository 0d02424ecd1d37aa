// TEST INFRASTRUCTURE — not product code. Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load the library built from this file.
//
// Reference harness: compiles the reference's own `src/emitters/sunsky.cpp` and
// `include/mitsuba/render/sunsky/sunsky.h` IN PLACE from /root/reference (nothing is copied into
// this repository) against
//   * the reference's real header-only math: drjit/*.h, mitsuba/core/{warp,frame,vector,transform,
//     quad,distr_1d,spectrum,bsphere,bbox,ray,math}.h
//   * the small stand-ins under oracle/ref_shadow/ for the parts of Mitsuba that live in libmitsuba
//     (Object, Properties, Texture, Emitter base, Scene, FileStream, logging)
// and exposes SunskyEmitter<Float, Spectrum> through a C ABI for ctypes.
//
// Two arithmetic flavours are instantiated:
//   flavour 0  Float = float                     == scalar_rgb / scalar_spectral
//   flavour 1  Float = dr::DynamicArray<float>   == the array code path llvm_ad_rgb / llvm_ad_spectral
//              take (array `erf` branch, FMA-chain dot/norm, quadrature for the sky/sun weight,
//              JIT DiscreteDistribution semantics, float->double promotions); per-lane IEEE ops are
//              the ones the LLVM backend emits (explicit fma, no contraction: -mfma -ffp-contract=off)
//
// Build: oracle/build_ref.sh  ->  oracle/_ref/libsunsky_ref.so   (git-ignored, travels with gpurun)

#include <cstdint>
#include <cstring>
#include <cstdio>
#include <string>
#include <vector>
#include <map>
#include <array>
#include <thread>
#include <sstream>
#include <filesystem>
#include <stdexcept>
#include <atomic>
#include <algorithm>
#include <memory>
#include <tuple>

#include <mitsuba/core/bsphere.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/core/distr_1d.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/scene.h>
#include <drjit/dynamic.h>
#include <drjit/tensor.h>
#include <drjit/sphere.h>
#include <drjit/random.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/fstream.h>

// ---- link-time symbols the real headers declare but libmitsuba would define ------------------
namespace mitsuba {
namespace string {
std::string indent(const std::string &s, size_t) { return s; }
std::string indent(const Object *, size_t) { return "<object>"; }
} // namespace string

namespace detail {
CIE1932Tables<float> color_space_tables_scalar;
} // namespace detail

// CIE 1931 tables: the numeric block is extracted from /root/reference/src/core/spectrum.cpp at
// build time by oracle/build_ref.sh into oracle/_ref/gen/cie1931_tbl.inc (never committed).
using Float = float;
#include "cie1931_tbl.inc"
#undef Float
static struct CieInit { CieInit() { detail::color_space_tables_scalar.initialize(cie1931_tbl); } } cie_init;
} // namespace mitsuba

// ---- shim: horizontal sums of array-of-array spectra -----------------------------------------
// `dr::sum_inner(Color<Float,3>)` yields width-1 *arrays* for JIT floats but plain scalars for
// dr::DynamicArray, which sunsky.cpp:815,849 (written against the JIT types) cannot consume.
// These overloads return width-1 arrays and add in the order the LLVM backend does
// (drjit-core src/llvm_ts.cpp:232-316 + llvm_red.h:78-116: sequential fp32 sums over 16384-wide
// chunks when size > 2*16384, then a sequential sum of the chunk partials).
namespace drjit {
inline float ssk_llvm_block_sum(const DynamicArray<float> &a) {
    const uint32_t size = (uint32_t) a.size(), ideal = 16384;
    uint32_t chunk = size;
    if (chunk > 2 * ideal) chunk = ideal;
    float total = 0.f;
    for (uint32_t start = 0; start < size; start += chunk) {
        uint32_t end = std::min(start + chunk, size);
        float accum = 0.f;
        for (uint32_t j = start; j < end; ++j) accum = accum + a[j];
        if (chunk == size) return accum;
        total = total + accum;
    }
    return total;
}
template <typename S> S ssk_sum_inner_llvm(const S &v) {
    S r;
    for (size_t i = 0; i < v.size(); ++i) r[i] = DynamicArray<float>(ssk_llvm_block_sum(v[i]));
    return r;
}
inline mitsuba::Color<DynamicArray<float>, 3> sum_inner(const mitsuba::Color<DynamicArray<float>, 3> &v) {
    return ssk_sum_inner_llvm(v);
}
inline mitsuba::Spectrum<DynamicArray<float>, 11> sum_inner(const mitsuba::Spectrum<DynamicArray<float>, 11> &v) {
    return ssk_sum_inner_llvm(v);
}
} // namespace drjit

// ---- the reference plugin itself, compiled where it lies -------------------------------------
#define private public
#define protected public
#include "src/emitters/sunsky.cpp"
#undef private
#undef protected

// ---- flavour 2: Float = dr::Packet<float, W>, the reference's query members on SIMD packets ----------------------
// Stand-in for the speed of llvm_ad_*: the same member bodies (eval, sample_direction, pdf_direction, sample_ray,
// sample_wavelengths: sunsky.cpp:303-480, 538-763) instantiated on AVX2 (W = 8) or AVX-512 (W = 16) packets, one
// instruction per 8 / 16 lanes as the LLVM backend's vectorised kernels issue them. The constructor and
// parameters_changed() are written for width-1 JIT arrays and do not instantiate on packets, so they are
// explicitly specialised here: the state comes from a flavour-1 emitter built from the same properties. Packet
// `rcp` is the hardware estimate plus one Newton step (drjit/packet_avx.h), so results sit within a few ulp of
// flavour 1 instead of matching bit for bit: flavour 2 is the TIMED baseline, flavour 1 stays the parity target.
#if defined(__AVX512F__)
#  define SSK_REF_PACKET_WIDTH 16
#else
#  define SSK_REF_PACKET_WIDTH 8
#endif
namespace mitsuba {
using PktF = drjit::Packet<float, SSK_REF_PACKET_WIDTH>;
using ArrF = drjit::DynamicArray<float>;

template <typename DstSpec, typename SrcSpec>
void ssk_packet_state_from(SunskyEmitter<PktF, DstSpec> &d, const SunskyEmitter<ArrF, SrcSpec> &s) {
    using Dst = SunskyEmitter<PktF, DstSpec>;
    d.m_bsphere = typename Dst::BoundingSphere3f(typename Dst::ScalarPoint3f(0.f), 1.f);
    d.m_turbidity = s.m_turbidity[0];
    d.m_sky_sampling_w = s.m_sky_sampling_w[0];
    d.m_sky_scale = s.m_sky_scale; d.m_sun_scale = s.m_sun_scale;
    d.m_sun_half_aperture = s.m_sun_half_aperture;
    d.m_active_record = s.m_active_record;
    for (int i = 0; i < 3; ++i) {
        d.m_sun_dir[i] = s.m_sun_dir[i][0];
        d.m_local_sun_frame.s[i] = s.m_local_sun_frame.s[i][0];
        d.m_local_sun_frame.t[i] = s.m_local_sun_frame.t[i][0];
        d.m_local_sun_frame.n[i] = s.m_local_sun_frame.n[i][0];
    }
    d.m_sun_angles[0] = s.m_sun_angles[0][0]; d.m_sun_angles[1] = s.m_sun_angles[1][0];
    d.m_sky_params = s.m_sky_params; d.m_sky_radiance = s.m_sky_radiance; d.m_sun_radiance = s.m_sun_radiance;
    d.m_gaussians = s.m_gaussians; d.m_sun_ld = s.m_sun_ld;
    d.m_gaussian_distr = DiscreteDistribution<PktF>(s.m_gaussian_distr.pmf());
    if constexpr (is_spectral_v<DstSpec>) {
        const Vector<float, 2> range = { WAVELENGTHS<float>[1], WAVELENGTHS<float>[NB_WAVELENGTHS - 1] };
        d.m_spectral_distr = ContinuousDistribution<typename Dst::Wavelength>(range, s.m_spectral_distr.pdf());
    }
    d.m_flags = s.m_flags;
}
#define SSK_PACKET_SPECIALISE(DstSpec, SrcSpec)                                                                    \
    template <> SunskyEmitter<PktF, DstSpec>::SunskyEmitter(const Properties &props) : Base(props) {              \
        SunskyEmitter<ArrF, SrcSpec> src(props);                                                                   \
        ssk_packet_state_from(*this, src);                                                                         \
    }                                                                                                              \
    template <> void SunskyEmitter<PktF, DstSpec>::parameters_changed(const std::vector<std::string> &) {          \
        Throw("flavour 2 (packets) is a timing baseline: build a new emitter instead of updating one");            \
    }                                                                                                              \
    template <> void SunskyEmitter<PktF, DstSpec>::traverse(TraversalCallback *) { }
#define SSK_COMMA ,
SSK_PACKET_SPECIALISE(Color<PktF SSK_COMMA 3>, Color<ArrF SSK_COMMA 3>)
SSK_PACKET_SPECIALISE(Spectrum<PktF SSK_COMMA 4>, Spectrum<ArrF SSK_COMMA 4>)
} // namespace mitsuba

namespace dr = drjit;
using namespace mitsuba;

// ---- C ABI -----------------------------------------------------------------------------------
extern "C" {

struct ssk_ref_params {
    float turbidity;
    int32_t n_albedo;          // 0 -> reference default (0.3), 1 -> uniform, 3 / 11 -> per channel
    float albedo[11];
    int32_t use_record;        // 0 -> sun_direction, 1 -> location/time record
    float sun_direction[3];
    float latitude, longitude, timezone;
    int32_t year, month, day;
    float hour, minute, second;
    float sun_scale, sky_scale;
    float sun_aperture_deg;    // <= 0 -> reference default
    int32_t has_to_world;
    float to_world[16];        // row-major
    int32_t record_defaults;   // 1 -> leave every record field at the reference default except `hour`
    int32_t has_to_local;      // 1 -> `to_local` is the inverse the Transform carries (a caller-built Transform4f,
    float to_local[16];        //      transform.h:59-62); 0 -> dr::inverse_transpose in fp64 as Properties does
};

struct ssk_ref_state {
    int32_t n_channels;
    float sky_params[99];
    float sky_radiance[11];
    float sun_radiance[3240];
    int32_t n_sun_radiance;
    float sun_ld[66];
    float gaussians[100];
    float pmf[20], cdf[20];
    float cdf_sum, cdf_normalization;
    float sky_sampling_w;
    float sun_angles[2];
    float frame_s[3], frame_t[3], frame_n[3];
    float sun_half_aperture;
    float sun_dir_world[3];
    float bsphere_center[3], bsphere_radius;
    float spectral_pdf[10], spectral_cdf[10];
    float spectral_integral, spectral_normalization;
    int32_t n_spectral;
    int32_t flags;
    float to_world[16], to_local[16];   // m_to_world.scalar().matrix and its carried inverse, row-major
};

} // extern "C"

namespace {

thread_local std::string g_err;

struct RefBase {
    virtual ~RefBase() { }
    virtual void update(const ssk_ref_params &p) = 0;
    virtual void state(ssk_ref_state &s) const = 0;
    virtual void set_scene(const float *bmin, const float *bmax) = 0;
    /// planes are `stride` floats apart (stride >= n): threads work in place on slices of the caller's arrays
    virtual void eval(size_t n, const float *wi, const float *wav, float *out, size_t stride = 0) const = 0;
    virtual void sample_direction(size_t n, const float *u, const float *itp, const float *wav, float *d,
                                  float *pdf, float *weight, float *p, float *dist) const = 0;
    virtual void pdf_direction(size_t n, const float *d, float *pdf) const = 0;
    virtual void sample_ray(size_t n, const float *wl_sample, const float *s2, const float *s3, float *o,
                            float *d, float *wav, float *weight) const = 0;
    virtual void sample_wavelengths(size_t n, const float *wi, const float *sample, float *wav, float *weight) const = 0;
    virtual std::vector<std::string> traverse_names() = 0;
    virtual std::string to_string() const = 0;
    virtual int channels() const = 0;
};

Properties make_props(const ssk_ref_params &p, bool spectral) {
    Properties props;
    props.set_float("turbidity", p.turbidity);
    props.set_float("sun_scale", p.sun_scale);
    props.set_float("sky_scale", p.sky_scale);
    if (p.sun_aperture_deg > 0.f)
        props.set_float("sun_aperture", p.sun_aperture_deg);
    if (p.n_albedo > 0)
        props.set_albedo(p.albedo, p.n_albedo);
    if (!p.use_record) {
        props.set_vec3("sun_direction", p.sun_direction[0], p.sun_direction[1], p.sun_direction[2]);
    } else if (p.record_defaults) {
        props.set_float("hour", p.hour);
    } else {
        props.set_float("latitude", p.latitude);
        props.set_float("longitude", p.longitude);
        props.set_float("timezone", p.timezone);
        props.set_float("year", p.year);
        props.set_float("month", p.month);
        props.set_float("day", p.day);
        props.set_float("hour", p.hour);
        props.set_float("minute", p.minute);
        props.set_float("second", p.second);
    }
    if (p.has_to_world)
        props.set_matrix("to_world", p.to_world, p.has_to_local ? p.to_local : nullptr);
    (void) spectral;
    return props;
}

template <typename Float> struct lanes {
    static constexpr bool IsArray = dr::is_array_v<Float>;
};

// helpers to move between SoA float buffers and Float
template <typename Float> Float load_f(const float *ptr, size_t off, size_t n) {
    if constexpr (dr::is_packed_array_v<Float> && !dr::is_dynamic_array_v<Float>) {
        if (n == Float::Size) return dr::load<Float>(ptr + off);
        Float r = dr::zeros<Float>();                       // tail: zero-padded lanes
        for (size_t i = 0; i < n; ++i) r[i] = ptr[off + i];
        return r;
    } else if constexpr (dr::is_array_v<Float>) return dr::load<Float>(ptr + off, n);
    else { (void) n; return ptr[off]; }
}
template <typename Float> void store_f(float *ptr, size_t off, size_t n, const Float &v) {
    if constexpr (dr::is_packed_array_v<Float> && !dr::is_dynamic_array_v<Float>) {
        if (n == Float::Size) dr::store(ptr + off, v);
        else for (size_t i = 0; i < n; ++i) ptr[off + i] = v[i];
    } else if constexpr (dr::is_array_v<Float>) {
        if (v.size() == n) memcpy(ptr + off, v.data(), n * sizeof(float));
        else if (v.size() == 1) for (size_t i = 0; i < n; ++i) ptr[off + i] = v[0];
        else throw std::runtime_error("store_f: unexpected array size");
    } else { (void) n; ptr[off] = v; }
}

template <typename Float, typename Spectrum> struct RefImpl final : RefBase {
    using Emitter_ = SunskyEmitter<Float, Spectrum>;
    MI_IMPORT_TYPES(Scene, Texture)
    static constexpr bool IsArray = dr::is_array_v<Float>;
    static constexpr bool IsPacket = dr::is_packed_array_v<Float> && !dr::is_dynamic_array_v<Float>;
    static constexpr bool IsRGB = is_rgb_v<Spectrum>;
    static constexpr size_t Chunk = IsPacket ? dr::size_v<Float> : (IsArray ? 16384 : 1);
    static constexpr int C = IsRGB ? 3 : 4;

    std::unique_ptr<Emitter_> em;
    mutable Scene scene;

    RefImpl(const ssk_ref_params &p) {
        Properties props = make_props(p, !IsRGB);
        em.reset(new Emitter_(props));
    }

    void update(const ssk_ref_params &p) override {
        // what mi.traverse(emitter)[...] = v; params.update() does (sunsky.cpp:220-285)
        em->m_turbidity = p.turbidity;
        em->m_sky_scale = p.sky_scale;
        em->m_sun_scale = p.sun_scale;
        if (p.n_albedo > 0)
            em->m_albedo = new Texture(std::vector<float>(p.albedo, p.albedo + p.n_albedo));
        if (p.has_to_world) {
            Properties tmp; tmp.set_matrix("to_world", p.to_world, p.has_to_local ? p.to_local : nullptr);
            em->m_to_world = tmp.get<ScalarTransform4f>("to_world");
        }
        if (em->m_active_record) {
            if (!p.use_record) throw std::runtime_error("emitter was built from a date/location record");
            em->m_location.latitude = p.latitude; em->m_location.longitude = p.longitude;
            em->m_location.timezone = p.timezone;
            em->m_time.year = p.year; em->m_time.month = p.month; em->m_time.day = p.day;
            em->m_time.hour = p.hour; em->m_time.minute = p.minute; em->m_time.second = p.second;
        } else {
            if (p.use_record) throw std::runtime_error("emitter was built from a sun direction");
            em->m_sun_dir = Vector3f(p.sun_direction[0], p.sun_direction[1], p.sun_direction[2]);
        }
        em->parameters_changed({});
    }

    static float sc(const Float &v) { if constexpr (IsArray) return v[0]; else return v; }

    void state(ssk_ref_state &s) const override {
        memset(&s, 0, sizeof(s));
        s.n_channels = IsRGB ? 3 : 11;
        for (size_t i = 0; i < em->m_sky_params.size(); ++i) s.sky_params[i] = em->m_sky_params[i];
        for (size_t i = 0; i < em->m_sky_radiance.size(); ++i) s.sky_radiance[i] = em->m_sky_radiance[i];
        s.n_sun_radiance = (int32_t) em->m_sun_radiance.size();
        for (size_t i = 0; i < em->m_sun_radiance.size(); ++i) s.sun_radiance[i] = em->m_sun_radiance[i];
        if constexpr (!IsRGB)
            for (size_t i = 0; i < em->m_sun_ld.size(); ++i) s.sun_ld[i] = em->m_sun_ld[i];
        for (size_t i = 0; i < em->m_gaussians.size(); ++i) s.gaussians[i] = em->m_gaussians[i];
        for (size_t i = 0; i < 20; ++i) {
            s.pmf[i] = em->m_gaussian_distr.pmf()[i];
            s.cdf[i] = em->m_gaussian_distr.cdf()[i];
        }
        s.cdf_sum = sc(Float(em->m_gaussian_distr.sum()));
        s.cdf_normalization = sc(Float(em->m_gaussian_distr.normalization()));
        s.sky_sampling_w = sc(em->m_sky_sampling_w);
        s.sun_angles[0] = sc(em->m_sun_angles.x()); s.sun_angles[1] = sc(em->m_sun_angles.y());
        for (int i = 0; i < 3; ++i) {
            s.frame_s[i] = sc(em->m_local_sun_frame.s[i]);
            s.frame_t[i] = sc(em->m_local_sun_frame.t[i]);
            s.frame_n[i] = sc(em->m_local_sun_frame.n[i]);
            s.sun_dir_world[i] = sc(em->m_sun_dir[i]);
            s.bsphere_center[i] = sc(em->m_bsphere.center[i]);
        }
        s.bsphere_radius = sc(em->m_bsphere.radius);
        s.sun_half_aperture = em->m_sun_half_aperture;
        s.flags = (int32_t) em->m_flags;
        {
            const auto &T = em->m_to_world.scalar();
            for (int r = 0; r < 4; ++r)
                for (int c = 0; c < 4; ++c) {
                    s.to_world[r * 4 + c] = T.matrix(r, c);
                    s.to_local[r * 4 + c] = T.inverse_transpose(c, r);
                }
        }
        if constexpr (!IsRGB) {
            const auto &sd = em->m_spectral_distr;
            s.n_spectral = (int32_t) sd.pdf().size();
            for (size_t i = 0; i < sd.pdf().size() && i < 10; ++i) s.spectral_pdf[i] = sd.pdf()[i];
            for (size_t i = 0; i < sd.cdf().size() && i < 10; ++i) s.spectral_cdf[i] = sd.cdf()[i];
            s.spectral_integral = sc(Float(sd.integral()));
            s.spectral_normalization = sc(Float(sd.normalization()));
        }
    }

    void set_scene(const float *bmin, const float *bmax) override {
        scene.m_bbox = ScalarBoundingBox3f();
        if (bmin && bmax)
            scene.m_bbox = ScalarBoundingBox3f(ScalarPoint3f(bmin[0], bmin[1], bmin[2]),
                                               ScalarPoint3f(bmax[0], bmax[1], bmax[2]));
        em->set_scene(&scene);
    }

    Wavelength load_wav(const float *wav, size_t n_total, size_t off, size_t n) const {
        Wavelength w = dr::zeros<Wavelength>();
        if constexpr (!IsRGB) {
            if (wav)
                for (int c = 0; c < 4; ++c)
                    w[c] = load_f<Float>(wav + c * n_total, off, n);
        }
        (void) wav; (void) n_total; (void) off; (void) n;
        return w;
    }

    void eval(size_t n, const float *wi, const float *wav, float *out, size_t stride) const override {
        const size_t S = stride ? stride : n;
        SurfaceInteraction3f si = dr::zeros<SurfaceInteraction3f>();
        for (size_t off = 0; off < n; off += Chunk) {
            size_t m = std::min(Chunk, n - off);
            si.wi = Vector3f(load_f<Float>(wi, off, m), load_f<Float>(wi + S, off, m), load_f<Float>(wi + 2 * S, off, m));
            si.wavelengths = load_wav(wav, S, off, m);
            Spectrum r = em->eval(si, Mask(true));
            for (int c = 0; c < C; ++c) store_f<Float>(out + c * S, off, m, r[c]);
        }
    }

    void sample_direction(size_t n, const float *u, const float *itp, const float *wav, float *d, float *pdf,
                          float *weight, float *p, float *dist) const override {
        for (size_t off = 0; off < n; off += Chunk) {
            size_t m = std::min(Chunk, n - off);
            Interaction3f it = dr::zeros<Interaction3f>();
            if (itp)
                it.p = Point3f(load_f<Float>(itp, off, m), load_f<Float>(itp + n, off, m), load_f<Float>(itp + 2 * n, off, m));
            it.wavelengths = load_wav(wav, n, off, m);
            Point2f uu(load_f<Float>(u, off, m), load_f<Float>(u + n, off, m));
            auto [ds, w] = em->sample_direction(it, uu, Mask(true));
            for (int c = 0; c < 3; ++c) store_f<Float>(d + c * n, off, m, ds.d[c]);
            store_f<Float>(pdf, off, m, ds.pdf);
            for (int c = 0; c < C; ++c) store_f<Float>(weight + c * n, off, m, w[c]);
            if (p) for (int c = 0; c < 3; ++c) store_f<Float>(p + c * n, off, m, ds.p[c]);
            if (dist) store_f<Float>(dist, off, m, ds.dist);
        }
    }

    void pdf_direction(size_t n, const float *d, float *pdf) const override {
        for (size_t off = 0; off < n; off += Chunk) {
            size_t m = std::min(Chunk, n - off);
            Interaction3f it = dr::zeros<Interaction3f>();
            DirectionSample3f ds = dr::zeros<DirectionSample3f>();
            ds.d = Vector3f(load_f<Float>(d, off, m), load_f<Float>(d + n, off, m), load_f<Float>(d + 2 * n, off, m));
            Float r = em->pdf_direction(it, ds, Mask(true));
            store_f<Float>(pdf, off, m, r);
        }
    }

    void sample_ray(size_t n, const float *wl_sample, const float *s2, const float *s3, float *o, float *d,
                    float *wav, float *weight) const override {
        for (size_t off = 0; off < n; off += Chunk) {
            size_t m = std::min(Chunk, n - off);
            Point2f a(load_f<Float>(s2, off, m), load_f<Float>(s2 + n, off, m)),
                    b(load_f<Float>(s3, off, m), load_f<Float>(s3 + n, off, m));
            auto [ray, w] = em->sample_ray(Float(0.f), load_f<Float>(wl_sample, off, m), a, b, Mask(true));
            for (int c = 0; c < 3; ++c) {
                store_f<Float>(o + c * n, off, m, ray.o[c]);
                store_f<Float>(d + c * n, off, m, ray.d[c]);
            }
            if constexpr (!IsRGB)
                for (int c = 0; c < 4; ++c) store_f<Float>(wav + c * n, off, m, ray.wavelengths[c]);
            for (int c = 0; c < C; ++c) store_f<Float>(weight + c * n, off, m, w[c]);
        }
    }

    void sample_wavelengths(size_t n, const float *wi, const float *sample, float *wav, float *weight) const override {
        for (size_t off = 0; off < n; off += Chunk) {
            size_t m = std::min(Chunk, n - off);
            SurfaceInteraction3f si = dr::zeros<SurfaceInteraction3f>();
            si.wi = Vector3f(load_f<Float>(wi, off, m), load_f<Float>(wi + n, off, m), load_f<Float>(wi + 2 * n, off, m));
            auto [wl, w] = em->sample_wavelengths(si, load_f<Float>(sample, off, m), Mask(true));
            if constexpr (!IsRGB)
                for (int c = 0; c < 4; ++c) store_f<Float>(wav + c * n, off, m, wl[c]);
            for (int c = 0; c < C; ++c) store_f<Float>(weight + c * n, off, m, w[c]);
        }
    }

    std::vector<std::string> traverse_names() override {
        TraversalCallback cb;
        em->traverse(&cb);
        return cb.names;
    }
    std::string to_string() const override { return em->to_string(); }
    int channels() const override { return C; }
};

using DynF = dr::DynamicArray<float>;

template <typename F> int guarded(F &&f) {
    try { f(); return 0; }
    catch (const std::exception &e) { g_err = e.what(); return 1; }
    catch (...) { g_err = "unknown exception"; return 1; }
}

// static partition of [0, n) over `threads` std::threads; each thread works on contiguous SoA slices
template <typename F> void parallel_ranges(size_t n, int threads, F &&f) {
    if (threads <= 1 || n < 2) { f(0, n); return; }
    std::vector<std::thread> pool;
    size_t per = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        size_t b = std::min(n, per * t), e = std::min(n, b + per);
        if (b < e) pool.emplace_back([=, &f] { f(b, e); });
    }
    for (auto &t : pool) t.join();
}

} // namespace

extern "C" {

const char *ssk_ref_last_error() { return g_err.c_str(); }

void ssk_ref_set_resource_root(const char *root) { FileResolver::root() = root ? root : ""; }

void *ssk_ref_create(int flavour, int spectral, const ssk_ref_params *p) {
    RefBase *r = nullptr;
    int rc = guarded([&] {
        if (flavour == 0 && !spectral) r = new RefImpl<float, Color<float, 3>>(*p);
        else if (flavour == 0 && spectral) r = new RefImpl<float, Spectrum<float, 4>>(*p);
        else if (flavour == 1 && !spectral) r = new RefImpl<DynF, Color<DynF, 3>>(*p);
        else if (flavour == 1 && spectral) r = new RefImpl<DynF, Spectrum<DynF, 4>>(*p);
        else if (flavour == 2 && !spectral) r = new RefImpl<PktF, Color<PktF, 3>>(*p);
        else if (flavour == 2 && spectral) r = new RefImpl<PktF, Spectrum<PktF, 4>>(*p);
        else throw std::runtime_error("bad flavour");
    });
    return rc ? nullptr : r;
}
void ssk_ref_destroy(void *h) { delete (RefBase *) h; }
int ssk_ref_update(void *h, const ssk_ref_params *p) { return guarded([&] { ((RefBase *) h)->update(*p); }); }
int ssk_ref_get_state(void *h, ssk_ref_state *s) { return guarded([&] { ((RefBase *) h)->state(*s); }); }
int ssk_ref_set_scene(void *h, const float *bmin, const float *bmax) {
    return guarded([&] { ((RefBase *) h)->set_scene(bmin, bmax); });
}
int ssk_ref_channels(void *h) { return ((RefBase *) h)->channels(); }
int ssk_ref_packet_width() { return SSK_REF_PACKET_WIDTH; }
int ssk_ref_warn_count() { return mitsuba::detail::oracle_warn_count(); }

// All buffers are SoA: component c of element i lives at ptr[c * n + i].
int ssk_ref_eval(void *h, size_t n, const float *wi, const float *wav, float *out, int threads) {
    RefBase *r = (RefBase *) h;
    int C = r->channels();
    return guarded([&] {
        if (threads <= 1) { r->eval(n, wi, wav, out); return; }
        // static partition; every thread reads and writes its slice of the caller's planes in place
        (void) C;
        parallel_ranges(n, threads, [&](size_t b, size_t e) {
            r->eval(e - b, wi + b, wav ? wav + b : nullptr, out + b, n);
        });
    });
}

int ssk_ref_sample_direction(void *h, size_t n, const float *u, const float *itp, const float *wav, float *d,
                             float *pdf, float *weight, float *p, float *dist, int threads) {
    RefBase *r = (RefBase *) h;
    int C = r->channels();
    return guarded([&] {
        if (threads <= 1) { r->sample_direction(n, u, itp, wav, d, pdf, weight, p, dist); return; }
        parallel_ranges(n, threads, [&](size_t b, size_t e) {
            size_t m = e - b;
            std::vector<float> uu(2 * m), ip(itp ? 3 * m : 0), wv(wav ? 4 * m : 0), dd(3 * m), pp(m),
                ww((size_t) C * m), po(p ? 3 * m : 0), di(dist ? m : 0);
            for (int c = 0; c < 2; ++c) memcpy(&uu[c * m], u + c * n + b, m * 4);
            if (itp) for (int c = 0; c < 3; ++c) memcpy(&ip[c * m], itp + c * n + b, m * 4);
            if (wav) for (int c = 0; c < 4; ++c) memcpy(&wv[c * m], wav + c * n + b, m * 4);
            r->sample_direction(m, uu.data(), itp ? ip.data() : nullptr, wav ? wv.data() : nullptr, dd.data(),
                                pp.data(), ww.data(), p ? po.data() : nullptr, dist ? di.data() : nullptr);
            for (int c = 0; c < 3; ++c) memcpy(d + c * n + b, &dd[c * m], m * 4);
            memcpy(pdf + b, pp.data(), m * 4);
            for (int c = 0; c < C; ++c) memcpy(weight + c * n + b, &ww[c * m], m * 4);
            if (p) for (int c = 0; c < 3; ++c) memcpy(p + c * n + b, &po[c * m], m * 4);
            if (dist) memcpy(dist + b, di.data(), m * 4);
        });
    });
}

int ssk_ref_pdf_direction(void *h, size_t n, const float *d, float *pdf, int threads) {
    RefBase *r = (RefBase *) h;
    return guarded([&] {
        if (threads <= 1) { r->pdf_direction(n, d, pdf); return; }
        parallel_ranges(n, threads, [&](size_t b, size_t e) {
            size_t m = e - b;
            std::vector<float> dd(3 * m);
            for (int c = 0; c < 3; ++c) memcpy(&dd[c * m], d + c * n + b, m * 4);
            r->pdf_direction(m, dd.data(), pdf + b);
        });
    });
}

int ssk_ref_sample_ray(void *h, size_t n, const float *wl_sample, const float *s2, const float *s3, float *o,
                       float *d, float *wav, float *weight) {
    return guarded([&] { ((RefBase *) h)->sample_ray(n, wl_sample, s2, s3, o, d, wav, weight); });
}

int ssk_ref_sample_wavelengths(void *h, size_t n, const float *wi, const float *sample, float *wav, float *weight) {
    return guarded([&] { ((RefBase *) h)->sample_wavelengths(n, wi, sample, wav, weight); });
}

int ssk_ref_traverse_names(void *h, char *buf, size_t len) {
    return guarded([&] {
        std::string s;
        for (auto &n : ((RefBase *) h)->traverse_names()) { s += n; s += ";"; }
        snprintf(buf, len, "%s", s.c_str());
    });
}

int ssk_ref_to_string(void *h, char *buf, size_t len) {
    return guarded([&] { snprintf(buf, len, "%s", ((RefBase *) h)->to_string().c_str()); });
}

// ---- direct access to the Dr.Jit math the path is built from (unit tests of the restatements) ----
//   flavour 0 evaluates on `float`, flavour 1 on dr::DynamicArray<float> (array branch)
enum { SSK_M_SIN = 0, SSK_M_COS, SSK_M_ASIN, SSK_M_ACOS, SSK_M_ATAN2, SSK_M_EXP, SSK_M_EXP2, SSK_M_LOG,
       SSK_M_LOG2, SSK_M_CBRT, SSK_M_ERF, SSK_M_ERFINV, SSK_M_POW15, SSK_M_TAN, SSK_M_SQRT, SSK_M_RCP };

int ssk_ref_math(int op, int flavour, size_t n, const float *a, const float *b, float *out) {
    return guarded([&] {
        auto apply = [&](auto x, auto y) {
            using T = decltype(x);
            switch (op) {
                case SSK_M_SIN: return T(dr::sin(x));
                case SSK_M_COS: return T(dr::cos(x));
                case SSK_M_ASIN: return T(dr::asin(x));
                case SSK_M_ACOS: return T(dr::acos(x));
                case SSK_M_ATAN2: return T(dr::atan2(x, y));
                case SSK_M_EXP: return T(dr::exp(x));
                case SSK_M_EXP2: return T(dr::exp2(x));
                case SSK_M_LOG: return T(dr::log(x));
                case SSK_M_LOG2: return T(dr::log2(x));
                case SSK_M_CBRT: return T(dr::cbrt(x));
                case SSK_M_ERF: return T(dr::erf(x));
                case SSK_M_ERFINV: return T(dr::erfinv(x));
                case SSK_M_POW15: return T(dr::pow(x, 1.5));
                case SSK_M_TAN: return T(dr::tan(x));
                case SSK_M_SQRT: return T(dr::sqrt(x));
                case SSK_M_RCP: return T(dr::rcp(x));
                default: throw std::runtime_error("bad math op");
            }
        };
        if (flavour == 0) {
            for (size_t i = 0; i < n; ++i) out[i] = apply(a[i], b ? b[i] : 0.f);
        } else {
            DynF x = dr::load<DynF>(a, n), y = b ? dr::load<DynF>(b, n) : dr::zeros<DynF>(n);
            DynF r = apply(x, y);
            store_f<DynF>(out, 0, n, r);
        }
    });
}

// Blanco-Muriel sun position, exactly as the plugin calls it (sunsky.h:283-374)
int ssk_ref_sun_coordinates(int flavour, int year, int month, int day, float hour, float minute, float second,
                            float latitude, float longitude, float timezone, float *out3) {
    return guarded([&] {
        auto run = [&](auto tag) {
            using F = decltype(tag);
            DateTimeRecord<F> t; LocationRecord<F> l;
            t.year = year; t.month = month; t.day = day; t.hour = hour; t.minute = minute; t.second = second;
            l.latitude = latitude; l.longitude = longitude; l.timezone = timezone;
            auto v = compute_sun_coordinates<F>(t, l);
            for (int i = 0; i < 3; ++i) {
                if constexpr (dr::is_array_v<F>) out3[i] = v[i][0]; else out3[i] = v[i];
            }
        };
        if (flavour == 0) run(float()); else run(DynF(0.f));
    });
}

// PCG32 streams as sampling_test.py / the survey's cfg 2-3 use them: PCG32(size = n) draws one
// float per stream j with initstate + j, initseq + j (drjit/random.h:44-65, 131-133)
int ssk_ref_pcg32_floats(size_t n, uint64_t stream_offset, float *out) {
    return guarded([&] {
        for (size_t j = 0; j < n; ++j) {
            dr::PCG32<uint32_t> rng(1, PCG32_DEFAULT_STATE + stream_offset + j, PCG32_DEFAULT_STREAM + stream_offset + j);
            out[j] = rng.next_float32();
        }
    });
}

// quad::gauss_legendre<DynF>(n) (quad.h:27-86)
int ssk_ref_gauss_legendre(int n, float *nodes, float *weights) {
    return guarded([&] {
        auto [x, w] = quad::gauss_legendre<DynF>(n);
        for (int i = 0; i < n; ++i) { nodes[i] = x[i]; weights[i] = w[i]; }
    });
}

// dr::normalize on a ScalarVector3f, as init_from_props applies it to the sun_direction property (sunsky.cpp:922)
void ssk_ref_normalize3(const float *in3, float *out3) {
    using ScalarVector3f = Vector<float, 3>;
    ScalarVector3f v = dr::normalize(ScalarVector3f(in3[0], in3[1], in3[2]));
    for (int i = 0; i < 3; ++i) out3[i] = v[i];
}
// Transform<Point<double, 4>>(m) -> ScalarTransform4f: the inverse a to_world property ends up with, row-major
void ssk_ref_invert4(const float *m16, float *inv16) {
    Properties tmp; tmp.set_matrix("m", m16);
    auto T = tmp.get<Transform<Point<float, 4>>>("m");
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) inv16[r * 4 + c] = T.inverse_transpose(c, r);
}

double ssk_ref_area_ratio_numerator() { return 1 - dr::cos(SUN_HALF_APERTURE); }
float ssk_ref_area_ratio(float half_aperture) { return get_area_ratio(half_aperture); }

} // extern "C"
