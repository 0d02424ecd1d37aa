// Mitsuba 3 adapter plugin for libsunsky_b200.so -- the reference-side binding of INTEGRATION.md section 3.
//
// Goes into the mitsuba3-sunsky tree as src/emitters/sunsky_b200.cpp, next to the plugin it stands in for
// (src/emitters/sunsky.cpp), registered with `add_plugin(sunsky_b200 sunsky_b200.cpp)` after
// src/emitters/CMakeLists.txt:11 and linked with `target_link_libraries(sunsky_b200 PRIVATE sunsky_b200)`.
// It is meant for the cuda_* variants: a Dr.Jit CUDA array already is one device plane per component, so every
// Emitter call maps 1:1 onto one C-ABI entry point of include/sunsky_b200.h, enqueued on Dr.Jit's own stream.
// The same source also instantiates for CPU array types whose `.data()` is host memory (it then uses the *_host
// entry points); that is what tests/test_host_cpu.py::test_adapter_plugin_type_checks type-checks against the reference's headers.
//
// Member-by-member correspondence (reference: src/emitters/sunsky.cpp):
//   constructor / init_from_props  :162-218, :889-948   -> ssk_params + ssk_create
//   traverse / parameters_changed  :220-285             -> same parameter names, ssk_update
//   set_scene                      :287-301             -> ssk_set_scene
//   eval                           :303-352             -> ssk_eval_*
//   sample_ray                     :354-397             -> ssk_sample_ray_*
//   sample_direction               :399-441             -> ssk_sample_direction_*
//   pdf_direction / eval_direction :443-461             -> ssk_pdf_direction_* / ssk_eval_direction_*
//   sample_wavelengths             :463-480             -> ssk_sample_wavelengths_*
//   sample_position / bbox / to_string :482-519         -> as the reference
#include <mitsuba/core/bsphere.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/texture.h>

#include <sunsky_b200.h>

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum>
class SunskyB200Emitter final : public Emitter<Float, Spectrum> {
public:
    MI_IMPORT_BASE(Emitter, m_flags, m_to_world)
    MI_IMPORT_TYPES(Scene, Texture)

    static constexpr bool IsCUDA = dr::is_cuda_v<Float>;
    static constexpr size_t Channels = is_spectral_v<Spectrum> ? 4 : 3;

    SunskyB200Emitter(const Properties &props) : Base(props) {
        if constexpr (!(is_rgb_v<Spectrum> || is_spectral_v<Spectrum>))
            Throw("Unsupported spectrum type, can only render in Spectral or RGB modes!");   // sunsky.cpp:164-165
        ssk_params_default(&m_params);
        m_params.arith = props.get<bool>("fast_arithmetic", false) ? SSK_ARITH_FAST : SSK_ARITH_EXACT;
        read_properties(props);
        m_albedo = props.texture<Texture>("albedo", 0.3f);                                  // :908-910
        if (m_albedo->is_spatially_varying())
            Log(Error, "Expected a non-spatially varying radiance spectra!");
        fill_albedo();
        fill_to_world();

        FileResolver *fr = Thread::thread()->file_resolver();
        std::string dir = fr->resolve("resources/sunsky/datasets").string();                 // sunsky.h:65
        int device = -1;
        if constexpr (IsCUDA)
            device = jit_cuda_device_raw();
        if (ssk_create(&m_params, dir.c_str(), device, is_spectral_v<Spectrum> ? SSK_MODE_SPECTRAL : SSK_MODE_RGB, &m_ctx))
            Throw("%s", ssk_last_error());
        m_flags = +EmitterFlags::Infinite | +EmitterFlags::SpatiallyVarying;                // :217
    }

    ~SunskyB200Emitter() { ssk_destroy(m_ctx); }

    // ---- properties (names, defaults and the sun_direction XOR date/location rule of :889-948) ---------
    void read_properties(const Properties &props) {
        m_params.sun_scale = props.get<ScalarFloat>("sun_scale", 1.f);
        m_params.sky_scale = props.get<ScalarFloat>("sky_scale", 1.f);
        m_params.turbidity = props.get<ScalarFloat>("turbidity", 3.f);
        m_params.sun_aperture_deg = props.get<ScalarFloat>("sun_aperture", 0.5358f);
        if (props.has_property("sun_direction")) {
            for (const char *k : { "latitude", "longitude", "timezone", "year", "month", "day", "hour", "minute", "second" })
                if (props.has_property(k))
                    Log(Error, "Both the 'sun_direction' and parameters for time/location "
                               "were provided, both information cannot be given at the same time!");
            ScalarVector3f d = props.get<ScalarVector3f>("sun_direction");
            m_params.use_record = 0;
            m_sun_direction = d;
            for (int i = 0; i < 3; ++i) m_params.sun_direction[i] = d[i];
        } else {
            m_params.use_record = 1;
            m_params.latitude  = props.get<ScalarFloat>("latitude", 35.6894f);
            m_params.longitude = props.get<ScalarFloat>("longitude", 139.6917f);
            m_params.timezone  = props.get<ScalarFloat>("timezone", 9);
            m_params.year      = props.get<ScalarInt32>("year", 2010);
            m_params.month     = props.get<ScalarUInt32>("month", 7);
            m_params.day       = props.get<ScalarUInt32>("day", 10);
            m_params.hour      = props.get<ScalarFloat>("hour", 15.0f);
            m_params.minute    = props.get<ScalarFloat>("minute", 0.0f);
            m_params.second    = props.get<ScalarFloat>("second", 0.0f);
        }
    }

    /// extract_albedo (:956-978): 3 RGB values or the spectrum at the 11 dataset wavelengths 320..720 nm
    void fill_albedo() {
        SurfaceInteraction3f si = dr::zeros<SurfaceInteraction3f>();
        if constexpr (is_rgb_v<Spectrum>) {
            auto a = m_albedo->eval(si);
            m_params.n_albedo = 3;
            for (int c = 0; c < 3; ++c) m_params.albedo[c] = dr::slice(a[c], 0);
        } else {
            m_params.n_albedo = 11;
            for (int c = 0; c < 11; ++c) {
                si.wavelengths = Float(320.f + 40.f * c);
                m_params.albedo[c] = dr::slice(m_albedo->eval(si)[0], 0);
            }
        }
    }

    void fill_to_world() {
        // a Transform carries matrix AND inverse_transpose (transform.h:44-45); the reference's per-lane
        // world -> local transform is m_to_world.value().inverse() = transpose(inverse_transpose) (:77-79),
        // so hand the library exactly that inverse instead of letting it invert the matrix again
        const auto &T = m_to_world.scalar();
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                m_params.to_world[r * 4 + c] = T.matrix(r, c);
                m_params.to_local[r * 4 + c] = T.inverse_transpose(c, r);
            }
        m_params.flags |= SSK_FLAG_TO_LOCAL_GIVEN;
    }

    void traverse(TraversalCallback *cb) override {                                         // :220-240
        Base::traverse(cb);
        cb->put_parameter("turbidity", m_params.turbidity, +ParamFlags::Differentiable);
        cb->put_parameter("sky_scale", m_params.sky_scale, +ParamFlags::NonDifferentiable);
        cb->put_parameter("sun_scale", m_params.sun_scale, +ParamFlags::NonDifferentiable);
        cb->put_object("albedo", m_albedo.get(), +ParamFlags::Differentiable);
        if (m_params.use_record) {
            cb->put_parameter("latitude", m_params.latitude, +ParamFlags::NonDifferentiable);
            cb->put_parameter("longitude", m_params.longitude, +ParamFlags::NonDifferentiable);
            cb->put_parameter("timezone", m_params.timezone, +ParamFlags::NonDifferentiable);
            cb->put_parameter("year", m_params.year, +ParamFlags::NonDifferentiable);
            cb->put_parameter("day", m_params.day, +ParamFlags::NonDifferentiable);
            cb->put_parameter("month", m_params.month, +ParamFlags::NonDifferentiable);
            cb->put_parameter("hour", m_params.hour, +ParamFlags::NonDifferentiable);
            cb->put_parameter("minute", m_params.minute, +ParamFlags::NonDifferentiable);
            cb->put_parameter("second", m_params.second, +ParamFlags::NonDifferentiable);
        } else {
            cb->put_parameter("sun_direction", m_sun_direction, +ParamFlags::Differentiable);
        }
        cb->put_parameter("to_world", *m_to_world.ptr(), +ParamFlags::NonDifferentiable);
    }

    void parameters_changed(const std::vector<std::string> &keys) override {                // :242-285
        if (!m_params.use_record)
            for (int i = 0; i < 3; ++i) m_params.sun_direction[i] = m_sun_direction[i];
        fill_albedo();
        fill_to_world();
        if (ssk_update(m_ctx, &m_params))
            Throw("%s", ssk_last_error());
        Base::parameters_changed(keys);
    }

    void set_scene(const Scene *scene) override {                                           // :287-301
        ScalarBoundingBox3f b = scene->bbox();
        if (b.valid()) {
            float lo[3] = { (float) b.min[0], (float) b.min[1], (float) b.min[2] },
                  hi[3] = { (float) b.max[0], (float) b.max[1], (float) b.max[2] };
            ssk_set_scene(m_ctx, lo, hi);
        } else {
            ssk_set_scene(m_ctx, nullptr, nullptr);
        }
    }

    // ---- plane plumbing ----------------------------------------------------------------------------------
    // A Dr.Jit array of width n is one plane of n floats (device memory for CUDAArray, host memory for CPU array
    // types). Inputs are evaluated (and broadcast to width n if they are literals) before their pointer is taken.
    static Float plane(const Float &v, size_t n) {
        Float r = dr::width(v) == n ? v : v + dr::zeros<Float>(n);
        dr::eval(r);
        return r;
    }
    static void *stream() {
        if constexpr (IsCUDA) return (void *) jit_cuda_stream();
        else return nullptr;
    }
    void check(int rc) const { if (rc) Throw("%s", ssk_last_error()); }

    Spectrum eval_planes(const Vector3f &dir, const Wavelength &wavelengths, bool direction, size_t n) const {
        Float x = plane(dir.x(), n), y = plane(dir.y(), n), z = plane(dir.z(), n), w[4];
        const float *in[3] = { x.data(), y.data(), z.data() }, *wav[4] = { nullptr, nullptr, nullptr, nullptr };
        if constexpr (is_spectral_v<Spectrum>)
            for (size_t k = 0; k < 4; ++k) { w[k] = plane(wavelengths[k], n); wav[k] = w[k].data(); }
        Spectrum result;
        float *out[4] = { nullptr, nullptr, nullptr, nullptr };
        for (size_t c = 0; c < Channels; ++c) { result[c] = dr::empty<Float>(n); out[c] = result[c].data(); }
        if constexpr (IsCUDA)
            check((direction ? ssk_eval_direction_device : ssk_eval_device)(m_ctx, in, wav, n, out, stream()));
        else
            check((direction ? ssk_eval_direction_host : ssk_eval_host)(m_ctx, in, wav, n, out));
        return result;
    }

    // ---- Emitter interface -------------------------------------------------------------------------------
    Spectrum eval(const SurfaceInteraction3f &si, Mask active) const override {             // :303-352
        MI_MASKED_FUNCTION(ProfilerPhase::EndpointEvaluate, active);
        size_t n = dr::maximum(dr::width(si.wi), dr::width(active));
        return eval_planes(si.wi, si.wavelengths, false, n) & active;
    }

    std::pair<Ray3f, Spectrum> sample_ray(Float time, Float wavelength_sample, const Point2f &sample2,
                                          const Point2f &sample3, Mask active) const override {   // :354-397
        MI_MASKED_FUNCTION(ProfilerPhase::EndpointSampleRay, active);
        size_t n = dr::maximum(dr::maximum(dr::width(sample2), dr::width(sample3)), dr::width(wavelength_sample));
        Float wl = plane(wavelength_sample, n), a0 = plane(sample2.x(), n), a1 = plane(sample2.y(), n),
              b0 = plane(sample3.x(), n), b1 = plane(sample3.y(), n);
        const float *s2[2] = { a0.data(), a1.data() }, *s3[2] = { b0.data(), b1.data() };
        Point3f o; Vector3f d; Wavelength wavelengths = dr::zeros<Wavelength>(); Spectrum weight;
        float *po[3], *pd[3], *pw[4] = { nullptr, nullptr, nullptr, nullptr }, *pq[4] = { nullptr, nullptr, nullptr, nullptr };
        for (size_t k = 0; k < 3; ++k) { o[k] = dr::empty<Float>(n); d[k] = dr::empty<Float>(n); po[k] = o[k].data(); pd[k] = d[k].data(); }
        for (size_t c = 0; c < Channels; ++c) { weight[c] = dr::empty<Float>(n); pq[c] = weight[c].data(); }
        if constexpr (is_spectral_v<Spectrum>)
            for (size_t k = 0; k < 4; ++k) { wavelengths[k] = dr::empty<Float>(n); pw[k] = wavelengths[k].data(); }
        if constexpr (IsCUDA)
            check(ssk_sample_ray_device(m_ctx, is_spectral_v<Spectrum> ? wl.data() : nullptr, s2, s3, n, po, pd, pw, pq, stream()));
        else
            check(ssk_sample_ray_host(m_ctx, is_spectral_v<Spectrum> ? wl.data() : nullptr, s2, s3, n, po, pd, pw, pq));
        return { Ray3f(o, d, time, wavelengths), weight & active };
    }

    std::pair<DirectionSample3f, Spectrum> sample_direction(const Interaction3f &it, const Point2f &sample,
                                                            Mask active) const override {   // :399-441
        MI_MASKED_FUNCTION(ProfilerPhase::EndpointSampleDirection, active);
        size_t n = dr::maximum(dr::width(sample), dr::width(it.p));
        Float u0 = plane(sample.x(), n), u1 = plane(sample.y(), n), p0 = plane(it.p.x(), n), p1 = plane(it.p.y(), n),
              p2 = plane(it.p.z(), n), w[4];
        const float *u[2] = { u0.data(), u1.data() }, *itp[3] = { p0.data(), p1.data(), p2.data() },
                    *wav[4] = { nullptr, nullptr, nullptr, nullptr };
        if constexpr (is_spectral_v<Spectrum>)
            for (size_t k = 0; k < 4; ++k) { w[k] = plane(it.wavelengths[k], n); wav[k] = w[k].data(); }
        DirectionSample3f ds = dr::zeros<DirectionSample3f>();
        Spectrum weight;
        float *pd[3], *pp[3], *pq[4] = { nullptr, nullptr, nullptr, nullptr };
        for (size_t k = 0; k < 3; ++k) { ds.d[k] = dr::empty<Float>(n); ds.p[k] = dr::empty<Float>(n); pd[k] = ds.d[k].data(); pp[k] = ds.p[k].data(); }
        for (size_t c = 0; c < Channels; ++c) { weight[c] = dr::empty<Float>(n); pq[c] = weight[c].data(); }
        ds.pdf = dr::empty<Float>(n); ds.dist = dr::empty<Float>(n);
        if constexpr (IsCUDA)
            check(ssk_sample_direction_device(m_ctx, u, itp, wav, n, pd, ds.pdf.data(), pq, pp, ds.dist.data(), stream()));
        else
            check(ssk_sample_direction_host(m_ctx, u, itp, wav, n, pd, ds.pdf.data(), pq, pp, ds.dist.data()));
        ds.n = -ds.d; ds.uv = sample; ds.time = it.time; ds.delta = false;                  // :416-428
        ds.emitter = this;
        return { ds, weight & active };
    }

    Float pdf_direction(const Interaction3f &, const DirectionSample3f &ds, Mask active) const override {   // :443-451
        MI_MASKED_FUNCTION(ProfilerPhase::EndpointEvaluate, active);
        size_t n = dr::width(ds.d);
        Float x = plane(ds.d.x(), n), y = plane(ds.d.y(), n), z = plane(ds.d.z(), n), pdf = dr::empty<Float>(n);
        const float *d[3] = { x.data(), y.data(), z.data() };
        if constexpr (IsCUDA) check(ssk_pdf_direction_device(m_ctx, d, n, pdf.data(), stream()));
        else check(ssk_pdf_direction_host(m_ctx, d, n, pdf.data()));
        return dr::select(active, pdf, 0.f);
    }

    Spectrum eval_direction(const Interaction3f &it, const DirectionSample3f &ds, Mask active) const override {   // :453-461
        MI_MASKED_FUNCTION(ProfilerPhase::EndpointEvaluate, active);
        return eval_planes(ds.d, it.wavelengths, true, dr::width(ds.d)) & active;
    }

    std::pair<Wavelength, Spectrum> sample_wavelengths(const SurfaceInteraction3f &si, Float sample,
                                                       Mask active) const override {        // :463-480
        size_t n = dr::maximum(dr::width(si.wi), dr::width(sample));
        if constexpr (!is_spectral_v<Spectrum>) {
            return { Wavelength(0.f), eval_planes(si.wi, si.wavelengths, false, n) & active };
        } else {
            Float x = plane(si.wi.x(), n), y = plane(si.wi.y(), n), z = plane(si.wi.z(), n), s = plane(sample, n);
            const float *wi[3] = { x.data(), y.data(), z.data() };
            Wavelength wavelengths; Spectrum weight;
            float *pw[4], *pq[4];
            for (size_t k = 0; k < 4; ++k) {
                wavelengths[k] = dr::empty<Float>(n); weight[k] = dr::empty<Float>(n);
                pw[k] = wavelengths[k].data(); pq[k] = weight[k].data();
            }
            if constexpr (IsCUDA) check(ssk_sample_wavelengths_device(m_ctx, wi, s.data(), n, pw, pq, stream()));
            else check(ssk_sample_wavelengths_host(m_ctx, wi, s.data(), n, pw, pq));
            return { wavelengths, weight & active };
        }
    }

    std::pair<PositionSample3f, Float> sample_position(Float, const Point2f &, Mask) const override {   // :482-495
        if constexpr (dr::is_jit_v<Float>)
            return { dr::zeros<PositionSample3f>(), dr::NaN<Float> };
        else
            NotImplementedError("sample_position");
    }

    ScalarBoundingBox3f bbox() const override { return ScalarBoundingBox3f(); }             // :497-500

    std::string to_string() const override {                                                // :502-519
        char buf[4096];
        ssk_to_string(m_ctx, buf, sizeof(buf));
        return buf;
    }

    MI_DECLARE_CLASS()
private:
    ssk_ctx *m_ctx = nullptr;
    ssk_params m_params;
    ref<Texture> m_albedo;
    ScalarVector3f m_sun_direction = ScalarVector3f(0.f, 0.f, 1.f);
};

MI_IMPLEMENT_CLASS_VARIANT(SunskyB200Emitter, Emitter)
MI_EXPORT_PLUGIN(SunskyB200Emitter, "B200 sun and sky emitter")
NAMESPACE_END(mitsuba)
