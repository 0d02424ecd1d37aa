// TEST INFRASTRUCTURE (oracle/): stand-ins for the interaction / sample records, reduced to the
// fields the sunsky emitter reads or writes (reference: include/mitsuba/render/interaction.h:157,
// 528-530; include/mitsuba/render/records.h:88,200).
#pragma once
#include <mitsuba/render/fwd.h>
#include <mitsuba/core/vector.h>
#include <mitsuba/core/spectrum.h>
#include <mitsuba/core/ray.h>
#include <drjit/array_traverse.h>

NAMESPACE_BEGIN(mitsuba)

template <typename Float_, typename Spectrum_> struct Interaction {
    using Float = Float_; using Spectrum = Spectrum_;
    MI_IMPORT_RENDER_BASIC_TYPES()
    Float t, time;
    Wavelength wavelengths;
    Point3f p;
    Normal3f n;
    DRJIT_STRUCT(Interaction, t, time, wavelengths, p, n)
};

template <typename Float_, typename Spectrum_> struct SurfaceInteraction : Interaction<Float_, Spectrum_> {
    using Float = Float_; using Spectrum = Spectrum_;
    MI_IMPORT_RENDER_BASIC_TYPES()
    using Base = Interaction<Float_, Spectrum_>;
    using Base::t; using Base::time; using Base::wavelengths; using Base::p; using Base::n;
    Point2f uv;
    Vector3f wi;
    DRJIT_STRUCT(SurfaceInteraction, t, time, wavelengths, p, n, uv, wi)
};

template <typename Float_, typename Spectrum_> struct PositionSample {
    using Float = Float_; using Spectrum = Spectrum_;
    MI_IMPORT_RENDER_BASIC_TYPES()
    Point3f p; Normal3f n; Point2f uv; Float time, pdf; Mask delta;
    DRJIT_STRUCT(PositionSample, p, n, uv, time, pdf, delta)
};

template <typename Float_, typename Spectrum_> struct DirectionSample : PositionSample<Float_, Spectrum_> {
    using Float = Float_; using Spectrum = Spectrum_;
    MI_IMPORT_RENDER_BASIC_TYPES()
    using Base = PositionSample<Float_, Spectrum_>;
    using Base::p; using Base::n; using Base::uv; using Base::time; using Base::pdf; using Base::delta;
    using EmitterPtr = typename RenderAliases::EmitterPtr;
    Vector3f d; Float dist; EmitterPtr emitter;
    DRJIT_STRUCT(DirectionSample, p, n, uv, time, pdf, delta, d, dist, emitter)
};

NAMESPACE_END(mitsuba)
