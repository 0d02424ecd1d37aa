// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/render/emitter.h> / endpoint.h: the virtual
// Emitter surface the sunsky plugin overrides (reference: include/mitsuba/render/endpoint.h:99-373,
// include/mitsuba/render/emitter.h:52-97), without anything that lives in libmitsuba.
#pragma once
#include <mitsuba/core/object.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/core/field.h>
#include <mitsuba/core/transform.h>
#include <mitsuba/core/bbox.h>
#include <mitsuba/render/interaction.h>
#include <mitsuba/render/texture.h>
#include <mitsuba/render/scene.h>

NAMESPACE_BEGIN(mitsuba)

enum class EmitterFlags : uint32_t {
    Empty = 0x0, DeltaPosition = 0x1, DeltaDirection = 0x2, Infinite = 0x4, Surface = 0x8,
    SpatiallyVarying = 0x10
};
constexpr uint32_t operator+(EmitterFlags f) { return (uint32_t) f; }

enum class ProfilerPhase : int { EndpointEvaluate, EndpointSampleRay, EndpointSampleDirection };
#undef MI_MASKED_FUNCTION
#define MI_MASKED_FUNCTION(phase, mask) (void) mask;

template <typename Float, typename Spectrum> class Emitter : public Object {
public:
    MI_IMPORT_TYPES(Scene, Texture)

    Emitter(const Properties &props) {
        m_to_world = props.get<ScalarTransform4f>("to_world", ScalarTransform4f());
    }

    virtual void traverse(TraversalCallback *) { }
    virtual void parameters_changed(const std::vector<std::string> & = {}) { }
    virtual void set_scene(const Scene *) { }
    virtual Spectrum eval(const SurfaceInteraction3f &, Mask) const = 0;
    virtual std::pair<Ray3f, Spectrum> sample_ray(Float, Float, const Point2f &, const Point2f &, Mask) const = 0;
    virtual std::pair<DirectionSample3f, Spectrum> sample_direction(const Interaction3f &, const Point2f &, Mask) const = 0;
    virtual Float pdf_direction(const Interaction3f &, const DirectionSample3f &, Mask) const = 0;
    virtual Spectrum eval_direction(const Interaction3f &, const DirectionSample3f &, Mask) const = 0;
    virtual std::pair<Wavelength, Spectrum> sample_wavelengths(const SurfaceInteraction3f &, Float, Mask) const = 0;
    virtual std::pair<PositionSample3f, Float> sample_position(Float, const Point2f &, Mask) const = 0;
    virtual ScalarBoundingBox3f bbox() const = 0;
    virtual std::string to_string() const = 0;

    uint32_t m_flags = 0;
    field<Transform4f, ScalarTransform4f> m_to_world;
};

NAMESPACE_END(mitsuba)
