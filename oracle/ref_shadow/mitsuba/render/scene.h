// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/render/scene.h>: only bbox() is used
// (reference: src/emitters/sunsky.cpp:287-301).
#pragma once
#include <mitsuba/render/fwd.h>
#include <mitsuba/core/bbox.h>
NAMESPACE_BEGIN(mitsuba)
template <typename Float, typename Spectrum> class Scene {
public:
    MI_IMPORT_CORE_TYPES()
    ScalarBoundingBox3f m_bbox;
    const ScalarBoundingBox3f &bbox() const { return m_bbox; }
};
NAMESPACE_END(mitsuba)
