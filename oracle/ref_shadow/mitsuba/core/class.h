// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/core/class.h>. The RTTI / plugin
// registration macros (reference: include/mitsuba/core/class.h:152-225) collapse to
// nothing; the reference harness instantiates SunskyEmitter directly.
#pragma once
#include <mitsuba/core/fwd.h>
NAMESPACE_BEGIN(mitsuba)
class Class { };
NAMESPACE_END(mitsuba)
#define MI_DECLARE_CLASS()
#define MI_IMPLEMENT_CLASS(Name, Parent, ...)
#define MI_IMPLEMENT_CLASS_VARIANT(Name, Parent, ...)
#define MI_EXPORT_PLUGIN(Name, Descr)
