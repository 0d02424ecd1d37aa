// TEST INFRASTRUCTURE: type-checks adapter/sunsky_b200.cpp by instantiating it for the reference's CPU array
// types against the reference's real headers (Dr.Jit, mitsuba/core, mitsuba/render records) plus the oracle's
// stand-ins for the libmitsuba classes (oracle/ref_shadow). Compiled with -fsyntax-only by
// tests/test_host_cpu.py::test_adapter_plugin_type_checks when /root/reference is present.
#include <drjit/dynamic.h>
#include "../../adapter/sunsky_b200.cpp"

using FloatA = drjit::DynamicArray<float>;
template class mitsuba::SunskyB200Emitter<FloatA, mitsuba::Color<FloatA, 3>>;
template class mitsuba::SunskyB200Emitter<FloatA, mitsuba::Spectrum<FloatA, 4>>;
