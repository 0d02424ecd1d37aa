// TEST INFRASTRUCTURE: type-checks adapter/sunsky_b200.cpp by instantiating it for the reference's CPU array
// types and for its CUDA JIT array type (dr::CUDAArray<float> = the Float of cuda_rgb / cuda_spectral) against the reference's real headers (Dr.Jit, mitsuba/core, mitsuba/render records) plus the oracle's
// stand-ins for the libmitsuba classes (oracle/ref_shadow). Compiled with -fsyntax-only by
// tests/test_host_cpu.py::test_adapter_plugin_type_checks when /root/reference is present.
#include <drjit/jit.h>
#include <drjit/dynamic.h>
#include "../../adapter/sunsky_b200.cpp"

using FloatA = drjit::DynamicArray<float>;
template class mitsuba::SunskyB200Emitter<FloatA, mitsuba::Color<FloatA, 3>>;
template class mitsuba::SunskyB200Emitter<FloatA, mitsuba::Spectrum<FloatA, 4>>;

// the cuda_* branch (jit_cuda_stream / jit_cuda_device_raw, *_device entry points on Dr.Jit's stream): the JIT array
// type itself, against the reference's drjit/jit.h and drjit-core/jit.h declarations
using FloatC = drjit::CUDAArray<float>;
template class mitsuba::SunskyB200Emitter<FloatC, mitsuba::Color<FloatC, 3>>;
template class mitsuba::SunskyB200Emitter<FloatC, mitsuba::Spectrum<FloatC, 4>>;
