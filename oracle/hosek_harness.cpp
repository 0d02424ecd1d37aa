// TEST INFRASTRUCTURE -- second, independent fp64 oracle: the original Hosek-Wilkie sky model as the reference ships
// it (include/mitsuba/render/sunsky/ArHosekSkyModel.c, compiled IN PLACE from /root/reference with the reference's
// own Dr.Jit math for the three functions it was edited to call: dr::sin, dr::sqrt, dr::deg_to_rad, :758-763).
// The reference exposes one function of it to Python, mi.hosek_sun_rad (src/render/python/sunsky_v.cpp:19); this
// harness exports the same function for ctypes. Build: oracle/build_ref.sh -> oracle/_ref/libhosek_ref.so
#include <drjit/math.h>
#include <drjit/sphere.h>
namespace dr = drjit;
#include "include/mitsuba/render/sunsky/ArHosekSkyModel.c"

extern "C" double ssk_ref_hosek_sun_rad(double turbidity, double wavelength, double elevation, double gamma) {
    return arhosekskymodel_solar_radiance_internal2(turbidity, wavelength, elevation, gamma);
}
