"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref/libsunsky_ref.so (the reference's own
sunsky.cpp compiled in place by oracle/build_ref.sh). Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libsunsky_ref.so")               # AVX2 + FMA: the parity oracle
LIB_PATH_AVX512 = os.path.join(_HERE, "_ref", "libsunsky_ref_avx512.so")  # 16-wide packets for the timed baseline
REPO = os.path.dirname(_HERE)
# dataset root: a directory that contains resources/sunsky/datasets/*.bin
DEFAULT_ROOTS = ["/root/reference", os.path.join(REPO, "mitsuba3-sunsky_b200", "data")]


class RefParams(C.Structure):
    _fields_ = [
        ("turbidity", C.c_float), ("n_albedo", C.c_int32), ("albedo", C.c_float * 11),
        ("use_record", C.c_int32), ("sun_direction", C.c_float * 3),
        ("latitude", C.c_float), ("longitude", C.c_float), ("timezone", C.c_float),
        ("year", C.c_int32), ("month", C.c_int32), ("day", C.c_int32),
        ("hour", C.c_float), ("minute", C.c_float), ("second", C.c_float),
        ("sun_scale", C.c_float), ("sky_scale", C.c_float), ("sun_aperture_deg", C.c_float),
        ("has_to_world", C.c_int32), ("to_world", C.c_float * 16), ("record_defaults", C.c_int32),
        ("has_to_local", C.c_int32), ("to_local", C.c_float * 16),
    ]


class RefState(C.Structure):
    _fields_ = [
        ("n_channels", C.c_int32), ("sky_params", C.c_float * 99), ("sky_radiance", C.c_float * 11),
        ("sun_radiance", C.c_float * 3240), ("n_sun_radiance", C.c_int32), ("sun_ld", C.c_float * 66),
        ("gaussians", C.c_float * 100), ("pmf", C.c_float * 20), ("cdf", C.c_float * 20),
        ("cdf_sum", C.c_float), ("cdf_normalization", C.c_float), ("sky_sampling_w", C.c_float),
        ("sun_angles", C.c_float * 2), ("frame_s", C.c_float * 3), ("frame_t", C.c_float * 3),
        ("frame_n", C.c_float * 3), ("sun_half_aperture", C.c_float), ("sun_dir_world", C.c_float * 3),
        ("bsphere_center", C.c_float * 3), ("bsphere_radius", C.c_float),
        ("spectral_pdf", C.c_float * 10), ("spectral_cdf", C.c_float * 10),
        ("spectral_integral", C.c_float), ("spectral_normalization", C.c_float),
        ("n_spectral", C.c_int32), ("flags", C.c_int32),
        ("to_world", C.c_float * 16), ("to_local", C.c_float * 16),
    ]


_lib = None
_libs = {}
FP = C.POINTER(C.c_float)


def host_has_avx512():
    try:
        flags = open("/proc/cpuinfo").read()
    except OSError:
        return False
    return all((" " + f) in flags for f in ("avx512f", "avx512vl", "avx512dq", "avx512bw", "avx512cd"))


def best_isa():
    """ISA build for the timed CPU baseline: AVX-512 when the host has it and the library was built."""
    return "avx512" if host_has_avx512() and os.path.exists(LIB_PATH_AVX512) else "avx2"


def available():
    return os.path.exists(LIB_PATH)


def lib(isa="avx2"):
    """isa "avx2": the parity oracle (every golden vector comes from it). "avx512": same harness, 16-wide packets,
    for timing only. "best": best_isa()."""
    global _lib
    if isa == "best":
        isa = best_isa()
    if isa != "avx2":
        if isa not in _libs:
            _libs[isa] = _load(LIB_PATH_AVX512)
        return _libs[isa]
    if _lib is None:
        _lib = _load(LIB_PATH)
    return _lib


def _load(path):
    if True:
        L = C.CDLL(path)
        L.ssk_ref_last_error.restype = C.c_char_p
        L.ssk_ref_create.restype = C.c_void_p
        L.ssk_ref_create.argtypes = [C.c_int, C.c_int, C.POINTER(RefParams)]
        L.ssk_ref_destroy.argtypes = [C.c_void_p]
        L.ssk_ref_update.argtypes = [C.c_void_p, C.POINTER(RefParams)]
        L.ssk_ref_get_state.argtypes = [C.c_void_p, C.POINTER(RefState)]
        L.ssk_ref_set_scene.argtypes = [C.c_void_p, FP, FP]
        L.ssk_ref_channels.argtypes = [C.c_void_p]
        L.ssk_ref_eval.argtypes = [C.c_void_p, C.c_size_t, FP, FP, FP, C.c_int]
        L.ssk_ref_sample_direction.argtypes = [C.c_void_p, C.c_size_t, FP, FP, FP, FP, FP, FP, FP, FP, C.c_int]
        L.ssk_ref_pdf_direction.argtypes = [C.c_void_p, C.c_size_t, FP, FP, C.c_int]
        L.ssk_ref_sample_ray.argtypes = [C.c_void_p, C.c_size_t, FP, FP, FP, FP, FP, FP, FP]
        L.ssk_ref_sample_wavelengths.argtypes = [C.c_void_p, C.c_size_t, FP, FP, FP, FP]
        L.ssk_ref_traverse_names.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        L.ssk_ref_to_string.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        L.ssk_ref_math.argtypes = [C.c_int, C.c_int, C.c_size_t, FP, FP, FP]
        L.ssk_ref_sun_coordinates.argtypes = [C.c_int] + [C.c_int] * 3 + [C.c_float] * 6 + [FP]
        L.ssk_ref_pcg32_floats.argtypes = [C.c_size_t, C.c_uint64, FP]
        L.ssk_ref_gauss_legendre.argtypes = [C.c_int, FP, FP]
        L.ssk_ref_area_ratio_numerator.restype = C.c_double
        L.ssk_ref_area_ratio.restype = C.c_float
        L.ssk_ref_area_ratio.argtypes = [C.c_float]
        L.ssk_ref_set_resource_root.argtypes = [C.c_char_p]
        L.ssk_ref_normalize3.argtypes = [FP, FP]
        L.ssk_ref_invert4.argtypes = [FP, FP]
        for r in DEFAULT_ROOTS:
            if os.path.exists(os.path.join(r, "resources", "sunsky", "datasets", "tgmm_tables.bin")):
                L.ssk_ref_set_resource_root(r.encode())
                break
        L.ssk_ref_packet_width.restype = C.c_int
    return L


def _fp(a):
    return None if a is None else a.ctypes.data_as(FP)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


MATH_OPS = dict(sin=0, cos=1, asin=2, acos=3, atan2=4, exp=5, exp2=6, log=7, log2=8, cbrt=9, erf=10,
                erfinv=11, pow15=12, tan=13, sqrt=14, rcp=15)


def math(op, x, y=None, flavour=1):
    x = _f32(x); y = _f32(y)
    out = np.empty_like(x)
    rc = lib().ssk_ref_math(MATH_OPS[op], flavour, x.size, _fp(x), _fp(y), _fp(out))
    if rc:
        raise RuntimeError(lib().ssk_ref_last_error().decode())
    return out


def make_params(turbidity=3.0, albedo=None, sun_direction=None, record=None, hour=None, sun_scale=1.0,
                sky_scale=1.0, sun_aperture=None, to_world=None, to_local=None):
    p = RefParams()
    p.turbidity = turbidity
    if albedo is None:
        p.n_albedo = 0
    else:
        a = np.atleast_1d(np.asarray(albedo, dtype=np.float32))
        p.n_albedo = a.size
        for i, v in enumerate(a):
            p.albedo[i] = v
    if sun_direction is not None:
        p.use_record = 0
        for i in range(3):
            p.sun_direction[i] = sun_direction[i]
    elif record is not None:
        p.use_record = 1
        (p.latitude, p.longitude, p.timezone, p.year, p.month, p.day, p.hour, p.minute, p.second) = record
    else:
        p.use_record = 1
        p.record_defaults = 1
        p.latitude, p.longitude, p.timezone = 35.6894, 139.6917, 9
        p.year, p.month, p.day = 2010, 7, 10
        p.hour = 15.0 if hour is None else hour
    p.sun_scale, p.sky_scale = sun_scale, sky_scale
    p.sun_aperture_deg = -1.0 if sun_aperture is None else sun_aperture
    if to_world is not None:
        p.has_to_world = 1
        m = np.asarray(to_world, dtype=np.float32).reshape(16)
        for i in range(16):
            p.to_world[i] = m[i]
    if to_local is not None:
        p.has_to_local = 1
        m = np.asarray(to_local, dtype=np.float32).reshape(16)
        for i in range(16):
            p.to_local[i] = m[i]
    return p


class RefEmitter:
    """SunskyEmitter of the reference. flavour 0 = scalar_* arithmetic, 1 = array arithmetic (llvm_ad_*-like; the
    parity target), 2 = the query members on SIMD packets (timed baseline; within a few ulp of flavour 1)."""

    def __init__(self, spectral=False, flavour=1, isa="avx2", **kw):
        self.L = lib(isa)
        self.spectral = spectral
        self.params = make_params(**kw)
        self.h = self.L.ssk_ref_create(flavour, int(spectral), C.byref(self.params))
        if not self.h:
            raise RuntimeError(self.L.ssk_ref_last_error().decode())
        self.C = self.L.ssk_ref_channels(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ssk_ref_destroy(self.h)
            self.h = None

    def _chk(self, rc):
        if rc:
            raise RuntimeError(self.L.ssk_ref_last_error().decode())

    def update(self, **kw):
        self.params = make_params(**kw)
        self._chk(self.L.ssk_ref_update(self.h, C.byref(self.params)))

    def state(self):
        s = RefState()
        self._chk(self.L.ssk_ref_get_state(self.h, C.byref(s)))
        return s

    def set_scene(self, bmin=None, bmax=None):
        a = _f32(bmin); b = _f32(bmax)
        self._chk(self.L.ssk_ref_set_scene(self.h, _fp(a), _fp(b)))

    def eval(self, wi, wavelengths=None, threads=1):
        """wi: (3, n) SoA; wavelengths: (4, n) for spectral. Returns (C, n)."""
        wi = _f32(wi); wav = _f32(wavelengths)
        n = wi.shape[1]
        out = np.empty((self.C, n), dtype=np.float32)
        self._chk(self.L.ssk_ref_eval(self.h, n, _fp(wi), _fp(wav), _fp(out), threads))
        return out

    def sample_direction(self, u, it_p=None, wavelengths=None, threads=1, want_p=False):
        u = _f32(u); itp = _f32(it_p); wav = _f32(wavelengths)
        n = u.shape[1]
        d = np.empty((3, n), np.float32); pdf = np.empty(n, np.float32)
        w = np.empty((self.C, n), np.float32)
        p = np.empty((3, n), np.float32) if want_p else None
        dist = np.empty(n, np.float32) if want_p else None
        self._chk(self.L.ssk_ref_sample_direction(self.h, n, _fp(u), _fp(itp), _fp(wav), _fp(d), _fp(pdf),
                                                  _fp(w), _fp(p), _fp(dist), threads))
        return (d, pdf, w, p, dist) if want_p else (d, pdf, w)

    def pdf_direction(self, d, threads=1):
        d = _f32(d)
        n = d.shape[1]
        pdf = np.empty(n, np.float32)
        self._chk(self.L.ssk_ref_pdf_direction(self.h, n, _fp(d), _fp(pdf), threads))
        return pdf

    def sample_ray(self, wl_sample, s2, s3):
        wl = _f32(wl_sample); s2 = _f32(s2); s3 = _f32(s3)
        n = wl.size
        o = np.empty((3, n), np.float32); d = np.empty((3, n), np.float32)
        wav = np.zeros((4, n), np.float32); w = np.empty((self.C, n), np.float32)
        self._chk(self.L.ssk_ref_sample_ray(self.h, n, _fp(wl), _fp(s2), _fp(s3), _fp(o), _fp(d), _fp(wav), _fp(w)))
        return o, d, wav, w

    def sample_wavelengths(self, wi, sample):
        wi = _f32(wi); s = _f32(sample)
        n = s.size
        wav = np.zeros((4, n), np.float32); w = np.empty((self.C, n), np.float32)
        self._chk(self.L.ssk_ref_sample_wavelengths(self.h, n, _fp(wi), _fp(s), _fp(wav), _fp(w)))
        return wav, w

    def traverse_names(self):
        buf = C.create_string_buffer(1024)
        self._chk(self.L.ssk_ref_traverse_names(self.h, buf, 1024))
        return [s for s in buf.value.decode().split(";") if s]

    def to_string(self):
        buf = C.create_string_buffer(4096)
        self._chk(self.L.ssk_ref_to_string(self.h, buf, 4096))
        return buf.value.decode()


def sun_coordinates(year, month, day, hour, minute, second, latitude, longitude, timezone, flavour=1):
    out = np.empty(3, np.float32)
    rc = lib().ssk_ref_sun_coordinates(flavour, year, month, day, hour, minute, second, latitude, longitude,
                                       timezone, _fp(out))
    if rc:
        raise RuntimeError(lib().ssk_ref_last_error().decode())
    return out


def normalize3(v):
    """dr::normalize(ScalarVector3f) as init_from_props applies it to `sun_direction` (sunsky.cpp:922)."""
    a = _f32(v); out = np.empty(3, np.float32)
    lib().ssk_ref_normalize3(_fp(a), _fp(out))
    return out


def invert4(m):
    """The inverse a `to_world` property carries once read back as ScalarTransform4f (fp64 inverse, cast)."""
    a = _f32(np.asarray(m, np.float32).reshape(16)); out = np.empty(16, np.float32)
    lib().ssk_ref_invert4(_fp(a), _fp(out))
    return out.reshape(4, 4)


_hosek = None


def hosek_sun_rad(turbidity, wavelength, elevation, gamma):
    """mi.hosek_sun_rad of the reference: the original fp64 Hosek-Wilkie model compiled in place
    (oracle/hosek_harness.cpp -> oracle/_ref/libhosek_ref.so). The second, independent oracle of SURVEY.md 8(c)."""
    global _hosek
    if _hosek is None:
        H = C.CDLL(os.path.join(_HERE, "_ref", "libhosek_ref.so"))
        H.ssk_ref_hosek_sun_rad.restype = C.c_double
        H.ssk_ref_hosek_sun_rad.argtypes = [C.c_double] * 4
        _hosek = H
    return _hosek.ssk_ref_hosek_sun_rad(turbidity, wavelength, elevation, gamma)


def hosek_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libhosek_ref.so"))


def pcg32_floats(n, stream_offset=0):
    out = np.empty(n, np.float32)
    lib().ssk_ref_pcg32_floats(n, stream_offset, _fp(out))
    return out


def gauss_legendre(n):
    x = np.empty(n, np.float32); w = np.empty(n, np.float32)
    lib().ssk_ref_gauss_legendre(n, _fp(x), _fp(w))
    return x, w
