"""Host-side mirror of the reference's `sunsky` emitter plugin on top of the C ABI
(include/sunsky_b200.h -> libsunsky_b200.so).

`SunskyEmitter` keeps the plugin's surface (reference: src/emitters/sunsky.cpp):
properties `turbidity, albedo, sun_direction | latitude..second, sun_scale, sky_scale, sun_aperture,
to_world` (:19-104), `eval` (:303), `sample_direction` (:399), `pdf_direction` (:443),
`eval_direction` (:453), `traverse`/`parameters_changed` (:220-285, exposed as `traverse()` /
`update()`), `set_scene` (:287) and `to_string` (:502).

Array conventions: planar float32, shape (components, n). numpy arrays go through the `*_host`
entry points (H2D + kernel + D2H inside the call); torch CUDA tensors go through the `*_device`
entry points on torch's current stream and return torch tensors. Nothing here computes radiance on
the CPU: without the built library and a CUDA device the compute calls raise.
"""
import ctypes as C
import numpy as np

from . import _capi
from ._capi import (MODE_RGB, MODE_SPECTRAL, FLAG_SUN_DIR_NORMALIZED, FLAG_TO_LOCAL_GIVEN, ARITH_EXACT, ARITH_FAST, DATASET_DIR,
                    Params, State)

__all__ = ["SunskyEmitter", "SunskyError", "sun_coordinates", "hosek_sun_rad", "array_from_file", "array_to_file",
           "launch_count", "MODE_RGB", "MODE_SPECTRAL"]


class SunskyError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"[ssk {code}] {msg}")
        self.code = code


def _check(rc):
    if rc != 0:
        raise SunskyError(rc, _capi.lib().ssk_last_error().decode())


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _planes_np(a, k, name):
    a = np.asarray(a, dtype=np.float32)
    if a.ndim != 2 or a.shape[0] != k:
        raise ValueError(f"{name}: expected shape ({k}, n), got {a.shape}")
    return [np.ascontiguousarray(a[i]) for i in range(k)]


def _pp(planes):
    arr = (_capi.FP * len(planes))()
    for i, p in enumerate(planes):
        arr[i] = p.ctypes.data_as(_capi.FP)
    return arr


def _pp_t(t, k, name, device=None):
    import torch
    if t.dtype != torch.float32 or t.dim() != 2 or t.shape[0] != k:
        raise ValueError(f"{name}: expected float32 tensor of shape ({k}, n), got {tuple(t.shape)} {t.dtype}")
    if not t.is_cuda or (device is not None and t.device.index != device):
        # a host pointer handed to a kernel is a sticky illegal-address fault: refuse it here
        raise ValueError(f"{name}: expected a CUDA tensor on device {device}, got {t.device}")
    if t.stride(1) != 1:
        t = t.contiguous()
    arr = (_capi.FP * k)()
    for i in range(k):
        arr[i] = C.cast(t[i].data_ptr(), _capi.FP)
    return arr, t


_PARAM_KEYS = ("turbidity", "albedo", "sun_direction", "latitude", "longitude", "timezone", "year", "month",
               "day", "hour", "minute", "second", "sun_scale", "sky_scale", "sun_aperture", "to_world", "to_local")


class SunskyEmitter:
    """Sun and sky dome background emitter (Hosek-Wilkie sky + sun, TGMM importance sampling)."""

    def __init__(self, spectral=False, device=0, dataset_dir=None, sun_direction_normalized=False, arith="exact",
                 **props):
        """arith: "exact" replays the reference's llvm_ad_* op sequence (bit-identical results); "fast" keeps
        the geometry and the sampled directions on that sequence and moves the per-channel / per-Gaussian
        transcendentals to the MUFU unit (radiance / pdf / weight within 1e-5 relative)."""
        if arith not in ("exact", "fast"):
            raise ValueError("arith must be 'exact' or 'fast'")
        self.arith = arith
        unknown = set(props) - set(_PARAM_KEYS)
        if unknown:
            raise TypeError(f"unknown sunsky properties: {sorted(unknown)}")
        record_keys = {"latitude", "longitude", "timezone", "year", "month", "day", "hour", "minute", "second"}
        if "sun_direction" in props and record_keys & set(props):
            # same rule as init_from_props (sunsky.cpp:912-920)
            raise SunskyError(_capi.SSK_ERR_INVALID,
                              "Both the 'sun_direction' and parameters for time/location were provided, "
                              "both information cannot be given at the same time!")
        self._L = _capi.lib()
        self.spectral = bool(spectral)
        self.device = device
        self._normalized = bool(sun_direction_normalized)
        self._props = {}
        p = Params()
        self._L.ssk_params_default(C.byref(p))
        self._fill(p, props, creating=True)
        self._params = p
        ctx = _capi.CTX()
        _check(self._L.ssk_create(C.byref(p), (dataset_dir or DATASET_DIR).encode(), int(device),
                                  MODE_SPECTRAL if spectral else MODE_RGB, C.byref(ctx)))
        self._ctx = ctx
        self.channels = self._L.ssk_channels(ctx)
        if not p.use_record:
            # the constructor normalises sun_direction (sunsky.cpp:922); parameters_changed() then works on the
            # stored m_sun_dir as is, so later updates must start from the normalised vector, not the raw property
            sd = self.state().sun_dir_world
            for i in range(3):
                self._params.sun_direction[i] = sd[i]

    # -- property surface ---------------------------------------------------------------------
    def _fill(self, p, props, creating):
        for k, v in props.items():
            if k == "albedo" and isinstance(v, (int, float)):
                p.n_albedo = 1
                p.albedo[0] = float(v)
            elif k == "albedo":
                a = np.atleast_1d(np.asarray(v, dtype=np.float32)).ravel()
                if a.size not in (1, 3, 11):
                    raise SunskyError(_capi.SSK_ERR_INVALID, "albedo needs 1, 3 (RGB) or 11 (spectral) values")
                p.n_albedo = a.size
                for i, x in enumerate(a):
                    p.albedo[i] = float(x)
            elif k == "sun_direction":
                p.use_record = 0
                for i in range(3):
                    p.sun_direction[i] = float(v[i])
            elif k == "sun_aperture":
                p.sun_aperture_deg = float(v)
            elif k in ("to_world", "to_local"):
                # to_local: the inverse the caller's Transform4f carries (transform.h:44-45); without it the
                # library inverts to_world the way a `to_world` property is inverted (fp64, rounded)
                m = np.asarray(v, dtype=np.float32).reshape(16)
                dst = p.to_world if k == "to_world" else p.to_local
                for i in range(16):
                    dst[i] = float(m[i])
            elif k in ("year", "month", "day"):
                setattr(p, k, int(v))
            else:
                setattr(p, k, float(v))
        if creating and "sun_direction" not in props:
            p.use_record = 1
        self._props.update(props)
        if "to_world" in props and "to_local" not in props:
            self._props.pop("to_local", None)          # a new to_world without its inverse: back to the computed one
        p.flags = (FLAG_SUN_DIR_NORMALIZED if self._normalized else 0) | \
                  (FLAG_TO_LOCAL_GIVEN if "to_local" in self._props else 0)
        p.arith = ARITH_FAST if self.arith == "fast" else ARITH_EXACT

    def traverse(self):
        """Names and current values of the parameters `traverse()` exposes (sunsky.cpp:220-240)."""
        p = self._params
        out = {"turbidity": p.turbidity, "sky_scale": p.sky_scale, "sun_scale": p.sun_scale,
               "albedo": [p.albedo[i] for i in range(p.n_albedo)]}
        if p.use_record:
            out.update(latitude=p.latitude, longitude=p.longitude, timezone=p.timezone, year=p.year, day=p.day,
                       month=p.month, hour=p.hour, minute=p.minute, second=p.second)
        else:
            out["sun_direction"] = [p.sun_direction[i] for i in range(3)]
        out["to_world"] = np.array(p.to_world, dtype=np.float32).reshape(4, 4)
        return out

    def update(self, **props):
        """`params[...] = v; params.update()` -> parameters_changed() (sunsky.cpp:242-285)."""
        unknown = set(props) - set(_PARAM_KEYS) | ({"sun_aperture"} & set(props))
        if unknown:
            raise TypeError(f"not a traversable sunsky parameter: {sorted(unknown)}")
        record_keys = {"latitude", "longitude", "timezone", "year", "month", "day", "hour", "minute", "second"}
        if (self._params.use_record and "sun_direction" in props) or \
                (not self._params.use_record and record_keys & set(props)):
            # traverse() only exposes the parametrisation the emitter was created with (sunsky.cpp:226-238)
            raise SunskyError(_capi.SSK_ERR_INVALID, "an emitter keeps the sun parametrisation (direction or "
                              "date/location) it was created with")
        p = Params.from_buffer_copy(self._params)
        self._fill(p, props, creating=False)
        _check(self._L.ssk_update(self._ctx, C.byref(p)))
        self._params = p

    parameters_changed = update

    def set_scene(self, bbox_min=None, bbox_max=None):
        a = None if bbox_min is None else np.ascontiguousarray(bbox_min, dtype=np.float32)
        b = None if bbox_max is None else np.ascontiguousarray(bbox_max, dtype=np.float32)
        _check(self._L.ssk_set_scene(self._ctx, None if a is None else a.ctypes.data_as(_capi.FP),
                                     None if b is None else b.ctypes.data_as(_capi.FP)))

    def state(self):
        s = State()
        _check(self._L.ssk_get_state(self._ctx, C.byref(s)))
        return s

    def to_string(self):
        buf = C.create_string_buffer(2048)
        _check(self._L.ssk_to_string(self._ctx, buf, 2048))
        return buf.value.decode()

    __repr__ = to_string

    def close(self):
        if getattr(self, "_ctx", None):
            self._L.ssk_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- Emitter interface --------------------------------------------------------------------
    def _stream(self):
        """torch's current stream ON THE CONTEXT'S DEVICE (not on whatever device happens to be current)."""
        import torch
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _pp_t(self, t, k, name):
        return _pp_t(t, k, name, self.device)

    def _eval_impl(self, dirs, wavelengths, direction):
        Cn = self.channels
        if _is_torch(dirs):
            import torch
            d_pp, dirs = self._pp_t(dirs, 3, "directions")
            n = dirs.shape[1]
            w_pp = None
            if self.spectral:
                if wavelengths is None:
                    raise ValueError("spectral mode needs wavelengths of shape (4, n)")
                w_pp, wavelengths = self._pp_t(wavelengths, 4, "wavelengths")
            out = torch.empty((Cn, n), dtype=torch.float32, device=dirs.device)
            o_pp, _ = self._pp_t(out, Cn, "out")
            fn = self._L.ssk_eval_direction_device if direction else self._L.ssk_eval_device
            _check(fn(self._ctx, d_pp, w_pp, n, o_pp, self._stream()))
            return out
        planes = _planes_np(dirs, 3, "directions")
        n = planes[0].size
        wp = None
        if self.spectral:
            if wavelengths is None:
                raise ValueError("spectral mode needs wavelengths of shape (4, n)")
            wpl = _planes_np(wavelengths, 4, "wavelengths")
            wp = _pp(wpl)
        out = np.empty((Cn, n), dtype=np.float32)
        fn = self._L.ssk_eval_direction_host if direction else self._L.ssk_eval_host
        _check(fn(self._ctx, _pp(planes), wp, n, _pp([out[c] for c in range(Cn)])))
        return out

    def eval(self, wi, wavelengths=None):
        """Radiance arriving along -wi. wi: (3, n); wavelengths: (4, n) in nm (spectral). -> (C, n)"""
        return self._eval_impl(wi, wavelengths, False)

    def eval_direction(self, d, wavelengths=None):
        """eval() for a direction sample: wi = -d."""
        return self._eval_impl(d, wavelengths, True)

    def pdf_direction(self, d):
        """Solid-angle density of sample_direction() at world direction d: (3, n) -> (n,)"""
        if _is_torch(d):
            import torch
            d_pp, d = self._pp_t(d, 3, "d")
            n = d.shape[1]
            out = torch.empty(n, dtype=torch.float32, device=d.device)
            _check(self._L.ssk_pdf_direction_device(self._ctx, d_pp, n, C.cast(out.data_ptr(), _capi.FP), self._stream()))
            return out
        planes = _planes_np(d, 3, "d")
        n = planes[0].size
        out = np.empty(n, dtype=np.float32)
        _check(self._L.ssk_pdf_direction_host(self._ctx, _pp(planes), n, out.ctypes.data_as(_capi.FP)))
        return out

    def sample_direction(self, u, it_p=None, wavelengths=None, want_p=False):
        """Importance-sample a direction. u: (2, n) in [0,1)^2; it_p: (3, n) reference points or None;
        wavelengths: (4, n) (spectral). Returns (d (3,n), pdf (n,), weight (C,n)) and, with want_p,
        also (p (3,n), dist (n,)) = ds.p, ds.dist."""
        Cn = self.channels
        if _is_torch(u):
            import torch
            u_pp, u = self._pp_t(u, 2, "u")
            n = u.shape[1]
            dev = u.device
            p_pp = w_pp = None
            if it_p is not None:
                p_pp, it_p = self._pp_t(it_p, 3, "it_p")
            if self.spectral:
                if wavelengths is None:
                    raise ValueError("spectral mode needs wavelengths of shape (4, n)")
                w_pp, wavelengths = self._pp_t(wavelengths, 4, "wavelengths")
            d = torch.empty((3, n), dtype=torch.float32, device=dev)
            pdf = torch.empty(n, dtype=torch.float32, device=dev)
            wgt = torch.empty((Cn, n), dtype=torch.float32, device=dev)
            d_pp, _ = self._pp_t(d, 3, "d"); q_pp, _ = self._pp_t(wgt, Cn, "weight")
            op_pp = dist_p = None
            if want_p:
                op = torch.empty((3, n), dtype=torch.float32, device=dev)
                dist = torch.empty(n, dtype=torch.float32, device=dev)
                op_pp, _ = self._pp_t(op, 3, "p"); dist_p = C.cast(dist.data_ptr(), _capi.FP)
            _check(self._L.ssk_sample_direction_device(self._ctx, u_pp, p_pp, w_pp, n, d_pp,
                                                       C.cast(pdf.data_ptr(), _capi.FP), q_pp, op_pp, dist_p,
                                                       self._stream()))
            return (d, pdf, wgt, op, dist) if want_p else (d, pdf, wgt)
        up = _planes_np(u, 2, "u")
        n = up[0].size
        pp = _pp(_planes_np(it_p, 3, "it_p")) if it_p is not None else None
        wp = None
        if self.spectral:
            if wavelengths is None:
                raise ValueError("spectral mode needs wavelengths of shape (4, n)")
            wp = _pp(_planes_np(wavelengths, 4, "wavelengths"))
        d = np.empty((3, n), np.float32); pdf = np.empty(n, np.float32); wgt = np.empty((Cn, n), np.float32)
        op = np.empty((3, n), np.float32) if want_p else None
        dist = np.empty(n, np.float32) if want_p else None
        _check(self._L.ssk_sample_direction_host(
            self._ctx, _pp(up), pp, wp, n, _pp([d[i] for i in range(3)]), pdf.ctypes.data_as(_capi.FP),
            _pp([wgt[c] for c in range(Cn)]), _pp([op[i] for i in range(3)]) if want_p else None,
            dist.ctypes.data_as(_capi.FP) if want_p else None))
        return (d, pdf, wgt, op, dist) if want_p else (d, pdf, wgt)


    def eval_jvp(self, wi, wavelengths=None, d_turbidity=0.0, d_albedo=None, d_sun_direction=None):
        """Forward derivative of eval() along (d_turbidity, d_albedo, d_sun_direction): what dr.forward_from(params[...])
        gives on plugin.eval(si) (sunsky.cpp:222-237, sunsky-testing/traversal_test.py:94-144).
        d_albedo: None, a scalar, or one tangent per channel (3 RGB / 11 spectral bins); d_sun_direction: None or 3
        floats (world space). Device tensors only."""
        import torch
        if not _is_torch(wi):
            raise TypeError("eval_jvp takes CUDA tensors (float32, plane-major)")
        d_pp, wi = self._pp_t(wi, 3, "wi")
        n = wi.shape[1]
        w_pp = None
        if self.spectral:
            if wavelengths is None:
                raise ValueError("spectral mode needs wavelengths of shape (4, n)")
            w_pp, wavelengths = self._pp_t(wavelengths, 4, "wavelengths")
            if wavelengths.shape[1] != n:
                raise ValueError("wavelengths: expected (4, n)")
        da = np.zeros(0, np.float32) if d_albedo is None else np.ascontiguousarray(d_albedo, dtype=np.float32).reshape(-1)
        out = torch.empty((self.channels, n), dtype=torch.float32, device=wi.device)
        o_pp, _ = self._pp_t(out, self.channels, "out")
        ds = None if d_sun_direction is None else np.ascontiguousarray(d_sun_direction, dtype=np.float32).reshape(3)
        _check(self._L.ssk_eval_jvp_device(self._ctx, float(d_turbidity), da.ctypes.data_as(_capi.FP) if da.size else None,
                                           int(da.size), None if ds is None else ds.ctypes.data_as(_capi.FP), d_pp, w_pp, n,
                                           o_pp, self._stream()))
        return out

    def sample_wavelengths(self, wi, sample=None):
        """Emitter.sample_wavelengths(si, sample) (sunsky.cpp:463-480). wi: (3, n); sample: (n,) in [0, 1).
        Spectral: returns (wavelengths (4, n), weight (4, n) = eval / pdf). RGB: (None, eval (3, n))."""
        if not self.spectral:
            return None, self.eval(wi)
        if sample is None:
            raise ValueError("spectral mode needs the wavelength sample of shape (n,)")
        if _is_torch(wi):
            import torch
            d_pp, wi = self._pp_t(wi, 3, "wi")
            n = wi.shape[1]
            sample = sample.contiguous()
            if sample.dtype != torch.float32 or sample.numel() != n or sample.device != wi.device:
                raise ValueError("sample: expected float32 CUDA tensor of n elements on the emitter's device")
            wav = torch.empty((4, n), dtype=torch.float32, device=wi.device)
            wgt = torch.empty((4, n), dtype=torch.float32, device=wi.device)
            w_pp, _ = self._pp_t(wav, 4, "wavelengths"); q_pp, _ = self._pp_t(wgt, 4, "weight")
            _check(self._L.ssk_sample_wavelengths_device(self._ctx, d_pp, C.cast(sample.data_ptr(), _capi.FP), n,
                                                         w_pp, q_pp, self._stream()))
            return wav, wgt
        planes = _planes_np(wi, 3, "wi")
        n = planes[0].size
        sample = np.ascontiguousarray(sample, dtype=np.float32).reshape(-1)
        if sample.size != n:
            raise ValueError(f"sample: expected {n} elements, got {sample.size}")
        wav = np.empty((4, n), np.float32); wgt = np.empty((4, n), np.float32)
        _check(self._L.ssk_sample_wavelengths_host(self._ctx, _pp(planes), sample.ctypes.data_as(_capi.FP), n,
                                                   _pp([wav[c] for c in range(4)]), _pp([wgt[c] for c in range(4)])))
        return wav, wgt

    def sample_ray(self, wavelength_sample, sample2, sample3):
        """Emitter.sample_ray(time, wavelength_sample, sample2, sample3) (sunsky.cpp:354-397).
        wavelength_sample: (n,) (ignored for RGB, may be None); sample2, sample3: (2, n).
        Returns (o (3, n), d (3, n), wavelengths (4, n) or None, weight (C, n))."""
        Cn = self.channels
        if _is_torch(sample2):
            import torch
            a_pp, sample2 = self._pp_t(sample2, 2, "sample2")
            b_pp, sample3 = self._pp_t(sample3, 2, "sample3")
            n = sample2.shape[1]
            dev = sample2.device
            wl_p = w_pp = wav = None
            if self.spectral:
                if wavelength_sample is None:
                    raise ValueError("spectral mode needs wavelength_sample of shape (n,)")
                wavelength_sample = wavelength_sample.contiguous()
                if wavelength_sample.dtype != torch.float32 or wavelength_sample.numel() != n or wavelength_sample.device != dev:
                    raise ValueError("wavelength_sample: expected float32 CUDA tensor of n elements on the emitter's device")
                wl_p = C.cast(wavelength_sample.data_ptr(), _capi.FP)
                wav = torch.empty((4, n), dtype=torch.float32, device=dev)
                w_pp, _ = self._pp_t(wav, 4, "wavelengths")
            o = torch.empty((3, n), dtype=torch.float32, device=dev)
            d = torch.empty((3, n), dtype=torch.float32, device=dev)
            wgt = torch.empty((Cn, n), dtype=torch.float32, device=dev)
            o_pp, _ = self._pp_t(o, 3, "o"); d_pp, _ = self._pp_t(d, 3, "d"); q_pp, _ = self._pp_t(wgt, Cn, "weight")
            _check(self._L.ssk_sample_ray_device(self._ctx, wl_p, a_pp, b_pp, n, o_pp, d_pp, w_pp, q_pp, self._stream()))
            return o, d, wav, wgt
        a = _planes_np(sample2, 2, "sample2"); b = _planes_np(sample3, 2, "sample3")
        n = a[0].size
        wl_p = w_pp = wav = None
        if self.spectral:
            if wavelength_sample is None:
                raise ValueError("spectral mode needs wavelength_sample of shape (n,)")
            wl = np.ascontiguousarray(wavelength_sample, dtype=np.float32).reshape(-1)
            if wl.size != n:
                raise ValueError(f"wavelength_sample: expected {n} elements, got {wl.size}")
            wl_p = wl.ctypes.data_as(_capi.FP)
            wav = np.empty((4, n), np.float32)
            w_pp = _pp([wav[c] for c in range(4)])
        o = np.empty((3, n), np.float32); d = np.empty((3, n), np.float32); wgt = np.empty((Cn, n), np.float32)
        _check(self._L.ssk_sample_ray_host(self._ctx, wl_p, _pp(a), _pp(b), n, _pp([o[i] for i in range(3)]),
                                           _pp([d[i] for i in range(3)]), w_pp, _pp([wgt[c] for c in range(Cn)])))
        return o, d, wav, wgt


def update_batch(emitters, props_list):
    """parameters_changed for several emitters of one device and mode at once (ssk_update_batch): props_list[i] are
    the keyword arguments `emitters[i].update()` would take. All or nothing: on an error no emitter changes."""
    if len(emitters) != len(props_list):
        raise ValueError("one property dict per emitter")
    if not emitters:
        return
    n = len(emitters)
    arr = (Params * n)()
    new = []
    for i, (e, props) in enumerate(zip(emitters, props_list)):
        unknown = set(props) - set(_PARAM_KEYS) | ({"sun_aperture"} & set(props))
        if unknown:
            raise TypeError(f"not a traversable sunsky parameter: {sorted(unknown)}")
        p = Params.from_buffer_copy(e._params)
        e._fill(p, props, creating=False)
        arr[i] = p
        new.append(p)
    ctxs = (_capi.CTX * n)(*[e._ctx for e in emitters])
    _check(_capi.lib().ssk_update_batch(ctxs, arr, n))
    for e, p in zip(emitters, new):
        e._params = p


def envmaps(emitters, width, height, wavelengths4=None, out=None):
    """Equirectangular radiance maps, one per emitter (all on one device, same mode), written by one
    kernel launch straight into a torch CUDA tensor of shape (n_maps, C, height, width)."""
    import torch
    L = _capi.lib()
    e0 = emitters[0]
    Cn = e0.channels
    n = len(emitters)
    dev = torch.device("cuda", e0.device)
    if out is None:
        out = torch.empty((n, Cn, height, width), dtype=torch.float32, device=dev)
    ctxs = (_capi.CTX * n)(*[e._ctx for e in emitters])
    wl = None
    if e0.spectral:
        wl_np = np.ascontiguousarray(np.broadcast_to(np.asarray(wavelengths4, dtype=np.float32), (n, 4)))
        wl = wl_np.ctypes.data_as(_capi.FP)
    with torch.cuda.device(dev):
        _check(L.ssk_envmap_device(ctxs, n, width, height, wl, C.cast(out.data_ptr(), _capi.FP),
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def envmap_directions(emitter, width, height):
    """Device-generated pixel directions of `envmaps` (3, height*width), for parity checks."""
    import torch
    L = _capi.lib()
    d = torch.empty((3, height * width), dtype=torch.float32, device=torch.device("cuda", emitter.device))
    pp, _ = _pp_t(d, 3, "d")
    _check(L.ssk_envmap_directions_device(emitter._ctx, width, height, pp,
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return d


def sun_coordinates(year, month, day, hour, minute, second, latitude, longitude, timezone):
    out = np.empty(3, np.float32)
    _check(_capi.lib().ssk_sun_coordinates(year, month, day, hour, minute, second, latitude, longitude, timezone,
                                           out.ctypes.data_as(_capi.FP)))
    return out


def hosek_sun_rad(turbidity, wavelength, elevation, gamma, dataset_dir=None):
    """`mi.hosek_sun_rad` (src/render/python/sunsky_v.cpp:19): direct solar radiance of the original fp64
    Hosek-Wilkie model."""
    out = C.c_double()
    _check(_capi.lib().ssk_hosek_sun_rad((dataset_dir or DATASET_DIR).encode(), float(turbidity), float(wavelength),
                                         float(elevation), float(gamma), C.byref(out)))
    return out.value


def array_from_file(path, file_is_f64=True):
    """`mi.array_from_file_d / _f` (src/render/python/sunsky_v.cpp:11-20): dataset -> (fp32 array, shape)"""
    L = _capi.lib()
    cnt = C.c_size_t(); nd = C.c_size_t(); shape = (C.c_size_t * 8)()
    _check(L.ssk_array_from_file(path.encode(), int(file_is_f64), None, 0, C.byref(cnt), shape, C.byref(nd)))
    out = np.empty(cnt.value, np.float32)
    _check(L.ssk_array_from_file(path.encode(), int(file_is_f64), out.ctypes.data_as(_capi.FP), out.size,
                                 C.byref(cnt), shape, C.byref(nd)))
    return out.reshape([shape[i] for i in range(nd.value)])


def array_to_file(path, data, as_f64=False):
    a = np.ascontiguousarray(data, dtype=np.float32)
    shape = (C.c_size_t * a.ndim)(*a.shape)
    _check(_capi.lib().ssk_array_to_file(path.encode(), a.ctypes.data_as(_capi.FP), shape, a.ndim, int(as_f64)))


def launch_count():
    return int(_capi.lib().ssk_launch_count())
