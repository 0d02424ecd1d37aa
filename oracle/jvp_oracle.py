"""TEST INFRASTRUCTURE -- independent oracle for the forward derivative of SunskyEmitter::eval.

The reference differentiates eval() with Dr.Jit's AD (traverse() flags turbidity, albedo and sun_direction as
Differentiable, src/emitters/sunsky.cpp:222-237; sunsky-testing/traversal_test.py:94-144 drives it with
dr.forward_from), which cannot run here (no JIT). This module restates the same formulas -- state interpolation
(include/mitsuba/render/sunsky/sunsky.h:157-231, 403-419), eval / render_sky / render_sun / compute_sun_ld
(sunsky.cpp:303-352, 538-650), unit_angle / unit_angle_z (drjit/sphere.h:53-81) -- in numpy float64 on forward-mode
DUAL NUMBERS: every operation propagates (value, tangent) exactly as forward-mode AD does, including the
non-differentiable pieces AD treats as constants (floor, masks, the segment / bin indices). It is neither finite
differences nor the CUDA kernel. It is pinned twice in tests/test_jvp_oracle_cpu.py: its VALUE against the
reference library (oracle/_ref, fp32, <= 2e-5) and its TANGENT against central differences of itself in fp64
(<= 1e-6) and of the reference library (fp32 noise level).

Only tests/ may import this module."""
import os

import numpy as np

HALF_PI = 0.5 * np.pi
SUN_APERTURE_DEG = 0.5358
NB_SEG = 45


class Dual:
    """value + eps * tangent, elementwise on numpy arrays (float64)."""
    __array_priority__ = 1000

    def __init__(self, v, t=None):
        self.v = np.asarray(v, dtype=np.float64)
        self.t = np.zeros_like(self.v) if t is None else np.asarray(t, dtype=np.float64) + np.zeros_like(self.v)

    @staticmethod
    def lift(x):
        return x if isinstance(x, Dual) else Dual(x)

    def __add__(self, o):
        o = Dual.lift(o); return Dual(self.v + o.v, self.t + o.t)
    __radd__ = __add__

    def __sub__(self, o):
        o = Dual.lift(o); return Dual(self.v - o.v, self.t - o.t)

    def __rsub__(self, o):
        return Dual.lift(o) - self

    def __neg__(self):
        return Dual(-self.v, -self.t)

    def __mul__(self, o):
        o = Dual.lift(o); return Dual(self.v * o.v, self.t * o.v + self.v * o.t)
    __rmul__ = __mul__

    def __truediv__(self, o):
        o = Dual.lift(o); return Dual(self.v / o.v, (self.t * o.v - self.v * o.t) / (o.v * o.v))

    def __rtruediv__(self, o):
        return Dual.lift(o) / self

    def __pow__(self, k):          # integer or real constant exponent
        return Dual(self.v ** k, k * self.v ** (k - 1) * self.t)


def dexp(x):
    e = np.exp(x.v); return Dual(e, e * x.t)


def dsqrt(x):
    s = np.sqrt(np.maximum(x.v, 0)); return Dual(s, np.where(s > 0, 0.5 * x.t / np.where(s > 0, s, 1), 0.0))


def dsin(x):
    return Dual(np.sin(x.v), np.cos(x.v) * x.t)


def dcos(x):
    return Dual(np.cos(x.v), -np.sin(x.v) * x.t)


def dasin(x):
    return Dual(np.arcsin(x.v), x.t / np.sqrt(np.maximum(1 - x.v * x.v, 1e-300)))


def dcbrt(x):
    c = np.cbrt(x.v); return Dual(c, x.t / (3 * c * c))


def where(m, a, b):
    a, b = Dual.lift(a), Dual.lift(b)
    return Dual(np.where(m, a.v, b.v), np.where(m, a.t, b.t))


def lerp(a, b, t):
    return a + (b - a) * t


def unit_angle(a, b):
    """drjit/sphere.h:53-59: 2 asin(|b - mulsign(a, a.b)| / 2), pi - that when a.b < 0. a, b: 3-lists of Dual."""
    dot = a[0] * b[0] + a[1] * b[1] + a[2] * b[2]
    sgn = np.where(dot.v >= 0, 1.0, -1.0)
    d = [b[i] - a[i] * sgn for i in range(3)]
    temp = 2 * dasin(0.5 * dsqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]))
    return where(dot.v >= 0, temp, np.pi - temp)


def unit_angle_z(v):
    """drjit/sphere.h:73-81"""
    sgn = np.where(v[2].v >= 0, 1.0, -1.0)
    dz = v[2] - sgn
    temp = 2 * dasin(0.5 * dsqrt(v[0] * v[0] + v[1] * v[1] + dz * dz))
    return where(v[2].v >= 0, temp, np.pi - temp)


def read_dataset(path, f64=True):
    raw = open(path, "rb").read()
    assert raw[:3] in (b"SKY", b"SUN")
    ndim = int(np.frombuffer(raw, np.uint64, 1, 7)[0])
    shape = np.frombuffer(raw, np.uint64, ndim, 15).astype(int)
    return np.frombuffer(raw, np.float64 if f64 else np.float32, int(np.prod(shape)), 15 + 8 * ndim).astype(np.float64).reshape(shape)


class JvpOracle:
    def __init__(self, dataset_dir, spectral):
        t = "_spec_" if spectral else "_rgb_"
        f32 = lambda a: a.astype(np.float32).astype(np.float64)         # the plugin keeps the datasets in fp32
        self.spectral = spectral
        self.C = 11 if spectral else 3
        self.sky_params = f32(read_dataset(os.path.join(dataset_dir, f"sky{t}params.bin")))   # (10, 2, 6, C, 9)
        self.sky_rad = f32(read_dataset(os.path.join(dataset_dir, f"sky{t}rad.bin")))         # (10, 2, 6, C)
        self.sun_rad = f32(read_dataset(os.path.join(dataset_dir, f"sun{t}rad.bin")))         # (10, 45, C, 4[, 6])
        self.sun_ld = f32(read_dataset(os.path.join(dataset_dir, "sun_spec_ld.bin"))) if spectral else None   # (11, 6)

    # ---- state (sunsky.h:157-231, 403-419) as duals -----------------------------------------------------
    def state(self, turbidity, albedo, sun_dir):
        """turbidity: Dual scalar; albedo: list of C Duals; sun_dir: 3-list of Duals (local frame, to_world = I)."""
        theta = unit_angle_z(sun_dir)
        eta = HALF_PI - theta
        x = dcbrt(2 / np.pi * eta)
        t_high = int(np.floor(turbidity.v)); t_low = t_high - 1
        t_rem = turbidity - float(t_high)
        coefs = [1, 5, 10, 10, 5, 1]
        w = [coefs[i] * (x ** i if i else Dual(1.0)) * ((1 - x) ** (5 - i) if i < 5 else Dual(1.0)) for i in range(6)]

        def bezier(data, t_idx, a_idx):          # data (10, 2, 6, ...) -> Dual array of shape data.shape[3:]
            if not (0 <= t_idx < 10):
                return Dual(np.zeros(data.shape[3:]))
            acc = Dual(np.zeros(data.shape[3:]))
            for i in range(6):
                acc = acc + Dual(data[t_idx, a_idx, i]) * Dual(np.full(data.shape[3:], w[i].v), np.full(data.shape[3:], w[i].t))
            return acc

        def interp(data):
            lo = lerp(bezier(data, t_low, 0), bezier(data, t_high, 0), Dual(np.array(t_rem.v), np.array(t_rem.t)))
            hi = lerp(bezier(data, t_low, 1), bezier(data, t_high, 1), Dual(np.array(t_rem.v), np.array(t_rem.t)))
            av = np.array([a.v for a in albedo]); at = np.array([a.t for a in albedo])
            shape = (self.C,) + (1,) * (lo.v.ndim - 1)
            return lerp(lo, hi, Dual(av.reshape(shape), at.reshape(shape)))

        params, rad = interp(self.sky_params), interp(self.sky_rad)
        ok = 0.0 <= float(eta.v) <= HALF_PI
        if not ok:
            params, rad = Dual(np.zeros_like(params.v)), Dual(np.zeros_like(rad.v))
        s_lo = self.sun_rad[t_low] if 0 <= t_low < 10 else np.zeros_like(self.sun_rad[0])
        s_hi = self.sun_rad[t_high] if 0 <= t_high < 10 else np.zeros_like(self.sun_rad[0])
        sun = lerp(Dual(s_lo), Dual(s_hi), Dual(np.array(t_rem.v), np.array(t_rem.t)))
        return dict(params=params, rad=rad, sun=sun, n=sun_dir, eta=eta)

    # ---- eval (sunsky.cpp:303-352) -------------------------------------------------------------------------
    def eval(self, st, wo, wavelengths=None, sky_scale=1.0, sun_scale=1.0, area_ratio=1.0):
        """wo: (3, n) float64 local directions TOWARDS the sky (= -si.wi). Returns Dual of shape (C_out, n)."""
        n = wo.shape[1]
        b = [Dual(wo[i]) for i in range(3)]
        a = [Dual(np.full(n, c.v), np.full(n, c.t)) for c in st["n"]]
        cos_theta = wo[2]
        gamma = unit_angle(a, b)
        active = cos_theta >= 0
        half_ap = np.float32(0.5) * np.float32(SUN_APERTURE_DEG) * np.float32(np.pi / 180)
        dotv = a[0].v * wo[0] + a[1].v * wo[1] + a[2].v * wo[2]
        hit = active & (dotv >= np.cos(np.float64(half_ap)))
        # area_ratio: get_area_ratio (sunsky.h:98-101) is an fp32 artefact (1.00216 at the default aperture): a constant
        # handed in by the caller (ref_oracle.lib().ssk_ref_area_ratio(half_aperture))
        cg = dcos(gamma)
        cg2 = cg * cg
        ct = np.maximum(cos_theta, 0.0)

        def sky(ch):          # ch: int array (n,) -> Dual (n,)
            P = lambda k: Dual(st["params"].v[ch, k], st["params"].t[ch, k])
            L = Dual(st["rad"].v[ch], st["rad"].t[ch])
            c1 = 1 + P(0) * dexp(P(1) / (ct + 0.01))
            chi = (1 + cg2) / ((1 + P(8) * P(8) - 2 * P(8) * cg) ** 1.5)
            c2 = P(2) + P(3) * dexp(P(4) * gamma) + P(5) * cg2 + P(6) * chi + P(7) * np.sqrt(ct)
            return c1 * c2 * L

        # sun (render_sun, sunsky.cpp:572-614): the segment depends on the RAY's elevation only
        elev = HALF_PI - np.arccos(np.clip(cos_theta, -1, 1))
        pos = np.minimum(np.floor(np.cbrt(2 * elev / np.pi) * NB_SEG).astype(int), NB_SEG - 1)
        pos = np.where(active, pos, 0)
        brk = HALF_PI * (pos / NB_SEG) ** 3
        xs = elev - brk
        sg = dsin(gamma)
        ar2 = 1.0 / np.sin(np.float64(half_ap)) ** 2
        cos_psi = dsqrt(1 - ar2 * sg * sg)

        def sun_rgb(ch):
            acc = Dual(np.zeros(n))
            for k in range(4):
                for j in range(6):
                    S = Dual(st["sun"].v[pos, ch, k, j], st["sun"].t[pos, ch, k, j])
                    acc = acc + (xs ** k) * (cos_psi ** j if j else Dual(np.ones(n))) * S
            return acc

        def sun_spec(ch):     # ch int array; channel 11 reads as 0
            chc = np.minimum(ch, 10)
            acc = Dual(np.zeros(n))
            for k in range(4):
                S = Dual(np.where(ch <= 10, st["sun"].v[pos, chc, k], 0.0), np.where(ch <= 10, st["sun"].t[pos, chc, k], 0.0))
                acc = acc + (xs ** k) * S
            return acc

        out = []
        if not self.spectral:
            for c in range(3):
                chv = np.full(n, c)
                res = sky_scale * sky(chv)
                res = res + where(hit, sun_scale * sun_rgb(chv) * area_ratio * 467.069280386, Dual(np.zeros(n)))
                res = res * (1.0 / 106.7502593994140625)
                out.append(where(active, res, Dual(np.zeros(n))))
        else:
            Pv = np.concatenate([st["params"].v, np.zeros((1, 9))]); Pt = np.concatenate([st["params"].t, np.zeros((1, 9))])
            Rv = np.concatenate([st["rad"].v, [0.0]]); Rt = np.concatenate([st["rad"].t, [0.0]])
            st = dict(st, params=Dual(Pv, Pt), rad=Dual(Rv, Rt))
            for wq in range(4):
                nw = (wavelengths[wq] - 320.0) / 40.0
                valid = (nw >= 0) & (nw <= 10)
                lo = np.where(valid, np.floor(nw), 0).astype(int); hi = lo + 1
                f = nw - np.floor(nw)
                res = sky_scale * lerp(sky(lo), sky(hi), f)
                ld = Dual(np.zeros(n))
                ldt = np.concatenate([self.sun_ld, np.zeros((1, 6))])
                for j in range(6):
                    ld = ld + (cos_psi ** j if j else Dual(np.ones(n))) * (ldt[lo, j] + (ldt[hi, j] - ldt[lo, j]) * f)
                sun = lerp(sun_spec(lo), sun_spec(hi), f)
                res = res + where(hit, sun_scale * sun * ld * area_ratio, Dual(np.zeros(n)))
                out.append(where(active & valid, res, Dual(np.zeros(n))))
        return Dual(np.stack([o.v for o in out]), np.stack([o.t for o in out])), hit

    def eval_jvp(self, turbidity, albedo, sun_dir, wo, wavelengths=None, d_turbidity=0.0, d_albedo=None, d_sun_dir=None,
                 sky_scale=1.0, sun_scale=1.0, area_ratio=1.0):
        """(value, tangent, hit_sun) of eval along (d_turbidity, d_albedo[C], d_sun_dir[3]); to_world = identity."""
        alb = np.broadcast_to(np.asarray(albedo, np.float64), (self.C,))
        da = np.zeros(self.C) if d_albedo is None else np.broadcast_to(np.asarray(d_albedo, np.float64), (self.C,))
        ds = np.zeros(3) if d_sun_dir is None else np.asarray(d_sun_dir, np.float64)
        st = self.state(Dual(float(turbidity), float(d_turbidity)), [Dual(alb[c], da[c]) for c in range(self.C)],
                        [Dual(float(sun_dir[i]), float(ds[i])) for i in range(3)])
        r, hit = self.eval(st, np.asarray(wo, np.float64), None if wavelengths is None else np.asarray(wavelengths, np.float64),
                           sky_scale, sun_scale, area_ratio)
        return r.v, r.t, hit
