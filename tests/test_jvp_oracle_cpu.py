"""Pins oracle/jvp_oracle.py (fp64 dual-number restatement of SunskyEmitter::eval, the independent oracle of the
eval_jvp kernel): its value against the reference library, its tangent against central differences of itself
(fp64: the dual arithmetic is right) and of the reference library (fp32: the restated formulas are the reference's)."""
import os

import numpy as np
import pytest

import common

DATA = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mitsuba3-sunsky_b200", "data", "resources",
                    "sunsky", "datasets")


def _inputs(spectral, sun_dir, n=1500, seed=5):
    rng = np.random.default_rng(seed)
    wo = np.concatenate([common.hemisphere_dirs(n, rng), common.around(sun_dir, 0.2, 64, rng)], axis=1).astype(np.float32)
    wav = (330.0 + 380.0 * rng.random((4, wo.shape[1]), dtype=np.float32)).astype(np.float32) if spectral else None
    return wo, wav


def _area_ratio(ref):
    half = np.float32(0.5) * np.float32(0.5358) * np.float32(np.pi / 180)
    return float(ref.lib().ssk_ref_area_ratio(half))


CASES = [(False, 3.4, 0.3, 30.0), (True, 6.7, 0.55, 50.2), (False, 9.2, 0.9, 8.0), (True, 1.6, 0.1, 75.0)]


@pytest.mark.parametrize("spectral,t,a,eta", CASES)
def test_value_matches_the_reference(ref, spectral, t, a, eta):
    import jvp_oracle
    sd = np.array(common.sun_dir_from_elevation(eta, 0.9), np.float32)
    r = ref.RefEmitter(spectral=spectral, flavour=1, turbidity=t, albedo=a, sun_direction=list(sd))
    sdn = np.array(r.state().sun_dir_world, np.float64)
    wo, wav = _inputs(spectral, sdn)
    want = r.eval(-wo, wav).astype(np.float64)
    o = jvp_oracle.JvpOracle(DATA, spectral)
    val, _, hit = o.eval_jvp(t, a, sdn, wo, wav, area_ratio=_area_ratio(ref))
    assert np.array_equal(val == 0, want == 0)
    nz = want != 0
    rel = np.abs(val[nz] - want[nz]) / np.abs(want[nz])
    # fp32 reference vs fp64 restatement: cancellation-prone lanes (low sun, horizon) carry up to ~1e-4 of fp32 noise
    assert np.quantile(rel, 0.99) < 2e-5 and rel.max() < 2e-3, (np.quantile(rel, 0.99), rel.max())
    assert hit.sum() >= 32                                              # the sun-disc term is exercised


@pytest.mark.parametrize("spectral,t,a,eta", CASES)
def test_tangent_matches_central_differences_of_itself(spectral, t, a, eta):
    import jvp_oracle
    o = jvp_oracle.JvpOracle(DATA, spectral)
    sd = np.array(common.sun_dir_from_elevation(eta, 0.9), np.float64)
    wo, wav = _inputs(spectral, sd, n=600)
    C = 11 if spectral else 3
    rng = np.random.default_rng(1)
    for d_t, d_a, d_s in ((1.0, None, None), (0.0, rng.normal(size=C), None), (0.0, None, rng.normal(size=3)),
                          (0.7, rng.normal(size=C), rng.normal(size=3))):
        val, tan, hit = o.eval_jvp(t, a, sd, wo, wav, d_turbidity=d_t, d_albedo=d_a, d_sun_dir=d_s)
        h = 1e-6
        alb = np.full(C, a)
        da = np.zeros(C) if d_a is None else d_a
        ds = np.zeros(3) if d_s is None else d_s
        vp, _, hp = o.eval_jvp(t + h * d_t, alb + h * da, sd + h * ds, wo, wav)
        vm, _, hm = o.eval_jvp(t - h * d_t, alb - h * da, sd - h * ds, wo, wav)
        fd = (vp - vm) / (2 * h)
        stable = (hp == hit) & (hm == hit)                              # lanes that do not cross the sun-disc edge
        scale = np.abs(tan[:, stable]).max()
        tol = 2e-6 * scale + 1e-9 + 4e-16 * np.abs(val[:, stable]) / h          # + the rounding noise of the difference quotient
        assert np.all(np.abs(fd[:, stable] - tan[:, stable]) <= tol), (d_t, d_a is not None, d_s is not None)


@pytest.mark.parametrize("spectral,t,a,eta", CASES[:2])
def test_tangent_matches_central_differences_of_the_reference(ref, spectral, t, a, eta):
    """The fp32 reference library differentiated numerically (h chosen against its rounding noise) agrees with the
    dual-number tangent: the restated formulas are the reference's, parameter by parameter."""
    import jvp_oracle
    o = jvp_oracle.JvpOracle(DATA, spectral)
    sd = np.array(common.sun_dir_from_elevation(eta, 0.9), np.float32)
    sdn = np.array(ref.RefEmitter(spectral=spectral, flavour=1, turbidity=t, albedo=a, sun_direction=list(sd)).state().sun_dir_world, np.float64)
    wo, wav = _inputs(spectral, sdn, n=600)
    sky_only = dict(sun_scale=0.0)                                      # the sun edge moves with the sun direction
    C = 11 if spectral else 3
    perp = np.cross(sdn, [0.0, 0.0, 1.0]); perp /= np.linalg.norm(perp)
    for d_t, d_a, d_s, h in ((1.0, None, None, 1 / 64), (0.0, np.ones(C), None, 1 / 64), (0.0, None, perp, 1 / 256)):
        _, tan, _ = o.eval_jvp(t, a, sdn, wo, wav, d_turbidity=d_t, d_albedo=d_a, d_sun_dir=d_s, sun_scale=0.0)
        da = 0.0 if d_a is None else 1.0
        ds = np.zeros(3) if d_s is None else d_s

        def ref_eval(sign):
            s = sdn + sign * h * ds
            r = ref.RefEmitter(spectral=spectral, flavour=1, turbidity=t + sign * h * d_t, albedo=a + sign * h * da,
                               sun_direction=[float(v) for v in s], **sky_only)
            return r.eval(-wo, wav).astype(np.float64)
        fd = (ref_eval(+1) - ref_eval(-1)) / (2 * h)
        # the sun_direction property is re-normalised by the constructor: compare along a tangent perpendicular to it.
        # Within a few degrees of the sun a step of 1/256 rad is as large as gamma itself (the fp64 self-check above
        # covers those lanes with h = 1e-6): keep the lanes more than 5 degrees away
        far = (wo.astype(np.float64) * sdn[:, None]).sum(0) < np.cos(np.radians(5.0)) if d_s is not None else np.ones(wo.shape[1], bool)
        scale = np.abs(tan[:, far]).max()
        err = np.abs(fd - tan)[:, far].max()
        assert err <= 4e-3 * scale, (d_t, d_a is not None, d_s is not None, err / scale)
