"""TEST INFRASTRUCTURE shared by tests/test_reference_suite_{gpu,cpu}.py: the grids, emitters and error measures of
the reference's src/emitters/tests/test_sunsky.py, written against any object with the plugin's surface
(`make(spectral=..., **properties)` -> emitter with .eval(wi, wavelengths)): the product emitter on the GPU, or
the reference library itself on the CPU (which pins the oracle to the reference's shipped golden data)."""
import os

import numpy as np

import common

EPS = 1e-4
SUN_HALF_APERTURE_ANGLE = np.radians(0.5388 / 2.0)                       # test_sunsky.py:10 (sic: the plugin uses 0.5358)
SPECIAL_ALBEDO = [0.56, 0.21, 0.58, 0.24, 0.92, 0.42, 0.53, 0.75, 0.54, 0.20, 0.46]   # :12-16, at 320..720 nm
RGB_CASES = [(9.5, 2, 0.2), (12.25, 5.2, 0.0), (18.3, 9.8, 0.5)]         # :115-119 (hour, turbidity, albedo)
SPEC_CASES = [(np.radians(2), 2, 0.0), (np.radians(20), 5.2, 0.0), (np.radians(45), 9.8, 0.0)]   # :129-133


def load_refdata():
    return np.load(os.path.join(common.GOLDEN_DIR, "reference_test_data.npz"))


def sun_direction(sun_phi, sun_theta):
    """make_emitter_angles (test_sunsky.py:28-39): Python-float sincos, handed over as a property."""
    return [float(np.cos(sun_phi) * np.sin(sun_theta)), float(np.sin(sun_phi) * np.sin(sun_theta)), float(np.cos(sun_theta))]


def render_error(make, refdata, spectral, render_params, ref_key):
    """generate_and_compare (test_sunsky.py:62-110): mean(|render - ref| / (|ref| + 0.001)) on the 32 x 64 grid."""
    H, W = 32, 64
    if not spectral:
        hour, turb, albedo = render_params
        e = make(spectral=False, hour=hour, turbidity=turb, albedo=albedo, sun_scale=0.0, sky_scale=1.0)
    else:
        sun_eta, turb, albedo = render_params
        e = make(spectral=True, sun_direction=sun_direction(0.0, np.pi / 2 - sun_eta), turbidity=turb, albedo=albedo,
                 sun_scale=0.0, sky_scale=1.0)
    wi = common.latlong_grid(W, H)                                       # :93-100
    if not spectral:
        rendered = np.asarray(e.eval(wi)).T.reshape(H, W, 3)
    else:
        start, end = 360, 830
        step = (end - start) / 10
        bands = []
        for i in range(10):
            lbda = np.float32(start + step / 2 + i * step)
            bands.append(np.asarray(e.eval(wi, np.full((4, wi.shape[1]), lbda, np.float32)))[0])   # plugin.eval(si)[0], :57
        rendered = np.stack(bands, axis=-1).reshape(H, W, 10)
    reference = refdata["render/" + ref_key]
    return float(np.mean(np.abs(rendered - reference) / (np.abs(reference) + 0.001)))


def rgb_key(p):
    return f"sky_rgb_hour{p[0]:.2f}_t{p[1]:.3f}_a{p[2]:.3f}"


def spec_key(p):
    return f"sky_spec_eta{p[0]:.3f}_t{p[1]:.3f}_a{p[2]:.3f}"


def sun_spectrum_errors(make, refdata):
    """test04_sun_radiance (test_sunsky.py:154-196) over all 5 x 4 x 4 = 80 .spd spectra -> (5, 4, 4) array of
    mean(|res - ref| / (ref + 1e-6))."""
    wavelengths = np.linspace(310, 800, 15)
    assert np.array_equal(wavelengths, refdata["sun_wavelengths"])
    turbs = np.linspace(1, 10, 5)
    etas = np.linspace(EPS, np.pi / 2 - EPS, 4)
    gammas = np.linspace(0, SUN_HALF_APERTURE_ANGLE - EPS, 4)
    err = np.zeros((5, 4, 4))
    for i, turb in enumerate(turbs):
        for j, eta_ray in enumerate(etas):
            for k, gamma in enumerate(gammas):
                phi = np.pi / 5
                theta_ray = np.pi / 2 - eta_ray
                sun_theta = theta_ray - gamma
                if sun_theta < 0:
                    sun_theta = theta_ray + gamma
                e = make(spectral=True, sun_direction=sun_direction(phi, sun_theta), turbidity=float(turb), albedo=0.0,
                         sun_scale=1.0, sky_scale=0.0)
                d = np.array([np.cos(phi) * np.sin(theta_ray), np.sin(phi) * np.sin(theta_ray), np.cos(theta_ray)])
                wi = np.repeat((-d).astype(np.float32)[:, None], 15, axis=1)
                wav = np.repeat(wavelengths.astype(np.float32)[None, :], 4, axis=0)
                res = np.asarray(e.eval(wi, wav))[0].astype(np.float64)
                ref_rad = refdata["sun_spectra"][i, j, k]
                err[i, j, k] = float(np.mean(np.abs(res - ref_rad) / (ref_rad + 1e-6)))
                if hasattr(e, "close"):
                    e.close()
    return err
