"""Shared workload definitions for the tests, the golden-vector generator and the bench: the emitter
configurations of BASELINE.json / SURVEY.md section 8(d) and deterministic input sets."""
import numpy as np

GOLDEN_DIR = __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.abspath(__file__)), "golden")

PHI_SUN = -4.0 * np.pi / 5.0          # azimuth used throughout the reference's sunsky-testing/ scripts


def sphdir(theta, phi):
    return [float(np.cos(phi) * np.sin(theta)), float(np.sin(phi) * np.sin(theta)), float(np.cos(theta))]


def sun_dir_from_elevation(eta_deg, phi=PHI_SUN):
    return sphdir(np.pi / 2 - np.radians(eta_deg), phi)


# name -> (spectral, emitter keyword arguments)
CONFIGS = {
    # BASELINE configs[0] / [1]: t=3, a=0.3, sun elevation 30 deg
    "cfg1_rgb": (False, dict(turbidity=3.0, albedo=0.3, sun_direction=sun_dir_from_elevation(30.0))),
    "cfg2_spec": (True, dict(turbidity=3.0, albedo=0.3, sun_direction=sun_dir_from_elevation(30.0))),
    # BASELINE configs[2] = sunsky-testing/sampling_test.py:7-25 (sky only) and its sun+sky twin
    "cfg3_rgb_sky": (False, dict(turbidity=6.0, albedo=0.5, sun_direction=sun_dir_from_elevation(50.2), sun_scale=0.0)),
    "cfg3_rgb_sunsky": (False, dict(turbidity=6.0, albedo=0.5, sun_direction=sun_dir_from_elevation(50.2))),
    "cfg3_spec_sunsky": (True, dict(turbidity=6.0, albedo=0.5, sun_direction=sun_dir_from_elevation(50.2))),
    # edges: fractional turbidity + per-channel albedo, turbidity limits, low / high sun, wide aperture
    "rgb_t4p2_alb3": (False, dict(turbidity=4.2, albedo=[0.1, 0.5, 0.9], sun_direction=sun_dir_from_elevation(12.0))),
    "spec_t10_zenith": (True, dict(turbidity=10.0, albedo=0.0, sun_direction=sun_dir_from_elevation(88.0, 0.3))),
    "rgb_t1_low": (False, dict(turbidity=1.0, albedo=1.0, sun_direction=sun_dir_from_elevation(1.5, 2.0))),
    "spec_alb11": (True, dict(turbidity=7.7, albedo=[0.2, 0.1, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0, 0.0],
                              sun_direction=sun_dir_from_elevation(60.0, -1.0))),
    "rgb_aperture30": (False, dict(turbidity=5.0, albedo=0.3, sun_direction=sun_dir_from_elevation(40.0),
                                   sun_aperture=30.0)),
    # date/time record (test_sunsky.py:115-124 uses hour = 9.5 / 12.25 / 18.3 with the Tokyo defaults)
    "rgb_record_h9p5": (False, dict(turbidity=2.0, albedo=0.2, hour=9.5)),
}


def latlong_grid(width, height):
    """The lat-long grid of src/emitters/tests/test_sunsky.py:93-100: phi in [0, 2pi], theta from pi to 0."""
    phis = np.linspace(0, 2 * np.pi, width, dtype=np.float32)
    thetas = np.linspace(np.pi, 0, height, dtype=np.float32)
    P, T = np.meshgrid(phis, thetas)
    st = np.sin(T)
    return np.stack([np.cos(P) * st, np.sin(P) * st, np.cos(T)]).reshape(3, -1).astype(np.float32)


def hemisphere_dirs(n, rng):
    """wo uniform on the upper hemisphere (SURVEY 8d cfg 2): cos(theta) = u1, phi = 2 pi u2."""
    u1 = rng.random(n, dtype=np.float32); u2 = rng.random(n, dtype=np.float32)
    ct = u1; st = np.sqrt(np.maximum(0, 1 - ct * ct)); ph = (2 * np.pi * u2).astype(np.float32)
    return np.stack([st * np.cos(ph), st * np.sin(ph), ct]).astype(np.float32)


def around(direction, max_deg, n, rng):
    """Unit vectors within max_deg of `direction` (exercises the sun-disc branch)."""
    d = np.asarray(direction, dtype=np.float64); d /= np.linalg.norm(d)
    a = np.array([1.0, 0, 0]) if abs(d[0]) < 0.9 else np.array([0, 1.0, 0])
    s = np.cross(d, a); s /= np.linalg.norm(s); t = np.cross(d, s)
    ang = np.radians(max_deg) * np.sqrt(rng.random(n)); az = 2 * np.pi * rng.random(n)
    v = (np.cos(ang)[None] * d[:, None] + np.sin(ang)[None] * (np.cos(az)[None] * s[:, None] + np.sin(az)[None] * t[:, None]))
    return v.astype(np.float32)


def parity_inputs(sun_dir_world, spectral, seed=1234, n_rand=2048):
    """Deterministic inputs for a parity case: world directions `wo` (towards the sky), wavelengths, samples."""
    rng = np.random.default_rng(seed)
    special = np.array([[0, 0, 1], [1, 0, 0], [0, 1, 0], [0, 0, -1], [0.6, 0.8, 0.0], [-0.6, 0.0, 0.8],
                        [1e-4, 0, 1], [0, -1e-7, 1]], dtype=np.float32).T
    special = special / np.linalg.norm(special, axis=0, keepdims=True)
    wo = np.concatenate([
        -latlong_grid(32, 16),                       # eval(si.wi = -wo): both hemispheres
        hemisphere_dirs(n_rand, rng),
        around(sun_dir_world, 0.25, 256, rng),       # inside the default 0.2679 deg disc
        around(sun_dir_world, 1.0, 256, rng),        # straddling its edge
        special.astype(np.float32),
    ], axis=1).astype(np.float32)
    pad = (-wo.shape[1]) % 4 + 1                     # deliberately not a multiple of 4: tail path
    wo = np.concatenate([wo, hemisphere_dirs(pad, rng)], axis=1)
    n = wo.shape[1]
    wav = None
    if spectral:
        wav = np.minimum(360.0 + 360.0 * rng.random((4, n), dtype=np.float32), 719.99).astype(np.float32)
        wav[:, :12] = np.array([[320.0, 360.0, 719.999, 719.99, 300.0, 730.0, 400.0, 440.0, 480.0, 520.0, 680.0, 321.0]] * 4,
                               dtype=np.float32)
        wav[1, :4] = [359.9, 500.0, 640.0, 700.0]
    u = rng.random((2, n), dtype=np.float32)
    u[:, :6] = np.array([[0.0, 0.0], [0.999999, 0.999999], [0.5, 0.0], [0.0, 0.5], [1e-7, 0.3], [0.35, 0.999]],
                        dtype=np.float32).T
    return dict(wo=wo, wav=wav, u=u)


SCENE_BBOX = ([-1.0, -2.0, -3.0], [4.0, 3.0, 2.0])     # set_scene box of the sample_ray parity cases


def ray_inputs(n, seed):
    """Extra deterministic inputs of sample_wavelengths / sample_ray: the wavelength sample and sample2
    (sample3 reuses `u` of parity_inputs). Includes the edges 0, 1, the shift wrap points and the disc centre."""
    rng = np.random.default_rng(seed + 977)
    wl = rng.random(n, dtype=np.float32)
    # not 0.25 / 0.5 / 0.75 / 1.0 themselves: a shifted sample of exactly 1 maps to 720 nm, where the reference
    # gathers one past the end of its coefficient table (the oracle's arrays bounds-check)
    wl[:8] = np.array([0.0, 0.9999, 0.2499, 0.5001, 0.7501, 0.999, 1e-7, 0.2501], dtype=np.float32)
    s2 = rng.random((2, n), dtype=np.float32)
    s2[:, :5] = np.array([[0.5, 0.5], [0.0, 0.0], [1.0, 0.5], [0.5, 1.0], [0.25, 0.75]], dtype=np.float32).T
    return dict(wl_sample=wl, s2=s2)


def rel_err(a, b, floor=0.0):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floor if floor > 0 else np.finfo(np.float64).tiny)


def max_rel_err(a, b):
    """max |a-b| / |b| over entries where b != 0; entries where b == 0 must match exactly."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    nz = b != 0
    zero_ok = np.all(a[~nz] == 0) if np.any(~nz) else True
    both_nan = np.isnan(a) & np.isnan(b)
    e = np.abs(a[nz & ~both_nan] - b[nz & ~both_nan]) / np.abs(b[nz & ~both_nan])
    return (float(e.max()) if e.size else 0.0), bool(zero_ok)


def bit_equal(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32); b = np.ascontiguousarray(b, dtype=np.float32)
    return np.array_equal(a.view(np.uint32), b.view(np.uint32))


def ulp_diff(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, dtype=np.float32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7fffffff), a); b = np.where(b < 0, -(b & 0x7fffffff), b)
    return np.abs(a - b)
