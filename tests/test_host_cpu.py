"""CPU-only tests (`-m "not gpu"`): the C ABI loads and exports every declared symbol, host-side setup
(dataset loading, coefficient interpolation, TGMM mixture, sun position, sampling weight) reproduces
the reference's emitter state bit for bit, property validation mirrors the plugin, and the compute
entry points fail loudly without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import common
import emitters

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_capi_exports_every_declared_symbol(ssk):
    from sunsky_b200 import _capi
    hdr = open(os.path.join(REPO, "include", "sunsky_b200.h")).read()
    declared = set(re.findall(r"\b(ssk_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    L = _capi.lib()
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in include/sunsky_b200.h but not exported"
    assert declared == set(_capi.SYMBOLS), (declared ^ set(_capi.SYMBOLS))
    assert b"sm_100a" in L.ssk_version()


STATE_EXACT = ["sky_params", "sky_radiance", "sun_radiance", "sun_ld", "gaussians", "pmf", "cdf", "cdf_sum",
               "cdf_normalization", "sky_sampling_w", "sun_angles", "frame_s", "frame_t", "frame_n",
               "sun_half_aperture"]


@pytest.mark.parametrize("name", list(common.CONFIGS))
def test_state_matches_reference_bitwise(ssk, name):
    """Every interpolated coefficient, the TGMM mixture + CDF, the sun frame and the quadrature-based
    sampling weight are bit-identical to the reference's private members (golden from oracle/_ref)."""
    g = emitters.golden(name)
    e = emitters.make(ssk, name, device=-1, g=g)
    s = e.state()
    nc = s.n_channels
    assert nc == int(g["state_n_channels"])
    sizes = dict(sky_params=nc * 9, sky_radiance=nc, sun_radiance=s.n_sun_radiance, sun_ld=66 if nc == 11 else 0)
    for f in STATE_EXACT:
        mine = np.atleast_1d(np.array(getattr(s, f), dtype=np.float32))
        ref = np.atleast_1d(g["state_" + f])
        k = sizes.get(f, min(mine.size, ref.size))
        assert common.bit_equal(mine[:k], ref[:k]), f"{name}: {f} differs (max abs {np.abs(mine[:k] - ref[:k]).max()})"


@pytest.mark.parametrize("name", [n for n, (sp, _) in common.CONFIGS.items() if sp])
def test_wavelength_distribution_matches_reference_bitwise(ssk, name):
    """m_spectral_distr (sunsky.cpp:869-881): node densities, cumulative trapezoid sums, integral, normalisation."""
    g = emitters.golden(name)
    s = emitters.make(ssk, name, device=-1, g=g).state()
    assert common.bit_equal(np.array(s.spectral_pdf[:10], np.float32), g["state_spectral_pdf"][:10])
    assert common.bit_equal(np.array(s.spectral_cdf[:9], np.float32), g["state_spectral_cdf"][:9])
    assert common.bit_equal(np.float32(s.spectral_integral), g["state_spectral_integral"])
    assert common.bit_equal(np.float32(s.spectral_normalization), g["state_spectral_normalization"])


def test_state_against_live_reference(ssk, ref):
    """Same check against the reference library itself on configurations that are not in the fixtures."""
    rng = np.random.default_rng(5)
    for spectral in (False, True):
        for _ in range(4):
            t = float(rng.uniform(1, 10)); a = float(rng.uniform(0, 1))
            sd = common.sun_dir_from_elevation(float(rng.uniform(0.5, 89.5)), float(rng.uniform(-3, 3)))
            scale = float(rng.uniform(0.1, 30.0))                 # the property need not be unit length
            sd = [float(np.float32(v) * np.float32(scale)) for v in sd]
            r = ref.RefEmitter(spectral=spectral, flavour=1, turbidity=t, albedo=a, sun_direction=sd)
            rs = r.state()
            e = ssk.SunskyEmitter(spectral=spectral, device=-1, turbidity=t, albedo=a, sun_direction=sd)
            s = e.state()
            for f in ("sun_dir_world", "gaussians", "cdf", "sun_angles", "frame_s", "frame_t"):
                assert common.bit_equal(np.array(getattr(s, f)), np.array(getattr(rs, f))), f
            assert common.bit_equal(np.array(s.sky_params[:s.n_channels * 9]), np.array(rs.sky_params[:s.n_channels * 9]))
            assert s.sky_sampling_w == rs.sky_sampling_w


def test_sun_coordinates_match_reference(ssk):
    g = emitters.golden("drjit_math")
    for rec, want in zip(g["sun_coord_in"], g["sun_coord_out"]):
        got = ssk.sun_coordinates(int(rec[0]), int(rec[1]), int(rec[2]), *[float(v) for v in rec[3:]])
        assert common.bit_equal(got, want), (rec, got, want)


def test_default_record_and_traverse_surface(ssk):
    e = ssk.SunskyEmitter(device=-1)
    names = list(e.traverse())
    # traverse() order of the reference (sunsky.cpp:220-240)
    assert names == ["turbidity", "sky_scale", "sun_scale", "albedo", "latitude", "longitude", "timezone", "year",
                     "day", "month", "hour", "minute", "second", "to_world"]
    e2 = ssk.SunskyEmitter(device=-1, sun_direction=[0, 0, 1])
    assert list(e2.traverse()) == ["turbidity", "sky_scale", "sun_scale", "albedo", "sun_direction", "to_world"]
    assert "SunskyEmitter[" in e.to_string() and "turbidity = 3" in e.to_string()
    assert abs(e.state().sun_half_aperture - np.radians(0.5358 / 2)) < 1e-9


def test_traverse_names_match_reference(ssk, ref):
    r = ref.RefEmitter(spectral=False, flavour=0, turbidity=3.0, albedo=0.3, sun_direction=[0, 0, 1])
    assert r.traverse_names() == list(ssk.SunskyEmitter(device=-1, sun_direction=[0, 0, 1]).traverse())
    r = ref.RefEmitter(spectral=False, flavour=0, turbidity=3.0, albedo=0.3, hour=12.0)
    assert r.traverse_names() == list(ssk.SunskyEmitter(device=-1, hour=12.0).traverse())


@pytest.mark.parametrize("bad", [dict(turbidity=0.5), dict(turbidity=10.5), dict(sun_scale=-1.0), dict(sky_scale=-0.1),
                                 dict(albedo=1.5), dict(albedo=-0.1), dict(sun_aperture=0.0), dict(sun_aperture=180.0),
                                 dict(sun_direction=[0, 0, 1], hour=3.0)])
def test_invalid_properties_are_rejected(ssk, bad):
    """Same validation as init_from_props / parameters_changed (sunsky.cpp:243-252, 889-948, 974)."""
    with pytest.raises(ssk.SunskyError):
        ssk.SunskyEmitter(device=-1, **bad)


def test_invalid_properties_match_reference(ssk, ref):
    for bad in (dict(turbidity=0.5), dict(turbidity=11.0), dict(sun_scale=-1.0), dict(albedo=1.5), dict(sun_aperture=190.0)):
        kw = dict(turbidity=3.0, albedo=0.3, sun_direction=[0, 0, 1]); kw.update(bad)
        with pytest.raises(RuntimeError):
            ref.RefEmitter(spectral=False, flavour=0, **kw)
        with pytest.raises(ssk.SunskyError):
            ssk.SunskyEmitter(device=-1, **kw)


def test_to_world_inverse_and_normalisation_match_reference(ssk, ref):
    """Host state for rotated / scaled / sheared / translated to_world matrices and non-unit sun directions:
    the inverse (fp64 cofactor expansion in the order of drjit/matrix.h:440-500, rounded to fp32 as
    Properties -> ScalarTransform4f does) and dr::normalize(ScalarVector3f) are bit-identical to the reference's,
    and so is everything derived from the local sun direction."""
    rng = np.random.default_rng(17)
    for it in range(40):
        M = np.eye(4, dtype=np.float64)
        if it % 2 == 0:                       # rotation x uniform scale + translation
            q = rng.normal(size=4); q /= np.linalg.norm(q); w, x, y, z = q
            R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                          [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                          [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
            M[:3, :3] = R * rng.uniform(0.3, 3.0)
        else:                                 # general affine
            M[:3, :3] = rng.normal(size=(3, 3))
        M[:3, 3] = rng.normal(size=3) * 5
        M = M.astype(np.float32)
        sd = (rng.normal(size=3) * 10 ** rng.uniform(-2, 2)).astype(np.float32)
        kw = dict(turbidity=3.5, albedo=0.4, sun_direction=list(sd), to_world=M)
        try:
            r = ref.RefEmitter(spectral=False, flavour=1, **kw)
        except RuntimeError as err:           # sheared to_world: local sun direction not unit length -> NaN weights
            assert "non-negative" in str(err)
            with pytest.raises(ssk.SunskyError, match="non-negative"):
                ssk.SunskyEmitter(spectral=False, device=-1, **kw)
            continue
        e = ssk.SunskyEmitter(spectral=False, device=-1, **kw)
        rs, s = r.state(), e.state()
        assert common.bit_equal(np.array(s.to_local).reshape(3, 3), ref.invert4(M)[:3, :3])
        assert common.bit_equal(np.array(s.to_local).reshape(3, 3), np.array(rs.to_local).reshape(4, 4)[:3, :3])
        assert common.bit_equal(np.array(s.sun_dir_world), ref.normalize3(sd))
        for f in ("sun_dir_world", "sun_angles", "frame_s", "frame_t", "frame_n", "gaussians", "cdf"):
            assert common.bit_equal(np.array(getattr(s, f)), np.array(getattr(rs, f))), (it, f)
        assert common.bit_equal(np.array(s.sky_params[:27]), np.array(rs.sky_params[:27]))


def test_update_is_parameters_changed(ssk):
    g = emitters.golden("cfg3_rgb_sunsky")
    e = emitters.make(ssk, "cfg1_rgb", device=-1, g=emitters.golden("cfg1_rgb"))
    e.update(turbidity=6.0, albedo=0.5, sun_direction=[float(v) for v in g["state_sun_dir_world"]])
    s = e.state()
    assert common.bit_equal(np.array(s.sky_params[:27]), g["state_sky_params"][:27])
    assert common.bit_equal(np.array(s.gaussians), g["state_gaussians"])
    assert s.sky_sampling_w == float(g["state_sky_sampling_w"])
    with pytest.raises(ssk.SunskyError):
        e.update(turbidity=42.0)
    assert e.state().sky_sampling_w == float(g["state_sky_sampling_w"])      # failed update leaves state intact
    with pytest.raises(ssk.SunskyError):
        e.update(hour=3.0)                                                    # cannot switch parametrisation


def test_missing_dataset_is_an_io_error(ssk, tmp_path):
    with pytest.raises(ssk.SunskyError) as ei:
        ssk.SunskyEmitter(device=-1, dataset_dir=str(tmp_path))
    assert ei.value.code == -2 and "does not exist" in str(ei.value)
    bad = tmp_path / "d"; bad.mkdir()
    for f in os.listdir(ssk.DATASET_DIR):
        (bad / f).write_bytes(b"XYZ" + open(os.path.join(ssk.DATASET_DIR, f), "rb").read()[3:])
    with pytest.raises(ssk.SunskyError):
        ssk.SunskyEmitter(device=-1, dataset_dir=str(bad))


def test_dataset_file_roundtrip(ssk, tmp_path):
    """array_from_file / array_to_file (sunsky.h:515-597): shapes of the shipped datasets, write/read."""
    a = ssk.array_from_file(os.path.join(ssk.DATASET_DIR, "sky_rgb_params.bin"), True)
    assert a.shape == (10, 2, 6, 3, 9)
    t = ssk.array_from_file(os.path.join(ssk.DATASET_DIR, "tgmm_tables.bin"), False)
    assert t.shape == (9, 30, 5, 5) and abs(t[..., 4].sum(axis=-1) - 1).max() < 1e-4
    p = str(tmp_path / "x.bin")
    ssk.array_to_file(p, t, as_f64=False)
    assert common.bit_equal(ssk.array_from_file(p, False), t)
    ssk.array_to_file(p, a, as_f64=True)
    assert common.bit_equal(ssk.array_from_file(p, True), a)


def test_no_cpu_fallback(ssk):
    """Host-only contexts answer state queries but every compute entry point refuses to run."""
    e = ssk.SunskyEmitter(device=-1, sun_direction=[0, 0, 1])
    wi = np.zeros((3, 8), np.float32); wi[2] = -1
    for call in (lambda: e.eval(wi), lambda: e.pdf_direction(wi), lambda: e.sample_direction(np.zeros((2, 8), np.float32)),
                 lambda: e.eval_direction(wi)):
        with pytest.raises(ssk.SunskyError) as ei:
            call()
        assert ei.value.code == -3 and "no CPU fallback" in str(ei.value)


def test_cuda_device_missing_is_loud(ssk):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(ssk.SunskyError) as ei:
        ssk.SunskyEmitter(device=0, sun_direction=[0, 0, 1])
    assert ei.value.code == -3


def test_reference_scalar_and_array_flavours_differ(ref):
    """Documents SURVEY fact 4: scalar_* and llvm_ad_*-like arithmetic are different functions
    (sampling weight 0.5 vs quadrature; erf branch), which is why parity targets the array flavour."""
    kw = dict(turbidity=3.0, albedo=0.3, sun_direction=common.sun_dir_from_elevation(30.0))
    a = ref.RefEmitter(spectral=False, flavour=1, **kw).state().sky_sampling_w
    s = ref.RefEmitter(spectral=False, flavour=0, **kw).state().sky_sampling_w
    assert s == 0.5 and abs(a - 0.352) < 1e-3
    x = np.linspace(-3, 3, 2001).astype(np.float32)
    assert not common.bit_equal(ref.math("erf", x, flavour=0), ref.math("erf", x, flavour=1))


def test_golden_reproducible_from_reference(ref):
    """The committed fixtures are what the reference library produces here (guards against stale goldens)."""
    for name in ("cfg1_rgb", "cfg2_spec", "rgb_record_h9p5"):
        spectral, kw = common.CONFIGS[name]
        g = emitters.golden(name)
        r = ref.RefEmitter(spectral=spectral, flavour=1, **kw)
        wav = g["wav"] if spectral else None
        assert common.bit_equal(r.eval(-g["wo"], wav), g["eval"])
        assert common.bit_equal(r.pdf_direction(g["wo"]), g["pdf"])
        d, pdf, w = r.sample_direction(g["u"], None, wav)
        assert common.bit_equal(d, g["s_d"]) and common.bit_equal(pdf, g["s_pdf"]) and common.bit_equal(w, g["s_weight"])
        sw_wav, sw_w = r.sample_wavelengths(-g["wo"], g["wl_sample"])
        assert common.bit_equal(sw_w, g["sw_weight"]) and (not spectral or common.bit_equal(sw_wav, g["sw_wav"]))
        r.set_scene(*common.SCENE_BBOX)
        ro, rd, rwav, rw = r.sample_ray(g["wl_sample"], g["s2"], g["u"])
        assert common.bit_equal(ro, g["ray_o"]) and common.bit_equal(rd, g["ray_d"]) and common.bit_equal(rw, g["ray_weight"])


def test_adapter_plugin_type_checks():
    """adapter/sunsky_b200.cpp (the Mitsuba-side binding of INTEGRATION.md) instantiates for the reference's array
    types against the reference's own headers: every C-ABI call it makes matches include/sunsky_b200.h."""
    import os
    import shutil
    import subprocess
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref = os.environ.get("SSK_REFERENCE", "/root/reference")
    gen = os.path.join(repo, "oracle", "_ref", "gen")
    if not os.path.isdir(os.path.join(ref, "include", "mitsuba")) or not os.path.isdir(os.path.join(gen, "include")):
        pytest.skip("needs the reference tree and oracle/_ref/gen (container only)")
    cxx = shutil.which("g++")
    inc = [os.path.join(repo, "include"), os.path.join(repo, "oracle", "ref_shadow"), gen, os.path.join(gen, "include"), ref,
           os.path.join(ref, "include"), os.path.join(ref, "ext/drjit/include"),
           os.path.join(ref, "ext/drjit/ext/drjit-core/include"),
           os.path.join(ref, "ext/drjit/ext/drjit-core/ext/nanothread/include"), os.path.join(ref, "ext/tinyformat"),
           os.path.join(ref, "ext/fastfloat/include"), os.path.join(ref, "ext/nanobind/include")]
    cmd = [cxx, "-std=c++17", "-fsyntax-only", "-w", "-mavx2", "-mfma"] + [f"-I{d}" for d in inc] + \
          [os.path.join(repo, "tests", "adapter", "instantiate.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


def test_dataset_generators_reproduce_shipped_files(tmp_path):
    """tools/make_datasets.py (the reference's commented-out write_* generators, sunsky.h:609-895, and
    mk_sampling_dataset.py) rebuilds all eight dataset files byte for byte from the Hosek-Wilkie headers and the
    TGMM fit table."""
    import filecmp
    import os
    import subprocess
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref = os.environ.get("SSK_REFERENCE", "/root/reference")
    hosek = os.path.join(ref, "include", "mitsuba", "render", "sunsky")
    csv = os.path.join(ref, "sunsky-testing", "res", "datasets", "model_hosek.csv")
    if not (os.path.exists(os.path.join(hosek, "ArHosekSkyModelData_Spectral.h")) and os.path.exists(csv)):
        pytest.skip("needs the Hosek-Wilkie data headers and model_hosek.csv (container only)")
    subprocess.run([sys.executable, os.path.join(repo, "tools", "make_datasets.py"), "--hosek-dir", hosek, "--tgmm-csv", csv,
                    "--out", str(tmp_path)], check=True)
    shipped = os.path.join(repo, "mitsuba3-sunsky_b200", "data", "resources", "sunsky", "datasets")
    names = sorted(os.listdir(shipped))
    assert len(names) == 8
    for n in names:
        assert filecmp.cmp(os.path.join(shipped, n), os.path.join(str(tmp_path), n), shallow=False), f"{n} differs"


def test_update_batch_on_device_less_contexts(ssk):
    """ssk_update_batch on contexts without a device runs the host setup per emitter; errors leave the failing
    emitter untouched and are reported like ssk_update's."""
    es = [ssk.SunskyEmitter(spectral=False, device=-1, turbidity=3.0, albedo=0.3, sun_direction=[0.3, 0.2, 0.6]) for _ in range(3)]
    ref = ssk.SunskyEmitter(spectral=False, device=-1, turbidity=4.5, albedo=0.1, sun_direction=[0.3, 0.2, 0.6])
    ssk.update_batch(es, [dict(turbidity=4.5, albedo=0.1)] * 3)
    for e in es:
        assert e.state().sky_sampling_w == ref.state().sky_sampling_w
        assert common.bit_equal(np.array(e.state().sky_params, np.float32), np.array(ref.state().sky_params, np.float32))
    with pytest.raises(ssk.SunskyError):
        ssk.update_batch(es, [dict(turbidity=0.5)] * 3)
    with pytest.raises(ValueError):
        ssk.update_batch(es, [dict(turbidity=2.0)])
