"""The oracle against the reference's own golden vectors (`-m "not gpu"`): the sky renders and the 80 sun spectra
under resources/sunsky/test_data, at the tolerances of src/emitters/tests/test_sunsky.py.
Also keeps the packed fixture honest: when the reference tree is present it is re-read and compared."""
import os

import numpy as np
import pytest

import reference_suite as rs


@pytest.fixture(scope="module")
def refdata():
    return rs.load_refdata()


def maker(ref, flavour):
    return lambda spectral, **kw: ref.RefEmitter(spectral=spectral, flavour=flavour, **kw)


@pytest.mark.parametrize("flavour", [0, 1])
def test_oracle_reproduces_reference_sky_renders(ref, refdata, flavour):
    for p in rs.RGB_CASES:                                              # test01, <= 0.017
        assert rs.render_error(maker(ref, flavour), refdata, False, p, rs.rgb_key(p)) <= 0.017
    for p in rs.SPEC_CASES:                                             # test02, <= 0.037
        assert rs.render_error(maker(ref, flavour), refdata, True, p, rs.spec_key(p)) <= 0.037
    assert rs.render_error(maker(ref, flavour), refdata, True, (np.radians(60), 4.2, rs.SPECIAL_ALBEDO),
                           "sky_spectrum_special") <= 0.03             # test03


def test_oracle_reproduces_reference_sun_spectra(ref, refdata):
    err = rs.sun_spectrum_errors(maker(ref, 1), refdata)                # test04, <= 1e-2 each
    assert err.max() <= 1e-2
    assert err.max() <= 2e-3                                            # measured 1.63e-3 (fp32 sun polynomial)


def test_packed_fixture_matches_the_reference_tree(refdata):
    data = "/root/reference/resources/sunsky/test_data"
    if not os.path.isdir(data):
        pytest.skip("reference tree not present")
    import exr_reader
    import glob
    files = sorted(glob.glob(os.path.join(data, "renders", "*.exr")))
    assert len(files) == 7 and len(glob.glob(os.path.join(data, "spectrum", "*.spd"))) == 80
    for f in files:
        ch, names = exr_reader.read_exr(f)
        order = ["R", "G", "B"] if set(names) == {"R", "G", "B"} else names
        img = np.stack([ch[c] for c in order], axis=-1)
        assert np.array_equal(img, refdata["render/" + os.path.splitext(os.path.basename(f))[0]])
    t, eta, g = 5.5, np.linspace(1e-4, np.pi / 2 - 1e-4, 4)[1], np.linspace(0, rs.SUN_HALF_APERTURE_ANGLE - 1e-4, 4)[2]
    rows = open(os.path.join(data, "spectrum", f"sun_spectrum_t{t:.1f}_eta{eta:.2f}_gamma{g:.3e}.spd")).read().split()
    assert np.array_equal(np.array(rows[1::2], float), refdata["sun_spectra"][2, 1, 2])


def test_pcg32_streams_match_the_reference():
    """sunsky_b200.pcg32 (input generation for configs[1..2], test05) against dr::PCG32 (drjit/random.h)."""
    import ref_oracle
    if not ref_oracle.available():
        pytest.skip("oracle/_ref not built")
    from sunsky_b200 import pcg32
    assert np.array_equal(pcg32.floats(50_000)[0].numpy(), ref_oracle.pcg32_floats(50_000))
    assert np.array_equal(pcg32.floats(999, 1, 7_000_000_000)[0].numpy(), ref_oracle.pcg32_floats(999, 7_000_000_000))
    pairs = pcg32.point2_pairs(1000).numpy()
    flat = ref_oracle.pcg32_floats(2000)
    assert np.array_equal(pairs[0], flat[0::2]) and np.array_equal(pairs[1], flat[1::2])


@pytest.mark.parametrize("name", ["cfg1_rgb", "cfg2_spec", "cfg3_rgb_sunsky", "spec_alb11"])
def test_scalar_and_array_flavours_agree(ref, name):
    """Both arithmetic flavours of the oracle are the same formulas (they differ in erf's coefficient set, the
    dot-product order and FMA use: SURVEY.md section 0.4), so they agree to a few ulp on well-conditioned lanes. Guards
    the build recipe: without -fno-strict-aliasing (which Dr.Jit requires, ext/drjit/CMakeLists.txt:71) g++ -O2
    miscompiled the scalar instantiation -- its sky term came out as 0 -- and round 1 timed that as the CPU arm."""
    import common
    import emitters
    g = emitters.golden(name)
    spectral, kw = common.CONFIGS[name]
    wav = g["wav"] if spectral else None
    a = ref.RefEmitter(spectral=spectral, flavour=0, **kw).eval(-g["wo"], wav)
    b = g["eval"]
    assert np.array_equal(a == 0, b == 0)                               # same lanes lit
    nz = b != 0
    rel = np.abs(a[nz] - b[nz]) / np.abs(b[nz])
    assert np.median(rel) < 1e-6 and np.quantile(rel, 0.99) < 1e-4, (np.median(rel), rel.max())
    assert np.array_equal(a, g["eval_scalar"])


@pytest.mark.parametrize("isa", ["avx2", "best"])
@pytest.mark.parametrize("name", ["cfg1_rgb", "cfg2_spec", "cfg3_rgb_sunsky", "cfg3_spec_sunsky"])
def test_packet_flavour_tracks_the_parity_flavour(ref, name, isa):
    """Flavour 2 (the reference's query members on dr::Packet<float, 8 | 16>, the timed CPU baseline) against
    flavour 1: sampled directions and pdfs are bit-identical; radiance differs by the packet `rcp` (hardware
    estimate + one Newton step instead of 1 / x) and stays within 4e-6 relative."""
    import common
    import emitters
    g = emitters.golden(name)
    spectral, kw = common.CONFIGS[name]
    wav = g["wav"] if spectral else None
    r = ref.RefEmitter(spectral=spectral, flavour=2, isa=isa, **kw)
    a, b = r.eval(-g["wo"], wav), g["eval"]
    assert np.array_equal(a == 0, b == 0)
    nz = b != 0
    assert (np.abs(a[nz] - b[nz]) / np.abs(b[nz])).max() < 4e-6
    assert common.bit_equal(r.pdf_direction(g["wo"]), g["pdf"])
    d, pdf, w = r.sample_direction(g["u"], None, wav)
    assert common.bit_equal(d, g["s_d"]) and common.bit_equal(pdf, g["s_pdf"])
    # threaded driver = single-threaded driver
    assert common.bit_equal(r.eval(-g["wo"], wav, threads=3), a)


def test_hosek_sun_rad_matches_the_fp64_model_and_the_spd_fixtures(ssk, ref, refdata):
    """ssk_hosek_sun_rad (the library's mi.hosek_sun_rad) against the reference's ArHosekSkyModel.c compiled in place
    (second oracle, fp64) over a (turbidity, wavelength, elevation, gamma) grid, and against the 80 shipped .spd
    spectra, which are output of that model (fp64 restatement within 2.4e-8)."""
    if not ref.hosek_available():
        pytest.skip("oracle/_ref/libhosek_ref.so not built")
    rng = np.random.default_rng(3)
    worst = 0.0
    for _ in range(4000):
        t, wl = rng.uniform(1, 10), rng.uniform(315, 725)
        el, g = rng.uniform(1e-4, np.pi / 2), rng.uniform(0, 0.006)
        a, b = ssk.hosek_sun_rad(t, wl, el, g), ref.hosek_sun_rad(t, wl, el, g)
        assert (a == 0) == (b == 0)
        if b != 0:
            worst = max(worst, abs(a - b) / abs(b))
    assert worst < 1e-12, worst
    for t, wl in ((10.0, 500.0), (3.0, 720.0), (1.0, 320.0), (0.5, 500.0), (3.0, 319.0)):        # edges and out of range
        assert ssk.hosek_sun_rad(t, wl, 0.7, 0.001) == pytest.approx(ref.hosek_sun_rad(t, wl, 0.7, 0.001), rel=1e-12, abs=0)
    wavelengths = refdata["sun_wavelengths"]
    turbs = np.linspace(1, 10, 5); etas = np.linspace(1e-4, np.pi / 2 - 1e-4, 4)
    gammas = np.linspace(0, rs.SUN_HALF_APERTURE_ANGLE - 1e-4, 4)
    worst = 0.0
    for i, t in enumerate(turbs):
        for j, eta in enumerate(etas):
            for k, g in enumerate(gammas):
                got = np.array([ssk.hosek_sun_rad(t, wl, eta, g) for wl in wavelengths])
                want = refdata["sun_spectra"][i, j, k]
                assert np.array_equal(got == 0, want == 0)
                nz = want != 0
                if nz.any():
                    worst = max(worst, float((np.abs(got[nz] - want[nz]) / want[nz]).max()))
    assert worst < 1e-6, worst          # the .spd files carry 8 significant digits
