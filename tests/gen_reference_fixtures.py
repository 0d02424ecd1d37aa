"""Packs the reference's own golden data for the sunsky path (/root/reference/resources/sunsky/test_data:
7 EXR sky renders, 80 .spd sun spectra; used by src/emitters/tests/test_sunsky.py:115-196) into
tests/golden/reference_test_data.npz, so that the restated reference tests (tests/test_reference_suite_gpu.py)
can run on the GPU box, where /root/reference does not exist.

Run where the reference tree is present:  python tests/gen_reference_fixtures.py
Keys: render/<file stem> -> float32 (32, 64, C) in the channel order mi.Bitmap hands to TensorXf (R, G, B for the
RGB files; ascending wavelength band for the 10-band files); sun_spectra -> float64 (5 turbidities, 4 ray
elevations, 4 gammas, 15 wavelengths), sun_wavelengths -> the 15 wavelengths of column 0."""
import glob
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import exr_reader  # noqa: E402

REF = os.environ.get("SSK_REFERENCE", "/root/reference")
DATA = os.path.join(REF, "resources", "sunsky", "test_data")
EPS = 1e-4
SUN_HALF_APERTURE_ANGLE = np.radians(0.5388 / 2.0)          # test_sunsky.py:10


def spd_grid():
    """Parametrisation of test04_sun_radiance (test_sunsky.py:154-156)."""
    return (np.linspace(1, 10, 5), np.linspace(EPS, np.pi / 2 - EPS, 4),
            np.linspace(0, SUN_HALF_APERTURE_ANGLE - EPS, 4))


def main():
    out = {}
    for f in sorted(glob.glob(os.path.join(DATA, "renders", "*.exr"))):
        ch, names = exr_reader.read_exr(f)
        order = ["R", "G", "B"] if set(names) == {"R", "G", "B"} else names
        out["render/" + os.path.splitext(os.path.basename(f))[0]] = np.stack([ch[c] for c in order], axis=-1)
    turbs, etas, gammas = spd_grid()
    spectra = np.zeros((5, 4, 4, 15), np.float64)
    wavelengths = None
    for i, t in enumerate(turbs):
        for j, eta in enumerate(etas):
            for k, gamma in enumerate(gammas):
                path = os.path.join(DATA, "spectrum", f"sun_spectrum_t{t:.1f}_eta{eta:.2f}_gamma{gamma:.3e}.spd")
                rows = [line.split() for line in open(path).read().splitlines() if line.strip()]
                spectra[i, j, k] = [float(r[1]) for r in rows]
                wl = np.array([float(r[0]) for r in rows])
                assert wavelengths is None or np.array_equal(wl, wavelengths)
                wavelengths = wl
    out["sun_spectra"] = spectra
    out["sun_wavelengths"] = wavelengths
    dst = os.path.join(HERE, "golden", "reference_test_data.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
