#!/usr/bin/env python
"""Regenerates the eight sunsky dataset files from their sources:

  python tools/make_datasets.py --hosek-dir <dir with ArHosekSkyModelData_{Spectral,RGB}.h> \
                                --tgmm-csv <model_hosek.csv> --out <dir>

Counterpart of the reference's (commented-out) generators: write_sky_data, write_sun_data_spectral,
write_sun_data_rgb, write_limb_darkening_data (include/mitsuba/render/sunsky/sunsky.h:609-895) and
sunsky-testing/mk_sampling_dataset.py. The Hosek-Wilkie coefficient headers are third-party data
(https://cgg.mff.cuni.cz/projects/SkylightModelling/); they are parsed where they lie, never copied.

File format ("SKY"/"SUN", u32 version, u64 ndim, u64 shape[], row-major payload): sunsky.h:573-597."""
import argparse
import os
import re
import struct
from fractions import Fraction

import numpy as np

NB_WAVELENGTHS, NB_TURBIDITY, NB_ALBEDO, NB_CTRL, NB_PARAMS = 11, 10, 2, 6, 9
NB_SUN_SEGMENTS, NB_SUN_CTRL, NB_LD = 45, 4, 6
WAVELENGTHS = [320 + 40 * i for i in range(11)]

_NUM = re.compile(r"[-+]?(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?")


def parse_arrays(path):
    """{name: float64 array} for every `double name[] = { ... };` of a Hosek data header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    out = {}
    for m in re.finditer(r"double\s+(\w+)\s*\[\s*\]\s*=\s*\{(.*?)\}\s*;", text, flags=re.S):
        out[m.group(1)] = np.array([float(v) for v in _NUM.findall(m.group(2))], dtype=np.float64)
    return out


def write_file(path, tag, shape, data, dtype):
    data = np.ascontiguousarray(data, dtype=dtype)
    assert data.size == int(np.prod(shape)), (path, data.size, shape)
    with open(path, "wb") as f:
        f.write(tag.encode()); f.write(struct.pack("<I", 0)); f.write(struct.pack("<Q", len(shape)))
        for s in shape:
            f.write(struct.pack("<Q", s))
        f.write(data.tobytes())


def sky_data(per_channel, with_params):
    """write_sky_data (sunsky.h:820-876): [channel][albedo][turbidity][ctrl][param] -> [turbidity][albedo][ctrl][channel][param]"""
    n_p = NB_PARAMS if with_params else 1
    src = np.stack([a.reshape(NB_ALBEDO, NB_TURBIDITY, NB_CTRL, n_p) for a in per_channel])      # (C, a, t, k, p)
    dst = src.transpose(2, 1, 3, 0, 4)                                                            # (t, a, k, C, p)
    shape = (NB_TURBIDITY, NB_ALBEDO, NB_CTRL, len(per_channel)) + ((NB_PARAMS,) if with_params else ())
    return shape, dst.reshape(-1)


def sun_spectral(solar):
    """write_sun_data_spectral (sunsky.h:780-813): [lambda][turbidity][segment][ctrl reversed] -> [turbidity][segment][lambda][ctrl]"""
    src = np.stack([a.reshape(NB_TURBIDITY, NB_SUN_SEGMENTS, NB_SUN_CTRL) for a in solar])[..., ::-1]   # (l, t, s, k)
    return (NB_TURBIDITY, NB_SUN_SEGMENTS, NB_WAVELENGTHS, NB_SUN_CTRL), src.transpose(1, 2, 0, 3).reshape(-1)


# linear_rgb_rec(lambda) at the 11 dataset wavelengths (include/mitsuba/core/spectrum.h:332-365): CIE 1931 2-degree
# colour matching functions (5 nm table of src/core/spectrum.cpp; the wavelengths sit on table nodes, 320 nm lies
# outside [360, 830] and gives 0) pushed through xyz_to_srgb (spectrum.h:405-412), all in fp32 like the reference's
# `tables.srgb`.
CIE_XYZ = {
    360: (0.0001299, 0.000003917, 0.0006061), 400: (0.01431, 0.000396, 0.06785001),
    440: (0.34828, 0.023, 1.74706), 480: (0.09564, 0.13902, 0.8129501), 520: (0.06327, 0.71, 0.07824999),
    560: (0.5945, 0.995, 0.0039), 600: (1.0622, 0.631, 0.0008), 640: (0.4479, 0.175, 0.00002),
    680: (0.04677, 0.017, 0.0), 720: (0.002899327, 0.001047, 0.0),
}
XYZ_TO_SRGB = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556],
                        [0.055648, -0.204043, 1.057311]], dtype=np.float32)


def linear_rgb_rec(wavelength):
    if wavelength not in CIE_XYZ:
        return np.zeros(3, dtype=np.float64)
    xyz = np.array(CIE_XYZ[wavelength], dtype=np.float32)
    rgb = np.empty(3, dtype=np.float32)
    for r in range(3):       # Matrix * vector in Dr.Jit: column-wise fmadd chain (matrix.h) -- col0*x, then +col1*y, +col2*z
        acc = np.float32(XYZ_TO_SRGB[r, 0] * xyz[0])
        acc = np.float32(np.float64(XYZ_TO_SRGB[r, 1]) * np.float64(xyz[1]) + np.float64(acc))
        acc = np.float32(np.float64(XYZ_TO_SRGB[r, 2]) * np.float64(xyz[2]) + np.float64(acc))
        rgb[r] = acc
    return rgb.astype(np.float64)


def fma64(a, b, c):
    """Correctly rounded fp64 fused multiply-add (exact rational arithmetic, one rounding)."""
    return float(Fraction(a) * Fraction(b) + Fraction(c))


def sun_rgb(solar, ld):
    """write_sun_data_rgb (sunsky.h:713-772): buffer += rectifier * solar * limb-darkening over lambda, / 11, in
    fp64. The shipped file was produced by a build that contracts the accumulation into a hardware FMA
    (buffer = fma(rectifier * solar, ld, buffer), the x86-64 default with FMA enabled); reproducing that single
    rounding makes the output byte-identical."""
    src = np.stack([a.reshape(NB_TURBIDITY, NB_SUN_SEGMENTS, NB_SUN_CTRL) for a in solar])[..., ::-1]   # (l, t, s, k)
    rect = np.stack([linear_rgb_rec(w) for w in WAVELENGTHS])                                              # (l, 3)
    prod = rect[:, None, None, :, None] * src[:, :, :, None, :]                                            # (l, t, s, 3, k) fp64
    out = np.zeros((NB_TURBIDITY, NB_SUN_SEGMENTS, 3, NB_SUN_CTRL, NB_LD), dtype=np.float64)
    flat = out.reshape(-1, NB_LD)
    pf = prod.reshape(NB_WAVELENGTHS, -1)
    for i in range(flat.shape[0]):
        for j in range(NB_LD):
            acc = 0.0
            for l in range(NB_WAVELENGTHS):
                acc = fma64(pf[l, i], ld[l][j], acc) if pf[l, i] != 0.0 else acc + 0.0 * ld[l][j]
            flat[i, j] = acc / NB_WAVELENGTHS
    return (NB_TURBIDITY, NB_SUN_SEGMENTS, 3, NB_SUN_CTRL, NB_LD), out.reshape(-1)


def tgmm_tables(csv_path):
    """mk_sampling_dataset.py: drop RMSE / MAE / Volume / Normalization / Azimuth, sort by (turbidity, elevation),
    keep (mean x, mean y, sigma x, sigma y, weight), elevation mean -> zenith angle; fp32, shape (9, 30, 5, 5)."""
    with open(csv_path) as f:
        header = f.readline().strip().split(",")
        rows = np.array([[float(v) for v in line.split(",")] for line in f if line.strip()], dtype=np.float64)
    keep = [header.index(k) for k in ("Turbidity", "Elevation", "Mean X", "Mean Y", "Sigma X", "Sigma Y", "Weight")]
    arr = rows[:, keep]
    order = np.lexsort([arr[:, 1], arr[:, 0]])
    arr = arr[order, 2:]
    arr[:, 1] = np.pi / 2 - arr[:, 1]
    return (9, 30, 5, 5), arr.reshape(-1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hosek-dir", required=True)
    ap.add_argument("--tgmm-csv", required=True)
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    spec = parse_arrays(os.path.join(a.hosek_dir, "ArHosekSkyModelData_Spectral.h"))
    rgb = parse_arrays(os.path.join(a.hosek_dir, "ArHosekSkyModelData_RGB.h"))
    p = lambda name: os.path.join(a.out, name)
    write_file(p("sky_spec_params.bin"), "SKY", *sky_data([spec[f"dataset{w}"] for w in WAVELENGTHS], True), np.float64)
    write_file(p("sky_spec_rad.bin"), "SKY", *sky_data([spec[f"datasetRad{w}"] for w in WAVELENGTHS], False), np.float64)
    write_file(p("sky_rgb_params.bin"), "SKY", *sky_data([rgb[f"datasetRGB{c}"] for c in (1, 2, 3)], True), np.float64)
    write_file(p("sky_rgb_rad.bin"), "SKY", *sky_data([rgb[f"datasetRGBRad{c}"] for c in (1, 2, 3)], False), np.float64)
    solar = [spec[f"solarDataset{w}"] for w in WAVELENGTHS]
    ld = [spec[f"limbDarkeningDataset{w}"] for w in WAVELENGTHS]
    write_file(p("sun_spec_rad.bin"), "SUN", *sun_spectral(solar), np.float64)
    write_file(p("sun_rgb_rad.bin"), "SUN", *sun_rgb(solar, ld), np.float64)
    write_file(p("sun_spec_ld.bin"), "SUN", (NB_WAVELENGTHS, NB_LD), np.concatenate(ld), np.float64)
    write_file(p("tgmm_tables.bin"), "SKY", *tgmm_tables(a.tgmm_csv), np.float32)


if __name__ == "__main__":
    main()
