"""Generates tests/golden/*.npz from the reference's own code (oracle/_ref, built by
oracle/build_ref.sh from /root/reference). Run in a container that has /root/reference:

    oracle/build_ref.sh && python tests/gen_golden.py

Each file holds the inputs and the reference's llvm_ad-like ("array flavour") outputs for one emitter
configuration of tests/common.py, plus the interpolated emitter state. The scalar_* flavour outputs of
eval are stored as well (they differ: different erf branch / dot order / w = 0.5)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
import common  # noqa: E402
import ref_oracle as R  # noqa: E402

STATE_FIELDS = ["sky_params", "sky_radiance", "sun_radiance", "sun_ld", "gaussians", "pmf", "cdf", "cdf_sum",
                "cdf_normalization", "sky_sampling_w", "sun_angles", "frame_s", "frame_t", "frame_n",
                "sun_half_aperture", "sun_dir_world", "spectral_pdf", "spectral_cdf", "spectral_integral",
                "spectral_normalization", "n_sun_radiance", "n_channels"]


def state_dict(s):
    out = {}
    for f in STATE_FIELDS:
        v = getattr(s, f)
        out["state_" + f] = np.array(v, dtype=np.float32) if hasattr(v, "__len__") else np.float32(v)
    return out


def main():
    os.makedirs(common.GOLDEN_DIR, exist_ok=True)
    for name, (spectral, kw) in common.CONFIGS.items():
        e = R.RefEmitter(spectral=spectral, flavour=1, **kw)
        s = e.state()
        inp = common.parity_inputs(list(s.sun_dir_world), spectral, seed=abs(hash(name)) % 2**31 if False else sum(map(ord, name)))
        wo, wav, u = inp["wo"], inp["wav"], inp["u"]
        out = dict(wo=wo, u=u)
        if spectral:
            out["wav"] = wav
        out["eval"] = e.eval(-wo, wav)                      # eval(si.wi = -wo)
        out["pdf"] = e.pdf_direction(wo)
        d, pdf, w, p, dist = e.sample_direction(u, None, wav, want_p=True)
        out.update(s_d=d, s_pdf=pdf, s_weight=w, s_p=p, s_dist=dist)
        out["s_pdf_requery"] = e.pdf_direction(d)
        # sample_wavelengths(si.wi = -wo) and sample_ray (after set_scene with a non-trivial bounding sphere)
        seed = sum(map(ord, name))
        ri = common.ray_inputs(wo.shape[1], seed)
        out.update(wl_sample=ri["wl_sample"], s2=ri["s2"])
        sw_wav, sw_weight = e.sample_wavelengths(-wo, ri["wl_sample"])
        out.update(sw_wav=sw_wav, sw_weight=sw_weight)
        e.set_scene(*common.SCENE_BBOX)
        r_o, r_d, r_wav, r_w = e.sample_ray(ri["wl_sample"], ri["s2"], u)
        out.update(ray_o=r_o, ray_d=r_d, ray_wav=r_wav, ray_weight=r_w)
        out["state_bsphere"] = np.array(list(e.state().bsphere_center) + [e.state().bsphere_radius], dtype=np.float32)
        es = R.RefEmitter(spectral=spectral, flavour=0, **kw)
        out["eval_scalar"] = es.eval(-wo, wav)
        out.update(state_dict(s))
        np.savez_compressed(os.path.join(common.GOLDEN_DIR, name + ".npz"), **out)
        print(name, "n =", wo.shape[1], "w =", s.sky_sampling_w, "max eval", float(out["eval"].max()),
              "sun-lit lanes", int((out["eval"].max(axis=0) > 1e4).sum()))

    # Dr.Jit math on dense inputs (array branch), for the unit tests of the CUDA/C restatements
    rng = np.random.default_rng(7)
    m = {}
    x = np.concatenate([rng.uniform(-8, 8, 4000), np.linspace(-1, 1, 501), [0.0, -0.0, 1e-30, 100.0, -100.0]]).astype(np.float32)
    for op in ("sin", "cos", "exp", "exp2", "erf", "cbrt", "tan"):
        m[op + "_x"] = x; m[op] = R.math(op, x)
    xu = np.concatenate([rng.uniform(-1, 1, 4000), np.linspace(-1, 1, 501)]).astype(np.float32)
    for op in ("asin", "acos"):
        m[op + "_x"] = xu; m[op] = R.math(op, xu)
    xe = np.clip(xu, -0.9999999, 0.9999999).astype(np.float32)
    m["erfinv_x"] = xe; m["erfinv"] = R.math("erfinv", xe)
    xp = np.concatenate([rng.uniform(1e-6, 50, 4000), np.logspace(-30, 30, 200)]).astype(np.float32)
    for op in ("log", "log2", "pow15", "sqrt"):
        m[op + "_x"] = xp; m[op] = R.math(op, xp)
    ya = rng.uniform(-3, 3, 4000).astype(np.float32); xa = rng.uniform(-3, 3, 4000).astype(np.float32)
    m["atan2_y"] = ya; m["atan2_x"] = xa; m["atan2"] = R.math("atan2", ya, xa)
    gx, gw = R.gauss_legendre(200)
    m["gl200_x"] = gx; m["gl200_w"] = gw
    m["pcg32_first_4096"] = R.pcg32_floats(4096)
    rec = [(2010, 7, 10, 9.5, 0, 0, 35.6894, 139.6917, 9.0), (2024, 12, 21, 12.0, 30.0, 15.0, -33.9, 151.2, 11.0),
           (1999, 1, 1, 6.25, 0, 0, 64.1, -21.9, 0.0), (2010, 7, 10, 18.3, 0, 0, 35.6894, 139.6917, 9.0)]
    m["sun_coord_in"] = np.array(rec, dtype=np.float64)
    m["sun_coord_out"] = np.stack([R.sun_coordinates(int(r[0]), int(r[1]), int(r[2]), *[float(v) for v in r[3:]]) for r in rec])
    np.savez_compressed(os.path.join(common.GOLDEN_DIR, "drjit_math.npz"), **m)
    print("math golden written")


if __name__ == "__main__":
    main()
