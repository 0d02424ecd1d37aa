#!/usr/bin/env python
"""Launches each kernel of the sunsky hot path a few times on bench-sized inputs, for `ncu`:

  python profiles/prof_driver.py [--kernels eval_spec,eval_rgb,pdf,sample_rgb,sample_spec,envmap] [--reps 3] [--n-log2 24]

Same emitters and input distributions as bench.py (configs[1] for eval/pdf, configs[2] for sampling).
It also prints CUDA-event times per kernel, so the plain run doubles as a quick timing check."""
import argparse
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(REPO, "mitsuba3-sunsky_b200"), os.path.join(REPO, "tests")):
    sys.path.insert(0, p)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kernels", default="eval_spec,eval_rgb,pdf,sample_rgb")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--n-log2", type=int, default=24)
    ap.add_argument("--arith", default="exact")
    args = ap.parse_args()
    import torch
    import common
    import sunsky_b200 as ssk
    dev = torch.device("cuda", 0)
    n = 1 << args.n_log2
    gen = torch.Generator(device=dev); gen.manual_seed(1234)
    u1 = torch.rand(n, device=dev, generator=gen); u2 = torch.rand(n, device=dev, generator=gen)
    st = torch.sqrt(torch.clamp(1 - u1 * u1, min=0)); ph = 2 * np.pi * u2
    wo = torch.stack([st * torch.cos(ph), st * torch.sin(ph), u1]).contiguous()
    wi = (-wo).contiguous()
    wav = (360.0 + 360.0 * torch.rand((4, n), device=dev, generator=gen)).clamp_(max=719.99).contiguous()
    u = torch.rand((2, n), device=dev, generator=gen)
    extra = {"arith": args.arith}
    kw = dict(turbidity=3.0, albedo=0.3, sun_direction=common.sun_dir_from_elevation(30.0), **extra)
    kw3 = dict(turbidity=6.0, albedo=0.5, sun_direction=common.sun_dir_from_elevation(50.2), **extra)
    es, er = ssk.SunskyEmitter(spectral=True, **kw), ssk.SunskyEmitter(spectral=False, **kw)
    ss, sr = ssk.SunskyEmitter(spectral=True, **kw3), ssk.SunskyEmitter(spectral=False, **kw3)
    sky = ssk.SunskyEmitter(spectral=False, sun_scale=0.0, **kw3)
    sky_s = ssk.SunskyEmitter(spectral=True, sun_scale=0.0, **kw3)
    runs = {
        "eval_spec": lambda: es.eval(wi, wav),
        "eval_rgb": lambda: er.eval(wi),
        "pdf": lambda: er.pdf_direction(wo),
        "sample_rgb": lambda: sr.sample_direction(u),
        "sample_sky": lambda: sky.sample_direction(u),
        "sample_spec": lambda: ss.sample_direction(u, None, wav),
        "sample_spec_sky": lambda: sky_s.sample_direction(u, None, wav),
        "envmap": lambda: ssk.envmaps([er], 4096, 2048),
        "envmap_spec": lambda: ssk.envmaps([es], 4096, 2048, (440.0, 520.0, 600.0, 680.0)),
        "wavelengths": lambda: es.sample_wavelengths(wi, u2),
        "ray_rgb": lambda: sr.sample_ray(None, u, u),
        "ray_spec": lambda: ss.sample_ray(u1, u, u),
        "jvp_rgb": lambda: er.eval_jvp(wi, None, d_turbidity=1.0, d_albedo=0.5),
        "jvp_spec": lambda: es.eval_jvp(wi, wav, d_turbidity=1.0, d_albedo=0.5),
    }
    if "setup" in args.kernels.split(","):
        import time
        for label, e in (("setup_rgb", er), ("setup_spec", es)):
            e.update(turbidity=3.1); torch.cuda.synchronize()
            t0 = time.perf_counter()
            n_up = 50 if args.reps >= 10 else args.reps
            for k in range(n_up):
                e.update(turbidity=3.0 + 0.01 * k)
            dt = (time.perf_counter() - t0) / n_up
            e.update(turbidity=3.0)
            print(f"{label}: {dt * 1e3:.3f} ms per parameters_changed (host interpolation + device quadrature + block upload)")
    for name in [k for k in args.kernels.split(",") if k != "setup"]:
        fn = runs[name]
        fn(); torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / max(args.reps, 1)
        units = n if not name.startswith("envmap") else 4096 * 2048
        print(f"{name}: {ms:.4f} ms per launch incl. output allocation, {units / ms / 1e6:.2f} G/s")


if __name__ == "__main__":
    main()
