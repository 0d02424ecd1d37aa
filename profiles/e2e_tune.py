#!/usr/bin/env python
"""End-to-end rate of ssk_eval_host (pinned host planes in and out, spectral, 2^24 rays) for the staging
configuration in the environment (SSK_HOST_CHUNK_LOG2, SSK_HOST_STREAMS)."""
import os, sys, time
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(REPO, "mitsuba3-sunsky_b200"), os.path.join(REPO, "tests")):
    sys.path.insert(0, p)
import ctypes as C
import numpy as np, torch
import sunsky_b200 as ssk, common
from sunsky_b200 import _capi
L = _capi.lib(); n = 1 << 24
e = ssk.SunskyEmitter(spectral=True, arith="fast", turbidity=3.0, albedo=0.3, sun_direction=common.sun_dir_from_elevation(30.0))
g = torch.Generator(); g.manual_seed(1)
u1 = torch.rand(n, generator=g); u2 = torch.rand(n, generator=g)
st = torch.sqrt(1 - u1 * u1); ph = 2 * np.pi * u2
wi = torch.empty((3, n), pin_memory=True); wi.copy_(-torch.stack([st * torch.cos(ph), st * torch.sin(ph), u1]))
wav = torch.empty((4, n), pin_memory=True); wav.copy_(360 + 359 * torch.rand((4, n), generator=g))
out = torch.empty((4, n), pin_memory=True)
pl = lambda t: (_capi.FP * t.shape[0])(*[C.cast(t[i].data_ptr(), _capi.FP) for i in range(t.shape[0])])
a, b, c = pl(wi), pl(wav), pl(out)
for _ in range(2): assert L.ssk_eval_host(e._ctx, a, b, n, c) == 0
t0 = time.perf_counter(); k = 10
for _ in range(k): assert L.ssk_eval_host(e._ctx, a, b, n, c) == 0
dt = (time.perf_counter() - t0) / k
print(f"chunk 2^{os.environ.get('SSK_HOST_CHUNK_LOG2', '20')} streams {os.environ.get('SSK_HOST_STREAMS', '3')}: "
      f"{dt * 1e3:.2f} ms/step, {n / dt / 1e9:.3f} Gevals/s, PCIe {(11 * 4 * n) / dt / 1e9:.1f} GB/s")
