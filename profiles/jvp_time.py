"""Throughput of ssk_eval_jvp_device (forward derivative of eval) on 2^24 rays, CUDA events, 20 launches."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mitsuba3-sunsky_b200"))
import sunsky_b200 as ssk

n = 1 << 24
g = torch.Generator(device="cuda").manual_seed(1)
u = torch.rand((2, n), device="cuda", generator=g)
ct = u[0]; st = (1 - ct * ct).clamp_min(0).sqrt(); ph = 6.2831853 * u[1]
wi = -torch.stack([st * ph.cos(), st * ph.sin(), ct]).contiguous()
wav = (360 + 360 * torch.rand((4, n), device="cuda", generator=g)).contiguous()
for spectral in (False, True):
    e = ssk.SunskyEmitter(spectral=spectral, turbidity=3.25, albedo=0.3, sun_direction=[0.3, 0.2, 0.6])
    w = wav if spectral else None
    for _ in range(3):
        e.eval_jvp(wi, w, d_turbidity=1.0, d_albedo=0.5)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(20):
        e.eval_jvp(wi, w, d_turbidity=1.0, d_albedo=0.5)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print(f"eval_jvp {'spectral' if spectral else 'rgb'}: {ms:.3f} ms / 2^24 rays = {n / ms / 1e6:.2f} Grays/s (includes the host state tangent per call)")
