"""Environment-map sweeps of BASELINE configs[3] and configs[4]: many emitters (a time-of-day sweep, a
turbidity / albedo animation), one equirectangular radiance map each, sharded over the GPUs of one box.

A map is rendered by `ssk_envmap_device`: directions come from the pixel index (no direction buffer is read),
the kernel only writes. Each rank renders the frames `sharding.frames_for_rank` assigns to it; nothing is
exchanged unless the caller asks for the frames on rank 0 (`gather=True`)."""
import numpy as np

from . import SunskyEmitter, envmaps, update_batch
from . import sharding


def sun_direction(eta_deg, phi=-4.0 * np.pi / 5.0):
    """Unit vector at elevation eta_deg, azimuth phi (the azimuth used throughout the reference's sunsky-testing/)."""
    th = np.pi / 2 - np.radians(eta_deg)
    return [float(np.cos(phi) * np.sin(th)), float(np.sin(phi) * np.sin(th)), float(np.cos(th))]


def make_emitters(params, spectral, device, arith="exact"):
    return [SunskyEmitter(spectral=spectral, device=device, arith=arith, turbidity=p["turbidity"], albedo=p["albedo"],
                          sun_direction=sun_direction(p["eta_deg"])) for p in params]


def render_sweep(config, width, height, rank=0, world_size=1, device=0, arith="exact", gather=False,
                 wavelengths4=(440.0, 520.0, 600.0, 680.0), chunk=8, out=None, frames=None):
    """Render this rank's share of a sweep.

    config: 'time_of_day' (RGB, 256 maps, blocked over ranks) or 'animation' (spectral, 128 frames,
    round-robin); `frames` overrides the frame list (tests use a subset). Returns (frame_indices, maps) with
    maps of shape (k, C, height, width) on the GPU; with gather=True rank 0 gets all frames in order instead
    (frame_indices = every frame), the other ranks (their indices, None).
    Emitter contexts are created `chunk` at a time so a 256-map sweep does not hold 256 contexts alive."""
    import torch
    spectral = config == "animation"
    mode = "round_robin" if spectral else "blocked"
    params = sharding.sweep_parameters(config)
    if frames is not None:
        params = [params[f] for f in frames]
    n_frames = len(params)
    mine = sharding.frames_for_rank(n_frames, rank, world_size, mode)
    C = 4 if spectral else 3
    dev = torch.device("cuda", device)
    if out is None:
        out = torch.empty((len(mine), C, height, width), dtype=torch.float32, device=dev)
    # Two pools of `chunk` emitter contexts, re-parametrised in batches (ssk_update_batch: one chain-sum launch and
    # one synchronisation per chunk): while pool A's maps are being written, pool B's emitters are set up.
    pools, events = [None, None], [None, None]
    for k, s in enumerate(range(0, len(mine), chunk)):
        idx = mine[s:s + chunk]
        slot = k & 1
        if events[slot] is not None:
            events[slot].synchronize()                    # the contexts must outlive their previous launches
        props = [dict(turbidity=params[f]["turbidity"], albedo=params[f]["albedo"],
                      sun_direction=sun_direction(params[f]["eta_deg"])) for f in idx]
        if pools[slot] is None or len(pools[slot]) != len(idx):
            for e in pools[slot] or []:
                e.close()
            pools[slot] = make_emitters([params[f] for f in idx], spectral, device, arith)
        else:
            update_batch(pools[slot], props)
        envmaps(pools[slot], width, height, wavelengths4 if spectral else None, out=out[s:s + len(idx)])
        events[slot] = torch.cuda.Event(); events[slot].record(torch.cuda.current_stream(dev))
    for slot in (0, 1):
        if events[slot] is not None:
            events[slot].synchronize()
        for e in pools[slot] or []:
            e.close()
    if gather:
        full = sharding.gather_frames(out, n_frames, mode=mode, dst=0)
        return (list(range(n_frames)), full) if full is not None else (mine, None)
    return mine, out
