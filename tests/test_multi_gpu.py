"""Multi-GPU entry points of the C ABI (csrc/ssk_multi.cu). The single-process tests run on one GPU too (a group
may list a device twice: the sharding, threading and peer paths are the same); the two-rank test needs two GPUs
(`gpurun --gpus 2`) and is skipped otherwise."""
import os
import subprocess
import sys

import numpy as np
import pytest

import common
import emitters

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _devices():
    import torch
    n = torch.cuda.device_count()
    return [0, 1] if n >= 2 else [0, 0]


@pytest.mark.parametrize("spectral", [False, True])
def test_group_eval_and_sample_equal_single_context(ssk, spectral):
    from sunsky_b200 import multi
    name = "cfg2_spec" if spectral else "cfg3_rgb_sunsky"
    g = emitters.golden(name)
    kw = common.CONFIGS[name][1]
    grp = multi.Group(_devices(), spectral=spectral, **kw)
    rng = np.random.default_rng(3)
    n = 100_003                                             # ragged: the last device takes the tail
    wo = common.hemisphere_dirs(n, rng)
    wav = (360 + 359 * rng.random((4, n), dtype=np.float32)).astype(np.float32) if spectral else None
    u = rng.random((2, n), dtype=np.float32)
    e = emitters.make(ssk, name, g=g)
    assert common.bit_equal(grp.eval(-wo, wav), e.eval(-wo, wav))
    d, pdf, w = grp.sample_direction(u, wav)
    d1, pdf1, w1 = e.sample_direction(u, None, wav)
    assert common.bit_equal(d, d1) and common.bit_equal(pdf, pdf1) and common.bit_equal(w, w1)
    grp.close()


@pytest.mark.parametrize("config", ["time_of_day", "animation"])
@pytest.mark.parametrize("gathered", [False, True])
def test_group_sweep_equals_envmaps(ssk, config, gathered):
    import torch
    from sunsky_b200 import multi, sharding, sweep as sweep_mod
    spectral = config == "animation"
    allp = sharding.sweep_parameters(config)
    frames = list(range(0, len(allp), 29))
    props = [dict(turbidity=allp[f]["turbidity"], albedo=allp[f]["albedo"], sun_direction=sweep_mod.sun_direction(allp[f]["eta_deg"]))
             for f in frames]
    devs = _devices()
    grp = multi.Group(devs, spectral=spectral, **props[0])
    W, H = 256, 128
    wl = (440.0, 520.0, 600.0, 680.0)
    outs = grp.sweep(props, W, H, mode=multi.SHARD_ROUND_ROBIN if spectral else multi.SHARD_BLOCKED,
                     gather_device=devs[0] if gathered else None, wavelengths4=wl, chunk=2)
    ems = sweep_mod.make_emitters([allp[f] for f in frames], spectral, 0, "exact")
    want = ssk.envmaps(ems, W, H, wl if spectral else None)
    for f, t in enumerate(outs):
        assert t.device.index == (devs[0] if gathered else t.device.index)
        assert torch.equal(t.to("cuda:0").view(torch.int32), want[f].view(torch.int32)), (config, f)
    grp.close()


def test_peer_frames_in_one_process(ssk):
    """ssk_peer_alloc'ed memory as sweep destination and as a torch view (the root rank's side of the gather)."""
    import torch
    from sunsky_b200 import multi, sharding, sweep as sweep_mod
    allp = sharding.sweep_parameters("time_of_day")[:5]
    props = [dict(turbidity=p["turbidity"], albedo=p["albedo"], sun_direction=sweep_mod.sun_direction(p["eta_deg"])) for p in allp]
    W, H = 128, 64
    peer = multi.PeerFrames(5 * 3 * W * H, 0)
    multi.sweep_render(multi.frame_params(props, False), False, W, H, 0, peer.frame_pointers(3 * W * H, 5), chunk=2)
    torch.cuda.synchronize()
    want = ssk.envmaps(sweep_mod.make_emitters(allp, False, 0, "exact"), W, H)
    assert torch.equal(peer.tensor((5, 3, H, W)).view(torch.int32), want.view(torch.int32))
    peer.close()


def test_sweep_render_reuses_contexts_while_maps_are_queued(ssk):
    """The sweep loop re-parametrises its pooled contexts while their previous maps are still queued (chunk 1, large
    maps): every frame must still show its own emitter (device-block ring of ssk_update_batch)."""
    import torch
    from sunsky_b200 import multi, sharding, sweep as sweep_mod
    allp = sharding.sweep_parameters("time_of_day")[10:200:31]
    props = [dict(turbidity=p["turbidity"], albedo=p["albedo"], sun_direction=sweep_mod.sun_direction(p["eta_deg"])) for p in allp]
    W, H = 2048, 1024
    n = len(props)
    out = torch.empty((n, 3, H, W), device="cuda")
    ptrs = (multi._capi.FP * n)(*[multi.C.cast(out[f].data_ptr(), multi._capi.FP) for f in range(n)])
    multi.sweep_render(multi.frame_params(props, False), False, W, H, 0, ptrs, chunk=1)
    torch.cuda.synchronize()
    want = ssk.envmaps(sweep_mod.make_emitters(allp, False, 0, "exact"), W, H)
    assert torch.equal(out.view(torch.int32), want.view(torch.int32))


def test_two_rank_peer_store_and_nccl_gather():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    port = 29500 + os.getpid() % 2000
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(REPO, "tests", "multi_worker.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "time_of_day" in r.stdout and "animation" in r.stdout
