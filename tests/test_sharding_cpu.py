"""CPU tests of the multi-GPU host logic (sunsky_b200.sharding): partition properties and a world_size-2
`gloo` run of the frame gather, the only exchange step of the path."""
import os
import socket
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "mitsuba3-sunsky_b200"))

from sunsky_b200 import sharding  # noqa: E402


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 1023, 1 << 24, (1 << 26) + 7])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_ray_ranges_tile_the_batch(n, world):
    ranges = [sharding.ray_range(n, r, world) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    for (b0, e0), (b1, e1) in zip(ranges, ranges[1:]):
        assert e0 == b1 and b0 <= e0
    for b, e in ranges[:-1]:
        assert b % 4 == 0 and e % 4 == 0                      # shards keep 16-byte aligned planes
    sizes = [e - b for b, e in ranges]
    assert max(sizes) - min(sizes) <= 4 + 3                   # balanced up to one aligned group + ragged tail


@pytest.mark.parametrize("mode", ["round_robin", "blocked"])
@pytest.mark.parametrize("n_frames,world", [(128, 8), (256, 8), (5, 2), (3, 4), (1, 8), (0, 2)])
def test_frames_partition(mode, n_frames, world):
    owned = [sharding.frames_for_rank(n_frames, r, world, mode) for r in range(world)]
    flat = sorted(f for o in owned for f in o)
    assert flat == list(range(n_frames))
    assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1
    for r, o in enumerate(owned):
        for f in o:
            assert sharding.owner_of_frame(f, n_frames, world, mode) == r


def test_bad_arguments():
    with pytest.raises(ValueError):
        sharding.ray_range(10, 2, 2)
    with pytest.raises(ValueError):
        sharding.frames_for_rank(4, 0, 2, mode="zigzag")
    with pytest.raises(ValueError):
        sharding.sweep_parameters("nope")


def test_sweep_parameters_match_baseline_configs():
    tod = sharding.sweep_parameters("time_of_day")
    assert len(tod) == 256 and tod[0]["eta_deg"] == 0.0 and tod[63]["eta_deg"] == 90.0
    assert sorted({p["turbidity"] for p in tod}) == [2.0, 4.0, 6.0, 8.0]
    ani = sharding.sweep_parameters("animation")
    assert len(ani) == 128 and ani[0] == dict(turbidity=2.0, albedo=0.0, eta_deg=30.0)
    assert abs(ani[-1]["turbidity"] - 8.0) < 1e-12 and ani[-1]["albedo"] == 1.0


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _gather_worker(rank, world, port, n_frames, mode, ret):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = sharding.frames_for_rank(n_frames, rank, world, mode)
        # frame f = a (2, 3, 5) "image" filled with f + small per-pixel ramp, as a rank would have rendered it
        base = torch.arange(30, dtype=torch.float32).reshape(2, 3, 5) / 64.0
        local = torch.stack([base + f for f in mine]) if mine else torch.empty((0, 2, 3, 5))
        out = sharding.gather_frames(local, n_frames, mode=mode, dst=0)
        if rank == 0:
            want = torch.stack([base + f for f in range(n_frames)]) if n_frames else torch.empty((0, 2, 3, 5))
            ret["ok"] = bool(torch.equal(out, want))
        else:
            ret[f"none{rank}"] = out is None
        # ray ranges: every rank reduces the number of rays it owns; the sum is the batch
        b, e = sharding.ray_range(1000003, rank, world)
        t = torch.tensor([e - b], dtype=torch.int64)
        dist.all_reduce(t)
        ret[f"sum{rank}"] = int(t.item())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode,n_frames", [("round_robin", 7), ("blocked", 7), ("round_robin", 1)])
def test_gather_frames_world_size_2_gloo(mode, n_frames):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_gather_worker, args=(world, port, n_frames, mode, ret), nprocs=world, join=True)
        assert ret["ok"] is True and ret["none1"] is True
        assert ret["sum0"] == ret["sum1"] == 1000003


def test_gather_frames_single_process_is_identity():
    import torch
    x = torch.arange(24, dtype=torch.float32).reshape(4, 6)
    assert sharding.gather_frames(x, 4) is x
    with pytest.raises(ValueError):
        sharding.gather_frames(x, 5)


def test_world_reads_torchrun_environment(monkeypatch):
    monkeypatch.setenv("RANK", "3"); monkeypatch.setenv("WORLD_SIZE", "8"); monkeypatch.setenv("LOCAL_RANK", "3")
    assert sharding.world() == (3, 8, 3)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k)
    assert sharding.world() == (0, 1, 0)
    assert isinstance(np.int64(1), np.integer)
