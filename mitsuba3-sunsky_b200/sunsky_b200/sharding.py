"""Multi-GPU sharding of sunsky query batches: one process per GPU, one emitter context per process.

Every ray / sample / pixel is independent given the 3.4 KB per-emitter block, which each rank builds for
itself from the same properties, so the data path needs NO collective. Work is split as SURVEY.md 8(e) says:

  * eval / pdf / sample_direction batches (BASELINE configs 2, 3): contiguous ray ranges, `ray_range`
  * environment-map sweeps (configs 4, 5): whole maps / frames per rank, `frames_for_rank`
    (blocked = consecutive maps stay together; round-robin = frame f on rank f % world)

The only optional exchange is collecting finished frames on one rank (`gather_frames`), over
`torch.distributed` -- NCCL between GPUs, gloo in the CPU tests. The reference has no multi-device layer at
all (no NCCL / MPI anywhere under src/ or include/), so this module has no counterpart there."""
import os


def world():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) outside of it."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def ray_range(n, rank, world_size, align=4):
    """Contiguous [begin, end) of an n-ray batch for `rank` (ssk_shard_ray_range). Boundaries are multiples of
    `align` rays so that every shard's planes stay 16-byte aligned (128-bit kernel variant); the last rank takes the
    ragged tail. Shards differ by at most `align` rays; empty shards (n < world_size * align) are legal."""
    from . import multi
    if n < 0 or align < 1:
        raise ValueError("n must be >= 0 and align >= 1")
    return multi.ray_range(n, rank, world_size, align)


def frames_for_rank(n_frames, rank, world_size, mode="round_robin"):
    """Frame / map indices rendered by `rank` (ssk_shard_frames). round_robin: f % world == rank (config 5: an
    animation stays load-balanced however long it is); blocked: consecutive maps (config 4: 256 maps, 32 per GPU)."""
    from . import multi
    if mode not in ("round_robin", "blocked"):
        raise ValueError("mode must be 'round_robin' or 'blocked'")
    return multi.frames_for_rank(n_frames, rank, world_size,
                                 multi.SHARD_ROUND_ROBIN if mode == "round_robin" else multi.SHARD_BLOCKED)


def owner_of_frame(f, n_frames, world_size, mode="round_robin"):
    if mode == "round_robin":
        return f % world_size
    for r in range(world_size):
        b, e = ray_range(n_frames, r, world_size, align=1)
        if b <= f < e:
            return r
    raise ValueError("frame out of range")


def gather_frames(local_frames, n_frames, mode="round_robin", dst=0, group=None):
    """Collect per-rank frames on rank `dst`, in frame order.

    local_frames: tensor (k, ...) holding this rank's frames in the order of frames_for_rank().
    Returns a (n_frames, ...) tensor on `dst`, None elsewhere. Point-to-point sends posted as one batch
    (`batch_isend_irecv`): NCCL groups them into one launch and NVSwitch gives every sender full bandwidth
    to the root, so there is no reason to stage through an all-gather. Ranks may own different frame counts."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized():
        if n_frames != local_frames.shape[0]:
            raise ValueError("single process: local_frames must hold every frame")
        return local_frames
    rank, ws = dist.get_rank(group), dist.get_world_size(group)
    mine = frames_for_rank(n_frames, rank, ws, mode)
    if local_frames.shape[0] != len(mine):
        raise ValueError(f"rank {rank} owns {len(mine)} frames, got {local_frames.shape[0]}")
    local_frames = local_frames.contiguous()
    if rank != dst:
        if len(mine):
            for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, local_frames, dst, group)]):
                w.wait()
        return None
    out = torch.empty((n_frames,) + tuple(local_frames.shape[1:]), dtype=local_frames.dtype, device=local_frames.device)
    ops, staged = [], []
    for r in range(ws):
        idx = frames_for_rank(n_frames, r, ws, mode)
        if not idx:
            continue
        if r == dst:
            out[idx] = local_frames
            continue
        buf = torch.empty((len(idx),) + tuple(local_frames.shape[1:]), dtype=local_frames.dtype, device=local_frames.device)
        ops.append(dist.P2POp(dist.irecv, buf, r, group))
        staged.append((idx, buf))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    for idx, buf in staged:
        out[idx] = buf
    return out


def sweep_parameters(config):
    """Emitter property sets of the BASELINE env-map sweeps.
    'time_of_day' (configs[3]): 64 sun elevations 90 deg * k / 63 x turbidity {2, 4, 6, 8}, albedo 0.3, RGB;
    'animation'   (configs[4]): 128 frames, turbidity 2 + 6 k / 127, albedo k / 127, sun elevation 30 deg, spectral.
    Returns a list of dicts (turbidity, albedo, sun elevation in degrees) in frame order."""
    if config == "time_of_day":
        return [dict(turbidity=float(t), albedo=0.3, eta_deg=90.0 * k / 63.0) for t in (2, 4, 6, 8) for k in range(64)]
    if config == "animation":
        return [dict(turbidity=2.0 + 6.0 * k / 127.0, albedo=k / 127.0, eta_deg=30.0) for k in range(128)]
    raise ValueError("config must be 'time_of_day' or 'animation'")
