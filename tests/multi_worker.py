"""Worker of tests/test_multi_gpu.py::test_two_rank_*: launched by torch.distributed.run with one process per GPU.
Renders a small sweep sharded over the ranks with the frames collected on rank 0 (a) by peer stores straight from
envmap_kernel into rank 0's IPC-mapped buffer and (b) by the library's NCCL send/recv gather on a second stream,
and compares both, bit for bit, with rank 0 rendering every frame alone."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(os.path.dirname(HERE), "mitsuba3-sunsky_b200"), HERE]
import ctypes as C  # noqa: E402

import sunsky_b200 as ssk  # noqa: E402
from sunsky_b200 import _capi, multi, sharding, sweep as sweep_mod  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")                      # host channel only: handles and ids are host data
    L = _capi.lib()
    for config, W, H, frames in (("time_of_day", 256, 128, list(range(0, 256, 23))), ("animation", 128, 64, list(range(0, 128, 19)))):
        spectral = config == "animation"
        mode = multi.SHARD_ROUND_ROBIN if spectral else multi.SHARD_BLOCKED
        Cn = 4 if spectral else 3
        allp = sharding.sweep_parameters(config)
        props = [dict(turbidity=allp[f]["turbidity"], albedo=allp[f]["albedo"],
                      sun_direction=sweep_mod.sun_direction(allp[f]["eta_deg"])) for f in frames]
        n = len(props)
        mine = multi.frames_for_rank(n, rank, world, mode)
        wl = (440.0, 520.0, 600.0, 680.0)
        frame_floats = Cn * W * H
        # (a) peer stores
        peer = multi.PeerFrames(n * frame_floats, local, rank, 0, dist)
        ptrs = peer.frame_pointers(frame_floats, n)
        my_ptrs = (_capi.FP * max(len(mine), 1))(*[ptrs[f] for f in mine])
        multi.sweep_render(multi.frame_params([props[f] for f in mine], spectral, "exact"), spectral, W, H, local, my_ptrs,
                           wavelengths4=wl, chunk=3)
        torch.cuda.synchronize()
        dist.barrier()
        # (b) NCCL gather inside the library, double-buffered
        comm = multi.Comm(rank, world, local, dist)
        chunk = 2
        staging = torch.empty((2 * chunk, Cn, H, W), device="cuda")
        full = torch.zeros((n, Cn, H, W), device="cuda") if rank == 0 else None
        fp = multi.frame_params([props[f] for f in mine], spectral, "exact")
        wl_np = np.ascontiguousarray(np.broadcast_to(np.asarray(wl, np.float32), (max(len(mine), 1), 4)))
        rc = L.ssk_sweep_render_gather_nccl(comm.h, fp, n, mode, _capi.DATASET_DIR.encode(),
                                            _capi.MODE_SPECTRAL if spectral else _capi.MODE_RGB, W, H,
                                            wl_np.ctypes.data_as(_capi.FP) if spectral else None,
                                            C.cast(staging.data_ptr(), _capi.FP),
                                            C.cast(full.data_ptr(), _capi.FP) if rank == 0 else None, 0, chunk,
                                            C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, L.ssk_multi_last_error()
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            ems = sweep_mod.make_emitters([allp[f] for f in frames], spectral, local, "exact")
            want = ssk.envmaps(ems, W, H, wl if spectral else None)
            got_peer = peer.tensor((n, Cn, H, W))
            assert torch.equal(got_peer.view(torch.int32), want.view(torch.int32)), f"{config}: peer-store gather differs"
            assert torch.equal(full.view(torch.int32), want.view(torch.int32)), f"{config}: NCCL gather differs"
            assert float(want.abs().max()) > 0
            print(f"{config}: {n} frames over {world} ranks, peer stores and NCCL {L.ssk_comm_version()} gather bit-identical to one GPU")
        dist.barrier()
        comm.close(); peer.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
