"""Python face of the library's multi-GPU entry points (include/sunsky_b200.h, "multi-GPU"; csrc/ssk_multi.cu).

All device work -- sharding arithmetic, the sweep loop, peer-mapped frame buffers, the NCCL send/recv gather, the
one-process N-GPU group driver -- is C++ in libsunsky_b200.so. What is left here is the host channel that a
one-process-per-GPU launch needs to hand the 64-byte buffer handle / 128-byte communicator id from the root rank
to the others (torch.distributed broadcast; gloo or nccl, it is host data) and torch views of the result."""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import Params, SHARD_BLOCKED, SHARD_ROUND_ROBIN, PEER_HANDLE_BYTES, COMM_ID_BYTES, DATASET_DIR


class MultiError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        L = _capi.lib()
        msg = L.ssk_multi_last_error().decode() or L.ssk_last_error().decode()
        raise MultiError(f"[ssk {rc}] {msg}")


def ray_range(n, rank, world_size, align=4):
    b, e = C.c_size_t(), C.c_size_t()
    if n < 0:
        raise ValueError("n must be >= 0")
    rc = _capi.lib().ssk_shard_ray_range(n, rank, world_size, max(align, 0), C.byref(b), C.byref(e))
    if rc:
        raise ValueError(_capi.lib().ssk_multi_last_error().decode())
    return b.value, e.value


def frames_for_rank(n_frames, rank, world_size, mode=SHARD_ROUND_ROBIN):
    cnt = C.c_int()
    buf = (C.c_int * max(n_frames, 1))()
    rc = _capi.lib().ssk_shard_frames(n_frames, rank, world_size, mode, buf, C.byref(cnt))
    if rc:
        raise ValueError(_capi.lib().ssk_multi_last_error().decode())
    return [buf[i] for i in range(cnt.value)]


def frame_params(props_list, spectral, arith="exact"):
    """ctypes array of ssk_params, one per frame, from property dicts (the keyword arguments of SunskyEmitter)."""
    from . import SunskyEmitter
    arr = (Params * len(props_list))()
    L = _capi.lib()
    shell = SunskyEmitter.__new__(SunskyEmitter)               # only _fill is used: no context is created
    for i, props in enumerate(props_list):
        shell.arith = arith; shell._normalized = False; shell._props = {}
        L.ssk_params_default(C.byref(arr[i]))
        shell._fill(arr[i], props, creating=True)
    return arr


class DeviceBuffer:
    """Raw device memory as a torch tensor (no copy) through __cuda_array_interface__."""
    def __init__(self, ptr, shape, device):
        self.ptr, self.shape, self.device = ptr, tuple(shape), device
        self.__cuda_array_interface__ = {"shape": self.shape, "typestr": "<f4", "data": (ptr, False), "version": 2}

    def tensor(self):
        import torch
        with torch.cuda.device(self.device):
            return torch.as_tensor(self, device=torch.device("cuda", self.device))


class PeerFrames:
    """A frame buffer on the root rank's GPU that every rank's kernels can store into (ssk_peer_alloc / _open).
    `dist` = torch.distributed (initialised) or None for a single process."""
    def __init__(self, n_floats, device, rank=0, root=0, dist=None):
        L = _capi.lib()
        self.device, self.rank, self.root, self.n_floats = device, rank, root, n_floats
        self.owner = rank == root
        handle = C.create_string_buffer(PEER_HANDLE_BYTES)
        p = C.c_void_p()
        if self.owner:
            _check(L.ssk_peer_alloc(device, n_floats * 4, C.byref(p), handle))
        if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
            box = [handle.raw if self.owner else None]
            dist.broadcast_object_list(box, src=root)
            if not self.owner:
                _check(L.ssk_peer_open(device, box[0], C.byref(p)))
        self.ptr = p.value

    def frame_pointers(self, frame_floats, n_frames):
        arr = (_capi.FP * n_frames)()
        for f in range(n_frames):
            arr[f] = C.cast(self.ptr + 4 * frame_floats * f, _capi.FP)
        return arr

    def tensor(self, shape):
        if not self.owner:
            raise MultiError("only the owning rank can view the buffer")
        return DeviceBuffer(self.ptr, shape, self.device).tensor()

    def close(self):
        if self.ptr:
            L = _capi.lib()
            (L.ssk_peer_free if self.owner else L.ssk_peer_close)(self.device, C.c_void_p(self.ptr))
            self.ptr = None


class Comm:
    """NCCL communicator owned by the library (ssk_comm_*): the id is created on rank 0 and broadcast as host data."""
    def __init__(self, rank, world, device, dist):
        L = _capi.lib()
        ident = C.create_string_buffer(COMM_ID_BYTES)
        if rank == 0:
            _check(L.ssk_comm_unique_id(ident))
        box = [ident.raw if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        h = C.c_void_p()
        _check(L.ssk_comm_create(box[0], rank, world, device, C.byref(h)))
        self.h, self.rank, self.world, self.device = h, rank, world, device

    def gather_frames(self, local, n_frames, mode, full=None, root=0, stream=None):
        """local: (k, ...) this rank's frames in shard order; full: (n_frames, ...) on the root."""
        import torch
        frame_floats = int(np.prod(local.shape[1:]))
        s = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream if stream is None else stream)
        _check(_capi.lib().ssk_comm_gather_frames(
            self.h, C.cast(local.data_ptr(), _capi.FP), frame_floats, n_frames, mode, 0, max(local.shape[0], 1) if n_frames else 0,
            C.cast(full.data_ptr(), _capi.FP) if full is not None else None, root, s))

    def close(self):
        if self.h:
            _capi.lib().ssk_comm_destroy(self.h)
            self.h = None


def sweep_render(params, spectral, width, height, device, out_ptrs, wavelengths4=None, chunk=8, stream=None,
                 dataset_dir=None):
    """ssk_sweep_render: params = ctypes array from frame_params (this rank's frames), out_ptrs = FP array with one
    destination per frame (local or peer-mapped)."""
    import torch
    n = len(params)
    wl = None
    if spectral:
        wl_np = np.ascontiguousarray(np.broadcast_to(np.asarray(wavelengths4, dtype=np.float32), (n, 4)))
        wl = wl_np.ctypes.data_as(_capi.FP)
    s = C.c_void_p(torch.cuda.current_stream(device).cuda_stream if stream is None else stream)
    _check(_capi.lib().ssk_sweep_render(params, n, (dataset_dir or DATASET_DIR).encode(), device,
                                        _capi.MODE_SPECTRAL if spectral else _capi.MODE_RGB, width, height, wl, out_ptrs,
                                        chunk, s))


class Group:
    """One process, N GPUs (ssk_group_*): numpy planes in host memory in, numpy planes out."""
    def __init__(self, devices, spectral=False, arith="exact", dataset_dir=None, **props):
        L = _capi.lib()
        self.spectral, self.devices = bool(spectral), list(devices)
        p = frame_params([props], spectral, arith)
        h = C.c_void_p()
        dv = (C.c_int * len(devices))(*devices)
        _check(L.ssk_group_create(dv, len(devices), p, (dataset_dir or DATASET_DIR).encode(),
                                  _capi.MODE_SPECTRAL if spectral else _capi.MODE_RGB, C.byref(h)))
        self.h = h
        self.channels = 4 if spectral else 3

    def eval(self, wi, wavelengths=None):
        from . import _planes_np, _pp
        planes = _planes_np(wi, 3, "wi")
        n = planes[0].size
        wp = _pp(_planes_np(wavelengths, 4, "wavelengths")) if self.spectral else None
        out = np.empty((self.channels, n), np.float32)
        _check(_capi.lib().ssk_group_eval_host(self.h, _pp(planes), wp, n, _pp([out[c] for c in range(self.channels)])))
        return out

    def sample_direction(self, u, wavelengths=None):
        from . import _planes_np, _pp
        up = _planes_np(u, 2, "u")
        n = up[0].size
        wp = _pp(_planes_np(wavelengths, 4, "wavelengths")) if self.spectral else None
        d = np.empty((3, n), np.float32); pdf = np.empty(n, np.float32); w = np.empty((self.channels, n), np.float32)
        _check(_capi.lib().ssk_group_sample_direction_host(self.h, _pp(up), wp, n, _pp([d[i] for i in range(3)]),
                                                           pdf.ctypes.data_as(_capi.FP), _pp([w[c] for c in range(self.channels)])))
        return d, pdf, w

    def sweep(self, props_list, width, height, mode=SHARD_BLOCKED, gather_device=None, wavelengths4=None, chunk=8, arith="exact"):
        """Returns a list of torch tensors (C, height, width), one per frame, on the GPU that holds it: the owner
        (gather_device None) or gather_device (the other GPUs store into it over NVLink)."""
        import torch
        n = len(props_list)
        fp = frame_params(props_list, self.spectral, arith)
        Cn = self.channels
        outs = []
        ptrs = (_capi.FP * n)()
        W = len(self.devices)
        owner = {}
        for r in range(W):
            for f in frames_for_rank(n, r, W, mode):
                owner[f] = self.devices[r]
        for f in range(n):
            dev = gather_device if gather_device is not None else owner[f]
            t = torch.empty((Cn, height, width), dtype=torch.float32, device=torch.device("cuda", dev))
            outs.append(t); ptrs[f] = C.cast(t.data_ptr(), _capi.FP)
        wl = None
        if self.spectral:
            wl_np = np.ascontiguousarray(np.broadcast_to(np.asarray(wavelengths4, dtype=np.float32), (n, 4)))
            wl = wl_np.ctypes.data_as(_capi.FP)
        for d in self.devices:
            torch.cuda.synchronize(d)
        _check(_capi.lib().ssk_group_sweep(self.h, fp, n, mode, width, height, wl, ptrs, chunk))
        return outs

    def close(self):
        if self.h:
            _capi.lib().ssk_group_destroy(self.h)
            self.h = None
