"""Ray-range sharding of a frame across ranks (one process per GPU) and the single gather to rank 0.

The reference parallelises over independent 8x8 pixel blocks with rayon (src/lib.rs:533-550); every ray costs
exactly n_coarse + (n_coarse + n_fine) MLP samples, so equal contiguous pixel-index ranges are perfectly
balanced.  Weights are replicated; the only exchange is the gather of the rgb ranges at the end of the frame.
"""


def ray_range(rank, world, n_rays):
    """Pixel-index range [begin, end) of `rank` (pixel_index = i * W + j, src/lib.rs:371)."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    return rank * n_rays // world, (rank + 1) * n_rays // world


def ray_counts(world, n_rays):
    return [ray_range(r, world, n_rays)[1] - ray_range(r, world, n_rays)[0] for r in range(world)]


def view_pixels(part, parts, stripe, n_rays):
    """Pixel indices of part `part` when the frame's n_rays pixel indices are dealt out in stripes of `stripe` to
    `parts` renderers in turn (stripe k belongs to part k % parts) - the numpy statement of RayMap::pixel /
    view_size (csrc/common.cuh), in view order.  Interleaved stripes balance the data-dependent cost of rays (empty
    rows are cheaper than rows through the object) that contiguous ranges do not."""
    import numpy as np
    if not 0 <= part < parts or stripe <= 0:
        raise ValueError("need stripe > 0 and 0 <= part < parts")
    starts = np.arange(part * stripe, n_rays, stripe * parts, dtype=np.int64)
    if len(starts) == 0:
        return np.zeros(0, np.int64)
    idx = (starts[:, None] + np.arange(stripe, dtype=np.int64)[None, :]).ravel()
    return idx[idx < n_rays]


def gather_frame(local, frame, rank, world, dist=None, group=None):
    """Gathers each rank's [n_local, 3] rgb rows into `frame` [n_rays, 3] on rank 0 (NCCL over NVLink on GPUs,
    gloo in the CPU tests).  `frame` is only read on rank 0; with world == 1 it must alias `local`."""
    if world == 1:
        return frame
    n_rays = frame.shape[0] if rank == 0 else None
    if rank == 0:
        counts = ray_counts(world, n_rays)
        offs = [sum(counts[:k]) for k in range(world + 1)]
        parts = [frame[offs[k]:offs[k + 1]] for k in range(world)]
        if parts[0].data_ptr() != local.data_ptr():
            parts[0].copy_(local)
        # ranges may differ by one ray, so this is a group of point-to-point receives (one NCCL group call)
        ops = [dist.P2POp(dist.irecv, parts[k], k, group) for k in range(1, world) if counts[k] > 0]
        for w in (dist.batch_isend_irecv(ops) if ops else []):
            w.wait()
        return frame
    if local.shape[0] > 0:
        for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, local, 0, group)]):
            w.wait()
    return None


class SharedFrame:
    """The frame buffer of a multi-GPU render, in rank 0's HBM, mapped into every rank: each rank renders its ray range
    with `d_rgb = frame.row_ptr(ray_begin)`, so the compositing kernel's stores go straight to rank 0 over NVLink and no
    collective moves pixels.  `complete()` is the only synchronisation: every rank drains its stream, then a barrier.

    how = "vmm" (default): the buffer alone is shared, through the CUDA virtual-memory API
    (nerf_shared_frame_*_fd); rank 0 hands its file descriptor to each rank over a Unix socket (SCM_RIGHTS).
    how = "ipc": cudaIpc handles (nerf_shared_frame_create / open); simpler, but opening one turns on device-wide peer
    access, which was measured to slow the MLP kernels of every rank by 1-3 %."""

    def __init__(self, renderer, n_rays, rank, world, dist=None, group=None, channels=3, itemsize=4, how="vmm"):
        import ctypes as C
        from . import _check, lib
        self.renderer, self.rank, self.world, self.dist, self.group = renderer, rank, world, dist, group
        self.n_rays, self.row_bytes = n_rays, channels * itemsize
        self.channels, self.how = channels, how
        ptr = C.c_void_p()
        if how == "vmm":
            self.ptr = self._init_vmm(ptr)
            self.handle = None
            return
        handle, err = [None], None
        if rank == 0:
            buf = C.create_string_buffer(64)
            rc = lib().nerf_shared_frame_create(renderer._h, n_rays * self.row_bytes, C.byref(ptr), buf)
            if rc == 0:
                handle = [buf.raw]
            else:
                err = lib().nerf_last_error().decode()
        if world > 1:
            dist.broadcast_object_list(handle, src=0, group=group)  # 64 bytes over the control plane
        if handle[0] is None:  # every rank learns of rank 0's failure through the broadcast: nobody is left waiting
            raise RuntimeError("shared frame: rank 0 could not create it" + (f" ({err})" if err else ""))
        if rank != 0:
            _check(lib().nerf_shared_frame_open(renderer._h, handle[0], C.byref(ptr)))
        self.handle = handle[0]
        self.ptr = ptr.value

    def _init_vmm(self, ptr):
        """Rank 0 creates the allocation and serves its file descriptor on a Unix socket; the others connect, receive
        it (SCM_RIGHTS) and map it.  The socket path and the mapped size travel through broadcast_object_list."""
        import ctypes as C
        import os
        import socket
        import tempfile
        from . import _check, lib
        L, h = lib(), self.renderer._h
        info, srv, fd = [None], None, -1
        if self.rank == 0:
            size, cfd = C.c_uint64(), C.c_int(-1)
            rc = L.nerf_shared_frame_create_fd(h, self.n_rays * self.row_bytes, C.byref(ptr), C.byref(size), C.byref(cfd))
            if rc == 0:
                fd = cfd.value
                path = None
                if self.world > 1:
                    path = os.path.join(tempfile.mkdtemp(prefix="nerf_frame_"), "fd.sock")
                    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                    srv.bind(path)
                    srv.listen(self.world)
                    srv.settimeout(60.0)  # a rank that died before connecting must not hang the job: accept() raises,
                    #                       the caller's all_reduce(MIN) then moves every rank to the NCCL gather
                info = [(path, size.value)]
            else:
                info = [("error: " + L.nerf_last_error().decode(), 0)]
        if self.world > 1:
            self.dist.broadcast_object_list(info, src=0, group=self.group)
        path, size = info[0]
        if size == 0:
            raise RuntimeError("shared frame: rank 0 could not create it (" + str(path) + ")")
        if self.rank == 0:
            try:
                for _ in range(self.world - 1):
                    conn, _ = srv.accept()
                    socket.send_fds(conn, [b"f"], [fd])
                    conn.close()
            finally:
                if srv is not None:
                    srv.close()
                    os.unlink(path)
                    os.rmdir(os.path.dirname(path))
                os.close(fd)  # the mapping keeps the allocation alive
        else:
            cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            cli.settimeout(60.0)
            cli.connect(path)
            _, fds, _, _ = socket.recv_fds(cli, 16, 1)
            cli.close()
            try:
                _check(L.nerf_shared_frame_open_fd(h, fds[0], size, C.byref(ptr)))
            finally:
                os.close(fds[0])
        return ptr.value

    def row_ptr(self, ray):
        return self.ptr + ray * self.row_bytes

    def complete(self, stream=None):
        """Returns once every rank's pixels are in rank 0's buffer."""
        if stream is not None:
            stream.synchronize()
        if self.world > 1:
            self.dist.barrier(group=self.group)

    def tensor(self):
        """Rank 0: the frame as a torch tensor aliasing the buffer (float32 [n_rays, channels])."""
        import torch

        class _Raw:
            pass
        raw = _Raw()
        raw.__cuda_array_interface__ = {"shape": (self.n_rays, self.channels), "typestr": "<f4",
                                        "data": (self.ptr, False), "version": 2}
        return torch.as_tensor(raw, device=torch.device("cuda", self.renderer.device))

    def close(self):
        from . import _check, lib
        if self.ptr:
            if self.how == "vmm":
                _check(lib().nerf_shared_frame_close_fd(self.renderer._h, self.ptr))
            else:
                _check(lib().nerf_shared_frame_close(self.renderer._h, self.ptr, 1 if self.rank == 0 else 0))
            self.ptr = 0
