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
