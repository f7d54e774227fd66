"""Rank-side logic of the two multi-GPU shardings; one process per GPU.

1. independent blobs, no collective:   blob_slice(rank, world, batch)
2. point-range split of one large MSM: point_range(rank, world, n) + gather_partials(...) — every rank ends
   with all `world` projective partial sums (4*L words each) in rank order and folds them locally
   (kzg_gpu_g1_sum_partials); EC addition is not an NCCL reduction op, hence all-gather + local fold.
"""


def blob_slice(rank, world, batch):
    """Contiguous share of `batch` independent blobs owned by `rank` (sizes differ by at most one)."""
    base, extra = divmod(batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def point_range(rank, world, n):
    """Contiguous range of points/scalars of an n-term MSM owned by `rank`."""
    return blob_slice(rank, world, n)


def gather_partials(dist, partial_bytes, device=None):
    """All-gather one fixed-size byte string per rank (NCCL on `device`, or gloo on the CPU when device is None).
    Returns the list of byte strings in rank order."""
    import torch
    t = torch.frombuffer(bytearray(partial_bytes), dtype=torch.uint8)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [bytes(o.cpu().numpy().tobytes()) for o in out]


class _DevBytes:
    """A raw device allocation presented through __cuda_array_interface__ so torch can wrap it without a copy."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def gather_partials_device(dist, ctx, device):
    """NCCL all-gather of this rank's projective partial straight from the context's device buffer, then the fold
    on the device (kzg_gpu_g1_sum_partials_device).  Either the caller has already waited for the MSM on ctx.stream,
    or it runs this with ctx.stream as torch's current stream, which orders MSM -> all-gather -> fold on the device."""
    import torch
    ptr, nb = ctx.msm_partial_device()
    mine = torch.as_tensor(_DevBytes(ptr, nb), device=device)
    world = dist.get_world_size()
    allp = torch.empty(world * nb, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(allp, mine)
    cur = torch.cuda.current_stream(device)
    if cur.cuda_stream != ctx.stream:  # under `with torch.cuda.stream(ExternalStream(ctx.stream))` the fold is already
        cur.synchronize()              # ordered after the collective on the context's own stream
    return ctx.g1_sum_partials_device(allp, world)
