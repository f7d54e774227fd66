"""N > 1 host logic on the CPU: two gloo ranks run the point-range split and the blob sharding with the ORACLE
standing in for the device (the GPU kernels are covered by the -m gpu tests; this covers rank arithmetic, the
all-gather of partial sums in rank order and the fold)."""
import os
import random
import sys

import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kzg-commitments_b200"))
import sharding  # noqa: E402


def test_slices_cover_everything_once():
    for world in (1, 2, 3, 4, 8):
        for total in (0, 1, 7, 8, 4096, 4097):
            cuts = [sharding.blob_slice(r, world, total) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from oracle.kzg_oracle import get_curve
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    C = get_curve("bn158")
    rng = random.Random(99)
    s = rng.randrange(C.r)
    n = 24
    g1, _ = C.setup(s, n)
    scalars = [rng.randrange(C.r) for _ in range(n)]
    lo, hi = sharding.point_range(rank, world, n)
    part = C.msm(scalars[lo:hi], g1[lo:hi])          # this rank's partial sum (affine point or None)
    raw = C.g1_to_octet(part)                          # fixed-size bytes; infinity = zeros
    parts = sharding.gather_partials(dist, raw)
    acc = None
    for p in parts:                                    # fold in rank order, as kzg_gpu_g1_sum_partials does
        body = p[1:]
        pt = None if body == bytes(len(body)) else (int.from_bytes(body[:len(body) // 2], "big"),
                                                    int.from_bytes(body[len(body) // 2:], "big"))
        acc = C.add(acc, pt)
    full = C.msm(scalars, g1)
    # blob sharding: every rank commits to its own blobs, results gathered by rank order
    blobs = [[rng.randrange(C.r) for _ in range(6)] for _ in range(5)]
    b0, b1 = sharding.blob_slice(rank, world, len(blobs))
    mine = b"".join(C.g1_to_octet(C.msm(b, g1[:6])) for b in blobs[b0:b1])
    mine = mine.ljust(3 * len(C.g1_to_octet(None)), b"\0")   # fixed-size message (<= 3 blobs per rank here)
    gathered = sharding.gather_partials(dist, mine)
    q.put((rank, acc == full, len(parts), gathered[0][:len(C.g1_to_octet(None))] == C.g1_to_octet(C.msm(blobs[0], g1[:6]))))
    dist.destroy_process_group()


def test_point_range_split_and_fold_two_ranks_gloo():
    world, port = 2, 29500 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_fold, nparts, ok_blob in results:
        assert ok_fold and nparts == world and ok_blob, (rank, ok_fold, nparts, ok_blob)
