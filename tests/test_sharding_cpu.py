"""Multi-process (world_size 2 and 3, gloo, CPU) tests of the frame/stream sharding logic:
disjoint + complete partitions, ragged and empty shards, order-preserving gather, no collective when
gather=False."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from depth_estimation_b200 import sharding


@pytest.mark.parametrize("n", [0, 1, 2, 7, 8, 31, 32, 33, 256])
@pytest.mark.parametrize("G", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("mode", ["slab", "stream"])
def test_partitions_are_disjoint_and_complete(n, G, mode):
    parts = [sharding.owned_indices(n, r, G, mode) for r in range(G)]
    allidx = np.concatenate(parts) if parts else np.zeros(0, np.int64)
    assert sorted(allidx.tolist()) == list(range(n))
    sizes = [len(p) for p in parts]
    assert max(sizes) - min(sizes) <= 1                       # balanced
    if mode == "slab":
        for r in range(G):
            lo, hi = sharding.slab_range(n, r, G)
            assert parts[r].tolist() == list(range(lo, hi))
    else:
        for r in range(G):
            assert all(i % G == r for i in parts[r])


def test_bad_arguments():
    with pytest.raises(ValueError):
        sharding.slab_range(4, 2, 2)
    with pytest.raises(ValueError):
        sharding.owned_indices(4, 0, 2, "diagonal")
    with pytest.raises(ValueError):
        sharding.ShardedEstimator(object(), mode="nope")


class _FakeEstimator:
    """predict_batch stand-in: the map of a frame is a deterministic function of its pixels only, so any
    mix-up of frame order or ownership shows in the gathered result."""

    def __init__(self):
        self.calls = 0
        self.frames_seen = 0

    def predict_batch(self, frames):
        self.calls += 1
        self.frames_seen += frames.shape[0]
        return frames.astype(np.float32).sum(axis=3) * 0.5 + 1.0


def _frames(n, H=6, W=5):
    rng = np.random.default_rng(7)
    return rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, G, port, n, mode, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(G))
    dist.init_process_group("gloo", rank=rank, world_size=G)
    try:
        est = _FakeEstimator()
        sh = sharding.ShardedEstimator(est, mode)
        frames = _frames(n)
        out = sh.predict_sharded(frames, gather=True, dst=0)
        idx, mine = sh.predict_sharded(frames, gather=False)
        q.put((rank, None if out is None else out.copy(), idx.copy(), mine.copy(), est.frames_seen))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("G,n,mode", [(2, 8, "slab"), (2, 7, "stream"), (2, 1, "slab"), (3, 5, "stream"), (2, 0, "slab")])
def test_sharded_predict_gloo(G, n, mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, G, port, n, mode, q)) for r in range(G)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(G)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort(key=lambda t: t[0])
    want = _FakeEstimator().predict_batch(_frames(n)) if n else np.zeros((0, 0, 0), np.float32)
    seen = 0
    for rank, out, idx, mine, frames_seen in res:
        if rank == 0:
            assert out is not None and out.shape[0] == n
            if n:
                assert np.array_equal(out, want)             # original frame order, bit-identical
        else:
            assert out is None
        assert idx.tolist() == sharding.owned_indices(n, rank, G, mode).tolist()
        if len(idx):
            assert np.array_equal(mine, want[idx])
        seen += frames_seen
    assert seen == 2 * n                                     # every frame computed exactly once per call


def test_single_process_passthrough():
    est = _FakeEstimator()
    out = sharding.ShardedEstimator(est).predict_sharded(_frames(4))
    assert np.array_equal(out, _FakeEstimator().predict_batch(_frames(4)))
    t = torch.zeros(3, 2, 2)
    assert sharding.gather_frames(t, 3) is t
    with pytest.raises(ValueError):
        sharding.gather_frames(t, 4)
