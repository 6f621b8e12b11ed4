"""Frame / stream data parallelism: every frame is an independent forward pass
(/root/reference/depth_midas.py:21-35 holds no cross-frame state), so ranks own disjoint frames, weights are
replicated and there is NO collective on the hot path.  torch.distributed (NCCL on the GPU box, gloo in the CPU
tests) is used only to bring per-frame outputs back to one rank when the caller asks for that.

Two partitions, both order-preserving after `gather_frames`:
  * slab   - contiguous B/G blocks of one batch (batch mode, BASELINE config #3)
  * stream - frame i -> rank i mod G (stream mode, configs #4/#5: keeps per-rank latency even on a live feed)
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    """(rank, world_size) from an initialised process group, else from the torchrun environment, else (0, 1)."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def slab_range(n_frames: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) of the frames rank owns; ragged counts give the first n % G ranks one extra frame."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    if n_frames < 0:
        raise ValueError("n_frames must be >= 0")
    q, r = divmod(n_frames, world_size)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def stream_indices(n_frames: int, rank: int, world_size: int) -> np.ndarray:
    """Indices i with i mod G == rank (round-robin over a stream)."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return np.arange(rank, max(n_frames, 0), world_size, dtype=np.int64)


def owned_indices(n_frames: int, rank: int, world_size: int, mode: str = "slab") -> np.ndarray:
    if mode == "slab":
        lo, hi = slab_range(n_frames, rank, world_size)
        return np.arange(lo, hi, dtype=np.int64)
    if mode == "stream":
        return stream_indices(n_frames, rank, world_size)
    raise ValueError(f"unknown partition mode {mode!r}")


def gather_frames(local: torch.Tensor, n_frames: int, mode: str = "slab", dst: int = 0):
    """Collects the per-rank output maps [n_local, H, W] on rank `dst` in original frame order.

    Ragged shards are padded to the largest shard so that ONE all_gather moves everything (NCCL all_gather needs
    equal sizes; over NVSwitch a padded frame costs microseconds).  Returns the [n_frames, H, W] tensor on `dst`
    and None elsewhere.  With a single process it returns `local` unchanged."""
    rank, G = world()
    if G == 1:
        if local.shape[0] != n_frames:
            raise ValueError("single-process gather: local shard must hold every frame")
        return local
    counts = [len(owned_indices(n_frames, r, G, mode)) for r in range(G)]
    if local.shape[0] != counts[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} frames, its shard is {counts[rank]}")
    cap = max(counts)
    if cap == 0:
        return local.new_zeros((0,) + tuple(local.shape[1:])) if rank == dst else None
    send = local
    if local.shape[0] < cap:
        send = local.new_zeros((cap,) + tuple(local.shape[1:]))
        send[: local.shape[0]] = local
    send = send.contiguous()
    bucket: List[torch.Tensor] = [torch.empty_like(send) for _ in range(G)]
    dist.all_gather(bucket, send)
    if rank != dst:
        return None
    out = local.new_empty((n_frames,) + tuple(local.shape[1:]))
    for r in range(G):
        idx = owned_indices(n_frames, r, G, mode)
        if len(idx):
            out[torch.from_numpy(idx).to(out.device)] = bucket[r][: len(idx)]
    return out


class ShardedEstimator:
    """Runs any object with `predict_batch(frames[B,H,W,3]) -> [B,H,W]` on this rank's share of a batch.

    `estimator` is a DepthEstimator on this rank's GPU in production; the CPU tests pass a stand-in, which is all
    this class needs (it never touches CUDA itself)."""

    def __init__(self, estimator, mode: str = "slab"):
        if mode not in ("slab", "stream"):
            raise ValueError(f"unknown partition mode {mode!r}")
        self.estimator = estimator
        self.mode = mode

    def predict_sharded(self, frames: Sequence[np.ndarray] | np.ndarray, gather: bool = True, dst: int = 0):
        """frames: the WHOLE batch (every rank sees the same list / array, e.g. a shared decode queue).
        Each rank computes only the frames it owns; with gather=True rank `dst` gets np.float32 [n,H,W] in the
        original order (others None), otherwise every rank gets (indices, its maps)."""
        rank, G = world()
        n = len(frames)
        idx = owned_indices(n, rank, G, self.mode)
        if len(idx):
            mine = np.ascontiguousarray(np.stack([np.asarray(frames[int(i)]) for i in idx]))
            maps = np.asarray(self.estimator.predict_batch(mine), dtype=np.float32)
        else:
            H, W = np.asarray(frames[0]).shape[:2] if n else (0, 0)
            maps = np.zeros((0, H, W), np.float32)
        if not gather:
            return idx, maps
        t = torch.from_numpy(maps)
        if G > 1 and dist.get_backend() == "nccl":
            t = t.cuda()
        out = gather_frames(t, n, self.mode, dst)
        return None if out is None else out.cpu().numpy()
