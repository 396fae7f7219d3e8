"""Frame sharding of a sequence across the ranks of one box: frames are independent given
(triplane, pose, camera), so rank r renders a contiguous block and the only exchange is the gather of finished
frames (NCCL over NVLink on GPUs; the same code runs over gloo in the CPU tests)."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def frame_shard(num_frames: int, rank: int, world: int) -> Tuple[int, int]:
    """[start, stop) of rank's block: the first ``num_frames % world`` ranks get one extra frame."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, extra = divmod(num_frames, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_sizes(num_frames: int, world: int) -> List[int]:
    return [frame_shard(num_frames, r, world)[1] - frame_shard(num_frames, r, world)[0] for r in range(world)]


def gather_frames(local: torch.Tensor, num_frames: int, dst: int = 0, group=None,
                  out: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """local [f_r, H, W, 3] -> on ``dst`` the whole sequence [num_frames, H, W, 3] in frame order (None elsewhere).
    Ragged shards are sent point-to-point into their slice of the output (no padding, no extra copy)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = shard_sizes(num_frames, world)
    if local.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} frames, its shard has {sizes[rank]}")
    if rank != dst:
        if sizes[rank]:
            dist.send(local.contiguous(), dst, group=group)
        return None
    if out is None:
        out = torch.empty((num_frames,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    reqs = []
    for r in range(world):
        s, e = frame_shard(num_frames, r, world)
        if r == dst:
            out[s:e].copy_(local)
        elif e > s:
            reqs.append(dist.irecv(out[s:e], r, group=group))
    for q in reqs:
        q.wait()
    return out


def shard_inputs(tokens: torch.Tensor, cams: dict, params: dict, rank: int, world: int):
    """Slice [B=1, T, ...] sequence inputs to this rank's frames."""
    T = tokens.shape[1]
    s, e = frame_shard(T, rank, world)
    cut = lambda t: t[:, s:e]
    return cut(tokens), {k: cut(v) for k, v in cams.items()}, {k: cut(v) for k, v in params.items()}


# ------------------------------------------------------------------------------------------------
# training: the (subject, view) render jobs are split over ranks by VIEW; every rank decodes
# the same subjects' Gaussians, renders its own views, and the gradients that the views share -- dL/dtriplane tokens
# and the decoder's dL/dW, dL/db -- are summed with ONE all-reduce over a flat bucket.
# ------------------------------------------------------------------------------------------------
def view_shard(num_views: int, rank: int, world: int) -> Tuple[int, int]:
    """[start, stop) of the views rank renders (same rule as frame_shard)."""
    return frame_shard(num_views, rank, world)


def allreduce_gradients(tensors: List[Optional[torch.Tensor]], group=None, average: bool = False) -> None:
    """Sum (or average) the given gradient tensors over the ranks in place: one flat bucket, one collective.
    ``None`` entries are skipped (all ranks must pass the same structure)."""
    live = [t for t in tensors if t is not None]
    if not live or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    flat = torch.cat([t.reshape(-1) for t in live])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    if average:
        flat /= dist.get_world_size(group)
    off = 0
    for t in live:
        n = t.numel()
        t.copy_(flat[off:off + n].view_as(t))
        off += n
