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


def chunk_bounds(size: int, chunks: int) -> List[int]:
    """chunks + 1 boundaries cutting ``size`` frames into near-equal pieces, the larger (and, for size < chunks, the
    non-empty) ones first (same rule on every rank)."""
    chunks = max(1, int(chunks))
    return [(i * size + chunks - 1) // chunks for i in range(chunks + 1)]


def halving_bounds(size: int, min_piece: int = 16, max_pieces: int = 4) -> List[int]:
    """Piece boundaries size/2, size/4, ... for a pipeline whose LAST piece is the exposed one (render-bound staged
    gathers): 120 -> [0, 60, 90, 105, 120]; at most ``max_pieces`` pieces, none cut below ``min_piece`` frames."""
    cuts, left = [0], int(size)
    while left > min_piece and len(cuts) < max_pieces:
        cuts.append(cuts[-1] + left // 2)
        left -= left // 2
    cuts.append(int(size))
    return cuts


def render_and_gather(render, tokens: torch.Tensor, cams: dict, params: dict, num_frames: int, chunks: int = 4,
                      dst: int = 0, group=None, out: Optional[torch.Tensor] = None,
                      defer=None) -> Optional[torch.Tensor]:
    """Render this rank's shard ([1, f_r, ...] inputs as ``shard_inputs`` cuts them) in ``chunks`` pieces and gather
    the frames on ``dst`` while the next piece renders: ``render(tokens, cams, None, params)`` is ``Renderer.forward``.

    The gather is what limits scaling when it runs after the render: ``dst`` takes in (world - 1) x f_r frames over
    one NVLink port (900 GB/s), 10.6 GB = 12 ms for 8 x 120 frames of 1024^2 next to an 18 ms render.  Posting the
    send of piece k as soon as it is enqueued lets NCCL's stream move it while the main stream renders piece k+1,
    which leaves only the last piece's transfer exposed.  ``defer``: ``renderer.deferred_overflow_checks`` -- the
    pieces then skip their per-call overflow read (a host sync each); the flags are read once at the end, combined
    over the ranks, and on an overflow every rank renders and sends again with the checks on.
    Returns [num_frames, H, W, 3] on ``dst``, None elsewhere."""
    if defer is not None:
        with defer() as chk:
            res = render_and_gather(render, tokens, cams, params, num_frames, chunks, dst, group, out, None)
        flag = torch.tensor([1.0 if chk.overflowed() else 0.0], device=tokens.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=group)
        if flag.item() > 0:
            res = render_and_gather(render, tokens, cams, params, num_frames, 1, dst, group, res if res is not None else out, None)
        return res
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = shard_sizes(num_frames, world)
    if tokens.shape[1] != sizes[rank]:
        raise ValueError(f"rank {rank} holds {tokens.shape[1]} frames, its shard has {sizes[rank]}")
    cuts = [chunk_bounds(sz, chunks) for sz in sizes]
    starts = [frame_shard(num_frames, r, world)[0] for r in range(world)]
    reqs, keep = [], []
    for k in range(max(1, int(chunks))):
        a, b = cuts[rank][k], cuts[rank][k + 1]
        img = None
        if b > a:
            with torch.no_grad():
                img = render(tokens[:, a:b], {n: v[:, a:b] for n, v in cams.items()}, None,
                             {n: v[:, a:b] for n, v in params.items()})[0][0]
        if rank == dst:
            if out is None:
                if img is None:                                  # rank dst holds no frame at all: nobody does
                    break
                out = torch.empty((num_frames,) + tuple(img.shape[1:]), dtype=img.dtype, device=img.device)
            if img is not None:
                out[starts[rank] + a:starts[rank] + b].copy_(img)
            ops = [dist.P2POp(dist.irecv, out[starts[r] + cuts[r][k]:starts[r] + cuts[r][k + 1]], r, group)
                   for r in range(world) if r != dst and cuts[r][k + 1] > cuts[r][k]]
            if ops:
                reqs += dist.batch_isend_irecv(ops)              # one grouped call: all peers' pieces move concurrently
        elif img is not None:
            img = img.contiguous()
            keep.append(img)                                     # alive until the send has completed
            reqs.append(dist.isend(img, dst, group=group))
    for q in reqs:
        q.wait()
    return out if rank == dst else None


class PeerFrameGather:
    """Gather by construction: the destination rank's frame buffer is symmetric memory mapped into every rank of
    the box, and each rank's blend kernel stores its pixels straight into its slice of that buffer over
    NVLink / NVSwitch (plain global stores to a peer address).  There is no separate gather step, no second copy of
    the frames and no staging buffer: the transfer is spread over the 12 ms the blend kernel runs (1.5 GB per rank
    = 130 GB/s per sender at 120 frames of 1024^2) instead of following it.  One device-side barrier at the end
    tells the destination that every rank's stores have landed.

    Needs the NCCL backend and peer access between the GPUs (one NVSwitch box); ``available()`` says whether the
    symmetric-memory allocator is usable, and ``render_and_gather`` is the NCCL send/recv alternative."""

    def __init__(self, num_frames: int, height: int, width: int, device, dst: int = 0, group=None):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank, self.dst = dist.get_world_size(group), dist.get_rank(group), dst
        self.num_frames = num_frames
        self.shape = (num_frames, height, width, 3)
        self.local = symm.empty(self.shape, dtype=torch.float32, device=device)       # same size on every rank
        self.handle = symm.rendezvous(self.local, self.group)
        self.target = self.handle.get_buffer(dst, self.shape, torch.float32)          # dst's buffer, as seen from here
        self.start, self.stop = frame_shard(num_frames, self.rank, self.world)

    @staticmethod
    def available() -> bool:
        try:
            import torch.distributed._symmetric_memory as symm  # noqa: F401
            return dist.is_initialized() and dist.get_backend() == "nccl" and torch.cuda.is_available()
        except Exception:
            return False

    def render(self, render, tokens: torch.Tensor, cams: dict, params: dict) -> Optional[torch.Tensor]:
        """``render`` = ``Renderer.forward``; inputs are this rank's shard ([1, f_r, ...]).  Returns the whole
        sequence [num_frames, H, W, 3] on ``dst`` (valid once the current stream has passed the barrier), None
        elsewhere."""
        if tokens.shape[1] != self.stop - self.start:
            raise ValueError(f"rank {self.rank} holds {tokens.shape[1]} frames, its shard has {self.stop - self.start}")
        if self.stop > self.start:
            with torch.no_grad():
                render(tokens, cams, None, params, out=self.target[self.start:self.stop])
        self.handle.barrier()                      # every rank's kernels (and their peer stores) are complete
        return self.local if self.rank == self.dst else None

    def render_staged(self, render, tokens: torch.Tensor, cams: dict, params: dict, chunks: int = 4) -> Optional[torch.Tensor]:
        """The same gather with the transfer taken OFF the blend kernel: the shard is rendered in ``chunks`` pieces into
        local staging buffers (per-pixel stores, the fast local kernel) and every finished piece is pushed into the
        destination's buffer by a peer copy on a second stream while the next piece renders.  Seven senders storing
        384-byte row segments from inside the blend kernel reach ~520 GB/s into rank 0; plain peer copies reach ~795 GB/s
        (tools/ingress_probe.py), and only the last piece's copy is exposed."""
        n = self.stop - self.start
        if tokens.shape[1] != n:
            raise ValueError(f"rank {self.rank} holds {tokens.shape[1]} frames, its shard has {n}")
        dev = tokens.device
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(dev)
            self._staging = [None, None]
            self._staged_free = [None, None]
        cuts = chunk_bounds(n, chunks)
        if self.rank == self.dst:
            # the destination's own frames need no transfer: rendered straight into its buffer
            if n:
                with torch.no_grad():
                    render(tokens, cams, None, params, out=self.local[self.start:self.stop])
        else:
            piece = max((cuts[i + 1] - cuts[i] for i in range(len(cuts) - 1)), default=0)
            for k in range(len(cuts) - 1):
                a, b = cuts[k], cuts[k + 1]
                if b <= a:
                    continue
                slot = k & 1
                if self._staging[slot] is None or self._staging[slot].shape[0] < piece:
                    self._staging[slot] = torch.empty((piece,) + self.shape[1:], dtype=torch.float32, device=dev)
                if self._staged_free[slot] is not None:
                    main.wait_event(self._staged_free[slot])           # the copy of piece k - 2 has drained this buffer
                stage = self._staging[slot][:b - a]
                with torch.no_grad():
                    render(tokens[:, a:b], {m: v[:, a:b] for m, v in cams.items()}, None, {m: v[:, a:b] for m, v in params.items()},
                           out=stage)
                done = torch.cuda.Event()
                done.record(main)
                with torch.cuda.stream(self._copy_stream):
                    self._copy_stream.wait_event(done)
                    self.target[self.start + a:self.start + b].copy_(stage, non_blocking=True)
                    freed = torch.cuda.Event()
                    freed.record(self._copy_stream)
                self._staged_free[slot] = freed
            main.wait_stream(self._copy_stream)
        self.handle.barrier()
        return self.local if self.rank == self.dst else None

    def enable_u8(self) -> None:
        """Allocate the uint8 destination buffer (symmetric memory: a collective call, every rank makes it)."""
        if getattr(self, "local_u8", None) is None:
            import torch.distributed._symmetric_memory as symm
            self.local_u8 = symm.empty(self.shape, dtype=torch.uint8, device=self.local.device)
            self.handle_u8 = symm.rendezvous(self.local_u8, self.group)
            self.target_u8 = self.handle_u8.get_buffer(self.dst, self.shape, torch.uint8)

    def render_staged_u8(self, render, tokens: torch.Tensor, cams: dict, params: dict, chunks=4,
                         bgr: bool = False) -> Optional[torch.Tensor]:
        """``render_staged`` for a consumer that wants 8-bit frames (the demo's video sink, main2.py:226-239,332-363:
        ``(frame * 255).astype(np.uint8)``, optionally RGB -> BGR): every piece is converted on the rank that rendered it
        (amav_frames_to_u8) and only the uint8 frames cross NVLink -- 3.1 MB per 1024^2 frame into the destination instead
        of 12.6 MB, which takes the destination's ingress (the limiter of the fp32 gather at 8 GPUs) off the critical
        path.  Returns uint8 [num_frames, H, W, 3] on ``dst``, None elsewhere.  Opt-in: the fp32 gather stays the default."""
        from .stages import frames_to_u8
        self.enable_u8()
        n = self.stop - self.start
        if tokens.shape[1] != n:
            raise ValueError(f"rank {self.rank} holds {tokens.shape[1]} frames, its shard has {n}")
        dev = tokens.device
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(dev)
            self._staging = [None, None]
            self._staged_free = [None, None]
        if getattr(self, "_staging_u8", None) is None:
            self._staging_u8 = [None, None]
            self._staged_u8_free = [None, None]
        # ``chunks``: a number of equal pieces, or "halving": n/2, n/4, ... frames -- the render is the bottleneck of this
        # variant, so only the LAST piece's copy is exposed and it should be the smallest
        cuts = halving_bounds(n) if chunks == "halving" else chunk_bounds(n, chunks)
        piece = max((cuts[i + 1] - cuts[i] for i in range(len(cuts) - 1)), default=0)
        if piece and (self._staging[0] is None or self._staging[0].shape[0] < piece):
            self._staging[0] = torch.empty((piece,) + self.shape[1:], dtype=torch.float32, device=dev)
        for k in range(len(cuts) - 1):
            a, b = cuts[k], cuts[k + 1]
            if b <= a:
                continue
            stage = self._staging[0][:b - a]                  # fp32 piece: consumed by the conversion on the same stream
            with torch.no_grad():
                render(tokens[:, a:b], {m: v[:, a:b] for m, v in cams.items()}, None, {m: v[:, a:b] for m, v in params.items()},
                       out=stage)
            if self.rank == self.dst:
                frames_to_u8(stage, bgr=bgr, out=self.local_u8, out_first=self.start + a)
                continue
            slot = k & 1
            if self._staging_u8[slot] is None or self._staging_u8[slot].shape[0] < piece:
                self._staging_u8[slot] = torch.empty((piece,) + self.shape[1:], dtype=torch.uint8, device=dev)
            if self._staged_u8_free[slot] is not None:
                main.wait_event(self._staged_u8_free[slot])
            st8 = self._staging_u8[slot][:b - a]
            frames_to_u8(stage, bgr=bgr, out=st8)
            done = torch.cuda.Event()
            done.record(main)
            with torch.cuda.stream(self._copy_stream):
                self._copy_stream.wait_event(done)
                self.target_u8[self.start + a:self.start + b].copy_(st8, non_blocking=True)
                freed = torch.cuda.Event()
                freed.record(self._copy_stream)
            self._staged_u8_free[slot] = freed
        main.wait_stream(self._copy_stream)
        self.handle_u8.barrier()
        return self.local_u8 if self.rank == self.dst else None

    def capture(self, renderer, tokens: torch.Tensor, cams: dict, params: dict):
        """CUDA-graph this rank's shard render with the peer stores inside the graph (``Renderer.capture`` with
        ``out`` = this rank's slice of the destination buffer).  A 15-frame shard (120 frames over 8 GPUs) is ~2 ms of
        device work behind ~15 launches plus glue: as a graph it is one launch.  ``render_captured()`` replays it and
        runs the barrier (the barrier stays outside the graph: it is a cross-rank wait)."""
        if tokens.shape[1] != self.stop - self.start:
            raise ValueError(f"rank {self.rank} holds {tokens.shape[1]} frames, its shard has {self.stop - self.start}")
        self._graph = renderer.capture(tokens, cams, params, out=self.target[self.start:self.stop]) \
            if self.stop > self.start else None
        return self._graph

    def render_captured(self) -> Optional[torch.Tensor]:
        if self._graph is not None:
            self._graph.replay()
        self.handle.barrier()
        return self.local if self.rank == self.dst else None


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
