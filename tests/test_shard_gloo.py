"""Frame sharding + gather over a world_size-2 (and 3, ragged) gloo group on the CPU."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from audio_motion_avatar_b200.shard import frame_shard, shard_sizes


def test_frame_shard_partitions_every_frame_once():
    for F in (0, 1, 7, 120, 1800):
        for world in (1, 2, 3, 4, 8):
            cover = []
            for r in range(world):
                s, e = frame_shard(F, r, world)
                assert 0 <= s <= e <= F
                cover += list(range(s, e))
            assert cover == list(range(F))
            sz = shard_sizes(F, world)
            assert sum(sz) == F and max(sz) - min(sz) <= 1
    with pytest.raises(ValueError):
        frame_shard(10, 2, 2)


def _worker(rank, world, port, F, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from audio_motion_avatar_b200.shard import gather_frames, shard_inputs
        tokens = torch.arange(F * 6, dtype=torch.float32).reshape(1, F, 2, 3)
        cams = {"intrinsic": torch.arange(F * 9, dtype=torch.float32).reshape(1, F, 3, 3)}
        params = {"transl": torch.arange(F * 3, dtype=torch.float32).reshape(1, F, 3)}
        tk, cm, pr = shard_inputs(tokens, cams, params, rank, world)
        s, e = frame_shard(F, rank, world)
        assert tk.shape[1] == e - s and torch.equal(pr["transl"][0], params["transl"][0, s:e])
        # "render": frame f -> an image filled with f
        local = torch.stack([torch.full((4, 5, 3), float(f)) for f in range(s, e)]) if e > s else torch.empty(0, 4, 5, 3)
        out = gather_frames(local, F, dst=0)
        # the pipelined variant: "render" pieces of the shard and gather them as they come (3 pieces, ragged)
        from audio_motion_avatar_b200.shard import render_and_gather
        calls = []

        def fake_render(tok, cm2, _none, pr2):
            calls.append(tok.shape[1])
            frames = pr2["transl"][0, :, 0] / 3.0                       # transl[f] = (3f, 3f+1, 3f+2) -> frame index
            return torch.stack([torch.full((4, 5, 3), float(v)) for v in frames])[None], None

        out2 = render_and_gather(fake_render, tk, cm, pr, F, chunks=3, dst=0)
        if rank == 0:
            ok = out is not None and out.shape == (F, 4, 5, 3) and all(torch.all(out[f] == f) for f in range(F))
            ok = ok and out2 is not None and out2.shape == (F, 4, 5, 3) and all(torch.all(out2[f] == f) for f in range(F))
            q.put(bool(ok))
        else:
            assert out is None and out2 is None
        assert sum(calls) == e - s
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,F", [(2, 120), (3, 7), (2, 1)])
def test_gather_frames_gloo(world, F):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + world * 17 + F) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, F, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True


def _grad_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from audio_motion_avatar_b200.shard import allreduce_gradients, view_shard
        # a "loss" that is a sum over views of w * view_weight: rank r owns views [s, e)
        views = 5
        s, e = view_shard(views, rank, world)
        w = torch.arange(6, dtype=torch.float32).reshape(2, 3).requires_grad_(True)
        b = torch.ones(4, requires_grad=True)
        loss = sum((w * (v + 1)).sum() + (b * (v + 1) ** 2).sum() for v in range(s, e))
        if e > s:
            loss.backward()
        else:
            w.grad, b.grad = torch.zeros_like(w), torch.zeros_like(b)
        allreduce_gradients([w.grad, None, b.grad])
        ok = torch.allclose(w.grad, torch.full((2, 3), float(sum(v + 1 for v in range(views))))) and \
            torch.allclose(b.grad, torch.full((4,), float(sum((v + 1) ** 2 for v in range(views)))))
        q.put(bool(ok))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_allreduce_gradients_view_sharded_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world = 2
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_grad_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True and q.get(timeout=10) is True
