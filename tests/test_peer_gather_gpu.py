"""Frame gather by peer stores (shard.PeerFrameGather): every rank's blend kernel writes its frames into rank 0's
symmetric-memory buffer over NVLink.  Needs 2 GPUs (skipped on a 1-GPU box); the NCCL alternative
(shard.render_and_gather) and the sharding logic are covered on the CPU by tests/test_shard_gloo.py."""
import os
from types import SimpleNamespace

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from audio_motion_avatar_b200 import synthetic as S
        from audio_motion_avatar_b200.renderer import Renderer
        from audio_motion_avatar_b200.shard import PeerFrameGather, frame_shard, render_and_gather, shard_inputs
        H, W, C, R, T = 72, 64, 64, 16, 5                              # 5 frames over 2 ranks: ragged 3 + 2; W % 32 == 0: row stores
        cfg = SimpleNamespace(device=f"cuda:{rank}", triplane_resolution=R, triplane_feature_dim=C, radius=1.4,
                              smplx_model_path=None, flat_hand_mean=True, num_expression_coeffs=10, upsample_triplane=False,
                              densify_smplx_verts=False, subdivide_steps=2, predict_smplx_params=False, no_point_refiner=True,
                              image_size=(H, W), rgb=True, sh_degree=3, num_gaussians=0)
        r = Renderer(cfg, body_model=S.make_body_model(seed=0))
        r.gaussian_decoder.load_state_dict(S.make_decoder_weights(C, seed=0))
        tokens = S.make_triplane_tokens(T, C, R, seed=0).to(dev)        # the same sequence on every rank
        params = {k: v.to(dev) for k, v in S.make_pose_sequence(1, T, seed=0).items()}
        cams = {k: v.to(dev) for k, v in S.make_cameras(1, T, H, W).items()}
        with torch.no_grad():
            full, _ = r(tokens, cams, None, params)                     # all frames rendered locally: the expected result
        tk, cm, pr = shard_inputs(tokens, cams, params, rank, world)
        peer = PeerFrameGather(T, H, W, dev, dst=0)
        out = peer.render(r, tk, cm, pr)
        ok = True
        if rank == 0:
            torch.cuda.synchronize()
            ok = out is not None and out.shape == (T, H, W, 3) and torch.equal(out, full[0])
        out2 = render_and_gather(r, tk, cm, pr, T, chunks=2, dst=0)
        torch.cuda.synchronize()
        if rank == 0:
            ok = ok and out2 is not None and torch.equal(out2, full[0])
        # staged variant: local render in pieces + peer copies on a second stream
        if rank == 0:
            peer.local.zero_()
        dist.barrier()
        for _ in range(2):
            out4 = peer.render_staged(r, tk, cm, pr, chunks=2)
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            ok = ok and out4 is not None and torch.equal(out4, full[0])
        # uint8 gather (the demo's video sink): converted on the rendering rank, only the 8-bit frames cross NVLink
        from audio_motion_avatar_b200.stages import frames_to_u8
        want8 = frames_to_u8(full[0], bgr=True)
        for ch in (2, "halving"):
            out5 = peer.render_staged_u8(r, tk, cm, pr, chunks=ch, bgr=True)
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            ok = ok and out5 is not None and out5.dtype == torch.uint8 and torch.equal(out5, want8)
        else:
            assert out5 is None
        # the same shard step as one CUDA graph with the peer stores inside (PeerFrameGather.capture)
        if rank == 0:
            peer.local.zero_()
        dist.barrier()
        gf = peer.capture(r, tk, cm, pr)
        for _ in range(2):
            out3 = peer.render_captured()
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            ok = ok and out3 is not None and torch.equal(out3, full[0]) and gf.launches > 0 and not gf.overflowed()
            q.put(bool(ok))
        else:
            assert out is None and out2 is None and out3 is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_peer_and_nccl_gather_two_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(rk, 2, port, q)) for rk in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert q.get(timeout=5) is True
