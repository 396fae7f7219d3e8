"""Host-side logic of the wrappers that needs no GPU: argument validation, the decoder-head packing of the 6D pose
entry, and that every product entry refuses CPU tensors instead of falling back (there is no CPU path)."""
import pytest
import torch

from audio_motion_avatar_b200 import _lib, stages


def test_pose_6d_from_decoder_heads_follows_smplx_joint_order():
    """SMPLX.forward order: global, 21 body, jaw, leye, reye, 15 left hand, 15 right hand; SMPLXDecoder's hand head holds
    the left hand first (smplx_decoder.py:114-115)."""
    Fr = 3
    root = torch.full((Fr, 6), 1.0)
    body = torch.arange(21 * 6, dtype=torch.float32).repeat(Fr, 1) + 100
    hand = torch.arange(30 * 6, dtype=torch.float32).repeat(Fr, 1) + 1000
    jaw, leye, reye = torch.full((Fr, 6), 2.0), torch.full((Fr, 6), 3.0), torch.full((Fr, 6), 4.0)
    p = stages.pose_6d_from_decoder_heads(root, body, hand, jaw, leye, reye)
    assert p.shape == (Fr, 55, 6) and p.is_contiguous()
    assert torch.equal(p[:, 0], root) and torch.equal(p[:, 1:22].reshape(Fr, -1), body)
    assert torch.equal(p[:, 22], jaw) and torch.equal(p[:, 23], leye) and torch.equal(p[:, 24], reye)
    assert torch.equal(p[:, 25:40].reshape(Fr, -1), hand[:, :90]) and torch.equal(p[:, 40:].reshape(Fr, -1), hand[:, 90:])


def test_frames_to_u8_validates_arguments_and_refuses_cpu_tensors():
    fr = torch.rand(2, 8, 8, 3)
    with pytest.raises(ValueError):
        stages.frames_to_u8(fr[..., :2])                                  # not 3 channels
    with pytest.raises(ValueError):
        stages.frames_to_u8(fr.double())
    with pytest.raises(ValueError):
        stages.frames_to_u8(fr, out_first=1)                              # placement needs an explicit buffer
    with pytest.raises(ValueError):
        stages.frames_to_u8(fr, out=torch.empty(2, 8, 8, 3, dtype=torch.uint8), out_first=0, out_step=2)   # frame 1 would land at index 2
    with pytest.raises(ValueError):
        stages.frames_to_u8(fr, out=torch.empty(2, 8, 8, 3, dtype=torch.float32))
    with pytest.raises(_lib.AmavError):
        stages.frames_to_u8(fr)                                           # CPU tensor: no CPU path


def test_triplane_project_validates_layout_and_refuses_cpu_tensors():
    C, R = 64, 16
    w = torch.zeros(14, 3 + 3 * C)
    with pytest.raises(ValueError):
        stages.triplane_project(torch.zeros(1, C, 3 * R * R), w, C, R, token_major=True)   # channel-major given as token-major
    with pytest.raises(ValueError):
        stages.triplane_project(torch.zeros(1, 3 * R * R, C), w, C, R)
    with pytest.raises(_lib.AmavError):
        stages.triplane_project(torch.zeros(1, 3 * R * R, C), w, C, R, token_major=True)
    with pytest.raises(_lib.AmavError):
        stages.triplane_project(torch.zeros(1, C, 3 * R * R), w, C, R)


def test_every_header_symbol_has_a_ctypes_signature():
    import os
    import re
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "amav.h")).read()
    declared = set(re.findall(r"\b(amav_[a-z0-9_]+)\s*\(", src))
    special = {"amav_last_error", "amav_version", "amav_launch_count", "amav_splat_scan_tmp_elems", "amav_splat_scan_order_offset",
               "amav_triplane_project_tc_scratch_elems"}
    assert declared - special == set(_lib._SIGS), (declared - special) ^ set(_lib._SIGS)
    for name in ("amav_lbs_forward_6d", "amav_lbs_backward_6d", "amav_triplane_project_tm", "amav_triplane_project_backward_tm",
                 "amav_frames_to_u8"):
        assert name in declared


def test_training_jobs_split_covers_every_render_once():
    """bench.train_jobs: the 16 (subject, view) jobs of configs[3] over 1/2/4/8/16 ranks -- contiguous subject-major blocks,
    every job exactly once, equal shares."""
    import bench
    B, Tv = 4, 4
    for world in (1, 2, 4, 8, 16):
        seen = []
        for rank in range(world):
            subj, views = bench.train_jobs(rank, world, B, Tv)
            assert subj == list(range(subj[0], subj[-1] + 1)) and views == list(range(views[0], views[-1] + 1))
            seen += [(s, v) for s in subj for v in views]
            assert len(subj) * len(views) == B * Tv // world
        assert sorted(seen) == [(s, v) for s in range(B) for v in range(Tv)]
    assert bench.train_jobs(3, 8, B, Tv, per_rank_all=True) == (list(range(B)), list(range(Tv)))


def test_strong_split_frames_are_the_same_sequence():
    """bench.synthetic_inputs: frames [f0, f1) cut by any rank are the frames of the one 120-frame sequence (poses and
    cameras; the tokens are seeded per frame index on the device, covered by the N > 1 gather_check)."""
    import bench
    import torch
    from audio_motion_avatar_b200 import synthetic as S
    from audio_motion_avatar_b200.shard import frame_shard
    whole = S.make_pose_sequence(1, 120, seed=0)
    for world in (2, 8):
        for rank in (0, world - 1):
            f0, f1 = frame_shard(120, rank, world)
            part = {k: v[:, f0:f1] for k, v in S.make_pose_sequence(1, 120, seed=0).items()}
            assert all(torch.equal(part[k], whole[k][:, f0:f1]) for k in whole)
    assert sum(frame_shard(120, r, 8)[1] - frame_shard(120, r, 8)[0] for r in range(8)) == 120


def test_halving_piece_schedule_covers_the_shard_once():
    """shard.halving_bounds: the piece schedule of the uint8 staged gather -- increasing bounds from 0 to n, pieces shrink,
    the last piece is the smallest, no empty piece except for an empty shard."""
    from audio_motion_avatar_b200.shard import halving_bounds
    assert halving_bounds(120) == [0, 60, 90, 105, 120]
    assert halving_bounds(60) == [0, 30, 45, 60]
    for n in (0, 1, 2, 15, 16, 17, 33, 225, 1800):
        cuts = halving_bounds(n)
        assert cuts[0] == 0 and cuts[-1] == n and len(cuts) <= 5
        pieces = [b - a for a, b in zip(cuts[:-1], cuts[1:])]
        assert sum(pieces) == n and (n == 0 or all(p > 0 for p in pieces))
        assert pieces == sorted(pieces, reverse=True) or pieces[-1] - pieces[-2] <= 1      # non-increasing up to the odd remainder
