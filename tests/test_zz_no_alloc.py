"""amav_triplane_project_tc_static: the allocation-free variant of the tensor-core triplane projection (split-weight scratch
in a file-scope device array of the library, include/amav.h) against the existing amav_triplane_project_tc with
caller-owned scratch, through the C ABI, on two calls in a row that share the static array.

No counterpart in the reference.  Collected last on purpose
(file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 2e-5          # the tolerance tests/test_stages_gpu.py uses for the tensor-core projection (the two paths run the same
                    # kernels on the same split weights, so the expected difference is 0)


def _inputs(F, C, R, seed):
    from audio_motion_avatar_b200 import synthetic as S
    g = torch.Generator().manual_seed(seed)
    tokens = torch.randn(F, C, 3 * R * R, generator=g).cuda()
    wd = S.make_decoder_weights(C, seed=seed)
    W = torch.cat([wd[f"{n}_layer.weight"] for n in ("xyz", "rotation", "scaling", "opacity", "shs")], 0).contiguous().cuda()
    return tokens, W


def _existing(tokens, W, F, C, R):
    from audio_motion_avatar_b200._lib import call, ptr, stream_ptr
    out = torch.full((F, 3, R * R, 16), float("nan"), device="cuda")
    scratch = torch.empty(3 * (C // 32) * 1024, device="cuda")
    call("amav_triplane_project_tc", F, C, R, ptr(tokens), C * 3 * R * R, ptr(W), ptr(scratch), ptr(out), stream_ptr())
    return out


def _static(tokens, W, F, C, R):
    from audio_motion_avatar_b200._lib import call, ptr, stream_ptr
    out = torch.full((F, 3, R * R, 16), float("nan"), device="cuda")
    call("amav_triplane_project_tc_static", F, C, R, ptr(tokens), C * 3 * R * R, ptr(W), ptr(out), stream_ptr())
    return out


def test_static_scratch_projection_matches_caller_scratch_on_two_calls_in_a_row():
    # call 1: the headline shape (C = 512 fills the whole static array); call 2: other weights and a smaller C, which
    # reuses a prefix of the array that still holds call 1's split weights
    cases = [(2, 512, 64, 11), (3, 128, 32, 12)]
    prepared = [(_inputs(F, C, R, seed), (F, C, R)) for F, C, R, seed in cases]
    want = [_existing(tok, W, *dims) for (tok, W), dims in prepared]
    torch.cuda.synchronize()
    got = [_static(tok, W, *dims) for (tok, W), dims in prepared]          # two calls in a row, same stream
    torch.cuda.synchronize()
    for a, b in zip(got, want):
        assert torch.isfinite(a).all() and torch.isfinite(b).all()
        assert float(b.abs().max()) > 0.1
        assert float((a - b).abs().max()) <= TOL
    # and once more in the opposite order (small C first, then the full array again)
    again = [_static(tok, W, *dims) for (tok, W), dims in reversed(prepared)]
    torch.cuda.synchronize()
    for a, b in zip(again, reversed(want)):
        assert float((a - b).abs().max()) <= TOL


def test_static_scratch_projection_rejects_sizes_beyond_the_maximum():
    from audio_motion_avatar_b200._lib import AmavError, call, ptr, stream_ptr
    C, R, F = 544, 16, 1                                                    # a multiple of 32 above AMAV_TC_STATIC_MAX_C = 512
    tokens = torch.zeros(F, C, 3 * R * R, device="cuda")
    W = torch.zeros(14, 3 + 3 * C, device="cuda")
    out = torch.zeros(F, 3, R * R, 16, device="cuda")
    with pytest.raises(AmavError, match="512"):
        call("amav_triplane_project_tc_static", F, C, R, ptr(tokens), C * 3 * R * R, ptr(W), ptr(out), stream_ptr())
