"""Times the projection kernels alone (FP32-FMA vs tcgen05) on 120 frames of triplane tokens."""
import sys, torch
sys.path.insert(0, '.')
from audio_motion_avatar_b200 import stages, synthetic as S
C, R, F = 512, 64, 120
dev = torch.device("cuda")
tok = torch.randn(F, C, 3 * R * R, device=dev)
wd = S.make_decoder_weights(C, seed=0)
W = torch.cat([wd[f"{n}_layer.weight"] for n in ("xyz", "rotation", "scaling", "opacity", "shs")], 0).to(dev)
out = torch.empty(F, 3, R * R, 16, device=dev)
for impl in ("ffma", "tc"):
    for _ in range(3):
        stages.triplane_project(tok, W, C, R, out=out, impl=impl)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        stages.triplane_project(tok, W, C, R, out=out, impl=impl)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    by = tok.numel() * 4 + out.numel() * 4
    print(impl, round(ms, 4), "ms", round(by / ms / 1e6, 1), "GB/s")
x = stages.triplane_project(tok, W, C, R, impl="tc"); y = stages.triplane_project(tok, W, C, R, impl="ffma")
print("max diff", float((x - y).abs().max()))
