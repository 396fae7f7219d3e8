import sys, torch
sys.path.insert(0, '.')
import bench
from audio_motion_avatar_b200.renderer import _workspace, deferred_overflow_checks
class A: pass
ctx = bench.Ctx()
r = bench.make_renderer(2048, 500000, ctx)
dev = ctx.dev
tok, par, cam = bench.synthetic_inputs(0, 16, 16, 2048, dev, seed=1)
par = {k: v.to(dev) for k, v in par.items()}; cam = {k: v.to(dev) for k, v in cam.items()}
out = torch.empty(15, 2048, 2048, 3, device=dev)
for it in range(3):
    with deferred_overflow_checks() as chk, torch.no_grad():
        for a, b in ((0, 15), (15, 16)):
            r(tok[:, a:b], {k: v[:, a:b] for k, v in cam.items()}, None, {k: v[:, a:b] for k, v in par.items()}, out=out[:b-a])
            ws = _workspace(dev)
            torch.cuda.synchronize()
            nt = (b - a) * 128 * 128
            print(it, (a, b), "cap", ws.capacity, "last_D", ws.last_D, "tile_start[nt]", int(ws.bufs["tile_start"][nt].item()), "flag", int(ws.overflow.item()), flush=True)
    print("chk", chk.overflowed())
