"""Where does Renderer.render_sequence's time go?  Host enqueue time vs device time for several chunk sizes."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench as B
from audio_motion_avatar_b200 import synthetic as S
from audio_motion_avatar_b200.renderer import Renderer

dev = torch.device("cuda:0")
Fr, N, size = 120, 100000, 1024
r = Renderer(B.make_cfg(size, N, "cuda:0"), body_model=S.make_body_model(seed=0))
r.gaussian_decoder.load_state_dict(S.make_decoder_weights(B.C, seed=0))
r.fix_site_subset(seed=0)
tokens, params, cams = B.synthetic_inputs(Fr, size, dev, 0)
tok_h = torch.empty(tokens.shape, dtype=torch.float32, pin_memory=True); tok_h.copy_(tokens)
out_h = torch.empty(1, Fr, size, size, 3, dtype=torch.float32, pin_memory=True)
with torch.no_grad():
    r(tokens, {k: v.to(dev) for k, v in cams.items()}, None, {k: v.to(dev) for k, v in params.items()})
for chunk in (2, 4, 8, 12, 20):
    for _ in range(2):
        r.render_sequence(tok_h, cams, params, chunk=chunk, out=out_h)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        r.render_sequence(tok_h, cams, params, chunk=chunk, out=out_h)
    torch.cuda.synchronize()
    print(f"chunk {chunk:3d}: {(time.perf_counter() - t0) / 3 * 1e3:7.1f} ms per 120 frames")
# copies only, same chunking
h2d, d2h = torch.cuda.Stream(), torch.cuda.Stream()
dt = torch.empty(8, B.C, 3 * B.R * B.R, device=dev); di = torch.empty(8, size, size, 3, device=dev)
tf = tok_h.reshape(Fr, B.C, -1); of = out_h.reshape(Fr, size, size, 3)
torch.cuda.synchronize(); t0 = time.perf_counter()
for k in range(15):
    with torch.cuda.stream(h2d): dt.copy_(tf[8 * k:8 * k + 8], non_blocking=True)
    with torch.cuda.stream(d2h): of[8 * k:8 * k + 8].copy_(di, non_blocking=True)
torch.cuda.synchronize(); print("copies only (both directions, chunk 8):", (time.perf_counter() - t0) * 1e3, "ms")
# host enqueue cost of one device-resident forward
torch.cuda.synchronize(); t0 = time.perf_counter()
with torch.no_grad():
    r(tokens, {k: v.to(dev) for k, v in cams.items()}, None, {k: v.to(dev) for k, v in params.items()})
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"forward(120 frames): host enqueue {1e3 * (t1 - t0):.2f} ms, total {1e3 * (t2 - t0):.2f} ms")
