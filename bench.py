#!/usr/bin/env python
"""bench.py -- rendered frames/s of the render path (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Headline (`value`): BASELINE.json configs[1] -- a 120-frame audio-driven sequence (one 3x64x64x512 fp32 triplane per
frame), 100k Gaussians on the synthetic SMPL-X-shaped body (55 joints, level-2 densified), 1024x1024.  One STEP = one pass
of the whole path (LBS -> densify -> triplane project -> sample/decode -> splat) over the sequence.  N > 1 is configs[2] AS
WRITTEN: the SAME 120 frames cut into N contiguous shards ("scaling": "strong"; frames are independent, no data-path
collective), every rank's blend kernel storing its frames straight into rank 0's buffer over NVLink (frame gather).
The 120-frames-PER-RANK run of round 1 is kept under the key `weak`.

Other keys of the line (all measured in the same process, after the headline's timed region):
  `graph`        the same step as ONE CUDA-graph launch per rank (Renderer.capture);
  `e2e`          frames/s through Renderer.render_sequence with HOST (pinned) triplanes in and HOST frames out;
  `parity`       frame 0 of the workload on the GPU against the CPU oracle on the same inputs (N = 1);
  `latency_T6`   one Renderer.forward call at the reference's native clip length T = 6, eager and as a CUDA graph;
  `train`        BASELINE.json configs[3]: 4 subjects x 4 views, 100k Gaussians, 512x512, forward + loss + backward + SGD;
                 N > 1: the 16 (subject, view) renders are split over the ranks, gradients all-reduced inside the step;
  `stress`       BASELINE.json configs[4] sizes: 500k Gaussians, 2048x2048; 1800/8 = 225 frames per rank at N = 8, 16 else;
  `cpu_baseline` the oracle port timed on the host cores (N = 1);  `roofline` per-stage table + dominant kernel.
`--impl reference`: the reference's CPU implementation of the path (oracle port: torch CPU ops for LBS / grid_sample /
Linear heads exactly as renderer.py calls them, C rasteriser -O3 with OpenMP), all host threads, one frame per step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch

METRIC = "rendered frames/sec @100k Gaussians 1024^2"
UNIT = "frames/s"
FRAMES, NGAUSS, SIZE, C, R = 120, 100_000, 1024, 512, 64


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=FRAMES)
    ap.add_argument("--gaussians", type=int, default=NGAUSS)
    ap.add_argument("--size", type=int, default=SIZE)
    ap.add_argument("--chunk", type=int, default=4, help="frames per pipeline chunk of the e2e path")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gather", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline only: no graph / parity / latency_T6 / train / stress / weak blocks")
    ap.add_argument("--only", default="", help="comma list of secondary blocks to keep (graph,parity,latency,train,stress,weak,sink,yardsticks)")
    ap.add_argument("--gather-chunks", type=int, default=0,
                    help="N > 1, --gather nccl: pieces the shard is rendered and gathered in (0: 1 for N <= 2, else 4)")
    ap.add_argument("--gather", default="auto", choices=["auto", "staged", "peer", "nccl"],
                    help="N > 1: peer = the blend kernels store into rank 0's symmetric buffer themselves (one launch sequence per "
                         "shard); staged = local render in >= 30-frame pieces + peer copies on a second stream (many senders, long "
                         "shards: 7 -> 1 row-segment stores reach ~520 GB/s, plain peer copies ~795 GB/s); auto = staged when "
                         "N >= 8 and a shard has >= 60 frames, else peer; nccl = NCCL send/recv")
    ap.add_argument("--workload", default="sequence", choices=["sequence", "train"],
                    help="sequence = the default line; train = configs[3] alone as its own line")
    ap.add_argument("--loss", default="l1_ssim", choices=["l1_ssim", "l1_torch"],
                    help="train workload: fused l1 + 0.1 (1 - ssim) kernels (lightning_model_wrapper.py:138-140) or a plain torch L1")
    ap.add_argument("--subjects", type=int, default=4)
    ap.add_argument("--views", type=int, default=4)
    ap.add_argument("--per-rank-16", action="store_true", help="train: every rank renders all 16 (subject, view) jobs (round-1 weak variant)")
    return ap.parse_args()


def subdivide_steps(ngauss):
    """Level-2 subdivision of the 10,475-vertex body has 167,420 vertices; the 500k-Gaussian stress configuration
    (BASELINE.json configs[4]) draws its sites from level 3 (669,644)."""
    return 2 if ngauss <= 167_420 else 3


def make_cfg(size, ngauss, device):
    return SimpleNamespace(device=device, triplane_resolution=R, triplane_feature_dim=C, radius=1.4,
                           smplx_model_path=None, flat_hand_mean=True, num_expression_coeffs=10,
                           upsample_triplane=False, densify_smplx_verts=ngauss != 10475, subdivide_steps=subdivide_steps(ngauss),
                           predict_smplx_params=False, no_point_refiner=True, use_gaussian_splatting=True,
                           image_size=(size, size), rgb=True, sh_degree=3, num_gaussians=ngauss)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def synthetic_tokens(f0, f1, device, seed=0):
    """Frames [f0, f1) of the synthetic sequence, device-resident: frame f = base + 0.05 * noise_f with
    noise_f drawn from a generator seeded by (seed, f), so that any rank can produce any frame of the SAME sequence."""
    g = torch.Generator(device=device).manual_seed(seed)
    base = torch.randn(C, 3 * R * R, device=device, generator=g)
    tokens = torch.empty(1, f1 - f0, C, 3 * R * R, device=device)
    for f in range(f0, f1):
        g.manual_seed(seed * 1_000_003 + 7919 * (f + 1))
        tokens[0, f - f0] = base + 0.05 * torch.randn(C, 3 * R * R, device=device, generator=g)
    return tokens


def synthetic_inputs(f0, f1, total, size, device, seed=0):
    """(tokens, params, cams) of frames [f0, f1) of a `total`-frame sequence: smooth poses over the whole sequence,
    E = I cameras (the TED loader's, dataset_speech_vid.py:314-317)."""
    from audio_motion_avatar_b200 import synthetic as S
    params = {k: v[:, f0:f1].contiguous() for k, v in S.make_pose_sequence(1, total, seed=seed).items()}
    cams = {k: v[:, f0:f1].contiguous() for k, v in S.make_cameras(1, total, size, size).items()}
    return synthetic_tokens(f0, f1, device, seed), params, cams


def stage_bytes(name, Fr, N, V, J, S, P, H, W, D):
    """ALGORITHMIC HBM bytes of one C-ABI stage call over Fr frames (SURVEY.md 8d / DESIGN.md)."""
    R2 = R * R
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    if name == "amav_triplane_project":
        return Fr * (3 * C * R2 * 4 + 3 * R2 * 64) + 14 * (3 * C + 3) * 4
    if name == "amav_sample_decode":
        return Fr * (N * (12 + 56) + 3 * R2 * 64)
    if name == "amav_lbs_forward":
        const = (J - 1) * 9 * 3 * V * 4 + V * 3 * S * 4 + J * V * 4 + V * 12
        return const + Fr * (J * 12 + S * 4 + 2 * V * 12 + 2 * V * 12)
    if name == "amav_densify_gather":
        return N * P * 8 + Fr * (V * 12 + N * 12)
    if name == "amav_splat_preprocess":
        return Fr * N * (56 + 48 + 8 + 4) + Fr * tiles * 4
    if name == "amav_splat_scan":
        return Fr * tiles * 8
    if name == "amav_splat_bin":
        return Fr * N * (8 + 4) + D * (8 + 8 + 4) + Fr * tiles * 8
    if name == "amav_splat_blend":
        return D * 48 + Fr * H * W * 12                     # 48 D + 12 H W per frame
    # backward stages (Fr = frames the stage sees; for decode / LBS stages that is the number of Gaussian sets)
    if name == "amav_splat_blend_backward":
        return D * 48 + Fr * H * W * (12 + 8) + Fr * N * 48
    if name == "amav_splat_preprocess_backward":
        return Fr * N * (16 + 36) + (Fr // max(P, 1)) * N * (56 + 56)          # P = views per set here
    if name == "amav_sample_decode_backward":
        return Fr * (N * (12 + 56 + 12) + 2 * 3 * R2 * 64)
    if name == "amav_triplane_project_backward":
        return Fr * (2 * 3 * C * R2 * 4 + 2 * 3 * R2 * 64) + 14 * (3 * C + 3) * 4
    if name == "amav_lbs_backward":
        const = (J - 1) * 9 * 3 * V * 4 + V * 3 * S * 4 + J * V * 4
        return const + Fr * (4 * V * 12 + J * 48 * 2)
    if name == "amav_densify_gather_backward":
        return N * P * 8 + Fr * (V * 12 + N * 12)
    if name == "amav_camera_setup":
        return Fr * (36 + 64 + 64 + 64 + 8)
    if name in ("amav_image_loss_forward", "amav_image_loss_backward"):
        return Fr * H * W * 3 * 4 * (2 + 3) + (Fr * H * W * 12 if name.endswith("backward") else 0)
    return 0


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_facts(Fr, N, size):
    """Counters of the committed ncu --set full capture (profiles/ncu_traffic.json): DRAM bytes per launch and, for the
    issue-bound blend kernel, its issue-slot and pipe utilisation.  Empty when the capture is of another workload."""
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(tpath):
        return {}
    with open(tpath) as fh:
        tj = json.load(fh)
    if tj.get("workload") != {"frames": Fr, "gaussians": N, "size": size}:
        return {}
    return tj


# ----------------------------------------------------------------------------------------------------------------
# CPU legs (the oracle port; the ONLY places bench.py touches oracle/)
# ----------------------------------------------------------------------------------------------------------------
def cpu_reference(frames_per_step, steps, warmup, size, ngauss, seed=0, inputs=None, variant="fast"):
    """The reference's CPU implementation of the path (oracle port), all host threads: `frames_per_step` frames of
    the workload per step.  `inputs` = (tokens [F,C,3RR], flat params, K, E) on the CPU to render instead of the
    seeded default.  Returns (frames/s, cores, seconds per step, images [F,H,W,3], gaussians)."""
    from audio_motion_avatar_b200 import synthetic as S
    from oracle import ref_path as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = S.make_body_model(seed=0)
    w = S.make_decoder_weights(C, seed=0)
    levels, idx = None, None
    if ngauss != 10475:
        levels = S.subdivision_levels(10475, model["faces"].numpy(), subdivide_steps(ngauss))
        nd = 10475 + sum(len(e) for e in levels)
        idx = torch.randperm(nd, generator=torch.Generator().manual_seed(seed))[:ngauss]
    if inputs is None:
        tokens = S.make_triplane_tokens(frames_per_step, C, R, seed=seed)[0]
        params = S.make_pose_sequence(1, frames_per_step, seed=seed)
        flat = {k: v.reshape(frames_per_step, *v.shape[2:]) for k, v in params.items()}
        cams = S.make_cameras(1, frames_per_step, size, size)
        K, E = cams["intrinsic"][0], cams["extrinsic"][0]
    else:
        tokens, flat, K, E = inputs
    times, img, g = [], None, None
    with O.oracle_variant(variant):
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            with torch.no_grad():
                img, g = O.reference_frames(model, w, tokens, flat, K, E, size, size, C, R, 1.4, levels, idx)
            if it >= warmup:
                times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return frames_per_step / sec, cores, sec, img, g


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    warm = max(0, min(args.warmup, 2))
    fps, cores, sec, _, _ = cpu_reference(1, max(1, args.steps), warm, args.size, args.gaussians)
    sample = f"1 frame of the {args.frames}-frame sequence per step ({args.gaussians} Gaussians, {args.size}x{args.size})"
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": max(1, args.steps),
            "warmup": warm, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args), "frames_per_step": 1,
                       "note": "CPU port of the reference path (kind 'port': smplx / pytorch3d / diff-gaussian-rasterization are not "
                               "installable offline): torch CPU LBS, F.grid_sample twice + five Linear as renderer.py:136-171, "
                               "C rasteriser built -O3 -march=native with OpenMP on preprocess, key emit, sort and blend"},
            "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_name(args):
    which = "configs[1]" if (args.frames, args.gaussians, args.size) == (FRAMES, NGAUSS, SIZE) else \
        "configs[4] sizes, secondary line" if (args.gaussians, args.size) == (500_000, 2048) else "non-default sizes, secondary line"
    return (f"{args.frames}-frame audio-driven sequence, {args.gaussians} Gaussians, {args.size}x{args.size}, "
            f"triplane 3x{R}x{R}x{C} fp32 per frame, 55-joint synthetic SMPL-X body (BASELINE.json {which})")


# ----------------------------------------------------------------------------------------------------------------
# shared pieces of the GPU arm
# ----------------------------------------------------------------------------------------------------------------
class Ctx:
    """Process-wide state of the GPU arm: rank / world, device, barrier, max-over-ranks reduction."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a GPU (there is no CPU path); use --impl reference for the CPU baseline")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            # NCCL_DEBUG is left as the caller set it.  NCCL logs to stdout by default and keeps logging until the library
            # unloads (after Python's last line), so its log goes to stderr: stdout then ends with the JSON line
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return x
        t = torch.tensor([x], device=self.dev, dtype=torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, step, steps, warmup, clocks=True):
        """`warmup` untimed steps, then exactly `steps` steps between barrier + synchronize on both sides, CUDA events on
        the launching stream, max over ranks.  Returns (ms per step, clocks summary, library launches per step)."""
        from audio_motion_avatar_b200 import _lib
        for _ in range(warmup):
            step()
        self.barrier()
        n0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(self.local) if clocks else None
        if sampler:
            sampler.__enter__()
        self.barrier()
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        self.barrier()
        if sampler:
            sampler.__exit__()
        ms = self.max_over_ranks(e0.elapsed_time(e1)) / steps
        return ms, (sampler.summary() if sampler else None), (_lib.launch_count() - n0) // steps


def make_renderer(size, N, ctx):
    from audio_motion_avatar_b200 import synthetic as S
    from audio_motion_avatar_b200.renderer import Renderer
    r = Renderer(make_cfg(size, N, f"cuda:{ctx.local}"), body_model=S.make_body_model(seed=0))
    r.gaussian_decoder.load_state_dict(S.make_decoder_weights(C, seed=0))
    if N != 10475:
        r.fix_site_subset(seed=0)
    return r


def stage_table(r, tokens, cams_d, params_d, Fr, N, size, repeats=2):
    """Per-stage device time of the eager forward (CUDA events around every C-ABI call) -> per-stage roofline table.
    Measured in its own eager passes right after the timed region (a CUDA-graph replay cannot carry timing events)."""
    from audio_motion_avatar_b200 import _lib
    from audio_motion_avatar_b200.renderer import _workspace
    with torch.no_grad():
        r(tokens, cams_d, None, params_d)
        with _lib.StageTimer() as timer:
            for _ in range(repeats):
                r(tokens, cams_d, None, params_d)
        stages_ms = {k: v[0] / repeats for k, v in timer.summary().items()}
    D = int(_workspace(r._device).last_D)
    body = r.smplx_model
    P = int(r._dense_par.shape[1]) if r._dense_par is not None else 1
    peak, peak_src = peaks()
    per_stage = {}
    for name, sms in stages_ms.items():
        by = stage_bytes(name, Fr, N, body.V, body.J, body.S, P, size, size, D)
        per_stage[name] = {"ms": round(sms, 4), "algorithmic_bytes": by, "gbs": round(by / (sms * 1e-3) / 1e9, 1),
                           "frac": round(by / (sms * 1e-3) / 1e9 / peak, 4)}
    return per_stage, stages_ms, D, peak, peak_src


def roofline_block(per_stage, stages_ms, D, peak, peak_src, Fr, N, size, ms_per_step):
    top = max(stages_ms, key=stages_ms.get)
    facts = ncu_facts(Fr, N, size)
    traffic = facts.get("bytes_per_launch", {}).get(top)
    roof = {"kernel": top, "bound": "hbm", "achieved": per_stage[top]["gbs"], "peak": peak, "unit": "GB/s",
            "frac": per_stage[top]["frac"], "traffic": traffic, "peak_source": peak_src,
            "share_of_step": round(stages_ms[top] / ms_per_step, 4), "stages": per_stage,
            "entries_D": D, "entries_per_frame": D // max(Fr, 1),
            "stages_timing": "CUDA events around every C-ABI call in 2 eager passes right after the timed region"}
    if top == "amav_splat_blend":
        # The dominant kernel is bound by instruction issue / the FP32 pipe, not by HBM (its DRAM traffic is below its
        # algorithmic bytes: L2 serves the record gather).  The contract's `bound` takes hbm|tensor, so the HBM fraction
        # (SURVEY 8d bytes: 48 D + 12 H W F) stays in `frac` and the true limiter is named here with its ncu counters.
        roof["limiter"] = {"bound": "issue", "note": "FP32 instruction issue + FMA pipe (per-pair exact arithmetic); ncu counters "
                           "of the committed capture under profiles/", **facts.get("blend_issue", {})}
        roof["blend_pixel_entry_pairs_per_s"] = round(D * 256 / (stages_ms[top] * 1e-3), 0)
    return roof


# ----------------------------------------------------------------------------------------------------------------
# secondary blocks
# ----------------------------------------------------------------------------------------------------------------
def parity_block(ctx, r, tokens, params, cams, N, size):
    """Frame 0 of the workload: GPU arm vs the CPU oracle on the same inputs (north_star: lists bit-exact, attributes and
    images max-abs <= 1e-3 / PSNR >= 60 dB).  Also: how far the oracle's shared polynomial exp / FMA nesting is from
    upstream's literal arithmetic (libm expf, no FMA), counted in blend decisions."""
    import math
    import numpy as np
    from audio_motion_avatar_b200.renderer import camera_matrices
    from audio_motion_avatar_b200.splat import SplatWorkspace, splat_forward
    from oracle import ref_path as O
    dev = ctx.dev
    tok0 = tokens[:, :1]
    par0 = {k: v[:, :1] for k, v in params.items()}
    cam0 = {k: v[:, :1] for k, v in cams.items()}
    with torch.no_grad():
        img, g = r(tok0, {k: v.to(dev) for k, v in cam0.items()}, None, {k: v.to(dev) for k, v in par0.items()})
    flat = {k: v.reshape(1, *v.shape[2:]).cpu() for k, v in par0.items()}
    _, cores, sec, ref_img, ref_g = cpu_reference(1, 1, 0, size, N, inputs=(tok0[0].cpu(), flat, cam0["intrinsic"][0].cpu(),
                                                                               cam0["extrinsic"][0].cpu()), variant="exact")
    a, b = img[0, 0].cpu().numpy().astype(np.float64), ref_img[0].numpy().astype(np.float64)
    d = np.abs(a - b)
    mse = float((d ** 2).mean())
    attr = {k: float((g[k].cpu() - ref_g[k]).abs().max()) for k in ("xyz", "scale", "rot", "opacity", "color")}
    # stage-isolated: the GPU's own Gaussians, activated on the CPU, through both rasterisers
    gc = {k: v[0].detach().cpu() for k, v in g.items()}
    sc, op, col = O.activate_for_raster(gc)
    view, full, tan = camera_matrices(cam0["intrinsic"].reshape(1, 3, 3).to(dev), cam0["extrinsic"].reshape(1, 4, 4).to(dev), size, size)
    out = splat_forward(gc["xyz"][None].to(dev), sc[None].to(dev), gc["rot"][None].to(dev), op[None].to(dev), col[None].to(dev),
                        view, full, tan, torch.ones(3, device=dev), size, size, want_aux=True, workspace=SplatWorkspace())
    torch.cuda.synchronize()
    pre, binned, ora = O.rasterize(gc["xyz"], sc, gc["rot"], op, col, view[0].cpu(), full[0].cpu(), float(tan[0, 0]), float(tan[0, 1]),
                                   (1.0, 1.0, 1.0), size, size)
    Dn = binned["D"]
    tiles = ((size + 15) // 16) ** 2
    ts = out["tile_start"][:tiles + 1].cpu().numpy().astype(np.int64)
    lists_ok = bool(Dn == out["D"] and np.array_equal(np.diff(ts), binned["ranges"][:, 1] - binned["ranges"][:, 0]) and
                    np.array_equal(out["point_list"][:Dn].cpu().numpy(), binned["point_list"]))
    blend_ok = bool(np.array_equal(out["n_contrib"][0].cpu().numpy(), ora["n_contrib"]) and
                    np.array_equal(out["final_T"][0].cpu().numpy(), ora["final_T"]) and
                    np.array_equal(out["color"][0].cpu().numpy(), ora["color"]))
    with O.oracle_variant("upstream"):
        up = O.raster_blend(pre, binned, col, (1.0, 1.0, 1.0), size, size)
    flips = int((up["n_contrib"] != ora["n_contrib"]).sum())
    dcol = np.abs(up["color"] - ora["color"])
    return {"config": "configs[1] frame 0" if (N, size) == (NGAUSS, SIZE) else f"frame 0 ({N} Gaussians, {size}x{size})",
            "max_abs": float(d.max()), "psnr_db": 999.0 if mse == 0 else 10.0 * math.log10(1.0 / mse),
            "pixels_off_by_more_than_1e-3": int((d.max(axis=-1) > 1e-3).sum()), "pixels": int(size * size),
            "attr_max_abs": attr, "lists_bit_exact": lists_ok, "blend_bit_exact_on_same_gaussians": blend_ok,
            "entries_D": int(Dn),
            "oracle_vs_upstream_arithmetic": {"n_contrib_pixels_differ": flips, "colour_pixels_differ_gt_1e-5": int((dcol.max(axis=0) > 1e-5).sum()),
                                              "max_abs_colour": float(dcol.max()),
                                              "note": "same sorted lists blended with libm expf and upstream's literal power expression "
                                                      "(no FMA) instead of the polynomial exp / FMA nesting shared by oracle and kernels"},
            "oracle_seconds": round(sec, 2), "cores": cores}


def latency_block(ctx, r, size, N, T=6, iters=30):
    """One Renderer.forward call at the reference's native clip length (T = 6 frames per call,
    configs/model/triplane_audio_net.yaml:4; the per-frame loop it replaces is renderer.py:475-477), device-resident."""
    from audio_motion_avatar_b200 import _lib
    dev = ctx.dev
    tokens, params, cams = synthetic_inputs(0, T, FRAMES, size, dev)
    params_d = {k: v.to(dev) for k, v in params.items()}
    cams_d = {k: v.to(dev) for k, v in cams.items()}
    res = {"frames_per_call": T, "size": size, "gaussians": N}

    def time_loop(fn, n):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, (time.perf_counter() - t0) * 1e3 / n

    def eager():
        with torch.no_grad():
            r(tokens, cams_d, None, params_d)

    res["eager_ms_per_call"], res["eager_wall_ms_per_call"] = time_loop(eager, iters)
    with torch.no_grad():
        with _lib.StageTimer() as timer:
            for _ in range(3):
                r(tokens, cams_d, None, params_d)
        res["kernel_ms_per_call"] = sum(v[0] for v in timer.summary().values()) / 3
    gf = r.capture(tokens, cams_d, params_d)
    res["graph_ms_per_call"], res["graph_wall_ms_per_call"] = time_loop(gf.replay, iters)
    res["graph_overflowed"] = bool(gf.overflowed())
    res["graph_library_launches"] = int(gf.launches)
    res["graph_over_kernel_time"] = res["graph_ms_per_call"] / res["kernel_ms_per_call"]
    res["eager_over_kernel_time"] = res["eager_ms_per_call"] / res["kernel_ms_per_call"]
    res["frames_per_s_graph"] = T / (res["graph_ms_per_call"] * 1e-3)
    res["note"] = ("eager: Renderer.forward incl. its one host read of the overflow flag per call; graph: Renderer.capture(...).replay(), "
                   "flag read once after the loop; kernel_ms = sum of CUDA-event times of the C-ABI calls")
    return res


def train_jobs(rank, world, B, Tv, per_rank_all=False):
    """The (subject, view) render jobs of one training step split over the ranks in contiguous blocks, subject-major:
    16 jobs over 8 ranks = 2 views of one subject per rank; over 4 ranks one subject per rank; over 2 ranks two subjects."""
    total = B * Tv
    if per_rank_all or world == 1:
        return list(range(B)), list(range(Tv))
    if total % world != 0:
        raise SystemExit(f"{total} (subject, view) jobs do not split over {world} ranks")
    per = total // world
    j0 = rank * per
    if per >= Tv:
        return list(range(j0 // Tv, (j0 + per) // Tv)), list(range(Tv))
    return [j0 // Tv], list(range(j0 % Tv, j0 % Tv + per))


def train_block(ctx, args, steps=3, warmup=3, size=512, clocks=True, graph=True):
    """BASELINE.json configs[3]: `subjects` Gaussian sets (own triplane and pose) x `views` orbit cameras, 100k Gaussians,
    512x512, l1 + 0.1 (1 - ssim), full backward through splat, decode / triplane query and LBS, SGD on the decoder.
    N > 1: the subjects x views jobs are split over the ranks (train_jobs); dL/dtokens and the decoder gradients are
    all-reduced inside the timed step.  (lightning_model_wrapper.py:51,132-145.)"""
    from audio_motion_avatar_b200 import _lib, losses, synthetic as S
    from audio_motion_avatar_b200.renderer import render_multi_view
    from audio_motion_avatar_b200.shard import allreduce_gradients
    world, rank, dev = ctx.world, ctx.rank, ctx.dev
    B, Tv, N = args.subjects, args.views, NGAUSS if args.workload != "train" else args.gaussians
    r = make_renderer(size, N, ctx)
    subj, views = train_jobs(rank, world, B, Tv, args.per_rank_16)
    g = torch.Generator(device=dev).manual_seed(0)                                 # same subjects on every rank
    tokens = torch.randn(B, 1, C, 3 * R * R, device=dev, generator=g).requires_grad_(True)
    params = {k: v.reshape(B, 1, *v.shape[2:]).to(dev) for k, v in S.make_pose_sequence(1, B, seed=0).items()}
    params["transl"] = torch.tensor([0.0, 0.35, 0.0], device=dev).expand(B, 1, 3).contiguous()   # body at the orbit centre
    pose_leaves = {k: params[k].clone().requires_grad_(True) for k in ("global_orient", "body_pose", "betas", "expression", "transl")}
    params = {**params, **pose_leaves}
    cams = {k: v.to(dev) for k, v in S.make_cameras(B, Tv, size, size, orbit=True).items()}   # 4 cameras on a 2.4 m circle
    target = torch.rand(B, Tv, size, size, 3, device=dev, generator=g)
    opt = torch.optim.SGD(r.gaussian_decoder.parameters(), lr=1e-4)
    leaves = [tokens] + list(pose_leaves.values())
    s0, s1 = subj[0], subj[-1] + 1
    v0, v1 = views[0], views[-1] + 1
    total_jobs = B * Tv if not (args.per_rank_16 and world > 1) else B * Tv * world
    my_jobs = len(subj) * len(views)
    if args.loss == "l1_ssim":
        image_loss = lambda a, b: losses.l1_ssim_loss(a, b)[0]
    else:
        image_loss = lambda a, b: (a - b).abs().mean()
    sub = lambda d: {k: v[s0:s1] for k, v in d.items()}

    def step():
        opt.zero_grad(set_to_none=True)
        for t in leaves:
            t.grad = None
        first = {k: v[s0:s1, v0:v0 + 1] for k, v in cams.items()}
        img, gs = r(tokens[s0:s1], first, None, sub(params))                          # decode + first view of the block
        loss = image_loss(img, target[s0:s1, v0:v0 + 1]) * len(subj)
        if v1 - v0 > 1:
            mv = render_multi_view(gs, cams["intrinsic"][s0:s1, v0 + 1:v1], cams["extrinsic"][s0:s1, v0 + 1:v1], r.cfg)
            loss = loss + image_loss(mv, target[s0:s1, v0 + 1:v1]) * (len(subj) * (v1 - v0 - 1))
        loss = loss / (B * Tv)                                                        # mean over the step's 16 renders
        loss.backward()
        if world > 1:
            allreduce_gradients([tokens.grad] + [p.grad for p in r.gaussian_decoder.parameters()])
        opt.step()
        return loss

    ms, clk, launches = ctx.timed(step, steps, warmup, clocks=clocks)
    loss_value = float(step().item())          # (no reference to the loss tensor is kept: it would pin this step's autograd graph)
    with _lib.StageTimer() as timer:
        step()
    stages = {k: {"ms": round(v[0], 4), "calls": v[1]} for k, v in timer.summary().items()}
    kernel_ms = sum(v["ms"] for v in stages.values())
    blk = {"metric": "training renders/sec (fwd+bwd) @100k Gaussians 512^2", "value": total_jobs / (ms * 1e-3),
           "unit": "(subject,view) renders/s", "ms_per_step": ms, "steps": steps, "warmup": warmup,
           "scaling": "weak" if (args.per_rank_16 and world > 1) else "strong",
           "config": {"workload": f"training step: {B} subjects x {Tv} views = {B * Tv} renders, {N} Gaussians, {size}x{size}, loss {args.loss}, "
                                  "forward + backward through splat, decode/triplane query and LBS, SGD on the decoder (BASELINE.json configs[3])",
                      "jobs_total": total_jobs, "jobs_this_rank": my_jobs, "rank0_subjects": subj, "rank0_views": views,
                      "parallelism": ("1 GPU" if world == 1 else f"(subject, view) jobs split over {world} ranks + NCCL all-reduce of "
                                      "dL/dtokens (100 MB) and decoder gradients inside the timed step")},
           "clocks": clk, "gpu_launches": int(launches), "loss": loss_value, "stages": stages,
           "library_kernel_ms": round(kernel_ms, 4), "step_over_library_kernel_time": round(ms / kernel_ms, 3) if kernel_ms else None}
    if graph:
        try:
            blk["graph"] = train_graph(ctx, step, steps, kernel_ms)
            if "ms_per_step" in blk["graph"] and not blk["graph"].get("overflowed"):
                # the captured step is the way to run training (host-proof); the eager numbers stay beside it
                blk["eager"] = {"value": blk["value"], "ms_per_step": blk["ms_per_step"]}
                blk["ms_per_step"] = blk["graph"]["ms_per_step"]
                blk["value"] = total_jobs / (blk["ms_per_step"] * 1e-3)
                blk["mode"] = "one CUDA graph per step (forward + loss + backward + all-reduce + SGD); `eager` = the same step launched kernel by kernel"
        except Exception as exc:                                     # a capture failure must not cost the eager numbers
            import traceback
            traceback.print_exc(file=sys.stderr)
            blk["graph"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
    del r
    return blk


def train_graph(ctx, step, steps, kernel_ms):
    """The same training step captured as ONE CUDA graph (forward + loss + backward + SGD): replay time next to the sum
    of its kernels' times -- what makes the step independent of the host's launch rate."""
    from audio_motion_avatar_b200 import splat
    s = torch.cuda.Stream(ctx.dev)
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with splat.fixed_training_capacity():
            for _ in range(3):
                step()
            s.synchronize()
            g = torch.cuda.CUDAGraph()
            with splat.fixed_training_capacity():                # fresh flag list: only the captured calls' flags
                with torch.cuda.graph(g, stream=s):
                    step()
                flags = splat.training_overflow_flags()
    torch.cuda.current_stream().wait_stream(s)
    n = max(steps, 5)
    ms, _, _ = ctx.timed(g.replay, n, 3, clocks=False)          # barrier + synchronize on both sides, max over ranks
    over = bool(torch.stack([f.reshape(()) for f in flags]).any().item()) if flags else False
    return {"ms_per_step": ms, "over_library_kernel_time": round(ms / kernel_ms, 3) if kernel_ms else None, "overflowed": over,
            "note": "torch.cuda.graph over forward + fused loss + backward + SGD; bucket capacities fixed from the eager steps, "
                    "overflow flags read after the replays"}


def stress_block(ctx, steps=2, warmup=1, N=500_000, size=2048, chunk=15):
    """BASELINE.json configs[4] sizes: 500k Gaussians (level-3 subdivision), 2048x2048.  At 8 GPUs the 1800-frame sequence
    is 225 frames per rank, rendered in `chunk`-frame calls (a call's bucket buffer is ~118 MB per frame) and kept
    sharded (90 GB of frames: no gather); below 8 GPUs 16 frames per rank."""
    from audio_motion_avatar_b200.renderer import deferred_overflow_checks
    world, rank, dev = ctx.world, ctx.rank, ctx.dev
    per_rank = 1800 // world if world >= 8 else 16
    total = per_rank * world
    r = make_renderer(size, N, ctx)
    f0 = rank * per_rank
    tokens, params, cams = synthetic_inputs(f0, f0 + per_rank, total, size, dev, seed=1)
    params_d = {k: v.to(dev) for k, v in params.items()}
    cams_d = {k: v.to(dev) for k, v in cams.items()}
    nc = min(chunk, per_rank)
    out = torch.empty(nc, size, size, 3, device=dev)
    cuts = list(range(0, per_rank, chunk)) + [per_rank]
    state = {}

    def step():
        with deferred_overflow_checks() as chk, torch.no_grad():
            for a, b in zip(cuts[:-1], cuts[1:]):
                r(tokens[:, a:b], {k: v[:, a:b] for k, v in cams_d.items()}, None, {k: v[:, a:b] for k, v in params_d.items()},
                  out=out[:b - a])
        state["chk"] = chk

    ms, clk, launches = ctx.timed(step, steps, warmup)
    over = state["chk"].overflowed()
    per_stage, stages_ms, D, peak, _ = stage_table(r, tokens[:, :nc], {k: v[:, :nc] for k, v in cams_d.items()},
                                                   {k: v[:, :nc] for k, v in params_d.items()}, nc, N, size, repeats=1)
    blk = {"metric": "rendered frames/sec @500k Gaussians 2048^2", "value": total / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
           "steps": steps, "warmup": warmup, "scaling": "strong" if world >= 8 else "weak",
           "config": {"workload": f"{total}-frame sequence, {N} Gaussians (level-3 subdivision), {size}x{size} (BASELINE.json configs[4]" +
                                  (")" if world >= 8 else f" sizes; the full 1800 frames need 8 GPUs, {per_rank} frames per rank here)"),
                      "frames_per_rank": per_rank, "frames_total": total, "frames_per_call": chunk,
                      "parallelism": f"frame-sharded x{world}, frames stay on the rank that rendered them (1800 x 50 MB = 90 GB: no gather)"},
           "clocks": clk, "gpu_launches": int(launches), "overflow": bool(over), "entries_per_frame": D // max(nc, 1),
           "stages_one_call": {k: {"ms": v["ms"], "frac": v["frac"]} for k, v in per_stage.items()}}
    del r, tokens, out
    torch.cuda.empty_cache()
    return blk


# ----------------------------------------------------------------------------------------------------------------
# the default line
# ----------------------------------------------------------------------------------------------------------------
def run_ours(args):
    from audio_motion_avatar_b200.renderer import deferred_overflow_checks
    from audio_motion_avatar_b200.shard import PeerFrameGather, frame_shard, render_and_gather
    ctx = Ctx()
    world, rank, dev, dist = ctx.world, ctx.rank, ctx.dev, ctx.dist
    warmup, steps = max(3, args.warmup), max(1, args.steps)
    Ftot, N, size = args.frames, args.gaussians, args.size
    only = set(x for x in args.only.split(",") if x)
    want = lambda name: (not args.no_extras) and (not only or name in only)
    default_wl = (Ftot, N, size) == (FRAMES, NGAUSS, SIZE)

    r = make_renderer(size, N, ctx)
    f0, f1 = frame_shard(Ftot, rank, world)                                 # STRONG split: this rank's frames of the SAME sequence
    Fr = f1 - f0
    tokens, params, cams = synthetic_inputs(f0, f1, Ftot, size, dev)
    params_d = {k: v.to(dev) for k, v in params.items()}
    cams_d = {k: v.to(dev) for k, v in cams.items()}

    def make_gather(frames_total):
        """(peer, buffer, mode) for a gather of `frames_total` frames on rank 0."""
        peer, buf, mode = None, None, "none"
        want_mode = args.gather
        if want_mode == "auto":
            # fused row stores take ~520 GB/s into rank 0 from many senders: they stay hidden behind a shard's render up to
            # N = 4 (3 x 1.5 GB = 8.7 ms < 14 ms); at N = 8 (10.6 GB = 20 ms) the staged copies (~795 GB/s) win
            want_mode = "staged" if world >= 8 and frames_total // world >= 60 else "peer"
        if world > 1 and not args.no_gather:
            mode = "nccl"
            if want_mode in ("peer", "staged") and PeerFrameGather.available():
                try:
                    peer = PeerFrameGather(frames_total, size, size, dev, dst=0)
                except Exception as exc:                 # no peer access / allocator unavailable: NCCL send/recv instead
                    print(f"[bench] symmetric memory unavailable ({type(exc).__name__}: {exc}); using the NCCL gather", file=sys.stderr)
                ok = torch.tensor([1.0 if peer is not None else 0.0], device=dev)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN)            # every rank takes the same path
                if ok.item() > 0:
                    mode = want_mode
                else:
                    peer = None
            if peer is None and rank == 0:
                buf = torch.empty(frames_total, size, size, 3, device=dev)
        return peer, buf, mode

    chunks = args.gather_chunks if args.gather_chunks > 0 else (1 if world <= 2 else 4)
    stage_chunks = lambda frames: args.gather_chunks if args.gather_chunks > 0 else max(1, frames // 30)

    def make_step(peer, buf, tok, cm, pr, frames_total, chk_box, mode):
        def step():
            with deferred_overflow_checks() as chk:      # overflow flags are read once, after the timed region
                if peer is not None and mode == "staged":
                    # frame gather (north_star): pieces rendered locally, pushed to rank 0 by peer copies while the next renders
                    peer.render_staged(r, tok, cm, pr, chunks=stage_chunks(tok.shape[1]))
                elif peer is not None:
                    # frame gather fused into the blend kernel: pixels are stored into rank 0's buffer over NVLink
                    peer.render(r, tok, cm, pr)
                elif world > 1 and not args.no_gather:
                    render_and_gather(r, tok, cm, pr, frames_total, chunks=chunks, dst=0, out=buf)
                else:
                    with torch.no_grad():
                        r(tok, cm, None, pr)
            chk_box["chk"] = chk
        return step

    # ---- headline: configs[1] (N = 1) / configs[2] strong split (N > 1) ----
    peer, gather_buf, gather_mode = make_gather(Ftot)
    box = {}
    step = make_step(peer, gather_buf, tokens, cams_d, params_d, Ftot, box, gather_mode)
    ms_per_step, clocks, launches = ctx.timed(step, steps, warmup)
    overflow = bool(box["chk"].overflowed())
    if world > 1:
        t = torch.tensor([1.0 if overflow else 0.0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        overflow = bool(t.item() > 0)
    value = Ftot / (ms_per_step * 1e-3)
    per_stage, stages_ms, D, peak, peak_src = stage_table(r, tokens, cams_d, params_d, Fr, N, size)
    par = f"frame-sharded x{world}" + {"none": "", "peer": " + frame gather to rank 0 by peer stores from the blend kernel (symmetric memory over NVLink)",
                                        "staged": " + frame gather to rank 0: pieces rendered locally, peer copies into rank 0's symmetric buffer on a second stream",
                                        "nccl": f" + NCCL frame gather to rank 0 overlapped with the render ({chunks} pieces)"}[gather_mode]
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args), "frames_per_rank": Fr, "frames_total": Ftot, "parallelism": par,
                       "api": "Renderer.forward (eager), bucket-overflow flags read once after the timed region",
                       "l2": "inputs (25 MB of triplane per frame, 3 GB per 120 frames) exceed the 126 MB L2; no flush needed"},
            "clocks": clocks, "gpu_launches": int(launches), "overflow": overflow,
            "roofline": roofline_block(per_stage, stages_ms, D, peak, peak_src, Fr, N, size, ms_per_step)}

    # the same step as ONE CUDA-graph launch per rank (Renderer.capture); at N > 1 the peer stores are inside the graph
    if want("graph") and (world == 1 or (peer is not None and gather_mode == "peer")):
        try:
            if peer is not None:
                gobj = peer.capture(r, tokens, cams_d, params_d)
                gstep = peer.render_captured
            else:
                gobj = r.capture(tokens, cams_d, params_d)
                gstep = gobj.replay
            gms, _, _ = ctx.timed(gstep, steps, warmup, clocks=False)
            line["graph"] = {"value": Ftot / (gms * 1e-3), "unit": UNIT, "ms_per_step": gms,
                             "library_launches_per_replay": int(gobj.launches) if gobj is not None else 0,
                             "overflow": bool(gobj.overflowed()) if gobj is not None else False,
                             "api": "Renderer.capture(...).replay(): the whole shard step as one CUDA-graph launch"}
            del gobj
            if peer is not None:
                peer._graph = None
        except Exception as exc:
            line["graph"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}

    # N > 1: the data plane checks itself -- rank 0 re-renders rank 1's frames locally and compares its gathered buffer
    if world > 1 and not args.no_gather:
        step()
        ctx.barrier()
        if rank == 0:
            a1, b1 = frame_shard(Ftot, 1, world)
            n = min(b1 - a1, 4)
            tk, pr, cm = synthetic_inputs(a1, a1 + n, Ftot, size, dev)
            with torch.no_grad():
                ref = r(tk, {k: v.to(dev) for k, v in cm.items()}, None, {k: v.to(dev) for k, v in pr.items()})[0][0]
            got = (peer.local if peer is not None else gather_buf)[a1:a1 + n]
            line["gather_check"] = {"rank1_frames_checked": n, "bit_exact_vs_local_rerender": bool(torch.equal(got, ref)),
                                    "max_abs": float((got - ref).abs().max().item())}
            del tk, ref
        ctx.barrier()
        # the same strong split with the opt-in 8-bit gather (a separate line; fp32 above is the headline): what the step
        # costs once rank 0's ingress carries 3 bytes per pixel instead of 12
        if peer is not None and want("weak"):
            def s8_step():
                with deferred_overflow_checks() as chk:
                    peer.render_staged_u8(r, tokens, cams_d, params_d, chunks=1, bgr=True)
                box["chk8"] = chk
            s8ms, _, _ = ctx.timed(s8_step, steps, warmup, clocks=False)
            line["u8_gather"] = {"value": Ftot / (s8ms * 1e-3), "unit": UNIT, "ms_per_step": s8ms, "frames_total": Ftot,
                                 "out": "uint8 BGR [F,H,W,3] on rank 0", "overflow": bool(box["chk8"].overflowed())}
    del peer, gather_buf
    torch.cuda.empty_cache()

    # ---- weak variant (round 1's line): 120 frames PER RANK, gathered ----
    if world > 1 and want("weak") and default_wl:
        wtok, wpar, wcam = synthetic_inputs(rank * Ftot, (rank + 1) * Ftot, world * Ftot, size, dev)
        wpar_d, wcam_d = {k: v.to(dev) for k, v in wpar.items()}, {k: v.to(dev) for k, v in wcam.items()}
        wpeer, wbuf, wmode = make_gather(world * Ftot)
        wbox = {}
        wstep = make_step(wpeer, wbuf, wtok, wcam_d, wpar_d, world * Ftot, wbox, wmode)
        wms, wclk, _ = ctx.timed(wstep, steps, warmup)
        line["weak"] = {"value": world * Ftot / (wms * 1e-3), "unit": UNIT, "ms_per_step": wms, "frames_per_rank": Ftot,
                        "frames_total": world * Ftot, "scaling": "weak", "gather": wmode, "clocks": wclk,
                        "overflow": bool(wbox["chk"].overflowed())}
        # opt-in 8-bit gather (the demo's video sink wants uint8): converted where rendered, 4x fewer bytes into rank 0;
        # a separate line -- the fp32 gather above stays the headline
        if wpeer is not None:
            from audio_motion_avatar_b200.stages import frames_to_u8

            def u8_step():
                with deferred_overflow_checks() as chk:
                    wpeer.render_staged_u8(r, wtok, wcam_d, wpar_d, chunks="halving" if world >= 8 else max(1, wtok.shape[1] // 60), bgr=True)   # only the last piece's copy is exposed
                wbox["chk8"] = chk
            ums, uclk, _ = ctx.timed(u8_step, steps, warmup)
            blk8 = {"value": world * Ftot / (ums * 1e-3), "unit": UNIT, "ms_per_step": ums, "frames_total": world * Ftot,
                    "out": "uint8 BGR [F,H,W,3] on rank 0 (amav_frames_to_u8 on the rendering rank, staged peer copies)",
                    "bytes_into_rank0": (world - 1) * Ftot * size * size * 3, "clocks": uclk,
                    "overflow": bool(wbox["chk8"].overflowed())}
            if rank == 0:                                # rank 1's first frame in the gathered buffer == a local render of it
                tk1, pr1, cm1 = synthetic_inputs(Ftot, Ftot + 1, world * Ftot, size, dev)
                with torch.no_grad():
                    one, _ = r(tk1, {k: v.to(dev) for k, v in cm1.items()}, None, {k: v.to(dev) for k, v in pr1.items()})
                torch.cuda.synchronize()
                blk8["rank1_frame0_bit_exact"] = bool(torch.equal(wpeer.local_u8[Ftot], frames_to_u8(one[0], bgr=True)[0]))
                del tk1, one
            ctx.barrier()
            line["weak"]["u8_gather"] = blk8
        del wtok, wpeer, wbuf
        torch.cuda.empty_cache()

    # ---- e2e: host (pinned) triplanes in, host frames out, through the public Renderer.render_sequence ----
    if not args.no_e2e:
        tok_h = torch.empty(tokens.shape, dtype=torch.float32, pin_memory=True)
        tok_h.copy_(tokens)
        out_h = torch.empty(1, Fr, size, size, 3, dtype=torch.float32, pin_memory=True)
        small_h = sum(v.numel() * 4 for v in params.values()) + sum(v.numel() * 4 for v in cams.values())
        for _ in range(2):
            r.render_sequence(tok_h, cams, params, chunk=args.chunk, out=out_h)
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            with torch.no_grad():
                r.render_sequence(tok_h, cams, params, chunk=args.chunk, out=out_h)
        torch.cuda.synchronize()
        sec = ctx.max_over_ranks(time.perf_counter() - t0)
        line["e2e"] = {"value": Ftot * steps / sec, "unit": UNIT, "h2d_bytes_per_step": tok_h.numel() * 4 + small_h,
                       "d2h_bytes_per_step": out_h.numel() * 4, "ms_per_step": sec / steps * 1e3,
                       "bytes_are": "per rank" if world > 1 else "whole job",
                       "api": f"Renderer.render_sequence(chunk={args.chunk}), pinned host buffers, 3 streams"}
        with torch.no_grad():
            img, _ = r(tokens, cams_d, None, params_d)               # this rank's shard, device-resident path
        torch.cuda.synchronize()
        line["e2e"]["max_abs_vs_device_path"] = float((out_h.to(dev) - img).abs().max().item())
        del out_h
        if want("sink") and world == 1:
            sink_and_layout_blocks(line, r, tokens, tok_h, img, cams, params, Fr, size, steps, args, ctx)
        del tok_h, img
        torch.cuda.empty_cache()

    # ---- secondary blocks ----
    if want("parity") and world == 1 and rank == 0:
        line["parity"] = parity_block(ctx, r, tokens, params, cams, N, size)
    if want("latency") and default_wl:
        lb = latency_block(ctx, r, size, N)
        if rank == 0:
            line["latency_T6"] = lb
    if want("yardsticks") and world == 1 and default_wl:
        try:
            from baseline.gpu_yardsticks import yardsticks_block
            line["gpu_yardsticks"] = yardsticks_block(ctx, r, size, N)
        except ImportError:
            pass
        except Exception as exc:
            line["gpu_yardsticks"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
    del tokens
    torch.cuda.empty_cache()
    if want("train") and default_wl:
        line["train"] = train_block(ctx, args)
        torch.cuda.empty_cache()
    if want("stress") and default_wl:
        line["stress"] = stress_block(ctx)

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        fps, cores, sec, _, _ = cpu_reference(1, 2, 1, size, N)
        line["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"1 frame of the workload per step, 1 warm-up + 2 timed ({sec:.2f} s each): torch CPU LBS/grid_sample/Linear + "
                                          "C rasteriser (-O3 -march=native, OpenMP on preprocess, key emit, sort, blend)"}
    finish(ctx, line)


def sink_and_layout_blocks(line, r, tokens, tok_h, img, cams, params, Fr, size, steps, args, ctx):
    """SURVEY 8f-4 (uint8 video sink on the device) and 8f-2 (producer-side token-major layout), kernels alone."""
    from audio_motion_avatar_b200 import stages as _st
    dev = ctx.dev
    out8 = torch.empty(1, Fr, size, size, 3, dtype=torch.uint8, pin_memory=True)
    with torch.no_grad():
        r.render_sequence(tok_h, cams, params, chunk=args.chunk, out=out8, out_dtype=torch.uint8)
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            r.render_sequence(tok_h, cams, params, chunk=args.chunk, out=out8, out_dtype=torch.uint8)
        torch.cuda.synchronize()
        sec8 = time.perf_counter() - t0
    flat = img.reshape(Fr, size, size, 3)
    dst8 = torch.empty(Fr, size, size, 3, dtype=torch.uint8, device=dev)
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        _st.frames_to_u8(flat, out=dst8)
    ea.record()
    for _ in range(10):
        _st.frames_to_u8(flat, out=dst8)
    eb.record()
    torch.cuda.synchronize()
    sink_ms = ea.elapsed_time(eb) / 10
    sink_bytes = flat.numel() * 5                                   # 4 B read + 1 B written per value
    line["sink"] = {"e2e_u8_frames_per_s": Fr * steps / sec8, "d2h_bytes_per_step": out8.numel(),
                    "bit_exact_vs_float_frames": bool(torch.equal(out8.reshape(Fr, size, size, 3).to(dev), _st.frames_to_u8(flat))),
                    "kernel_ms": sink_ms, "gbs": sink_bytes / (sink_ms * 1e-3) / 1e9,
                    "frac_of_peak": sink_bytes / (sink_ms * 1e-3) / 1e9 / peaks()[0],
                    "note": "frames (1.5 GB) exceed L2"}
    del out8, dst8
    Cc, Rr = int(r.cfg.triplane_feature_dim), int(r.cfg.triplane_resolution)
    tok_cm = tokens.reshape(Fr, Cc, 3 * Rr * Rr)
    tok_tm = tok_cm.transpose(1, 2).contiguous()
    Wd = r._decoder_wb()[0].detach().contiguous()
    pj = torch.empty(Fr, 3, Rr * Rr, 16, device=dev)
    res = {}
    for name, tk, tmf in (("channel_major", tok_cm, False), ("token_major", tok_tm, True)):
        for _ in range(3):
            _st.triplane_project(tk, Wd, Cc, Rr, out=pj, token_major=tmf)
        ea.record()
        for _ in range(10):
            _st.triplane_project(tk, Wd, Cc, Rr, out=pj, token_major=tmf)
        eb.record()
        torch.cuda.synchronize()
        ms = ea.elapsed_time(eb) / 10
        by = tk.numel() * 4 + pj.numel() * 4
        res[name] = {"kernel_ms": ms, "gbs": by / (ms * 1e-3) / 1e9, "frac_of_peak": by / (ms * 1e-3) / 1e9 / peaks()[0]}
    res["max_abs_difference"] = float((_st.triplane_project(tok_tm, Wd, Cc, Rr, token_major=True) -
                                       _st.triplane_project(tok_cm, Wd, Cc, Rr)).abs().max().item())
    line["producer_layout"] = res


def finish(ctx, line):
    """The JSON line is the LAST thing on stdout: NCCL's own INFO lines (NCCL_DEBUG is left to the caller) come before."""
    if ctx.world > 1:
        ctx.barrier()
        ctx.dist.destroy_process_group()
    sys.stdout.flush()
    if ctx.rank == 0:
        print(json.dumps(line), flush=True)


def run_train(args):
    """`--workload train`: BASELINE.json configs[3] alone as its own line (the default run carries it as `train`)."""
    ctx = Ctx()
    blk = train_block(ctx, args, steps=max(1, args.steps), warmup=max(3, args.warmup), size=args.size if args.size != SIZE else 512)
    line = {"metric": blk["metric"], "value": blk["value"], "unit": blk["unit"], "n_gpus": ctx.world, "steps": blk["steps"],
            "warmup": blk["warmup"], "ms_per_step": blk["ms_per_step"], "higher_is_better": True, "scaling": blk["scaling"],
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", **{k: blk[k] for k in
            ("config", "clocks", "gpu_launches", "loss", "stages", "library_kernel_ms", "step_over_library_kernel_time") if k in blk}}
    if "graph" in blk:
        line["graph"] = blk["graph"]
    finish(ctx, line)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "train":
        run_train(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
