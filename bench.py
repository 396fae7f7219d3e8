#!/usr/bin/env python
"""bench.py -- rendered frames/s of the render path (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1]): a 120-frame audio-driven sequence (one triplane per frame, 3x64x64x512 fp32),
100k Gaussians on the synthetic SMPL-X-shaped body (55 joints, level-2 densified), 1024x1024.  One STEP = one pass
of the whole path (LBS -> densify -> triplane project -> sample/decode -> splat) over the rank's 120-frame shard.
N > 1: every rank renders its own 120-frame shard of a 120*N-frame sequence (weak scaling; frames are independent,
no data-path collective) and rank 0 gathers the frames over NCCL.

`value`  : frames/s with all inputs resident in HBM (device timing, max over ranks).
`e2e`    : frames/s through Renderer.render_sequence with HOST (pinned) triplanes in and HOST frames out.
`--impl reference`: the reference's CPU implementation of the path (oracle port: torch CPU ops for
           LBS / grid_sample / Linear heads exactly as renderer.py calls them, C rasteriser with OpenMP), all host
           threads, one frame of the same workload per step.
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
    ap.add_argument("--gather-chunks", type=int, default=0,
                    help="N > 1, --gather nccl: pieces the shard is rendered and gathered in (0: 1 for N <= 2, else 4)")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N > 1: blend kernels store into rank 0's buffer over NVLink (symmetric memory), or NCCL send/recv")
    ap.add_argument("--workload", default="sequence", choices=["sequence", "train"],
                    help="sequence = BASELINE.json configs[1] (the headline metric); train = configs[3], a secondary line")
    ap.add_argument("--loss", default="l1_ssim", choices=["l1_ssim", "l1_torch"],
                    help="train workload: fused l1 + 0.1 (1 - ssim) kernels (lightning_model_wrapper.py:138-140) or a plain torch L1")
    ap.add_argument("--subjects", type=int, default=4)
    ap.add_argument("--views", type=int, default=4)
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


def synthetic_inputs(frames, size, device, seed):
    """Device-resident synthetic sequence: per-frame triplane = base + 0.05*noise, smooth poses,
    E = I cameras.  Tokens are generated on the device (3 GB of host RNG would take longer than the bench)."""
    from audio_motion_avatar_b200 import synthetic as S
    g = torch.Generator(device=device).manual_seed(seed)
    base = torch.randn(C, 3 * R * R, device=device, generator=g)
    tokens = torch.empty(1, frames, C, 3 * R * R, device=device)
    for f in range(frames):
        tokens[0, f] = base + 0.05 * torch.randn(C, 3 * R * R, device=device, generator=g)
    params = S.make_pose_sequence(1, frames, seed=seed)
    cams = S.make_cameras(1, frames, size, size)
    return tokens, params, cams


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
        return D * (4 + 48) + Fr * H * W * 12 + Fr * tiles * 4
    # backward stages (Fr = frames the stage sees; for decode / LBS stages that is the number of Gaussian sets)
    if name == "amav_splat_blend_backward":
        return D * (4 + 48) + Fr * H * W * (12 + 8) + Fr * N * 48 + Fr * tiles * 4
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


def cpu_reference(frames_per_step, steps, warmup, size, ngauss, seed=0):
    """The reference's CPU implementation of the path (oracle port), all host threads: `frames_per_step` frames of
    the workload per step.  Returns (frames/s, cores, seconds per step)."""
    from audio_motion_avatar_b200 import synthetic as S
    from oracle import ref_path as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    O.build_c_oracle()
    model = S.make_body_model(seed=0)
    w = S.make_decoder_weights(C, seed=0)
    levels, idx = None, None
    if ngauss != 10475:
        levels = S.subdivision_levels(10475, model["faces"].numpy(), subdivide_steps(ngauss))
        nd = 10475 + sum(len(e) for e in levels)
        idx = torch.randperm(nd, generator=torch.Generator().manual_seed(seed))[:ngauss]
    tokens = S.make_triplane_tokens(frames_per_step, C, R, seed=seed)[0]
    params = S.make_pose_sequence(1, frames_per_step, seed=seed)
    flat = {k: v.reshape(frames_per_step, *v.shape[2:]) for k, v in params.items()}
    cams = S.make_cameras(1, frames_per_step, size, size)
    K, E = cams["intrinsic"][0], cams["extrinsic"][0]
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        with torch.no_grad():
            O.reference_frames(model, w, tokens, flat, K, E, size, size, C, R, 1.4, levels, idx)
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return frames_per_step / sec, cores, sec


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    fps, cores, sec = cpu_reference(1, max(1, args.steps), max(0, min(args.warmup, 1)), args.size, args.gaussians)
    sample = f"1 frame of the {args.frames}-frame sequence per step ({args.gaussians} Gaussians, {args.size}x{args.size})"
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args), "frames_per_step": 1},
            "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_name(args):
    which = "configs[1]" if (args.frames, args.gaussians, args.size) == (FRAMES, NGAUSS, SIZE) else \
        "configs[4] sizes, secondary line" if (args.gaussians, args.size) == (500_000, 2048) else "non-default sizes, secondary line"
    return (f"{args.frames}-frame audio-driven sequence, {args.gaussians} Gaussians, {args.size}x{args.size}, "
            f"triplane 3x{R}x{R}x{C} fp32 per frame, 55-joint synthetic SMPL-X body (BASELINE.json {which})")


def run_ours(args):
    import torch.distributed as dist
    from audio_motion_avatar_b200 import _lib, synthetic as S
    from audio_motion_avatar_b200.renderer import Renderer, _workspace

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (there is no CPU path); use --impl reference for the CPU baseline")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.pop("NCCL_DEBUG", None)          # NCCL prints its version banner on stdout; stdout carries the JSON line
        dist.init_process_group("nccl", device_id=dev)
    warmup = max(3, args.warmup)
    steps = max(1, args.steps)
    Fr, N, size = args.frames, args.gaussians, args.size

    model = S.make_body_model(seed=0)
    r = Renderer(make_cfg(size, N, f"cuda:{local}"), body_model=model)
    r.gaussian_decoder.load_state_dict(S.make_decoder_weights(C, seed=0))
    if N != 10475:
        r.fix_site_subset(seed=0)
    tokens, params, cams = synthetic_inputs(Fr, size, dev, seed=rank)        # each rank: its own shard of the sequence
    params_d = {k: v.to(dev) for k, v in params.items()}
    cams_d = {k: v.to(dev) for k, v in cams.items()}
    from audio_motion_avatar_b200.shard import PeerFrameGather, render_and_gather
    from audio_motion_avatar_b200.renderer import deferred_overflow_checks
    gather_buf, peer, gather_mode = None, None, "none"
    chunks = args.gather_chunks if args.gather_chunks > 0 else (1 if world <= 2 else 4)
    if world > 1 and not args.no_gather:
        gather_mode = "nccl"
        if args.gather == "peer" and PeerFrameGather.available():
            try:
                peer = PeerFrameGather(world * Fr, size, size, dev, dst=0)
            except Exception as exc:                 # no peer access / allocator unavailable: NCCL send/recv instead
                print(f"[bench] symmetric memory unavailable ({type(exc).__name__}: {exc}); using the NCCL gather", file=sys.stderr)
            ok = torch.tensor([1.0 if peer is not None else 0.0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)            # every rank takes the same path
            if ok.item() > 0:
                gather_mode = "peer"
            else:
                peer = None
        if peer is None and rank == 0:
            gather_buf = torch.empty(world * Fr, size, size, 3, device=dev)

    def step():
        if peer is not None:
            # frame gather (north_star) fused into the blend kernel: pixels are stored into rank 0's buffer over NVLink
            return peer.render(r, tokens, cams_d, params_d)
        if world > 1 and not args.no_gather:
            # NCCL frame gather, overlapped with the render piece by piece
            return render_and_gather(r, tokens, cams_d, params_d, world * Fr, chunks=chunks, dst=0, out=gather_buf,
                                     defer=deferred_overflow_checks)
        with torch.no_grad():
            img, _ = r(tokens, cams_d, None, params_d)
        return img

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    launches0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks, _lib.StageTimer() as timer:
        barrier()
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    launches = (_lib.launch_count() - launches0) // steps
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / steps
    value = world * Fr / (ms_per_step * 1e-3)

    # per-stage device time inside the timed region -> the dominant stage's roofline
    stages_ms = {k: v[0] / steps for k, v in timer.summary().items()}
    D = int(_workspace(dev).last_D)
    body = r.smplx_model
    P = int(r._dense_par.shape[1]) if r._dense_par is not None else 1
    peak, peak_src = peaks()
    per_stage = {}
    for name, sms in stages_ms.items():
        by = stage_bytes(name, Fr, N, body.V, body.J, body.S, P, size, size, D)
        per_stage[name] = {"ms": round(sms, 4), "algorithmic_bytes": by, "gbs": round(by / (sms * 1e-3) / 1e9, 1),
                           "frac": round(by / (sms * 1e-3) / 1e9 / peak, 4)}
    top = max(stages_ms, key=stages_ms.get)
    traffic = None                                   # DRAM bytes per launch of the dominant kernel from the committed ncu capture
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as fh:
            tj = json.load(fh)
        if tj["workload"] == {"frames": Fr, "gaussians": N, "size": size}:
            traffic = tj["bytes_per_launch"].get(top)
    roofline = {"kernel": top, "bound": "hbm", "achieved": per_stage[top]["gbs"], "peak": peak, "unit": "GB/s",
                "frac": per_stage[top]["frac"], "traffic": traffic, "peak_source": peak_src,
                "share_of_step": round(stages_ms[top] / ms_per_step, 4), "stages": per_stage,
                "entries_D": D, "entries_per_frame": D // max(Fr, 1)}
    if "amav_splat_blend" in stages_ms and stages_ms["amav_splat_blend"] > 0:
        # the blend's true limiter is instruction issue, so its rate is also given in (pixel, list entry) pairs:
        # every list entry of a tile is a candidate for its 256 pixels (upstream evaluates exactly these pairs until the
        # tile saturates); ncu's issue-slot utilisation of the kernel is in profiles/ (83 % of peak, IPC 3.3 of 4)
        roofline["blend_pixel_entry_pairs_per_s"] = round(D * 256 / (stages_ms["amav_splat_blend"] * 1e-3), 0)
        roofline["blend_note"] = "compute-bound (FP32 issue), HBM fraction reported because the contract asks for hbm|tensor"

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args), "frames_per_rank": Fr, "frames_total": world * Fr,
                       "parallelism": f"frame-sharded x{world}" + {"none": "", "peer": " + frame gather to rank 0 by peer stores from the blend kernel (symmetric memory over NVLink)",
                                                                           "nccl": f" + NCCL frame gather to rank 0 overlapped with the render ({chunks} pieces)"}[gather_mode],
                       "l2": "inputs (3 GB of triplanes per step) exceed the 126 MB L2; no flush needed"},
            "clocks": clocks.summary(), "gpu_launches": int(launches), "roofline": roofline}

    # e2e: host (pinned) triplanes in, host frames out, through the public Renderer.render_sequence
    if not args.no_e2e:
        tok_h = torch.empty(tokens.shape, dtype=torch.float32, pin_memory=True)
        tok_h.copy_(tokens)
        out_h = torch.empty(1, Fr, size, size, 3, dtype=torch.float32, pin_memory=True)
        small_h = sum(v.numel() * 4 for v in params.values()) + sum(v.numel() * 4 for v in cams.values())
        for _ in range(2):
            r.render_sequence(tok_h, cams, params, chunk=args.chunk, out=out_h)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            with torch.no_grad():
                r.render_sequence(tok_h, cams, params, chunk=args.chunk, out=out_h)
        torch.cuda.synchronize()
        sec = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([sec], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            sec = float(t.item())
        line["e2e"] = {"value": world * Fr * steps / sec, "unit": UNIT, "h2d_bytes_per_step": tok_h.numel() * 4 + small_h,
                       "d2h_bytes_per_step": out_h.numel() * 4, "ms_per_step": sec / steps * 1e3,
                       "api": f"Renderer.render_sequence(chunk={args.chunk}), pinned host buffers, 3 streams"}
        # the pipelined host path must produce the frames of the device-resident path
        with torch.no_grad():
            img, _ = r(tokens, cams_d, None, params_d)               # this rank's shard, device-resident path
        torch.cuda.synchronize()
        line["e2e"]["max_abs_vs_device_path"] = float((out_h.to(dev) - img).abs().max().item())
        del out_h
        # the same call with the video sink on the device (uint8 frames leave the GPU), and the sink
        # kernel alone over this shard's frames (HBM-bound: 12 bytes in + 3 out per pixel)
        from audio_motion_avatar_b200 import stages as _st
        out8 = torch.empty(1, Fr, size, size, 3, dtype=torch.uint8, pin_memory=True)
        with torch.no_grad():
            r.render_sequence(tok_h, cams, params, chunk=args.chunk, out=out8, out_dtype=torch.uint8)
            barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                r.render_sequence(tok_h, cams, params, chunk=args.chunk, out=out8, out_dtype=torch.uint8)
            torch.cuda.synchronize()
            sec8 = time.perf_counter() - t0
        flat = img.reshape(Fr, size, size, 3)
        dst8 = torch.empty(Fr, size, size, 3, dtype=torch.uint8, device=dev)
        for _ in range(3):
            _st.frames_to_u8(flat, out=dst8)
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ea.record()
        for _ in range(10):
            _st.frames_to_u8(flat, out=dst8)
        eb.record()
        torch.cuda.synchronize()
        sink_ms = ea.elapsed_time(eb) / 10
        sink_bytes = flat.numel() * 5                                   # 4 B read + 1 B written per value
        line["sink"] = {"e2e_u8_frames_per_s": Fr * steps / sec8, "d2h_bytes_per_step": out8.numel(),
                        "bit_exact_vs_float_frames": bool(torch.equal(out8.reshape(Fr, size, size, 3).to(dev),
                                                                      _st.frames_to_u8(flat))),
                        "kernel_ms": sink_ms, "gbs": sink_bytes / (sink_ms * 1e-3) / 1e9,
                        "frac_of_peak": sink_bytes / (sink_ms * 1e-3) / 1e9 / peaks()[0],
                        "note": "this rank only; frames (1.5 GB) exceed L2"}
        del tok_h, out8, dst8
        # the projection fed in the producer's token-major layout [F, 3*R*R, C] (no permute+contiguous
        # upstream) next to the channel-major one, same tokens, kernels alone
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
        del tok_tm, pj

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        fps, cores, sec = cpu_reference(1, 1, 0, size, N)
        line["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"1 frame of the workload ({sec:.1f} s): torch CPU LBS/grid_sample/Linear + C rasteriser (OpenMP)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_train(args):
    """BASELINE.json configs[3]: one training step = `subjects` Gaussian sets (each its own triplane and pose) decoded
    once and rendered from `views` cameras, 100k Gaussians, 512x512, l1 + 0.1 (1 - ssim) image loss, full backward through the splat,
    the decode / triplane query and LBS, SGD step on the decoder.  N > 1: views are sharded over ranks (each rank
    renders `views` views of the same subjects: weak scaling) and dL/dtokens + decoder gradients are all-reduced."""
    import torch.distributed as dist
    from audio_motion_avatar_b200 import _lib, losses, synthetic as S
    from audio_motion_avatar_b200.renderer import Renderer, render_multi_view
    from audio_motion_avatar_b200.shard import allreduce_gradients
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (there is no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.pop("NCCL_DEBUG", None)          # NCCL prints its version banner on stdout; stdout carries the JSON line
        dist.init_process_group("nccl", device_id=dev)
    warmup, steps = max(3, args.warmup), max(1, args.steps)
    B, Tv, N = args.subjects, args.views, args.gaussians
    size = args.size if args.size != SIZE else 512
    model = S.make_body_model(seed=0)
    r = Renderer(make_cfg(size, N, f"cuda:{local}"), body_model=model)
    r.gaussian_decoder.load_state_dict(S.make_decoder_weights(C, seed=0))
    if N != 10475:
        r.fix_site_subset(seed=0)
    g = torch.Generator(device=dev).manual_seed(0)                                 # same subjects on every rank
    tokens = torch.randn(B, 1, C, 3 * R * R, device=dev, generator=g).requires_grad_(True)
    params = {k: v.reshape(B, 1, *v.shape[2:]).to(dev) for k, v in S.make_pose_sequence(1, B, seed=0).items()}
    params["transl"] = torch.tensor([0.0, 0.35, 0.0], device=dev).expand(B, 1, 3).contiguous()   # body at the orbit centre
    pose_leaves = {k: params[k].clone().requires_grad_(True) for k in ("global_orient", "body_pose", "betas", "expression", "transl")}
    params = {**params, **pose_leaves}
    cams = S.make_cameras(B, Tv, size, size, orbit=True)                              # 4 cameras on a 2.4 m circle
    cams = {k: v.to(dev) for k, v in cams.items()}
    target = torch.rand(B, Tv, size, size, 3, device=dev, generator=g)
    opt = torch.optim.SGD(r.gaussian_decoder.parameters(), lr=1e-4)
    leaves = [tokens] + list(pose_leaves.values())

    if args.loss == "l1_ssim":
        image_loss = lambda a, b: losses.l1_ssim_loss(a, b)[0]
    else:
        image_loss = lambda a, b: (a - b).abs().mean()

    def step():
        opt.zero_grad(set_to_none=True)
        for t in leaves:
            t.grad = None
        first = {k: v[:, :1] for k, v in cams.items()}
        img, gs = r(tokens, first, None, params)                                     # decode + view 0
        loss = image_loss(img, target[:, :1])
        if Tv > 1:
            mv = render_multi_view(gs, cams["intrinsic"][:, 1:], cams["extrinsic"][:, 1:], r.cfg)
            loss = (loss + image_loss(mv, target[:, 1:]) * (Tv - 1)) / Tv
        loss.backward()
        if world > 1:
            allreduce_gradients([tokens.grad] + [p.grad for p in r.gaussian_decoder.parameters()])
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    launches0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks, _lib.StageTimer() as timer:
        barrier()
        e0.record()
        for _ in range(steps):
            loss = step()
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    launches = (_lib.launch_count() - launches0) // steps
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / steps
    stages_ms = {k: (v[0] / steps, v[1] // steps) for k, v in timer.summary().items()}
    peak, peak_src = peaks()
    body = r.smplx_model
    P = int(r._dense_par.shape[1]) if r._dense_par is not None else 1
    per_stage = {k: {"ms": round(v[0], 4), "calls": v[1]} for k, v in stages_ms.items()}
    line = {"metric": "training renders/sec (fwd+bwd) @100k Gaussians 512^2", "value": world * B * Tv / (ms_per_step * 1e-3),
            "unit": "(subject,view) renders/s", "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"training step: {B} subjects x {Tv} views per rank, {N} Gaussians, {size}x{size}, loss {args.loss}, forward + "
                                   "backward through splat, decode/triplane query and LBS, SGD on the decoder "
                                   "(BASELINE.json configs[3])",
                       "parallelism": f"view-sharded x{world}" + ("" if world == 1 else " + NCCL all-reduce of dL/dtokens and decoder grads")},
            "clocks": clocks.summary(), "gpu_launches": int(launches), "loss": float(loss.item()),
            "stages": per_stage, "peak_hbm_gbs": peak, "peak_source": peak_src,
            "body": {"V": body.V, "J": body.J, "P": P}}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


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
