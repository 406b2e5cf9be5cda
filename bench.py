#!/usr/bin/env python
"""bench.py -- stereo frames/s of the per-frame VO front end at KITTI shape (BASELINE.json).

  python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W

Workload (config.workload): BASELINE.json configs[1] -- synthetic KITTI-shape stereo 1241x376 uint8,
kitti00.yaml calibration, ~2000 tracked features per frame, one sequence per GPU.  A "step" is
FRAMES_PER_STEP consecutive stereo frames through VisualOdometry::stereo_callback.
  value : frames/s with every frame already resident in HBM (svo_process_frame_batch_dev, stream
          ordered, no host sync inside the timed region), CUDA events on the context's stream.
  e2e   : frames/s through the reference-facing C ABI with pinned HOST images (svo_submit_frame / svo_collect,
          <= 3 frames in flight; H2D copy of every frame, kernels, D2H of every result inside the timed region);
          e2e.sync_per_frame_value is the same through the synchronous svo_process_frame (one sync per frame).
  N > 1 : BASELINE.json configs[4] -- 64 independent sequences sharded s mod N, 64/N batched per GPU, strong
          scaling, the NCCL gather of the transforms inside the timed region (run_config5).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 1241, 376
PITCH = 1280   # row pitch of the device-resident frames (a multiple of 128 bytes)
FRAMES_PER_STEP = 400
E2E_DEPTH = int(os.environ.get("SVO_BENCH_DEPTH", "3"))   # frames in flight of the e2e arm
# dram__bytes_read.sum + dram__bytes_write.sum of the five launches of one 64-frame ingest set, from the round-2
# ncu --set full capture profiles/r02_ingest_full.raw.csv (rows 0..4): pyrDown 62.02 + 52.13, 15.44 + 0.37, 4.70 + 0.00,
# FAST 35.53 + 0.07, gated FAST 0.03 MB
NCU_INGEST_TRAFFIC_BYTES_PER_FRAME = 170.29e6 / 64
REF_FRAMES_PER_STEP = 8
METRIC = "stereo frames/s at KITTI 1241x376"


def pyramid_bytes(w, h, levels=3):
    tot, lw, lh = 0, w, h
    for _ in range(levels):
        lw, lh = (lw + 1) // 2, (lh + 1) // 2
        tot += lw * lh
    return tot


# each new u8 pixel read once + pyramid levels 1..3 written once, both images
B_ALG_FRAME = 2 * W * H + 2 * pyramid_bytes(W, H)


def render_frames(n_unique, seed, out_l=None, out_r=None, threads=None):
    """n_unique stereo frames of synthetic sequence `seed` (synthdata.py; cv2 releases the GIL, so threads scale)."""
    from concurrent.futures import ThreadPoolExecutor
    import synthdata as synth
    seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=n_unique, seed=seed)
    synth.texture_atlas(1234, seq.density)   # build the shared atlas once, before the threads start
    L = np.empty((n_unique, H, W), np.uint8) if out_l is None else out_l
    R = np.empty((n_unique, H, W), np.uint8) if out_r is None else out_r

    def one(i):
        L[i], R[i] = seq.render(i)

    with ThreadPoolExecutor(max_workers=threads or min(16, os.cpu_count() or 1)) as ex:
        list(ex.map(one, range(n_unique)))
    return seq, L, R


def pingpong(i, n):
    """Endless physically valid playback of n rendered frames: forward, then backward, ..."""
    p = i % (2 * n - 2) if n > 1 else 0
    return p if p < n else 2 * n - 2 - p


class ClockSampler:
    def __init__(self, dev):
        self.dev, self.rows, self.proc = dev, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.dev), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for k, nm in enumerate(names):
                    if r[2 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU path (OpenCV kernels via the cv2 oracle restatement,
    oracle/ref_cv2.py; the C++ reference cannot be built here: no OpenCV C++ headers/libs)."""
    if rank != 0:
        return
    import cv2
    from oracle import ref_cv2 as O
    cores = os.cpu_count() or 1
    cv2.setNumThreads(cores)
    n_unique = min(64, 2 + REF_FRAMES_PER_STEP * 2)
    seq, L, R = render_frames(n_unique, 0)
    vo = O.RefVO(seq.Pl, seq.Pr)
    vo.stereo_callback(L[0], R[0])
    k = 1

    def step():
        nonlocal k
        for _ in range(REF_FRAMES_PER_STEP):
            j = pingpong(k, n_unique)
            vo.stereo_callback(L[j], R[j])
            k += 1

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    fps = args.steps * REF_FRAMES_PER_STEP / dt
    sample = f"{args.steps} steps x {REF_FRAMES_PER_STEP} frames of the same synthetic KITTI-shape sequence"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "synthetic KITTI-shape stereo 1241x376 uint8, kitti00.yaml, ~2000 features (configs[1])",
                   "frames_per_step": REF_FRAMES_PER_STEP},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample,
                         "note": "cv2 (OpenCV 4.13) kernels = the same library calls the reference makes"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def cpu_baseline(seq, L, R, budget_s=12.0):
    import cv2
    from oracle import ref_cv2 as O
    out = {}
    for cores in (1, os.cpu_count() or 1):
        cv2.setNumThreads(cores)
        vo = O.RefVO(seq.Pl, seq.Pr)
        vo.stereo_callback(L[0], R[0])
        vo.stereo_callback(L[1], R[1])
        t0 = time.perf_counter()
        n, k = 0, 2
        stages = {}
        while time.perf_counter() - t0 < budget_s / 2 and n < 200:
            j = pingpong(k, len(L))
            vo.stereo_callback(L[j], R[j])
            for kk, v in vo.taps.times.items():
                stages[kk] = stages.get(kk, 0.0) + v
            k += 1
            n += 1
        dt = time.perf_counter() - t0
        out[cores] = (n / dt, n, {kk: 1e3 * v / n for kk, v in stages.items()})
    best = max(out, key=lambda c: out[c][0])
    return {"value": out[best][0], "unit": "frames/s", "cores": best, "kind": "port",
            "sample": f"{out[best][1]} frames of the same sequence, cv2 oracle (OpenCV kernels the reference calls)",
            "single_thread_value": out[1][0], "stage_ms_single_thread": out[1][2]}


N_SEQ5 = 64            # BASELINE.json configs[4]: batch of 64 independent sequences
FRAMES5_PER_STEP = 25  # frames of every sequence per step
T_OFF = 16             # svo_frame_result as float64[32]: transform = words 16..31, success = int32 word 0


def run_config5(args, rank, world, local_rank):
    """N > 1: BASELINE.json configs[4] -- 64 independent synthetic KITTI-shape sequences (generator seed = sequence
    id), sequence s on rank s mod N (dist.shard_sequences), every rank batches its 64 / N sequences through one
    context (sequence = grid dimension), no data-path collective.  All ranks time their shard between barriers; the
    one exchange of the path -- the gather of the [n_local, frames, 16] transforms over NCCL -- is INSIDE the timed
    region.  value = 64 x frames / max-over-ranks time (strong scaling: the total work is fixed)."""
    import torch
    import torch.distributed as dist
    from stereo_visual_odometry_b200 import dist as svo_dist
    from stereo_visual_odometry_b200._lib import FrameResult, check, default_params, lib, ptr

    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L_ = lib()
    fsz = C.sizeof(FrameResult)
    mine = svo_dist.shard_sequences(N_SEQ5, rank, world)
    n_local, U = len(mine), args.frames5
    t_r = time.perf_counter()
    Lh = torch.empty((U, n_local, H, W), dtype=torch.uint8).pin_memory()
    Rh = torch.empty((U, n_local, H, W), dtype=torch.uint8).pin_memory()
    Ln, Rn = Lh.numpy(), Rh.numpy()
    seq0 = None
    thr = max(1, (os.cpu_count() or 1) // world)
    for q, sid in enumerate(mine):
        seq, _, _ = render_frames(U, sid, Ln[:, q], Rn[:, q], threads=thr)
        seq0 = seq0 or seq
    render_s = time.perf_counter() - t_r
    Ld = torch.zeros((U, n_local, H, PITCH), dtype=torch.uint8, device="cuda")   # 128-byte aligned row pitch
    Rd = torch.zeros((U, n_local, H, PITCH), dtype=torch.uint8, device="cuda")
    Ld[..., :W].copy_(Lh, non_blocking=True)
    Rd[..., :W].copy_(Rh, non_blocking=True)
    torch.cuda.synchronize()

    def make_ctx(n_seq):
        p = default_params(W, H, n_seq=n_seq, device=local_rank)
        ctx_ = L_.svo_create(C.byref(p))
        if not ctx_:
            raise RuntimeError("svo_create: " + L_.svo_last_error().decode())
        Pl, Pr = np.ascontiguousarray(seq0.Pl), np.ascontiguousarray(seq0.Pr)
        check(L_.svo_set_projection(ctx_, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
        return ctx_

    n_total = (args.warmup + args.steps) * FRAMES5_PER_STEP + 1
    res_dev = torch.zeros((n_total, n_local, fsz // 8), dtype=torch.float64, device="cuda")
    ctx = make_ctx(n_local)
    stream = torch.cuda.ExternalStream(L_.svo_stream(ctx))
    Lptr = [Ld[j].data_ptr() for j in range(U)]
    Rptr = [Rd[j].data_ptr() for j in range(U)]
    rp = res_dev.data_ptr()

    def enqueue(i):
        j = pingpong(i, U)
        check(L_.svo_process_frame_batch_dev(ctx, Lptr[j], Rptr[j], PITCH, PITCH * H, rp + i * n_local * fsz))

    k = 0
    for _ in range(args.warmup * FRAMES5_PER_STEP + 1):
        enqueue(k)
        k += 1
    check(L_.svo_synchronize(ctx))
    # warm the NCCL communicator up (its first collective builds the rings): not part of the path's steady state
    svo_dist.gather_results(torch.zeros((n_local, 4, 16), dtype=torch.float64, device="cuda"), dst=0)
    dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = L_.svo_launch_count(ctx)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0 = k
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(args.steps * FRAMES5_PER_STEP):
            enqueue(k)
            k += 1
        check(L_.svo_stream_join(ctx))   # the primary stream waits for the context's other streams
        # the path's only exchange: transforms (+ success flags in word 0) of every sequence, one NCCL gather
        local = res_dev[k0:k, :, T_OFF:T_OFF + 16].transpose(0, 1).contiguous()           # [n_local, frames, 16]
        flags = res_dev[k0:k, :, 0].transpose(0, 1).contiguous()                          # int32 pair as one f64 word
        gathered = svo_dist.gather_results(torch.cat([local, flags.unsqueeze(-1)], -1), dst=0)
        e1.record(stream)
    check(L_.svo_synchronize(ctx))
    torch.cuda.synchronize()
    ms = svo_dist.max_over_ranks(e0.elapsed_time(e1), device="cuda")
    launches = L_.svo_launch_count(ctx) - launches0
    clocks = sampler.stop()
    n_frames = args.steps * FRAMES5_PER_STEP
    value = N_SEQ5 * n_frames / (ms * 1e-3)

    # ---------------- e2e: pinned HOST frames through svo_submit_frame_batch / svo_collect (H2D of every frame of
    # every local sequence and D2H of every result inside the timed region), same sharding, max over ranks
    ctx2 = make_ctx(n_local)
    resv = (FrameResult * n_local)()
    Lhp = [Lh[j].data_ptr() for j in range(U)]
    Rhp = [Rh[j].data_ptr() for j in range(U)]
    kk = 0

    def run_stream(n):
        nonlocal kk
        ok = 0
        for _ in range(n):
            j = pingpong(kk, U)
            check(L_.svo_submit_frame_batch(ctx2, Lhp[j], Rhp[j], W, W * H))
            kk += 1
            if L_.svo_pending(ctx2) > E2E_DEPTH:
                ok += check(L_.svo_collect(ctx2, resv))
        while L_.svo_pending(ctx2) > 0:
            ok += check(L_.svo_collect(ctx2, resv))
        return ok

    run_stream(8)
    dist.barrier()
    n_e2e = min(n_frames, 4 * FRAMES5_PER_STEP)
    t0 = time.perf_counter()
    ok_e2e = run_stream(n_e2e)
    dt = svo_dist.max_over_ranks(time.perf_counter() - t0, device="cuda")
    L_.svo_destroy(ctx2)
    e2e = {"value": N_SEQ5 * n_e2e / dt, "unit": "frames/s", "h2d_bytes_per_step": 2 * W * H * N_SEQ5 * FRAMES5_PER_STEP,
           "d2h_bytes_per_step": fsz * N_SEQ5 * FRAMES5_PER_STEP, "frames_timed": N_SEQ5 * n_e2e,
           "success_rate_rank0": ok_e2e / (n_e2e * n_local), "frames_in_flight": E2E_DEPTH,
           "timing": "host wall clock, max over ranks, around svo_submit_frame_batch / svo_collect with pinned host "
                     "frames (H2D of every frame and D2H of every result inside the timed region)"}

    # ---------------- the same thing one sequence per GPU (weak scaling of configs[1]; round-1's N > 1 number)
    ctx1 = make_ctx(1)
    res1 = torch.zeros((n_total * 4, fsz // 8), dtype=torch.float64, device="cuda")
    s1 = torch.cuda.ExternalStream(L_.svo_stream(ctx1))
    L1 = [Ld[j, 0].data_ptr() for j in range(U)]
    R1 = [Rd[j, 0].data_ptr() for j in range(U)]
    for i in range(60):
        check(L_.svo_process_frame_batch_dev(ctx1, L1[pingpong(i, U)], R1[pingpong(i, U)], PITCH, PITCH * H, None))
    check(L_.svo_synchronize(ctx1))
    dist.barrier()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record(s1)
    n_one = 800
    for i in range(60, 60 + n_one):
        check(L_.svo_process_frame_batch_dev(ctx1, L1[pingpong(i, U)], R1[pingpong(i, U)], PITCH, PITCH * H, None))
    check(L_.svo_stream_join(ctx1))
    a1.record(s1)
    check(L_.svo_synchronize(ctx1))
    ms_one = svo_dist.max_over_ranks(a0.elapsed_time(a1), device="cuda")
    L_.svo_destroy(ctx1)

    out = None
    if rank == 0:
        full = svo_dist.interleave_sharded([g.cpu() for g in gathered], N_SEQ5)            # [64, frames, 17]
        succ = np.ascontiguousarray(full[:, :, 16].numpy()).view(np.int32).reshape(N_SEQ5, n_frames, 2)[:, :, 0]
        # bit-for-bit check of sequences that other ranks own against a local re-run (a sequence's result does not
        # depend on which GPU or which batch it runs in)
        check_ids = sorted({1 % N_SEQ5, N_SEQ5 - 1} - set(mine)) or [mine[-1]]
        Lc = torch.empty((U, len(check_ids), H, W), dtype=torch.uint8)
        Rc = torch.empty((U, len(check_ids), H, W), dtype=torch.uint8)
        for q, sid in enumerate(check_ids):
            render_frames(U, sid, Lc.numpy()[:, q], Rc.numpy()[:, q])
        Lcd = torch.zeros((U, len(check_ids), H, PITCH), dtype=torch.uint8, device="cuda")
        Rcd = torch.zeros((U, len(check_ids), H, PITCH), dtype=torch.uint8, device="cuda")
        Lcd[..., :W].copy_(Lc)
        Rcd[..., :W].copy_(Rc)
        ctxc = make_ctx(len(check_ids))
        resc = torch.zeros((k, len(check_ids), fsz // 8), dtype=torch.float64, device="cuda")
        for i in range(k):
            j = pingpong(i, U)
            check(L_.svo_process_frame_batch_dev(ctxc, Lcd[j].data_ptr(), Rcd[j].data_ptr(), PITCH, PITCH * H,
                                                 resc.data_ptr() + i * len(check_ids) * fsz))
        check(L_.svo_synchronize(ctxc))
        torch.cuda.synchronize()
        L_.svo_destroy(ctxc)
        ref = resc[k0:k, :, T_OFF:T_OFF + 16].transpose(0, 1).cpu()
        equal = all(bool(torch.equal(ref[q], full[sid, :, :16])) for q, sid in enumerate(check_ids))
        out = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "batch of 64 independent synthetic KITTI-shape sequences 1241x376 uint8, kitti00.yaml "
                                   "(BASELINE.json configs[4]), sequence s on GPU s mod N, 64/N sequences batched per GPU",
                       "sequences": N_SEQ5, "sequences_per_gpu": n_local, "frames_per_step": FRAMES5_PER_STEP,
                       "unique_frames_per_sequence": U, "playback": "ping-pong", "seed": "sequence id",
                       "l2": "inputs larger than L2: %d MB of distinct frames per GPU cycle through a 126 MB L2" % (2 * U * n_local * W * H >> 20),
                       "success_rate": float(succ.mean())},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "gather": {"inside_timed_region": True, "collective": "one NCCL gather of [n_local, frames, 17] float64 per rank",
                       "bytes_per_rank": int(n_local * n_frames * 17 * 8)},
            "verify": {"sequences_rerun_on_rank0": check_ids, "bit_equal_to_the_gathered_transforms": bool(equal)},
            "one_sequence_per_gpu": {"value": world * n_one / (ms_one * 1e-3), "unit": "frames/s", "scaling": "weak",
                                     "note": "configs[1] replicated: one single-sequence streaming context per GPU"},
            "reference_arm_note": "--impl reference is ONE CPU process (all host threads) whatever --gpus says",
            "render_s": render_s,
        }
        print(json.dumps(out))
    L_.svo_destroy(ctx)
    dist.barrier()
    dist.destroy_process_group()


def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from stereo_visual_odometry_b200 import _lib
    from stereo_visual_odometry_b200._lib import FrameResult, check, default_params, lib, ptr

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L_ = lib()
    n_unique = args.frames
    t_r = time.perf_counter()
    seq, Lh, Rh = render_frames(n_unique, seed=rank)
    render_s = time.perf_counter() - t_r
    Lp, Rp = torch.from_numpy(Lh).pin_memory(), torch.from_numpy(Rh).pin_memory()
    # device-resident frames: rows padded to a 128-byte multiple (the caller's stride is a parameter of the C ABI), so
    # that the source planes qualify for the TMA tile loads
    Ld = torch.zeros((n_unique, H, PITCH), dtype=torch.uint8, device="cuda")
    Rd = torch.zeros((n_unique, H, PITCH), dtype=torch.uint8, device="cuda")
    Ld[:, :, :W].copy_(Lp, non_blocking=True)
    Rd[:, :, :W].copy_(Rp, non_blocking=True)
    torch.cuda.synchronize()

    def make_ctx(n_seq=1):
        p = default_params(W, H, n_seq=n_seq, device=local_rank)
        ctx = L_.svo_create(C.byref(p))
        if not ctx:
            raise RuntimeError("svo_create: " + L_.svo_last_error().decode())
        Pl, Pr = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
        check(L_.svo_set_projection(ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
        return ctx

    n_total = (args.warmup + args.steps) * FRAMES_PER_STEP + 1
    res_dev = torch.zeros((n_total, C.sizeof(FrameResult)), dtype=torch.uint8, device="cuda")

    # ---------------- value: frames resident in HBM, stream-ordered, no host sync in the timed region
    ctx = make_ctx()
    stream = torch.cuda.ExternalStream(L_.svo_stream(ctx))
    fsz = C.sizeof(FrameResult)
    k = 0

    # buffer addresses of the frames (device-resident and pinned-host copies), looked up once
    Ld_ptr = [Ld[j].data_ptr() for j in range(n_unique)]
    Rd_ptr = [Rd[j].data_ptr() for j in range(n_unique)]
    Lp_ptr = [Lp[j].data_ptr() for j in range(n_unique)]
    Rp_ptr = [Rp[j].data_ptr() for j in range(n_unique)]
    res_ptr = res_dev.data_ptr()

    def enqueue(ctx_, i):
        j = pingpong(i, n_unique)
        check(L_.svo_process_frame_batch_dev(ctx_, Ld_ptr[j], Rd_ptr[j], PITCH, PITCH * H, res_ptr + i * fsz))

    enqueue(ctx, 0)
    k = 1
    for _ in range(args.warmup):
        for _ in range(FRAMES_PER_STEP):
            enqueue(ctx, k)
            k += 1
    check(L_.svo_synchronize(ctx))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = L_.svo_launch_count(ctx)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    k_timed0 = k
    for _ in range(args.steps):
        for _ in range(FRAMES_PER_STEP):
            enqueue(ctx, k)
            k += 1
    check(L_.svo_stream_join(ctx))   # the primary stream waits for the context's other streams
    e1.record(stream)
    check(L_.svo_synchronize(ctx))
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    launches = L_.svo_launch_count(ctx) - launches0
    # trajectories: gather over NCCL (the only exchange of the path), inside the step budget
    traj = res_dev[k_timed0:k].clone()
    from stereo_visual_odometry_b200 import dist as svo_dist
    gathered = svo_dist.gather_results(traj, dst=0)      # one NCCL gather: the path's only exchange
    ms = svo_dist.max_over_ranks(ms, device="cuda")      # multi-GPU time = max over ranks
    clocks = sampler.stop()
    res = np.frombuffer(traj.cpu().numpy().tobytes(), dtype=np.dtype([
        ("i", np.int32, 8), ("R", np.float64, 9), ("t", np.float64, 3), ("T", np.float64, 16)]))
    n_frames_timed = args.steps * FRAMES_PER_STEP
    success_rate = float(res["i"][:, 0].mean())
    mean_matched = float(res["i"][:, 3].mean())
    mean_bucketed = float(res["i"][:, 2].mean())
    mean_iters = float(res["i"][:, 7].mean())
    value = world * n_frames_timed / (ms * 1e-3)

    # accuracy vs ground truth over the timed frames (independent of the oracle)
    terr = []
    for q in range(min(n_frames_timed, 300)):
        i = k_timed0 + q
        a, b = pingpong(i - 1, n_unique), pingpong(i, n_unique)
        T0 = np.eye(4); T0[:3, :3], T0[:3, 3] = seq.R_wc[a], seq.C[a]
        T1 = np.eye(4); T1[:3, :3], T1[:3, 3] = seq.R_wc[b], seq.C[b]
        G = np.linalg.inv(T0) @ T1
        if res["i"][q, 0]:
            terr.append(np.linalg.norm(res["T"][q].reshape(4, 4)[:3, 3] - G[:3, 3]) / max(np.linalg.norm(G[:3, 3]), 1e-9))

    out = None
    if rank == 0:
        # ---------------- per-stage device times (separate pass, CUDA events between stages on the ctx stream)
        check(L_.svo_enable_timers(ctx, 1))
        check(L_.svo_reset_timers(ctx))
        n_t = 200
        for _ in range(n_t):
            enqueue(ctx, min(k, n_total - 1) if False else k % n_total)
            k += 1
        names = ["ingest_fast_pyramid", "bucket", "klt", "select_triangulate", "pnp_hypotheses", "pnp_refit"]
        stages = {}
        for s_, nm in enumerate(names):
            msv, ln = C.c_double(0), C.c_longlong(0)
            L_.svo_get_stage_time(ctx, s_, C.byref(msv), C.byref(ln))
            stages[nm] = {"us_per_frame": 1e3 * msv.value / n_t, "launches_per_frame": ln.value / n_t}
        check(L_.svo_enable_timers(ctx, 0))
        # ---------------- look-ahead ingest of ONE trajectory: K frames per launch set
        lookahead = {}
        for K in (8, 16):
            pk = default_params(W, H, device=local_rank, lookahead=K)
            ck = L_.svo_create(C.byref(pk))
            Plk, Prk = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
            check(L_.svo_set_projection(ck, ptr(Plk, C.c_float), ptr(Prk, C.c_float)))
            check(L_.svo_enable_timers(ck, 1))
            frk = FrameResult()
            jj = 0
            for rep in range(6):
                if rep == 2:
                    check(L_.svo_reset_timers(ck))
                lp = (C.c_void_p * K)(*[Lp_ptr[pingpong(jj + q, n_unique)] for q in range(K)])
                rp = (C.c_void_p * K)(*[Rp_ptr[pingpong(jj + q, n_unique)] for q in range(K)])
                check(L_.svo_submit_frames(ck, K, lp, rp, W))
                jj += K
                while L_.svo_pending(ck) > 0:
                    check(L_.svo_collect(ck, C.byref(frk)))
            msv, ln = C.c_double(0), C.c_longlong(0)
            L_.svo_get_stage_time(ck, 0, C.byref(msv), C.byref(ln))
            lookahead[K] = 1e3 * msv.value / 4   # us per K-frame launch set
            L_.svo_destroy(ck)
    L_.svo_destroy(ctx)

    e2e = None
    if True:
        # ---------------- e2e: the reference-facing C-ABI calls with pinned HOST images; every frame's H2D copy
        # and the D2H of its result struct are inside the timed region.
        #   streaming : svo_submit_frame / svo_collect, <= E2E_DEPTH frames in flight (a camera that delivers
        #               frames back to back; results identical to the synchronous call, in order)
        #   sync      : svo_process_frame, one host sync per frame (pure latency mode)
        ctx2 = make_ctx()
        T = np.zeros(16, np.float64)
        fr = FrameResult()
        n_e2e = min(n_frames_timed, 1000)
        kk = 0

        def run_stream(n):
            nonlocal kk
            ok = 0
            for _ in range(n):
                j = pingpong(kk, n_unique)
                check(L_.svo_submit_frame(ctx2, Lp_ptr[j], Rp_ptr[j], W))
                kk += 1
                if L_.svo_pending(ctx2) > E2E_DEPTH:
                    ok += check(L_.svo_collect(ctx2, C.byref(fr)))
            while L_.svo_pending(ctx2) > 0:
                ok += check(L_.svo_collect(ctx2, C.byref(fr)))
            return ok

        run_stream(51)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        ok_stream = run_stream(n_e2e)
        dt = time.perf_counter() - t0
        dt = svo_dist.max_over_ranks(dt, device="cuda")
        t0 = time.perf_counter()
        n_sync = min(n_e2e, 300)
        for _ in range(n_sync):
            j = pingpong(kk, n_unique)
            check(L_.svo_process_frame(ctx2, Lp_ptr[j], Rp_ptr[j], W, ptr(T, C.c_double)))
            kk += 1
        dt_sync = time.perf_counter() - t0
        e2e = {"value": world * n_e2e / dt, "unit": "frames/s", "h2d_bytes_per_step": 2 * W * H * FRAMES_PER_STEP,
               "d2h_bytes_per_step": C.sizeof(FrameResult) * FRAMES_PER_STEP, "frames_timed": n_e2e,
               "success_rate": ok_stream / n_e2e, "frames_in_flight": E2E_DEPTH,
               "timing": "host wall clock around svo_submit_frame/svo_collect with pinned host images "
                         "(H2D of every frame and D2H of every result inside the timed region)",
               "sync_per_frame_value": n_sync / dt_sync}
        L_.svo_destroy(ctx2)

    if rank == 0:
        # ---------------- batched ingest/pipeline (config 5 shape on this rank): 64/N sequences per launch
        n_b = max(1, 64 // world)
        batch = None
        try:
            nfb, U5 = 40, args.frames5
            ctxb = make_ctx(n_seq=n_b)
            # 64 distinct sequences (generator seed = sequence id), U5 rendered frames each, ping-pong playback
            Lb5 = torch.empty((U5, n_b, H, W), dtype=torch.uint8)
            Rb5 = torch.empty((U5, n_b, H, W), dtype=torch.uint8)
            for s_ in range(n_b):
                render_frames(U5, s_, Lb5.numpy()[:, s_], Rb5.numpy()[:, s_])
            Lb5d = torch.zeros((U5, n_b, H, PITCH), dtype=torch.uint8, device="cuda")
            Rb5d = torch.zeros((U5, n_b, H, PITCH), dtype=torch.uint8, device="cuda")
            Lb5d[..., :W].copy_(Lb5)
            Rb5d[..., :W].copy_(Rb5)
            sb = torch.cuda.ExternalStream(L_.svo_stream(ctxb))
            check(L_.svo_enable_timers(ctxb, 0))
            Lb = [Lb5d[pingpong(f, U5)] for f in range(nfb)]
            Rb = [Rb5d[pingpong(f, U5)] for f in range(nfb)]
            torch.cuda.synchronize()
            for f in range(5):
                check(L_.svo_process_frame_batch_dev(ctxb, Lb[f].data_ptr(), Rb[f].data_ptr(), PITCH, PITCH * H, None))
            check(L_.svo_synchronize(ctxb))
            b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            b0.record(sb)
            for f in range(5, nfb):
                check(L_.svo_process_frame_batch_dev(ctxb, Lb[f].data_ptr(), Rb[f].data_ptr(), PITCH, PITCH * H, None))
            check(L_.svo_stream_join(ctxb))
            b1.record(sb)
            check(L_.svo_synchronize(ctxb))
            bms = b0.elapsed_time(b1)
            check(L_.svo_enable_timers(ctxb, 1))
            check(L_.svo_reset_timers(ctxb))
            for f in range(nfb - 1, nfb - 11, -1):
                check(L_.svo_process_frame_batch_dev(ctxb, Lb[f].data_ptr(), Rb[f].data_ptr(), PITCH, PITCH * H, None))
            msv, ln = C.c_double(0), C.c_longlong(0)
            L_.svo_get_stage_time(ctxb, 0, C.byref(msv), C.byref(ln))
            ing_us = 1e3 * msv.value / 10
            fr = FrameResult()
            L_.svo_get_result(ctxb, 0, C.byref(fr))
            batch = {"sequences_per_launch": n_b, "frames_per_s": n_b * (nfb - 5) / (bms * 1e-3),
                     "ingest_us_per_launch": ing_us,
                     "ingest_GBps": B_ALG_FRAME * n_b / (ing_us * 1e-6) / 1e9, "last_matched": fr.n_matched}
            L_.svo_destroy(ctxb)
        except Exception as ex:   # keep the headline line alive
            batch = {"error": str(ex)[:200]}

        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (of fallback)"
        # roofline of the HBM-side stage: ingest = pyrDown L0->L1 with the fused level-0 copy, L1->L2,
        # L2->L3, FAST/NMS/bucket keys.  The figure that can approach the HBM pipe at all is the batched launch
        # (64 frames = 79 MB algorithmic per launch set); the single-frame launch is reported beside it.
        # traffic: dram__bytes_read + dram__bytes_write of the same launches from the committed ncu capture
        # (profiles/r01_ncu_summary.md), per launch set like `achieved`; the kernels are issue bound, not HBM bound.
        ing = stages["ingest_fast_pyramid"]
        ach1 = B_ALG_FRAME / (ing["us_per_frame"] * 1e-6) / 1e9
        roofline_single = {"bound": "hbm", "kernel": "ingest, 1 frame per launch (5 launches)", "achieved": ach1,
                           "peak": peak, "unit": "GB/s", "frac": ach1 / peak, "us_per_launch_set": ing["us_per_frame"],
                           "algorithmic_bytes_per_launch_set": B_ALG_FRAME}
        if batch and "ingest_GBps" in batch:
            nfl = batch["sequences_per_launch"]
            roofline = {"bound": "hbm",
                        "kernel": "ingest: pyrdown_tile_kernel x3 (level-0 copy fused) + fast_kernel<128> x2 (thr-20 pass, gated low-threshold pass), %d frames per launch" % nfl,
                        "achieved": batch["ingest_GBps"], "peak": peak, "unit": "GB/s",
                        "frac": batch["ingest_GBps"] / peak, "traffic": (NCU_INGEST_TRAFFIC_BYTES_PER_FRAME * nfl) if NCU_INGEST_TRAFFIC_BYTES_PER_FRAME else None,
                        "traffic_source": "ncu --set full capture of the same launches: profiles/r02_ingest_full.raw.csv (summary: profiles/r02_ncu_summary.md)",
                        "peak_source": peak_src, "frames_per_launch": nfl,
                        "algorithmic_bytes_per_launch_set": B_ALG_FRAME * nfl,
                        "us_per_launch_set": batch["ingest_us_per_launch"],
                        "limiter": "SM issue (FAST 62 %, pyrDown 50-61 % issue-slot utilisation; FAST is 89 of the 149 us), not HBM: see profiles/r02_ncu_summary.md"}
        else:
            roofline = dict(roofline_single, traffic=None, peak_source=peak_src)
        # one trajectory with look-ahead: K = 8 frames per ingest launch set (svo_submit_frames)
        roofline_lookahead = {}
        for K_, usk in lookahead.items():
            ak = K_ * B_ALG_FRAME / (usk * 1e-6) / 1e9
            roofline_lookahead["K%d" % K_] = {"bound": "hbm", "kernel": "ingest of ONE trajectory, %d frames per launch set (svo_submit_frames)" % K_,
                                              "achieved": ak, "peak": peak, "unit": "GB/s", "frac": ak / peak, "us_per_launch_set": usk,
                                              "us_per_frame": usk / K_, "algorithmic_bytes_per_launch_set": K_ * B_ALG_FRAME}
        roofline_batch = roofline_single   # kept under the old key for continuity: the 1-frame figure
        cpu = cpu_baseline(seq, Lh, Rh) if not args.no_cpu_baseline else None
        out = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "synthetic KITTI-shape stereo 1241x376 uint8, kitti00.yaml, ~2000 features, "
                                   "one sequence per GPU (BASELINE.json configs[1])",
                       "frames_per_step": FRAMES_PER_STEP, "unique_frames": n_unique, "playback": "ping-pong",
                       "l2": "inputs larger than L2: %d MB of distinct frames per GPU cycle through a 126 MB L2 (every frame is read once per pass; no flush needed)" % (2 * n_unique * W * H >> 20),
                       "features_bucketed_mean": mean_bucketed, "features_matched_mean": mean_matched,
                       "lk_iters_per_frame": mean_iters, "success_rate": success_rate,
                       "rel_translation_err_vs_gt_median": float(np.median(terr)) if terr else None},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
            "roofline_single_frame": roofline_batch, "roofline_lookahead": roofline_lookahead, "stages": stages, "batch64": batch, "cpu_baseline": cpu,
            "render_s": render_s,
        }
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=1000, help="unique rendered frames of the N = 1 sequence (configs[1]: 1000)")
    ap.add_argument("--frames5", type=int, default=24, help="unique rendered frames per sequence of the 64-sequence batch")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif world > 1:
        run_config5(args, rank, world, local_rank)
    else:
        run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
