#!/usr/bin/env python
"""bench.py — headline benchmark of the per-frame 2x upscaler hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cnn|bicubic|stream|wide|rrdb] [--impl reference]

One JSON line on stdout (rank 0).  Metric (BASELINE.json): 2x-SR frames/sec at 1080p->2160p.

* ``value``   device-resident throughput: inputs already in HBM, CUDA events on the launching stream,
              barrier + synchronize on both sides, max over ranks; frames of ALL ranks / that time.
* ``e2e``     same metric through the public host-buffer call (``vex_upscale_host`` behind
              ``B200VideoEnhancer``): pinned host frames -> H2D -> kernels -> D2H, copies inside the timed region.
* ``roofline``  dominant kernel against MEASURED_PEAKS.json (HBM GB/s for bicubic, bf16 TFLOP/s for the CNN).
* ``cpu_baseline``  the oracle (the reference's own torch-CPU code path) on this box's host cores, bounded sample.
* ``--impl reference``  times only that CPU path, on all host threads (rank 0 only).

A step = one pass of the hot path over one batch of synthetic frames (``config.frames_per_step`` per GPU).
Inputs are larger than L2 (126 MB) for both workloads, so no explicit L2 flush is needed (``config.l2``).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W = 1080, 1920
BICUBIC_BYTES_PER_FRAME = H * W * 3 * 5                 # read HW3 + write 4HW3
CNN_FLOP_PER_PX = 1_508_352
DEFAULT_WORKLOAD = "cnn"                                # the north-star metric; bicubic rides along as a secondary object
METRIC = "2x-SR frames/sec at 1080p->2160p"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


PROFILE_CNN_CSV = ("profiles/r2_ncu_full_cnn.csv", "profiles/r1_v7_ncu_full_cnn.csv")        # newest first
PROFILE_BIC_CSV = ("profiles/r2_ncu_full_bicubic.csv", "profiles/r1_v7_ncu_full_bicubic.csv")
PROFILE_WIDE_CSV = ("profiles/r2_ncu_full_wide.csv",)          # conv3x3_tc launches of ONE 2160p frame of the configs[4] plan
_UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def ncu_traffic(candidates, kernel_regex):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (mean over the matching launches) from the newest committed
    `ncu --set full` summary (profiles/*.csv, written by tools/ncu_summary.sh from the .ncu-rep of HEAD).  Read at run
    time so the number follows the capture, not a literal typed into this file.  Returns (bytes, source, n_launches,
    per-kernel dict) or (None, None, 0, {})."""
    import csv
    import re
    for rel in candidates:
        path = os.path.join(ROOT, rel)
        if not os.path.exists(path):
            continue
        rows = list(csv.reader(open(path)))
        if len(rows) < 3:
            continue
        head, units = rows[0], rows[1]
        try:
            ki, ri, wi = head.index("Kernel Name"), head.index("dram__bytes_read.sum"), head.index("dram__bytes_write.sum")
        except ValueError:
            continue
        per, vals = {}, []
        for r in rows[2:]:
            if len(r) <= max(ki, ri, wi) or not re.search(kernel_regex, r[ki]):
                continue
            b = float(r[ri]) * _UNIT.get(units[ri], 1.0) + float(r[wi]) * _UNIT.get(units[wi], 1.0)
            vals.append(b)
            per.setdefault(r[ki].split("(")[0].replace("void ", ""), []).append(b)
        if vals:
            return float(np.mean(vals)), rel, len(vals), {k: float(np.mean(v)) for k, v in per.items()}
    return None, None, 0, {}


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period=0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "sw_power_cap": 0x4, "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10,
                 "applications_clocks_setting": 0x2}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# --------------------------------------------------------------------------------------- CPU reference arm

def cpu_reference(workload: str, budget_s: float):
    """The reference's own CPU implementation of the path (oracle = app.py code on torch CPU), all host threads.
    Bounded sample: as many 1080p frames as fit in ~budget_s (at least 1)."""
    import torch
    from oracle import upscaler_ref as O
    from video_enhancer_proto_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    frames = synth.mixed_clip(3, H, W, seed=1234)
    if workload == "bicubic":
        fn = O.ref_bicubic_frame
    else:
        model = O.make_model(0)
        fn = lambda f: O.ref_cnn_frame(model, f)
    n, t0 = 0, time.perf_counter()
    if workload == "bicubic":
        fn(frames[0])                                     # warm the thread pool (cheap for bicubic only)
        t0 = time.perf_counter()
    while True:
        fn(frames[n % len(frames)])
        n += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or n >= 300:
            break
    return {"value": n / dt, "unit": "frames/s", "cores": cores, "kind": "port",
            "sample": f"{n} frame(s) 1080p->2160p, {workload}, torch {torch.__version__} CPU fp32, {torch.get_num_threads()} threads, {dt:.1f}s"}


def run_reference_rrdb(args):
    """--impl reference --workload rrdb: the oracle's Real-ESRGAN (the reference's network on torch CPU, all host threads) on
    padded tiles of a 1080p frame, one tile per step, scaled to frames by pixel count."""
    import torch
    from oracle import realesrgan_ref as R
    from video_enhancer_proto_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = R.make_network(0, 4, 23)
    frame = synth.random_frame(H, W, np.random.default_rng(1234))
    plan = R.tile_plan(H, W, 512, 10)
    px = sum((py1 - py0) * (px1 - px0) for (py0, py1, px0, px1, *_r) in plan)
    vals, tpx_total, t_total = [], 0, 0.0
    for k in range(max(1, min(args.steps, 3))):
        py0, py1, px0, px1 = plan[(5 + k) % len(plan)][:4]
        tile = np.ascontiguousarray(frame[py0:py1, px0:px1])
        t0 = time.perf_counter()
        R.process_single(model, tile)
        dt = time.perf_counter() - t0
        tpx = (py1 - py0) * (px1 - px0)
        vals.append(1.0 / (dt * px / tpx)); tpx_total += tpx; t_total += dt
    v = float(np.mean(vals))
    base = {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
            "sample": f"{len(vals)} padded tile(s) of the 12 of a 1080p frame ({tpx_total / px:.2f} frames' worth of pixels) in {t_total:.1f} s, "
                      f"torch {torch.__version__} CPU fp32, {torch.get_num_threads()} threads, scaled by pixel count"}
    print(json.dumps({"impl": "reference", "metric": "Real-ESRGAN x4 frames/sec at 1080p->4320p (reference tile loop)", "value": v,
                      "unit": "frames/s", "n_gpus": args.gpus, "steps": len(vals), "warmup": 0, "ms_per_step": 1000.0 / v,
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": "Real-ESRGAN RRDBNet 23 blocks x4, 1080p->4320p, tile 512 pad 10 (SURVEY 8(f)4)",
                                 "note": "reference CPU path (oracle port), one padded tile per step"},
                      "cpu_baseline": base,
                      "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


def run_reference(args, workload):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if workload == "rrdb":
        return run_reference_rrdb(args)
    if workload not in ("cnn", "bicubic"):                  # stream / wide: the reference arm is the CNN at configs[2]
        workload = "cnn"
    per_step = 4.0 if workload == "bicubic" else 30.0
    vals, base = [], None
    for _ in range(max(1, min(args.steps, 3))):
        base = cpu_reference(workload, per_step)
        vals.append(base["value"])
    v = float(np.mean(vals))
    base["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": len(vals), "warmup": 0, "ms_per_step": 1000.0 / v, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(workload), "note": "reference CPU path (oracle port), bounded sample per step"},
            "cpu_baseline": base,
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_name(w):
    return {"bicubic": "bicubic fallback 2x 1080p->2160p uint8 (BASELINE configs[1])",
            "cnn": "AdvancedUpscaler 2x CNN bf16 1080p->2160p, batch 8 frames/GPU (BASELINE configs[2]/[3])"}[w]


# --------------------------------------------------------------------------------------- GPU arm

def gpu_workload(workload, enh, args, world, rank, local, warmup, peaks, torch, dist, synth, _lib):
    """One workload on this rank's GPU: device-resident steps (value), host-buffer steps (e2e), roofline."""
    fps_step = args.frames or (300 if workload == "bicubic" else 8)
    # synthetic pool: 12 distinct frames (random / demo pattern / edges, seed 1234+rank), tiled to the batch
    pool = synth.mixed_clip(12, H, W, seed=1234 + rank)
    dev_pool = torch.from_numpy(np.stack(pool)).cuda()
    idx = torch.arange(fps_step, device="cuda") % len(pool)
    dev_in = dev_pool[idx].contiguous()                     # [F,1080,1920,3] resident in HBM
    dev_out = torch.empty((fps_step, 2 * H, 2 * W, 3), dtype=torch.uint8, device="cuda")
    del dev_pool
    mode = "bicubic" if workload == "bicubic" else "cnn"
    steps = args.steps

    def step():
        enh.upscale_device(dev_in, mode=mode, out=dev_out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    if mode == "cnn":
        enh.set_option("profile", 1)                        # per-layer events on the launching stream
    l0 = enh.kernel_launches()
    sampler = ClockSampler(local)
    sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    layer_ms = []
    barrier()
    ev[0].record()
    for i in range(steps):
        step()
        ev[i + 1].record()
        if mode == "cnn":
            layer_ms.append(None)                           # read back after the timed region
    barrier()
    clocks = sampler.stop()
    launches = enh.kernel_launches() - l0
    total_ms = ev[0].elapsed_time(ev[-1])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    prof = prof_stats = None
    if mode == "cnn":
        prof_stats = enh.cnn_profile_stats()                # every timed step (the warm-up ran with profile off)
        prof = prof_stats["mean"]
        enh.set_option("profile", 0)
    t = torch.tensor([total_ms], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * fps_step * steps / (total_ms / 1e3)

    # ---- end to end: pinned host pool -> vex_upscale_host -> pinned host output
    e2e_frames = min(fps_step, 96) if workload == "bicubic" else max(fps_step, 48)   # CNN: 6 batches of 8 per streaming call
    fin = H * W * 3
    pin_in = enh._pinned_buf("in", e2e_frames * fin)
    pin_out = enh._pinned_buf("out", e2e_frames * fin * 4)
    vin = pin_in.numpy()[: e2e_frames * fin].reshape(e2e_frames, H, W, 3)
    for i in range(e2e_frames):
        vin[i] = pool[i % len(pool)]
    mode_id = _lib.VEX_MODE_BICUBIC if workload == "bicubic" else _lib.VEX_MODE_CNN
    lib = _lib.load()

    def e2e_step():
        rc = lib.vex_upscale_host(enh._ctx, mode_id, pin_in.data_ptr(), pin_out.data_ptr(), e2e_frames, H, W, 8, 0)
        _lib.check(enh._ctx, rc, "vex_upscale_host")

    for _ in range(2):
        e2e_step()
    barrier()
    e2e_steps = max(3, min(steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()                                           # returns when every output frame is on the host
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = world * e2e_frames * e2e_steps / float(dt.item())

    # ---- the same frames as a continuous stream (what process_video does): vex_submit / vex_wait, batches of 8 frames from
    # rotating pinned buffers, two batches in flight, so the copies of one batch overlap the kernels of its neighbours ACROSS
    # calls and nothing drains between them.  Every batch is copied H2D and its result D2H inside the timed region.
    import ctypes as C
    nbuf, bsz = 3, 8
    sin = [enh._pinned_buf(f"sq_in{i}", bsz * fin) for i in range(nbuf)]
    sout = [enh._pinned_buf(f"sq_out{i}", bsz * fin * 4) for i in range(nbuf)]
    for i in range(nbuf):
        v = sin[i].numpy()[: bsz * fin].reshape(bsz, H, W, 3)
        for j in range(bsz):
            v[j] = pool[(i * bsz + j) % len(pool)]

    def stream_batches(nb):
        tickets = []
        for i in range(nb):
            if i >= 2:
                _lib.check(enh._ctx, lib.vex_wait(enh._ctx, tickets[i - 2]), "vex_wait")    # frees buffer (i - 2) % 3 = (i + 1) % 3
            t = C.c_ulonglong()
            rc = lib.vex_submit(enh._ctx, mode_id, sin[i % nbuf].data_ptr(), sout[i % nbuf].data_ptr(), bsz, H, W, 0, C.byref(t))
            _lib.check(enh._ctx, rc, "vex_submit")
            tickets.append(t.value)
        for t in tickets[-2:]:
            _lib.check(enh._ctx, lib.vex_wait(enh._ctx, t), "vex_wait")
    stream_batches(4)
    barrier()
    nb = e2e_frames * e2e_steps // bsz
    t0 = time.perf_counter()
    stream_batches(nb)                                       # returns when the last batch's output is on the host
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_stream_value = world * nb * bsz / float(dt.item())

    if workload == "bicubic":
        kern_ms = float(np.mean(step_ms))                    # the step IS one launch of bicubic2x_strip_kernel
        achieved = BICUBIC_BYTES_PER_FRAME * fps_step / (kern_ms / 1e3) / 1e9
        # dram bytes of one launch from the committed ncu --set full capture (300 frames), scaled to this launch's frames
        tr, tr_src, _, _ = ncu_traffic(PROFILE_BIC_CSV, r"bicubic2x_strip")
        roof = {"bound": "hbm", "kernel": "bicubic2x_strip_kernel", "achieved": achieved, "peak": peaks["hbm_gbs"],
                "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                "traffic": int(tr * fps_step / 300) if tr else None, "traffic_source": tr_src,
                "peak_source": peaks["source"],
                "algorithmic_bytes_per_launch": BICUBIC_BYTES_PER_FRAME * fps_step, "launch_ms": kern_ms,
                "launch_ms_min_max": [float(np.min(step_ms)), float(np.max(step_ms))]}
        dtype = "u8"
    else:
        # dominant kernel: conv3x3_tc_kernel (conv2+conv3+conv4 = 3 launches per step, 97.8 % of the FLOPs).  Its time is
        # the MEAN over every timed step of the per-layer CUDA events on the launching stream (vex_cnn_profile_stats).
        px = H * W * fps_step
        tc_flop = 2 * 9 * (128 * 128 + 128 * 256 + 256 * 128) * px
        tc_ms = prof["conv2"] + prof["conv3"] + prof["conv4"]
        achieved = tc_flop / (tc_ms / 1e3) / 1e12
        timed_s = total_ms / 1e3
        # MEASURED_PEAKS.json: the burst figure is the peak for a kernel timed alone or in a short window, the sustained one
        # for a kernel inside a long (seconds) power-capped step.  The window decides which one `frac` is quoted against;
        # both fractions are always reported.
        sustained = timed_s >= 2.0
        peak = peaks["bf16_tflops_sustained"] if sustained else peaks["bf16_tflops"]
        mean_step = float(np.mean(step_ms))
        whole = CNN_FLOP_PER_PX * px / (mean_step / 1e3) / 1e12
        tr, tr_src, tr_n, tr_per = ncu_traffic(PROFILE_CNN_CSV, r"conv3x3_tc_kernel")
        hbm = peaks["hbm_gbs"]

        def thin(ms, bytes_per_px):
            gbs = bytes_per_px * px / (ms / 1e3) / 1e9 if ms > 0 else 0.0
            return {"ms": ms, "algorithmic_bytes": bytes_per_px * px, "achieved_gbs": gbs, "frac_of_hbm": gbs / hbm}
        roof = {"bound": "tensor", "kernel": "conv3x3_tc_kernel (conv2+conv3+conv4, 3 launches/step)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                # DRAM bytes per launch (mean of the conv3x3_tc launches of one 8-frame step in the ncu --set full capture),
                # scaled to this step's frame count
                "traffic": int(tr * fps_step / 8) if tr else None, "traffic_source": tr_src, "traffic_launches": tr_n,
                "peak_source": peaks["source"] + (" (sustained cuBLAS bf16: timed region %.1f s >= 2 s of back-to-back steps)" % timed_s
                                                   if sustained else " (burst cuBLAS bf16: timed region %.2f s < 2 s)" % timed_s),
                "frac_of_burst_peak": achieved / peaks["bf16_tflops"],
                "frac_of_sustained_peak": achieved / peaks["bf16_tflops_sustained"],
                "algorithmic_flop_per_step": tc_flop, "kernel_ms_per_step": tc_ms,
                "kernel_share_of_step": tc_ms / mean_step,
                "layer_ms": prof, "layer_ms_min": prof_stats["min"], "layer_ms_max": prof_stats["max"],
                "layer_ms_steps": prof_stats["n"],
                "whole_net_tflops": whole, "whole_net_frac": whole / peak,
                "whole_net_frac_of_burst_peak": whole / peaks["bf16_tflops"],
                "whole_net_frac_of_sustained_peak": whole / peaks["bf16_tflops_sustained"],
                # the thin (HBM-bound) layers against the measured HBM bandwidth: algorithmic bytes per input pixel =
                # conv1 3 (u8) + 256 (x1 bf16); conv5 256 (x4) + 16 (x5 fp32x4); fused tail 16 (x5) + 12 (u8 out)
                "thin": {"conv1": thin(prof["conv1"], 259), "conv5": thin(prof["conv5"], 272),
                         "tail": thin(prof["upconv1"] + prof["upconv2"], 28),
                         "ms_per_step": prof["conv1"] + prof["conv5"] + prof["upconv1"] + prof["upconv2"]}}
        dtype = "bf16"
    return {"value": value, "ms_per_step": total_ms / steps, "fps_step": fps_step, "clocks": clocks,
            "launches": int(launches), "roofline": roof, "dtype": dtype,
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": e2e_frames * fin,
                    "d2h_bytes_per_step": e2e_frames * fin * 4, "frames_per_step": e2e_frames, "steps": e2e_steps,
                    "streaming": {"value": e2e_stream_value, "unit": "frames/s", "batches": nb, "frames_per_batch": bsz,
                                  "h2d_bytes_per_batch": bsz * fin, "d2h_bytes_per_batch": bsz * fin * 4,
                                  "api": "vex_submit / vex_wait: pinned host batches of 8 frames, two in flight (the loop of process_video)"},
                    "api": "vex_upscale_host (pinned host in/out, H2D/compute/D2H streams, 2 device slots, chunks 1, 2, 4, 8, ..., 8, 1 frames; the call returns when every output frame is on the host)"}}


def gpu_plugin_e2e(enh, world, rank, steps, torch, dist, synth):
    """`e2e_plugin`: the same metric through the METHOD the reference's video loop calls (app.py:463,475):
    `_process_frame_batch_gpu(frames_batch)` - a Python list of 8 separately allocated, pageable RGB ndarrays in, a list of
    8 fresh BGR ndarrays out.  Host staging, H2D and D2H are inside the timed region (wall clock around the calls)."""
    pool = synth.mixed_clip(8, H, W, seed=4321 + rank)
    frames = [f.copy() for f in pool]                          # plain numpy (pageable) arrays, one allocation each
    out = None
    for _ in range(4):                                         # the previous list stays alive until the next one is bound, like
        out = enh._process_frame_batch_gpu(frames)             # `enhanced_batch = ...` in the reference's loop (app.py:463): two
    if world > 1:                                              # pinned result blocks alternate; both exist after the warm-up
        dist.barrier()
    torch.cuda.synchronize()
    n_calls = max(5, min(steps, 20))
    t0 = time.perf_counter()
    for _ in range(n_calls):
        out = enh._process_frame_batch_gpu(frames)             # returns when all 8 outputs are on the host
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    assert len(out) == 8 and out[0].shape == (2 * H, 2 * W, 3)
    fin = H * W * 3
    return {"value": world * 8 * n_calls / float(dt.item()), "unit": "frames/s", "calls": n_calls, "frames_per_call": 8,
            "h2d_bytes_per_step": 8 * fin, "d2h_bytes_per_step": 8 * fin * 4,
            "api": "B200VideoEnhancer._process_frame_batch_gpu(list of 8 pageable ndarrays) -> list of 8 ndarrays "
                   "(vex_upscale_frames: chunks 1,2,4,1 staged into pinned memory while the previous chunk computes; "
                   "outputs land in a fresh pinned tensor, no host copy out)"}


def gpu_file_e2e(enh, n_frames, torch):
    """`e2e_file`: `process_video` (app.py:388-519) file to file on a synthetic 1080p mp4v clip written in this run, with
    the decode-only and encode-only rates of the same OpenCV codecs beside it, so the bound is named."""
    import shutil
    import tempfile
    import cv2
    from video_enhancer_proto_b200 import synth
    tmp = tempfile.mkdtemp(prefix="vex_bench_")
    try:
        src, dst, enc = os.path.join(tmp, "in.mp4"), os.path.join(tmp, "out.mp4"), os.path.join(tmp, "enc.mp4")
        wr = cv2.VideoWriter(src, cv2.VideoWriter_fourcc(*"mp4v"), 30.0, (W, H))
        if not wr.isOpened():
            return {"unavailable": "cv2.VideoWriter cannot open an mp4v file on this box"}
        rng = np.random.default_rng(7)
        base = [synth.demo_frame(i, size=(W, H), rng=rng, bgr=True) for i in range(10)]
        for i in range(n_frames):
            wr.write(base[i % 10])
        wr.release()
        # decode only
        cap = cv2.VideoCapture(src)
        buf = np.empty((H, W, 3), np.uint8)
        t0, n = time.perf_counter(), 0
        while True:
            ok, _ = cap.read(buf)
            if not ok:
                break
            n += 1
        dec_fps = n / (time.perf_counter() - t0)
        cap.release()
        # encode only: the frames this very call will produce (the enhancer's output for the first 10 frames), bounded sample
        outs = enh._run_host(1, base, 0b11).copy()
        wr = cv2.VideoWriter(enc, cv2.VideoWriter_fourcc(*"mp4v"), 30.0, (2 * W, 2 * H))
        n_enc = min(n_frames, 60)
        t0 = time.perf_counter()
        for i in range(n_enc):
            wr.write(outs[i % 10])
        wr.release()
        enc_fps = n_enc / (time.perf_counter() - t0)
        ok, msg = enh.process_video(src, dst)                  # warm-up of buffers is part of a real call: time the second
        t0 = time.perf_counter()
        ok, msg = enh.process_video(src, dst)
        dt = time.perf_counter() - t0
        if not ok:
            return {"unavailable": msg}
        fps = n / dt
        bound = min((("decode", dec_fps), ("encode", enc_fps)), key=lambda kv: kv[1])
        return {"value": fps, "unit": "frames/s", "frames": n, "seconds": dt,
                "decode_only_fps": dec_fps, "encode_only_fps": enc_fps, "encode_sample_frames": n_enc,
                "bound": f"{bound[0]} ({bound[1]:.0f} fps alone; OpenCV mp4v on the host CPU, one thread each side of the GPU)",
                "fraction_of_bound": fps / bound[1],
                "api": "B200VideoEnhancer.process_video(in.mp4, out.mp4): cv2 decode thread -> pinned batch -> vex_submit/"
                       "vex_wait (2 batches in flight) -> pinned batch -> cv2 encode thread"}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


DEMO_SHAPES = ((480, 640), (360, 480))     # (h, w): utils/create_demo_video.py:159 / .github/workflows/ci.yml:97, app.py:1487


def gpu_demo480(enh, cpu_too, torch):
    """BASELINE.json configs[0]: the 30-frame 480p demo clip (both demo shapes) through the CNN, host buffers in and out,
    with the reference CPU path (oracle) timed beside it on a bounded sample of the same frames (N=1 only)."""
    from video_enhancer_proto_b200 import synth
    out = {}
    for (h, w) in DEMO_SHAPES:
        rng = np.random.default_rng(0)
        frames = [synth.demo_frame(i, size=(w, h), rng=rng, bgr=False) for i in range(30)]
        fin = h * w * 3
        pin_in = enh._pinned_buf("demo_in", 30 * fin)
        pin_out = enh._pinned_buf("demo_out", 30 * fin * 4)
        pin_in.numpy()[: 30 * fin] = np.stack(frames).ravel()
        lib = enh._lib

        def call():
            rc = lib.vex_upscale_host(enh._ctx, 1, pin_in.data_ptr(), pin_out.data_ptr(), 30, h, w, 8, 0)
            if rc:
                raise RuntimeError(lib.vex_last_error(enh._ctx))
        call()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            call()
        dt = (time.perf_counter() - t0) / reps
        rec = {"gpu_e2e_fps": 30 / dt, "clip_ms": dt * 1e3, "frames": 30}
        if cpu_too:
            from oracle import upscaler_ref as O
            model = O.make_model(0)
            torch.set_num_threads(os.cpu_count() or 1)
            O.ref_cnn_frame(model, frames[0])
            t0, k = time.perf_counter(), 0
            while k < 30 and (k < 3 or time.perf_counter() - t0 < 6.0):
                O.ref_cnn_frame(model, frames[k])
                k += 1
            rec.update(cpu_reference_fps=k / (time.perf_counter() - t0), cpu_sample_frames=k,
                       cpu_cores=os.cpu_count() or 1)
        out[f"{w}x{h}"] = rec
    return {"workload": "AdvancedUpscaler 2x CNN on the 30-frame 480p demo clip shapes (BASELINE configs[0])",
            "api": "vex_upscale_host, pinned host buffers, copies inside", **out}


def gpu_stream(enh, args, world, rank, torch, dist, synth, _lib):
    """BASELINE.json configs[3]: stream N synthetic frames (default 3000) through the CNN, frame-sharded as contiguous
    ranges over the ranks, host buffers in and out, outputs re-ordered on the host.  Strong scaling: the total is fixed.
    A feeder thread (the decoder's stand-in) copies the next 48 frames from a 64-frame pool into the second pinned input
    buffer while the GPU works on the current one.  Order check: every output frame's sparse checksum must equal the
    checksum of the pool frame that was fed at that stream position, gathered to rank 0 in frame order."""
    from video_enhancer_proto_b200 import shard
    total = args.frames or 3000
    lo, hi = shard.frame_range(rank, world, total)
    pool_n, call_n = 64, 48
    pool = np.stack(synth.mixed_clip(pool_n, H, W, seed=1234))
    fin = H * W * 3
    pins = [enh._pinned_buf(f"s_in{i}", call_n * fin) for i in range(2)]
    pout = enh._pinned_buf("s_out", call_n * fin * 4)
    vins = [p.numpy()[: call_n * fin].reshape(call_n, H, W, 3) for p in pins]
    vout = pout.numpy()[: call_n * fin * 4].reshape(call_n, 2 * H, 2 * W, 3)
    lib = enh._lib
    enh._ensure_weights()

    def run(buf, n):
        rc = lib.vex_upscale_host(enh._ctx, _lib.VEX_MODE_CNN, pins[buf].data_ptr(), pout.data_ptr(), n, H, W, 8, 0)
        _lib.check(enh._ctx, rc, "vex_upscale_host")

    def csum(frame):
        return int(frame[::97, ::89].astype(np.uint32).sum())

    # expected checksum per pool frame (also the warm-up)
    want = []
    for p0 in range(0, pool_n, call_n):
        n = min(call_n, pool_n - p0)
        vins[0][:n] = pool[p0:p0 + n]
        run(0, n)
        want += [csum(vout[i]) for i in range(n)]

    def fill(buf, f0, n):
        for i in range(n):
            vins[buf][i] = pool[(f0 + i) % pool_n]

    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    got = []
    calls = [(f0, min(call_n, hi - f0)) for f0 in range(lo, hi, call_n)]
    t0 = time.perf_counter()
    if calls:
        fill(0, *calls[0])
    for k, (f0, n) in enumerate(calls):
        feeder = None
        if k + 1 < len(calls):
            feeder = threading.Thread(target=fill, args=((k + 1) & 1, *calls[k + 1]))
            feeder.start()
        run(k & 1, n)
        got += [csum(vout[i]) for i in range(n)]
        if feeder:
            feeder.join()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        merged = shard.gather_to_rank0(got, total)
    else:
        merged = got
    ok = None
    if rank == 0:
        ok = len(merged) == total and all(merged[i] == want[i % pool_n] for i in range(total))
    return {"value": total / float(dt.item()), "seconds": float(dt.item()), "frames": total, "order_ok": ok,
            "frames_this_rank": hi - lo}


WIDE_WIDTHS, WIDE_EXTRA, WIDE_H, WIDE_W = (256, 256, 512, 256), 2, 2160, 3840


def gpu_wide(args, world, rank, local, warmup, peaks, torch, dist, synth, _lib):
    """BASELINE.json configs[4]: the widened + deepened random-init variant (3-256-256(+skip)-512-512-512-256-3, same
    head) at 2160p -> 4320p, frames sharded over the ranks (weak scaling: fixed frames per GPU per step).  Same
    measurements as the default workload: device-resident steps timed with CUDA events, the tensor-core layers'
    live per-layer times against the sustained bf16 peak, and host-buffer steps through vex_upscale_host."""
    from video_enhancer_proto_b200 import B200VideoEnhancer, weights as Wt
    Hh, Ww = WIDE_H, WIDE_W
    enh = B200VideoEnhancer(device=f"cuda:{local}", batch_size=2,
                            state_dict=Wt.init_state(0, widths=WIDE_WIDTHS, n_extra=WIDE_EXTRA))
    enh._ensure_ctx()
    enh._ensure_weights()
    fps_step = args.frames or 2
    pool = synth.mixed_clip(3, Hh, Ww, seed=1234 + rank)
    dev_in = torch.from_numpy(np.stack([pool[i % 3] for i in range(fps_step)])).cuda()
    dev_out = torch.empty((fps_step, 2 * Hh, 2 * Ww, 3), dtype=torch.uint8, device="cuda")
    steps = args.steps

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        enh.upscale_device(dev_in, mode="cnn", out=dev_out)
    barrier()
    enh.set_option("profile", 1)
    l0 = enh.kernel_launches()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        enh.upscale_device(dev_in, mode="cnn", out=dev_out)
    e1.record()
    barrier()
    clocks = sampler.stop()
    launches = enh.kernel_launches() - l0
    prof = enh.cnn_profile_ms()
    enh.set_option("profile", 0)
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * fps_step * steps / (total_ms / 1e3)

    c1, c2, c3, c4 = WIDE_WIDTHS
    tc_macs = c1 * c2 + c2 * c3 + WIDE_EXTRA * c3 * c3 + c3 * c4
    tc_flop = 2 * 9 * tc_macs * Hh * Ww * fps_step
    all_flop = (2 * 9 * (3 * c1 + tc_macs + c4 * 3) + 2 * (3072 + 6912)) * Hh * Ww * fps_step
    tc_ms = prof["conv2"] + prof["conv3"] + prof["conv4"]
    achieved = tc_flop / (tc_ms / 1e3) / 1e12
    peak = peaks["bf16_tflops_sustained"]

    fin = Hh * Ww * 3
    pin_in = enh._pinned_buf("in", fps_step * fin)
    pin_out = enh._pinned_buf("out", fps_step * fin * 4)
    vin = pin_in.numpy()[: fps_step * fin].reshape(fps_step, Hh, Ww, 3)
    for i in range(fps_step):
        vin[i] = pool[i % 3]
    lib = _lib.load()

    def e2e_step():
        rc = lib.vex_upscale_host(enh._ctx, _lib.VEX_MODE_CNN, pin_in.data_ptr(), pin_out.data_ptr(), fps_step, Hh, Ww, 1, 0)
        _lib.check(enh._ctx, rc, "vex_upscale_host")

    e2e_step()
    barrier()
    e2e_steps = max(3, min(steps, 5))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = world * fps_step * e2e_steps / float(dt.item())
    enh.close()
    if rank != 0:
        return
    tr, wide_src, wide_n, _ = ncu_traffic(PROFILE_WIDE_CSV, r"conv3x3_tc_kernel")   # mean DRAM bytes per conv3x3_tc launch, one frame
    wide_traffic = int(tr) if tr else None
    print(json.dumps({
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": total_ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "widened+deepened CNN variant 2x, 2160p->4320p (BASELINE configs[4])",
                   "plan": {"widths": list(WIDE_WIDTHS), "extra_blocks_c3": WIDE_EXTRA, "up_ch": 64,
                            "layers": "conv3x3 3-256-256(+skip)-512-512-512-256-3, ConvT k4s2 3-64, ConvT k3 64-3"},
                   "frames_per_step_per_gpu": fps_step, "height": Hh, "width": Ww,
                   "weights": "random-init (reference init procedure, seed 0)",
                   "flop_per_frame": all_flop / fps_step,
                   "l2": "25 GB of activations per frame > 126 MB L2, no flush",
                   "sharding": "contiguous frame ranges per GPU, no collective"},
        "clocks": clocks, "gpu_launches": int(launches),
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": fps_step * fin,
                "d2h_bytes_per_step": fps_step * fin * 4, "frames_per_step": fps_step, "steps": e2e_steps,
                "api": "vex_upscale_host (pinned host in/out, one frame per chunk)"},
        "roofline": {"bound": "tensor", "kernel": "conv3x3_tc_kernel (conv2, conv3 + 2 extra blocks, conv4: 5 launches/frame)",
                     "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                     "traffic": wide_traffic, "traffic_source": wide_src, "traffic_launches": wide_n,
                     "frac_of_burst_peak": achieved / peaks["bf16_tflops"],
                     "peak_source": peaks["source"] + " (sustained cuBLAS bf16)",
                     "algorithmic_flop_per_step": tc_flop, "kernel_ms_per_step": tc_ms,
                     "kernel_share_of_step": tc_ms / (total_ms / steps), "layer_ms": prof,
                     "whole_net_tflops": all_flop / (total_ms / steps / 1e3) / 1e12}}), flush=True)


def gpu_rrdb(args, world, rank, local, warmup, peaks, torch, dist, synth):
    """Real-ESRGAN RRDBNet (23 blocks, x4, seeded random weights - the handler's own fallback when no
    checkpoint is found, realesrgan_handler.py:153-156) on 1080p frames -> 4320p, each frame through the reference's tile
    loop (tile 512, pad 10: 12 tiles, :324-372), frames sharded over the ranks.  value: frames resident in HBM; e2e: the
    handler method `_enhance_frame` (host ndarray in, host ndarray out).  The CPU baseline is the oracle on ONE padded tile,
    scaled to a frame by pixel count (stated in `sample`)."""
    from oracle import realesrgan_ref as R
    from video_enhancer_proto_b200 import B200RealESRGAN
    hd = B200RealESRGAN(device=f"cuda:{local}", scale=4, seed=0, num_block=23)
    hd._ensure_model_loaded()
    fps_step = args.frames or 2
    steps = max(2, min(args.steps, 5))
    rng = np.random.default_rng(1234 + rank)
    frames = [synth.random_frame(H, W, rng) if i % 2 == 0 else synth.edges_frame(H, W, rng) for i in range(fps_step)]
    dev = [torch.from_numpy(f).cuda() for f in frames]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        for d in dev:
            hd.frame_device(d, in_bgr=True, out_bgr=True)
    for _ in range(max(1, warmup // 3)):
        step()
    barrier()
    l0 = hd.kernel_launches()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    launches = hd.kernel_launches() - l0
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * fps_step * steps / (total_ms / 1e3)
    plan = R.tile_plan(H, W, 512, 10)
    px = sum((py1 - py0) * (px1 - px0) for (py0, py1, px0, px1, *_r) in plan)       # what the reference's loop computes on
    flop = R.flops_per_input_pixel(4, 23) * px
    achieved = flop * fps_step * steps / (total_ms / 1e3) / 1e12
    # warm-up keeps two results alive at once, as the loop below does: the second pinned output block is then already in
    # torch's host allocator cache and its first cudaHostAlloc (tens of ms for 100 MB) stays out of the timed region
    w1 = hd._enhance_frame(frames[0])
    w2 = hd._enhance_frame(frames[1 % fps_step])
    del w1, w2
    barrier()
    t0 = time.perf_counter()
    n_e2e = max(4, fps_step)
    for i in range(n_e2e):
        out = hd._enhance_frame(frames[i % fps_step])
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    assert out.shape == (4 * H, 4 * W, 3)
    line = {"metric": "Real-ESRGAN x4 frames/sec at 1080p->4320p (reference tile loop)", "value": value, "unit": "frames/s",
            "n_gpus": world, "steps": steps, "warmup": max(1, warmup // 3), "ms_per_step": total_ms / steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": {"workload": "Real-ESRGAN RRDBNet 23 blocks x4, 1080p->4320p, tile 512 pad 10 (SURVEY 8(f)4)",
                       "frames_per_step_per_gpu": fps_step, "tiles_per_frame": len(plan), "pixels_per_frame_incl_padding": px,
                       "weights": "random-init (torch default Conv2d init, seed 0)", "flop_per_frame": flop,
                       "l2": "20 GB of activations per frame (12 canvases) > 126 MB L2, no flush"},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": world * n_e2e / float(dt.item()), "unit": "frames/s", "h2d_bytes_per_step": H * W * 3,
                    "d2h_bytes_per_step": H * W * 3 * 16, "api": "B200RealESRGAN._enhance_frame(ndarray BGR) -> ndarray BGR"},
            "roofline": {"bound": "tensor", "kernel": "rconv3x3_kernel (418 of the 424 launches of a pass)", "achieved": achieved,
                         "peak": peaks["bf16_tflops"], "unit": "TFLOP/s", "frac": achieved / peaks["bf16_tflops"],
                         "traffic": None, "peak_source": peaks["source"] + " (burst cuBLAS bf16; fp16 has the same dense rate)",
                         "algorithmic_flop_per_step": flop * fps_step,
                         "note": "whole pass incl. the 6 thin launches; the 12 tiles of a frame run as one batch of canvases; wide layers wait for HBM, narrow ones for the epilogue (DESIGN.md 4)"}}
    if world == 1 and not args.no_cpu_baseline:
        model = R.make_network(0, 4, 23)
        torch.set_num_threads(os.cpu_count() or 1)
        py0, py1, px0, px1 = plan[5][:4]                                  # an interior tile: 532 x 532 with padding
        tile = np.ascontiguousarray(frames[0][py0:py1, px0:px1])
        t0 = time.perf_counter()
        R.process_single(model, tile)
        dtc = time.perf_counter() - t0
        tpx = (py1 - py0) * (px1 - px0)
        line["cpu_baseline"] = {"value": 1.0 / (dtc * px / tpx), "unit": "frames/s", "cores": os.cpu_count() or 1, "kind": "port",
                                "sample": f"one padded tile {py1 - py0}x{px1 - px0} of the 12 ({tpx / px:.3f} of a frame's pixels) "
                                          f"through the oracle in {dtc:.1f} s, torch CPU fp32, scaled by pixel count"}
    hd.close()
    if rank == 0:
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=["cnn", "bicubic", "stream", "wide", "rrdb"])
    ap.add_argument("--frames", type=int, default=0, help="frames per step per GPU (0 = workload default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary objects (bicubic, e2e_plugin, stream, demo480, e2e_file)")
    ap.add_argument("--no-file", action="store_true", help="skip e2e_file (process_video on an mp4v clip written here)")
    ap.add_argument("--stream-frames", type=int, default=960, help="total frames of the `stream` secondary object (configs[3])")
    ap.add_argument("--file-frames", type=int, default=300, help="frames of the e2e_file clip")
    args = ap.parse_args()
    workload = args.workload
    if args.impl == "reference":
        return run_reference(args, workload)

    import torch
    import torch.distributed as dist
    from video_enhancer_proto_b200 import B200VideoEnhancer, synth, _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        # stdout carries exactly one JSON line.  NCCL prints its "NCCL version ..." banner to stdout when the environment
        # sets NCCL_DEBUG=VERSION (this image does), so file descriptor 1 points at stderr while the communicator comes up.
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    warmup = max(args.warmup, 3)
    peaks = load_peaks()
    if workload == "rrdb":                                  # SURVEY 8(f)4; a reporting workload, not the driver's default
        gpu_rrdb(args, world, rank, local, warmup, peaks, torch, dist, synth)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    if workload == "wide":                                  # configs[4]; a reporting workload, not the driver's default
        gpu_wide(args, world, rank, local, warmup, peaks, torch, dist, synth, _lib)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    enh = B200VideoEnhancer(device=f"cuda:{local}", seed=0, batch_size=8)
    enh._ensure_ctx()
    if workload == "stream":                                # configs[3]; a reporting workload, not the driver's default
        sampler = ClockSampler(local); sampler.start()
        r = gpu_stream(enh, args, world, rank, torch, dist, synth, _lib)
        clocks = sampler.stop()
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": r["value"], "unit": "frames/s", "n_gpus": world, "steps": 1,
                              "warmup": 1, "ms_per_step": r["seconds"] * 1e3, "higher_is_better": True, "scaling": "strong",
                              "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "clocks": clocks,
                              "config": {"workload": f"AdvancedUpscaler 2x CNN 1080p->2160p streaming {r['frames']} frames, "
                                                     "frame-sharded contiguous ranges, host buffers in/out (BASELINE configs[3])",
                                         "order_check": r["order_ok"], "frames_rank0": r["frames_this_rank"]},
                              "e2e": {"value": r["value"], "unit": "frames/s", "h2d_bytes_per_step": r["frames"] * H * W * 3,
                                      "d2h_bytes_per_step": r["frames"] * H * W * 12}}), flush=True)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        enh.close()
        return
    r = gpu_workload(workload, enh, args, world, rank, local, warmup, peaks, torch, dist, synth, _lib)
    secondary = None
    if workload == "cnn" and not args.no_secondary:
        sargs = argparse.Namespace(**vars(args))
        sargs.frames, sargs.steps = 0, max(3, min(args.steps, 10))
        b = gpu_workload("bicubic", enh, sargs, world, rank, local, warmup, peaks, torch, dist, synth, _lib)
        secondary = {"workload": workload_name("bicubic"), "value": b["value"], "unit": "frames/s",
                     "ms_per_step": b["ms_per_step"], "frames_per_step_per_gpu": b["fps_step"], "dtype": "u8",
                     "roofline": b["roofline"], "e2e": b["e2e"], "gpu_launches": b["launches"]}
    extras = {}
    if workload == "cnn" and not args.no_secondary:
        extras["e2e_plugin"] = gpu_plugin_e2e(enh, world, rank, args.steps, torch, dist, synth)
        extras["e2e_plugin"]["vs_e2e"] = extras["e2e_plugin"]["value"] / r["e2e"]["value"]
        sargs = argparse.Namespace(**vars(args))
        sargs.frames = args.stream_frames                       # configs[3], fixed total: strong scaling over the ranks
        st = gpu_stream(enh, sargs, world, rank, torch, dist, synth, _lib)
        extras["stream"] = {"workload": "AdvancedUpscaler 2x CNN 1080p->2160p streaming, frame-sharded contiguous ranges, "
                                        "host buffers in/out (BASELINE configs[3])",
                            "value": st["value"], "unit": "frames/s", "frames_total": st["frames"], "scaling": "strong",
                            "seconds": st["seconds"], "order_check": st["order_ok"], "frames_rank0": st["frames_this_rank"]}
        if rank == 0:
            extras["demo480"] = gpu_demo480(enh, world == 1 and not args.no_cpu_baseline, torch)
            if world == 1 and not args.no_file:
                extras["e2e_file"] = gpu_file_e2e(enh, args.file_frames, torch)
        if world > 1:
            dist.barrier()
    fin = H * W * 3
    if rank == 0:
        line = {"metric": METRIC, "value": r["value"], "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": r["dtype"], "data": "synthetic",
                "config": {"workload": workload_name(workload), "frames_per_step_per_gpu": r["fps_step"], "height": H, "width": W,
                           "weights": "random-init (reference init procedure, seed 0)" if workload == "cnn" else None,
                           "l2": f"inputs per step {r['fps_step'] * fin / 1e6:.0f} MB + outputs {r['fps_step'] * fin * 4 / 1e6:.0f} MB"
                                 + (" + multi-GB activations" if workload == "cnn" else "") + " > 126 MB L2, no flush",
                           "sharding": "contiguous frame ranges per GPU, no collective"},
                "clocks": r["clocks"], "gpu_launches": r["launches"], "e2e": r["e2e"], "roofline": r["roofline"]}
        if secondary:
            line["bicubic"] = secondary
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference(workload, 6.0 if workload == "bicubic" else 25.0)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    enh.close()


if __name__ == "__main__":
    main()
