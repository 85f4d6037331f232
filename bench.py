#!/usr/bin/env python3
"""bench.py — converted frames/sec of the fused pre-encode stage on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A *step* is one pass of the hot path over one batch of synthetic input: S
independent 3840x2160 desktop sessions (BASELINE.json configs[1], the largest
single-GPU config the metric is quoted on) converted BGRx→NV12 with damage
maps by ONE kernel launch (sk_batch_submit).  Each session keeps two captured
frames resident in HBM and alternates between them, so every step reads
S*33.2 MB of BGRx + S*12.4 MB of previous output and writes S*12.4 MB: the
working set (≈1.9 GB at S=32) is far larger than the 126 MB L2.

Lines printed (one JSON object, rank 0 only):
  value     frames/s with inputs already in HBM (device-event timed, max over ranks)
  e2e       same metric through the C ABI with HOST buffers: pinned staging → H2D →
            kernel → D2H of the NV12 frame + map, every step, wall-clock between syncs
  roofline  algorithmic bytes per launch / CUDA-event kernel time vs MEASURED_PEAKS.json
  cpu_baseline  the C oracle (restatement of videoconvert, "port") on the host cores

`--impl reference` times that CPU path alone on the same workload (GStreamer
itself cannot be installed here).
Under torchrun (N>1) sessions are sharded S per GPU with no data-path
collective; torch.distributed (gloo) is used only for the barrier and the
max-over-ranks of the timing.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 3840, 2160
WORKLOAD = "3840x2160 BGRx -> NV12 fused convert + damage map on 1xB200"
METRIC = "converted frames/sec @4K per GPU (fused BGRx->NV12 + 16x16 damage map)"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([time.perf_counter()] + [c.strip() for c in line.split(",")])

    def stop(self, t_begin=None, t_end=None):
        """Summary of the samples; t_begin/t_end (perf_counter) mark the timed region inside the sampled span."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, inside = [], [], set(), 0
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if t_begin is not None and r[0] < t_begin:
                continue                       # before the timed region (warm-up)
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            if t_begin is not None and t_begin <= r[0] <= t_end + 0.06:
                inside += 1
            for n, v in zip(names, r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "samples_in_timed_region": inside, "reasons": sorted(reasons),
                "span": "timed region, plus the same step repeated untimed when the region is shorter than 0.4 s"}


# ----------------------------------------------------------------------------- CPU arm
def make_inputs(n_sessions: int, rank: int):
    """Two captured frames per session: desktop content, then 5 % of its macroblocks changed."""
    from selkies_gstreamer_b200 import synth
    base = synth.desktop(W, H, seed=42 + rank)
    out = []
    for i in range(n_sessions):
        a = np.roll(base, shift=(37 * i, 101 * i), axis=(0, 1))
        b = synth.mutate(a, 0.05, seed=1000 * rank + i, aligned=False)
        out.append((a, b))
    return out


def cpu_pass(frames, prevs, threads):
    """One step on the host: convert + damage for every session, sessions in parallel.

    Sessions are independent pipelines in the reference, each videoconvert with its own
    n-threads row bands (gstwebrtc_app.py:612), so the host baseline runs them concurrently:
    `workers` sessions at a time, each with threads/workers row-band threads (ctypes releases the GIL)."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import cpu_oracle as o
    workers = max(1, min(len(frames), threads // 4 if threads >= 8 else 1))
    per = max(1, threads // workers)

    def one(fp):
        f, p = fp
        cur = o.convert(f, out_fmt=o.OUT_NV12, matrix=o.MATRIX_BT709, chroma_site=o.SITE_H_COSITED, n_threads=per)
        o.damage(W, H, o.OUT_NV12, cur, p)
        return cur

    if workers == 1:
        return [one(fp) for fp in zip(frames, prevs)]
    with ThreadPoolExecutor(max_workers=workers) as ex:
        return list(ex.map(one, zip(frames, prevs)))


def cpu_baseline(budget_s: float = 12.0):
    """Bounded sample of the same workload on this box's host cores (C oracle, 'port')."""
    from oracle import cpu_oracle as o
    o.build()
    threads = len(os.sched_getaffinity(0))
    s = max(1, min(8, threads // 4))
    inputs = make_inputs(s, 0)
    prevs = cpu_pass([p[0] for p in inputs], [None] * s, threads)
    n, t0 = 0, time.perf_counter()
    while True:
        prevs = cpu_pass([p[(n + 1) & 1] for p in inputs], prevs, threads)
        n += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or n >= 2000:
            break
    return {"value": n * s / dt, "unit": "frames/s", "cores": threads, "kind": "port",
            "sample": f"{n * s} 4K frames ({s} sessions in parallel x {n} passes), convert+damage, C oracle, "
                      f"{threads} threads in total, {dt:.1f} s"}


def run_reference(args, rank, world):
    """--impl reference: the CPU path on the host cores, same workload/metric."""
    if rank != 0:
        return
    from oracle import cpu_oracle as o
    o.build()
    threads = len(os.sched_getaffinity(0))
    s = min(args.sessions, max(1, min(8, threads // 4)))        # bounded sample of the S-session batch
    inputs = make_inputs(s, 0)
    prevs = cpu_pass([p[0] for p in inputs], [None] * s, threads)
    for k in range(args.warmup):
        prevs = cpu_pass([p[(k + 1) & 1] for p in inputs], prevs, threads)
    t0 = time.perf_counter()
    for k in range(args.steps):
        prevs = cpu_pass([p[(k + args.warmup + 1) & 1] for p in inputs], prevs, threads)
    dt = time.perf_counter() - t0
    fps = s * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sessions_per_step": s,
                   "note": "CPU restatement of GStreamer videoconvert (+damage) — GStreamer itself is not installable here"},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": "port",
                         "sample": f"{s} 4K frames per step x {args.steps} steps, {threads} threads"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- GPU arm
def run_b200(args, rank, world, local_rank, dist):
    import selkies_gstreamer_b200 as sk
    from selkies_gstreamer_b200 import _abi, synth

    lib = _abi.load()
    if lib.sk_device_count() <= local_rank:
        raise SystemExit(f"rank {rank}: CUDA device {local_rank} not available — the product path has no fallback")
    if args.variant is not None:
        lib.sk_set_kernel_variant(args.variant)
    if args.no_snake:
        lib.sk_set_kernel_variant(102)
    S = args.sessions
    inputs = make_inputs(S, rank)

    # ---- device-resident arm: S sessions, one launch per step
    sess = [sk.PreencodeSession(W, H, out_format="NV12", colorimetry="bt709", chroma_site="mpeg2",
                                device_id=local_rank, n_inflight=2, host_output=False, timing=(i == 0))
            for i in range(S)]
    for s, (a, b) in zip(sess, inputs):
        s.upload(0, a); s.wait(s.submit(0))
        s.upload(1, b); s.wait(s.submit(1))
    batch = sk.SessionBatch(sess)
    sets = [batch.frame_set([s.device_input(k)[0] for s in sess]) for k in (0, 1)]
    stride = sess[0].device_input(0)[1]
    leader = sess[0]
    fid = 1000
    counts = []

    def step(k, wait_prev=True):
        nonlocal fid, counts
        batch.submit(sets[k & 1], stride, fid)
        if wait_prev and fid > 1000:
            counts = batch.wait(fid - 1)      # the host consumes every session's damage count each step
        fid += 1

    def sync_all():
        nonlocal counts
        counts = batch.wait(fid - 1)

    sampler = ClockSampler(local_rank)
    sampler.start()                            # nvidia-smi needs a moment to come up (longer on an 8-GPU box)
    for k in range(args.warmup):
        step(k)
    sync_all()
    if dist:
        dist.barrier()
    launches0 = sk.launch_count()
    kernel_us = []
    leader.marker(0)
    t0 = time.perf_counter()
    for k in range(args.steps):
        step(k + args.warmup)
        if k > 0:
            kernel_us.append(leader.last_kernel_us())
    leader.marker(1)
    sync_all()
    wall = time.perf_counter() - t0
    kernel_us.append(leader.last_kernel_us())
    dev_us = leader.marker_elapsed_us()
    launches = sk.launch_count() - launches0
    t_end = time.perf_counter()
    k = args.steps + args.warmup
    def samples_under_load():
        return sum(1 for r in sampler.rows if r[0] >= t0)

    # nvidia-smi samples every 50 ms: keep the same load up, untimed, until a few samples exist
    while (time.perf_counter() - t0 < 0.4 or samples_under_load() < 3) and time.perf_counter() - t0 < 5.0:
        step(k)
        k += 1
    sync_all()
    clocks = sampler.stop(t0, t_end)
    # last result sanity: ~5 % of blocks damaged, never 0 and never all
    r = sess[1].wait(fid - 1)
    nblk = r.damage_map.size
    assert 0 < r.damaged_count < nblk, (r.damaged_count, nblk)

    elapsed = max(dev_us * 1e-6, 1e-9)
    if dist:
        import torch
        t = torch.tensor([elapsed, wall], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed, wall = float(t[0]), float(t[1])
    frames_total = S * args.steps * world
    value = frames_total / elapsed

    # ---- end-to-end arm: host staging → H2D → kernel → D2H frame + map
    for s in sess[4:]:
        s.close()
    for s in sess[:4]:
        s.close()
    Se = min(args.e2e_sessions, S)
    esess = [sk.PreencodeSession(W, H, out_format="NV12", colorimetry="bt709", chroma_site="mpeg2",
                                 device_id=local_rank, n_inflight=2, host_output=True, timing=(i == 0))
             for i in range(Se)]
    for s, (a, b) in zip(esess, inputs):
        s.upload(0, a)
        s.upload(1, b)
    esteps = max(3, min(args.steps, args.e2e_steps))

    pending = []

    def estep(k):
        """Submit step k for every session (H2D + kernel + D2H queued), then collect step k-1:
        two staging slots per session keep the copy engines busy while the host reads results."""
        nonlocal pending
        ids = [s.submit(k & 1) for s in esess]
        out = 0
        for s, i in zip(esess, pending):
            out += s.wait(i).damaged_count
        pending = ids
        return out

    def edrain():
        nonlocal pending
        for s, i in zip(esess, pending):
            s.wait(i)
        pending = []

    for k in range(3):
        estep(k)
    edrain()
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    for k in range(esteps):
        estep(k + 3)
    edrain()
    e2e_t = time.perf_counter() - t0
    if dist:
        import torch
        t = torch.tensor([e2e_t], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_t = float(t[0])
    e2e_fps = Se * esteps * world / e2e_t
    # Same pipeline fed from PAGEABLE capture buffers (what a GstBuffer of ximagesrc is) through
    # sk_submit_host: the driver stages the copy; reported beside the pinned-staging figure.
    pg_steps = max(3, esteps // 4)
    pend = []
    t0 = time.perf_counter()
    for k in range(pg_steps):
        ids = [s.submit_host(k & 1, pair[k & 1]) for s, pair in zip(esess, inputs)]
        for s, i in zip(esess, pend):
            s.wait(i)
        pend = ids
    for s, i in zip(esess, pend):
        s.wait(i)
    e2e_pageable_fps = Se * pg_steps / (time.perf_counter() - t0)
    frame_bytes = esess[0].frame_bytes
    # PCIe in isolation (north_star: the copies are reported separately from the HBM roofline):
    # one session, one frame at a time, CUDA events around its H2D and D2H.
    h2d, d2h = [], []
    for k in range(8):
        rr = esess[0].wait(esess[0].submit(k & 1))
        h2d.append(rr.h2d_us); d2h.append(rr.d2h_us)
    pcie = {"h2d_gbs_isolated": 4 * W * H / (statistics.median(h2d) * 1e-6) / 1e9,
            "d2h_gbs_isolated": (frame_bytes + 2 * (W // 16) * (H // 16)) / (statistics.median(d2h) * 1e-6) / 1e9,
            "how": "median of 8 single-frame submits, CUDA events around cudaMemcpyAsync from/to pinned memory"}
    meta_bytes = 2 * (((r.damage_map.size + 3) // 4) * 4)      # damage map + policy map
    for s in esess:
        s.close()

    if rank != 0:
        return
    peak, peak_src = measured_peak()
    bytes_per_frame = synth.algorithmic_bytes(W, H, W, H, damage=True)
    k_us = statistics.mean(kernel_us)
    achieved = bytes_per_frame * S / (k_us * 1e-6) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = int(json.load(open(tp))["dram_bytes_per_frame"] * S)
        except Exception:
            traffic = None
    cpu = cpu_baseline() if not args.no_cpu_baseline else None
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": elapsed / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sessions_per_gpu": S, "step": "one sk_batch_submit launch over all sessions",
                   "matrix": "bt709", "chroma_site": "h-cosited", "l2": "inputs larger than L2 "
                   f"({S}x(33.2+12.4+12.4) MB per step, two resident frames per session alternated)",
                   "kernel_variant": args.variant if args.variant is not None else 0},
        "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": world * Se * 4 * W * H,
                "d2h_bytes_per_step": world * Se * (frame_bytes + meta_bytes), "sessions_per_gpu": Se, "steps": esteps,
                "h2d_gbs_per_gpu": e2e_fps / world * 4 * W * H / 1e9,
                "d2h_gbs_per_gpu": e2e_fps / world * (frame_bytes + meta_bytes) / 1e9,
                "timing": "wall clock between host syncs, max over ranks",
                "pageable_input_fps_rank0": e2e_pageable_fps},
        "pcie": pcie,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "frac_of_8TBs_nominal": achieved / 8000.0,
                     "kernel_us_per_launch": k_us, "algorithmic_bytes_per_launch": bytes_per_frame * S},
        "wall_s_timed_region": wall,
    }
    if cpu:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sessions", type=int, default=32, help="4K sessions per GPU (one launch per step)")
    ap.add_argument("--e2e-sessions", type=int, default=4)
    ap.add_argument("--e2e-steps", type=int, default=40)
    ap.add_argument("--variant", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-snake", action="store_true", help="tuning: do not alternate the traversal direction per launch")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        if args.steps > 20:
            args.steps = 20          # bounded: each step is already seconds of CPU work
        run_reference(args, rank, world)
        return
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist_mod.init_process_group(backend="gloo", rank=rank, world_size=world)
        dist = dist_mod
    try:
        run_b200(args, rank, world, local_rank, dist)
    finally:
        if dist:
            dist.barrier()
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
