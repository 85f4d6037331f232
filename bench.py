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

Keys of the one JSON line (rank 0 only):
  value     frames/s with inputs already in HBM (device-event timed, max over ranks)
  e2e       the same metric through the ELEMENT's per-frame call (PreEncodeStage.transform, what
            GstBaseTransform::transform runs): E stage instances on E host threads, each handed
            pageable host buffers from a recycled pool (input and output) every step inside the
            timed region — H2D from the input buffer, kernel, D2H into the output buffer.
            Side keys: the C-ABI-only figure on pre-filled pinned staging, un-pinnable one-off
            buffers, device consumers (no frame D2H), damage-gated read-back, coalesced launches.
  roofline  algorithmic bytes per launch / CUDA-event kernel time vs MEASURED_PEAKS.json
  configs   every other BASELINE.json config, device-resident (N=1 only): frames/s, kernel µs,
            roofline fraction by the BASELINE.md §4 formula, parity of one member vs the oracle
  pcie      isolated copy rates and the bare-copy ceiling with all ranks copying at once
  cpu_baseline  the C oracle (restatement of videoconvert, "port") on the host cores

`--impl reference` times that CPU path alone on the same workload (GStreamer
itself cannot be installed here).
Under torchrun (N>1) sessions are sharded S per GPU with no data-path
collective; torch.distributed (gloo) is used only for the barrier and the
max-over-ranks of the timing.  Session i lives on GPU i mod N (the reference's
only multi-GPU notion is `--gpu_id` → `cuda-device-id`,
src/selkies/legacy/webrtc.py:459-461, gstwebrtc_app.py:264-272).
"""
from __future__ import annotations

import argparse
import glob
import hashlib
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
KERNEL_SOURCE = os.path.join(ROOT, "selkies_gstreamer_b200", "csrc", "sk_kernels.cu")


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


# ----------------------------------------------------------------------------- placement
def gpu_for_session(session_index: int, n_gpus: int) -> int:
    """Round-robin: session i → GPU i mod N."""
    if n_gpus < 1:
        raise ValueError("n_gpus must be >= 1")
    if session_index < 0:
        raise ValueError("session_index must be >= 0")
    return session_index % n_gpus


def sessions_of_rank(rank: int, world_size: int, total_sessions: int) -> list[int]:
    """Global session ids owned by `rank` (one rank per GPU)."""
    if not 0 <= rank < world_size:
        raise ValueError("rank out of range")
    return [i for i in range(total_sessions) if gpu_for_session(i, world_size) == rank]


def placement(total_sessions: int, n_gpus: int) -> list[list[int]]:
    return [sessions_of_rank(r, n_gpus, total_sessions) for r in range(n_gpus)]


def aggregate_fps(frames_per_rank: list[int], seconds_per_rank: list[float]) -> float:
    """Whole-job throughput: all frames over the slowest rank's time (weak scaling)."""
    if not frames_per_rank or len(frames_per_rank) != len(seconds_per_rank):
        raise ValueError("one frame count and one time per rank")
    return sum(frames_per_rank) / max(seconds_per_rank)


def reduce_over_ranks(dist, values: list[float], op: str = "max") -> list[float]:
    """max / sum / min over ranks of host floats through torch.distributed (gloo); identity without it."""
    if dist is None:
        return [float(v) for v in values]
    import torch
    t = torch.tensor(values, dtype=torch.float64)
    dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "sum": dist.ReduceOp.SUM, "min": dist.ReduceOp.MIN}[op])
    return [float(v) for v in t]


def bind_to_gpu_numa_node(gpu_index: int) -> dict:
    """Pin this process to the CPUs of the NUMA node its GPU hangs off, when the box exposes more than
    one node — before the library allocates any pinned memory, so that staging lands on that node."""
    try:
        nodes = sorted(glob.glob("/sys/devices/system/node/node[0-9]*"))
        if len(nodes) < 2:
            return {"bound": False, "numa_nodes": len(nodes), "reason": "one NUMA node exposed"}
        bdf = subprocess.run(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        bdf = bdf[-12:]                              # sysfs uses a 4-digit PCI domain
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return {"bound": False, "numa_nodes": len(nodes), "reason": f"GPU {bdf} reports no NUMA node"}
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return {"bound": False, "numa_nodes": len(nodes), "reason": f"no allowed CPU on node {node}"}
        os.sched_setaffinity(0, cpus)
        return {"bound": True, "numa_nodes": len(nodes), "node": node, "cpus": len(cpus)}
    except Exception as e:                           # noqa: BLE001 - affinity is best effort
        return {"bound": False, "reason": f"{type(e).__name__}: {e}"}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def code_sha(path: str) -> str:
    """sha of a CUDA/C++ source with comments and white space taken out: what the compiler sees, so that an
    edited comment does not make a measurement look stale (tools/make_traffic.py stamps with the same function)."""
    import re
    src = open(path, "r", encoding="utf-8", errors="replace").read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    return hashlib.sha256("".join(src.split()).encode()).hexdigest()[:16]


def kernel_sha() -> str:
    return code_sha(KERNEL_SOURCE)


def recorded_traffic(S: int):
    """ncu DRAM bytes per frame from profiles/traffic.json — only if that capture was taken with the
    kernel source this run was built from; otherwise the figure is stale and is not reported."""
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        j = json.load(open(tp))
        if j.get("kernel_sha") != kernel_sha():
            return None, f"profiles/traffic.json is for kernel source {j.get('kernel_sha')}, this is {kernel_sha()}: stale, not reported"
        return int(j["dram_bytes_per_frame"] * S), f"profiles/traffic.json @ sk_kernels.cu sha {j['kernel_sha']} (ncu --set full, per frame x {S})"
    except Exception as e:                           # noqa: BLE001
        return None, f"no usable profiles/traffic.json ({type(e).__name__})"


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


# ----------------------------------------------------------------------------- inputs
def make_inputs(n_sessions: int, rank: int, w: int = W, h: int = H, fraction: float = 0.05):
    """Two captured frames per session: desktop content, then `fraction` of its macroblocks changed."""
    from selkies_gstreamer_b200 import synth
    base = synth.desktop(w, h, seed=42 + rank)
    out = []
    for i in range(n_sessions):
        a = np.roll(base, shift=(37 * i, 101 * i), axis=(0, 1))
        b = a if fraction <= 0 else synth.mutate(a, fraction, seed=1000 * rank + i, aligned=False)
        out.append((a, b))
    return out


# ----------------------------------------------------------------------------- CPU arm
def cpu_pass(frames, prevs, threads, w=W, h=H):
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
        o.damage(w, h, o.OUT_NV12, cur, p)
        return cur

    if workers == 1:
        return [one(fp) for fp in zip(frames, prevs)]
    with ThreadPoolExecutor(max_workers=workers) as ex:
        return list(ex.map(one, zip(frames, prevs)))


def cpu_baseline(budget_s: float = 10.0):
    """Bounded sample of the same workload on this box's host cores (C oracle, 'port'):
    all cores, and the reference's own single-pipeline setting n-threads = min(4, ncpu-1)."""
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
    # one pipeline as the reference configures it (gstwebrtc_app.py:612)
    nt = min(4, max(1, threads - 1))
    a, b = inputs[0]
    prev = o.convert(a, out_fmt=o.OUT_NV12, matrix=o.MATRIX_BT709, chroma_site=o.SITE_H_COSITED, n_threads=nt)
    m, t1 = 0, time.perf_counter()
    while True:
        cur = o.convert((b, a)[m & 1], out_fmt=o.OUT_NV12, matrix=o.MATRIX_BT709, chroma_site=o.SITE_H_COSITED, n_threads=nt)
        o.damage(W, H, o.OUT_NV12, cur, prev)
        prev = cur
        m += 1
        dt1 = time.perf_counter() - t1
        if dt1 >= budget_s / 3 or m >= 500:
            break
    return {"value": n * s / dt, "unit": "frames/s", "cores": threads, "kind": "port",
            "sample": f"{n * s} 4K frames ({s} sessions in parallel x {n} passes), convert+damage, C oracle, "
                      f"{threads} threads in total, {dt:.1f} s",
            "reference_setting": {"value": m / dt1, "unit": "frames/s", "cores": nt,
                                  "what": f"ONE pipeline with n-threads = min(4, ncpu-1) = {nt} row bands "
                                          f"(gstwebrtc_app.py:612), {m} 4K frames in {dt1:.1f} s"}}


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


# ----------------------------------------------------------------------------- GPU arm: helpers
def oracle_check(sess, frames_ab, k_last, fmt_name, out_w=None, out_h=None, scale_variant=0):
    """Outside every timed region: the last two steps of `sess` (a batch member) against the oracle.
    Step k converted frames_ab[k & 1]; the map of step k_last is the diff of steps k_last and k_last-1."""
    from oracle import cpu_oracle as o
    fmt = {"NV12": o.OUT_NV12, "I420": o.OUT_I420}[fmt_name]
    ow, oh = out_w or sess.width, out_h or sess.height
    res = sess.wait(k_last)
    got = sess.read_frame(k_last)
    kw = dict(out_fmt=fmt, matrix=sess.matrix, chroma_site=sess.site, scale_variant=scale_variant, n_threads=8)
    if (ow, oh) != (sess.width, sess.height):
        cur = o.convert(frames_ab[1], ow, oh, **kw)
        prv = o.convert(frames_ab[0], ow, oh, **kw)
    else:
        cur = o.convert(frames_ab[1], **kw)
        prv = o.convert(frames_ab[0], **kw)
    ok = bool((got == cur).all())
    if res.damage_map is not None:
        em, en = o.damage(ow, oh, fmt, cur, prv)
        ok = ok and bool((res.damage_map == em).all()) and res.damaged_count == en
    return ok


def device_config(sk, name, in_w, in_h, out_w, out_h, fmt, S, steps, local_rank, peak, *, fraction=0.05,
                  damage=True, colorimetry="bt709", site="mpeg2", single=False, inputs=None, check=True):
    """One BASELINE.json config, device-resident: S sessions, one launch per step (or one session per
    launch with single=True).  Returns frames/s, kernel µs, roofline fraction (BASELINE.md §4 formula)."""
    from selkies_gstreamer_b200 import synth
    if inputs is None:
        inputs = make_inputs(S, 7, in_w, in_h, fraction)
    sess = [sk.PreencodeSession(in_w, in_h, out_width=out_w, out_height=out_h, out_format=fmt, colorimetry=colorimetry,
                                chroma_site=site, device_id=local_rank, n_inflight=2, host_output=False,
                                damage=damage, timing=(i == 0), spin_wait=True) for i in range(S)]
    try:
        for s, (a, b) in zip(sess, inputs):
            s.upload(0, a); s.wait(s.submit(0))
            s.upload(1, b); s.wait(s.submit(1))
        stride = sess[0].device_input(0)[1]
        kus = []
        fid = 100
        if single:
            s = sess[0]
            ptrs = [s.device_input(k)[0] for k in (0, 1)]
            for k in range(5):
                s.wait(s.submit_device(ptrs[k & 1], stride, fid)); fid += 1
            s.marker(0)
            t0 = time.perf_counter()
            for k in range(steps):
                s.submit_device(ptrs[(fid) & 1], stride, fid)
                if k:
                    s.wait(fid - 1); kus.append(s.last_kernel_us())
                fid += 1
            s.marker(1)
            s.wait(fid - 1); kus.append(s.last_kernel_us())
            wall = time.perf_counter() - t0
            dev_us = s.marker_elapsed_us()
            frames = steps
        else:
            batch = sk.SessionBatch(sess)
            batch.prepare()
            sets = [batch.frame_set([s.device_input(k)[0] for s in sess]) for k in (0, 1)]
            for k in range(5):
                batch.submit(sets[fid & 1], stride, fid); batch.wait(fid); fid += 1
            sess[0].marker(0)
            t0 = time.perf_counter()
            for k in range(steps):
                batch.submit(sets[fid & 1], stride, fid)
                if k:
                    batch.wait(fid - 1); kus.append(sess[0].last_kernel_us())
                fid += 1
            sess[0].marker(1)
            batch.wait(fid - 1); kus.append(sess[0].last_kernel_us())
            wall = time.perf_counter() - t0
            dev_us = sess[0].marker_elapsed_us()
            frames = steps * S
        k_last = fid - 1
        member = min(1, S - 1)
        ab = inputs[member]
        pair = (ab[(k_last - 1) & 1], ab[k_last & 1])
        parity = oracle_check(sess[member], pair, k_last, fmt, out_w, out_h) if check else None
        n_per = 1 if single else S
        bpf = synth.algorithmic_bytes(in_w, in_h, out_w, out_h, damage=damage)
        k_us = statistics.mean(kus)
        achieved = bpf * n_per / (k_us * 1e-6) / 1e9
        return {"workload": name, "frames_per_s": frames / (dev_us * 1e-6), "sessions_per_launch": n_per,
                "kernel_us_per_launch": k_us, "kernel_us_per_frame": k_us / n_per, "steps": steps,
                "algorithmic_bytes_per_frame": bpf,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "frac_of_8TBs_nominal": achieved / 8000.0},
                "wall_fps": frames / wall, "damage_fraction": fraction if damage else None,
                "parity_vs_oracle": parity}
    finally:
        for s in sess[1:]:
            s.close()
        sess[0].close()


class StageRig:
    """E PreEncodeStage instances, each on its own host thread, each cycling through a small pool of
    pageable host buffers — what E pipelines of one process look like to the stage (buffer pools of
    ximagesrc upstream and of the encoder downstream, gstwebrtc_app.py:210-255, :941-994)."""

    def __init__(self, E, inputs, local_rank, pool=3, w=W, h=H, **props):
        from selkies_gstreamer_b200.stage import PreEncodeStage
        self.E, self.pool = E, pool
        self.stages, self.inbufs, self.outbufs = [], [], []
        for i in range(E):
            st = PreEncodeStage(f"bench{i}")
            st.set_property("cuda-device-id", local_rank)
            for k, v in props.items():
                st.set_property(k, v)
            inc = {"name": "video/x-raw", "format": "BGRx", "width": w, "height": h}
            outc = {"name": "video/x-raw", "format": "NV12", "width": w, "height": h, "colorimetry": "bt709",
                    "chroma-site": "mpeg2"}
            st.start()
            if not st.set_caps(inc, outc):
                raise SystemExit(f"stage set_caps failed: {st.last_error}")
            self.stages.append(st)
            a, b = inputs[i % len(inputs)]
            # pageable pool buffers (plain numpy allocations), pre-filled like a capture would have
            self.inbufs.append([np.array((a, b)[k & 1], copy=True) for k in range(pool)])
            self.outbufs.append([np.empty(st.session.frame_bytes, np.uint8) for _ in range(pool)])
        self.frame_bytes = self.stages[0].session.frame_bytes
        self.errors = []

    def _worker(self, i, k0, steps, barrier):
        st, ins, outs = self.stages[i], self.inbufs[i], self.outbufs[i]
        barrier.wait()
        for k in range(k0, k0 + steps):
            src, dst = ins[k % self.pool], outs[k % self.pool]
            rc = st.transform(src.ctypes.data, src.nbytes, dst.ctypes.data, dst.nbytes)
            if rc != 0:
                self.errors.append(st.last_error)
                return

    def run(self, k0, steps):
        """All stages convert `steps` frames each, concurrently; returns the wall time."""
        barrier = threading.Barrier(self.E + 1)
        ths = [threading.Thread(target=self._worker, args=(i, k0, steps, barrier)) for i in range(self.E)]
        for t in ths:
            t.start()
        barrier.wait()
        t0 = time.perf_counter()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        if self.errors:
            raise SystemExit(f"stage transform failed: {self.errors[0]}")
        return dt

    def check(self, inputs):
        """Output buffer of the last step of stage 0 against the oracle (outside the timed region)."""
        from oracle import cpu_oracle as o
        st = self.stages[0]
        a = self.inbufs[0][(self._last_k) % self.pool]
        exp = o.convert(a, out_fmt=o.OUT_NV12, matrix=st.session.matrix, chroma_site=st.session.site, n_threads=8)
        return bool((self.outbufs[0][self._last_k % self.pool] == exp).all())

    def close(self):
        for st in self.stages:
            st.stop()


def timed_rig(rig, warm, steps, dist, reps=1):
    """Warm up, then `reps` timed regions of `steps` frames per stage (barrier before each); returns the list of
    max-over-ranks times.  The host side of this path (page-locked DMA, thread wake-ups) is noisy on a shared
    box, so the headline takes the median region."""
    rig.run(0, warm)
    times, k0 = [], warm
    for _ in range(reps):
        if dist:
            dist.barrier()
        dt = rig.run(k0, steps)
        times.append(reduce_over_ranks(dist, [dt], "max")[0])
        k0 += steps
    rig._last_k = k0 - 1
    return times


# ----------------------------------------------------------------------------- GPU arm
def run_b200(args, rank, world, local_rank, dist):
    affinity = bind_to_gpu_numa_node(local_rank) if not args.no_affinity else {"bound": False, "reason": "--no-affinity"}
    import selkies_gstreamer_b200 as sk
    from selkies_gstreamer_b200 import _abi, synth
    import ctypes

    lib = _abi.load()
    if lib.sk_device_count() <= local_rank:
        raise SystemExit(f"rank {rank}: CUDA device {local_rank} not available — the product path has no fallback")
    if args.variant is not None or args.no_snake:
        if not hasattr(lib, "sk_tuning_set_variant"):
            raise SystemExit("--variant/--no-snake need the tuning build (make -C selkies_gstreamer_b200/csrc tuning; "
                             "SELKIES_PREENCODE_LIB=.../libselkies_preencode_tuning.so)")
        if args.variant is not None:
            lib.sk_tuning_set_variant(args.variant)
        if args.no_snake:
            lib.sk_tuning_set_variant(102)
    S = args.sessions
    inputs = make_inputs(S, rank)
    peak, peak_src = measured_peak()

    # ---- device-resident arm: S sessions, one launch per step
    sess = [sk.PreencodeSession(W, H, out_format="NV12", colorimetry="bt709", chroma_site="mpeg2",
                                device_id=local_rank, n_inflight=2, host_output=False, timing=(i == 0), spin_wait=True)
            for i in range(S)]
    for s, (a, b) in zip(sess, inputs):
        s.upload(0, a); s.wait(s.submit(0))
        s.upload(1, b); s.wait(s.submit(1))
    batch = sk.SessionBatch(sess)
    batch.prepare()
    sets = [batch.frame_set([s.device_input(k)[0] for s in sess]) for k in (0, 1)]
    stride = sess[0].device_input(0)[1]
    leader = sess[0]
    fid = 1000
    counts = []

    def step(wait_prev=True):
        nonlocal fid, counts
        batch.submit(sets[fid & 1], stride, fid)
        if wait_prev and fid > 1000:
            counts = batch.wait(fid - 1)      # the host consumes every session's damage count each step
        fid += 1

    def sync_all():
        nonlocal counts
        counts = batch.wait(fid - 1)

    sampler = ClockSampler(local_rank)
    sampler.start()                            # nvidia-smi needs a moment to come up (longer on an 8-GPU box)
    for k in range(args.warmup):
        step()
    sync_all()
    if dist:
        dist.barrier()
    launches0 = sk.launch_count()
    kernel_us = []
    leader.marker(0)
    t0 = time.perf_counter()
    for k in range(args.steps):
        step()
        if k > 0:
            kernel_us.append(leader.last_kernel_us())
    leader.marker(1)
    sync_all()
    wall = time.perf_counter() - t0
    kernel_us.append(leader.last_kernel_us())
    dev_us = leader.marker_elapsed_us()
    launches = sk.launch_count() - launches0
    t_end = time.perf_counter()

    def samples_under_load():
        return sum(1 for r in sampler.rows if r[0] >= t0)

    # nvidia-smi samples every 50 ms: keep the same load up, untimed, until a few samples exist
    while (time.perf_counter() - t0 < 0.4 or samples_under_load() < 3) and time.perf_counter() - t0 < 5.0:
        step()
    sync_all()
    clocks = sampler.stop(t0, t_end)
    # Parity of what was just benchmarked, on EVERY rank (so device_id > 0 is covered under torchrun):
    # one member of the 32-session launch, frame + map + count, against the oracle.
    k_last = fid - 1
    m = (rank + 1) % S
    ab = inputs[m]
    parity_ok = oracle_check(sess[m], (ab[(k_last - 1) & 1], ab[k_last & 1]), k_last, "NV12")
    if not parity_ok:
        raise SystemExit(f"rank {rank}: the benchmarked launch does not match the oracle (member {m})")

    elapsed = max(dev_us * 1e-6, 1e-9)
    elapsed, wall = reduce_over_ranks(dist, [elapsed, wall], "max")
    parity_ranks = int(round(reduce_over_ranks(dist, [1.0 if parity_ok else 0.0], "sum")[0]))
    frames_total = S * args.steps * world
    value = frames_total / elapsed

    # ---- damage-fraction sweep on the same sessions (N=1 only): 0 / 0.5 % / 100 % beside the headline's 5 %
    configs = {}
    if world == 1 and not args.no_configs:
        csteps = max(50, min(args.steps, 300))
        for frac, label in ((0.0, "4k_nv12_damage_0pct"), (0.005, "4k_nv12_damage_0.5pct"), (1.0, "4k_nv12_damage_100pct")):
            for i, (s, (a, _)) in enumerate(zip(sess, inputs)):
                b2 = a if frac == 0.0 else synth.mutate(a, frac, seed=77 + i, aligned=False)
                s.upload(1, b2); s.wait(s.submit(1))
            step(wait_prev=False)              # the uploads above went through the rings: step fid-1 is gone
            for k in range(3):
                step()
            sync_all()
            kus = []
            leader.marker(0)
            for k in range(csteps):
                step()
                if k > 0:
                    kus.append(leader.last_kernel_us())
            leader.marker(1)
            sync_all()
            kus.append(leader.last_kernel_us())
            k_us = statistics.mean(kus)
            bpf = synth.algorithmic_bytes(W, H, W, H, damage=True)
            ach = bpf * S / (k_us * 1e-6) / 1e9
            nblk = (W // 16) * (H // 16)
            configs[label] = {"workload": f"4K NV12 + damage, {frac * 100:g} % of macroblocks changed per frame",
                              "frames_per_s": S * csteps / (leader.marker_elapsed_us() * 1e-6), "sessions_per_launch": S,
                              "kernel_us_per_launch": k_us, "kernel_us_per_frame": k_us / S, "steps": csteps,
                              "damaged_fraction_measured": sum(counts) / (S * nblk),
                              "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                           "frac_of_8TBs_nominal": ach / 8000.0}}
    for s in sess[1:]:
        s.close()
    sess[0].close()

    # ---- end-to-end arm: the element's per-frame call on E stage instances, pageable pool buffers
    E = min(args.e2e_sessions, S)
    esteps = max(3, min(args.steps, args.e2e_steps))
    rig = StageRig(E, inputs, local_rank)
    e2e_times = timed_rig(rig, 2 * rig.pool, esteps, dist, reps=3)
    e2e_t = statistics.median(e2e_times)
    e2e_fps = E * esteps * world / e2e_t
    e2e_parity = rig.check(inputs)
    frame_bytes = rig.frame_bytes
    pins = rig.stages[0]._pins
    pin_stats = {"pinned_buffers_stage0": len(pins), "hits": pins.hits, "misses": pins.misses, "enabled": pins.enabled}
    rig.close()
    if not e2e_parity:
        raise SystemExit(f"rank {rank}: the element's output buffer does not match the oracle")
    side = {}
    if not args.no_side:
        # same call, launches shared between the stages (coalesce.py)
        rig = StageRig(E, inputs, local_rank, coalesce=True)
        t = timed_rig(rig, 2 * rig.pool, esteps, dist)[0]
        co = rig.stages[0]._coalescer
        side["coalesced_fps"] = E * esteps * world / t
        side["coalesced_frames_per_launch_rank0"] = (co.frames / co.launches) if co and co.launches else None
        rig.close()
        # same call, buffers that are never seen twice and so cannot be page-locked (driver-staged copies)
        rig = StageRig(E, inputs, local_rank, **{"pin-buffers": False})
        psteps = max(3, esteps // 4)
        t = timed_rig(rig, 2, psteps, dist)[0]
        side["unpinned_fps"] = E * psteps * world / t
        rig.close()

        # the same call at the reference's own default desktop size (BASELINE.json config 1: 1920x1080), with the
        # host cores' rate for that size beside it (bounded 3 s sample)
        if world == 1:
            w1, h1 = 1920, 1080
            in1 = make_inputs(E, rank, w1, h1)
            rig = StageRig(E, in1, local_rank, w=w1, h=h1)
            s1 = esteps * 4
            t = statistics.median(timed_rig(rig, 2 * rig.pool, s1, dist, reps=3))
            rig.close()
            threads = len(os.sched_getaffinity(0))
            prevs = cpu_pass([p[0] for p in in1], [None] * E, threads, w1, h1)
            n1, t0 = 0, time.perf_counter()
            while time.perf_counter() - t0 < 3.0:
                prevs = cpu_pass([p[(n1 + 1) & 1] for p in in1], prevs, threads, w1, h1)
                n1 += 1
            side["config1_1080p"] = {"fps": E * s1 / t, "h2d_gbs": E * s1 * 4 * w1 * h1 / t / 1e9, "stages": E,
                                     "cpu_port_fps": n1 * E / (time.perf_counter() - t0), "cpu_cores": threads,
                                     "what": "PreEncodeStage.transform on 1920x1080 BGRx -> NV12 frames, pageable pool buffers, "
                                             "median of 3 regions; C oracle on all host cores beside it"}

        # C ABI only, pre-filled pinned staging (round 1's e2e figure), frame + map read back
        esess = [sk.PreencodeSession(W, H, out_format="NV12", colorimetry="bt709", chroma_site="mpeg2",
                                     device_id=local_rank, n_inflight=2, host_output=True, timing=(i == 0), spin_wait=True)
                 for i in range(E)]
        for s, (a, b) in zip(esess, inputs):
            s.upload(0, a)
            s.upload(1, b)

        def abi_loop(sessions, n, submit):
            pending = []
            t0 = time.perf_counter()
            for k in range(n):
                ids = [submit(s, k) for s in sessions]
                for s, i in zip(sessions, pending):
                    s.wait(i)
                pending = ids
            for s, i in zip(sessions, pending):
                s.wait(i)
            return time.perf_counter() - t0

        abi_loop(esess, 3, lambda s, k: s.submit(k & 1))
        if dist:
            dist.barrier()
        t = reduce_over_ranks(dist, [abi_loop(esess, esteps, lambda s, k: s.submit(k & 1))], "max")[0]
        side["abi_pinned_staging_fps"] = E * esteps * world / t
        # PCIe in isolation: one session, one frame at a time, CUDA events around its H2D and D2H
        h2d, d2h = [], []
        for k in range(8):
            rr = esess[0].wait(esess[0].submit(k & 1))
            h2d.append(rr.h2d_us); d2h.append(rr.d2h_us)
        map_bytes = 2 * (W // 16) * (H // 16)
        pcie = {"h2d_gbs_isolated": 4 * W * H / (statistics.median(h2d) * 1e-6) / 1e9,
                "d2h_gbs_isolated": (frame_bytes + map_bytes) / (statistics.median(d2h) * 1e-6) / 1e9,
                "how": "median of 8 single-frame submits, CUDA events around cudaMemcpyAsync from/to pinned memory"}
        for s in esess:
            s.close()

        # device consumers (NVENC branch): H2D + kernel + the 32 KB maps only, no frame D2H
        dsess = [sk.PreencodeSession(W, H, out_format="NV12", colorimetry="bt709", chroma_site="mpeg2",
                                     device_id=local_rank, n_inflight=2, host_output=False) for i in range(E)]
        for s, (a, b) in zip(dsess, inputs):
            s.upload(0, a)
            s.upload(1, b)
        abi_loop(dsess, 3, lambda s, k: s.submit(k & 1))
        if dist:
            dist.barrier()
        t = reduce_over_ranks(dist, [abi_loop(dsess, esteps, lambda s, k: s.submit(k & 1))], "max")[0]
        side["device_consumer_fps"] = E * esteps * world / t
        for s in dsess:
            s.close()

        # damage-gated read-back: only the macroblocks that changed cross PCIe on the way back
        gsess = [sk.PreencodeSession(W, H, out_format="NV12", colorimetry="bt709", chroma_site="mpeg2",
                                     device_id=local_rank, n_inflight=2, host_output=True, gated_readback=True)
                 for i in range(E)]
        for s, (a, b) in zip(gsess, inputs):
            s.upload(0, a)
            s.upload(1, b)
        abi_loop(gsess, 4, lambda s, k: s.submit(k & 1))
        if dist:
            dist.barrier()
        t = reduce_over_ranks(dist, [abi_loop(gsess, esteps, lambda s, k: s.submit(k & 1))], "max")[0]
        rr = gsess[0].wait(gsess[0].submit(0))
        side["gated_readback_fps"] = E * esteps * world / t
        side["gated_readback_bytes_per_frame"] = rr.readback_bytes
        from oracle import cpu_oracle as o
        exp = o.convert(inputs[0][0], out_fmt=o.OUT_NV12, matrix=gsess[0].matrix, chroma_site=gsess[0].site, n_threads=8)
        side["gated_readback_parity_vs_oracle"] = bool((rr.frame == exp).all())
        for s in gsess:
            s.close()

        # bare-copy ceiling, every rank copying at once, no session machinery (sk_pcie_probe)
        up, down = ctypes.c_double(), ctypes.c_double()
        if dist:
            dist.barrier()
        lib.sk_pcie_probe(local_rank, 4 * W * H, 300, 2, ctypes.byref(up), ctypes.byref(down))
        both = (up.value, down.value)
        if dist:
            dist.barrier()
        lib.sk_pcie_probe(local_rank, 4 * W * H, 300, 0, ctypes.byref(up), ctypes.byref(down))
        only_up = up.value
        agg = reduce_over_ranks(dist, [both[0], both[1], only_up], "sum")
        mn = reduce_over_ranks(dist, [both[0], both[1], only_up], "min")
        pcie["aggregate_ceiling_gbs"] = {"h2d_with_d2h": agg[0], "d2h_with_h2d": agg[1], "h2d_alone": agg[2],
                                         "per_gpu_min": {"h2d_with_d2h": mn[0], "d2h_with_h2d": mn[1], "h2d_alone": mn[2]},
                                         "how": "SUSTAINED: 300 x 33 MB cudaMemcpyAsync per direction from/to 8 pinned buffers used in turn, all ranks at "
                                                "once (a 24-copy burst reads 2x higher on the 8-GPU box and is not what a stream of frames gets)"}
        # How close the element's path comes to that ceiling.  The probe's two-direction figure has D2H at full
        # rate beside the uploads; the path brings 0.38 bytes down per byte up, so its own ceiling lies between
        # that figure and the uploads-alone one: both fractions are reported (the first may exceed 1).
        h2d_rate = e2e_fps * 4 * W * H / 1e9
        pcie["e2e_h2d_fraction_of_ceiling"] = h2d_rate / agg[0] if agg[0] else None
        pcie["e2e_h2d_fraction_of_h2d_alone"] = h2d_rate / agg[2] if agg[2] else None
    else:
        pcie = {}

    # ---- the other BASELINE.json configs, device-resident, N=1 only
    if world == 1 and not args.no_configs:
        csteps = max(50, min(args.steps, 300))
        configs["1080p_i420"] = device_config(sk, "1920x1080 BGRx -> I420 + damage (config 1's conversion)", 1920, 1080,
                                              1920, 1080, "I420", 32, csteps, local_rank, peak)
        configs["1440p_to_1080p_i420_bt709"] = device_config(
            sk, "2560x1440 -> 1920x1080 bilinear -> I420, BT.709 + damage (config 3)", 2560, 1440, 1920, 1080, "I420",
            32, csteps, local_rank, peak)
        configs["8k_i420_sparse_damage"] = device_config(
            sk, "7680x4320 BGRx -> I420, 0.5 % of macroblocks changed (config 5)", 7680, 4320, 7680, 4320, "I420",
            8, max(30, csteps // 3), local_rank, peak, fraction=0.005)
        configs["4k_nv12_convert_only"] = device_config(
            sk, "4K NV12 convert only (B': no previous-frame read, no map)", W, H, W, H, "NV12", 32, csteps, local_rank,
            peak, damage=False, inputs=inputs)
        configs["4k_nv12_single_session"] = device_config(
            sk, "4K NV12 + damage, ONE session per launch (what a lone pipeline gets)", W, H, W, H, "NV12", 1, csteps * 2,
            local_rank, peak, single=True, inputs=inputs[:2])
        configs["1080p_i420_single_session"] = device_config(
            sk, "1080p I420 + damage, ONE session per launch", 1920, 1080, 1920, 1080, "I420", 1, csteps * 2, local_rank,
            peak, single=True)
        # not a BASELINE.json config (the reference never enlarges): the scale-behind-the-matrix chain, for the record
        configs["1080p_to_4k_nv12_enlarge"] = device_config(
            sk, "1920x1080 -> 3840x2160 bilinear (picture grows: scaled behind the matrix) -> NV12 + damage", 1920, 1080,
            W, H, "NV12", 8, max(30, csteps // 3), local_rank, peak)

    if rank != 0:
        return
    bytes_per_frame = synth.algorithmic_bytes(W, H, W, H, damage=True)
    k_us = statistics.mean(kernel_us)
    achieved = bytes_per_frame * S / (k_us * 1e-6) / 1e9
    traffic, traffic_src = recorded_traffic(S)
    cpu = cpu_baseline() if not args.no_cpu_baseline else None
    meta_bytes = 2 * (((W // 16) * (H // 16) + 3) // 4 * 4) + 128
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": elapsed / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sessions_per_gpu": S, "step": "one sk_batch_submit launch over all sessions",
                   "matrix": "bt709", "chroma_site": "h-cosited", "l2": "inputs larger than L2 "
                   f"({S}x(33.2+12.4+12.4) MB per step, two resident frames per session alternated)",
                   "kernel_variant": args.variant if args.variant is not None else 0,
                   "host_wait": "the device-resident arms busy-poll (SK_FLAG_SPIN_WAIT) so that the launch queue never "
                                "runs dry; the e2e arm's stages sleep on their frames (the library default)"},
        "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": world * E * 4 * W * H,
                "d2h_bytes_per_step": world * E * (frame_bytes + meta_bytes), "stages_per_gpu": E, "steps": esteps,
                "call": "PreEncodeStage.transform(in_ptr, out_ptr) on E stage instances x E host threads; pageable "
                        "numpy buffers from a pool of 3 per stage and direction, page-locked by the stage on first "
                        "sight (warm-up); input and output buffer handed over every step inside the timed region",
                "h2d_gbs_per_gpu": e2e_fps / world * 4 * W * H / 1e9,
                "d2h_gbs_per_gpu": e2e_fps / world * (frame_bytes + meta_bytes) / 1e9,
                "timing": "wall clock around the E threads' loops, max over ranks; median of 3 timed regions",
                "regions_fps": [E * esteps * world / t for t in e2e_times],
                "parity_vs_oracle": e2e_parity, "pin_cache": pin_stats, **side},
        "pcie": pcie,
        "affinity": affinity,
        "gpu_launches": int(launches),
        "parity_checked_ranks": parity_ranks,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "frac_of_8TBs_nominal": achieved / 8000.0,
                     "kernel_us_per_launch": k_us, "algorithmic_bytes_per_launch": bytes_per_frame * S},
        "wall_s_timed_region": wall,
    }
    if configs:
        line["configs"] = configs
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
    ap.add_argument("--e2e-sessions", type=int, default=4, help="stage instances (host threads) per GPU in the e2e arm")
    ap.add_argument("--e2e-steps", type=int, default=60)
    ap.add_argument("--variant", type=int, default=None, help="tuning build only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-config sweep (N=1 only anyway)")
    ap.add_argument("--no-side", action="store_true", help="skip the e2e side figures and the PCIe probes")
    ap.add_argument("--no-affinity", action="store_true", help="do not bind the rank to its GPU's NUMA node")
    ap.add_argument("--no-snake", action="store_true", help="tuning build: do not alternate the traversal direction per launch")
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
