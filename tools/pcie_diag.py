#!/usr/bin/env python3
"""Where does the end-to-end path lose PCIe rate when every GPU of the box copies at once?

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_diag.py

Per rank, all ranks at once after a barrier: the bare-copy probe (sk_pcie_probe) as a short burst and as a
sustained run, one direction and both; then the session path (pinned staging, sk_submit / sk_wait) with
1, 2, 4 sessions and with / without the frame coming back.  Rank 0 prints one markdown table (sum over ranks,
GB/s; per-GPU minimum in brackets)."""
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import selkies_gstreamer_b200 as sk  # noqa: E402
from selkies_gstreamer_b200 import _abi  # noqa: E402

W, H = 3840, 2160
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
if world > 1:
    dist.init_process_group("gloo", rank=rank, world_size=world)
lib = _abi.load()


def allred(vals):
    if world == 1:
        return list(vals), list(vals)
    t = torch.tensor(vals, dtype=torch.float64)
    s, m = t.clone(), t.clone()
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    dist.all_reduce(m, op=dist.ReduceOp.MIN)
    return [float(x) for x in s], [float(x) for x in m]


def barrier():
    if world > 1:
        dist.barrier()


rows = []


def probe(label, mode, iters):
    up, down = ctypes.c_double(), ctypes.c_double()
    barrier()
    lib.sk_pcie_probe(local, 4 * W * H, iters, mode, ctypes.byref(up), ctypes.byref(down))
    s, m = allred([up.value, down.value])
    rows.append((label, s[0], m[0], s[1], m[1]))


def sessions(label, n_sess, host_output, steps=40):
    frame = np.random.default_rng(rank).integers(0, 256, (H, W, 4), dtype=np.uint8)
    ss = [sk.PreencodeSession(W, H, out_format="NV12", colorimetry="bt709", chroma_site="mpeg2", device_id=local,
                              n_inflight=2, host_output=host_output, spin_wait=True) for _ in range(n_sess)]
    for s in ss:
        s.upload(0, frame); s.upload(1, frame)
    pend = []
    def loop(n):
        nonlocal pend
        for k in range(n):
            ids = [s.submit(k & 1) for s in ss]
            for s, i in zip(ss, pend):
                s.wait(i)
            pend = ids
        for s, i in zip(ss, pend):
            s.wait(i)
        pend = []
    loop(3)
    barrier()
    t0 = time.perf_counter()
    loop(steps)
    dt = time.perf_counter() - t0
    fb = ss[0].frame_bytes
    for s in ss:
        s.close()
    up = n_sess * steps * 4 * W * H / dt / 1e9
    down = (n_sess * steps * fb / dt / 1e9) if host_output else 0.0
    s_, m_ = allred([up, down])
    rows.append((label, s_[0], m_[0], s_[1], m_[1]))


probe("bare copies, both directions, burst (24 x 33 MB each way)", 2, 24)
probe("bare copies, both directions, sustained (400 x 33 MB each way)", 2, 400)
probe("bare copies, host->device only, sustained (400 x 33 MB)", 0, 400)
probe("bare copies, device->host only, sustained (400 x 33 MB)", 1, 400)
for n in (1, 2, 4):
    sessions(f"session path, {n} session(s) per GPU, frame + maps come back", n, True)
for n in (1, 4):
    sessions(f"session path, {n} session(s) per GPU, maps only (device consumer)", n, False)
if rank == 0:
    print(f"| what ({world} GPU(s), all at once) | H2D GB/s total (min per GPU) | D2H GB/s total (min per GPU) |")
    print("|---|---|---|")
    for (label, su, mu, sd, md) in rows:
        print(f"| {label} | {su:.1f} ({mu:.1f}) | {sd:.1f} ({md:.1f}) |")
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
