#!/usr/bin/env python3
"""Device time of sk_scale_frame_noalloc beside the session path, same frames (markdown to stdout).
    python tools/noalloc_timing.py > profiles/r2_noalloc_timing.md
The caller-owned buffers come from torch (standing in for an encoder's pool); timing is CUDA events
on the stream the calls are queued on."""
import ctypes
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from selkies_gstreamer_b200 import PreencodeSession, _abi, frame_layout, synth  # noqa: E402

lib = _abi.load()
print("| geometry | session path kernel us | sk_scale_frame_noalloc us (memset + tap upload + kernel) | ratio |")
print("|---|---|---|---|")
for (w, h, ow, oh, fmt) in [(2560, 1440, 1920, 1080, "I420"), (3840, 2160, 1920, 1080, "NV12"), (1920, 1080, 1920, 1080, "NV12"),
                            (3840, 2160, 3840, 2160, "NV12")]:
    f0 = synth.desktop(w, h)
    f1 = synth.mutate(f0, 0.05, seed=1, aligned=False)
    with PreencodeSession(w, h, out_width=ow, out_height=oh, out_format=fmt, colorimetry=1, chroma_site=1,
                          host_output=False, timing=True, spin_wait=True) as s:
        s.upload(0, f0); s.wait(s.submit(0))
        s.upload(1, f1); s.wait(s.submit(1))
        d = [s.device_input(k)[0] for k in (0, 1)]
        st = s.device_input(0)[1]
        ks = []
        for k in range(40):
            r = s.wait(s.submit_device(d[k & 1], st))
            if k >= 10:
                ks.append(r.kernel_us)
        sess_us = statistics.median(ks)
    dev = torch.device("cuda:0")
    nbytes = frame_layout(ow, oh, fmt)[0]
    din = [torch.from_numpy(f).to(dev) for f in (f0, f1)]
    dout = [torch.empty(nbytes, dtype=torch.uint8, device=dev) for _ in range(2)]
    dmap = torch.empty(((ow + 15) // 16) * ((oh + 15) // 16), dtype=torch.uint8, device=dev)
    dcnt = torch.zeros(1, dtype=torch.int32, device=dev)
    cfg = _abi.SkConfig(ctypes.sizeof(_abi.SkConfig), -1, w, h, 0, 0, ow, oh, {"NV12": 0, "I420": 1}[fmt], 1, 1, 1, 0, 0, 0)
    err = ctypes.create_string_buffer(256)
    stream = torch.cuda.current_stream().cuda_stream
    ts = []
    for k in range(40):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = lib.sk_scale_frame_noalloc(ctypes.byref(cfg), din[k & 1].data_ptr(), 0, dout[k & 1].data_ptr(),
                                        dout[(k & 1) ^ 1].data_ptr(), dmap.data_ptr(), dcnt.data_ptr(), stream, err, len(err))
        assert rc == 0, err.value
        b.record()
        torch.cuda.synchronize()
        if k >= 10:
            ts.append(a.elapsed_time(b) * 1e3)
    na_us = statistics.median(ts)
    print(f"| {w}x{h} -> {ow}x{oh} {fmt} | {sess_us:.1f} | {na_us:.1f} | {na_us / sess_us:.2f} |")
print("\nOne frame per call, B200, CUDA events; the session path runs the tiled (scaling) or the no-scale convert kernel,")
print("the allocation-free entry point always runs the generic gather kernel on caller-owned buffers.")
