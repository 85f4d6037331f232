#!/usr/bin/env python3
"""Host time of sk_batch_submit / sk_batch_wait per step (32 x 4K), to see whether the host keeps up with the GPU."""
import os, sys, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import selkies_gstreamer_b200 as sk
from selkies_gstreamer_b200 import synth
S = int(sys.argv[1]) if len(sys.argv) > 1 else 32
W, H = 3840, 2160
sess = [sk.PreencodeSession(W, H, host_output=False, colorimetry="bt709", chroma_site="mpeg2", timing=(i == 0)) for i in range(S)]
base = synth.desktop(W, H); nxt = synth.mutate(base, 0.05, seed=1, aligned=False)
for s in sess:
    s.upload(0, base); s.wait(s.submit(0)); s.upload(1, nxt); s.wait(s.submit(1))
b = sk.SessionBatch(sess)
sets = [b.frame_set([s.device_input(k)[0] for s in sess]) for k in (0, 1)]
stride = sess[0].device_input(0)[1]
ts, tw = [], []
fid = 100
for k in range(300):
    t0 = time.perf_counter(); b.submit(sets[k & 1], stride, fid); t1 = time.perf_counter()
    if k: b.wait(fid - 1)
    t2 = time.perf_counter()
    ts.append(t1 - t0); tw.append(t2 - t1); fid += 1
b.wait(fid - 1)
print(f"submit host us median {statistics.median(ts)*1e6:.1f}  wait (incl. blocking) us median {statistics.median(tw)*1e6:.1f}  kernel us {sess[0].last_kernel_us():.1f}")
# pure host cost of wait when everything is already complete
time.sleep(0.05)
t0 = time.perf_counter()
for _ in range(50): b.wait(fid - 1)
print(f"wait on a finished frame: {(time.perf_counter()-t0)/50*1e6:.1f} us")
