#!/usr/bin/env python3
"""Run one geometry device-resident for a few steps (ncu target).
    python tools/run_config.py W H OW OH FMT BATCH STEPS [VARIANT]
VARIANT needs the tuning build: make -C selkies_gstreamer_b200/csrc tuning;
SELKIES_PREENCODE_LIB=selkies_gstreamer_b200/libselkies_preencode_tuning.so"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import selkies_gstreamer_b200 as sk  # noqa: E402
from selkies_gstreamer_b200 import synth  # noqa: E402

w, h, ow, oh = (int(x) for x in sys.argv[1:5])
fmt, batch, steps = sys.argv[5], int(sys.argv[6]), int(sys.argv[7])
if len(sys.argv) > 8:
    from selkies_gstreamer_b200 import _abi
    _abi.load().sk_tuning_set_variant(int(sys.argv[8]))
sess = [sk.PreencodeSession(w, h, out_width=ow, out_height=oh, out_format=fmt, colorimetry="bt709",
                            chroma_site="mpeg2", host_output=False, timing=(i == 0), spin_wait=True) for i in range(batch)]
base = synth.desktop(w, h)
nxt = synth.mutate(base, 0.05, seed=1, aligned=False)
for s in sess:
    s.upload(0, base); s.wait(s.submit(0))
    s.upload(1, nxt); s.wait(s.submit(1))
b = sk.SessionBatch(sess)
sets = [b.frame_set([s.device_input(k)[0] for s in sess]) for k in (0, 1)]
stride = sess[0].device_input(0)[1]
for k in range(steps):
    b.submit(sets[k & 1], stride, 100 + k)
    b.wait(100 + k)
import statistics
ks = []
for k in range(steps, steps + 30):
    b.submit(sets[k & 1], stride, 100 + k)
    b.wait(100 + k)
    ks.append(sess[0].last_kernel_us())
print("kernel_us", round(statistics.median(ks), 2))
