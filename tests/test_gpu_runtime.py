"""GPU tests of the runtime around the kernels, all through the C ABI and all against the CPU oracle:
caller-owned output buffers, device-side consumers, host-buffer batches, the launch coalescer,
damage-gated read-back, and the exact shape bench.py's headline measures (32 x 4K in one launch)."""
import ctypes
import threading

import numpy as np
import pytest

from oracle import cpu_oracle as o
from selkies_gstreamer_b200 import (PreencodeError, PreencodeSession, SessionBatch, _abi, host_register, host_unregister,
                                   launch_count, synth)

pytestmark = pytest.mark.gpu


def rand_frame(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)


def expect(frame, fmt="NV12", matrix=1, site=1, ow=None, oh=None):
    return o.convert(frame, ow, oh, out_fmt={"NV12": 0, "I420": 1}[fmt], matrix=matrix, chroma_site=site, n_threads=8)


# ---- the benchmarked shape -------------------------------------------------------------------
def test_batch_of_32_4k_sessions_matches_the_oracle_for_every_member():
    """bench.py's headline step: ONE launch over 32 resident 4K sessions, two steps (the second one
    diffs against the first).  Frames, maps and counts of all 32 members are checked."""
    w, h, n = 3840, 2160, 32
    base = synth.desktop(w, h, seed=42)
    pairs = []
    for i in range(n):
        a = np.roll(base, shift=(37 * i, 101 * i), axis=(0, 1))
        pairs.append((a, synth.mutate(a, 0.05, seed=i, aligned=False)))
    sessions = [PreencodeSession(w, h, out_format="NV12", colorimetry="bt709", chroma_site="mpeg2", host_output=False)
                for _ in range(n)]
    try:
        for s, (a, b) in zip(sessions, pairs):
            s.upload(0, a); s.wait(s.submit(0))
            s.upload(1, b); s.wait(s.submit(1))
        batch = SessionBatch(sessions)
        batch.prepare()
        sets = [batch.frame_set([s.device_input(k)[0] for s in sessions]) for k in (0, 1)]
        stride = sessions[0].device_input(0)[1]
        before = launch_count()
        batch.submit(sets[0], stride, 100)
        batch.submit(sets[1], stride, 101)
        counts = batch.wait(101)
        assert launch_count() == before + 2
        buf = np.empty(sessions[0].frame_bytes, np.uint8)
        for i, (s, (a, b)) in enumerate(zip(sessions, pairs)):
            e0, e1 = expect(a), expect(b)
            assert (s.read_frame(100, buf) == e0).all(), f"member {i}, step 0"
            assert (s.read_frame(101, buf) == e1).all(), f"member {i}, step 1"
            em, en = o.damage(w, h, 0, e1, e0)
            r = s.wait(101)
            assert (r.damage_map == em).all() and r.damaged_count == en == counts[i], f"member {i}"
    finally:
        for s in sessions[1:]:
            s.close()
        sessions[0].close()


# ---- caller-owned buffers ----------------------------------------------------------------------
@pytest.mark.parametrize("pinned", [False, True])
def test_submit_io_writes_the_callers_output_buffer(pinned):
    w, h = 1920, 1080
    f0 = synth.desktop(w, h, seed=3)
    f1 = synth.mutate(f0, 0.02, seed=5, aligned=False)
    with PreencodeSession(w, h, out_format="I420", colorimetry=1, chroma_site=1, host_output=False) as s:
        outs = [np.full(s.frame_bytes + 64, 0xAB, np.uint8) for _ in range(2)]
        ins = [f0.copy(), f1.copy()]
        if pinned:
            for a in ins + outs:
                host_register(a)
        try:
            r0 = s.wait(s.submit_io(0, ins[0].ctypes.data, 0, outs[0].ctypes.data))
            r1 = s.wait(s.submit_io(1, ins[1].ctypes.data, 4 * w, outs[1].ctypes.data))
            e0, e1 = expect(f0, "I420"), expect(f1, "I420")
            assert r0.host_frame_ptr == outs[0].ctypes.data and r1.host_frame_ptr == outs[1].ctypes.data
            assert (outs[0][:s.frame_bytes] == e0).all() and (outs[1][:s.frame_bytes] == e1).all()
            assert (outs[0][s.frame_bytes:] == 0xAB).all()            # nothing written past frame_bytes
            assert (r1.frame == e1).all() and r1.readback_bytes == s.frame_bytes
            em, en = o.damage(w, h, 1, e1, e0)
            assert (r1.damage_map == em).all() and r1.damaged_count == en
            # out_ptr None: no frame comes back at all (host_output is off), maps still do
            r2 = s.wait(s.submit_io(0, ins[0].ctypes.data, 0, None))
            assert r2.frame is None and r2.damaged_count == en and r2.readback_bytes == 0
        finally:
            if pinned:
                for a in ins + outs:
                    host_unregister(a)


@pytest.mark.parametrize("spin", [False, True])
def test_wait_sleeps_by_default_and_can_spin(spin):
    f = rand_frame(640, 360, 9)
    with PreencodeSession(640, 360, spin_wait=spin) as s:
        for k in range(6):
            r = s.convert(f, slot=k & 1)
        assert (r.frame == o.convert(f, matrix=s.matrix, chroma_site=s.site)).all() and r.damaged_count == 0


def test_strides_that_overflow_32_bit_offsets_are_refused():
    with PreencodeSession(64, 4096) as s:
        d, st = s.device_input(0)
        with pytest.raises(PreencodeError) as e:
            s.submit_device(d, 1 << 20)
        assert e.value.code == _abi.SK_ERR_ARG and "4 GiB" in str(e.value)


# ---- device-side consumers (NVENC branch) --------------------------------------------------------
class _RawDeviceBytes:
    """A raw device address dressed up for torch.as_tensor (CUDA array interface v2: no stream sync)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def _device_view(torch, ptr, nbytes):
    return torch.as_tensor(_RawDeviceBytes(ptr, nbytes), device="cuda")


def test_wait_device_chains_a_consumer_stream_without_host_sync():
    """A consumer on its own CUDA stream waits for the frame with sk_wait_device (no host sync, no frame
    D2H: SK_FLAG_HOST_OUTPUT is off), reads dev_frame / dev_damage_map / dev_damaged_count there, and
    releases the ring entry with sk_release_device.  What it saw must be the oracle's frame."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    w, h = 1920, 1080
    frames = [synth.desktop(w, h, seed=1)]
    for k in range(1, 6):
        frames.append(synth.mutate(frames[-1], 0.03, seed=k, aligned=False))
    consumer = torch.cuda.Stream()
    with PreencodeSession(w, h, out_format="NV12", colorimetry=1, chroma_site=1, host_output=False, n_inflight=2) as s:
        nbytes = s.frame_bytes
        mbw, mbh = (w + 15) // 16, (h + 15) // 16
        copies, maps, cnts = [], [], []
        for k, f in enumerate(frames):
            fid = s.submit_host(k & 1, f)
            r = s.wait_device(fid, consumer.cuda_stream)
            assert r.frame is None and r.damage_map is None and r.dev_frame and r.dev_damaged_count
            with torch.cuda.stream(consumer):
                # the "encoder": reads the surface, the map and the count on ITS stream (device to device)
                dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
                dmap = torch.empty(mbw * mbh, dtype=torch.uint8, device="cuda")
                dcnt = torch.empty(4, dtype=torch.uint8, device="cuda")
                dst.copy_(_device_view(torch, r.dev_frame, nbytes), non_blocking=True)
                dmap.copy_(_device_view(torch, r.dev_damage_map, mbw * mbh), non_blocking=True)
                dcnt.copy_(_device_view(torch, r.dev_damaged_count, 4), non_blocking=True)
            s.release_device(fid, consumer.cuda_stream)
            copies.append(dst); maps.append(dmap); cnts.append(dcnt)
            # the host never waits here: with n_inflight = 2 the ring wraps at k = 2, and only the
            # release keeps frame k's surface intact until the consumer's copy has run
        consumer.synchronize()
        prev = None
        for k, f in enumerate(frames):
            e = expect(f)
            assert (copies[k].cpu().numpy() == e).all(), f"frame {k}"
            em, en = o.damage(w, h, 0, e, prev)
            assert (maps[k].cpu().numpy().reshape(mbh, mbw) == em).all() and int(cnts[k].cpu().numpy().view(np.uint32)[0]) == en, f"frame {k}"
            prev = e


# ---- host-buffer batches and the coalescer ------------------------------------------------------
def test_batch_submit_host_with_changing_membership():
    w, h, n = 1280, 720, 4
    sessions = [PreencodeSession(w, h, out_format="NV12", colorimetry=1, chroma_site=1, host_output=False) for _ in range(n)]
    try:
        cur = [synth.desktop(w, h, seed=20 + i) for i in range(n)]
        prev = [None] * n
        outs = [np.empty(sessions[0].frame_bytes, np.uint8) for _ in range(n)]
        lib = _abi.load()
        groups = [[0, 1, 2, 3], [1, 3], [2, 0, 1], [3], [0, 1, 2, 3], [3, 2]]       # leader and members change every step
        for step, g in enumerate(groups):
            for i in g:
                cur[i] = synth.mutate(cur[i], 0.02, seed=100 * step + i, aligned=False) if prev[i] is not None else cur[i]
            handles = (ctypes.c_void_p * len(g))(*[sessions[i].handle.value for i in g])
            ins = (ctypes.c_void_p * len(g))(*[cur[i].ctypes.data for i in g])
            ops = (ctypes.c_void_p * len(g))(*[outs[i].ctypes.data for i in g])
            before = launch_count()
            rc = lib.sk_batch_submit_host(handles, len(g), ins, 0, ops, 1000 + step)
            assert rc == 0, lib.sk_last_error(sessions[g[0]].handle)
            assert launch_count() == before + 1
            for i in g:
                r = sessions[i].wait(1000 + step)
                e = expect(cur[i])
                assert (outs[i] == e).all(), (step, i)
                em, en = o.damage(w, h, 0, e, prev[i])
                assert (r.damage_map == em).all() and r.damaged_count == en, (step, i)
                prev[i] = e
        # a member used alone in between keeps its history straight
        f = synth.mutate(cur[2], 0.01, seed=999, aligned=False)
        r = sessions[2].wait(sessions[2].submit_io(0, f.ctypes.data, 0, outs[2].ctypes.data))
        e = expect(f)
        em, en = o.damage(w, h, 0, e, prev[2])
        assert (outs[2] == e).all() and (r.damage_map == em).all() and r.damaged_count == en
    finally:
        for s in sessions:
            s.close()


@pytest.mark.parametrize("coalesce", [False, True])
def test_stages_on_threads_with_pool_buffers(coalesce):
    """What bench.py's e2e arm does: E stage instances on E host threads, recycled pageable buffers
    page-locked on first sight, output written straight into the caller's buffer."""
    from selkies_gstreamer_b200 import stage as st
    w, h, E, steps, pool = 1280, 720, 3, 9, 3
    stages, seqs, ins, outs = [], [], [], []
    inc = {"name": "video/x-raw", "format": "BGRx", "width": w, "height": h}
    outc = {"name": "video/x-raw", "format": "NV12", "width": w, "height": h, "colorimetry": "bt709", "chroma-site": "mpeg2"}
    for i in range(E):
        s = st.PreEncodeStage(f"t{i}")
        s.set_property("coalesce", coalesce)
        s.set_property("coalesce-window-us", 2000)
        assert s.start() and s.set_caps(inc, outc), s.last_error
        stages.append(s)
        fr = [synth.desktop(w, h, seed=50 + i)]
        for k in range(1, steps):
            fr.append(synth.mutate(fr[-1], 0.02, seed=10 * i + k, aligned=False))
        seqs.append(fr)
        ins.append([np.empty((h, w, 4), np.uint8) for _ in range(pool)])
        outs.append([np.empty(s.session.frame_bytes, np.uint8) for _ in range(pool)])
    results = [[] for _ in range(E)]
    barrier = threading.Barrier(E)

    def worker(i):
        s = stages[i]
        for k in range(steps):
            src, dst = ins[i][k % pool], outs[i][k % pool]
            src[...] = seqs[i][k]                        # "capture" into the recycled buffer
            barrier.wait()
            rc = s.transform(src.ctypes.data, src.nbytes, dst.ctypes.data, dst.nbytes)
            results[i].append((rc, dst.copy(), s.last_damage_map.copy(), s.last_damaged))

    co = stages[0]._coalescer
    co0 = (co.frames, co.launches) if co else None
    before = launch_count()
    ths = [threading.Thread(target=worker, args=(i,)) for i in range(E)]
    [t.start() for t in ths]
    [t.join() for t in ths]
    launches = launch_count() - before
    for i in range(E):
        prev = None
        assert len(results[i]) == steps
        for k, (rc, frame, dmap, cnt) in enumerate(results[i]):
            assert rc == st.FLOW_OK, stages[i].last_error
            e = expect(seqs[i][k])
            assert (frame == e).all(), (i, k)
            em, en = o.damage(w, h, 0, e, prev)
            assert (dmap == em).all() and cnt == en, (i, k)
            prev = e
        assert len(stages[i]._pins) == 2 * pool and stages[i]._pins.hits >= 2 * (steps - pool)
    if coalesce:
        assert launches < E * steps, "no launch was shared"
        assert co.frames - co0[0] == E * steps and co.launches - co0[1] == launches
    else:
        assert launches == E * steps
    for s in stages:
        assert s.stop()
        assert len(s._pins) == 0


# ---- damage-gated read-back ------------------------------------------------------------------------
@pytest.mark.parametrize("w,h,fmt,n_inflight", [(1920, 1080, "NV12", 2), (1280, 720, "I420", 3), (1366, 768, "NV12", 2),
                                                (258, 50, "I420", 2)])
def test_gated_readback_keeps_the_host_frame_exact(w, h, fmt, n_inflight):
    """Only macroblocks that changed since a host frame was last refreshed cross PCIe; the host frame
    must still equal the oracle's whole frame every time, whichever ring entry it lives in."""
    ofmt = {"NV12": 0, "I420": 1}[fmt]
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    cur = synth.desktop(w, h, seed=8)
    with PreencodeSession(w, h, out_format=fmt, colorimetry=1, chroma_site=1, host_output=True, gated_readback=True,
                          n_inflight=n_inflight) as s:
        maps = []
        for k in range(4 * n_inflight + 3):
            if k == 5:
                cur = synth.noise(w, h, seed=k)                                  # everything changes
            elif k == 7:
                pass                                                           # nothing changes
            elif k:
                cur = synth.mutate(cur, 0.02, seed=k, aligned=(k % 2 == 0))
            r = s.convert(cur, slot=k % n_inflight)
            e = expect(cur, fmt)
            assert (r.frame == e).all(), f"frame {k}"
            maps.append(r.damage_map.copy())
            # bytes that crossed = blocks damaged in any of the frames this host copy had missed
            miss = np.zeros((mbh, mbw), bool)
            for m in maps[-n_inflight:]:
                miss |= m.astype(bool)
            if k < n_inflight:
                miss[...] = True                                                # first use of each host frame
            assert r.readback_bytes == int(miss.sum()) * 384, f"frame {k}"
        assert maps[7].sum() == 0 and maps[5].all()
        # history reset: everything is damaged, everything comes back
        s.reset_history()
        r = s.convert(cur, slot=0)
        assert r.damage_map.all() and r.readback_bytes == mbw * mbh * 384 and (r.frame == expect(cur, fmt)).all()
        with pytest.raises(PreencodeError) as e:
            s.submit_io(0, cur.ctypes.data, 0, np.empty(s.frame_bytes, np.uint8).ctypes.data)
        assert e.value.code == _abi.SK_ERR_UNSUPPORTED


def test_pcie_probe_reports_rates():
    lib = _abi.load()
    up, down = ctypes.c_double(), ctypes.c_double()
    assert lib.sk_pcie_probe(0, 8 << 20, 4, 2, ctypes.byref(up), ctypes.byref(down)) == 0
    assert up.value > 1.0 and down.value > 1.0
    assert lib.sk_pcie_probe(0, 8 << 20, 4, 0, ctypes.byref(up), ctypes.byref(down)) == 0
    assert up.value > 1.0 and down.value == 0.0


def test_stage_with_cuda_memory_caps_keeps_the_frame_on_the_gpu():
    """src caps video/x-raw(memory:CUDAMemory) (the reference's NVENC branches): the stage converts, only the maps come
    back, a consumer stream reads the surface in HBM — and it is the oracle's frame."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from selkies_gstreamer_b200 import stage as st
    w, h = 1920, 1080
    s = st.PreEncodeStage()
    s.set_property("cuda-memory-output", True)
    sink = st.parse_caps(f"video/x-raw, format=(string)BGRx, width=(int){w}, height=(int){h}, framerate=(fraction)60/1")
    other = s.transform_caps(st.SINK, sink, st.parse_caps("video/x-raw(memory:CUDAMemory), format=(string)NV12, colorimetry=(string)bt709, chroma-site=(string)mpeg2"))
    fixed = s.fixate_caps(st.SINK, sink, other)
    assert s.start() and s.set_caps(sink[0], fixed), s.last_error
    assert s.device_output and s.session.flags & _abi.SK_FLAG_HOST_OUTPUT == 0
    consumer = torch.cuda.Stream()
    frames = [synth.desktop(w, h, seed=2)]
    frames.append(synth.mutate(frames[0], 0.03, seed=1, aligned=False))
    frames.append(synth.mutate(frames[1], 0.03, seed=2, aligned=False))
    seen, prev = [], None
    for f in frames:
        rc, r = s.transform_device(f.ctypes.data, f.nbytes, consumer.cuda_stream, pin=False)
        assert rc == st.FLOW_OK, s.last_error
        assert r.frame is None and r.dev_frame and r.readback_bytes == 0
        with torch.cuda.stream(consumer):
            dst = torch.empty(s.session.frame_bytes, dtype=torch.uint8, device="cuda")
            dst.copy_(_device_view(torch, r.dev_frame, s.session.frame_bytes), non_blocking=True)
        s.release_device(r.frame_id, consumer.cuda_stream)
        seen.append((dst, s.last_damage_map.copy(), s.last_damaged))
    consumer.synchronize()
    for k, (f, (dst, dmap, cnt)) in enumerate(zip(frames, seen)):
        e = expect(f)
        assert (dst.cpu().numpy() == e).all(), k
        em, en = o.damage(w, h, 0, e, prev)
        assert (dmap == em).all() and cnt == en, k
        prev = e
    assert s.stop()


# ---- resources come back ---------------------------------------------------------------------------
def test_sessions_give_their_device_memory_back():
    """Create, use and destroy sessions of every flavour many times over: the GPU's free memory must return to
    where it was (first cycle excluded: context, modules and the tensor-map encoder load lazily)."""
    torch = pytest.importorskip("torch")

    def cycle(seed):
        rng = np.random.default_rng(seed)
        flavours = [
            dict(w=640, h=360, kw={}),
            dict(w=1280, h=720, kw=dict(out_width=854, out_height=480, out_format="I420")),
            dict(w=640, h=360, kw=dict(out_width=1280, out_height=720)),                       # growing picture
            dict(w=1920, h=1080, kw=dict(gated_readback=True, n_inflight=3)),
            dict(w=512, h=288, kw=dict(policy=True, out_format="I420")),
            dict(w=800, h=600, kw=dict(host_output=False, scale_generic=True, out_width=400, out_height=300)),
        ]
        for fl in flavours:
            f = rng.integers(0, 256, (fl["h"], fl["w"], 4), dtype=np.uint8)
            with PreencodeSession(fl["w"], fl["h"], colorimetry=1, chroma_site=1, **fl["kw"]) as s:
                r = s.convert(f)
                assert r.damaged_count > 0
        members = [PreencodeSession(640, 360, colorimetry=1, chroma_site=1) for _ in range(4)]
        try:
            b = SessionBatch(members)
            b.prepare()
            frames = [rng.integers(0, 256, (360, 640, 4), dtype=np.uint8) for _ in members]
            outs = [np.empty(m.frame_bytes, np.uint8) for m in members]
            b.submit_host([f.ctypes.data for f in frames], 4 * 640, [o_.ctypes.data for o_ in outs], 1)
            b.wait(1)
            for m in members:
                m.wait(1)
        finally:
            for m in members:
                m.close()

    cycle(0)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for k in range(1, 13):
        cycle(k)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free0 - free1 < (4 << 20), f"{(free0 - free1) / 2**20:.1f} MiB of device memory did not come back"
