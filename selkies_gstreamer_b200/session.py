"""Per-desktop session: the Python face of one sk_session.

One session is what one instance of the reference's convert stage is
(src/selkies/legacy/gstwebrtc_app.py:611-617 videoconvert / :263-284
cudaupload+cudaconvert): fixed caps, frames in, NV12/I420 frames out — plus the
per-16x16 damage map.  All pixel work happens in libselkies_preencode.so on the
GPU; this file only moves pointers.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import numpy as np

from . import _abi, colorimetry as _col

IN_FORMATS = {"BGRX": _abi.SK_IN_BGRX, "BGRA": _abi.SK_IN_BGRX, "RGBX": _abi.SK_IN_RGBX, "RGBA": _abi.SK_IN_RGBX}
OUT_FORMATS = {"NV12": _abi.SK_OUT_NV12, "I420": _abi.SK_OUT_I420}
_ERR_NAMES = {-1: "SK_ERR_ARG", -2: "SK_ERR_CUDA", -3: "SK_ERR_ALIGN", -4: "SK_ERR_STATE",
              -5: "SK_ERR_UNSUPPORTED", -6: "SK_ERR_NOMEM"}


class PreencodeError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"{_ERR_NAMES.get(code, code)}: {message}")
        self.code = code


def frame_layout(width: int, height: int, out_format: str):
    """GstVideoInfo default layout: (size, offsets[3], strides[3])."""
    ru = lambda v, a: (v + a - 1) // a * a
    h2 = ru(height, 2)
    ys = ru(width, 4)
    if out_format.upper() == "I420":
        cs = ru(ru(width, 2) // 2, 4)
        off = [0, ys * h2, ys * h2 + cs * (h2 // 2)]
        return off[2] + cs * (h2 // 2), off, [ys, cs, cs]
    off = [0, ys * h2, 0]
    return off[1] + ys * (h2 // 2), off, [ys, ys, 0]


@dataclass
class FrameResult:
    frame_id: int
    frame: np.ndarray | None        # packed output frame (pinned host view) or None
    damage_map: np.ndarray | None   # uint8 [mb_h, mb_w]
    damaged_count: int
    dev_frame: int                  # device address of the packed frame
    dev_damage_map: int
    plane_offset: tuple
    plane_stride: tuple
    h2d_us: float
    kernel_us: float
    d2h_us: float
    policy_map: np.ndarray | None = None   # uint8 [mb_h, mb_w] of SK_POLICY_* bits
    dev_policy_map: int = 0
    dev_damaged_count: int = 0      # device address of the uint32 count
    readback_bytes: int = 0         # frame bytes that crossed PCIe for `frame`
    host_frame_ptr: int = 0         # address of the host copy (library ring, or the caller's buffer of submit_io)


class PreencodeSession:
    def __init__(self, width: int, height: int, *, in_format: str = "BGRx", out_format: str = "NV12",
                 out_width: int | None = None, out_height: int | None = None,
                 colorimetry=None, chroma_site=None, device_id: int = -1, n_inflight: int = 2,
                 in_stride: int = 0, damage: bool = True, host_output: bool = True, timing: bool = False,
                 scale_variant: int = _abi.SK_SCALE_VARIANT_GST_LQ, policy: bool = False, damage_block_threshold: int = 0,
                 damage_block_duration: int = 0, paint_over_trigger_frames: int = 0, scale_generic: bool = False,
                 spin_wait: bool = False, gated_readback: bool = False, parity_switches: int = 0):
        self._lib = _abi.load()
        self._h = ctypes.c_void_p()
        if in_format.upper() not in IN_FORMATS:
            raise ValueError(f"unsupported input format {in_format!r}")
        if out_format.upper() not in OUT_FORMATS:
            raise ValueError(f"unsupported output format {out_format!r}")
        self.width, self.height = int(width), int(height)
        self.out_width = int(out_width or width)
        self.out_height = int(out_height or height)
        self.in_format, self.out_format = in_format, out_format.upper()
        self.matrix, self.site = _col.resolve(self.out_height, colorimetry, chroma_site)
        self.in_stride = int(in_stride) or 4 * self.width
        self.n_inflight = n_inflight
        flags = (0 if damage else _abi.SK_FLAG_NO_DAMAGE) | (_abi.SK_FLAG_HOST_OUTPUT if host_output else 0) \
            | (_abi.SK_FLAG_TIMING if timing else 0) | (_abi.SK_FLAG_POLICY if policy else 0) \
            | (_abi.SK_FLAG_SCALE_GENERIC if scale_generic else 0) | (_abi.SK_FLAG_SPIN_WAIT if spin_wait else 0) \
            | (_abi.SK_FLAG_GATED_READBACK if gated_readback else 0)
        self.flags = flags
        self.device_index = max(0, int(device_id))
        self.scale_variant = int(scale_variant)
        self.parity_switches = int(parity_switches)
        self.policy_knobs = (int(damage_block_threshold) or 10, int(damage_block_duration) or 20,
                             int(paint_over_trigger_frames) or 15)
        cfg = _abi.SkConfig(ctypes.sizeof(_abi.SkConfig), device_id, self.width, self.height, self.in_stride,
                            IN_FORMATS[in_format.upper()], self.out_width, self.out_height,
                            OUT_FORMATS[self.out_format], self.matrix, self.site, _abi.SK_SCALE_BILINEAR,
                            scale_variant, n_inflight, flags, damage_block_threshold, damage_block_duration,
                            paint_over_trigger_frames, parity_switches)
        rc = self._lib.sk_session_create(ctypes.byref(cfg), ctypes.byref(self._h))
        if rc != _abi.SK_OK:
            self._h = ctypes.c_void_p()
            raise PreencodeError(rc, (self._lib.sk_last_error(None) or b"").decode())
        self.frame_bytes, self.plane_offset, self.plane_stride = frame_layout(self.out_width, self.out_height, self.out_format)
        self._next_id = 0

    # -- plumbing ---------------------------------------------------------
    def _check(self, rc: int):
        if rc != _abi.SK_OK:
            raise PreencodeError(rc, (self._lib.sk_last_error(self._h) or b"").decode())

    @property
    def handle(self):
        return self._h

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.sk_session_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- frame path -------------------------------------------------------
    def staging(self, slot: int) -> np.ndarray:
        """Pinned staging of `slot` as uint8 [H, in_stride] — capture straight into it."""
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        self._check(self._lib.sk_get_input_staging(self._h, slot, ctypes.byref(p), ctypes.byref(n)))
        buf = (ctypes.c_uint8 * n.value).from_address(p.value)
        return np.frombuffer(buf, dtype=np.uint8).reshape(self.height, self.in_stride)

    def device_input(self, slot: int) -> tuple[int, int]:
        p, st = ctypes.c_void_p(), ctypes.c_size_t()
        self._check(self._lib.sk_get_device_input(self._h, slot, ctypes.byref(p), ctypes.byref(st)))
        return p.value, st.value

    def upload(self, slot: int, frame: np.ndarray):
        """Copy a [H, W, 4] uint8 frame (row-strided views allowed) into staging."""
        if frame.dtype != np.uint8 or frame.ndim != 3 or frame.shape != (self.height, self.width, 4):
            raise ValueError(f"expected uint8 [{self.height}, {self.width}, 4], got {frame.dtype} {frame.shape}")
        if frame.strides[2] != 1 or frame.strides[1] != 4:
            frame = np.ascontiguousarray(frame)
        self._check(self._lib.sk_upload(self._h, slot, frame.ctypes.data, frame.strides[0]))

    def submit(self, slot: int, frame_id: int | None = None) -> int:
        fid = self._take_id(frame_id)
        self._check(self._lib.sk_submit(self._h, slot, fid))
        return fid

    def submit_host(self, slot: int, frame, stride: int = 0, frame_id: int | None = None) -> int:
        """H2D straight from `frame` (numpy [H, W, 4] uint8 or a raw address) — no staging memcpy.
        The buffer must stay alive and unchanged until wait(frame_id) returns."""
        if isinstance(frame, np.ndarray):
            if frame.dtype != np.uint8 or frame.shape != (self.height, self.width, 4) or frame.strides[1:] != (4, 1):
                raise ValueError("expected a uint8 [H, W, 4] array with contiguous pixels")
            ptr, stride = frame.ctypes.data, frame.strides[0]
        else:
            ptr = int(frame)
        fid = self._take_id(frame_id)
        self._check(self._lib.sk_submit_host(self._h, slot, ctypes.c_void_p(ptr), stride, fid))
        return fid

    def submit_io(self, slot: int, in_ptr: int, in_stride: int, out_ptr: int | None, frame_id: int | None = None) -> int:
        """sk_submit_io: H2D straight from the caller's input buffer and the converted frame copied back into
        the caller's output buffer (both raw addresses; pin recycled buffers with host_register)."""
        fid = self._take_id(frame_id)
        self._check(self._lib.sk_submit_io(self._h, slot, ctypes.c_void_p(in_ptr), in_stride,
                                           ctypes.c_void_p(out_ptr) if out_ptr else None, fid))
        return fid

    def submit_device(self, dptr: int, stride: int = 0, frame_id: int | None = None) -> int:
        fid = self._take_id(frame_id)
        self._check(self._lib.sk_submit_device(self._h, ctypes.c_void_p(dptr), stride, fid))
        return fid

    def _take_id(self, frame_id):
        if frame_id is None:
            frame_id = self._next_id
        self._next_id = int(frame_id) + 1
        return int(frame_id)

    def _result(self, r) -> FrameResult:
        frame = dmap = pmap = None
        if r.host_frame:
            buf = (ctypes.c_uint8 * r.frame_bytes).from_address(r.host_frame)
            frame = np.frombuffer(buf, dtype=np.uint8)
        if r.host_damage_map:
            buf = (ctypes.c_uint8 * (r.mb_w * r.mb_h)).from_address(r.host_damage_map)
            dmap = np.frombuffer(buf, dtype=np.uint8).reshape(r.mb_h, r.mb_w)
        if r.host_policy_map:
            buf = (ctypes.c_uint8 * (r.mb_w * r.mb_h)).from_address(r.host_policy_map)
            pmap = np.frombuffer(buf, dtype=np.uint8).reshape(r.mb_h, r.mb_w)
        return FrameResult(r.frame_id, frame, dmap, r.damaged_count, r.dev_frame or 0, r.dev_damage_map or 0,
                           tuple(r.plane_offset), tuple(r.plane_stride), r.h2d_us, r.kernel_us, r.d2h_us, pmap,
                           r.dev_policy_map or 0, r.dev_damaged_count or 0, int(r.readback_bytes), r.host_frame or 0)

    def wait(self, frame_id: int) -> FrameResult:
        """Block until `frame_id` is done. `frame` and `damage_map` are BORROWED views of the
        library's pinned buffers (or of the caller's output buffer): valid until n_inflight more
        frames are submitted or the session is closed — copy them if they must outlive that."""
        r = _abi.SkResult()
        self._check(self._lib.sk_wait(self._h, frame_id, ctypes.byref(r)))
        return self._result(r)

    def read_frame(self, frame_id: int, out: np.ndarray | None = None) -> np.ndarray:
        """sk_read_frame: blocking copy of the converted frame of `frame_id` out of HBM."""
        if out is None:
            out = np.empty(self.frame_bytes, np.uint8)
        if out.dtype != np.uint8 or out.nbytes < self.frame_bytes or not out.flags.c_contiguous:
            raise ValueError("need a contiguous uint8 buffer of frame_bytes")
        self._check(self._lib.sk_read_frame(self._h, frame_id, ctypes.c_void_p(out.ctypes.data)))
        return out

    def wait_device(self, frame_id: int, cuda_stream: int) -> FrameResult:
        """sk_wait_device: make `cuda_stream` (a cudaStream_t as an integer) wait for the kernel of
        `frame_id`; returns at once with the device pointers only."""
        r = _abi.SkResult()
        self._check(self._lib.sk_wait_device(self._h, frame_id, ctypes.c_void_p(cuda_stream), ctypes.byref(r)))
        return self._result(r)

    def release_device(self, frame_id: int, cuda_stream: int):
        """sk_release_device: the ring entry of `frame_id` is not overwritten before the work queued on
        `cuda_stream` so far has finished."""
        self._check(self._lib.sk_release_device(self._h, frame_id, ctypes.c_void_p(cuda_stream)))

    def convert(self, frame: np.ndarray, slot: int = 0, copy: bool = False) -> FrameResult:
        """Synchronous convenience: upload → submit → wait (see wait() about borrowed views)."""
        self.upload(slot, frame)
        r = self.wait(self.submit(slot))
        if copy:
            r.frame = None if r.frame is None else r.frame.copy()
            r.damage_map = None if r.damage_map is None else r.damage_map.copy()
            r.policy_map = None if r.policy_map is None else r.policy_map.copy()
        return r

    def reset_history(self):
        self._check(self._lib.sk_reset_history(self._h))

    def marker(self, which: int):
        self._check(self._lib.sk_marker_record(self._h, which))

    def marker_elapsed_us(self) -> float:
        us = ctypes.c_float()
        self._check(self._lib.sk_marker_elapsed_us(self._h, ctypes.byref(us)))
        return us.value

    def last_kernel_us(self) -> float:
        us = ctypes.c_float()
        self._check(self._lib.sk_last_kernel_us(self._h, ctypes.byref(us)))
        return us.value

    def planes(self, packed: np.ndarray):
        """Views (Y, U, V) of a packed output frame; for NV12 U and V are strided views of UV."""
        w, h = self.out_width, self.out_height
        h2, w2 = (h + 1) & ~1, (w + 1) // 2
        off, st = self.plane_offset, self.plane_stride
        y = packed[off[0]:off[0] + st[0] * h2].reshape(h2, st[0])[:h, :w]
        if self.out_format == "I420":
            u = packed[off[1]:off[1] + st[1] * (h2 // 2)].reshape(h2 // 2, st[1])[:, :w2]
            v = packed[off[2]:off[2] + st[2] * (h2 // 2)].reshape(h2 // 2, st[2])[:, :w2]
            return y, u, v
        uv = packed[off[1]:off[1] + st[1] * (h2 // 2)].reshape(h2 // 2, st[1])[:, :2 * w2]
        return y, uv[:, 0::2], uv[:, 1::2]


def batch_submit(sessions: list[PreencodeSession], dptrs: list[int], stride: int = 0, frame_id: int = 0):
    """One kernel launch over several sessions of identical configuration (same GPU)."""
    n = len(sessions)
    if n == 0 or n != len(dptrs):
        raise ValueError("sessions and dptrs must be non-empty and of equal length")
    lib = sessions[0]._lib
    hs = (ctypes.c_void_p * n)(*[s._h.value for s in sessions])
    ps = (ctypes.c_void_p * n)(*dptrs)
    rc = lib.sk_batch_submit(hs, n, ps, stride, frame_id)
    if rc != _abi.SK_OK:
        raise PreencodeError(rc, (lib.sk_last_error(sessions[0]._h) or b"").decode())
    for s in sessions:
        s._next_id = frame_id + 1
    return frame_id


class SessionBatch:
    """A fixed group of same-configuration sessions on one GPU, submitted with one launch per
    frame (throughput mode: many desktops per GPU).  ctypes arrays are built once."""

    def __init__(self, sessions: list[PreencodeSession]):
        if not sessions:
            raise ValueError("empty batch")
        if len(sessions) > _abi.SK_MAX_BATCH:
            raise ValueError(f"at most {_abi.SK_MAX_BATCH} sessions per batch")
        self.sessions = list(sessions)
        self._lib = sessions[0]._lib
        self._n = len(sessions)
        self._handles = (ctypes.c_void_p * self._n)(*[s._h.value for s in sessions])
        self._counts = (ctypes.c_uint32 * self._n)()
        self._frames = {}

    def frame_set(self, dptrs: list[int]):
        """Pre-build the pointer array for one set of resident frames (one per session)."""
        if len(dptrs) != self._n:
            raise ValueError("one device pointer per session")
        return (ctypes.c_void_p * self._n)(*dptrs)

    def submit(self, frame_set, stride: int, frame_id: int) -> int:
        rc = self._lib.sk_batch_submit(self._handles, self._n, frame_set, stride, frame_id)
        if rc != _abi.SK_OK:
            raise PreencodeError(rc, (self._lib.sk_last_error(self.sessions[0]._h) or b"").decode())
        return frame_id

    def prepare(self):
        """sk_batch_prepare: allocate the leader's map blocks now, not at the first submit."""
        rc = self._lib.sk_batch_prepare(self._handles, self._n)
        if rc != _abi.SK_OK:
            raise PreencodeError(rc, (self._lib.sk_last_error(self.sessions[0]._h) or b"").decode())

    def submit_host(self, in_ptrs: list[int], in_stride: int, out_ptrs: list[int] | None, frame_id: int) -> int:
        """sk_batch_submit_host: frames still in host memory, one upload per member, ONE launch."""
        ins = (ctypes.c_void_p * self._n)(*in_ptrs)
        outs = (ctypes.c_void_p * self._n)(*out_ptrs) if out_ptrs is not None else None
        rc = self._lib.sk_batch_submit_host(self._handles, self._n, ins, in_stride, outs, frame_id)
        if rc != _abi.SK_OK:
            raise PreencodeError(rc, (self._lib.sk_last_error(self.sessions[0]._h) or b"").decode())
        for s in self.sessions:
            s._next_id = frame_id + 1
        return frame_id

    def wait(self, frame_id: int) -> list[int]:
        rc = self._lib.sk_batch_wait(self._handles, self._n, frame_id, self._counts)
        if rc != _abi.SK_OK:
            raise PreencodeError(rc, (self._lib.sk_last_error(self.sessions[0]._h) or b"").decode())
        return list(self._counts)


def host_register(array, nbytes: int | None = None):
    """Pin a caller-owned buffer (numpy array, or raw address + nbytes: a capture segment, a pool
    buffer) for submit_host / submit_io. Undo with host_unregister."""
    lib = _abi.load()
    ptr, n = (array.ctypes.data, array.nbytes) if isinstance(array, np.ndarray) else (int(array), int(nbytes))
    rc = lib.sk_host_register(ctypes.c_void_p(ptr), n)
    if rc != _abi.SK_OK:
        raise PreencodeError(rc, (lib.sk_last_error(None) or b"").decode())


def host_unregister(array):
    lib = _abi.load()
    ptr = array.ctypes.data if isinstance(array, np.ndarray) else int(array)
    rc = lib.sk_host_unregister(ctypes.c_void_p(ptr))
    if rc != _abi.SK_OK:
        raise PreencodeError(rc, (lib.sk_last_error(None) or b"").decode())


def launch_count() -> int:
    return int(_abi.load().sk_launch_count())


def device_count() -> int:
    return int(_abi.load().sk_device_count())
