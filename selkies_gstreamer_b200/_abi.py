"""ctypes declarations for libselkies_preencode.so (include/selkies_preencode.h).

No PyTorch, no CPU fallback: if the library is missing or has no CUDA device to
run on, callers get an exception, never a silently different code path.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
# SELKIES_PREENCODE_LIB selects another build of the same library (A/B measurements); never a fallback.
LIB_PATH = os.environ.get("SELKIES_PREENCODE_LIB") or os.path.join(_PKG, "libselkies_preencode.so")
CSRC = os.path.join(_PKG, "csrc")

SK_OK = 0
SK_ERR_ARG, SK_ERR_CUDA, SK_ERR_ALIGN, SK_ERR_STATE, SK_ERR_UNSUPPORTED, SK_ERR_NOMEM = -1, -2, -3, -4, -5, -6
SK_IN_BGRX, SK_IN_RGBX = 0, 1
SK_OUT_NV12, SK_OUT_I420 = 0, 1
SK_MATRIX_BT601, SK_MATRIX_BT709, SK_MATRIX_BT2020 = 0, 1, 2
SK_SITE_NONE, SK_SITE_H_COSITED = 0, 1
SK_SCALE_BILINEAR = 1
SK_SCALE_VARIANT_GST_LQ, SK_SCALE_VARIANT_PLAIN_LQ = 0, 1
SK_FLAG_NO_DAMAGE, SK_FLAG_HOST_OUTPUT, SK_FLAG_TIMING, SK_FLAG_POLICY = 1, 2, 4, 8
SK_FLAG_SCALE_GENERIC, SK_FLAG_SPIN_WAIT, SK_FLAG_GATED_READBACK = 16, 32, 64
SK_PS_MATRIX_ROUND, SK_PS_COEFF_HALF_UP, SK_PS_CS_EDGE_KEEP, SK_PS_CHROMA_V_FIRST = 1, 2, 4, 8
SK_PS_CHROMA_BOX4, SK_PS_HINC_PLAIN, SK_PS_V_FIRST_ON_TIE = 16, 32, 64
SK_MAX_BATCH = 32
SK_ABI_VERSION = 3
SK_POLICY_DAMAGED, SK_POLICY_HOT, SK_POLICY_PAINT_OVER = 1, 2, 4
SK_NOALLOC_MAX_OUT_HEIGHT = 4320

# Every symbol include/selkies_preencode.h declares (tests check the .so exports all of them).
EXPORTS = (
    "sk_session_create", "sk_session_destroy", "sk_get_input_staging", "sk_get_device_input",
    "sk_upload", "sk_submit", "sk_submit_host", "sk_submit_io", "sk_host_register", "sk_host_unregister", "sk_submit_device",
    "sk_wait", "sk_read_frame", "sk_wait_device", "sk_release_device", "sk_reset_history",
    "sk_batch_prepare", "sk_batch_submit", "sk_batch_submit_host", "sk_batch_wait",
    "sk_last_kernel_us", "sk_marker_record", "sk_marker_elapsed_us", "sk_pcie_probe",
    "sk_scale_frame_noalloc", "sk_launch_count", "sk_last_error", "sk_version", "sk_device_count",
)


class SkConfig(ctypes.Structure):
    _fields_ = [
        ("struct_size", ctypes.c_uint32),
        ("device_id", ctypes.c_int32),
        ("in_width", ctypes.c_int32), ("in_height", ctypes.c_int32),
        ("in_stride", ctypes.c_int32),
        ("in_format", ctypes.c_int32),
        ("out_width", ctypes.c_int32), ("out_height", ctypes.c_int32),
        ("out_format", ctypes.c_int32),
        ("matrix", ctypes.c_int32),
        ("chroma_site", ctypes.c_int32),
        ("scale_method", ctypes.c_int32),
        ("scale_variant", ctypes.c_int32),
        ("n_inflight", ctypes.c_int32),
        ("flags", ctypes.c_uint32),
        ("damage_block_threshold", ctypes.c_int32),
        ("damage_block_duration", ctypes.c_int32),
        ("paint_over_trigger_frames", ctypes.c_int32),
        ("parity_switches", ctypes.c_uint32),
    ]


class SkResult(ctypes.Structure):
    _fields_ = [
        ("frame_id", ctypes.c_uint64),
        ("dev_frame", ctypes.c_void_p),
        ("host_frame", ctypes.c_void_p),
        ("frame_bytes", ctypes.c_size_t),
        ("plane_offset", ctypes.c_size_t * 3),
        ("plane_stride", ctypes.c_int32 * 3),
        ("host_damage_map", ctypes.c_void_p),
        ("dev_damage_map", ctypes.c_void_p),
        ("mb_w", ctypes.c_int32), ("mb_h", ctypes.c_int32),
        ("damaged_count", ctypes.c_uint32),
        ("h2d_us", ctypes.c_float), ("kernel_us", ctypes.c_float), ("d2h_us", ctypes.c_float),
        ("host_policy_map", ctypes.c_void_p),
        ("dev_policy_map", ctypes.c_void_p),
        ("dev_damaged_count", ctypes.c_void_p),
        ("readback_bytes", ctypes.c_uint64),
    ]


class PreencodeLibraryError(RuntimeError):
    """The native library is missing, stale or unusable. There is no fallback."""


def build(force: bool = False) -> str:
    """Compile the library in-tree with nvcc for sm_100a (seconds; works without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in ("sk_kernels.cu", "sk_session.cu", "sk_noalloc.cu", "sk_kernels.h", "sk_host_math.h")]
    srcs.append(os.path.join(_PKG, "..", "include", "selkies_preencode.h"))
    stale = (not os.path.exists(LIB_PATH)
             or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(s) for s in srcs))
    if force or stale:
        subprocess.check_call(["make", "-C", CSRC, "-B"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def load() -> ctypes.CDLL:
    """dlopen the library and declare every prototype. Raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PreencodeLibraryError(
            f"{LIB_PATH} not found: build it with `make -C {CSRC}` (or __graft_entry__.build()). "
            "This package has no CPU or PyTorch fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, vpp = ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)
    i, u64, sz = ctypes.c_int, ctypes.c_uint64, ctypes.c_size_t
    fp = ctypes.POINTER(ctypes.c_float)
    protos = {
        "sk_session_create": (i, [ctypes.POINTER(SkConfig), vpp]),
        "sk_session_destroy": (i, [vp]),
        "sk_get_input_staging": (i, [vp, i, vpp, ctypes.POINTER(sz)]),
        "sk_get_device_input": (i, [vp, i, vpp, ctypes.POINTER(sz)]),
        "sk_upload": (i, [vp, i, vp, sz]),
        "sk_submit": (i, [vp, i, u64]),
        "sk_submit_host": (i, [vp, i, vp, sz, u64]),
        "sk_submit_io": (i, [vp, i, vp, sz, vp, u64]),
        "sk_host_register": (i, [vp, sz]),
        "sk_host_unregister": (i, [vp]),
        "sk_submit_device": (i, [vp, vp, sz, u64]),
        "sk_wait": (i, [vp, u64, ctypes.POINTER(SkResult)]),
        "sk_read_frame": (i, [vp, u64, vp]),
        "sk_wait_device": (i, [vp, u64, vp, ctypes.POINTER(SkResult)]),
        "sk_release_device": (i, [vp, u64, vp]),
        "sk_reset_history": (i, [vp]),
        "sk_batch_prepare": (i, [vpp, i]),
        "sk_batch_submit": (i, [vpp, i, vpp, sz, u64]),
        "sk_batch_submit_host": (i, [vpp, i, vpp, sz, vpp, u64]),
        "sk_batch_wait": (i, [vpp, i, u64, ctypes.POINTER(ctypes.c_uint32)]),
        "sk_last_kernel_us": (i, [vp, fp]),
        "sk_marker_record": (i, [vp, i]),
        "sk_marker_elapsed_us": (i, [vp, fp]),
        "sk_pcie_probe": (i, [i, sz, i, i, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]),
        "sk_scale_frame_noalloc": (i, [ctypes.POINTER(SkConfig), vp, sz, vp, vp, vp, vp, vp, ctypes.c_char_p, sz]),
        "sk_launch_count": (u64, []),
        "sk_last_error": (ctypes.c_char_p, [vp]),
        "sk_version": (i, []),
        "sk_device_count": (i, []),
    }
    assert set(protos) == set(EXPORTS)
    for name, (res, args) in protos.items():
        try:
            fn = getattr(L, name)
        except AttributeError as e:
            raise PreencodeLibraryError(f"{LIB_PATH} does not export {name}; rebuild it") from e
        fn.restype = res
        fn.argtypes = args
    # The tuning build (make -C csrc tuning, selected with SELKIES_PREENCODE_LIB) also carries the
    # kernel-shape selector of the round-1 sweep; it is not part of the public header.
    if hasattr(L, "sk_tuning_set_variant"):
        L.sk_tuning_set_variant.restype = i
        L.sk_tuning_set_variant.argtypes = [i]
    if L.sk_version() != SK_ABI_VERSION:
        raise PreencodeLibraryError(f"ABI version mismatch: library {L.sk_version()}, binding {SK_ABI_VERSION}")
    _lib = L
    return L
