"""CPU checks of the drop-in boundary: the C-ABI library builds, loads, exports
every symbol include/selkies_preencode.h declares, and fails loudly (no
fallback) when there is no CUDA device.  No compute calls here."""
import ctypes
import os
import re

import pytest

from selkies_gstreamer_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "selkies_preencode.h")


@pytest.fixture(scope="module")
def lib():
    _abi.build()
    return _abi.load()


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sk_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(_abi.EXPORTS)


def test_library_exports_every_declared_symbol(lib):
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_version_and_struct_sizes(lib):
    assert lib.sk_version() == _abi.SK_ABI_VERSION
    assert ctypes.sizeof(_abi.SkConfig) == 19 * 4
    # the header's sk_result: u64, 2 ptr, size_t, 3 size_t, 3 i32 (+pad), 2 ptr, 2 i32, u32, 3 float, 2 ptr, ptr, u64
    # (the library static_asserts the same number)
    assert ctypes.sizeof(_abi.SkResult) == 8 + 16 + 8 + 24 + 16 + 16 + 8 + 4 + 12 + 16 + 8 + 8 == 144


def test_header_cites_reference_lines():
    src = open(HEADER).read()
    assert src.count("gstwebrtc_app.py:") >= 8


def test_bad_struct_size_is_rejected(lib):
    cfg = _abi.SkConfig()
    cfg.struct_size = 4
    h = ctypes.c_void_p()
    assert lib.sk_session_create(ctypes.byref(cfg), ctypes.byref(h)) == _abi.SK_ERR_ARG
    assert b"struct_size" in lib.sk_last_error(None)


@pytest.mark.parametrize("kw,code", [
    (dict(in_width=1921), _abi.SK_ERR_ARG),            # odd width (reference forces even: webrtc_input.py:677)
    (dict(in_height=0), _abi.SK_ERR_ARG),
    (dict(in_width=8194), _abi.SK_ERR_ARG),
    (dict(out_format=7), _abi.SK_ERR_ARG),
    (dict(matrix=9), _abi.SK_ERR_ARG),
    (dict(chroma_site=3), _abi.SK_ERR_ARG),
    (dict(n_inflight=1), _abi.SK_ERR_ARG),
    (dict(in_stride=100), _abi.SK_ERR_ARG),
    (dict(in_width=2, in_height=8192, in_stride=1 << 20), _abi.SK_ERR_ARG),   # 32-bit byte offsets in the kernels
    (dict(parity_switches=1 << 9), _abi.SK_ERR_ARG),                         # unknown switch bit
    (dict(flags=_abi.SK_FLAG_GATED_READBACK), _abi.SK_ERR_ARG),              # needs SK_FLAG_HOST_OUTPUT
    (dict(flags=_abi.SK_FLAG_GATED_READBACK | _abi.SK_FLAG_HOST_OUTPUT | _abi.SK_FLAG_NO_DAMAGE), _abi.SK_ERR_ARG),
    (dict(out_width=1280, out_height=720, scale_method=2), _abi.SK_ERR_UNSUPPORTED),   # only bilinear
    (dict(damage_block_threshold=300), _abi.SK_ERR_ARG),
    (dict(paint_over_trigger_frames=-1), _abi.SK_ERR_ARG),
])
def test_config_validation_happens_before_cuda(lib, kw, code):
    base = dict(struct_size=ctypes.sizeof(_abi.SkConfig), device_id=-1, in_width=1920, in_height=1080, in_stride=0,
                in_format=0, out_width=0, out_height=0, out_format=0, matrix=1, chroma_site=1, scale_method=1,
                scale_variant=0, n_inflight=2, flags=0)
    base.update(kw)
    cfg = _abi.SkConfig(**base)
    h = ctypes.c_void_p()
    assert lib.sk_session_create(ctypes.byref(cfg), ctypes.byref(h)) == code
    assert not h.value
    assert lib.sk_last_error(None)


def test_no_device_fails_loudly_not_silently(lib):
    if lib.sk_device_count() > 0:
        pytest.skip("a CUDA device is present")
    from selkies_gstreamer_b200 import PreencodeError, PreencodeSession
    with pytest.raises(PreencodeError) as e:
        PreencodeSession(1920, 1080)
    assert e.value.code == _abi.SK_ERR_CUDA and "no CPU path" in str(e.value)


def test_missing_library_raises(monkeypatch, tmp_path):
    monkeypatch.setattr(_abi, "LIB_PATH", str(tmp_path / "nope.so"))
    monkeypatch.setattr(_abi, "_lib", None)
    with pytest.raises(_abi.PreencodeLibraryError):
        _abi.load()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "selkies_gstreamer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower().replace("# oracle-free", ""), os.path.join(dirpath, f)


def test_null_handles_are_rejected(lib):
    assert lib.sk_session_destroy(None) == _abi.SK_ERR_ARG
    assert lib.sk_submit(None, 0, 0) == _abi.SK_ERR_ARG
    assert lib.sk_wait(None, 0, None) == _abi.SK_ERR_ARG
    assert lib.sk_reset_history(None) == _abi.SK_ERR_ARG
    assert lib.sk_submit_io(None, 0, None, 0, None, 0) == _abi.SK_ERR_ARG
    assert lib.sk_wait_device(None, 0, None, None) == _abi.SK_ERR_ARG
    assert lib.sk_release_device(None, 0, None) == _abi.SK_ERR_ARG
    assert lib.sk_read_frame(None, 0, None) == _abi.SK_ERR_ARG
    assert lib.sk_batch_prepare(None, 0) == _abi.SK_ERR_ARG
    assert lib.sk_batch_submit_host(None, 0, None, 0, None, 0) == _abi.SK_ERR_ARG
    assert lib.sk_pcie_probe(0, 0, 1, 0, None, None) == _abi.SK_ERR_ARG


def test_release_library_has_no_tuning_selector(lib):
    """The process-wide kernel-shape switch of the round-1 sweep lives in the tuning build only."""
    assert not hasattr(lib, "sk_set_kernel_variant")
    if "tuning" not in os.path.basename(_abi.LIB_PATH):
        assert not hasattr(lib, "sk_tuning_set_variant")
