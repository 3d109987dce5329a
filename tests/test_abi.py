"""The C-ABI shared library loads and exports every symbol include/b200attn.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "b200attn.h"


def _declared():
    text = HEADER.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"B2_API\s+[\w\s\*]+?\b(b2_\w+)\s*\(", text)))


def test_header_declares_expected_entry_points():
    names = _declared()
    for must in ("b2_version", "b2_last_error", "b2_reshape_and_cache", "b2_paged_decode",
                 "b2_paged_decode_workspace_bytes", "b2_bm_create", "b2_bm_plan_decode"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from b200attn import _lib
    assert _lib.LIB_PATH.exists(), "libb200attn.so not built (make -C vllm-backend-public_b200/csrc)"
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in b200attn.h but not exported"
    # and the Python binding covers the same set
    assert sorted(_lib.SIGNATURES) == _declared()


def test_version_and_error_text():
    from b200attn import _lib
    lib = _lib.load()
    assert lib.b2_version() == _lib.ABI_VERSION
    assert isinstance(lib.b2_last_error(), bytes)


def test_argument_validation_without_gpu():
    """Bad arguments are rejected on the host before any CUDA call."""
    from b200attn import _lib
    lib = _lib.load()
    rc = lib.b2_paged_decode(None, None, None, None, None, None, 4, 8, 2, 128, 16, 32, 100, 1024, 1024,
                             2 * 16 * 128, 16 * 128, 0.088, 0, None, None, None, 0, 0, None)
    assert rc == _lib.B2_EINVAL and b"null" in lib.b2_last_error()
    rc = lib.b2_paged_decode(16, 16, 16, 16, 16, 16, 4, 8, 2, 64, 16, 32, 100, 512, 512,
                             2 * 16 * 64, 16 * 64, 0.125, 0, None, None, None, 0, 0, None)
    assert rc == _lib.B2_EUNSUPPORTED and b"head_dim" in lib.b2_last_error()
    rc = lib.b2_reshape_and_cache(16, 16, 16, 16, 16, 4, 2, 128, 16, 256, 256, 4096, 2048, 7, None, None, None)
    assert rc == _lib.B2_EUNSUPPORTED
    assert lib.b2_paged_decode_workspace_bytes(64, 32, 8, 128, 16, 256, 1) == 0
    assert lib.b2_paged_decode_auto_splits(64, 32, 8, 128, 16, 256, 0) == 0      # 0 = persistent schedule
    assert lib.b2_paged_decode_workspace_bytes(64, 32, 8, 128, 16, 256, 4) == 64 * 32 * 4 * 129 * 4
    # dynamic persistent schedule: a 256-byte header (claim counter) + partial slots of 8 heads x (d + 1) floats; the slot count grows with the batch
    def hdr(B, ipb):
        return 256
    w = lib.b2_paged_decode_workspace_bytes(64, 32, 8, 128, 16, 256, 0)
    w4 = lib.b2_paged_decode_workspace_bytes(4, 32, 8, 128, 16, 256, 0)
    assert w > w4 > 256 and (w - hdr(64, 8)) % (8 * 129 * 4) == 0 and (w4 - hdr(4, 8)) % (8 * 129 * 4) == 0


def test_ops_refuse_cpu_tensors():
    import torch
    import b200attn
    q = torch.zeros(1, 8, 128, dtype=torch.bfloat16)
    kc = torch.zeros(4, 2, 16, 128, dtype=torch.bfloat16)
    with pytest.raises(ValueError, match="CUDA"):
        b200attn.paged_decode(q, kc, kc, torch.zeros(1, 2, dtype=torch.int32), torch.ones(1, dtype=torch.int32))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from b200attn import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(ImportError, match="no CPU"):
        _lib.load()
