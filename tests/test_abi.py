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


def test_workspace_bound_is_monotone_in_batch():
    """the automatic schedule cuts smaller batches into finer chunks, so the bytes ONE batch size
    needs are not monotone in B (B=24 needed more than B=32 at max_nb 512).  The exported size must serve every
    batch 1..B, i.e. never shrink as B grows -- an engine sizes the workspace once for max_num_seqs and captures CUDA
    graphs for every smaller bucket against that one buffer."""
    from b200attn import _lib
    lib = _lib.load()
    for (Hq, Hkv, bs, max_nb) in [(32, 8, 16, 512), (32, 8, 16, 256), (64, 8, 16, 512), (8, 1, 16, 512), (32, 8, 8, 1024),
                                  (32, 8, 32, 128), (4, 2, 16, 64)]:
        prev = 0
        for B in range(1, 301):
            w = lib.b2_paged_decode_workspace_bytes(B, Hq, Hkv, 128, bs, max_nb, 0)
            assert w >= prev, f"workspace bound shrinks at B={B} ({w} < {prev}) for Hq={Hq} Hkv={Hkv} bs={bs} max_nb={max_nb}"
            prev = w
    # the r1 counter-example: the bound for 32 must cover what 24 needs (finer chunks, more partial slots)
    assert lib.b2_paged_decode_workspace_bytes(32, 32, 8, 128, 16, 512, 0) >= 9308896


def test_decode_plan_rejects_bad_arguments_without_gpu():
    from b200attn import _lib
    lib = _lib.load()
    assert lib.b2_decode_plan_create(None, None, 8, 8, 2, 128, 16, 32, 100, 4096, 2048, 0.088, 0, None, None, None, 0,
                                     None) is None
    assert b"null" in lib.b2_last_error()
    assert lib.b2_decode_plan_create(16, 16, 0, 8, 2, 128, 16, 32, 100, 4096, 2048, 0.088, 0, None, None, None, 0,
                                     None) is None
    assert b"max_B" in lib.b2_last_error()
    # a workspace smaller than b2_paged_decode_workspace_bytes(max_B, ...) is refused before any CUDA call
    assert lib.b2_decode_plan_create(16, 16, 8, 8, 2, 128, 16, 32, 100, 4096, 2048, 0.088, 0, None, None, 16, 64,
                                     None) is None
    assert b"workspace" in lib.b2_last_error()
    assert lib.b2_decode_plan_run(None, 16, 16, 16, 16, 1, 1024, 1024, 0, None) == _lib.B2_EINVAL
    lib.b2_decode_plan_destroy(None)          # a no-op, like free(NULL)
    rc = lib.b2_reshape_and_cache_decode(16, 16, 16, 16, 16, None, None, 4, 2, 128, 16, 32, 256, 256, 4096, 2048, 0,
                                         None, None, None)
    assert rc == _lib.B2_EINVAL and b"block_tables" in lib.b2_last_error()


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
