"""Torch-tensor front end of the C ABI: the two device ops on the hot path.

Mirrors the call shape of vLLM's `ops.reshape_and_cache` / `ops.paged_attention_v1/v2` (the ops
the engine launched by reference benchmark-vllm/vllm_manager.py:371-374 runs each decode step),
but every argument is checked here and handed to libb200attn.so as raw pointers + sizes on the
current CUDA stream.  CUDA tensors only: there is no CPU path.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import _lib
from ._lib import KV_BF16, KV_FP8_E4M3


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# step inputs that may also live in PINNED host memory: the kernels then read them over PCIe (unified addressing),
# which replaces the H2D copy of an engine step (b200attn.step.DecodeStep(zero_copy_inputs=True))
_PINNED_OK = ("q", "k_new", "v_new")


def _need_cuda(**tensors) -> torch.device:
    dev = None
    for name, t in tensors.items():
        if t is None:
            continue
        if not t.is_cuda and name in _PINNED_OK and t.is_pinned():
            continue
        if not t.is_cuda:
            raise ValueError(f"{name} must be a CUDA tensor (b200attn has no CPU path)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise ValueError(f"{name} is on {t.device}, expected {dev}")
    return dev


def _stream(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


def _kv_dtype_of(cache: torch.Tensor) -> int:
    if cache.dtype == torch.bfloat16:
        return KV_BF16
    if cache.dtype in (torch.uint8, torch.float8_e4m3fn):
        return KV_FP8_E4M3
    raise ValueError(f"KV cache dtype {cache.dtype} unsupported (bf16 or uint8/float8_e4m3fn)")


def _check_cache(k_cache: torch.Tensor, v_cache: torch.Tensor):
    if k_cache.dim() != 4 or k_cache.shape != v_cache.shape or k_cache.dtype != v_cache.dtype:
        raise ValueError("k_cache / v_cache must both be [num_blocks, Hkv, bs, d] of one dtype")
    if k_cache.stride() != v_cache.stride():
        raise ValueError("k_cache / v_cache must share strides")
    nb, Hkv, bs, d = k_cache.shape
    s = k_cache.stride()
    if s[3] != 1 or s[2] != d:
        raise ValueError("each (block, head) tile of the KV cache must be contiguous [bs, d]")
    return nb, Hkv, bs, d, s[0], s[1]


def reshape_and_cache(k_new: torch.Tensor, v_new: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                      slot_mapping: torch.Tensor, k_scale: Optional[torch.Tensor] = None,
                      v_scale: Optional[torch.Tensor] = None) -> None:
    """k_cache[slot // bs, h, slot % bs, :] = k_new[t, h, :] (same for v); slot < 0 skips the token.

    k_new / v_new: [T, Hkv, d] bf16 (the last two dims contiguous; a token stride is allowed so a
    fused QKV projection can be written without a copy).  fp8 caches need 1-element fp32 device
    tensors k_scale / v_scale."""
    lib = _lib.load()
    dev = _need_cuda(k_new=k_new, v_new=v_new, k_cache=k_cache, v_cache=v_cache, slot_mapping=slot_mapping,
                     k_scale=k_scale, v_scale=v_scale)
    nb, Hkv, bs, d, bstride, hstride = _check_cache(k_cache, v_cache)
    kvd = _kv_dtype_of(k_cache)
    if k_new.dtype != torch.bfloat16 or v_new.dtype != torch.bfloat16:
        raise ValueError("k_new / v_new must be bf16")
    if k_new.dim() != 3 or k_new.shape != v_new.shape or k_new.shape[1:] != (Hkv, d):
        raise ValueError(f"k_new / v_new must be [T, {Hkv}, {d}]")
    for t in (k_new, v_new):
        if t.stride(2) != 1 or t.stride(1) != d:
            raise ValueError("k_new / v_new: head and dim axes must be contiguous")
    T = k_new.shape[0]
    if slot_mapping.dtype != torch.int64 or slot_mapping.shape != (T,) or not slot_mapping.is_contiguous():
        raise ValueError("slot_mapping must be a contiguous int64 [T]")
    if kvd == KV_FP8_E4M3:
        for s in (k_scale, v_scale):
            if s is None or s.dtype != torch.float32 or s.numel() != 1:
                raise ValueError("fp8 KV cache needs 1-element fp32 device tensors k_scale / v_scale")
    _lib.check(lib.b2_reshape_and_cache(
        _ptr(k_new), _ptr(v_new), _ptr(k_cache), _ptr(v_cache), _ptr(slot_mapping),
        T, Hkv, d, bs, k_new.stride(0) if T else Hkv * d, v_new.stride(0) if T else Hkv * d,
        bstride, hstride, kvd, _ptr(k_scale), _ptr(v_scale), _stream(dev)))


def reshape_and_cache_decode(k_new: torch.Tensor, v_new: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                             slot_mapping: torch.Tensor, block_tables: torch.Tensor, seq_lens: torch.Tensor,
                             k_scale: Optional[torch.Tensor] = None, v_scale: Optional[torch.Tensor] = None) -> None:
    """Decode-step KV write over DEVICE-RESIDENT block tables (b2_reshape_and_cache_decode): token b is the newest
    token of batch row b, and its slot also names the block of table column (seq_lens[b] - 1) // bs, which is
    written in the same pass -- so a step uploads seq_lens + slots instead of the [B, max_nb] table."""
    lib = _lib.load()
    dev = _need_cuda(k_new=k_new, v_new=v_new, k_cache=k_cache, v_cache=v_cache, slot_mapping=slot_mapping,
                     block_tables=block_tables, seq_lens=seq_lens, k_scale=k_scale, v_scale=v_scale)
    nb, Hkv, bs, d, bstride, hstride = _check_cache(k_cache, v_cache)
    kvd = _kv_dtype_of(k_cache)
    if k_new.dtype != torch.bfloat16 or v_new.dtype != torch.bfloat16:
        raise ValueError("k_new / v_new must be bf16")
    if k_new.dim() != 3 or k_new.shape != v_new.shape or k_new.shape[1:] != (Hkv, d):
        raise ValueError(f"k_new / v_new must be [B, {Hkv}, {d}]")
    for t in (k_new, v_new):
        if t.stride(2) != 1 or t.stride(1) != d:
            raise ValueError("k_new / v_new: head and dim axes must be contiguous")
    B = k_new.shape[0]
    if slot_mapping.dtype != torch.int64 or slot_mapping.shape != (B,) or not slot_mapping.is_contiguous():
        raise ValueError("slot_mapping must be a contiguous int64 [B]")
    if block_tables.dtype != torch.int32 or block_tables.dim() != 2 or block_tables.shape[0] != B \
            or block_tables.stride(1) != 1 or (B > 1 and block_tables.stride(0) != block_tables.shape[1]):
        raise ValueError("block_tables must be a contiguous int32 [B, max_nb]")
    if seq_lens.dtype != torch.int32 or seq_lens.shape != (B,) or not seq_lens.is_contiguous():
        raise ValueError("seq_lens must be a contiguous int32 [B]")
    if kvd == KV_FP8_E4M3:
        for s in (k_scale, v_scale):
            if s is None or s.dtype != torch.float32 or s.numel() != 1:
                raise ValueError("fp8 KV cache needs 1-element fp32 device tensors k_scale / v_scale")
    _lib.check(lib.b2_reshape_and_cache_decode(
        _ptr(k_new), _ptr(v_new), _ptr(k_cache), _ptr(v_cache), _ptr(slot_mapping), _ptr(block_tables), _ptr(seq_lens),
        B, Hkv, d, bs, block_tables.shape[1], k_new.stride(0) if B else Hkv * d, v_new.stride(0) if B else Hkv * d,
        bstride, hstride, kvd, _ptr(k_scale), _ptr(v_scale), _stream(dev)))


def decode_workspace_bytes(B: int, Hq: int, Hkv: int, d: int, bs: int, max_nb: int, num_splits: int = 0) -> int:
    return int(_lib.load().b2_paged_decode_workspace_bytes(B, Hq, Hkv, d, bs, max_nb, num_splits))


def decode_auto_splits(B: int, Hq: int, Hkv: int, d: int, bs: int, max_nb: int, kv_dtype: int = KV_BF16) -> int:
    return int(_lib.load().b2_paged_decode_auto_splits(B, Hq, Hkv, d, bs, max_nb, kv_dtype))


def plan_num_splits(seq_lens_host: torch.Tensor, Hq: int, Hkv: int, d: int = 128, bs: int = 16,
                    kv_dtype: int = KV_BF16) -> int:
    """Schedule choice from the scheduler's HOST copy of seq_lens (int32 CPU tensor): 1 = single-pass grid,
    0 = dynamic persistent schedule.  Pure host arithmetic (b2_paged_decode_plan_splits)."""
    if seq_lens_host.is_cuda or seq_lens_host.dtype != torch.int32 or not seq_lens_host.is_contiguous():
        raise ValueError("seq_lens_host must be a contiguous int32 CPU tensor")
    rc = _lib.load().b2_paged_decode_plan_splits(seq_lens_host.data_ptr(), seq_lens_host.numel(), Hq, Hkv, d, bs, kv_dtype)
    if rc < 0:
        _lib.check(rc)
    return int(rc)


def paged_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_tables: torch.Tensor,
                 seq_lens: torch.Tensor, scale: Optional[float] = None, k_scale: Optional[torch.Tensor] = None,
                 v_scale: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                 workspace: Optional[torch.Tensor] = None, num_splits: int = 0, impl: str = "tma") -> torch.Tensor:
    """out[b, hq] = softmax(scale * q[b, hq] . K[b, :seq_lens[b], hq // G]) . V[...]  -> bf16 [B, Hq, d].

    `workspace` is a uint8 device tensor of at least decode_workspace_bytes(...) bytes; one is
    allocated per call when omitted (the engine passes a persistent one).  `out` may be a pinned host tensor.  impl="simt" selects
    the CUDA-core A/B kernel."""
    lib = _lib.load()
    # `out` may also be a PINNED host tensor: the kernels then store the result rows straight into host memory
    # (unified addressing), which replaces the D2H copy of an engine step (b200attn.step.DecodeStep)
    out_on_host = out is not None and not out.is_cuda
    if out_on_host and not out.is_pinned():
        raise ValueError("out must be a CUDA tensor or a pinned host tensor (b200attn has no CPU path)")
    dev = _need_cuda(q=q, k_cache=k_cache, v_cache=v_cache, block_tables=block_tables, seq_lens=seq_lens,
                     k_scale=k_scale, v_scale=v_scale, out=None if out_on_host else out, workspace=workspace)
    nb, Hkv, bs, d, bstride, hstride = _check_cache(k_cache, v_cache)
    kvd = _kv_dtype_of(k_cache)
    if q.dtype != torch.bfloat16 or q.dim() != 3 or q.shape[2] != d:
        raise ValueError(f"q must be bf16 [B, Hq, {d}]")
    B, Hq, _ = q.shape
    if q.stride(2) != 1 or q.stride(1) != d:
        raise ValueError("q: head and dim axes must be contiguous")
    if block_tables.dtype != torch.int32 or block_tables.dim() != 2 or block_tables.shape[0] != B \
            or not block_tables.is_contiguous():
        raise ValueError("block_tables must be a contiguous int32 [B, max_nb]")
    if seq_lens.dtype != torch.int32 or seq_lens.shape != (B,) or not seq_lens.is_contiguous():
        raise ValueError("seq_lens must be a contiguous int32 [B]")
    max_nb = block_tables.shape[1]
    if out is None:
        out = torch.empty(B, Hq, d, dtype=torch.bfloat16, device=dev)
    elif out.dtype != torch.bfloat16 or out.shape != (B, Hq, d) or out.stride(2) != 1 or out.stride(1) != d:
        raise ValueError("out must be bf16 [B, Hq, d] with contiguous head/dim axes")
    if scale is None:
        scale = 1.0 / math.sqrt(d)
    if kvd == KV_FP8_E4M3:
        for s in (k_scale, v_scale):
            if s is None or s.dtype != torch.float32 or s.numel() != 1:
                raise ValueError("fp8 KV cache needs 1-element fp32 device tensors k_scale / v_scale")
    if B == 0:
        return out
    qs = q.stride(0) if B > 1 else Hq * d
    os_ = out.stride(0) if B > 1 else Hq * d
    if impl == "simt":
        _lib.check(lib.b2_paged_decode_simt(
            _ptr(out), _ptr(q), _ptr(k_cache), _ptr(v_cache), _ptr(block_tables), _ptr(seq_lens),
            B, Hq, Hkv, d, bs, max_nb, qs, os_, bstride, hstride, float(scale), kvd,
            _ptr(k_scale), _ptr(v_scale), _stream(dev)))
        return out
    if impl != "tma":
        raise ValueError(f"impl {impl!r}")
    need = lib.b2_paged_decode_workspace_bytes(B, Hq, Hkv, d, bs, max_nb, num_splits)
    if need and (workspace is None or workspace.numel() * workspace.element_size() < need):
        if workspace is not None:
            raise ValueError(f"workspace holds {workspace.numel() * workspace.element_size()} bytes, {need} required")
        workspace = torch.empty(need, dtype=torch.uint8, device=dev)
    wbytes = 0 if workspace is None else workspace.numel() * workspace.element_size()
    _lib.check(lib.b2_paged_decode(
        _ptr(out), _ptr(q), _ptr(k_cache), _ptr(v_cache), _ptr(block_tables), _ptr(seq_lens),
        B, Hq, Hkv, d, bs, max_nb, nb, qs, os_, bstride, hstride, float(scale), kvd,
        _ptr(k_scale), _ptr(v_scale), _ptr(workspace), wbytes, num_splits, _stream(dev)))
    return out


def paged_decode_append(q: torch.Tensor, k_new: torch.Tensor, v_new: torch.Tensor, slot_mapping: torch.Tensor,
                        k_cache: torch.Tensor, v_cache: torch.Tensor, block_tables: torch.Tensor, seq_lens: torch.Tensor,
                        scale: Optional[float] = None, k_scale: Optional[torch.Tensor] = None,
                        v_scale: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                        workspace: Optional[torch.Tensor] = None, num_splits: int = 0) -> torch.Tensor:
    """One decode step of one layer in one call (b2_paged_decode_append): reshape_and_cache(k_new, v_new, ...,
    slot_mapping) of the step's new token per row, then paged_decode over seq_lens (which count that token) -- same
    bits in `out` and in the caches as the two calls.  With num_splits = 1 and Hq / Hkv <= 16 it is ONE launch: the CTA
    that owns (row, kv head) writes the row's new K/V itself before it streams the row (append_is_fused)."""
    lib = _lib.load()
    dev = _need_cuda(q=q, k_new=k_new, v_new=v_new, slot_mapping=slot_mapping, k_cache=k_cache, v_cache=v_cache,
                     block_tables=block_tables, seq_lens=seq_lens, k_scale=k_scale, v_scale=v_scale, out=out,
                     workspace=workspace)
    nb, Hkv, bs, d, bstride, hstride = _check_cache(k_cache, v_cache)
    kvd = _kv_dtype_of(k_cache)
    if q.dtype != torch.bfloat16 or q.dim() != 3 or q.shape[2] != d:
        raise ValueError(f"q must be bf16 [B, Hq, {d}]")
    B, Hq, _ = q.shape
    if q.stride(2) != 1 or q.stride(1) != d:
        raise ValueError("q: head and dim axes must be contiguous")
    if k_new.dtype != torch.bfloat16 or v_new.dtype != torch.bfloat16:
        raise ValueError("k_new / v_new must be bf16")
    if k_new.shape != (B, Hkv, d) or v_new.shape != (B, Hkv, d):
        raise ValueError(f"k_new / v_new must be [{B}, {Hkv}, {d}] (one new token per row)")
    for t in (k_new, v_new):
        if t.stride(2) != 1 or t.stride(1) != d:
            raise ValueError("k_new / v_new: head and dim axes must be contiguous")
    if slot_mapping.dtype != torch.int64 or slot_mapping.shape != (B,) or not slot_mapping.is_contiguous():
        raise ValueError("slot_mapping must be a contiguous int64 [B]")
    if block_tables.dtype != torch.int32 or block_tables.dim() != 2 or block_tables.shape[0] != B \
            or not block_tables.is_contiguous():
        raise ValueError("block_tables must be a contiguous int32 [B, max_nb]")
    if seq_lens.dtype != torch.int32 or seq_lens.shape != (B,) or not seq_lens.is_contiguous():
        raise ValueError("seq_lens must be a contiguous int32 [B]")
    max_nb = block_tables.shape[1]
    if out is None:
        out = torch.empty(B, Hq, d, dtype=torch.bfloat16, device=dev)
    elif out.dtype != torch.bfloat16 or out.shape != (B, Hq, d) or out.stride(2) != 1 or out.stride(1) != d:
        raise ValueError("out must be bf16 [B, Hq, d] with contiguous head/dim axes")
    if scale is None:
        scale = 1.0 / math.sqrt(d)
    if kvd == KV_FP8_E4M3:
        for s in (k_scale, v_scale):
            if s is None or s.dtype != torch.float32 or s.numel() != 1:
                raise ValueError("fp8 KV cache needs 1-element fp32 device tensors k_scale / v_scale")
    if B == 0:
        return out
    need = lib.b2_paged_decode_workspace_bytes(B, Hq, Hkv, d, bs, max_nb, num_splits)
    if need and (workspace is None or workspace.numel() * workspace.element_size() < need):
        if workspace is not None:
            raise ValueError(f"workspace holds {workspace.numel() * workspace.element_size()} bytes, {need} required")
        workspace = torch.empty(need, dtype=torch.uint8, device=dev)
    wbytes = 0 if workspace is None else workspace.numel() * workspace.element_size()
    one = Hkv * d
    _lib.check(lib.b2_paged_decode_append(
        _ptr(out), _ptr(q), _ptr(k_new), _ptr(v_new), _ptr(slot_mapping),
        k_new.stride(0) if B > 1 else one, v_new.stride(0) if B > 1 else one,
        _ptr(k_cache), _ptr(v_cache), _ptr(block_tables), _ptr(seq_lens),
        B, Hq, Hkv, d, bs, max_nb, nb, q.stride(0) if B > 1 else Hq * d, out.stride(0) if B > 1 else Hq * d,
        bstride, hstride, float(scale), kvd, _ptr(k_scale), _ptr(v_scale), _ptr(workspace), wbytes, num_splits,
        _stream(dev)))
    return out


def append_is_fused(Hq: int, Hkv: int, num_splits: int) -> bool:
    """True when paged_decode_append runs as one launch for these heads and this schedule (b2_paged_decode_append_fused)."""
    rc = _lib.load().b2_paged_decode_append_fused(Hq, Hkv, num_splits)
    if rc < 0:
        _lib.check(rc)
    return rc == 1


class DecodePlan:
    """Per-engine decode handle (b2_decode_plan_*): caches, geometry, scales and workspace are fixed at creation, where
    the two TMA descriptors are encoded and the workspace header is zeroed; run() is then argument checks + launches
    (no driver call, no getenv; two launches for the dynamic schedule).  The caller keeps caches and workspace alive."""

    def __init__(self, k_cache: torch.Tensor, v_cache: torch.Tensor, max_batch: int, Hq: int, max_nb: int,
                 scale: Optional[float] = None, k_scale: Optional[torch.Tensor] = None,
                 v_scale: Optional[torch.Tensor] = None, workspace: Optional[torch.Tensor] = None):
        self._lib = _lib.load()
        self._h = None
        dev = _need_cuda(k_cache=k_cache, v_cache=v_cache, k_scale=k_scale, v_scale=v_scale, workspace=workspace)
        nb, Hkv, bs, d, bstride, hstride = _check_cache(k_cache, v_cache)
        kvd = _kv_dtype_of(k_cache)
        need = max(int(self._lib.b2_paged_decode_workspace_bytes(max_batch, Hq, Hkv, d, bs, max_nb, 0)), 256)
        if workspace is None:
            workspace = torch.zeros(need, dtype=torch.uint8, device=dev)
        elif workspace.numel() * workspace.element_size() < need:
            raise ValueError(f"workspace holds {workspace.numel() * workspace.element_size()} bytes, {need} required")
        self.dev, self.Hq, self.Hkv, self.d, self.bs, self.max_nb, self.max_batch = dev, Hq, Hkv, d, bs, max_nb, max_batch
        # the handle stores raw pointers: keep every tensor behind them alive
        self._keep = (k_cache, v_cache, k_scale, v_scale, workspace)
        self.workspace = workspace
        if scale is None:
            scale = 1.0 / math.sqrt(d)
        h = self._lib.b2_decode_plan_create(
            _ptr(k_cache), _ptr(v_cache), max_batch, Hq, Hkv, d, bs, max_nb, nb, bstride, hstride, float(scale), kvd,
            _ptr(k_scale), _ptr(v_scale), _ptr(workspace), workspace.numel() * workspace.element_size(), _stream(dev))
        if not h:
            raise _lib.B2Error(_lib.B2_EINVAL, self._lib.b2_last_error().decode("utf-8", "replace"))
        self._h = h
        self._run = self._lib.b2_decode_plan_run

    def run(self, q: torch.Tensor, block_tables: torch.Tensor, seq_lens: torch.Tensor, out: torch.Tensor,
            num_splits: int = 0) -> torch.Tensor:
        """q / out: bf16 [B, Hq, d] (row stride allowed; pinned host memory allowed); block_tables int32 [B, max_nb]
        and seq_lens int32 [B] on the device.  Shapes are the creator's; only the cheap checks are repeated here."""
        B = q.shape[0]
        if q.dtype != torch.bfloat16 or out.dtype != torch.bfloat16 or q.shape[1:] != (self.Hq, self.d) \
                or out.shape != q.shape or q.stride(1) != self.d or out.stride(1) != self.d:
            raise ValueError(f"q / out must be bf16 [B, {self.Hq}, {self.d}] with contiguous head/dim axes")
        if block_tables.dtype != torch.int32 or block_tables.shape != (B, self.max_nb) or not block_tables.is_cuda \
                or block_tables.stride(1) != 1 or (B > 1 and block_tables.stride(0) != self.max_nb):
            raise ValueError(f"block_tables must be a device int32 [B, {self.max_nb}]")
        if seq_lens.dtype != torch.int32 or seq_lens.shape != (B,) or not seq_lens.is_cuda:
            raise ValueError("seq_lens must be a device int32 [B]")
        for name, t in (("q", q), ("out", out)):
            if not t.is_cuda and not t.is_pinned():
                raise ValueError(f"{name} must be a CUDA tensor or a pinned host tensor (b200attn has no CPU path)")
        qs = q.stride(0) if B > 1 else self.Hq * self.d
        os_ = out.stride(0) if B > 1 else self.Hq * self.d
        _lib.check(self._run(self._h, out.data_ptr(), q.data_ptr(), block_tables.data_ptr(), seq_lens.data_ptr(), B,
                             qs, os_, num_splits, _stream(self.dev)))
        return out

    def run_append(self, q: torch.Tensor, k_new: torch.Tensor, v_new: torch.Tensor, slot_mapping: torch.Tensor,
                   block_tables: torch.Tensor, seq_lens: torch.Tensor, out: torch.Tensor,
                   num_splits: int = 0) -> torch.Tensor:
        """reshape_and_cache_decode(k_new, v_new, ..., slot_mapping, block_tables, seq_lens) + run(...) in one call
        (b2_decode_plan_run_append): one launch for the single-pass grid.  k_new / v_new: bf16 [B, Hkv, d] rows (strided
        views and pinned host memory allowed); slot_mapping int64 [B] and the RESIDENT block_tables int32 [B, max_nb]
        (updated in place) on the device."""
        B = q.shape[0]
        if q.dtype != torch.bfloat16 or out.dtype != torch.bfloat16 or q.shape[1:] != (self.Hq, self.d) \
                or out.shape != q.shape or q.stride(1) != self.d or out.stride(1) != self.d:
            raise ValueError(f"q / out must be bf16 [B, {self.Hq}, {self.d}] with contiguous head/dim axes")
        for name, t in (("k_new", k_new), ("v_new", v_new)):
            if t.dtype != torch.bfloat16 or t.shape != (B, self.Hkv, self.d) or t.stride(2) != 1 or t.stride(1) != self.d:
                raise ValueError(f"{name} must be bf16 [B, {self.Hkv}, {self.d}] with contiguous head/dim axes")
        if slot_mapping.dtype != torch.int64 or slot_mapping.shape != (B,) or not slot_mapping.is_cuda \
                or not slot_mapping.is_contiguous():
            raise ValueError("slot_mapping must be a contiguous device int64 [B]")
        if block_tables.dtype != torch.int32 or block_tables.shape != (B, self.max_nb) or not block_tables.is_cuda \
                or block_tables.stride(1) != 1 or (B > 1 and block_tables.stride(0) != self.max_nb):
            raise ValueError(f"block_tables must be a device int32 [B, {self.max_nb}]")
        if seq_lens.dtype != torch.int32 or seq_lens.shape != (B,) or not seq_lens.is_cuda:
            raise ValueError("seq_lens must be a device int32 [B]")
        for name, t in (("q", q), ("out", out), ("k_new", k_new), ("v_new", v_new)):
            if not t.is_cuda and not t.is_pinned():
                raise ValueError(f"{name} must be a CUDA tensor or a pinned host tensor (b200attn has no CPU path)")
        one = self.Hkv * self.d
        _lib.check(self._lib.b2_decode_plan_run_append(
            self._h, out.data_ptr(), q.data_ptr(), k_new.data_ptr(), v_new.data_ptr(), slot_mapping.data_ptr(),
            k_new.stride(0) if B > 1 else one, v_new.stride(0) if B > 1 else one, block_tables.data_ptr(),
            seq_lens.data_ptr(), B, q.stride(0) if B > 1 else self.Hq * self.d,
            out.stride(0) if B > 1 else self.Hq * self.d, num_splits, _stream(self.dev)))
        return out

    def close(self) -> None:
        if self._h:
            self._lib.b2_decode_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------------------------------------
# prompt-chunk attention and the element-wise ops of a decoder layer
# ------------------------------------------------------------------------------------------------
def dequant_kv_blocks(k_cache: torch.Tensor, v_cache: torch.Tensor, src_blocks: torch.Tensor,
                      k_out: Optional[torch.Tensor] = None, v_out: Optional[torch.Tensor] = None):
    """bf16 copies of the fp8-e4m3 cache blocks `src_blocks` (device int32 [n]): out block i = float(e4m3) of block
    src_blocks[i], exact in bf16 and still unscaled (b2_dequant_kv_blocks).  Returns (k_out, v_out),
    bf16 [n, Hkv, bs, d]; pass the cache's k_scale / v_scale to prefill_attention together with them."""
    lib = _lib.load()
    dev = _need_cuda(k_cache=k_cache, v_cache=v_cache, src_blocks=src_blocks, k_out=k_out, v_out=v_out)
    nb, Hkv, bs, d, bstride, hstride = _check_cache(k_cache, v_cache)
    if _kv_dtype_of(k_cache) != KV_FP8_E4M3:
        raise ValueError("dequant_kv_blocks needs an fp8 (uint8) cache")
    if src_blocks.dtype != torch.int32 or src_blocks.dim() != 1 or not src_blocks.is_contiguous():
        raise ValueError("src_blocks must be a contiguous int32 [n]")
    n = src_blocks.numel()
    shape = (n, Hkv, bs, d)
    outs = []
    for o in (k_out, v_out):
        if o is None:
            o = torch.empty(shape, dtype=torch.bfloat16, device=dev)
        elif o.dtype != torch.bfloat16 or tuple(o.shape) != shape or not o.is_contiguous():
            raise ValueError(f"k_out / v_out must be contiguous bf16 {list(shape)}")
        outs.append(o)
    _lib.check(lib.b2_dequant_kv_blocks(_ptr(outs[0]), _ptr(outs[1]), _ptr(k_cache), _ptr(v_cache), _ptr(src_blocks),
                                        n, Hkv, d, bs, bstride, hstride, _stream(dev)))
    return outs[0], outs[1]


def fp8_prefill_plan(block_tables: torch.Tensor, kv_lens_host, bs: int):
    """For prompt chunks over an fp8 cache: which cache blocks to dequantise and the block table of the scratch
    cache they land in.  kv_lens_host[c] = ctx_start[c] + len_c (host ints: the scheduler knows them).
    Returns (src_blocks device int32 [n], scratch_block_tables device int32 [n_chunks, max nb_c])."""
    n_chunks = block_tables.shape[0]
    nbs = [(int(L) + bs - 1) // bs for L in kv_lens_host]
    if len(nbs) != n_chunks or max(nbs, default=0) > block_tables.shape[1]:
        raise ValueError("kv_lens_host does not match the chunks / block_tables")
    dev = block_tables.device
    rows = torch.repeat_interleave(torch.arange(n_chunks), torch.tensor(nbs, dtype=torch.int64))
    cols = torch.cat([torch.arange(n) for n in nbs]) if nbs else torch.zeros(0, dtype=torch.int64)
    src = block_tables[rows.to(dev), cols.to(dev)].contiguous()
    bt = torch.zeros(n_chunks, max(max(nbs, default=1), 1), dtype=torch.int32)
    off = 0
    for c, n in enumerate(nbs):
        bt[c, :n] = torch.arange(off, off + n, dtype=torch.int32)
        off += n
    return src, bt.to(dev)


def _prefill_over_fp8(q, k_cache, v_cache, block_tables, cu_q, ctx_start, total_tokens, scale, out, k_scale, v_scale,
                      kv_lens_host):
    """fp8 cache: gather + dequantise the blocks each chunk can see into a bf16 scratch cache, then the bf16 kernel.
    Without kv_lens_host one device read sizes the scratch."""
    if kv_lens_host is None:
        kv_lens_host = (ctx_start.to(torch.int64) + (cu_q[1:] - cu_q[:-1]).to(torch.int64)).cpu().tolist()
    src, bt = fp8_prefill_plan(block_tables, kv_lens_host, k_cache.shape[2])
    ks, vs = dequant_kv_blocks(k_cache, v_cache, src)
    return prefill_attention(q, ks, vs, bt, cu_q, ctx_start, total_tokens, scale, out, k_scale, v_scale)


def prefill_attention(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_tables: torch.Tensor,
                      cu_q: torch.Tensor, ctx_start: torch.Tensor, total_tokens: Optional[int] = None,
                      scale: Optional[float] = None, out: Optional[torch.Tensor] = None,
                      k_scale: Optional[torch.Tensor] = None, v_scale: Optional[torch.Tensor] = None,
                      kv_lens_host=None) -> torch.Tensor:
    """Causal attention of prompt chunks over the paged cache (b2_prefill_attention).

    q / out: bf16 [T, Hq, d] (token stride allowed); chunk c = rows cu_q[c]:cu_q[c+1], positions
    ctx_start[c] + i, blocks block_tables[c].  The chunk's K/V must already be in the cache.
    `total_tokens` = cu_q[-1] as a host int (defaults to q.shape[0]; it sizes the grid without a device sync).
    fp8 caches (uint8 + k_scale / v_scale): the visible blocks are dequantised into a bf16 scratch cache first
    (b2_dequant_kv_blocks); pass kv_lens_host = [ctx_start[c] + len_c] to size it without a device read.
    bf16 caches with k_scale / v_scale: the cache is taken to hold K / k_scale and V / v_scale (that scratch)."""
    lib = _lib.load()
    dev = _need_cuda(q=q, k_cache=k_cache, v_cache=v_cache, block_tables=block_tables, cu_q=cu_q,
                     ctx_start=ctx_start, out=out, k_scale=k_scale, v_scale=v_scale)
    nb, Hkv, bs, d, bstride, hstride = _check_cache(k_cache, v_cache)
    kvd = _kv_dtype_of(k_cache)
    if kvd == KV_FP8_E4M3:
        if k_scale is None or v_scale is None:
            raise ValueError("fp8 KV cache needs 1-element fp32 device tensors k_scale / v_scale")
        if ctx_start.numel() == 0 or q.shape[0] == 0:
            return out if out is not None else torch.empty_like(q)
        return _prefill_over_fp8(q, k_cache, v_cache, block_tables, cu_q, ctx_start, total_tokens, scale, out,
                                 k_scale, v_scale, kv_lens_host)
    if q.dtype != torch.bfloat16 or q.dim() != 3 or q.shape[2] != d or q.stride(2) != 1 or q.stride(1) != d:
        raise ValueError(f"q must be bf16 [T, Hq, {d}] with contiguous head/dim axes")
    T, Hq, _ = q.shape
    n_chunks = ctx_start.numel()
    for name, t, n in (("cu_q", cu_q, n_chunks + 1), ("ctx_start", ctx_start, n_chunks)):
        if t.dtype != torch.int32 or t.shape != (n,) or not t.is_contiguous():
            raise ValueError(f"{name} must be a contiguous int32 [{n}]")
    if block_tables.dtype != torch.int32 or block_tables.dim() != 2 or block_tables.shape[0] != n_chunks \
            or not block_tables.is_contiguous():
        raise ValueError("block_tables must be a contiguous int32 [n_chunks, max_nb]")
    if out is None:
        out = torch.empty(T, Hq, d, dtype=torch.bfloat16, device=dev)
    elif out.dtype != torch.bfloat16 or out.shape != (T, Hq, d) or out.stride(2) != 1 or out.stride(1) != d:
        raise ValueError("out must be bf16 [T, Hq, d] with contiguous head/dim axes")
    if total_tokens is None:
        total_tokens = T
    if not 0 <= total_tokens <= T:
        raise ValueError(f"total_tokens={total_tokens} outside [0, {T}]")
    if scale is None:
        scale = 1.0 / math.sqrt(d)
    if n_chunks == 0 or total_tokens == 0:
        return out
    _lib.check(lib.b2_prefill_attention(
        _ptr(out), _ptr(q), _ptr(k_cache), _ptr(v_cache), _ptr(block_tables), _ptr(cu_q), _ptr(ctx_start),
        n_chunks, total_tokens, Hq, Hkv, d, bs, block_tables.shape[1], nb,
        q.stride(0) if T > 1 else Hq * d, out.stride(0) if T > 1 else Hq * d, bstride, hstride,
        float(scale), kvd, _ptr(k_scale), _ptr(v_scale), _stream(dev)))
    return out


def _rows(t: torch.Tensor, name: str):
    if t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be bf16 [T, H] with a contiguous last axis")
    return t.shape[0], t.shape[1], (t.stride(0) if t.shape[0] > 1 else t.shape[1])


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None,
             out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out = bf16(w * h * rsqrt(mean(h^2) + eps)); with `residual`, h = bf16(x + residual) is stored back into
    residual first (fused add).  b2_rms_norm."""
    lib = _lib.load()
    dev = _need_cuda(x=x, weight=weight, residual=residual, out=out)
    T, H, xs = _rows(x, "x")
    if weight.dtype != torch.bfloat16 or weight.shape != (H,) or not weight.is_contiguous():
        raise ValueError(f"weight must be a contiguous bf16 [{H}]")
    if out is None:
        out = torch.empty(T, H, dtype=torch.bfloat16, device=dev)
    To, Ho, os_ = _rows(out, "out")
    if (To, Ho) != (T, H):
        raise ValueError("out shape differs from x")
    rs = 0
    if residual is not None:
        Tr, Hr, rs = _rows(residual, "residual")
        if (Tr, Hr) != (T, H):
            raise ValueError("residual shape differs from x")
    _lib.check(lib.b2_rms_norm(_ptr(out), _ptr(x), _ptr(residual), _ptr(weight), T, H, xs, os_, rs, float(eps),
                               _stream(dev)))
    return out


def silu_and_mul(gate_up: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[t, i] = silu(gate_up[t, i]) * gate_up[t, I + i]  (b2_silu_and_mul)."""
    lib = _lib.load()
    dev = _need_cuda(gate_up=gate_up, out=out)
    T, H2, is_ = _rows(gate_up, "gate_up")
    if H2 % 2:
        raise ValueError("gate_up must be [T, 2*I]")
    I = H2 // 2
    if out is None:
        out = torch.empty(T, I, dtype=torch.bfloat16, device=dev)
    To, Io, os_ = _rows(out, "out")
    if (To, Io) != (T, I):
        raise ValueError("out must be [T, I]")
    _lib.check(lib.b2_silu_and_mul(_ptr(out), _ptr(gate_up), T, I, is_, os_, _stream(dev)))
    return out


def rope_and_cache(qkv: torch.Tensor, positions: torch.Tensor, cos_sin: torch.Tensor, k_cache: torch.Tensor,
                   v_cache: torch.Tensor, slot_mapping: torch.Tensor, Hq: int,
                   k_scale: Optional[torch.Tensor] = None, v_scale: Optional[torch.Tensor] = None,
                   write_k: bool = True) -> None:
    """Neox rotary on the q and k parts of qkv [T, (Hq + 2 Hkv) * d] in place, then K', V into the paged cache
    through slot_mapping (b2_rope_and_cache_ex).  cos_sin: fp32 [max_pos, d] = cos | sin halves per position.
    write_k=False leaves the k part of qkv unrotated (the rotated rows go to the cache only)."""
    lib = _lib.load()
    dev = _need_cuda(qkv=qkv, positions=positions, cos_sin=cos_sin, k_cache=k_cache, v_cache=v_cache,
                     slot_mapping=slot_mapping, k_scale=k_scale, v_scale=v_scale)
    nb, Hkv, bs, d, bstride, hstride = _check_cache(k_cache, v_cache)
    kvd = _kv_dtype_of(k_cache)
    T, W, ts = _rows(qkv, "qkv")
    if W != (Hq + 2 * Hkv) * d:
        raise ValueError(f"qkv rows must hold (Hq + 2 Hkv) * d = {(Hq + 2 * Hkv) * d} elements, got {W}")
    if positions.dtype != torch.int32 or positions.shape != (T,) or not positions.is_contiguous():
        raise ValueError("positions must be a contiguous int32 [T]")
    if slot_mapping.dtype != torch.int64 or slot_mapping.shape != (T,) or not slot_mapping.is_contiguous():
        raise ValueError("slot_mapping must be a contiguous int64 [T]")
    if cos_sin.dtype != torch.float32 or cos_sin.dim() != 2 or cos_sin.shape[1] != d or not cos_sin.is_contiguous():
        raise ValueError(f"cos_sin must be a contiguous fp32 [max_pos, {d}]")
    if kvd == KV_FP8_E4M3:
        for s in (k_scale, v_scale):
            if s is None or s.dtype != torch.float32 or s.numel() != 1:
                raise ValueError("fp8 KV cache needs 1-element fp32 device tensors k_scale / v_scale")
    _lib.check(lib.b2_rope_and_cache_ex(_ptr(qkv), _ptr(positions), _ptr(cos_sin), cos_sin.shape[0], _ptr(k_cache),
                                        _ptr(v_cache), _ptr(slot_mapping), T, Hq, Hkv, d, bs, ts, bstride, hstride, kvd,
                                        _ptr(k_scale), _ptr(v_scale), 1 if write_k else 0, _stream(dev)))


SKINNY_GEMM_MAX_M = 256


def skinny_gemm(x: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[M, N] = x[M, K] @ w[N, K]^T for the rows of a decode step (M <= 256): the weight-streaming tcgen05 kernel
    b2_skinny_gemm.  x bf16 [M, K] (row stride allowed), w bf16 [N, K] contiguous, out bf16 [M, N] (row stride allowed)."""
    lib = _lib.load()
    dev = _need_cuda(x=x, w=w, out=out)
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or x.dim() != 2 or w.dim() != 2 or x.shape[1] != w.shape[1]:
        raise ValueError("skinny_gemm: x [M, K] and w [N, K] must be bf16 with the same K")
    if x.stride(1) != 1 or not w.is_contiguous():
        raise ValueError("skinny_gemm: x rows and w must be contiguous")
    M, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    elif out.dtype != torch.bfloat16 or out.shape != (M, N) or out.stride(1) != 1:
        raise ValueError(f"skinny_gemm: out must be bf16 [{M}, {N}] with contiguous rows")
    _lib.check(lib.b2_skinny_gemm(_ptr(out), _ptr(x), _ptr(w), M, N, K, x.stride(0) if M > 1 else K,
                                  out.stride(0) if M > 1 else N, _stream(dev)))
    return out


def sample(logits: torch.Tensor, temperature: Optional[torch.Tensor] = None, seeds: Optional[torch.Tensor] = None,
           out: Optional[torch.Tensor] = None, top_k: Optional[torch.Tensor] = None,
           top_p: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Token per row of bf16 logits [B, V]: greedy where temperature <= 0 (or None), else one Gumbel-max draw from
    softmax(logits / T) keyed by seeds[b], restricted to the top_k[b] largest logits (int32 [B]; <= 0: no limit) and then
    to the smallest top set holding top_p[b] of that mass (fp32 [B]; >= 1: no limit) -- b2_sample_filtered.
    Returns int64 [B]."""
    lib = _lib.load()
    dev = _need_cuda(logits=logits, temperature=temperature, seeds=seeds, out=out, top_k=top_k, top_p=top_p)
    B, V, st = _rows(logits, "logits")
    if temperature is not None and (temperature.dtype != torch.float32 or temperature.shape != (B,)
                                    or not temperature.is_contiguous()):
        raise ValueError("temperature must be a contiguous fp32 [B]")
    if seeds is not None and (seeds.dtype != torch.int64 or seeds.shape != (B,) or not seeds.is_contiguous()):
        raise ValueError("seeds must be a contiguous int64 [B]")
    if out is None:
        out = torch.empty(B, dtype=torch.int64, device=dev)
    elif out.dtype != torch.int64 or out.shape != (B,) or not out.is_contiguous():
        raise ValueError("out must be a contiguous int64 [B]")
    if top_k is not None and (top_k.dtype != torch.int32 or top_k.shape != (B,) or not top_k.is_contiguous()):
        raise ValueError("top_k must be a contiguous int32 [B]")
    if top_p is not None and (top_p.dtype != torch.float32 or top_p.shape != (B,) or not top_p.is_contiguous()):
        raise ValueError("top_p must be a contiguous fp32 [B]")
    _lib.check(lib.b2_sample_filtered(_ptr(out), _ptr(logits), B, V, st, _ptr(temperature), _ptr(seeds), _ptr(top_k),
                                      _ptr(top_p), _stream(dev)))
    return out
