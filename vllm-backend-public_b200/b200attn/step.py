"""One decode step of one layer, driven from HOST buffers: the call an engine loop makes.

    step = DecodeStep(k_cache, v_cache, max_batch=B, Hq=32, max_nb=256)
    step.h_qkv[...] = ...; step.h_block_tables[...] = ...; step.h_seq_lens[...] = ...; step.h_slots[...] = ...
    out = step.run(B)            # pinned bf16 [B, Hq, d], valid when run() returns

What run() enqueues (captured once per (B, schedule) in a CUDA graph, then replayed):
    1 H2D copy  fused q|k|v rows of the step's tokens      pinned [B, Hq + 2 Hkv, d] bf16
    1 H2D copy  block_tables | seq_lens | slot_mapping     pinned int32 words
    b2_reshape_and_cache   (k, v are strided views of the fused buffer: no split copy)
    b2_paged_decode        (q is a strided view; `out` is the PINNED host tensor: the kernels store the result
                            rows straight into host memory over PCIe, so there is no separate D2H copy)
and then synchronises the stream.  The schedule (single-pass grid or dynamic persistent) is chosen on the host
from h_seq_lens by b2_paged_decode_plan_splits, as an engine's scheduler would.

Mirrors the per-step work of the engine behind the reference's launch contract
(reference benchmark-vllm/vllm_manager.py:371-374; sizes from benchmark-vllm/models.py:9,10,15).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import ops


class DecodeStep:
    def __init__(self, k_cache: torch.Tensor, v_cache: torch.Tensor, max_batch: int, Hq: int, max_nb: int,
                 k_scale: Optional[torch.Tensor] = None, v_scale: Optional[torch.Tensor] = None,
                 use_graph: bool = True, zero_copy_inputs: bool = False):
        if not k_cache.is_cuda:
            raise ValueError("DecodeStep needs CUDA caches (b200attn has no CPU path)")
        self.dev = k_cache.device
        self.k_cache, self.v_cache, self.k_scale, self.v_scale = k_cache, v_cache, k_scale, v_scale
        _, self.Hkv, self.bs, self.d = k_cache.shape
        self.Hq, self.max_nb, self.max_batch = Hq, max_nb, max_batch
        self.use_graph = use_graph
        # True: no H2D copy of the q|k|v rows; reshape_and_cache and paged_decode read them from the pinned buffer
        self.zero_copy_inputs = zero_copy_inputs
        B, H = max_batch, Hq + 2 * self.Hkv
        self.h_qkv = torch.zeros(B, H, self.d, dtype=torch.bfloat16).pin_memory()
        self.d_qkv = torch.zeros(B, H, self.d, dtype=torch.bfloat16, device=self.dev)
        # int32 words: block_tables [B*max_nb] | seq_lens [B] | pad to even | slots as int64 [B]
        self._off_sl = B * max_nb
        self._off_slot = (self._off_sl + B + 1) // 2 * 2
        n_words = self._off_slot + 2 * B
        self.h_meta = torch.zeros(n_words, dtype=torch.int32).pin_memory()
        self.d_meta = torch.zeros(n_words, dtype=torch.int32, device=self.dev)
        self.h_block_tables = self.h_meta[: self._off_sl].view(B, max_nb)
        self.h_seq_lens = self.h_meta[self._off_sl: self._off_sl + B]
        self.h_slots = self.h_meta[self._off_slot: self._off_slot + 2 * B].view(torch.int64)
        self.h_out = torch.zeros(B, Hq, self.d, dtype=torch.bfloat16).pin_memory()
        ws = max(ops.decode_workspace_bytes(B, Hq, self.Hkv, self.d, self.bs, max_nb, 0), 16)
        self.workspace = torch.zeros(ws, dtype=torch.uint8, device=self.dev)
        self._graphs: Dict[Tuple[int, int], torch.cuda.CUDAGraph] = {}
        self._stream = torch.cuda.Stream(self.dev)

    # ------------------------------------------------------------------------------------------
    @property
    def h2d_bytes(self) -> int:
        return self.h_qkv.numel() * 2 + self.h_meta.numel() * 4   # zero-copy moves the same bytes, inside the kernels

    @property
    def d2h_bytes(self) -> int:
        return self.h_out.numel() * 2

    def _views(self, B: int, zero_copy: bool):
        Hq, Hkv = self.Hq, self.Hkv
        src = self.h_qkv if zero_copy else self.d_qkv
        q = src[:B, :Hq]
        k = src[:B, Hq:Hq + Hkv]
        v = src[:B, Hq + Hkv:]
        bt = self.d_meta[: self._off_sl].view(self.max_batch, self.max_nb)[:B]
        sl = self.d_meta[self._off_sl: self._off_sl + B]
        slots = self.d_meta[self._off_slot: self._off_slot + 2 * self.max_batch].view(torch.int64)[:B]
        return q, k, v, bt, sl, slots

    def _enqueue(self, B: int, num_splits: int) -> None:
        # the dynamic schedule fetches a run's q rows with every run it starts: a PCIe round trip each would sit on
        # the pipeline's critical path, so zero-copy inputs are for the single-pass / grid-split schedules only
        zero_copy = self.zero_copy_inputs and num_splits != 0
        if not zero_copy:
            self.d_qkv.copy_(self.h_qkv, non_blocking=True)
        self.d_meta.copy_(self.h_meta, non_blocking=True)
        q, k, v, bt, sl, slots = self._views(B, zero_copy)
        ops.reshape_and_cache(k, v, self.k_cache, self.v_cache, slots, self.k_scale, self.v_scale)
        ops.paged_decode(q, self.k_cache, self.v_cache, bt, sl, k_scale=self.k_scale, v_scale=self.v_scale,
                         out=self.h_out[:B], workspace=self.workspace, num_splits=num_splits)

    def plan(self, B: int) -> int:
        return ops.plan_num_splits(self.h_seq_lens[:B], self.Hq, self.Hkv, self.d, self.bs,
                                   ops.KV_FP8_E4M3 if self.k_cache.dtype == torch.uint8 else ops.KV_BF16)

    def run(self, B: Optional[int] = None, num_splits: Optional[int] = None) -> torch.Tensor:
        """Run the step for the first B rows of the host buffers; returns the pinned output rows (valid on return)."""
        B = self.max_batch if B is None else B
        if not 0 < B <= self.max_batch:
            raise ValueError(f"B={B} outside (0, {self.max_batch}]")
        if num_splits is None:
            num_splits = self.plan(B)
        # everything runs on the caller's current stream (a stream context per step costs ~7 us of host time);
        # only the one-time capture uses the private stream
        cur = torch.cuda.current_stream(self.dev)
        if not self.use_graph:
            self._enqueue(B, num_splits)
        else:
            g = self._graphs.get((B, num_splits))
            if g is None:
                cur.synchronize()
                with torch.cuda.stream(self._stream):
                    self._enqueue(B, num_splits)          # warm-up outside capture (function attributes, lazy init)
                    self._stream.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=self._stream):
                        self._enqueue(B, num_splits)
                self._graphs[(B, num_splits)] = g
            g.replay()
        cur.synchronize()
        return self.h_out[:B]
