"""One decode step of one layer, driven from HOST buffers: the call an engine loop makes.

    step = DecodeStep(k_cache, v_cache, max_batch=B, Hq=32, max_nb=256)
    step.h_block_tables[...] = ...; step.upload_block_tables()      # when sequences are placed into batch rows
    step.h_qkv[...] = ...; step.h_seq_lens[...] = ...; step.h_slots[...] = ...      # every step
    out = step.run(B)            # pinned bf16 [B, Hq, d], valid when run() returns

The block tables are RESIDENT on the device.  A decode step appends one token per sequence, so at most one table
entry per row changes, and its value is already in the step's slot mapping: b2_reshape_and_cache_decode writes
block_tables[b][(seq_len - 1) // bs] = slot // bs in the same pass as the KV write.  A row is uploaded in full only
when a new sequence takes it (upload_block_tables).  Per-step host->device traffic is therefore the step's q|k|v rows
plus seq_lens and slots -- O(B) words, not the [B, max_nb] table.

What run() enqueues (captured once per (B, schedule) in a CUDA graph, then replayed):
    1 H2D copy  seq_lens | slot_mapping                    pinned int32 words, 12 B per sequence
    (dynamic schedule only: 1 H2D copy of the fused q|k|v rows; the grid schedules read them in place, below)
    b2_reshape_and_cache_decode   k, v are strided views of the PINNED fused q|k|v buffer: the kernel reads the rows
                                  over PCIe itself (the H2D leg, no staging copy)
    b2_decode_plan_run            q likewise; `out` is the PINNED host tensor: result rows leave as 16-byte vector
                                  stores straight into host memory (the D2H leg, no separate copy)
    (B2_STEP_FUSE_APPEND=1: both as one call, b2_decode_plan_run_append -- one launch for the single-pass grid; same
    bits, measured no faster on this path, see __init__)
and then synchronises the stream.  The schedule (single-pass grid or dynamic persistent) is chosen on the host
from h_seq_lens by b2_paged_decode_plan_splits, as an engine's scheduler would.

Mirrors the per-step work of the engine behind the reference's launch contract
(reference benchmark-vllm/vllm_manager.py:371-374; sizes from benchmark-vllm/models.py:9,10,15).
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import torch

from . import _lib, ops


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
        self.kv_dtype = ops.KV_FP8_E4M3 if k_cache.dtype == torch.uint8 else ops.KV_BF16
        B, H = max_batch, Hq + 2 * self.Hkv
        self.h_qkv = torch.zeros(B, H, self.d, dtype=torch.bfloat16).pin_memory()
        self.d_qkv = torch.zeros(B, H, self.d, dtype=torch.bfloat16, device=self.dev)
        # per-step words: seq_lens [B] int32 | pad to even | slots as int64 [B]
        self._off_slot = (B + 1) // 2 * 2
        n_words = self._off_slot + 2 * B
        self.h_meta = torch.zeros(n_words, dtype=torch.int32).pin_memory()
        self.d_meta = torch.zeros(n_words, dtype=torch.int32, device=self.dev)
        self.h_seq_lens = self.h_meta[:B]
        self.h_slots = self.h_meta[self._off_slot: self._off_slot + 2 * B].view(torch.int64)
        # resident tables: host staging (filled by the caller when sequences are placed) + the device copy
        self.h_block_tables = torch.zeros(B, max_nb, dtype=torch.int32).pin_memory()
        self.d_block_tables = torch.zeros(B, max_nb, dtype=torch.int32, device=self.dev)
        self.h_out = torch.zeros(B, Hq, self.d, dtype=torch.bfloat16).pin_memory()
        self.plan_handle = ops.DecodePlan(k_cache, v_cache, B, Hq, max_nb, k_scale=k_scale, v_scale=v_scale)
        self.workspace = self.plan_handle.workspace
        self._graphs: Dict[Tuple[int, int], torch.cuda.CUDAGraph] = {}
        self._stream = torch.cuda.Stream(self.dev)
        self._plan_fn = _lib.load().b2_paged_decode_plan_splits
        self._sl_ptr = self.h_seq_lens.data_ptr()
        # the schedule choice is a performance heuristic over lengths that grow by one token per step: it is refreshed
        # every `plan_every` steps (and whenever the batch size changes), not recomputed on the critical path of each
        self.plan_every = 8
        # B2_STEP_FUSE_APPEND=1: the step as ONE call (b2_decode_plan_run_append).  Off by default: measured on this path
        # (scripts/e2e_append_ab.py, config 2, back-to-back graph replays) 176.9-178.9 us fused vs 176.3-176.4 us for the
        # two calls with zero-copy inputs, 188.2 vs 188.2 us with the staged copy -- the step's cost over the resident
        # loop (164 us) is the 787 KB of q|k|v crossing PCIe before the first CTAs can start, not the kernel boundary
        self.fuse_append = os.environ.get("B2_STEP_FUSE_APPEND", "0") == "1"
        self._plan_cache = (0, 0, 0)            # (B, num_splits, steps left)

    # ------------------------------------------------------------------------------------------
    @property
    def h2d_bytes(self) -> int:
        """Per step: q|k|v rows + seq_lens + slots (zero-copy moves the same bytes, inside the kernels)."""
        return self.h_qkv.numel() * 2 + self.h_meta.numel() * 4

    @property
    def d2h_bytes(self) -> int:
        return self.h_out.numel() * 2

    def upload_block_tables(self, rows: Optional[slice] = None) -> None:
        """Copy h_block_tables (all rows, or `rows`) to the resident device tables.  Needed when a sequence is placed
        into a batch row; decode steps keep the tables current by themselves (see the module docstring)."""
        rows = slice(0, self.max_batch) if rows is None else rows
        self.d_block_tables[rows].copy_(self.h_block_tables[rows], non_blocking=True)

    def _views(self, B: int, zero_copy: bool):
        Hq, Hkv = self.Hq, self.Hkv
        src = self.h_qkv if zero_copy else self.d_qkv
        q = src[:B, :Hq]
        k = src[:B, Hq:Hq + Hkv]
        v = src[:B, Hq + Hkv:]
        sl = self.d_meta[:B]
        slots = self.d_meta[self._off_slot: self._off_slot + 2 * self.max_batch].view(torch.int64)[:B]
        return q, k, v, self.d_block_tables[:B], sl, slots

    def _enqueue(self, B: int, num_splits: int) -> None:
        # the dynamic schedule fetches a run's q rows with every run it starts: a PCIe round trip each would sit on
        # the pipeline's critical path, so zero-copy inputs are for the single-pass / grid-split schedules only
        zero_copy = self.zero_copy_inputs and num_splits != 0
        if not zero_copy:
            self.d_qkv.copy_(self.h_qkv, non_blocking=True)
        self.d_meta.copy_(self.h_meta, non_blocking=True)
        q, k, v, bt, sl, slots = self._views(B, zero_copy)
        if self.fuse_append:
            # one call; for the single-pass grid one launch (b2_decode_plan_run_append: the CTA that owns (row, kv head)
            # writes the row's new K/V and its table entry itself), else the KV write launched in front of the grid
            self.plan_handle.run_append(q, k, v, slots, bt, sl, self.h_out[:B], num_splits)
        else:
            ops.reshape_and_cache_decode(k, v, self.k_cache, self.v_cache, slots, bt, sl, self.k_scale, self.v_scale)
            self.plan_handle.run(q, bt, sl, self.h_out[:B], num_splits)

    def plan(self, B: int) -> int:
        # straight ctypes call on the pinned buffer's address: this runs on the host every step, before the launch
        rc = self._plan_fn(self._sl_ptr, B, self.Hq, self.Hkv, self.d, self.bs, self.kv_dtype)
        if rc < 0:
            _lib.check(rc)
        return rc

    def run(self, B: Optional[int] = None, num_splits: Optional[int] = None) -> torch.Tensor:
        """Run the step for the first B rows of the host buffers; returns the pinned output rows (valid on return)."""
        B = self.max_batch if B is None else B
        if not 0 < B <= self.max_batch:
            raise ValueError(f"B={B} outside (0, {self.max_batch}]")
        if num_splits is None:
            cb, cs, left = self._plan_cache
            if cb == B and left > 0:
                num_splits = cs
                self._plan_cache = (cb, cs, left - 1)
            else:
                num_splits = self.plan(B)
                self._plan_cache = (B, num_splits, self.plan_every - 1)
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
