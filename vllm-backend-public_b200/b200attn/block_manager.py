"""Python handle on the C++ block manager in libb200attn.so (include/b200attn.h, b2_bm_*).

Sizing knobs come from the reference's launch contract: block_size, max_num_seqs, max_model_len,
gpu_memory_utilization (reference benchmark-vllm/models.py:8-10,15)."""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import numpy as np

from . import _lib


class BlockManager:
    def __init__(self, num_blocks: int, block_size: int):
        self._lib = _lib.load()
        self._h = self._lib.b2_bm_create(num_blocks, block_size)
        if not self._h:
            raise _lib.B2Error(_lib.B2_EINVAL, self._lib.b2_last_error().decode())
        self.num_blocks = num_blocks
        self.block_size = block_size

    def close(self) -> None:
        if self._h:
            self._lib.b2_bm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_free(self) -> int:
        return int(self._lib.b2_bm_num_free(self._h))

    @property
    def num_seqs(self) -> int:
        return int(self._lib.b2_bm_num_seqs(self._h))

    def seq_len(self, seq_id: int) -> int:
        n = int(self._lib.b2_bm_seq_len(self._h, seq_id))
        if n < 0:
            raise _lib.B2Error(n, f"unknown sequence {seq_id}")
        return n

    def can_append(self, seq_id: int, n_tokens: int) -> bool:
        return bool(self._lib.b2_bm_can_append(self._h, seq_id, n_tokens))

    def alloc_seq(self, seq_id: int, n_tokens: int = 0) -> None:
        _lib.check(self._lib.b2_bm_alloc_seq(self._h, seq_id, n_tokens))

    def append(self, seq_id: int, n_new: int) -> np.ndarray:
        """Grow the sequence by n_new tokens; returns their slots (int64 [n_new])."""
        slots = np.empty(n_new, dtype=np.int64)
        _lib.check(self._lib.b2_bm_append(self._h, seq_id, n_new, slots.ctypes.data_as(C.c_void_p)))
        return slots

    def free_seq(self, seq_id: int) -> None:
        _lib.check(self._lib.b2_bm_free_seq(self._h, seq_id))

    def fill_tables(self, seq_ids: Sequence[int], max_nb: int, block_tables: np.ndarray = None,
                    seq_lens: np.ndarray = None):
        ids = np.ascontiguousarray(seq_ids, dtype=np.int64)
        B = ids.shape[0]
        if block_tables is None:
            block_tables = np.empty((B, max_nb), dtype=np.int32)
        if seq_lens is None:
            seq_lens = np.empty(B, dtype=np.int32)
        assert block_tables.dtype == np.int32 and block_tables.flags.c_contiguous and block_tables.shape[1] == max_nb
        assert seq_lens.dtype == np.int32 and seq_lens.flags.c_contiguous
        _lib.check(self._lib.b2_bm_fill_tables(self._h, ids.ctypes.data_as(C.c_void_p), B,
                                               block_tables.ctypes.data_as(C.c_void_p), max_nb,
                                               seq_lens.ctypes.data_as(C.c_void_p)))
        return block_tables, seq_lens

    def plan_decode(self, seq_ids: Sequence[int], max_nb: int, block_tables: np.ndarray = None,
                    seq_lens: np.ndarray = None, slot_mapping: np.ndarray = None):
        """One decode step: +1 token per sequence, then block_tables / seq_lens / slot_mapping."""
        ids = np.ascontiguousarray(seq_ids, dtype=np.int64)
        B = ids.shape[0]
        if block_tables is None:
            block_tables = np.empty((B, max_nb), dtype=np.int32)
        if seq_lens is None:
            seq_lens = np.empty(B, dtype=np.int32)
        if slot_mapping is None:
            slot_mapping = np.empty(B, dtype=np.int64)
        _lib.check(self._lib.b2_bm_plan_decode(self._h, ids.ctypes.data_as(C.c_void_p), B,
                                               block_tables.ctypes.data_as(C.c_void_p), max_nb,
                                               seq_lens.ctypes.data_as(C.c_void_p),
                                               slot_mapping.ctypes.data_as(C.c_void_p)))
        return block_tables, seq_lens, slot_mapping
