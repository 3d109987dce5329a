"""Python handle on the C++ continuous-batching scheduler in libb200attn.so (include/b200attn.h, b2_sched_*).

The limits are the reference's launch parameters: max_num_seqs, max_model_len, block_size and the pool sized from
gpu_memory_utilization (reference benchmark-vllm/models.py:8-10,15).  `step()` returns numpy views of the plan the
scheduler owns (valid until the next step): decode rows for b2_reshape_and_cache / b2_paged_decode, prefill chunks
for the prefill path, and the ids that were preempted / finished."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib


class _Plan(C.Structure):
    _fields_ = [
        ("n_decode", C.c_int), ("max_nb", C.c_int),
        ("decode_ids", C.POINTER(C.c_int64)), ("block_tables", C.POINTER(C.c_int32)),
        ("seq_lens", C.POINTER(C.c_int32)), ("slot_mapping", C.POINTER(C.c_int64)),
        ("n_prefill", C.c_int), ("n_prefill_tokens", C.c_int),
        ("prefill_ids", C.POINTER(C.c_int64)), ("prefill_start", C.POINTER(C.c_int32)),
        ("prefill_len", C.POINTER(C.c_int32)), ("prefill_last", C.POINTER(C.c_uint8)),
        ("prefill_slots", C.POINTER(C.c_int64)), ("prefill_block_tables", C.POINTER(C.c_int32)),
        ("n_preempted", C.c_int), ("preempted_ids", C.POINTER(C.c_int64)),
        ("n_finished", C.c_int), ("finished_ids", C.POINTER(C.c_int64)),
    ]


def _arr(ptr, n, dtype):
    if n <= 0:
        return np.empty(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).view(dtype)


@dataclass
class StepPlan:
    decode_ids: np.ndarray          # [n_decode] int64
    block_tables: np.ndarray        # [n_decode, max_nb] int32
    seq_lens: np.ndarray            # [n_decode] int32
    slot_mapping: np.ndarray        # [n_decode] int64
    prefill_ids: np.ndarray         # [n_prefill] int64
    prefill_start: np.ndarray       # [n_prefill] int32
    prefill_len: np.ndarray         # [n_prefill] int32
    prefill_last: np.ndarray        # [n_prefill] uint8: 0 more chunks, 1 sample token #1, 2 recompute done
    prefill_slots: np.ndarray       # [n_prefill_tokens] int64
    prefill_block_tables: np.ndarray  # [n_prefill, max_nb] int32
    preempted_ids: np.ndarray
    finished_ids: np.ndarray

    @property
    def empty(self) -> bool:
        return len(self.decode_ids) == 0 and len(self.prefill_ids) == 0


class Scheduler:
    def __init__(self, num_blocks: int, block_size: int, max_num_seqs: int, max_model_len: int,
                 max_num_batched_tokens: int = 8192):
        self._lib = _lib.load()
        self._h = self._lib.b2_sched_create(num_blocks, block_size, max_num_seqs, max_model_len, max_num_batched_tokens)
        if not self._h:
            raise _lib.B2Error(_lib.B2_EINVAL, self._lib.b2_last_error().decode())
        self.num_blocks, self.block_size = num_blocks, block_size
        self.max_num_seqs, self.max_model_len = max_num_seqs, max_model_len
        self.max_nb = (max_model_len + block_size - 1) // block_size
        self._plan = _Plan()

    def close(self) -> None:
        if self._h:
            self._lib.b2_sched_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add_request(self, req_id: int, prompt_len: int, max_new_tokens: int) -> None:
        _lib.check(self._lib.b2_sched_add_request(self._h, req_id, prompt_len, max_new_tokens))

    def finish(self, req_id: int) -> None:
        _lib.check(self._lib.b2_sched_finish(self._h, req_id))

    @property
    def num_waiting(self) -> int:
        return int(self._lib.b2_sched_num_waiting(self._h))

    @property
    def num_running(self) -> int:
        return int(self._lib.b2_sched_num_running(self._h))

    @property
    def num_free_blocks(self) -> int:
        return int(self._lib.b2_sched_num_free_blocks(self._h))

    def generated(self, req_id: int) -> int:
        return int(self._lib.b2_sched_generated(self._h, req_id))

    def step(self) -> StepPlan:
        p = self._plan
        _lib.check(self._lib.b2_sched_step(self._h, C.byref(p)))
        nd, npf, nb = p.n_decode, p.n_prefill, p.max_nb
        return StepPlan(
            decode_ids=_arr(p.decode_ids, nd, np.int64),
            block_tables=_arr(p.block_tables, nd * nb, np.int32).reshape(nd, nb),
            seq_lens=_arr(p.seq_lens, nd, np.int32),
            slot_mapping=_arr(p.slot_mapping, nd, np.int64),
            prefill_ids=_arr(p.prefill_ids, npf, np.int64),
            prefill_start=_arr(p.prefill_start, npf, np.int32),
            prefill_len=_arr(p.prefill_len, npf, np.int32),
            prefill_last=_arr(p.prefill_last, npf, np.uint8),
            prefill_slots=_arr(p.prefill_slots, p.n_prefill_tokens, np.int64),
            prefill_block_tables=_arr(p.prefill_block_tables, npf * nb, np.int32).reshape(npf, nb),
            preempted_ids=_arr(p.preempted_ids, p.n_preempted, np.int64),
            finished_ids=_arr(p.finished_ids, p.n_finished, np.int64),
        )
