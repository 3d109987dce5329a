"""Continuous-batching engine: C++ scheduler (b2_sched_*) -> token batch -> LlamaModel.forward -> b2_sample.

This is the process the reference deploys and probes (reference benchmark-vllm/vllm_manager.py:371-374,
888-983); its limits are the launch parameters of `VLLMConfig` (reference benchmark-vllm/models.py:8-15):
max_num_seqs, max_model_len, block_size, gpu_memory_utilization.  SURVEY.md section 8(f) ranks 1-3.

Per step:  plan = scheduler.step()  (decode rows first, then prompt chunks within the token budget)
  * decode-only plan  -> one CUDA-graph replay per (batch bucket, schedule): H2D of the step's int32/int64 rows from
    pinned memory, 32 layers, lm_head, b2_sample, D2H of the sampled ids.
  * plan with prompt chunks -> eager launch of the same ops over [decode rows | chunk rows].
The schedule of the decode kernel (single-pass grid or dynamic persistent) is chosen per step on the host from the
plan's seq_lens (b2_paged_decode_plan_splits).
"""
from __future__ import annotations

import itertools
import os
import threading
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np
import torch

from . import ops
from .model import LlamaModel, ModelConfig, PrefillBatch
from .scheduler import Scheduler, StepPlan


@dataclass
class SamplingParams:
    max_tokens: int = 16
    temperature: float = 1.0
    top_k: int = 0                    # <= 0: no limit
    top_p: float = 1.0                # >= 1: no limit
    seed: Optional[int] = None
    ignore_eos: bool = False
    stop_token_ids: Tuple[int, ...] = ()


@dataclass
class Request:
    id: int
    prompt_ids: List[int]
    params: SamplingParams
    output_ids: List[int] = field(default_factory=list)
    finished: bool = False
    finish_reason: Optional[str] = None
    t_arrival: float = 0.0
    t_first_token: Optional[float] = None
    t_done: Optional[float] = None
    token_times: List[float] = field(default_factory=list)
    on_token: Optional[Callable[["Request", Optional[int]], None]] = None   # called from the engine thread
    seed: int = 0
    n_preempted: int = 0


@dataclass
class EngineStats:
    steps: int = 0
    decode_steps: int = 0
    mixed_steps: int = 0
    prompt_tokens: int = 0
    generated_tokens: int = 0
    preemptions: int = 0
    gpu_launches: int = 0
    busy_s: float = 0.0
    device_ms: float = 0.0          # CUDA-event time of the decode-step replays (only with engine.time_device)
    device_steps: int = 0
    sync_wait_s: float = 0.0        # part of busy_s spent waiting for the GPU (stream synchronize): busy - wait = host work
    mixed_busy_s: float = 0.0


def _buckets(max_batch: int) -> List[int]:
    b = [1, 2, 4, 8]
    b += list(range(16, max_batch + 1, 8))
    b = sorted({x for x in b if x <= max_batch} | {max_batch})
    return b


class LLMEngine:
    def __init__(self, cfg: ModelConfig, device: Optional[torch.device] = None, *, max_num_seqs: int = 64,
                 max_model_len: int = 8192, block_size: int = 16, gpu_memory_utilization: float = 0.9,
                 max_num_batched_tokens: int = 8192, num_blocks: Optional[int] = None, seed: int = 0,
                 weights_dir=None, eos_token_id: Optional[int] = None, enforce_eager: bool = False,
                 kv_cache_dtype: str = "auto", tp=None):
        """`tp`: b200attn.tp.TPGroup when the engine runs tensor-parallel (one engine per rank, all built with the same
        arguments).  Rank 0 owns the request queue and publishes admissions / aborts before every step; the other ranks
        mirror its steps from `worker_loop()`."""
        if not torch.cuda.is_available():
            raise RuntimeError("b200attn engine needs a CUDA device: there is no CPU path")
        if kv_cache_dtype not in ("auto", "bfloat16", "bf16", "fp8", "fp8_e4m3"):
            raise ValueError(f"kv-cache-dtype {kv_cache_dtype!r}: supported are auto / bfloat16 / fp8 (= fp8_e4m3)")
        self.kv_dtype = torch.uint8 if kv_cache_dtype.startswith("fp8") else torch.bfloat16
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.dev)
        self.tp = tp if tp is not None and tp.size > 1 else None
        self.tp_size = self.tp.size if self.tp is not None else 1
        self.is_driver = self.tp is None or self.tp.rank == 0
        self._worker_msg = None
        self.cfg = cfg
        self.block_size = block_size
        self.max_num_seqs = max_num_seqs
        self.max_model_len = min(max_model_len, cfg.max_position) if max_model_len else cfg.max_position
        self.max_nb = (self.max_model_len + block_size - 1) // block_size
        self.eos_token_id = eos_token_id
        self.use_graphs = not enforce_eager
        self.max_num_batched_tokens = max(max_num_batched_tokens, max_num_seqs)

        # weights first, then size the block pool from what is left (gpu_memory_utilization of the whole device)
        free0, total = torch.cuda.mem_get_info(self.dev)
        self.model = LlamaModel(cfg, self.dev, num_blocks=1, block_size=block_size, seed=seed, weights_dir=weights_dir,
                                max_position=self.max_model_len, kv_dtype=self.kv_dtype, tp=self.tp)
        del self.model.kv
        if self.tp is not None:
            # peer-mapped exchange buffers (csrc/comm.cu): partial sums of up to max(max_num_seqs, 256) rows take the
            # one-shot path; the gathered logits hold one row per sequence
            self.tp.setup_comm(max(max_num_seqs, 256), cfg.hidden, max_num_seqs if self.model.vocab_parallel else 0,
                               cfg.vocab if self.model.vocab_parallel else 0)
        if num_blocks is None:
            util = gpu_memory_utilization if gpu_memory_utilization and gpu_memory_utilization > 0 else 0.9
            torch.cuda.synchronize(self.dev)
            free1, _ = torch.cuda.mem_get_info(self.dev)
            used = total - free1
            act = self.max_num_batched_tokens * (3 * cfg.inter + self.model.qkv_width + 6 * cfg.hidden) * 2 \
                + max_num_seqs * cfg.vocab * 2 * 2 + (1 << 30)
            budget = int(total * util) - used - act
            per_block = cfg.kv_bytes_per_token(1 if self.kv_dtype == torch.uint8 else 2, self.tp_size) * block_size
            want = max_num_seqs * self.max_nb + 64                   # every sequence at max_model_len
            num_blocks = max(min(budget // per_block, want), 0)
            if self.tp is not None:
                num_blocks = self.tp.min_int(num_blocks)             # every rank runs the same scheduler on the same pool
            if num_blocks < self.max_nb + 1:
                raise RuntimeError(f"KV pool of {num_blocks} blocks cannot hold one sequence of max_model_len "
                                   f"{self.max_model_len} (gpu_memory_utilization {util}, {free1 >> 20} MiB free)")
        if int(num_blocks) < self.max_nb + 1:
            # also for --num-gpu-blocks-override: a pool that cannot hold one max_model_len sequence would end long
            # requests early (the scheduler cuts their budget to the pool) -- refuse it at start-up instead
            raise RuntimeError(f"KV pool of {num_blocks} blocks cannot hold one sequence of max_model_len "
                               f"{self.max_model_len} ({self.max_nb + 1} blocks of {block_size} tokens needed)")
        self.num_blocks = int(num_blocks)
        self.model.num_blocks = self.num_blocks
        self.model.kv = torch.zeros(cfg.layers, 2, self.num_blocks, self.model.Hkv, block_size, cfg.d,
                                    dtype=self.kv_dtype, device=self.dev)
        self.sched = Scheduler(self.num_blocks, block_size, max_num_seqs, self.max_model_len, self.max_num_batched_tokens)

        # pinned staging + static device inputs of the decode graph
        B = max_num_seqs
        self._h_i64 = torch.zeros(3 * B, dtype=torch.int64).pin_memory()          # tokens | slots | seeds
        self._h_i32 = torch.zeros(2 * B, dtype=torch.int32).pin_memory()          # positions | seq_lens
        self._h_bt = torch.zeros(B, self.max_nb, dtype=torch.int32).pin_memory()
        self._h_temp = torch.zeros(2 * B, dtype=torch.float32).pin_memory()       # temperature | top_p
        self._h_topk = torch.zeros(B, dtype=torch.int32).pin_memory()
        self._h_out = torch.zeros(max(B, 1), dtype=torch.int64).pin_memory()
        self._d_i64 = torch.zeros(3 * B, dtype=torch.int64, device=self.dev)
        self._d_i32 = torch.zeros(2 * B, dtype=torch.int32, device=self.dev)
        self._d_bt = torch.zeros(B, self.max_nb, dtype=torch.int32, device=self.dev)
        self._d_temp = torch.zeros(2 * B, dtype=torch.float32, device=self.dev)
        self._d_topk = torch.zeros(B, dtype=torch.int32, device=self.dev)
        self._d_out = torch.zeros(B, dtype=torch.int64, device=self.dev)
        self._np_i64 = self._h_i64.numpy()
        self._np_i32 = self._h_i32.numpy()
        self._np_bt = self._h_bt.numpy()
        self._np_temp = self._h_temp.numpy()
        self._np_topk = self._h_topk.numpy()
        self._graphs: Dict[Tuple[int, int], torch.cuda.CUDAGraph] = {}
        self._buckets = _buckets(B)
        self._stream = torch.cuda.Stream(self.dev)
        self.time_device = False                 # bench: CUDA events around every decode replay -> stats.device_ms
        self._ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        self.model.prepare_decode(B, self.max_nb)     # workspace sized once for max_num_seqs + per-layer decode plans
        self._launches_per_layer = 7          # rms, rope+cache, attention(+merge), rms, silu, (GEMMs are cuBLAS)
        # what the decode-schedule planner sees: this rank's heads
        self._plan_heads = (self.model.Hq, self.model.Hkv)

        self.requests: Dict[int, Request] = {}
        self._ids = itertools.count(1)
        self._lock = threading.Lock()
        self._incoming: List[Request] = []
        self._aborts: List[int] = []
        self.stats = EngineStats()
        self.after_step: Optional[Callable[[], None]] = None    # engine thread: called once at the end of every step()
        self.metrics: Dict[str, int] = {}
        self._publish_metrics()
        torch.cuda.synchronize(self.dev)

    # ------------------------------------------------------------------------------------ requests
    def add_request(self, prompt_ids: List[int], params: SamplingParams,
                    on_token: Optional[Callable[[Request, Optional[int]], None]] = None) -> Request:
        """Thread-safe: queue a request; the engine thread admits it at its next step."""
        if not prompt_ids:
            raise ValueError("empty prompt")
        if len(prompt_ids) >= self.max_model_len:
            raise ValueError(f"prompt of {len(prompt_ids)} tokens leaves no room under max_model_len {self.max_model_len}")
        bad = [t for t in prompt_ids if not 0 <= t < self.cfg.vocab]
        if bad:
            raise ValueError(f"token id {bad[0]} outside the vocabulary [0, {self.cfg.vocab})")
        if params.max_tokens < 1:
            raise ValueError("max_tokens must be >= 1")
        rid = next(self._ids)
        r = Request(rid, list(prompt_ids), params, on_token=on_token, t_arrival=time.perf_counter(),
                    seed=params.seed if params.seed is not None else (rid * 0x9E3779B1) & 0x7FFFFFFF)
        with self._lock:
            self._incoming.append(r)
        return r

    def abort(self, rid: int) -> None:
        with self._lock:
            self._aborts.append(rid)

    def has_work(self) -> bool:
        with self._lock:
            if self._incoming:
                return True
        return self.sched.num_running > 0 or self.sched.num_waiting > 0

    def _finish(self, r: Request, reason: str) -> None:
        r.finished, r.finish_reason, r.t_done = True, reason, time.perf_counter()
        try:
            self.sched.finish(r.id)
        except Exception:
            pass                        # already released by the scheduler (length finish reported earlier)
        self.requests.pop(r.id, None)
        if r.on_token:
            r.on_token(r, None)

    def _publish_metrics(self) -> None:
        """Engine thread only: a fresh dict swapped in whole, read by the HTTP thread (/metrics)."""
        s = self.stats
        self.metrics = {"running": self.sched.num_running, "waiting": self.sched.num_waiting,
                        "free_blocks": self.sched.num_free_blocks, "prompt_tokens": s.prompt_tokens,
                        "generated_tokens": s.generated_tokens, "steps": s.steps, "preemptions": s.preemptions,
                        "mixed_steps": s.mixed_steps, "busy_s": round(s.busy_s, 4), "sync_wait_s": round(s.sync_wait_s, 4),
                        "mixed_busy_s": round(s.mixed_busy_s, 4)}

    # ------------------------------------------------------------------------------------ one step
    def _admit(self) -> None:
        if self.is_driver:
            with self._lock:
                incoming, self._incoming = self._incoming, []
                aborts, self._aborts = self._aborts, []
            if self.tp is not None:                       # tell the other ranks what this step admits / aborts
                from . import tp as _tp
                self.tp.channel.send(_tp.encode_step(incoming, aborts))
        else:
            incoming, aborts = self._worker_msg
            self._worker_msg = None
        for r in incoming:
            try:
                self.sched.add_request(r.id, len(r.prompt_ids), r.params.max_tokens)
                self.requests[r.id] = r
            except Exception as e:                                   # cannot ever fit the pool
                r.finished, r.finish_reason, r.t_done = True, f"error: {e}", time.perf_counter()
                if r.on_token:
                    r.on_token(r, None)
        for rid in aborts:
            r = self.requests.get(rid)
            if r is not None:
                self._finish(r, "abort")

    def _step_seed(self, r: Request) -> int:
        return (r.seed * 1000003 + len(r.output_ids)) & 0x7FFFFFFFFFFFFFFF

    def _ensure_mixed_buffers(self) -> None:
        if hasattr(self, "_hm_tok"):
            return
        Tm, B = self.max_num_batched_tokens + self.max_num_seqs, self.max_num_seqs
        pin = lambda n, dt: torch.zeros(n, dtype=dt).pin_memory()                      # noqa: E731
        self._hm_tok, self._hm_slot, self._hm_pos = pin(Tm, torch.int64), pin(Tm, torch.int64), pin(Tm, torch.int32)
        self._hm_aux = pin(2 * B, torch.int64)                 # seeds [B] | sample rows [B]
        self._hm_meta = pin(3 * B + 1, torch.int32)            # decode seq_lens [B] | cu_q [B + 1] | ctx_start [B]
        self._hm_temp = pin(2 * B, torch.float32)             # temperature [B] | top_p [B]
        self._hm_topk = pin(B, torch.int32)
        self._hm_pbt = torch.zeros(B, self.max_nb, dtype=torch.int32).pin_memory()

    def step(self) -> List[Tuple[Request, int]]:
        """Admit queued requests, run one scheduler plan on the GPU, return [(request, new token)]."""
        try:
            return self._step()
        finally:
            if self.after_step is not None:
                self.after_step()

    def _step(self) -> List[Tuple[Request, int]]:
        t0 = time.perf_counter()
        self._admit()
        plan = self.sched.step()
        for rid in plan.preempted_ids:
            r = self.requests.get(int(rid))
            if r is not None:
                r.n_preempted += 1
                self.stats.preemptions += 1
        for rid in plan.finished_ids:
            # normally the engine has already finished these itself (same length rule); a request the scheduler had
            # to end (its KV outgrew the whole pool) is still live here: close it
            r = self.requests.get(int(rid))
            if r is not None:
                r.finished, r.finish_reason, r.t_done = True, "length", time.perf_counter()
                self.requests.pop(r.id, None)
                if r.on_token:
                    r.on_token(r, None)
        if plan.empty:
            self._publish_metrics()
            return []
        nd, npf = len(plan.decode_ids), len(plan.prefill_ids)
        if npf == 0:
            sampled_rows, tokens = self._run_decode(plan)
            self.stats.decode_steps += 1
        else:
            tm = time.perf_counter()
            sampled_rows, tokens = self._run_mixed(plan)
            self.stats.mixed_steps += 1
            self.stats.mixed_busy_s += time.perf_counter() - tm
        now = time.perf_counter()
        out: List[Tuple[Request, int]] = []
        for r, tok in zip(sampled_rows, tokens):
            tok = int(tok)
            r.output_ids.append(tok)
            r.token_times.append(now)
            if r.t_first_token is None:
                r.t_first_token = now
            out.append((r, tok))
            self.stats.generated_tokens += 1
            if r.on_token:
                r.on_token(r, tok)
            p = r.params
            if not p.ignore_eos and ((self.eos_token_id is not None and tok == self.eos_token_id) or tok in p.stop_token_ids):
                self._finish(r, "stop")
            elif len(r.output_ids) >= p.max_tokens or len(r.prompt_ids) + len(r.output_ids) >= self.max_model_len:
                self._finish(r, "length")
        self.stats.steps += 1
        self.stats.busy_s += time.perf_counter() - t0
        self._publish_metrics()
        return out

    # ------------------------------------------------------------------------------------ decode-only step
    def _decode_graph(self, bucket: int, num_splits: int):
        key = (bucket, num_splits)
        g = self._graphs.get(key)
        if g is not None:
            return g
        B = self.max_num_seqs

        def run():
            logits = self.model.forward(self._d_i64[:bucket], self._d_i32[:bucket], self._d_i64[B:B + bucket], bucket,
                                        self._d_bt[:bucket], self._d_i32[B:B + bucket], None, None, num_splits)
            ops.sample(logits, self._d_temp[:bucket], self._d_i64[2 * B:2 * B + bucket], out=self._d_out[:bucket],
                       top_k=self._d_topk[:bucket], top_p=self._d_temp[B:B + bucket])

        if not self.use_graphs:
            self._graphs[key] = run
            return run
        with torch.cuda.stream(self._stream):
            # the warm-up run executes for real: point every row at "no slot, empty sequence" so it cannot touch the cache
            self._d_i64[:bucket].zero_()
            self._d_i64[B:B + bucket].fill_(-1)
            self._d_i32[:bucket].zero_()
            self._d_i32[B:B + bucket].zero_()
            run()                                   # warm-up outside capture (cuBLAS handles, function attributes)
            self._stream.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self._stream):
                run()
        self._graphs[key] = g.replay
        return g.replay

    def _run_decode(self, plan: StepPlan):
        nd, B = len(plan.decode_ids), self.max_num_seqs
        bucket = next(b for b in self._buckets if b >= nd)
        reqs = [self.requests[int(i)] for i in plan.decode_ids]
        i64, i32 = self._np_i64, self._np_i32
        sl = plan.seq_lens
        i64[:bucket] = 0
        i64[:nd] = [r.output_ids[-1] for r in reqs]
        i64[B:B + bucket] = -1
        i64[B:B + nd] = plan.slot_mapping
        i64[2 * B:2 * B + nd] = [self._step_seed(r) for r in reqs]
        i32[:bucket] = 0
        i32[:nd] = sl - 1
        i32[B:B + bucket] = 0
        i32[B:B + nd] = sl
        used = (int(sl.max()) + self.block_size - 1) // self.block_size
        self._np_bt[:nd, :used] = plan.block_tables[:, :used]
        self._np_temp[:bucket] = 0
        self._np_temp[:nd] = [r.params.temperature for r in reqs]
        self._np_temp[B:B + bucket] = 1.0
        self._np_temp[B:B + nd] = [r.params.top_p for r in reqs]
        self._np_topk[:bucket] = 0
        self._np_topk[:nd] = [r.params.top_k for r in reqs]
        num_splits = ops.plan_num_splits(self._h_i32[B:B + nd], *self._plan_heads, self.cfg.d, self.block_size,
                                         ops.KV_FP8_E4M3 if self.kv_dtype == torch.uint8 else ops.KV_BF16)
        replay = self._decode_graph(bucket, num_splits)
        with torch.cuda.stream(self._stream):
            self._d_i64.copy_(self._h_i64, non_blocking=True)
            self._d_i32.copy_(self._h_i32, non_blocking=True)
            self._d_bt[:nd, :used].copy_(self._h_bt[:nd, :used], non_blocking=True)
            self._d_temp.copy_(self._h_temp, non_blocking=True)
            self._d_topk.copy_(self._h_topk, non_blocking=True)
            if self.time_device:
                self._ev[0].record(self._stream)
            replay()
            if self.time_device:
                self._ev[1].record(self._stream)
            self._h_out[:bucket].copy_(self._d_out[:bucket], non_blocking=True)
        tw = time.perf_counter()
        self._stream.synchronize()
        self.stats.sync_wait_s += time.perf_counter() - tw
        if self.time_device:
            self.stats.device_ms += self._ev[0].elapsed_time(self._ev[1])
            self.stats.device_steps += 1
        self.stats.gpu_launches += self.cfg.layers * self._launches_per_layer + 3
        return reqs, self._h_out[:nd].tolist()

    # ------------------------------------------------------------------------------------ step with prompt chunks
    def _run_mixed(self, plan: StepPlan):
        self._ensure_mixed_buffers()
        nd, npf, B = len(plan.decode_ids), len(plan.prefill_ids), self.max_num_seqs
        Tp = int(plan.prefill_len.sum())
        T = nd + Tp
        tok, slot, pos = self._hm_tok.numpy(), self._hm_slot.numpy(), self._hm_pos.numpy()
        aux, meta = self._hm_aux.numpy(), self._hm_meta.numpy()
        dreqs = [self.requests[int(i)] for i in plan.decode_ids]
        preqs = [self.requests[int(i)] for i in plan.prefill_ids]
        if nd:
            tok[:nd] = [r.output_ids[-1] for r in dreqs]
            slot[:nd] = plan.slot_mapping
            pos[:nd] = plan.seq_lens - 1
            meta[:nd] = plan.seq_lens
        slot[nd:T] = plan.prefill_slots
        cu, ctx = meta[B:B + npf + 1], meta[2 * B + 1:2 * B + 1 + npf]
        cu[0] = 0
        sample_reqs, rows = list(dreqs), list(range(nd))
        off = nd
        for c, r in enumerate(preqs):
            st, n = int(plan.prefill_start[c]), int(plan.prefill_len[c])
            ids = r.prompt_ids if st + n <= len(r.prompt_ids) else r.prompt_ids + r.output_ids
            tok[off:off + n] = ids[st:st + n]
            pos[off:off + n] = np.arange(st, st + n, dtype=np.int32)
            off += n
            cu[c + 1] = off - nd
            ctx[c] = st
            if plan.prefill_last[c] == 1:
                sample_reqs.append(r)
                rows.append(off - 1)
            if st == 0 and r.n_preempted == 0:
                self.stats.prompt_tokens += len(r.prompt_ids)
        ns = len(rows)
        aux[:ns] = [self._step_seed(r) for r in sample_reqs]
        aux[B:B + ns] = rows
        if ns == 0:
            aux[B] = 0                       # no row to sample this step: lm_head runs on one dummy row
        self._hm_temp.numpy()[:ns] = [r.params.temperature for r in sample_reqs]
        self._hm_temp.numpy()[B:B + ns] = [r.params.top_p for r in sample_reqs]
        self._hm_topk.numpy()[:ns] = [r.params.top_k for r in sample_reqs]
        bsz = self.block_size
        used_p = (int((plan.prefill_start + plan.prefill_len).max()) + bsz - 1) // bsz
        self._hm_pbt.numpy()[:npf, :used_p] = plan.prefill_block_tables[:, :used_p]
        num_splits, used = 0, 0
        if nd:
            used = (int(plan.seq_lens.max()) + bsz - 1) // bsz
            self._np_bt[:nd, :used] = plan.block_tables[:, :used]
            num_splits = ops.plan_num_splits(self._hm_meta[:nd], *self._plan_heads, self.cfg.d, bsz,
                                                 ops.KV_FP8_E4M3 if self.kv_dtype == torch.uint8 else ops.KV_BF16)
        dev = self.dev
        with torch.cuda.stream(self._stream):
            d_tok = self._hm_tok[:T].to(dev, non_blocking=True)
            d_slot = self._hm_slot[:T].to(dev, non_blocking=True)
            d_pos = self._hm_pos[:T].to(dev, non_blocking=True)
            d_aux = self._hm_aux.to(dev, non_blocking=True)
            d_meta = self._hm_meta.to(dev, non_blocking=True)
            d_temp = self._hm_temp.to(dev, non_blocking=True)
            d_topk = self._hm_topk.to(dev, non_blocking=True)
            d_pbt = self._hm_pbt[:npf].to(dev, non_blocking=True)
            dec_bt = dec_sl = None
            if nd:
                self._d_bt[:nd, :used].copy_(self._h_bt[:nd, :used], non_blocking=True)
                dec_bt, dec_sl = self._d_bt[:nd], d_meta[:nd]
            pf = PrefillBatch(cu_q=d_meta[B:B + npf + 1], ctx_start=d_meta[2 * B + 1:2 * B + 1 + npf],
                              block_tables=d_pbt, total_tokens=Tp,
                              kv_lens_host=(plan.prefill_start + plan.prefill_len).tolist())
            logits = self.model.forward(d_tok, d_pos, d_slot, nd, dec_bt, dec_sl, pf, d_aux[B:B + max(ns, 1)], num_splits)
            if ns:
                out = ops.sample(logits, d_temp[:ns], d_aux[:ns], top_k=d_topk[:ns], top_p=d_temp[B:B + ns])
                self._h_out[:ns].copy_(out, non_blocking=True)
        tw = time.perf_counter()
        self._stream.synchronize()
        self.stats.sync_wait_s += time.perf_counter() - tw
        self.stats.gpu_launches += self.cfg.layers * (self._launches_per_layer + 1) + 3
        return sample_reqs, (self._h_out[:ns].tolist() if ns else [])

    def capture_all(self) -> int:
        """Capture the decode graph of every (batch bucket, schedule) now instead of at first use (server start-up).
        Tensor parallel: rank 0's call makes every rank capture (the warm-up runs exchange over NVLink in lockstep)."""
        if self.tp is not None and self.is_driver:
            from . import tp as _tp
            self.tp.channel.send(_tp.encode_step([], [], _tp.MSG_CAPTURE))
        n = 0
        if self.use_graphs:
            for b in self._buckets:
                for splits in (1, 0):
                    self._decode_graph(b, splits)
                    n += 1
        return n

    # ------------------------------------------------------------------------------------ tensor-parallel ranks
    def worker_loop(self) -> bool:
        """Ranks 1..tp-1: mirror rank 0 step by step.  Each message carries what rank 0 admitted / aborted before its
        step; the plan, the schedule, the CUDA graph and the sampled ids then follow from identical state.
        Returns True after rank 0's shutdown message (every rank then closes together), False when the launcher
        disappeared -- the caller must then leave without any collective (close() holds a barrier)."""
        from . import tp as _tp
        assert self.tp is not None and not self.is_driver
        parent = os.getppid()
        while True:
            try:
                payload = self.tp.channel.recv(timeout=1.0)
            except TimeoutError:
                if os.getppid() != parent:            # the launcher is gone: do not keep a GPU spinning
                    return False
                continue
            cmd, reqs, aborts = _tp.decode_step(payload)
            if cmd == _tp.MSG_SHUTDOWN:
                return True
            if cmd == _tp.MSG_CAPTURE:
                self.capture_all()
                continue
            incoming = [Request(rid, list(prompt), params, seed=seed) for rid, prompt, params, seed in reqs]
            self._worker_msg = (incoming, aborts)
            self.step()

    def shutdown_workers(self) -> None:
        """Rank 0: release the other ranks from worker_loop()."""
        if self.tp is not None and self.is_driver:
            from . import tp as _tp
            self.tp.channel.send(_tp.encode_step([], [], _tp.MSG_SHUTDOWN))

    def close(self) -> None:
        """Drop the CUDA graphs and the peer-mapped buffers (tensor parallel: call on every rank)."""
        self._graphs.clear()
        if self.tp is not None:
            self.tp.close()

    # ------------------------------------------------------------------------------------ loops
    def run_until_idle(self, max_steps: int = 10 ** 9) -> None:
        for _ in range(max_steps):
            if not self.has_work():
                return
            self.step()

    def generate(self, prompts: List[List[int]], params: SamplingParams) -> List[List[int]]:
        """Offline batch API: returns the generated ids of every prompt."""
        reqs = [self.add_request(p, params) for p in prompts]
        self.run_until_idle()
        return [r.output_ids for r in reqs]
