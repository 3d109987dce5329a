"""Llama-family decoder on the b200attn ops: the caller of the hot path.

What the engine launched by the reference (reference benchmark-vllm/vllm_manager.py:371-374, flags from
benchmark-vllm/vllm_templates.py:50-87) computes for every scheduled token.  The projections are plain library
GEMMs (torch.matmul -> cuBLAS); everything between them is this repo's sm_100a kernels behind the C ABI:
b2_rms_norm (fused residual add), b2_rope_and_cache (rotary + KV write in one pass over the fused QKV rows),
b2_paged_decode / b2_prefill_attention, b2_silu_and_mul, b2_sample.  There is no CPU path.

Weights: random-init from a seed (BASELINE.json configs[4] is random-init; the box has no network), or a local
HF directory (config.json + *.safetensors) when `model` is a path.  Geometry comes from the registry below or
from that config.json.  head_dim must be 128 (what the attention kernels are built for).
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass, field
from pathlib import Path
from typing import Dict, List, Optional

import torch

from . import ops


@dataclass
class ModelConfig:
    name: str
    hidden: int
    layers: int
    Hq: int
    Hkv: int
    inter: int
    vocab: int
    d: int = 128
    rope_theta: float = 500000.0
    rope_scaling: Optional[dict] = None
    eps: float = 1e-5
    qkv_bias: bool = False
    tie_embeddings: bool = False
    max_position: int = 131072

    @property
    def params(self) -> int:
        per_layer = self.hidden * (self.Hq + 2 * self.Hkv) * self.d + self.Hq * self.d * self.hidden \
            + 3 * self.hidden * self.inter + 2 * self.hidden
        emb = self.vocab * self.hidden * (1 if self.tie_embeddings else 2)
        return self.layers * per_layer + emb + self.hidden

    def kv_bytes_per_token(self, kv_elem_bytes: int = 2, tp: int = 1) -> int:
        return 2 * self.layers * (self.Hkv // tp) * self.d * kv_elem_bytes


_LLAMA3_SCALING = dict(factor=8.0, low_freq_factor=1.0, high_freq_factor=4.0, original_max_position_embeddings=8192)

REGISTRY: Dict[str, ModelConfig] = {
    "llama-3-8b": ModelConfig("llama-3-8b", 4096, 32, 32, 8, 14336, 128256, max_position=8192),
    "llama-3.1-8b": ModelConfig("llama-3.1-8b", 4096, 32, 32, 8, 14336, 128256, rope_scaling=_LLAMA3_SCALING),
    "llama-3-70b": ModelConfig("llama-3-70b", 8192, 80, 64, 8, 28672, 128256, max_position=8192),
    "qwen2-1.5b": ModelConfig("qwen2-1.5b", 1536, 28, 12, 2, 8960, 151936, rope_theta=1e6, eps=1e-6,
                              qkv_bias=True, tie_embeddings=True, max_position=32768),
    "qwen2-7b": ModelConfig("qwen2-7b", 3584, 28, 28, 4, 18944, 152064, rope_theta=1e6, eps=1e-6, qkv_bias=True,
                            max_position=32768),
    # test-sized geometries (same head_dim, same op sequence)
    "tiny": ModelConfig("tiny", 512, 2, 4, 2, 1024, 1024, max_position=8192),
    "small": ModelConfig("small", 1024, 4, 8, 2, 2816, 32000, max_position=32768),
    "tiny8": ModelConfig("tiny8", 1024, 2, 16, 8, 2048, 2048, max_position=8192),      # 8 kv heads: shards over tp <= 8
}

_ALIASES = {
    "meta-llama/meta-llama-3-8b": "llama-3-8b", "meta-llama/meta-llama-3-8b-instruct": "llama-3-8b",
    "meta-llama/llama-3.1-8b": "llama-3.1-8b", "meta-llama/llama-3.1-8b-instruct": "llama-3.1-8b",
    "meta-llama/meta-llama-3-70b": "llama-3-70b", "meta-llama/meta-llama-3-70b-instruct": "llama-3-70b",
    "qwen/qwen2-1.5b-instruct": "qwen2-1.5b", "qwen/qwen2-1.5b": "qwen2-1.5b",
    "qwen/qwen2-7b-instruct": "qwen2-7b", "qwen/qwen2-7b": "qwen2-7b",
}


def config_from_hf_dir(path: Path) -> ModelConfig:
    c = json.loads((path / "config.json").read_text())
    Hq = c["num_attention_heads"]
    d = c.get("head_dim") or c["hidden_size"] // Hq
    scaling = c.get("rope_scaling")
    if scaling and scaling.get("rope_type", scaling.get("type")) != "llama3":
        scaling = None
    return ModelConfig(
        name=path.name, hidden=c["hidden_size"], layers=c["num_hidden_layers"], Hq=Hq,
        Hkv=c.get("num_key_value_heads", Hq), inter=c["intermediate_size"], vocab=c["vocab_size"], d=d,
        rope_theta=float(c.get("rope_theta", 10000.0)), rope_scaling=scaling, eps=float(c.get("rms_norm_eps", 1e-5)),
        qkv_bias=bool(c.get("attention_bias", False)) or "qwen2" in c.get("model_type", ""),
        tie_embeddings=bool(c.get("tie_word_embeddings", False)),
        max_position=int(c.get("max_position_embeddings", 8192)))


def resolve_config(model: str) -> ModelConfig:
    """`--model` of the launch contract -> geometry.  A local HF directory wins; else the registry (by short name or
    by the hub id the reference's configs use); unknown names fail loudly (nothing can be downloaded here)."""
    p = Path(model)
    if p.is_dir() and (p / "config.json").exists():
        return config_from_hf_dir(p)
    key = model.lower()
    key = _ALIASES.get(key, key)
    if key in REGISTRY:
        return REGISTRY[key]
    raise ValueError(f"unknown model {model!r}: not a local HF directory and not one of {sorted(REGISTRY)} "
                     f"(or hub ids {sorted(_ALIASES)})")


def rope_table(max_pos: int, d: int, theta: float, scaling: Optional[dict], device) -> torch.Tensor:
    """fp32 [max_pos, d] = cos | sin halves per position (fp64 angles), Llama-3.1 frequency scaling when given.
    Same construction as the oracle's table (oracle/layer_ref.py::rope_table); transformers' llama3 init agrees."""
    inv = 1.0 / (theta ** (torch.arange(0, d, 2, dtype=torch.float64) / d))
    if scaling:
        f, lo, hi = scaling["factor"], scaling["low_freq_factor"], scaling["high_freq_factor"]
        old = scaling["original_max_position_embeddings"]
        wavelen = 2 * math.pi / inv
        smooth = ((old / wavelen) - lo) / (hi - lo)
        scaled = torch.where(wavelen > old / lo, inv / f, inv)
        mid = (wavelen <= old / lo) & (wavelen >= old / hi)
        inv = torch.where(mid, (1 - smooth) * inv / f + smooth * inv, scaled)
    ang = torch.arange(max_pos, dtype=torch.float64).unsqueeze(1) * inv.unsqueeze(0)
    return torch.cat([torch.cos(ang), torch.sin(ang)], dim=1).to(torch.float32).contiguous().to(device)


@dataclass
class PrefillBatch:
    """Device-side description of the prompt chunks of one step (rows n_decode.. of the token batch)."""
    cu_q: torch.Tensor            # int32 [n + 1], relative to the first prefill row
    ctx_start: torch.Tensor       # int32 [n]
    block_tables: torch.Tensor    # int32 [n, max_nb]
    total_tokens: int
    kv_lens_host: Optional[list] = None   # [ctx_start[c] + len_c] as host ints (fp8 cache: sizes the bf16 scratch)


class LlamaModel:
    """Weights + KV cache + forward over one scheduled token batch.  One instance per GPU."""

    def __init__(self, cfg: ModelConfig, device: torch.device, num_blocks: int, block_size: int = 16,
                 seed: int = 0, weights_dir: Optional[Path] = None, max_position: Optional[int] = None,
                 kv_dtype: torch.dtype = torch.bfloat16, tp=None):
        """`tp`: a b200attn.tp.TPGroup (or None).  With tp.size > 1 this instance holds rank tp.rank's shard: q / kv
        heads and MLP columns split over the ranks (column-parallel wqkv / wgu, row-parallel wo / wdown), the
        lm_head split over the vocabulary, embedding and norms replicated; the KV pool holds Hkv / tp heads."""
        if cfg.d != 128:
            raise ValueError(f"head_dim {cfg.d}: the attention kernels are built for 128")
        if device.type != "cuda":
            raise ValueError("LlamaModel needs a CUDA device (b200attn has no CPU path)")
        self.cfg, self.dev, self.bs, self.num_blocks = cfg, device, block_size, num_blocks
        self.tp = tp if tp is not None and tp.size > 1 else None
        self.tp_size, self.tp_rank = (tp.size, tp.rank) if self.tp is not None else (1, 0)
        n = self.tp_size
        if cfg.Hkv % n or cfg.Hq % n or cfg.inter % (8 * n):
            raise ValueError(f"tensor_parallel_size {n} must divide Hkv={cfg.Hkv}, Hq={cfg.Hq} and inter={cfg.inter} / 8")
        self.Hq, self.Hkv, self.inter = cfg.Hq // n, cfg.Hkv // n, cfg.inter // n          # this rank's share
        self.vocab_parallel = n > 1 and cfg.vocab % (8 * n) == 0
        self._exchange = os.environ.get("B2_TP_EXCHANGE", "push")         # push | pull | nccl (A/B knob)
        self.max_position = max_position or cfg.max_position
        self.qkv_width = (self.Hq + 2 * self.Hkv) * cfg.d
        self.w: Dict[str, torch.Tensor] = {}
        if weights_dir is not None and any(weights_dir.glob("*.safetensors")):
            self._load_hf(weights_dir)
        else:
            self._random_init(seed)
        self.cos_sin = rope_table(self.max_position, cfg.d, cfg.rope_theta, cfg.rope_scaling, device)
        # KV pool: [layers, 2, num_blocks, Hkv, bs, d]; one (block, head) tile = bs x d contiguous ("HND")
        self.kv = torch.zeros(cfg.layers, 2, num_blocks, self.Hkv, block_size, cfg.d, dtype=kv_dtype, device=device)
        self.k_scale = self.v_scale = None
        if kv_dtype != torch.bfloat16:
            self.k_scale = torch.ones(1, dtype=torch.float32, device=device)
            self.v_scale = torch.ones(1, dtype=torch.float32, device=device)
        self._ws: Optional[torch.Tensor] = None
        self._plans: Optional[list] = None      # per-layer b2_decode_plan handles (prepare_decode)

    # ------------------------------------------------------------------ weights
    def _put(self, name: str, full: torch.Tensor) -> None:
        """Store this rank's shard of the full tensor `name` (the whole tensor at tp = 1).  Shards are taken from the
        full tensor, so every tp size holds exactly the weights of the tp = 1 model (same seed or same checkpoint)."""
        from .parallel import shard_weight
        c = self.cfg
        full = shard_weight(name.rsplit(".", 1)[-1], full, self.tp_size, self.tp_rank, c.Hq, c.Hkv, c.d, c.inter,
                            self.vocab_parallel)
        self.w[name] = full

    def _random_init(self, seed: int) -> None:
        c, dev = self.cfg, self.dev
        g = torch.Generator(device=dev).manual_seed(seed)

        def rnd(*shape, std=0.02):
            return (torch.randn(*shape, generator=g, device=dev, dtype=torch.float32) * std).to(torch.bfloat16)

        put = self._put
        put("embed", rnd(c.vocab, c.hidden))
        full_qkv = (c.Hq + 2 * c.Hkv) * c.d
        for l in range(c.layers):
            put(f"l{l}.ln1", torch.ones(c.hidden, dtype=torch.bfloat16, device=dev))
            put(f"l{l}.wqkv", rnd(full_qkv, c.hidden))
            if c.qkv_bias:
                put(f"l{l}.bqkv", rnd(full_qkv))
            put(f"l{l}.wo", rnd(c.hidden, c.Hq * c.d))
            put(f"l{l}.ln2", torch.ones(c.hidden, dtype=torch.bfloat16, device=dev))
            put(f"l{l}.wgu", rnd(2 * c.inter, c.hidden))
            put(f"l{l}.wdown", rnd(c.hidden, c.inter))
        put("norm", torch.ones(c.hidden, dtype=torch.bfloat16, device=dev))
        put("lm_head", self.w["embed"] if c.tie_embeddings else rnd(c.vocab, c.hidden))

    def _load_hf(self, path: Path) -> None:
        from safetensors import safe_open
        c, dev = self.cfg, self.dev
        raw: Dict[str, torch.Tensor] = {}
        for f in sorted(path.glob("*.safetensors")):
            with safe_open(str(f), framework="pt", device="cpu") as sf:
                for k in sf.keys():
                    raw[k] = sf.get_tensor(k)

        def get(name):
            return raw[name].to(torch.bfloat16).to(dev)

        put = self._put
        put("embed", get("model.embed_tokens.weight"))
        for l in range(c.layers):
            p = f"model.layers.{l}."
            put(f"l{l}.ln1", get(p + "input_layernorm.weight"))
            put(f"l{l}.wqkv", torch.cat([get(p + f"self_attn.{x}_proj.weight") for x in "qkv"]).contiguous())
            if c.qkv_bias:
                put(f"l{l}.bqkv", torch.cat([get(p + f"self_attn.{x}_proj.bias") for x in "qkv"]).contiguous())
            put(f"l{l}.wo", get(p + "self_attn.o_proj.weight"))
            put(f"l{l}.ln2", get(p + "post_attention_layernorm.weight"))
            put(f"l{l}.wgu", torch.cat([get(p + "mlp.gate_proj.weight"), get(p + "mlp.up_proj.weight")]).contiguous())
            put(f"l{l}.wdown", get(p + "mlp.down_proj.weight"))
        put("norm", get("model.norm.weight"))
        put("lm_head", self.w["embed"] if c.tie_embeddings or "lm_head.weight" not in raw else get("lm_head.weight"))

    def weight_bytes(self) -> int:
        seen, n = set(), 0
        for t in self.w.values():
            if t.data_ptr() not in seen:
                seen.add(t.data_ptr())
                n += t.numel() * t.element_size()
        return n

    # ------------------------------------------------------------------ forward
    def decode_workspace(self, max_batch: int, max_nb: int) -> torch.Tensor:
        """The decode scratch, allocated ONCE for the largest batch (b2_paged_decode_workspace_bytes serves every batch
        size up to it).  Captured CUDA graphs hold its address, so it is never replaced: a larger request is a bug."""
        need = max(ops.decode_workspace_bytes(max_batch, self.Hq, self.Hkv, self.cfg.d, self.bs, max_nb, 0), 256)
        if self._ws is None:
            self._ws = torch.zeros(need, dtype=torch.uint8, device=self.dev)
        elif self._ws.numel() < need:
            raise RuntimeError(f"decode workspace holds {self._ws.numel()} bytes but batch {max_batch} x {max_nb} blocks "
                               f"needs {need}: size it once for max_num_seqs / max_model_len (prepare_decode)")
        return self._ws

    def prepare_decode(self, max_batch: int, max_nb: int) -> None:
        """One b2_decode_plan per layer over that layer's K/V pool, all sharing the one workspace (the layers run one
        after the other on one stream): the TMA descriptors are encoded here, not in every step."""
        ws = self.decode_workspace(max_batch, max_nb)
        self._plans = [ops.DecodePlan(self.kv[l, 0], self.kv[l, 1], max_batch, self.Hq, max_nb,
                                      k_scale=self.k_scale, v_scale=self.v_scale, workspace=ws)
                       for l in range(self.cfg.layers)]
        self._plan_shape = (max_batch, max_nb)

    def forward(self, tokens: torch.Tensor, positions: torch.Tensor, slots: torch.Tensor, n_decode: int,
                dec_block_tables: Optional[torch.Tensor], dec_seq_lens: Optional[torch.Tensor],
                prefill: Optional[PrefillBatch], sample_rows: Optional[torch.Tensor], num_splits: int = 0) -> torch.Tensor:
        """tokens int64 [T] (rows [0, n_decode) = one new token per running sequence, the rest = prompt chunks),
        positions int32 [T], slots int64 [T].  Returns bf16 logits [len(sample_rows) or T, vocab]."""
        c, w = self.cfg, self.w
        Hq, Hkv = self.Hq, self.Hkv                                      # this rank's heads (all of them at tp = 1)
        T = tokens.shape[0]
        # tensor parallel: partial sums of up to 256 rows (decode steps, small mixed steps) take the fused one-shot NVLink
        # exchange -- "push" (no barrier) by default, "pull" (flag barrier + peer loads) up to tp.max_rows rows or when
        # B2_TP_EXCHANGE=pull; larger prompt chunks are bandwidth-bound and go through NCCL + the plain fused norm
        one_shot = None
        if self.tp is not None and self._exchange != "nccl":
            if T <= self.tp.PUSH_MAX_ROWS and self._exchange == "push":
                one_shot = "push"
            elif T <= self.tp.max_rows:
                one_shot = "pull"
        resid = w["embed"].index_select(0, tokens)                       # [T, hidden]  (residual stream, bf16)
        x = ops.rms_norm(resid, w["l0.ln1"], c.eps)
        attn = torch.empty(T, Hq, c.d, dtype=torch.bfloat16, device=self.dev)
        fp8_plan = None
        if self.k_scale is not None and prefill is not None and prefill.total_tokens:
            kv_lens = prefill.kv_lens_host
            if kv_lens is None:      # one device read; the engine passes the scheduler's host copy instead
                kv_lens = (prefill.ctx_start.long() + (prefill.cu_q[1:] - prefill.cu_q[:-1]).long()).cpu().tolist()
            src, sbt = ops.fp8_prefill_plan(prefill.block_tables, kv_lens, self.bs)
            shape = (src.numel(), Hkv, self.bs, c.d)
            fp8_plan = (src, sbt, torch.empty(shape, dtype=torch.bfloat16, device=self.dev),
                        torch.empty(shape, dtype=torch.bfloat16, device=self.dev))
        part = None
        for l in range(c.layers):
            qkv = torch.nn.functional.linear(x, w[f"l{l}.wqkv"], w.get(f"l{l}.bqkv"))
            kc, vc = self.kv[l, 0], self.kv[l, 1]
            ops.rope_and_cache(qkv, positions, self.cos_sin, kc, vc, slots, Hq, self.k_scale, self.v_scale,
                               write_k=False)          # K is read from the cache only
            q = qkv.view(T, Hq + 2 * Hkv, c.d)[:, :Hq]
            if n_decode:
                if self._plans is not None and n_decode <= self._plan_shape[0] \
                        and dec_block_tables.shape[1] == self._plan_shape[1]:
                    self._plans[l].run(q[:n_decode], dec_block_tables, dec_seq_lens, attn[:n_decode], num_splits)
                else:
                    ops.paged_decode(q[:n_decode], kc, vc, dec_block_tables, dec_seq_lens, k_scale=self.k_scale,
                                     v_scale=self.v_scale, out=attn[:n_decode],
                                     workspace=self.decode_workspace(n_decode, dec_block_tables.shape[1]),
                                     num_splits=num_splits)
            if prefill is not None and prefill.total_tokens:
                if fp8_plan is None:
                    ops.prefill_attention(q[n_decode:], kc, vc, prefill.block_tables, prefill.cu_q, prefill.ctx_start,
                                          total_tokens=prefill.total_tokens, out=attn[n_decode:])
                else:
                    # fp8 cache: the blocks the chunks can see, as bf16, in a scratch cache shared by all layers
                    ops.dequant_kv_blocks(kc, vc, fp8_plan[0], k_out=fp8_plan[2], v_out=fp8_plan[3])
                    ops.prefill_attention(q[n_decode:], fp8_plan[2], fp8_plan[3], fp8_plan[1], prefill.cu_q,
                                          prefill.ctx_start, total_tokens=prefill.total_tokens, out=attn[n_decode:],
                                          k_scale=self.k_scale, v_scale=self.v_scale)
            o = self._row_parallel(attn.view(T, Hq * c.d), w[f"l{l}.wo"], 0, one_shot)
            x = self._reduce_norm(o, 0, one_shot, w[f"l{l}.ln2"], resid)
            gu = torch.nn.functional.linear(x, w[f"l{l}.wgu"])
            act = ops.silu_and_mul(gu)
            part = self._row_parallel(act, w[f"l{l}.wdown"], 1, one_shot)
            if l + 1 < c.layers:
                x = self._reduce_norm(part, 1, one_shot, w[f"l{l + 1}.ln1"], resid)
        # final norm over the rows that are sampled.  One-shot path: reduce every row (T is small), then pick -- the
        # partial-sum slots must strictly alternate; NCCL path: pick first, then reduce only those rows
        if one_shot:
            x = self._reduce_norm(part, 1, one_shot, w["norm"], resid)
            if sample_rows is not None:
                x = x.index_select(0, sample_rows)
        else:
            if sample_rows is not None:
                part = part.index_select(0, sample_rows)
                resid = resid.index_select(0, sample_rows)
            x = self._reduce_norm(part, 1, None, w["norm"], resid)
        logits = torch.nn.functional.linear(x, w["lm_head"])
        if self.vocab_parallel:
            logits = self.tp.allgather_cols(logits)       # [rows, vocab] view of the peer-written buffer (valid until the next gather)
        return logits

    def _row_parallel(self, inp: torch.Tensor, weight: torch.Tensor, slot: int, one_shot) -> torch.Tensor:
        """Row-parallel projection (o_proj / down_proj): this rank's PARTIAL [T, hidden] sum.  For the pull exchange the
        GEMM writes straight into the peer-mapped partial-sum buffer `slot`."""
        if one_shot == "pull":
            out = self.tp.inputs[slot][: inp.shape[0]]
            torch.mm(inp, weight.t(), out=out)
            return out
        return torch.nn.functional.linear(inp, weight)

    def _reduce_norm(self, part: torch.Tensor, slot: int, one_shot, weight: torch.Tensor,
                     resid: torch.Tensor) -> torch.Tensor:
        """resid += sum over ranks of `part`; returns RMSNorm(resid) * weight."""
        if self.tp is None:
            return ops.rms_norm(part, weight, self.cfg.eps, residual=resid, out=part)
        if one_shot == "push":
            return self.tp.push_allreduce_rms_norm(part, weight, resid, self.cfg.eps)
        if one_shot == "pull":
            return self.tp.allreduce_rms_norm(slot, part.shape[0], weight, resid, self.cfg.eps)
        self.tp.all_reduce_nccl(part)
        return ops.rms_norm(part, weight, self.cfg.eps, residual=resid, out=part)
