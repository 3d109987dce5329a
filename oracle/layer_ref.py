"""CPU oracle for the rows either side of the decode hot path:
prompt-chunk attention over the paged cache, RMSNorm, SwiGLU, rotary + KV write, token choice, and a whole
random-init Llama forward built from them.  TEST INFRASTRUCTURE ONLY.

    *** parity: unpinned by the reference ***
    The reference (ThakiCloud/vllm-backend-public) holds no code, test or fixture for any of these: they run
    inside the vLLM server image it launches (reference benchmark-vllm/vllm_manager.py:371-374).  This file
    restates the published semantics of the corresponding vLLM v0.9.1 ops / HF Llama modules in plain fp32
    PyTorch:
      prefill_attention_ref  flash_attn_varlen_func with block_table (chunked prefill / prefix attention)
      rms_norm_ref           csrc/layernorm_kernels.cu rms_norm / fused_add_rms_norm
      silu_and_mul_ref       csrc/activation_kernels.cu silu_and_mul
      rope_and_cache_ref     csrc/pos_encoding_kernels.cu rotary_embedding (neox) + cache_kernels.cu reshape_and_cache
      sample_ref             greedy / Gumbel-max draw (our own counter hash; documented in include/b200attn.h)
    rms_norm / silu / rotary are additionally checked against torch / transformers implementations in
    tests/test_layer_oracle.py (transformers' LlamaRMSNorm, apply_rotary_pos_emb, F.silu).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

import math
from typing import List, Optional

import numpy as np
import torch

from paged_attention_ref import (KV_BF16, KV_FP8_E4M3, reshape_and_cache_ref, _load_cache_f32,  # noqa: F401
                                 e4m3_decode, e4m3_encode)


def _bf16_round(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16)


# --------------------------------------------------------------------------------------
# prompt-chunk attention over the paged cache
# --------------------------------------------------------------------------------------
def prefill_attention_ref(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_tables: torch.Tensor,
                          cu_q: torch.Tensor, ctx_start: torch.Tensor, scale: Optional[float] = None,
                          return_f32: bool = False, kv_dtype: int = KV_BF16, k_scale: float = 1.0,
                          v_scale: float = 1.0) -> torch.Tensor:
    """q [T, Hq, d] bf16; chunk c = rows cu_q[c]:cu_q[c+1] at positions ctx_start[c] + i; token i attends cache
    positions j <= ctx_start[c] + i of block_tables[c].  fp32 softmax and contractions, one rounding to bf16.
    fp8-e4m3 caches (uint8) are read as float(e4m3) * scale, exactly (paged_attention_ref._load_cache_f32)."""
    T, Hq, d = q.shape
    nb, Hkv, bs, _ = k_cache.shape
    G = Hq // Hkv
    if scale is None:
        scale = 1.0 / math.sqrt(d)
    out = torch.zeros(T, Hq, d, dtype=torch.float32)
    kc = _load_cache_f32(k_cache, kv_dtype, k_scale)
    vc = _load_cache_f32(v_cache, kv_dtype, v_scale)
    for c in range(ctx_start.numel()):
        r0, r1 = int(cu_q[c]), int(cu_q[c + 1])
        n = r1 - r0
        if n == 0:
            continue
        start = int(ctx_start[c])
        total = start + n
        nblk = (total + bs - 1) // bs
        blocks = block_tables[c, :nblk].long()
        K = kc[blocks].permute(1, 0, 2, 3).reshape(Hkv, nblk * bs, d)[:, :total]      # [Hkv, total, d]
        V = vc[blocks].permute(1, 0, 2, 3).reshape(Hkv, nblk * bs, d)[:, :total]
        Q = q[r0:r1].to(torch.float32).permute(1, 0, 2)                                  # [Hq, n, d]
        K = K.repeat_interleave(G, dim=0)
        V = V.repeat_interleave(G, dim=0)
        S = torch.matmul(Q, K.transpose(1, 2)) * scale                                   # [Hq, n, total]
        pos = start + torch.arange(n).unsqueeze(1)
        mask = torch.arange(total).unsqueeze(0) > pos                                    # [n, total]
        S = S.masked_fill(mask.unsqueeze(0), float("-inf"))
        P = torch.softmax(S, dim=-1)
        out[r0:r1] = torch.matmul(P, V).permute(1, 0, 2)
    return out if return_f32 else out.to(torch.bfloat16)


# --------------------------------------------------------------------------------------
# element-wise ops of a decoder layer
# --------------------------------------------------------------------------------------
def rms_norm_ref(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None):
    """Returns (out bf16, new_residual bf16 or None).  h = bf16(x + residual) when residual is given."""
    h = x.to(torch.float32)
    new_res = None
    if residual is not None:
        new_res = _bf16_round(h + residual.to(torch.float32))
        h = new_res.to(torch.float32)
    var = (h * h).mean(dim=-1, keepdim=True)
    out = weight.to(torch.float32) * (h * torch.rsqrt(var + eps))
    return _bf16_round(out), new_res


def silu_and_mul_ref(gate_up: torch.Tensor) -> torch.Tensor:
    I = gate_up.shape[-1] // 2
    g = gate_up[..., :I].to(torch.float32)
    u = gate_up[..., I:].to(torch.float32)
    return _bf16_round(g / (1.0 + torch.exp(-g)) * u)


def rope_table(max_pos: int, d: int, theta: float = 500000.0, scaling: Optional[dict] = None) -> torch.Tensor:
    """fp32 [max_pos, d] = cos | sin halves; inv_freq = theta^(-2i/d), with Llama-3.1 style frequency scaling when
    `scaling` = {factor, low_freq_factor, high_freq_factor, original_max_position_embeddings}."""
    inv = 1.0 / (theta ** (torch.arange(0, d, 2, dtype=torch.float64) / d))
    if scaling:
        f, lo, hi = scaling["factor"], scaling["low_freq_factor"], scaling["high_freq_factor"]
        old = scaling["original_max_position_embeddings"]
        wavelen = 2 * math.pi / inv
        smooth = ((old / wavelen) - lo) / (hi - lo)
        scaled = torch.where(wavelen > old / lo, inv / f, inv)
        mid = (wavelen <= old / lo) & (wavelen >= old / hi)
        scaled = torch.where(mid, (1 - smooth) * inv / f + smooth * inv, scaled)
        inv = scaled
    ang = torch.arange(max_pos, dtype=torch.float64).unsqueeze(1) * inv.unsqueeze(0)
    return torch.cat([torch.cos(ang), torch.sin(ang)], dim=1).to(torch.float32).contiguous()


def rope_and_cache_ref(qkv: torch.Tensor, positions: torch.Tensor, cos_sin: torch.Tensor, k_cache: torch.Tensor,
                       v_cache: torch.Tensor, slot_mapping: torch.Tensor, Hq: int, kv_dtype: int = KV_BF16,
                       k_scale: float = 1.0, v_scale: float = 1.0) -> torch.Tensor:
    """Returns the rotated qkv (bf16, new tensor) and writes K', V into the caches (in place)."""
    T = qkv.shape[0]
    nb, Hkv, bs, d = k_cache.shape
    half = d // 2
    x = qkv.to(torch.float32).view(T, Hq + 2 * Hkv, d).clone()
    cs = cos_sin[positions.long()]                       # [T, d]
    cos = cs[:, :half].unsqueeze(1)
    sin = cs[:, half:].unsqueeze(1)
    rot = x[:, : Hq + Hkv]
    x1, x2 = rot[..., :half], rot[..., half:]
    o1 = x1 * cos - x2 * sin                             # each product and the sum rounded in fp32 (no fma on CPU)
    o2 = x2 * cos + x1 * sin
    x[:, : Hq + Hkv] = torch.cat([o1, o2], dim=-1)
    out = _bf16_round(x)
    k_new = out[:, Hq: Hq + Hkv].contiguous()
    v_new = out[:, Hq + Hkv:].contiguous()
    reshape_and_cache_ref(k_new, v_new, k_cache, v_cache, slot_mapping, kv_dtype, k_scale, v_scale)
    return out.view(T, -1)


# --------------------------------------------------------------------------------------
# token choice
# --------------------------------------------------------------------------------------
def _hash24(seed: int, idx: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + np.uint64(0x9E3779B97F4A7C15) * (idx.astype(np.uint64) + np.uint64(1))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(40)).astype(np.uint32)


def sample_scores_ref(logits: torch.Tensor, temperature: Optional[torch.Tensor], seeds: Optional[torch.Tensor]):
    """fp32 scores whose argmax (lowest index on ties) is the chosen token."""
    B, V = logits.shape
    s = logits.to(torch.float32).clone()
    for b in range(B):
        T = float(temperature[b]) if temperature is not None else 0.0
        if T > 0:
            seed = int(seeds[b]) if seeds is not None else 0
            u = (_hash24(seed, np.arange(V)).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 16777216.0)
            g = -np.log(-np.log(u.astype(np.float64)))
            s[b] = s[b] * np.float32(1.0 / T) + torch.from_numpy(g.astype(np.float32))
    return s


def sample_candidates_ref(logits: torch.Tensor, temperature, top_k=None, top_p=None, slack: float = 0.0):
    """bool [B, V]: the tokens a draw may return under top-k then top-p (vLLM's order; engine contract in
    include/b200attn.h::b2_sample_filtered).  Row b: t_k = the top_k-th largest logit (ties kept); masses
    w_i = exp((l_i - max) / T) over {l >= t_k}; the smallest set of top logit VALUES (ties kept) whose mass reaches
    top_p * (1 - slack) of that total.  `slack` > 0 narrows and < 0 widens the nucleus, so that a test can bound a
    kernel that sums the masses in another order from both sides."""
    B, V = logits.shape
    lf = logits.to(torch.float64)
    keep = torch.ones(B, V, dtype=torch.bool)
    for b in range(B):
        T = float(temperature[b]) if temperature is not None else 0.0
        if T <= 0:
            continue
        row = lf[b]
        k = int(top_k[b]) if top_k is not None else 0
        p = float(top_p[b]) if top_p is not None else 1.0
        m = torch.ones(V, dtype=torch.bool)
        if 0 < k < V:
            kth = torch.sort(row, descending=True).values[k - 1]
            m = row >= kth
        if p < 1.0:
            w = torch.where(m, torch.exp((row - row[m].max()) / T), torch.zeros_like(row))
            vals = torch.unique(row[m])                              # ascending
            mass_ge = torch.stack([w[row >= v].sum() for v in vals])  # non-increasing in v
            target = p * (1.0 - slack) * w.sum()
            ok = mass_ge >= target
            t = vals[ok].max()                                       # largest value whose top set still reaches the target
            m = m & (row >= t)
        keep[b] = m
    return keep


def sample_ref(logits, temperature=None, seeds=None, top_k=None, top_p=None) -> torch.Tensor:
    s = sample_scores_ref(logits, temperature, seeds)
    if top_k is not None or top_p is not None:
        s = torch.where(sample_candidates_ref(logits, temperature, top_k, top_p), s, torch.full_like(s, -float("inf")))
    # torch.argmax does not promise the first index on ties; do it explicitly
    mx = s.max(dim=-1, keepdim=True).values
    first = torch.where(s == mx, torch.arange(s.shape[1]).expand_as(s), torch.full_like(s, s.shape[1], dtype=torch.long))
    return first.min(dim=-1).values


# --------------------------------------------------------------------------------------
# whole-model restatement: a Llama decoder over ONE sequence, dense causal attention (no paging), fp32 math with the
# activations rounded to bf16 where the engine's tensors are bf16.  HF LlamaForCausalLM semantics
# (transformers/models/llama/modeling_llama.py); checked against that class in tests/test_layer_oracle.py.
# weights: the engine's dict (embed, l{i}.ln1/wqkv[/bqkv]/wo/ln2/wgu/wdown, norm, lm_head) as CPU tensors.
# --------------------------------------------------------------------------------------
def llama_forward_ref(w: dict, token_ids, *, layers: int, Hq: int, Hkv: int, d: int, eps: float,
                      cos_sin: torch.Tensor, kv_fp8_scales=None) -> torch.Tensor:
    """Returns fp32 logits [T, vocab] for positions 0..T-1 of one sequence.
    kv_fp8_scales = (k_scale, v_scale): K (after rotary) and V pass through the e4m3 cache encoding
    (sat_rne(x / scale), read back as float(e4m3) * scale), as an engine with an fp8 KV cache stores them."""
    f32 = lambda t: t.to(torch.float32)                              # noqa: E731
    ids = torch.as_tensor(token_ids, dtype=torch.long)
    T = ids.numel()
    half = d // 2
    G = Hq // Hkv
    resid = w["embed"][ids].to(torch.bfloat16)
    cos = cos_sin[:T, :half].unsqueeze(1)
    sin = cos_sin[:T, half:].unsqueeze(1)
    causal = torch.arange(T).unsqueeze(0) > torch.arange(T).unsqueeze(1)
    x, _ = rms_norm_ref(resid, w["l0.ln1"], eps)
    for l in range(layers):
        if l > 0:
            x, resid = rms_norm_ref(x, w[f"l{l}.ln1"], eps, resid)
        qkv = f32(x) @ f32(w[f"l{l}.wqkv"]).T
        if f"l{l}.bqkv" in w:
            qkv = qkv + f32(w[f"l{l}.bqkv"])
        qkv = f32(_bf16_round(qkv)).view(T, Hq + 2 * Hkv, d)
        rot = qkv[:, : Hq + Hkv]
        x1, x2 = rot[..., :half], rot[..., half:]
        rot = f32(_bf16_round(torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], dim=-1)))
        q = rot[:, :Hq].permute(1, 0, 2)
        k_tok, v_tok = rot[:, Hq:], qkv[:, Hq + Hkv:]
        if kv_fp8_scales is not None:
            ks, vs = kv_fp8_scales
            k_tok = e4m3_decode(e4m3_encode(k_tok / ks)) * ks
            v_tok = e4m3_decode(e4m3_encode(v_tok / vs)) * vs
        k = k_tok.permute(1, 0, 2).repeat_interleave(G, 0)
        v = v_tok.permute(1, 0, 2).repeat_interleave(G, 0)
        s = (q @ k.transpose(1, 2)) / math.sqrt(d)
        s = s.masked_fill(causal.unsqueeze(0), float("-inf"))
        attn = _bf16_round((torch.softmax(s, -1) @ v).permute(1, 0, 2).reshape(T, Hq * d))
        o = _bf16_round(f32(attn) @ f32(w[f"l{l}.wo"]).T)
        x, resid = rms_norm_ref(o, w[f"l{l}.ln2"], eps, resid)
        gu = _bf16_round(f32(x) @ f32(w[f"l{l}.wgu"]).T)
        act = silu_and_mul_ref(gu)
        x = _bf16_round(f32(act) @ f32(w[f"l{l}.wdown"]).T)
    x, _ = rms_norm_ref(x, w["norm"], eps, resid)
    return f32(x) @ f32(w["lm_head"]).T
