"""CPU oracle for the paged-attention decode hot path.  TEST INFRASTRUCTURE ONLY.

    *** parity: unpinned by the reference, pinned against its dependency ***
    The reference (ThakiCloud/vllm-backend-public) contains no implementation, test,
    golden vector or fixture for this path: the arithmetic lives in the third-party
    vLLM server image it launches (`vllm/vllm-openai:v0.9.1`,
    reference benchmark-vllm/vllm_manager.py:371-374; `vllm-cpu-env:latest`,
    benchmark-vllm/vllm_templates.py:124).  vLLM is not vendored under /root/reference,
    so this file restates vLLM's *published* paged-attention semantics (v0.9.1
    `csrc/cache_kernels.cu::reshape_and_cache`, `csrc/attention/attention_kernels.cuh`)
    in plain fp32 PyTorch, as BASELINE.json's north_star prescribes.  The parameters
    that shape it do come from the reference: `block_size` (benchmark-vllm/models.py:10),
    `max_num_seqs` / `max_model_len` (models.py:9,15), `tensor_parallel_size`
    (models.py:11), `dtype` (models.py:14).
    What pins it: outputs of vLLM's own CUDA kernels (`reshape_and_cache`,
    `paged_attention_v1/_v2`; the vllm 0.22.0 wheel of the image, run on a B200 by
    tests/golden/make_vllm_crosscheck.py) committed as tests/golden/vllm_crosscheck.pt --
    the KV write (bf16 and fp8-e4m3) is bit-identical, decode agrees within
    4e-3 + 1e-2|ref| (vLLM rounds softmax weights to the cache dtype); see
    tests/test_vllm_crosscheck.py.  That wheel is not the reference's pinned v0.9.1 image.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module -- as the checker or the timed CPU baseline, never as a product
path.  The product (`b200attn`) has no CPU fallback and never imports it.

Semantics:
  * slot -> (block, offset) = (slot // bs, slot % bs); slot < 0 means "skip this token".
  * caches are [num_blocks, Hkv, bs, d]; bf16, or uint8 holding e4m3 bit patterns.
  * GQA: q head hq reads kv head hq // (Hq // Hkv).
  * only positions j < seq_lens[b] attend; seq_len == 0 gives a zero row.
  * softmax and both contractions in fp32; output rounded to bf16 exactly once.
  * fp8: store sat_rne_e4m3(x / scale) (clamp to +-448, round-to-nearest-even, NaN -> 0x7f);
    load float(e4m3) * scale.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

KV_BF16 = 0
KV_FP8_E4M3 = 1

E4M3_MAX = 448.0


# --------------------------------------------------------------------------------------
# e4m3 ("fn" flavour: no inf, 0x7f/0xff = NaN, max finite 448) encode / decode, written out
# with integer arithmetic so it does not depend on torch's own (non-saturating) fp8 cast.
# Matches PTX `cvt.rn.satfinite.e4m3x2.f32`.
# --------------------------------------------------------------------------------------
def e4m3_encode(x: torch.Tensor) -> torch.Tensor:
    """fp32 tensor -> uint8 e4m3 bit patterns; saturating, round-to-nearest-even."""
    x = x.to(torch.float32)
    bits = x.contiguous().view(torch.int32)
    sign = ((bits >> 24) & 0x80).to(torch.int32)
    a = x.abs()
    nan = torch.isnan(x)
    a = torch.clamp(a, max=E4M3_MAX)  # satfinite (also maps +inf to 448)
    a = torch.where(nan, torch.zeros_like(a), a)
    # Quantum of the target grid at |a|: normals have 3 mantissa bits (2^(e-3)),
    # everything below 2^-6 sits on the subnormal grid 2^-9.
    e = torch.floor(torch.log2(torch.clamp(a, min=2.0 ** -30)))
    # log2/floor can be off by one ulp at exact powers of two: fix up exactly.
    e = torch.where(torch.pow(2.0, e) > a, e - 1, e)
    e = torch.where(torch.pow(2.0, e + 1) <= a, e + 1, e)
    e = torch.clamp(e, min=-6.0, max=8.0)
    q = torch.pow(2.0, e - 3.0)
    # a / q is exact in fp32 (power-of-two scaling); torch.round is half-to-even.
    n = torch.round(a / q)  # in [0, 16]
    # n == 16 means we rounded up into the next binade.
    up = n >= 16
    n = torch.where(up, torch.full_like(n, 8.0), n)
    e = torch.where(up, e + 1, e)
    n_i = n.to(torch.int32)
    e_i = e.to(torch.int32)
    # normal: n in [8, 15] -> exponent field e+7, mantissa n-8; subnormal (e == -6, n < 8).
    is_norm = n_i >= 8
    code = torch.where(is_norm, ((e_i + 7) << 3) | (n_i - 8), n_i)
    code = torch.clamp(code, max=0x7E)
    code = torch.where(nan, torch.full_like(code, 0x7F), code)
    return (code | sign).to(torch.uint8)


def e4m3_decode(b: torch.Tensor) -> torch.Tensor:
    """uint8 e4m3 bit patterns -> fp32 (exact)."""
    b = b.to(torch.int32)
    sign = torch.where((b & 0x80) != 0, -1.0, 1.0).to(torch.float32)
    ef = (b >> 3) & 0xF
    m = (b & 0x7).to(torch.float32)
    norm = torch.pow(2.0, (ef - 7).to(torch.float32)) * (1.0 + m / 8.0)
    sub = m * (2.0 ** -9)
    v = torch.where(ef == 0, sub, norm)
    v = torch.where((b & 0x7F) == 0x7F, torch.full_like(v, float("nan")), v)
    return sign * v


# --------------------------------------------------------------------------------------
# a5: reshape_and_cache
# --------------------------------------------------------------------------------------
def reshape_and_cache_ref(
    k_new: torch.Tensor,          # [T, Hkv, d] bf16
    v_new: torch.Tensor,          # [T, Hkv, d] bf16
    k_cache: torch.Tensor,        # [num_blocks, Hkv, bs, d] bf16 | uint8(e4m3), modified in place
    v_cache: torch.Tensor,
    slot_mapping: torch.Tensor,   # [T] int64, -1 = skip
    kv_dtype: int = KV_BF16,
    k_scale: float = 1.0,
    v_scale: float = 1.0,
) -> None:
    """vLLM `reshape_and_cache` semantics: k_cache[s // bs, h, s % bs, :] = k_new[t, h, :].

    Later tokens win when two tokens name the same slot (vLLM leaves this a race; callers
    never do it, and the tests do not either)."""
    bs = k_cache.shape[2]
    slots = slot_mapping.to(torch.int64)
    keep = slots >= 0
    t_idx = torch.nonzero(keep, as_tuple=False).flatten()
    s = slots[t_idx]
    blk = s // bs
    off = s % bs
    if kv_dtype == KV_BF16:
        k_cache[blk, :, off, :] = k_new[t_idx].to(k_cache.dtype)
        v_cache[blk, :, off, :] = v_new[t_idx].to(v_cache.dtype)
    elif kv_dtype == KV_FP8_E4M3:
        ks = torch.tensor(k_scale, dtype=torch.float32)
        vs = torch.tensor(v_scale, dtype=torch.float32)
        k_cache[blk, :, off, :] = e4m3_encode(k_new[t_idx].to(torch.float32) / ks)
        v_cache[blk, :, off, :] = e4m3_encode(v_new[t_idx].to(torch.float32) / vs)
    else:
        raise ValueError(f"kv_dtype {kv_dtype}")


# --------------------------------------------------------------------------------------
# a6-a8: paged-attention decode
# --------------------------------------------------------------------------------------
def _load_cache_f32(cache: torch.Tensor, kv_dtype: int, scale: float) -> torch.Tensor:
    if kv_dtype == KV_BF16:
        return cache.to(torch.float32)
    return e4m3_decode(cache) * torch.tensor(scale, dtype=torch.float32)


def paged_decode_ref(
    q: torch.Tensor,              # [B, Hq, d] bf16
    k_cache: torch.Tensor,        # [num_blocks, Hkv, bs, d]
    v_cache: torch.Tensor,
    block_tables: torch.Tensor,   # [B, max_nb] int32
    seq_lens: torch.Tensor,       # [B] int32
    scale: Optional[float] = None,
    kv_dtype: int = KV_BF16,
    k_scale: float = 1.0,
    v_scale: float = 1.0,
    return_f32: bool = False,
) -> torch.Tensor:
    """out[b, hq] = softmax_j(scale * <q[b,hq], K[b, j, hq // G]>) . V[b, j, hq // G], j < seq_lens[b].

    Everything in fp32; the result is rounded to bf16 once (or returned in fp32)."""
    B, Hq, d = q.shape
    nb_total, Hkv, bs, d2 = k_cache.shape
    assert d2 == d and Hq % Hkv == 0
    G = Hq // Hkv
    if scale is None:
        scale = 1.0 / math.sqrt(d)
    out = torch.zeros(B, Hq, d, dtype=torch.float32)
    qf = q.to(torch.float32)
    for b in range(B):
        L = int(seq_lens[b])
        if L <= 0:
            continue
        nb = (L + bs - 1) // bs
        blocks = block_tables[b, :nb].to(torch.int64)
        # [nb, Hkv, bs, d] -> [Hkv, nb*bs, d] -> first L tokens
        K = _load_cache_f32(k_cache[blocks], kv_dtype, k_scale).permute(1, 0, 2, 3).reshape(Hkv, nb * bs, d)[:, :L]
        V = _load_cache_f32(v_cache[blocks], kv_dtype, v_scale).permute(1, 0, 2, 3).reshape(Hkv, nb * bs, d)[:, :L]
        qb = qf[b].reshape(Hkv, G, d)
        s = torch.einsum("hgd,hld->hgl", qb, K) * scale
        p = torch.softmax(s, dim=-1)
        out[b] = torch.einsum("hgl,hld->hgd", p, V).reshape(Hq, d)
    return out if return_f32 else out.to(torch.bfloat16)


def paged_decode_partitioned_ref(
    q, k_cache, v_cache, block_tables, seq_lens, partition: int,
    scale: Optional[float] = None, kv_dtype: int = KV_BF16, k_scale: float = 1.0, v_scale: float = 1.0,
) -> torch.Tensor:
    """a7: split-K restatement -- per-partition (m, l, acc) merged with log-sum-exp weights.

    Mathematically identical to `paged_decode_ref`; exists so the tests can check the merge
    algebra (empty partitions, -inf maxima) independently of the GPU."""
    B, Hq, d = q.shape
    _, Hkv, bs, _ = k_cache.shape
    G = Hq // Hkv
    if scale is None:
        scale = 1.0 / math.sqrt(d)
    out = torch.zeros(B, Hq, d, dtype=torch.float32)
    qf = q.to(torch.float32)
    for b in range(B):
        L = int(seq_lens[b])
        if L <= 0:
            continue
        nb = (L + bs - 1) // bs
        blocks = block_tables[b, :nb].to(torch.int64)
        K = _load_cache_f32(k_cache[blocks], kv_dtype, k_scale).permute(1, 0, 2, 3).reshape(Hkv, nb * bs, d)[:, :L]
        V = _load_cache_f32(v_cache[blocks], kv_dtype, v_scale).permute(1, 0, 2, 3).reshape(Hkv, nb * bs, d)[:, :L]
        qb = qf[b].reshape(Hkv, G, d)
        ms, ls, accs = [], [], []
        for t0 in range(0, L, partition):
            s = torch.einsum("hgd,hld->hgl", qb, K[:, t0:t0 + partition]) * scale
            m = s.max(dim=-1).values
            p = torch.exp(s - m[..., None])
            ms.append(m)
            ls.append(p.sum(-1))
            accs.append(torch.einsum("hgl,hld->hgd", p, V[:, t0:t0 + partition]))
        m_all = torch.stack(ms).max(dim=0).values
        num = sum(torch.exp(m - m_all)[..., None] * a for m, a in zip(ms, accs))
        den = sum(torch.exp(m - m_all) * l for m, l in zip(ms, ls))
        out[b] = (num / den[..., None]).reshape(Hq, d)
    return out.to(torch.bfloat16)


# --------------------------------------------------------------------------------------
# Seeded synthetic inputs.  Pure CPU; used by tests and bench.
# --------------------------------------------------------------------------------------
def make_inputs(
    B: int, Hq: int, Hkv: int, d: int, bs: int, ctx: int,
    seed: int = 0, ragged: bool = False, slack_blocks: int = 64,
) -> dict:
    """q/k/v ~ N(0,1) in fp32 -> bf16; block tables are a random permutation of the pool."""
    g = torch.Generator().manual_seed(seed)
    if ragged:
        seq_lens = torch.randint(1, ctx + 1, (B,), generator=g, dtype=torch.int32)
    else:
        seq_lens = torch.full((B,), ctx, dtype=torch.int32)
    max_nb = (ctx + bs - 1) // bs
    num_blocks = B * max_nb + slack_blocks
    perm = torch.randperm(num_blocks, generator=g).to(torch.int32)
    block_tables = torch.zeros(B, max_nb, dtype=torch.int32)
    slots = []
    tok_seq = []
    cursor = 0
    for b in range(B):
        nb = (int(seq_lens[b]) + bs - 1) // bs
        block_tables[b, :nb] = perm[cursor:cursor + nb]
        cursor += nb
        j = torch.arange(int(seq_lens[b]), dtype=torch.int64)
        slots.append(block_tables[b, j // bs].to(torch.int64) * bs + (j % bs))
        tok_seq.append(torch.full((int(seq_lens[b]),), b, dtype=torch.int64))
    slot_mapping = torch.cat(slots)
    T = slot_mapping.numel()
    q = torch.randn(B, Hq, d, generator=g, dtype=torch.float32).to(torch.bfloat16)
    k = torch.randn(T, Hkv, d, generator=g, dtype=torch.float32).to(torch.bfloat16)
    v = torch.randn(T, Hkv, d, generator=g, dtype=torch.float32).to(torch.bfloat16)
    return dict(q=q, k=k, v=v, block_tables=block_tables, seq_lens=seq_lens,
                slot_mapping=slot_mapping, num_blocks=num_blocks, max_nb=max_nb,
                B=B, Hq=Hq, Hkv=Hkv, d=d, bs=bs, ctx=ctx)


def algorithmic_bytes(B: int, Hq: int, Hkv: int, d: int, bs: int, seq_lens, kv_bytes: int) -> int:
    """K+V read once, q read, out written, block table, seq_lens."""
    total = 0
    for L in seq_lens:
        L = int(L)
        total += L * 2 * Hkv * d * kv_bytes + ((L + bs - 1) // bs) * 4
    total += 2 * B * Hq * d * 2 + B * 4
    return total
