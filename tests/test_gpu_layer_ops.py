"""GPU parity for the rows either side of the decode path, through the C ABI:
prompt-chunk attention over the paged cache, RMSNorm (+ fused residual add), SwiGLU, rotary + KV write, token choice.

Bars (written here because the north star only states the decode bound):
  * rope_and_cache: bit-exact (bf16 KV and the fp8-e4m3 encode), q/k rotated in place bit-exact.
  * rms_norm / silu_and_mul: within 1 bf16 ulp of the fp32 oracle (rsqrt / exp are hardware approximations):
    |d| <= 2^-7 * |ref| + 1e-6; the fused residual (a bf16 add) is bit-exact.
  * prefill attention: |d| <= 4e-3 + 1e-2 * |ref| (softmax weights are rounded to bf16 before P.V, the
    FlashAttention convention; the decode kernel's hi+lo split is not used here because the op is tensor-bound).
  * sample: greedy is exact; a Gumbel draw must pick a token whose oracle score is within 1e-3 of the oracle max.
"""
import pytest
import torch

import b200attn
from b200attn import ops
import layer_ref as L
import paged_attention_ref as R

pytestmark = pytest.mark.gpu


def bits(t):
    return t.contiguous().view(torch.uint8).cpu()


def make_prefill_case(lens, starts, Hq, Hkv, bs, seed=0, poison=True, d=128):
    """Sequences with `starts[c]` cached tokens and a chunk of `lens[c]` new ones; caches are filled for
    [0, start + len) through the oracle's reshape_and_cache and poisoned (NaN) elsewhere."""
    g = torch.Generator().manual_seed(seed)
    n = len(lens)
    totals = [s + l for s, l in zip(starts, lens)]
    max_nb = max((t + bs - 1) // bs for t in totals) + 1
    num_blocks = sum((t + bs - 1) // bs for t in totals) + 17
    perm = torch.randperm(num_blocks, generator=g).to(torch.int32)
    bt = torch.zeros(n, max_nb, dtype=torch.int32)
    fill = float("nan") if poison else 0.0
    kc = torch.full((num_blocks, Hkv, bs, d), fill, dtype=torch.bfloat16)
    vc = torch.full((num_blocks, Hkv, bs, d), fill, dtype=torch.bfloat16)
    cur = 0
    for c, t in enumerate(totals):
        nb = (t + bs - 1) // bs
        bt[c, :nb] = perm[cur:cur + nb]
        cur += nb
        j = torch.arange(t)
        slots = bt[c, j // bs].long() * bs + (j % bs)
        k = torch.randn(t, Hkv, d, generator=g).to(torch.bfloat16)
        v = torch.randn(t, Hkv, d, generator=g).to(torch.bfloat16)
        R.reshape_and_cache_ref(k, v, kc, vc, slots)
    T = sum(lens)
    q = torch.randn(T, Hq, d, generator=g).to(torch.bfloat16)
    cu = torch.zeros(n + 1, dtype=torch.int32)
    cu[1:] = torch.tensor(lens).cumsum(0)
    return q, kc, vc, bt, cu, torch.tensor(starts, dtype=torch.int32)


def check_prefill(case, dev, atol=4e-3, rtol=1e-2):
    q, kc, vc, bt, cu, st = case
    ref = L.prefill_attention_ref(q, kc, vc, bt, cu, st, return_f32=True)
    out = ops.prefill_attention(q.to(dev), kc.to(dev), vc.to(dev), bt.to(dev), cu.to(dev), st.to(dev)).float().cpu()
    assert torch.isfinite(out).all(), "non-finite output (uninitialised cache bytes reached the MMA?)"
    err = (out - ref).abs()
    bad = err > atol + rtol * ref.abs()
    assert not bad.any(), f"{int(bad.sum())}/{bad.numel()} outside tolerance; max err {float(err.max()):.3e}"
    return float(err.max())


@pytest.mark.parametrize("lens,starts", [
    ([1], [0]), ([5], [0]), ([128], [0]), ([129], [0]), ([300], [0]),
    ([64], [100]), ([200], [333]), ([1], [777]),
    ([17, 130, 1, 256], [0, 64, 5, 1000]),
    ([100, 0, 50], [3, 0, 61]),
])
def test_prefill_attention_small(cuda, lens, starts):
    check_prefill(make_prefill_case(lens, starts, Hq=8, Hkv=2, bs=16, seed=sum(lens)), cuda)


@pytest.mark.parametrize("bs", [8, 32])
def test_prefill_attention_block_sizes(cuda, bs):
    check_prefill(make_prefill_case([150, 70], [40, 0], Hq=4, Hkv=4, bs=bs, seed=bs), cuda)


@pytest.mark.parametrize("Hq,Hkv", [(1, 1), (4, 1), (16, 2), (8, 1)])
def test_prefill_attention_shard_geometries(cuda, Hq, Hkv):
    """What one rank sees at --tensor-parallel-size 8 / 4 (Llama-3-8B: 4 q heads on 1 kv head; tiny8: 1 on 1; 70B at
    tp 8: 8 on 1): few heads make small launches, i.e. the 128-row split kernel, also with a prefix that leaves slot 1
    of a short row without any visible key tile (ragged chunk ends, prefix not a multiple of the 128-key tile)."""
    check_prefill(make_prefill_case([513, 1, 130, 255], [0, 1000, 77, 3000], Hq=Hq, Hkv=Hkv, bs=16, seed=Hq * 7 + Hkv), cuda)


def test_prefill_attention_llama_geometry(cuda):
    """Llama-3-8B heads, a 2k chunk after a 1k prefix and a whole 700-token prompt."""
    check_prefill(make_prefill_case([2048, 700], [1024, 0], Hq=32, Hkv=8, bs=16, seed=3), cuda)


def test_prefill_matches_decode_on_last_token(cuda):
    """Size-independent property: the last row of a chunk equals paged decode over the same sequence."""
    q, kc, vc, bt, cu, st = make_prefill_case([513, 1200], [0, 4000], Hq=32, Hkv=8, bs=16, seed=5, poison=False)
    dev = cuda
    out = ops.prefill_attention(q.to(dev), kc.to(dev), vc.to(dev), bt.to(dev), cu.to(dev), st.to(dev))
    last = (cu[1:] - 1).long()
    seq_lens = (st + (cu[1:] - cu[:-1])).to(torch.int32)
    dec = ops.paged_decode(q[last].contiguous().to(dev), kc.to(dev), vc.to(dev), bt.to(dev), seq_lens.to(dev))
    a, b = out[last.to(dev)].float().cpu(), dec.float().cpu()
    assert ((a - b).abs() <= 4e-3 + 1e-2 * b.abs()).all()


def test_prefill_chunked_equals_whole(cuda):
    """Chunking a prompt (the scheduler's token budget) does not change any row beyond rounding."""
    q, kc, vc, bt, cu, st = make_prefill_case([1000], [0], Hq=8, Hkv=2, bs=16, seed=9, poison=False)
    dev = cuda
    whole = ops.prefill_attention(q.to(dev), kc.to(dev), vc.to(dev), bt.to(dev), cu.to(dev), st.to(dev)).float().cpu()
    parts = []
    for a, b in ((0, 300), (300, 812), (812, 1000)):
        cu2 = torch.tensor([0, b - a], dtype=torch.int32)
        st2 = torch.tensor([a], dtype=torch.int32)
        parts.append(ops.prefill_attention(q[a:b].contiguous().to(dev), kc.to(dev), vc.to(dev), bt.to(dev),
                                           cu2.to(dev), st2.to(dev)).float().cpu())
    got = torch.cat(parts)
    assert ((got - whole).abs() <= 4e-3 + 1e-2 * whole.abs()).all()


def test_prefill_rejects_cpu_and_fp8_without_scales(cuda):
    q, kc, vc, bt, cu, st = make_prefill_case([16], [0], Hq=4, Hkv=2, bs=16, poison=False)
    with pytest.raises(ValueError):
        ops.prefill_attention(q, kc, vc, bt, cu, st)
    dev = cuda
    k8 = torch.zeros(kc.shape, dtype=torch.uint8, device=dev)
    with pytest.raises(ValueError):        # an fp8 cache needs its scales
        ops.prefill_attention(q.to(dev), k8, k8.clone(), bt.to(dev), cu.to(dev), st.to(dev))


@pytest.mark.parametrize("lens,starts,bs", [([300], [0], 16), ([64, 1, 130], [100, 777, 5], 16), ([200], [333], 8),
                                            ([17, 256], [0, 1000], 32)])
@pytest.mark.parametrize("host_lens", [True, False])
def test_prefill_attention_fp8_cache(cuda, lens, starts, bs, host_lens):
    """Prompt chunks over an fp8-e4m3 cache: b2_dequant_kv_blocks into a bf16 scratch cache + the bf16 kernel, against
    the oracle reading the same quantised cache exactly.  Untouched cache bytes are the e4m3 NaN (0x7f)."""
    dev = cuda
    q, kb, vb, bt, cu, st = make_prefill_case(lens, starts, Hq=8, Hkv=2, bs=bs, poison=False, seed=3)
    ks, vs = 0.013, 0.021
    # quantise the filled bf16 caches token by token through the oracle's cache write; NaN-poison the rest
    k8 = torch.full(kb.shape, 0x7F, dtype=torch.uint8)
    v8 = torch.full(vb.shape, 0x7F, dtype=torch.uint8)
    for c, (s0, n) in enumerate(zip(starts, lens)):
        j = torch.arange(s0 + n)
        blk, off = bt[c, j // bs].long(), j % bs
        slots = blk * bs + off
        R.reshape_and_cache_ref(kb[blk, :, off], vb[blk, :, off], k8, v8, slots, R.KV_FP8_E4M3, ks, vs)
    ref = L.prefill_attention_ref(q, k8, v8, bt, cu, st, return_f32=True, kv_dtype=R.KV_FP8_E4M3, k_scale=ks, v_scale=vs)
    ksd = torch.tensor([ks], dtype=torch.float32, device=dev)
    vsd = torch.tensor([vs], dtype=torch.float32, device=dev)
    kv_lens = [s0 + n for s0, n in zip(starts, lens)] if host_lens else None
    out = ops.prefill_attention(q.to(dev), k8.to(dev), v8.to(dev), bt.to(dev), cu.to(dev), st.to(dev), k_scale=ksd,
                                v_scale=vsd, kv_lens_host=kv_lens).float().cpu()
    assert torch.isfinite(out).all(), "non-finite output (uninitialised cache bytes reached the MMA?)"
    err = (out - ref).abs()
    # the scratch cache holds float(e4m3) exactly and the scales are folded in fp32: the bf16 path's bound applies
    bad = err > 4e-3 + 1e-2 * ref.abs()
    assert not bad.any(), f"{int(bad.sum())}/{bad.numel()} outside tolerance; max err {float(err.max()):.3e}"


def test_dequant_kv_blocks_bit_exact(cuda):
    g = torch.Generator().manual_seed(0)
    k8 = torch.randint(0, 256, (40, 2, 16, 128), generator=g, dtype=torch.uint8)
    v8 = torch.randint(0, 256, (40, 2, 16, 128), generator=g, dtype=torch.uint8)
    k8[(k8 & 0x7F) == 0x7F] = 0x10            # no NaN codes: compare values
    v8[(v8 & 0x7F) == 0x7F] = 0x10
    src = torch.tensor([7, 0, 39, 7, 21], dtype=torch.int32)
    ko, vo = ops.dequant_kv_blocks(k8.to(cuda), v8.to(cuda), src.to(cuda))
    k_ref, v_ref = R.e4m3_decode(k8[src.long()]), R.e4m3_decode(v8[src.long()])
    assert torch.equal(ko.cpu().float(), k_ref) and torch.equal(vo.cpu().float(), v_ref)     # exact, every code


# ------------------------------------------------------------------------------------------
# element-wise ops
# ------------------------------------------------------------------------------------------
def ulp_close(got, ref):
    got, ref = got.float().cpu(), ref.float().cpu()
    assert torch.isfinite(got).all()
    err = (got - ref).abs()
    assert (err <= 2.0 ** -7 * ref.abs() + 1e-6).all(), f"max err {float(err.max()):.3e}"


@pytest.mark.parametrize("T,H", [(1, 4096), (64, 4096), (7, 8192), (33, 256), (5, 2048), (3, 16384)])
@pytest.mark.parametrize("with_res", [False, True])
def test_rms_norm(cuda, T, H, with_res):
    g = torch.Generator().manual_seed(T + H)
    x = (torch.randn(T, H, generator=g) * 2).to(torch.bfloat16)
    w = (1 + 0.1 * torch.randn(H, generator=g)).to(torch.bfloat16)
    res = torch.randn(T, H, generator=g).to(torch.bfloat16) if with_res else None
    ref, ref_res = L.rms_norm_ref(x, w, 1e-5, res)
    d_res = res.to(cuda) if with_res else None
    out = ops.rms_norm(x.to(cuda), w.to(cuda), 1e-5, residual=d_res)
    ulp_close(out, ref)
    if with_res:
        assert torch.equal(bits(d_res), bits(ref_res)), "fused residual add is a bf16 add: must be bit-exact"


def test_rms_norm_strided_and_inplace(cuda):
    g = torch.Generator().manual_seed(1)
    big = torch.randn(9, 3 * 512, generator=g).to(torch.bfloat16).to(cuda)
    x = big[:, 512:1024]
    w = torch.ones(512, dtype=torch.bfloat16, device=cuda)
    ref, _ = L.rms_norm_ref(x.cpu(), w.cpu(), 1e-6)
    ops.rms_norm(x, w, 1e-6, out=x)
    ulp_close(x, ref)


@pytest.mark.parametrize("T,I", [(1, 14336), (64, 14336), (5, 64), (300, 1024)])
def test_silu_and_mul(cuda, T, I):
    g = torch.Generator().manual_seed(T)
    gu = (torch.randn(T, 2 * I, generator=g) * 3).to(torch.bfloat16)
    ulp_close(ops.silu_and_mul(gu.to(cuda)), L.silu_and_mul_ref(gu))


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("Hq,Hkv,bs", [(32, 8, 16), (8, 2, 8), (4, 4, 32)])
def test_rope_and_cache_bit_exact(cuda, Hq, Hkv, bs, fp8):
    d, T, nblk = 128, 75, 40
    g = torch.Generator().manual_seed(Hq + bs)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * d, generator=g).to(torch.bfloat16)
    pos = torch.randint(0, 5000, (T,), generator=g, dtype=torch.int32)
    slots = torch.randperm(nblk * bs, generator=g)[:T].to(torch.int64)
    slots[::9] = -1
    table = L.rope_table(5000, d, theta=500000.0,
                         scaling=dict(factor=8.0, low_freq_factor=1.0, high_freq_factor=4.0,
                                      original_max_position_embeddings=8192))
    dt = torch.uint8 if fp8 else torch.bfloat16
    kc_ref = torch.zeros(nblk, Hkv, bs, d, dtype=dt)
    vc_ref = torch.zeros_like(kc_ref)
    ks, vs = 0.013, 0.021
    ref = L.rope_and_cache_ref(qkv, pos, table, kc_ref, vc_ref, slots, Hq,
                               R.KV_FP8_E4M3 if fp8 else R.KV_BF16, ks, vs)
    kc = torch.zeros(nblk, Hkv, bs, d, dtype=dt, device=cuda)
    vc = torch.zeros_like(kc)
    dq = qkv.to(cuda)
    sc = dict(k_scale=torch.tensor([ks], device=cuda), v_scale=torch.tensor([vs], device=cuda)) if fp8 else {}
    ops.rope_and_cache(dq, pos.to(cuda), table.to(cuda), kc, vc, slots.to(cuda), Hq, **sc)
    assert torch.equal(bits(dq), bits(ref)), "rotated q/k (v untouched) must be bit-exact"
    assert torch.equal(bits(kc), bits(kc_ref)) and torch.equal(bits(vc), bits(vc_ref)), "cache write must be bit-exact"


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("Hq,Hkv,write_k", [(32, 8, True), (32, 8, False), (64, 8, True), (4, 1, False)])
def test_rope_and_cache_streaming_kernel_bit_exact(cuda, Hq, Hkv, write_k, fp8):
    """T large enough for the persistent, software-pipelined kernel (>= 2 x 4 CTAs x SMs tokens); (64, 8) takes the
    3-items-per-thread instantiation.  write_k=False: the k part of qkv stays as it was, the cache is the same."""
    d, bs, T = 128, 16, 2600
    nblk = (T + bs - 1) // bs + 8
    g = torch.Generator().manual_seed(Hq)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * d, generator=g).to(torch.bfloat16)
    pos = torch.randint(0, 5000, (T,), generator=g, dtype=torch.int32)
    slots = torch.randperm(nblk * bs, generator=g)[:T].to(torch.int64)
    slots[::11] = -1
    table = L.rope_table(5000, d, theta=500000.0, scaling=None)
    dt = torch.uint8 if fp8 else torch.bfloat16
    kc_ref = torch.zeros(nblk, Hkv, bs, d, dtype=dt)
    vc_ref = torch.zeros_like(kc_ref)
    ks, vs = 0.013, 0.021
    ref = L.rope_and_cache_ref(qkv, pos, table, kc_ref, vc_ref, slots, Hq, R.KV_FP8_E4M3 if fp8 else R.KV_BF16, ks, vs)
    kc = torch.zeros(nblk, Hkv, bs, d, dtype=dt, device=cuda)
    vc = torch.zeros_like(kc)
    dq = qkv.to(cuda)
    sc = dict(k_scale=torch.tensor([ks], device=cuda), v_scale=torch.tensor([vs], device=cuda)) if fp8 else {}
    ops.rope_and_cache(dq, pos.to(cuda), table.to(cuda), kc, vc, slots.to(cuda), Hq, write_k=write_k, **sc)
    want = ref.clone()
    if not write_k:
        want.view(T, Hq + 2 * Hkv, d)[:, Hq:Hq + Hkv] = qkv.view(T, Hq + 2 * Hkv, d)[:, Hq:Hq + Hkv]
    assert torch.equal(bits(dq), bits(want)), "rotated q (and k when asked) must be bit-exact"
    assert torch.equal(bits(kc), bits(kc_ref)) and torch.equal(bits(vc), bits(vc_ref)), "cache write must be bit-exact"


def test_sample_greedy_exact(cuda):
    g = torch.Generator().manual_seed(0)
    logits = torch.randn(9, 128256, generator=g).to(torch.bfloat16)
    logits[3, 100] = logits[3, 7000] = 50.0          # tie: lowest index wins
    logits[4, 128255] = 60.0                          # last element
    ref = L.sample_ref(logits)
    assert ref[3] == 100 and ref[4] == 128255
    assert torch.equal(ops.sample(logits.to(cuda)).cpu(), ref)
    small = torch.randn(3, 1003, generator=g).to(torch.bfloat16)      # V not a multiple of 8 needs an aligned stride
    padded = torch.zeros(3, 1008, dtype=torch.bfloat16)
    padded[:, :1003] = small
    got = ops.sample(padded.to(cuda)[:, :1003])
    assert torch.equal(got.cpu(), L.sample_ref(small))


def test_sample_gumbel(cuda):
    g = torch.Generator().manual_seed(1)
    B, V = 16, 32000
    logits = (torch.randn(B, V, generator=g) * 2).to(torch.bfloat16)
    temp = torch.full((B,), 0.8)
    temp[::4] = 0.0                                    # mixed greedy / sampled rows
    seeds = torch.arange(B, dtype=torch.int64) * 7919 + 13
    got = ops.sample(logits.to(cuda), temp.to(cuda), seeds.to(cuda)).cpu()
    scores = L.sample_scores_ref(logits, temp, seeds)
    picked = scores.gather(1, got.unsqueeze(1)).squeeze(1)
    assert (picked >= scores.max(dim=1).values - 1e-3).all()
    assert torch.equal(got[::4], L.sample_ref(logits)[::4])
    # a different seed gives a different draw somewhere; the same seed repeats
    again = ops.sample(logits.to(cuda), temp.to(cuda), seeds.to(cuda)).cpu()
    other = ops.sample(logits.to(cuda), temp.to(cuda), (seeds + 1).to(cuda)).cpu()
    assert torch.equal(again, got) and not torch.equal(other, got)


def test_sample_distribution(cuda):
    """Gumbel-max draws follow softmax(logits / T): chi-square-ish check on a 6-way distribution."""
    V = 8
    logits = torch.tensor([2.0, 1.0, 0.0, -1.0, 0.5, 1.5, -30.0, -30.0]).to(torch.bfloat16)
    N = 20000
    rows = logits.unsqueeze(0).repeat(N, 1).contiguous()
    temp = torch.full((N,), 1.0)
    seeds = torch.arange(N, dtype=torch.int64) * 2654435761 + 17
    got = ops.sample(rows.to(cuda), temp.to(cuda), seeds.to(cuda)).cpu()
    freq = torch.bincount(got, minlength=V).float() / N
    p = torch.softmax(logits.float(), 0)
    assert (freq - p).abs().max() < 0.015


def test_sample_top_k_top_p(cuda):
    """b2_sample_filtered against the oracle's candidate sets (oracle/layer_ref.py::sample_candidates_ref).  The kernel
    sums the softmax masses in another order than the oracle, so the nucleus is bounded from both sides: every draw lies
    in the oracle's set widened by 1e-3 of the mass, and is the best-scoring token of a set no smaller than the one
    narrowed by 1e-3; top-k alone (integer counts) is exact."""
    g = torch.Generator().manual_seed(5)
    B, V = 24, 32000
    logits = (torch.randn(B, V, generator=g) * 2.5).to(torch.bfloat16)
    temp = torch.full((B,), 0.9)
    temp[::6] = 0.0
    seeds = torch.arange(B, dtype=torch.int64) * 104729 + 3
    top_k = torch.tensor([0, 1, 5, 50, 1000, V, 7, 0] * 3, dtype=torch.int32)
    top_p = torch.tensor([1.0, 1.0, 0.5, 0.9, 0.3, 0.99, 1.0, 0.7] * 3, dtype=torch.float32)
    got = ops.sample(logits.to(cuda), temp.to(cuda), seeds.to(cuda), top_k=top_k.to(cuda), top_p=top_p.to(cuda)).cpu()
    scores = L.sample_scores_ref(logits, temp, seeds)
    loose = L.sample_candidates_ref(logits, temp, top_k, top_p, slack=-1e-3)      # nucleus target raised: a superset
    tight = L.sample_candidates_ref(logits, temp, top_k, top_p, slack=1e-3)
    assert loose.gather(1, got.unsqueeze(1)).all(), "a draw fell outside the candidate set"
    best_tight = torch.where(tight, scores, torch.full_like(scores, -float("inf"))).max(dim=1).values
    picked = scores.gather(1, got.unsqueeze(1)).squeeze(1)
    assert (picked >= best_tight - 1e-3).all(), "a better-scoring candidate was passed over"
    # top-k only: exact
    only_k = ops.sample(logits.to(cuda), temp.to(cuda), seeds.to(cuda), top_k=top_k.to(cuda)).cpu()
    ref_k = L.sample_ref(logits, temp, seeds, top_k=top_k)
    sk = scores.gather(1, only_k.unsqueeze(1)).squeeze(1)
    rk = scores.gather(1, ref_k.unsqueeze(1)).squeeze(1)
    assert (sk >= rk - 1e-3).all() and L.sample_candidates_ref(logits, temp, top_k, None).gather(1, only_k.unsqueeze(1)).all()
    # no filter = b2_sample bit for bit; top_k = 1 = greedy
    plain = ops.sample(logits.to(cuda), temp.to(cuda), seeds.to(cuda)).cpu()
    nofilt = ops.sample(logits.to(cuda), temp.to(cuda), seeds.to(cuda), top_k=torch.zeros(B, dtype=torch.int32, device=cuda),
                        top_p=torch.ones(B, device=cuda)).cpu()
    assert torch.equal(plain, nofilt)
    k1 = ops.sample(logits.to(cuda), torch.ones(B, device=cuda), seeds.to(cuda),
                    top_k=torch.ones(B, dtype=torch.int32, device=cuda)).cpu()
    assert torch.equal(k1, L.sample_ref(logits))


def test_sample_top_p_distribution(cuda):
    """Draws under top_p follow the renormalised nucleus: p = 0.8 over softmax([2, 1.5, 1, .5, 0, -1]) keeps 3 tokens (0.83)."""
    logits = torch.tensor([2.0, 1.0, 0.0, -1.0, 0.5, 1.5, -30.0, -30.0]).to(torch.bfloat16)
    N, V = 20000, 8
    rows = logits.unsqueeze(0).repeat(N, 1).contiguous()
    seeds = torch.arange(N, dtype=torch.int64) * 2654435761 + 17
    got = ops.sample(rows.to(cuda), torch.ones(N, device=cuda), seeds.to(cuda),
                     top_p=torch.full((N,), 0.8, device=cuda)).cpu()
    keep = L.sample_candidates_ref(logits.unsqueeze(0), torch.ones(1), None, torch.tensor([0.8]))[0]
    p = torch.softmax(logits.float(), 0) * keep
    p = p / p.sum()
    freq = torch.bincount(got, minlength=V).float() / N
    assert int(keep.sum()) == 3 and (freq[~keep] == 0).all() and (freq - p).abs().max() < 0.015
