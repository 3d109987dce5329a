"""CPU: the layer oracle (oracle/layer_ref.py) against independent implementations -- torch SDPA, transformers'
Llama modules (LlamaRMSNorm, apply_rotary_pos_emb, llama3 rope init), F.silu -- and against the decode oracle."""
import math

import pytest
import torch
import torch.nn.functional as F

import layer_ref as L
import paged_attention_ref as R


def _case(lens, starts, Hq=4, Hkv=2, bs=16, d=128, seed=0):
    g = torch.Generator().manual_seed(seed)
    n = len(lens)
    totals = [s + l for s, l in zip(starts, lens)]
    max_nb = max((t + bs - 1) // bs for t in totals) + 1
    num_blocks = sum((t + bs - 1) // bs for t in totals) + 5
    perm = torch.randperm(num_blocks, generator=g).to(torch.int32)
    bt = torch.zeros(n, max_nb, dtype=torch.int32)
    kc = torch.full((num_blocks, Hkv, bs, d), float("nan"), dtype=torch.bfloat16)
    vc = torch.full_like(kc, float("nan"))
    ks, vs, cur = [], [], 0
    for c, t in enumerate(totals):
        nb = (t + bs - 1) // bs
        bt[c, :nb] = perm[cur:cur + nb]
        cur += nb
        j = torch.arange(t)
        slots = bt[c, j // bs].long() * bs + (j % bs)
        k = torch.randn(t, Hkv, d, generator=g).to(torch.bfloat16)
        v = torch.randn(t, Hkv, d, generator=g).to(torch.bfloat16)
        R.reshape_and_cache_ref(k, v, kc, vc, slots)
        ks.append(k)
        vs.append(v)
    q = torch.randn(sum(lens), Hq, d, generator=g).to(torch.bfloat16)
    cu = torch.zeros(n + 1, dtype=torch.int32)
    cu[1:] = torch.tensor(lens).cumsum(0)
    return q, kc, vc, bt, cu, torch.tensor(starts, dtype=torch.int32), ks, vs


def test_prefill_oracle_matches_sdpa():
    lens, starts = [37, 130], [0, 45]
    q, kc, vc, bt, cu, st, ks, vs = _case(lens, starts)
    out = L.prefill_attention_ref(q, kc, vc, bt, cu, st, return_f32=True)
    assert torch.isfinite(out).all()
    G = q.shape[1] // ks[0].shape[1]
    for c, (n, s) in enumerate(zip(lens, starts)):
        Q = q[cu[c]:cu[c + 1]].float().permute(1, 0, 2)
        K = ks[c].float().permute(1, 0, 2).repeat_interleave(G, 0)
        V = vs[c].float().permute(1, 0, 2).repeat_interleave(G, 0)
        mask = torch.arange(s + n).unsqueeze(0) <= (s + torch.arange(n)).unsqueeze(1)
        ref = F.scaled_dot_product_attention(Q, K, V, attn_mask=mask).permute(1, 0, 2)
        assert torch.allclose(out[cu[c]:cu[c + 1]], ref, atol=2e-5, rtol=1e-5)


def test_prefill_oracle_last_row_is_decode_oracle():
    q, kc, vc, bt, cu, st, _, _ = _case([50, 200], [70, 0], seed=2)
    kc = torch.nan_to_num(kc)
    vc = torch.nan_to_num(vc)
    out = L.prefill_attention_ref(q, kc, vc, bt, cu, st, return_f32=True)
    last = (cu[1:] - 1).long()
    seq_lens = (st + (cu[1:] - cu[:-1])).to(torch.int32)
    dec = R.paged_decode_ref(q[last], kc, vc, bt, seq_lens, return_f32=True)
    assert torch.allclose(out[last], dec, atol=2e-5, rtol=1e-5)


def test_rms_norm_oracle_matches_transformers():
    from transformers.models.llama.modeling_llama import LlamaRMSNorm
    g = torch.Generator().manual_seed(0)
    x = torch.randn(5, 512, generator=g).to(torch.bfloat16)
    w = (1 + 0.1 * torch.randn(512, generator=g)).to(torch.bfloat16)
    m = LlamaRMSNorm(512, eps=1e-5).to(torch.bfloat16)
    with torch.no_grad():
        m.weight.copy_(w)
        hf = m(x)
    ours, _ = L.rms_norm_ref(x, w, 1e-5)
    # HF rounds the normalised value to bf16 before the weight multiply: at most 2 bf16 roundings apart
    assert ((ours.float() - hf.float()).abs() <= 2.0 ** -6 * hf.float().abs() + 1e-6).all()
    res = torch.randn(5, 512, generator=g).to(torch.bfloat16)
    ours2, new_res = L.rms_norm_ref(x, w, 1e-5, res)
    assert torch.equal(new_res, x + res)
    assert torch.equal(ours2, L.rms_norm_ref(x + res, w, 1e-5)[0])


def test_silu_oracle():
    g = torch.Generator().manual_seed(0)
    gu = (torch.randn(7, 64, generator=g) * 3).to(torch.bfloat16)
    ref = (F.silu(gu[:, :32].float()) * gu[:, 32:].float()).to(torch.bfloat16)
    got = L.silu_and_mul_ref(gu)
    assert ((got.float() - ref.float()).abs() <= 2.0 ** -7 * ref.float().abs() + 1e-7).all()


def test_rope_oracle_matches_transformers_llama3():
    from transformers import LlamaConfig
    from transformers.models.llama.modeling_llama import LlamaRotaryEmbedding, apply_rotary_pos_emb
    sc = dict(factor=8.0, low_freq_factor=1.0, high_freq_factor=4.0, original_max_position_embeddings=8192)
    cfg = LlamaConfig(hidden_size=512, num_attention_heads=4, num_key_value_heads=2, head_dim=128,
                      max_position_embeddings=131072, rope_theta=500000.0,
                      rope_scaling=dict(rope_type="llama3", **sc))
    cfg.rope_parameters = dict(rope_type="llama3", rope_theta=500000.0, **sc) if hasattr(cfg, "rope_parameters") else None
    rot = LlamaRotaryEmbedding(cfg)
    T, Hq, Hkv, d, bs = 6, 4, 2, 128, 16
    g = torch.Generator().manual_seed(0)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * d, generator=g).to(torch.bfloat16)
    pos = torch.tensor([0, 1, 17, 4000, 9000, 12345], dtype=torch.int32)
    table = L.rope_table(16384, d, 500000.0, sc)
    cos, sin = rot(qkv.float(), pos.long().unsqueeze(0))           # [1, T, d] fp32
    assert torch.allclose(table[pos.long(), :64], cos[0, :, :64].float(), atol=3e-3)   # HF forms pos*inv_freq in fp32; ours in fp64
    assert torch.allclose(table[pos.long(), 64:], sin[0, :, :64].float(), atol=3e-3)
    kc = torch.zeros(4, Hkv, bs, d, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    slots = torch.tensor([3, 20, -1, 40, 41, 63], dtype=torch.int64)
    out = L.rope_and_cache_ref(qkv, pos, table, kc, vc, slots, Hq).view(T, Hq + 2 * Hkv, d)
    x = qkv.view(T, Hq + 2 * Hkv, d).float()
    qh = x[:, :Hq].permute(1, 0, 2).unsqueeze(0)
    kh = x[:, Hq:Hq + Hkv].permute(1, 0, 2).unsqueeze(0)
    qr, kr = apply_rotary_pos_emb(qh, kh, cos.float(), sin.float())
    assert torch.allclose(out[:, :Hq].float(), qr[0].permute(1, 0, 2), atol=2e-2, rtol=8e-3)
    assert torch.allclose(out[:, Hq:Hq + Hkv].float(), kr[0].permute(1, 0, 2), atol=2e-2, rtol=8e-3)
    assert torch.equal(out[:, Hq + Hkv:], qkv.view(T, -1, d)[:, Hq + Hkv:])          # v untouched
    assert torch.equal(kc[1, :, 4], out[1, Hq:Hq + Hkv]) and torch.equal(vc[2, :, 8], out[3, Hq + Hkv:])
    assert not kc[2, :, 1].any()                                                      # slot -1 skipped... (slot 33 unused)


def test_sample_oracle():
    logits = torch.tensor([[0.0, 3.0, 3.0, 1.0], [5.0, 1.0, 0.0, 0.0]]).to(torch.bfloat16)
    assert L.sample_ref(logits).tolist() == [1, 0]
    N = 4000
    rows = torch.tensor([1.0, 0.0, -1.0, 2.0]).to(torch.bfloat16).repeat(N, 1)
    got = L.sample_ref(rows, torch.ones(N), torch.arange(N, dtype=torch.int64) * 977 + 5)
    freq = torch.bincount(got, minlength=4).float() / N
    assert (freq - torch.softmax(rows[0].float(), 0)).abs().max() < 0.03


def test_llama_forward_oracle_matches_transformers():
    """The whole-model oracle against transformers' LlamaForCausalLM (fp32 weights = the bf16 values)."""
    from transformers import LlamaConfig, LlamaForCausalLM
    hidden, layers, Hq, Hkv, d, inter, vocab = 256, 2, 2, 1, 128, 512, 300
    cfg = LlamaConfig(hidden_size=hidden, num_hidden_layers=layers, num_attention_heads=Hq, num_key_value_heads=Hkv,
                      head_dim=d, intermediate_size=inter, vocab_size=vocab, rms_norm_eps=1e-5, rope_theta=500000.0,
                      max_position_embeddings=512, tie_word_embeddings=False, attention_bias=False)
    torch.manual_seed(0)
    m = LlamaForCausalLM(cfg).eval()
    sd = {k: v.to(torch.bfloat16) for k, v in m.state_dict().items()}
    m.load_state_dict({k: v.float() for k, v in sd.items()})
    w = {"embed": sd["model.embed_tokens.weight"], "norm": sd["model.norm.weight"], "lm_head": sd["lm_head.weight"]}
    for l in range(layers):
        p = f"model.layers.{l}."
        w[f"l{l}.ln1"] = sd[p + "input_layernorm.weight"]
        w[f"l{l}.ln2"] = sd[p + "post_attention_layernorm.weight"]
        w[f"l{l}.wqkv"] = torch.cat([sd[p + f"self_attn.{x}_proj.weight"] for x in "qkv"])
        w[f"l{l}.wo"] = sd[p + "self_attn.o_proj.weight"]
        w[f"l{l}.wgu"] = torch.cat([sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"]])
        w[f"l{l}.wdown"] = sd[p + "mlp.down_proj.weight"]
    ids = torch.randint(0, vocab, (40,))
    with torch.no_grad():
        hf = m(ids.unsqueeze(0)).logits[0]
    ours = L.llama_forward_ref(w, ids, layers=layers, Hq=Hq, Hkv=Hkv, d=d, eps=1e-5,
                               cos_sin=L.rope_table(512, d, 500000.0))
    # ours rounds activations to bf16 between ops (as the engine does); HF here runs pure fp32
    assert (ours - hf).abs().max() < 3e-2 * hf.abs().max()
    assert (ours.argmax(-1) == hf.argmax(-1)).float().mean() > 0.9


def test_sample_candidates_oracle():
    """top-k then top-p (b2_sample_filtered's contract) on a hand-checkable row."""
    logits = torch.tensor([[3.0, 1.0, 2.0, 2.0, 0.0, -1.0]]).to(torch.bfloat16)
    T = torch.ones(1)
    k = lambda n: torch.tensor([n], dtype=torch.int32)                       # noqa: E731
    p = lambda x: torch.tensor([x], dtype=torch.float32)                     # noqa: E731
    assert L.sample_candidates_ref(logits, T, k(1), None)[0].tolist() == [True, False, False, False, False, False]
    assert L.sample_candidates_ref(logits, T, k(2), None)[0].tolist() == [True, False, True, True, False, False]   # tie kept
    assert L.sample_candidates_ref(logits, T, k(0), None)[0].all() and L.sample_candidates_ref(logits, T, k(6), None)[0].all()
    w = torch.exp(logits[0].double() - 3.0)
    frac0 = float(w[0] / w.sum())                                            # 0.44: mass of the top token alone
    assert L.sample_candidates_ref(logits, T, None, p(frac0 - 0.01))[0].tolist() == [True] + [False] * 5
    assert L.sample_candidates_ref(logits, T, None, p(frac0 + 0.01))[0].tolist() == [True, False, True, True, False, False]
    assert L.sample_candidates_ref(logits, T, None, p(1.0))[0].all()
    # top-p is taken over the top-k mass: k = 3 (-> {3, 2, 2}) then p just above the top token's share of THAT mass
    share = float(w[0] / (w[0] + w[2] + w[3]))
    assert L.sample_candidates_ref(logits, T, k(3), p(share - 0.01))[0].tolist() == [True] + [False] * 5
    assert L.sample_candidates_ref(logits, T, k(3), p(share + 0.01))[0].tolist() == [True, False, True, True, False, False]
    # greedy rows ignore the filters; a filtered draw is the best-scoring candidate
    assert L.sample_candidates_ref(logits, torch.zeros(1), k(1), p(0.1))[0].all()
    rows = logits.repeat(200, 1)
    got = L.sample_ref(rows, torch.ones(200), torch.arange(200, dtype=torch.int64), top_k=torch.full((200,), 2, dtype=torch.int32))
    assert set(got.tolist()) == {0, 2, 3}
