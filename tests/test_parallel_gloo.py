"""KV-head sharding on CPU: world_size-2 and -4 gloo groups.

Each rank owns Hkv/tp kv heads and the q heads that read them; block tables / seq_lens / slots are replicated.
The ranks compute their heads with the CPU oracle (stand-in for the CUDA kernel, which needs a GPU), then
`gather_heads` must reproduce the unsharded result exactly, and the sharded KV write must equal the matching
slice of the unsharded cache."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    for p in (ROOT / "vllm-backend-public_b200", ROOT / "oracle"):
        sys.path.insert(0, str(p))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import paged_attention_ref as R
        from b200attn import parallel
        B, Hq, Hkv, d, bs, ctx = 5, 16, 4, 128, 16, 200
        inp = R.make_inputs(B, Hq, Hkv, d, bs, ctx, seed=7, ragged=True)          # same seed: replicated inputs
        sh = parallel.from_env(Hq, Hkv)
        assert (sh.tp, sh.rank) == (world, rank)
        # unsharded reference
        kc = torch.zeros(inp["num_blocks"], Hkv, bs, d, dtype=torch.bfloat16)
        vc = torch.zeros_like(kc)
        R.reshape_and_cache_ref(inp["k"], inp["v"], kc, vc, inp["slot_mapping"])
        full = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"])
        # this rank: its kv heads only, same slot_mapping / block tables
        kc_l = torch.zeros(sh.cache_shape(inp["num_blocks"], bs, d), dtype=torch.bfloat16)
        vc_l = torch.zeros_like(kc_l)
        R.reshape_and_cache_ref(sh.kv(inp["k"]), sh.kv(inp["v"]), kc_l, vc_l, inp["slot_mapping"])
        assert torch.equal(kc_l, kc[:, sh.kv_heads]) and torch.equal(vc_l, vc[:, sh.kv_heads])
        out_l = R.paged_decode_ref(sh.q(inp["q"]), kc_l, vc_l, inp["block_tables"], inp["seq_lens"])
        assert out_l.shape == (B, sh.Hq_local, d)
        got = parallel.gather_heads(out_l, sh)
        assert got.shape == full.shape
        assert torch.equal(got.view(torch.int16), full.view(torch.int16)), "gathered heads differ from the unsharded result"
        ret[rank] = "ok"
    except Exception as e:  # noqa: BLE001
        ret[rank] = f"{type(e).__name__}: {e}"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_head_sharding_gather_matches_unsharded(world):
    port = _free_port()
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {r: "ok" for r in range(world)}, dict(ret)


def test_shard_arithmetic():
    sys.path.insert(0, str(ROOT / "vllm-backend-public_b200"))
    from b200attn.parallel import HeadShard
    s = HeadShard(8, 3, 64, 8)                       # BASELINE configs[3] at tp = 8
    assert (s.Hq_local, s.Hkv_local) == (8, 1) and s.q_heads == slice(24, 32) and s.kv_heads == slice(3, 4)
    assert s.cache_bytes_per_block(16, 128, 2) == 2 * 16 * 128 * 2
    for bad in [(3, 0, 64, 8), (2, 2, 64, 8), (2, 0, 30, 8)]:
        with pytest.raises(ValueError):
            HeadShard(*bad)


# ---------------------------------------------------------------------------------------------------------------
# tensor-parallel SERVING algebra (b200attn.tp / model.py) on CPU: every rank holds shard_weight(...) of a decoder
# layer, computes its heads and its MLP slice in plain fp32 torch, the row-parallel partial sums are all-reduced over
# gloo (the role csrc/comm.cu plays on GPUs), the vocabulary-parallel logits are all-gathered -- and the result must
# equal the unsharded layer.
# ---------------------------------------------------------------------------------------------------------------
def _layer(x, w, Hq, Hkv, d, reduce=None, gather=None):
    """One pre-norm decoder layer + lm_head in fp32 over T tokens of ONE sequence (dense causal attention, no rotary:
    the sharding algebra does not depend on it).  `reduce` sums row-parallel partials, `gather` concatenates logits."""
    def norm(h, g):
        return g * h * torch.rsqrt((h * h).mean(-1, keepdim=True) + 1e-5)
    T = x.shape[0]
    G = Hq // Hkv
    resid = x
    h = norm(resid, w["ln1"])
    qkv = (h @ w["wqkv"].T).view(T, Hq + 2 * Hkv, d)
    q, k, v = qkv[:, :Hq], qkv[:, Hq:Hq + Hkv], qkv[:, Hq + Hkv:]
    mask = torch.triu(torch.ones(T, T, dtype=torch.bool), 1)
    heads = []
    for hq in range(Hq):
        s = (q[:, hq] @ k[:, hq // G].T) * d ** -0.5
        heads.append(torch.softmax(s.masked_fill(mask, float("-inf")), -1) @ v[:, hq // G])
    o = torch.stack(heads, 1).reshape(T, Hq * d) @ w["wo"].T
    if reduce:
        o = reduce(o)
    resid = resid + o
    h = norm(resid, w["ln2"])
    gu = h @ w["wgu"].T
    I = gu.shape[1] // 2
    m = (torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]) @ w["wdown"].T
    if reduce:
        m = reduce(m)
    resid = resid + m
    logits = norm(resid, w["norm"]) @ w["lm_head"].T
    return gather(logits) if gather else logits


def _tp_worker(rank, world, port, ret):
    sys.path.insert(0, str(ROOT / "vllm-backend-public_b200"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from b200attn.parallel import shard_weight
        Hq, Hkv, d, hidden, inter, vocab, T = 8, 4, 16, 64, 96, 40, 9
        g = torch.Generator().manual_seed(3)                                 # same seed: every rank draws the full model
        full = {"ln1": torch.rand(hidden, generator=g) + 0.5, "ln2": torch.rand(hidden, generator=g) + 0.5,
                "norm": torch.rand(hidden, generator=g) + 0.5,
                "wqkv": torch.randn((Hq + 2 * Hkv) * d, hidden, generator=g) * 0.1,
                "wo": torch.randn(hidden, Hq * d, generator=g) * 0.1,
                "wgu": torch.randn(2 * inter, hidden, generator=g) * 0.1,
                "wdown": torch.randn(hidden, inter, generator=g) * 0.1,
                "lm_head": torch.randn(vocab, hidden, generator=g) * 0.1}
        x = torch.randn(T, hidden, generator=g)
        want = _layer(x, full, Hq, Hkv, d)
        mine = {k: shard_weight(k, v, world, rank, Hq, Hkv, d, inter) for k, v in full.items()}
        assert mine["wqkv"].shape[0] == (Hq + 2 * Hkv) * d // world and mine["lm_head"].shape[0] == vocab // world

        def reduce(t):
            t = t.clone()
            dist.all_reduce(t)
            return t

        def gather(t):
            parts = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(parts, t.contiguous())
            return torch.cat(parts, dim=1)

        got = _layer(x, mine, Hq // world, Hkv // world, d, reduce, gather)
        err = float((got - want).abs().max())
        assert got.shape == want.shape and err < 1e-4, err
        ret[rank] = "ok"
    except Exception as e:  # noqa: BLE001
        ret[rank] = f"{type(e).__name__}: {e}"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_tensor_parallel_layer_algebra_matches_unsharded(world):
    port = _free_port()
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_tp_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {r: "ok" for r in range(world)}, dict(ret)
