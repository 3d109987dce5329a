"""GPU parity tests proper: the sm_100a kernels, called through the C ABI (ctypes ->
libb200attn.so), against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star):
  * reshape_and_cache: bit-exact.
  * decode, bf16 KV: |delta| <= 2e-3 + 1e-2 * |ref|.
  * decode, fp8 KV vs the oracle reading the *same quantised cache*: same bound; vs the
    unquantised oracle the looser stated bound 5e-2 + 1e-1 * |ref| (quantisation error).
At BASELINE.json's full sizes the oracle is run on a subset of sequences and the rest is covered
by size-independent properties (split-count invariance, block-table permutation invariance,
agreement of two independent kernels)."""
import json
from pathlib import Path

import pytest
import torch

import b200attn
import paged_attention_ref as R

pytestmark = pytest.mark.gpu

ATOL, RTOL = 2e-3, 1e-2
GOLDEN = Path(__file__).parent / "golden"


def assert_close(got, ref, atol=ATOL, rtol=RTOL, what=""):
    got = got.float().cpu()
    ref = ref.float().cpu()
    assert torch.isfinite(got).all(), f"{what}: non-finite output"
    err = (got - ref).abs()
    bound = atol + rtol * ref.abs()
    bad = err > bound
    assert not bad.any(), (f"{what}: {int(bad.sum())}/{bad.numel()} outside {atol}+{rtol}|ref|; "
                           f"max err {float(err.max()):.3e} at ref {float(ref.flatten()[err.argmax()]):.3e}")


def build_caches(inp, dev, fp8=False, poison=False):
    """Fill device caches through our reshape_and_cache and the CPU caches through the oracle."""
    shape = (inp["num_blocks"], inp["Hkv"], inp["bs"], inp["d"])
    if fp8:
        ks = float(inp["k"].float().abs().max()) / 448.0
        vs = float(inp["v"].float().abs().max()) / 448.0
        fill8 = 0x7F if poison else 0     # 0x7f is the e4m3 NaN
        kc_ref = torch.full(shape, fill8, dtype=torch.uint8)
        vc_ref = torch.full(shape, fill8, dtype=torch.uint8)
        R.reshape_and_cache_ref(inp["k"], inp["v"], kc_ref, vc_ref, inp["slot_mapping"], R.KV_FP8_E4M3, ks, vs)
        kc = torch.full(shape, fill8, dtype=torch.uint8, device=dev)
        vc = torch.full(shape, fill8, dtype=torch.uint8, device=dev)
        ksd = torch.tensor([ks], dtype=torch.float32, device=dev)
        vsd = torch.tensor([vs], dtype=torch.float32, device=dev)
        b200attn.reshape_and_cache(inp["k"].to(dev), inp["v"].to(dev), kc, vc, inp["slot_mapping"].to(dev), ksd, vsd)
        return kc, vc, kc_ref, vc_ref, (ks, vs, ksd, vsd)
    fill = float("nan") if poison else 0.0
    kc_ref = torch.full(shape, fill, dtype=torch.bfloat16)
    vc_ref = torch.full(shape, fill, dtype=torch.bfloat16)
    R.reshape_and_cache_ref(inp["k"], inp["v"], kc_ref, vc_ref, inp["slot_mapping"])
    kc = torch.full(shape, fill, dtype=torch.bfloat16, device=dev)
    vc = torch.full(shape, fill, dtype=torch.bfloat16, device=dev)
    b200attn.reshape_and_cache(inp["k"].to(dev), inp["v"].to(dev), kc, vc, inp["slot_mapping"].to(dev))
    return kc, vc, kc_ref, vc_ref, None


def bits(t):
    return t.contiguous().view(torch.uint8).cpu()


# ------------------------------------------------------------------------------------------
# a5 reshape_and_cache: bit-exact
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("bs", [8, 16, 32])
def test_reshape_and_cache_bf16_bit_exact(cuda, bs):
    inp = R.make_inputs(5, 8, 2, 128, bs, 70, seed=bs, ragged=True, slack_blocks=5)
    slots = inp["slot_mapping"].clone()
    slots[::7] = -1
    inp["slot_mapping"] = slots
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    assert torch.equal(bits(kc), bits(kc_ref))
    assert torch.equal(bits(vc), bits(vc_ref))


def test_reshape_and_cache_token_stride_and_empty(cuda):
    """k/v taken as slices of a fused QKV projection (token stride > Hkv*d); T = 0 is a no-op."""
    Hq, Hkv, d, bs = 8, 2, 128, 16
    inp = R.make_inputs(3, Hq, Hkv, d, bs, 40, seed=11, ragged=True)
    T = inp["k"].shape[0]
    g = torch.Generator().manual_seed(1)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * d, generator=g).to(torch.bfloat16)
    qkv[:, Hq * d:(Hq + Hkv) * d] = inp["k"].reshape(T, -1)
    qkv[:, (Hq + Hkv) * d:] = inp["v"].reshape(T, -1)
    qkv_d = qkv.to(cuda)
    k_view = qkv_d[:, Hq * d:(Hq + Hkv) * d].unflatten(1, (Hkv, d))
    v_view = qkv_d[:, (Hq + Hkv) * d:].unflatten(1, (Hkv, d))
    shape = (inp["num_blocks"], Hkv, bs, d)
    kc = torch.zeros(shape, dtype=torch.bfloat16, device=cuda)
    vc = torch.zeros_like(kc)
    b200attn.reshape_and_cache(k_view, v_view, kc, vc, inp["slot_mapping"].to(cuda))
    kc_ref = torch.zeros(shape, dtype=torch.bfloat16)
    vc_ref = torch.zeros_like(kc_ref)
    R.reshape_and_cache_ref(inp["k"], inp["v"], kc_ref, vc_ref, inp["slot_mapping"])
    assert torch.equal(bits(kc), bits(kc_ref)) and torch.equal(bits(vc), bits(vc_ref))
    empty = torch.zeros(0, Hkv, d, dtype=torch.bfloat16, device=cuda)
    b200attn.reshape_and_cache(empty, empty, kc, vc, torch.zeros(0, dtype=torch.int64, device=cuda))
    assert torch.equal(bits(kc), bits(kc_ref))


def test_reshape_and_cache_fp8_bit_exact(cuda):
    inp = R.make_inputs(4, 8, 2, 128, 16, 90, seed=3, ragged=True)
    # stretch the range so saturation, subnormals and zero all occur
    inp["k"][::3] *= 300.0
    inp["k"][1::3] *= 1e-3
    inp["v"][::5] = 0.0
    kc_shape = (inp["num_blocks"], 2, 16, 128)
    ks, vs = 0.37, 0.0121
    kc_ref = torch.zeros(kc_shape, dtype=torch.uint8)
    vc_ref = torch.zeros_like(kc_ref)
    R.reshape_and_cache_ref(inp["k"], inp["v"], kc_ref, vc_ref, inp["slot_mapping"], R.KV_FP8_E4M3, ks, vs)
    kc = torch.zeros(kc_shape, dtype=torch.uint8, device=cuda)
    vc = torch.zeros_like(kc)
    b200attn.reshape_and_cache(inp["k"].to(cuda), inp["v"].to(cuda), kc, vc, inp["slot_mapping"].to(cuda),
                               torch.tensor([ks], device=cuda), torch.tensor([vs], device=cuda))
    assert torch.equal(kc.cpu(), kc_ref), f"{int((kc.cpu() != kc_ref).sum())} K bytes differ"
    assert torch.equal(vc.cpu(), vc_ref), f"{int((vc.cpu() != vc_ref).sum())} V bytes differ"
    assert int((kc_ref & 0x7F).max()) == 0x7E           # saturation was exercised


# ------------------------------------------------------------------------------------------
# a6/a7 decode, bf16 KV
# ------------------------------------------------------------------------------------------
def run_decode(inp, kc, vc, dev, **kw):
    return b200attn.paged_decode(inp["q"].to(dev), kc, vc, inp["block_tables"].to(dev), inp["seq_lens"].to(dev), **kw)


@pytest.mark.parametrize("impl,splits", [("tma", 1), ("tma", 2), ("tma", 0), ("tma", 7), ("simt", 1)])
def test_decode_config1(cuda, impl, splits):
    """BASELINE.json configs[0]: 8 q / 2 kv heads, d 128, bs 16, batch 4, ctx 512."""
    inp = R.make_inputs(4, 8, 2, 128, 16, 512, seed=0)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], return_f32=True)
    out = run_decode(inp, kc, vc, cuda, impl=impl, num_splits=splits)
    assert_close(out, ref, what=f"config1 {impl} splits={splits}")


def test_decode_config1_against_committed_golden(cuda):
    meta = json.loads((GOLDEN / "config1.json").read_text())
    inp = R.make_inputs(**meta["shape"], seed=meta["seed"])
    kc, vc, _, _, _ = build_caches(inp, cuda)
    out = run_decode(inp, kc, vc, cuda)
    assert_close(out, torch.load(GOLDEN / "config1_out.pt"), what="golden config1")


@pytest.mark.parametrize("Hq,Hkv", [(1, 1), (4, 4), (8, 2), (32, 8), (8, 1), (16, 1), (24, 2), (40, 2)])
def test_decode_gqa_group_sizes(cuda, Hq, Hkv):
    inp = R.make_inputs(3, Hq, Hkv, 128, 16, 200, seed=Hq, ragged=True)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], return_f32=True)
    assert_close(run_decode(inp, kc, vc, cuda), ref, what=f"G={Hq // Hkv}")
    assert_close(run_decode(inp, kc, vc, cuda, impl="simt"), ref, what=f"simt G={Hq // Hkv}")


@pytest.mark.parametrize("bs", [8, 16, 32])
@pytest.mark.parametrize("splits", [1, 3])
def test_decode_block_sizes_and_edge_lengths(cuda, bs, splits):
    """seq_len 0, 1, around block and stage (64-token) boundaries; untouched cache memory is NaN, so any
    read past seq_len that reaches the arithmetic shows up as a non-finite output."""
    lens = [0, 1, bs - 1, bs, bs + 1, 63, 64, 65, 127, 128, 129, 300]
    B = len(lens)
    inp = R.make_inputs(B, 8, 2, 128, bs, 320, seed=bs + splits)
    sl = torch.tensor(lens, dtype=torch.int32)
    # rebuild the token stream for the shortened sequences
    keep = []
    off = 0
    for b in range(B):
        keep.append(torch.arange(off, off + lens[b]))
        off += 320
    keep = torch.cat(keep)
    inp["k"], inp["v"], inp["slot_mapping"] = inp["k"][keep], inp["v"][keep], inp["slot_mapping"][keep]
    inp["seq_lens"] = sl
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda, poison=True)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], sl, return_f32=True)
    out = run_decode(inp, kc, vc, cuda, num_splits=splits)
    assert torch.all(out[0] == 0), "seq_len 0 must give a zero row"
    assert_close(out, ref, what=f"bs={bs} splits={splits}")


def test_decode_ragged_mixed_lengths(cuda):
    inp = R.make_inputs(16, 32, 8, 128, 16, 1500, seed=5, ragged=True)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], return_f32=True)
    for splits in (0, 1, 4):
        assert_close(run_decode(inp, kc, vc, cuda, num_splits=splits), ref, what=f"ragged splits={splits}")


def test_decode_q_and_out_strides(cuda):
    """q as a slice of a fused QKV row and out as a slice of a wider buffer (strided batch axis)."""
    inp = R.make_inputs(4, 8, 2, 128, 16, 100, seed=6)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], return_f32=True)
    wide = torch.zeros(4, 12 * 128, dtype=torch.bfloat16, device=cuda)
    wide[:, :8 * 128] = inp["q"].reshape(4, -1).to(cuda)
    qv = wide[:, :8 * 128].unflatten(1, (8, 128))
    obuf = torch.zeros(4, 10 * 128, dtype=torch.bfloat16, device=cuda)
    ov = obuf[:, :8 * 128].unflatten(1, (8, 128))
    b200attn.paged_decode(qv, kc, vc, inp["block_tables"].to(cuda), inp["seq_lens"].to(cuda), out=ov, num_splits=2)
    assert_close(ov, ref, what="strided q/out")
    assert torch.all(obuf[:, 8 * 128:] == 0)


def test_decode_large_magnitude_scores(cuda):
    """Scores far from zero: the running-max rescale must keep everything finite."""
    inp = R.make_inputs(2, 8, 2, 128, 16, 260, seed=8)
    inp["q"] = (inp["q"].float() * 6).to(torch.bfloat16)
    inp["k"] = (inp["k"].float() * 6).to(torch.bfloat16)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], return_f32=True)
    assert_close(run_decode(inp, kc, vc, cuda, num_splits=1), ref, atol=4e-3, rtol=2e-2, what="peaked softmax")


# ------------------------------------------------------------------------------------------
# a8 decode, fp8 KV
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("Hq,Hkv,splits", [(8, 2, 1), (32, 8, 0), (16, 1, 2)])
def test_decode_fp8(cuda, Hq, Hkv, splits):
    inp = R.make_inputs(4, Hq, Hkv, 128, 16, 400, seed=21, ragged=True)
    kc, vc, kc_ref, vc_ref, (ks, vs, ksd, vsd) = build_caches(inp, cuda, fp8=True)
    assert torch.equal(kc.cpu(), kc_ref) and torch.equal(vc.cpu(), vc_ref)
    ref_q = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"],
                               kv_dtype=R.KV_FP8_E4M3, k_scale=ks, v_scale=vs, return_f32=True)
    for impl in ("tma", "simt"):
        out = run_decode(inp, kc, vc, cuda, k_scale=ksd, v_scale=vsd, num_splits=splits, impl=impl)
        assert_close(out, ref_q, what=f"fp8 {impl} vs same-cache oracle")
    # against the unquantised oracle: the stated looser bound
    shape = kc.shape
    kb = torch.zeros(shape, dtype=torch.bfloat16)
    vb = torch.zeros_like(kb)
    R.reshape_and_cache_ref(inp["k"], inp["v"], kb, vb, inp["slot_mapping"])
    ref = R.paged_decode_ref(inp["q"], kb, vb, inp["block_tables"], inp["seq_lens"], return_f32=True)
    out = run_decode(inp, kc, vc, cuda, k_scale=ksd, v_scale=vsd)
    assert_close(out, ref, atol=5e-2, rtol=1e-1, what="fp8 vs unquantised oracle")


@pytest.mark.parametrize("bs", [8, 16, 32])
@pytest.mark.parametrize("splits", [0, 1, 3])
def test_decode_fp8_block_sizes_and_edge_lengths(cuda, bs, splits):
    """fp8 cache: seq_len 0, 1, around block, warp-slab (32-token) and stage (128-token) boundaries; untouched
    cache bytes are the e4m3 NaN (0x7f), so a read past seq_len that reaches an MMA gives a non-finite row."""
    lens = [0, 1, bs - 1, bs, bs + 1, 31, 32, 33, 63, 64, 65, 127, 128, 129, 255, 257, 300]
    B = len(lens)
    inp = R.make_inputs(B, 8, 2, 128, bs, 320, seed=40 + bs + splits)
    sl = torch.tensor(lens, dtype=torch.int32)
    keep = torch.cat([torch.arange(320 * b, 320 * b + lens[b]) for b in range(B)])
    inp["k"], inp["v"], inp["slot_mapping"] = inp["k"][keep], inp["v"][keep], inp["slot_mapping"][keep]
    inp["seq_lens"] = sl
    kc, vc, kc_ref, vc_ref, (ks, vs, ksd, vsd) = build_caches(inp, cuda, fp8=True, poison=True)
    assert torch.equal(kc.cpu(), kc_ref) and torch.equal(vc.cpu(), vc_ref)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], sl,
                             kv_dtype=R.KV_FP8_E4M3, k_scale=ks, v_scale=vs, return_f32=True)
    out = run_decode(inp, kc, vc, cuda, k_scale=ksd, v_scale=vsd, num_splits=splits)
    assert torch.all(out[0] == 0), "seq_len 0 must give a zero row"
    assert_close(out, ref, what=f"fp8 bs={bs} splits={splits}")


@pytest.mark.parametrize("qmul", [2.0 ** -12, 1.0, 37.0, 2.0 ** 9])
def test_decode_fp8_q_range(cuda, qmul):
    """The fp8 path feeds q to the e4m3 MMA as two terms after a per-head power-of-two scaling: q rows of very
    different magnitude (and one all-zero head) must come out as with exact bf16 q.  The softmax scale is
    divided by the same factor so that the scores stay O(1)."""
    inp = R.make_inputs(3, 16, 2, 128, 16, 500, seed=77, ragged=True)
    qf = inp["q"].float() * qmul
    qf[:, 3] *= 1e-3           # heads of one group with unlike ranges
    qf[:, 5] = 0.0
    qf[:, 9] *= 50.0
    inp["q"] = qf.to(torch.bfloat16)
    scale = 128 ** -0.5 / qmul
    kc, vc, kc_ref, vc_ref, (ks, vs, ksd, vsd) = build_caches(inp, cuda, fp8=True)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], scale=scale,
                             kv_dtype=R.KV_FP8_E4M3, k_scale=ks, v_scale=vs, return_f32=True)
    for splits in (0, 1, 2):
        out = run_decode(inp, kc, vc, cuda, k_scale=ksd, v_scale=vsd, num_splits=splits, scale=scale)
        assert_close(out, ref, atol=4e-3, rtol=2e-2, what=f"fp8 q x{qmul} splits={splits}")


# ------------------------------------------------------------------------------------------
# full-size configs: oracle on a subset + size-independent properties
# ------------------------------------------------------------------------------------------
def _full_size(cuda, B, Hq, Hkv, ctx, fp8, seed, n_oracle=8):
    inp = R.make_inputs(B, Hq, Hkv, 128, 16, ctx, seed=seed)
    dev_k, dev_v = inp["k"].to(cuda), inp["v"].to(cuda)
    shape = (inp["num_blocks"], Hkv, 16, 128)
    kw = {}
    if fp8:
        ks = float(inp["k"].float().abs().max()) / 448.0
        vs = float(inp["v"].float().abs().max()) / 448.0
        kw = dict(k_scale=torch.tensor([ks], device=cuda), v_scale=torch.tensor([vs], device=cuda))
        kc = torch.zeros(shape, dtype=torch.uint8, device=cuda)
    else:
        kc = torch.zeros(shape, dtype=torch.bfloat16, device=cuda)
    vc = torch.zeros_like(kc)
    b200attn.reshape_and_cache(dev_k, dev_v, kc, vc, inp["slot_mapping"].to(cuda), kw.get("k_scale"), kw.get("v_scale"))
    del dev_k, dev_v
    q, bt, sl = inp["q"].to(cuda), inp["block_tables"].to(cuda), inp["seq_lens"].to(cuda)
    out_auto = b200attn.paged_decode(q, kc, vc, bt, sl, **kw)
    out_one = b200attn.paged_decode(q, kc, vc, bt, sl, num_splits=1, **kw)
    out_many = b200attn.paged_decode(q, kc, vc, bt, sl, num_splits=16, **kw)
    # (1) split-count invariance: the merge is exact algebra, only bf16/fp32 rounding differs
    assert_close(out_one, out_auto.float(), what="splits 1 vs auto")
    assert_close(out_many, out_auto.float(), what="splits 16 vs auto")
    # (2) two independent kernels agree on every row
    out_simt = b200attn.paged_decode(q, kc, vc, bt, sl, impl="simt", **kw)
    assert_close(out_auto, out_simt.float(), what="tma vs simt")
    # (3) the oracle on n_oracle sequences drawn across the batch: the first and last rows, the row that owns the
    #     highest block id of the pool, and random ones.  The oracle builds ITS OWN cache from the same tokens
    #     (reshape_and_cache_ref), and the device cache is compared with it bit for bit on every block those rows own:
    #     the KV write is checked at full size against the oracle, not read back from the device.
    g = torch.Generator().manual_seed(seed + 100)
    top_row = int((inp["block_tables"].max(dim=1).values).argmax())
    rows = sorted({0, B - 1, top_row} | set(torch.randperm(B, generator=g)[:max(n_oracle - 3, 0)].tolist()))
    tok0 = torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(inp["seq_lens"].long(), 0)])
    tok = torch.cat([torch.arange(int(tok0[b]), int(tok0[b + 1])) for b in rows])
    kc_ref = torch.zeros(shape, dtype=kc.dtype)
    vc_ref = torch.zeros_like(kc_ref)
    okw = dict(kv_dtype=R.KV_FP8_E4M3, k_scale=ks, v_scale=vs) if fp8 else {}
    R.reshape_and_cache_ref(inp["k"][tok], inp["v"][tok], kc_ref, vc_ref, inp["slot_mapping"][tok],
                            *((R.KV_FP8_E4M3, ks, vs) if fp8 else ()))
    need = torch.unique(inp["block_tables"][rows].flatten()).long()
    assert torch.equal(bits(kc[need.to(cuda)]), bits(kc_ref[need])), "full-size KV write differs from the oracle (K)"
    assert torch.equal(bits(vc[need.to(cuda)]), bits(vc_ref[need])), "full-size KV write differs from the oracle (V)"
    ref = R.paged_decode_ref(inp["q"][rows], kc_ref, vc_ref, inp["block_tables"][rows], inp["seq_lens"][rows],
                             return_f32=True, **okw)
    assert_close(out_auto[torch.tensor(rows, device=cuda)], ref, what=f"oracle rows {rows}")
    # (4) permuting which sequence sits in which batch row permutes the output rows
    perm = torch.randperm(B, generator=torch.Generator().manual_seed(1)).to(cuda)
    out_p = b200attn.paged_decode(q[perm].contiguous(), kc, vc, bt[perm].contiguous(), sl[perm].contiguous(), **kw)
    # (the persistent schedule cuts segments by batch position, so rows agree to rounding, not bitwise)
    assert_close(out_p, out_auto[perm].float(), what="batch-row permutation")


def test_full_size_config2_llama3_8b_b64_ctx4k(cuda):
    """BASELINE.json configs[1]: 32 q / 8 kv heads, batch 64, ctx 4096, bf16 KV."""
    _full_size(cuda, 64, 32, 8, 4096, fp8=False, seed=0)


def test_full_size_config3_fp8_b128_ctx8k(cuda):
    """BASELINE.json configs[2]: fp8 e4m3 KV, batch 128, ctx 8192."""
    _full_size(cuda, 128, 32, 8, 8192, fp8=True, seed=1)


def test_full_size_config4_per_gpu_shard_tp8(cuda):
    """BASELINE.json configs[3] as seen by one of 8 ranks: 8 q / 1 kv head, batch 256, ctx 8192."""
    _full_size(cuda, 256, 8, 1, 8192, fp8=False, seed=2)


def test_full_size_config4_tp1_llama3_70b_b256_ctx8k(cuda):
    """BASELINE.json configs[3] unsharded: 64 q / 8 kv heads (G = 8), batch 256, ctx 8192, bf16 KV -- 8.6 GB of cache.
    The problem is generated on the device (bench.build_problem: per-sequence seeds); the oracle regenerates the tokens
    of 8 rows (first, last, random), builds its own cache from them and is compared on the KV bits and on the output."""
    import bench
    B, Hq, Hkv, ctx, seed = 256, 64, 8, 8192, 4
    prob = bench.build_problem(cuda, B, Hq, Hkv, ctx, False, seed=seed)
    kc, vc = prob["k_cache"], prob["v_cache"]
    bt, sl = prob["block_tables"].to(cuda), prob["seq_lens"].to(cuda)
    out_auto = b200attn.paged_decode(prob["q"], kc, vc, bt, sl)
    out_one = b200attn.paged_decode(prob["q"], kc, vc, bt, sl, num_splits=1)
    out_dyn = b200attn.paged_decode(prob["q"], kc, vc, bt, sl, num_splits=0)
    assert_close(out_one, out_auto.float(), what="config 4 tp1: single pass vs planner's schedule")
    assert_close(out_dyn, out_auto.float(), what="config 4 tp1: dynamic vs planner's schedule")
    g = torch.Generator().manual_seed(9)
    rows = sorted({0, B - 1, int(prob["block_tables"].max(dim=1).values.argmax())} | set(torch.randperm(B, generator=g)[:5].tolist()))
    shape = tuple(kc.shape)
    kc_ref = torch.zeros(shape[0], Hkv, 16, 128, dtype=torch.bfloat16)     # 8.6 GB of zeros on the host: lazily paged
    vc_ref = torch.zeros_like(kc_ref)
    pos = torch.arange(ctx)
    for b in rows:
        k, v = bench.gen_sequence_kv(cuda, seed, b, ctx, Hkv)
        slots = prob["block_tables"][b, pos // 16].long() * 16 + pos % 16
        R.reshape_and_cache_ref(k.cpu(), v.cpu(), kc_ref, vc_ref, slots)
    need = torch.unique(prob["block_tables"][rows].flatten()).long()
    assert torch.equal(bits(kc[need.to(cuda)]), bits(kc_ref[need])), "config 4 tp1: KV write differs from the oracle (K)"
    assert torch.equal(bits(vc[need.to(cuda)]), bits(vc_ref[need])), "config 4 tp1: KV write differs from the oracle (V)"
    ref = R.paged_decode_ref(prob["q"][rows].cpu(), kc_ref, vc_ref, prob["block_tables"][rows], prob["seq_lens"][rows],
                             return_f32=True)
    assert_close(out_auto[torch.tensor(rows, device=cuda)], ref, what=f"config 4 tp1 oracle rows {rows}")


# ------------------------------------------------------------------------------------------
# schedules of the persistent kernel: every (chunk size, static share) must give the same answer
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("env", [{}, {"B2_STATIC_PCT": "0", "B2_CHUNK_TOKENS": "64"},
                                 {"B2_STATIC_PCT": "50", "B2_CHUNK_TOKENS": "128"}, {"B2_STATIC_PCT": "100"},
                                 {"B2_STATIC_PCT": "85", "B2_CHUNK_TOKENS": "512"},
                                 {"B2_STATIC_PCT": "30", "B2_CHUNK_TOKENS": "256"}])
def test_decode_schedule_variants_ragged(cuda, env, monkeypatch):
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    inp = R.make_inputs(37, 16, 4, 128, 16, 1500, seed=11, ragged=True)
    inp["seq_lens"][3] = 0
    inp["seq_lens"][5] = 1          # (lengths may only shrink: the block tables were dealt for the original ones)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda, poison=True)
    ref = R.paged_decode_ref(inp["q"], torch.nan_to_num(kc_ref), torch.nan_to_num(vc_ref), inp["block_tables"],
                             inp["seq_lens"], return_f32=True)
    out = b200attn.paged_decode(inp["q"].to(cuda), kc, vc, inp["block_tables"].to(cuda), inp["seq_lens"].to(cuda),
                                num_splits=0)
    assert_close(out, ref, what=f"schedule {env}")
    assert float(out[3].float().abs().max()) == 0.0
    # the same problem through a plan handle, repeatedly: the claim counter in the workspace header must come back to
    # zero every time (the merge kernel re-arms it)
    plan = b200attn.ops.DecodePlan(kc, vc, 37, 16, inp["block_tables"].shape[1])
    o2 = torch.empty_like(out)
    for _ in range(3):
        o2.fill_(float("nan"))
        plan.run(inp["q"].to(cuda), inp["block_tables"].to(cuda), inp["seq_lens"].to(cuda), o2, 0)
        assert torch.equal(o2.view(torch.int16), out.view(torch.int16)), "plan-handle run differs from the stateless call"
    small = slice(0, 9)                       # a smaller batch on the same handle (header laid out for 37 sequences)
    o3 = torch.full((9, 16, 128), float("nan"), dtype=torch.bfloat16, device=cuda)
    plan.run(inp["q"][small].to(cuda), inp["block_tables"][small].to(cuda), inp["seq_lens"][small].to(cuda), o3, 0)
    assert_close(o3, ref[small], what=f"schedule {env}, 9 of 37 sequences on the same handle")
    plan.run(inp["q"].to(cuda), inp["block_tables"].to(cuda), inp["seq_lens"].to(cuda), o2, 0)
    assert torch.equal(o2.view(torch.int16), out.view(torch.int16))


def test_decode_many_sequences_grouped_prefix(cuda):
    """More sequences than the kernel's shared-memory prefix holds entries (512): grouped prefix + walk."""
    inp = R.make_inputs(700, 4, 1, 128, 16, 96, seed=12, ragged=True)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], return_f32=True)
    out = b200attn.paged_decode(inp["q"].to(cuda), kc, vc, inp["block_tables"].to(cuda), inp["seq_lens"].to(cuda))
    assert_close(out, ref, what="B=700")


# ------------------------------------------------------------------------------------------
# the engine-facing step: host buffers in, host rows out (CUDA graph, output stored to pinned memory)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("use_graph,zero_copy", [(False, False), (True, False), (False, True), (True, True)])
def test_decode_step_from_host_buffers(cuda, use_graph, zero_copy):
    B, Hq, Hkv, ctx = 6, 8, 2, 300
    inp = R.make_inputs(B, Hq, Hkv, 128, 16, ctx, seed=21, ragged=True)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    step = b200attn.DecodeStep(kc, vc, max_batch=8, Hq=Hq, max_nb=inp["max_nb"], use_graph=use_graph,
                               zero_copy_inputs=zero_copy)
    g = torch.Generator().manual_seed(5)
    for it in range(3):    # three decode steps: each rewrites the last token of every sequence, then attends
        k_new = torch.randn(B, Hkv, 128, generator=g).to(torch.bfloat16)
        v_new = torch.randn(B, Hkv, 128, generator=g).to(torch.bfloat16)
        q = torch.randn(B, Hq, 128, generator=g).to(torch.bfloat16)
        last = inp["seq_lens"].long() - 1
        slots = inp["block_tables"][torch.arange(B), last // 16].long() * 16 + last % 16
        step.h_qkv[:B, :Hq] = q
        step.h_qkv[:B, Hq:Hq + Hkv] = k_new
        step.h_qkv[:B, Hq + Hkv:] = v_new
        step.h_block_tables[:B] = inp["block_tables"]
        step.upload_block_tables()
        step.h_seq_lens[:B] = inp["seq_lens"]
        step.h_slots[:B] = slots
        out = step.run(B).clone()
        R.reshape_and_cache_ref(k_new, v_new, kc_ref, vc_ref, slots)
        ref = R.paged_decode_ref(q, kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], return_f32=True)
        assert_close(out, ref, what=f"DecodeStep graph={use_graph} step {it}")
    assert torch.equal(bits(kc), bits(kc_ref)) and torch.equal(bits(vc), bits(vc_ref))


@pytest.mark.parametrize("fuse", [False, True])
@pytest.mark.parametrize("fp8", [False, True])
def test_decode_step_resident_tables_follow_the_sequences(cuda, fp8, fuse, monkeypatch):
    """Device-resident block tables: they are uploaded ONCE; then every sequence grows by one token per step across
    several block boundaries, and the only per-step host->device words are seq_lens and slots.  The decode-step KV
    write must keep the resident tables current (b2_reshape_and_cache_decode), checked against the oracle, which
    is given the host-side tables the block manager would hold.  fuse: the step as ONE call, b2_decode_plan_run_append
    (captured in the step's CUDA graph; for the single-pass grid the owning CTA writes the row and its table entry and
    must take the new block's id from the slot, never from the poisoned entry)."""
    monkeypatch.setenv("B2_STEP_FUSE_APPEND", "1" if fuse else "0")
    B, Hq, Hkv, bs, ctx, steps = 5, 8, 2, 16, 70, 40
    inp = R.make_inputs(B, Hq, Hkv, 128, bs, ctx + steps + bs, seed=31, ragged=False, slack_blocks=8)
    lens = torch.tensor([1, 15, 16, 33, 70], dtype=torch.int32)                 # start lengths (tokens already cached)
    keep = torch.cat([torch.arange((ctx + steps + bs) * b, (ctx + steps + bs) * b + int(lens[b])) for b in range(B)])
    inp["k"], inp["v"], inp["slot_mapping"] = inp["k"][keep], inp["v"][keep], inp["slot_mapping"][keep]
    inp["seq_lens"] = lens.clone()
    kc, vc, kc_ref, vc_ref, sc = build_caches(inp, cuda, fp8=fp8)
    okw, skw = {}, {}
    if fp8:
        ks, vs, ksd, vsd = sc
        okw = dict(kv_dtype=R.KV_FP8_E4M3, k_scale=ks, v_scale=vs)
        skw = dict(k_scale=ksd, v_scale=vsd)
    full_bt = inp["block_tables"]                      # the blocks each sequence WILL own, in order
    step = b200attn.DecodeStep(kc, vc, max_batch=8, Hq=Hq, max_nb=inp["max_nb"], **skw)
    # the resident tables start with only the blocks in use; the rest of each row is poison the kernel must never follow
    nb0 = (lens.long() + bs - 1) // bs
    step.h_block_tables.fill_(inp["num_blocks"] + 12345)
    for b in range(B):
        step.h_block_tables[b, :int(nb0[b])] = full_bt[b, :int(nb0[b])]
    step.upload_block_tables()
    torch.cuda.synchronize()
    g = torch.Generator().manual_seed(7)
    cur = lens.clone()
    for it in range(steps):
        cur += 1                                       # the step's new token sits at position cur - 1
        last = cur.long() - 1
        slots = full_bt[torch.arange(B), last // bs].long() * bs + last % bs
        k_new = torch.randn(B, Hkv, 128, generator=g).to(torch.bfloat16)
        v_new = torch.randn(B, Hkv, 128, generator=g).to(torch.bfloat16)
        q = torch.randn(B, Hq, 128, generator=g).to(torch.bfloat16)
        step.h_qkv[:B, :Hq], step.h_qkv[:B, Hq:Hq + Hkv], step.h_qkv[:B, Hq + Hkv:] = q, k_new, v_new
        step.h_seq_lens[:B] = cur
        step.h_slots[:B] = slots
        out = step.run(B, num_splits=(0, 1, 3)[it % 3]).clone()
        if fp8:
            R.reshape_and_cache_ref(k_new, v_new, kc_ref, vc_ref, slots, R.KV_FP8_E4M3, ks, vs)
        else:
            R.reshape_and_cache_ref(k_new, v_new, kc_ref, vc_ref, slots)
        ref = R.paged_decode_ref(q, kc_ref, vc_ref, full_bt, cur, return_f32=True, **okw)
        assert_close(out, ref, what=f"resident tables step {it} fp8={fp8} fuse={fuse}")
    assert step.fuse_append == fuse
    nb1 = (cur.long() + bs - 1) // bs
    dt = step.d_block_tables.cpu()
    for b in range(B):
        assert torch.equal(dt[b, :int(nb1[b])], full_bt[b, :int(nb1[b])]), f"row {b}: resident table != block manager's"
    assert torch.equal(bits(kc), bits(kc_ref)) and torch.equal(bits(vc), bits(vc_ref))
    assert step.h2d_bytes == step.h_qkv.numel() * 2 + step.h_meta.numel() * 4 and step.h_meta.numel() <= 3 * 8 + 1


@pytest.mark.parametrize("fp8", [False, True])
def test_decode_plan_handle_matches_stateless_call(cuda, fp8):
    """b2_decode_plan_run == b2_paged_decode, call after call on one workspace (the dynamic schedule's claim counter
    must be back at zero after every call: a stale count would skip chunks), for every schedule and for batch sizes
    below the plan's max_B."""
    inp = R.make_inputs(12, 32, 8, 128, 16, 700, seed=41, ragged=True)
    kc, vc, kc_ref, vc_ref, sc = build_caches(inp, cuda, fp8=fp8, poison=True)
    kw = dict(k_scale=sc[2], v_scale=sc[3]) if fp8 else {}
    okw = dict(kv_dtype=R.KV_FP8_E4M3, k_scale=sc[0], v_scale=sc[1]) if fp8 else {}
    plan = b200attn.ops.DecodePlan(kc, vc, 16, 32, inp["max_nb"], **kw)
    q, bt, sl = inp["q"].to(cuda), inp["block_tables"].to(cuda), inp["seq_lens"].to(cuda)
    for B in (12, 5, 1):
        ref = R.paged_decode_ref(inp["q"][:B], kc_ref, vc_ref, inp["block_tables"][:B], inp["seq_lens"][:B],
                                 return_f32=True, **okw)     # (reads only tokens < seq_len: the poison stays out)
        for splits in (0, 0, 1, 0, 4, 0):
            out = torch.full((B, 32, 128), float("nan"), dtype=torch.bfloat16, device=cuda)
            plan.run(q[:B], bt[:B].contiguous(), sl[:B].contiguous(), out, num_splits=splits)
            assert_close(out, ref, what=f"plan handle B={B} splits={splits} fp8={fp8}")
            stateless = b200attn.paged_decode(q[:B], kc, vc, bt[:B].contiguous(), sl[:B].contiguous(), num_splits=splits, **kw)
            assert torch.equal(bits(out), bits(stateless)), "plan handle and stateless call differ bitwise"
    with pytest.raises(b200attn.B2Error):
        plan.run(inp["q"].repeat(2, 1, 1).to(cuda), bt.repeat(2, 1), sl.repeat(2),
                 torch.empty(24, 32, 128, dtype=torch.bfloat16, device=cuda))
    plan.close()


def test_decode_two_box_fallback_map(cuda, monkeypatch):
    """B2_TWO_BOX=1 builds the bf16 stage image from two TMA boxes per tile (the path taken when a driver refuses the
    one-box [64][row][half] map): same shared-memory layout, so the same bits."""
    inp = R.make_inputs(6, 16, 2, 128, 16, 500, seed=43, ragged=True)
    kc, vc, kc_ref, vc_ref, _ = build_caches(inp, cuda)
    ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], inp["seq_lens"], return_f32=True)
    outs = {}
    for two in ("0", "1"):
        monkeypatch.setenv("B2_TWO_BOX", two)
        for splits in (0, 1):
            outs[two, splits] = run_decode(inp, kc, vc, cuda, num_splits=splits)
            assert_close(outs[two, splits], ref, what=f"two_box={two} splits={splits}")
    for splits in (0, 1):
        assert torch.equal(bits(outs["0", splits]), bits(outs["1", splits]))


def test_workspace_for_max_batch_serves_every_smaller_batch(cuda):
    """one workspace sized for max_num_seqs = 32 at max_model_len 8192 must be accepted by every
    batch size below it under the dynamic schedule (B = 24 needs more partial slots than B = 32)."""
    Hq, Hkv, bs, max_nb = 32, 8, 16, 512
    ws = torch.empty(b200attn.decode_workspace_bytes(32, Hq, Hkv, 128, bs, max_nb, 0), dtype=torch.uint8, device=cuda)
    kc = torch.zeros(64, Hkv, bs, 128, dtype=torch.bfloat16, device=cuda)
    vc = torch.zeros_like(kc)
    for B in range(1, 33):
        q = torch.zeros(B, Hq, 128, dtype=torch.bfloat16, device=cuda)
        bt = torch.zeros(B, max_nb, dtype=torch.int32, device=cuda)
        sl = torch.full((B,), 40, dtype=torch.int32, device=cuda)
        out = b200attn.paged_decode(q, kc, vc, bt, sl, workspace=ws, num_splits=0)
        assert float(out.float().abs().max()) == 0.0
    torch.cuda.synchronize()


# ------------------------------------------------------------------------------------------
# our kernels against vLLM's own kernels' outputs (committed fixture; see tests/test_vllm_crosscheck.py)
# ------------------------------------------------------------------------------------------
def _vllm_cases():
    p = GOLDEN / "vllm_crosscheck.pt"
    return torch.load(p) if p.exists() else {}


@pytest.mark.parametrize("name", sorted(_vllm_cases()))
def test_kernels_match_vllm_fixture(cuda, name):
    import hashlib
    r = _vllm_cases()[name]
    kw, fp8 = r["make_inputs"], r["fp8"]
    inp = R.make_inputs(**kw)
    shape = (inp["num_blocks"], kw["Hkv"], kw["bs"], kw["d"])
    kc = torch.zeros(shape, dtype=torch.uint8 if fp8 else torch.bfloat16, device=cuda)
    vc = torch.zeros_like(kc)
    ksd = vsd = None
    if fp8:
        ksd = torch.tensor([r["k_scale"]], dtype=torch.float32, device=cuda)
        vsd = torch.tensor([r["v_scale"]], dtype=torch.float32, device=cuda)
    b200attn.reshape_and_cache(inp["k"].to(cuda), inp["v"].to(cuda), kc, vc, inp["slot_mapping"].to(cuda), ksd, vsd)
    sha = lambda t: hashlib.sha256(bits(t).numpy().tobytes()).hexdigest()  # noqa: E731
    assert sha(kc) == r["k_cache_sha"] and sha(vc) == r["v_cache_sha"], "KV write differs from vLLM's bits"
    for splits in (0, 1):
        out = b200attn.paged_decode(inp["q"].to(cuda), kc, vc, inp["block_tables"].to(cuda), inp["seq_lens"].to(cuda),
                                    k_scale=ksd, v_scale=vsd, num_splits=splits)
        assert_close(out, r["out_v1"].float(), atol=4e-3, rtol=1e-2, what=f"{name} vs vLLM v1 (splits {splits})")
