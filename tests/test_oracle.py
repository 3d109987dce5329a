"""CPU tests of the oracle itself (oracle/paged_attention_ref.py): hand-checkable known answers,
the e4m3 codec against an independent implementation (torch's own cast, in range), and the committed
golden fixture.  The reference holds no vectors for this path ("parity unpinned"), so these KATs are
what pins the oracle."""
import hashlib
import json
from pathlib import Path

import pytest
import torch

import paged_attention_ref as R

GOLDEN = Path(__file__).parent / "golden"


def _cache(inp, kv_dtype=R.KV_BF16, ks=1.0, vs=1.0):
    dt = torch.bfloat16 if kv_dtype == R.KV_BF16 else torch.uint8
    kc = torch.zeros(inp["num_blocks"], inp["Hkv"], inp["bs"], inp["d"], dtype=dt)
    vc = torch.zeros_like(kc)
    R.reshape_and_cache_ref(inp["k"], inp["v"], kc, vc, inp["slot_mapping"], kv_dtype, ks, vs)
    return kc, vc


def test_e4m3_decode_all_codes_match_torch():
    b = torch.arange(256, dtype=torch.uint8)
    ours = R.e4m3_decode(b)
    ref = b.view(torch.float8_e4m3fn).to(torch.float32)
    nan = torch.isnan(ref)
    assert torch.equal(torch.isnan(ours), nan)
    assert torch.equal(ours[~nan], ref[~nan])


def test_e4m3_encode_roundtrip_and_rne():
    b = torch.arange(256, dtype=torch.uint8)
    v = R.e4m3_decode(b)
    ok = ~torch.isnan(v)
    assert torch.equal(R.e4m3_encode(v)[ok], b[ok])
    # in-range values agree with torch's cast (an independent RNE implementation)
    g = torch.Generator().manual_seed(0)
    for s in (100.0, 1.0, 0.01, 0.001):
        x = (torch.randn(200000, generator=g) * s).clamp(-448, 448)
        assert torch.equal(R.e4m3_encode(x), x.to(torch.float8_e4m3fn).view(torch.uint8))
    # ties to even on the subnormal grid (quantum 2^-9) and at a binade edge
    x = torch.tensor([2.0 ** -10, 3 * 2.0 ** -10, 5 * 2.0 ** -10, 15.5 * 2.0 ** -9, 0.0, -0.0])
    assert R.e4m3_encode(x).tolist() == [0, 2, 2, 16, 0, 128]


def test_e4m3_encode_saturates():
    x = torch.tensor([448.0, 449.0, 464.0, 480.0, 1e9, float("inf"), -1e9, -float("inf")])
    assert R.e4m3_encode(x).tolist() == [0x7E, 0x7E, 0x7E, 0x7E, 0x7E, 0x7E, 0xFE, 0xFE]
    assert int(R.e4m3_encode(torch.tensor([float("nan")]))[0]) & 0x7F == 0x7F


def test_reshape_and_cache_semantics():
    inp = R.make_inputs(3, 4, 2, 128, 16, 40, seed=1, ragged=True, slack_blocks=3)
    slots = inp["slot_mapping"].clone()
    slots[::5] = -1                      # padding tokens are skipped
    kc = torch.full((inp["num_blocks"], 2, 16, 128), 7.0, dtype=torch.bfloat16)
    vc = torch.full_like(kc, 7.0)
    R.reshape_and_cache_ref(inp["k"], inp["v"], kc, vc, slots)
    for t, s in enumerate(slots.tolist()):
        if s < 0:
            continue
        assert torch.equal(kc[s // 16, :, s % 16, :], inp["k"][t])
        assert torch.equal(vc[s // 16, :, s % 16, :], inp["v"][t])
    written = torch.zeros(inp["num_blocks"] * 16, dtype=torch.bool)
    written[slots[slots >= 0]] = True
    untouched = kc.permute(0, 2, 1, 3).reshape(-1, 2, 128)[~written]
    assert torch.all(untouched == 7.0)


def test_decode_kat_seq_len_one_returns_v_row():
    inp = R.make_inputs(2, 8, 2, 128, 16, 1, seed=2)
    kc, vc = _cache(inp)
    out = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"])
    for b in range(2):
        for hq in range(8):
            assert torch.equal(out[b, hq], inp["v"][b, hq // 4])


def test_decode_kat_equal_keys_gives_mean_of_v():
    inp = R.make_inputs(1, 4, 1, 128, 16, 37, seed=3)
    inp["k"][:] = inp["k"][0:1]
    kc, vc = _cache(inp)
    out = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"], return_f32=True)
    mean = inp["v"].float().mean(dim=0)          # [Hkv=1, d]
    assert torch.allclose(out[0], mean.expand(4, 128), atol=1e-5)


def test_decode_kat_dominant_key():
    inp = R.make_inputs(1, 2, 1, 128, 16, 50, seed=4)
    inp["k"][:] = 0
    inp["k"][17, 0] = inp["q"][0, 0] * 8          # one key strongly aligned with q head 0
    kc, vc = _cache(inp)
    out = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"], return_f32=True)
    assert torch.allclose(out[0, 0], inp["v"][17, 0].float(), atol=1e-3)


def test_decode_zero_length_and_ragged_tail():
    inp = R.make_inputs(3, 8, 2, 128, 16, 33, seed=5)
    kc, vc = _cache(inp)
    sl = torch.tensor([0, 17, 33], dtype=torch.int32)
    out = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], sl, return_f32=True)
    assert torch.all(out[0] == 0)
    # positions >= seq_len do not attend: poisoning them changes nothing
    kc2, vc2 = kc.clone(), vc.clone()
    blk = int(inp["block_tables"][1, 1])
    kc2[blk, :, 1:, :] = float("nan")
    vc2[blk, :, 1:, :] = float("nan")
    out2 = R.paged_decode_ref(inp["q"], kc2, vc2, inp["block_tables"], sl, return_f32=True)
    assert torch.equal(out[1], out2[1])


def test_decode_block_table_permutation_invariance():
    inp = R.make_inputs(2, 8, 2, 128, 16, 100, seed=6)
    kc, vc = _cache(inp)
    out = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"])
    perm = torch.randperm(inp["num_blocks"], generator=torch.Generator().manual_seed(9))
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(inp["num_blocks"])
    kc2, vc2 = kc[perm], vc[perm]                 # physical block p now holds old block perm[p]
    bt2 = inv[inp["block_tables"].long()].to(torch.int32)
    out2 = R.paged_decode_ref(inp["q"], kc2, vc2, bt2, inp["seq_lens"])
    assert torch.equal(out, out2)


@pytest.mark.parametrize("partition", [16, 64, 500])
def test_split_k_merge_matches_single_pass(partition):
    inp = R.make_inputs(3, 8, 2, 128, 16, 300, seed=7, ragged=True)
    kc, vc = _cache(inp)
    a = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"], return_f32=True)
    b = R.paged_decode_partitioned_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"], partition).float()
    assert torch.allclose(a, b, atol=2e-3, rtol=1e-2)


def test_fp8_cache_decode_close_to_unquantised():
    inp = R.make_inputs(2, 8, 2, 128, 16, 200, seed=8)
    ks = float(inp["k"].float().abs().max()) / 448.0
    vs = float(inp["v"].float().abs().max()) / 448.0
    kc, vc = _cache(inp)
    kq, vq = _cache(inp, R.KV_FP8_E4M3, ks, vs)
    a = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"], return_f32=True)
    b = R.paged_decode_ref(inp["q"], kq, vq, inp["block_tables"], inp["seq_lens"], kv_dtype=R.KV_FP8_E4M3,
                           k_scale=ks, v_scale=vs, return_f32=True)
    # quantisation error bound vs the unquantised oracle (looser, stated): 5e-2 abs + 1e-1 rel
    assert torch.all((a - b).abs() <= 5e-2 + 1e-1 * a.abs())


def test_algorithmic_bytes_matches_baseline_md():
    # BASELINE.md section 3
    assert R.algorithmic_bytes(4, 8, 2, 128, 16, [512] * 4, 2) == 2_114_064
    assert R.algorithmic_bytes(64, 32, 8, 128, 16, [4096] * 64, 2) == 1_074_856_192
    assert R.algorithmic_bytes(128, 32, 8, 128, 16, [8192] * 128, 1) == 2_149_843_456
    assert R.algorithmic_bytes(256, 64, 8, 128, 16, [8192] * 256, 2) == 8_598_848_512


def test_golden_fixture_config1():
    """tests/golden/config1.json was produced by tests/golden/make_golden.py from this oracle;
    it freezes the seeded inputs and the oracle output so a silent change of either is caught."""
    meta = json.loads((GOLDEN / "config1.json").read_text())
    inp = R.make_inputs(**meta["shape"], seed=meta["seed"])
    kc, vc = _cache(inp)
    out = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"])

    def sha(t):
        return hashlib.sha256(t.contiguous().view(torch.uint8).numpy().tobytes()).hexdigest()

    assert sha(inp["q"]) == meta["sha256"]["q"]
    assert sha(kc) == meta["sha256"]["k_cache"]
    assert sha(inp["block_tables"]) == meta["sha256"]["block_tables"]
    ref = torch.load(GOLDEN / "config1_out.pt")
    assert torch.allclose(out.float(), ref.float(), atol=1e-3, rtol=0)   # within 1-2 bf16 ulp across BLAS builds
