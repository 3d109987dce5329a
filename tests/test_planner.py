"""Host-side schedule planner (b2_paged_decode_plan_splits): pure host arithmetic, no GPU."""
import pytest
import torch

import b200attn


def plan(lens, Hq=32, Hkv=8):
    return b200attn.plan_num_splits(torch.tensor(lens, dtype=torch.int32), Hq, Hkv)


def test_uniform_large_batches_take_the_single_pass_grid():
    assert plan([4096] * 64) == 1                 # BASELINE configs[1]: 512 items on 296 CTA slots
    assert plan([8192] * 128) == 1                # configs[2]
    assert plan([8192] * 256, 64, 8) == 1         # configs[3], tp = 1
    assert plan([8192] * 256, 8, 1) == 1          # configs[3] at tp = 8: 256 equal items fill 86 % of one wave (measured
                                                  # 6.7 TB/s single pass vs 6.1 dynamic)
    KV_FP8 = b200attn.KV_FP8_E4M3
    lens = torch.tensor([8192] * 128, dtype=torch.int32)
    assert b200attn.plan_num_splits(lens, 32, 8, 128, 16, KV_FP8) == 0   # fp8 (3 CTAs per SM): 2.3 waves lose to dynamic


def test_small_or_skewed_batches_take_the_dynamic_schedule():
    assert plan([512] * 4, 8, 2) == 0             # configs[0]: 8 items
    assert plan([8192] * 100, 8, 1) == 0          # 100 items on 296 slots: a third of the SMs would idle
    assert plan([32768] + [200] * 63) == 0        # one long prompt among short ones
    assert plan(list(range(100, 6500, 100))) == 0


def test_ragged_serving_batches_take_the_dynamic_schedule():
    """log-uniform 128..ctx lengths: the per-item pipeline ramp of the single-pass grid costs more than the dynamic
    schedule's ~4 % (measured +9 % / +7 % for configs[3] / configs[1] shapes, scripts/gpu_ragged.sh)."""
    import math
    g = torch.Generator().manual_seed(1234)
    for B, ctx, Hq, Hkv in ((256, 8192, 64, 8), (64, 4096, 32, 8)):
        lens = torch.exp(torch.empty(B).uniform_(math.log(128.0), math.log(float(ctx)), generator=g)).to(torch.int32)
        assert b200attn.plan_num_splits(lens.clamp_(128, ctx).contiguous(), Hq, Hkv) == 0


def test_bad_arguments():
    import pytest
    with pytest.raises(ValueError):
        b200attn.plan_num_splits(torch.tensor([1, 2], dtype=torch.int64), 8, 2)
    with pytest.raises(b200attn.B2Error):
        b200attn.plan_num_splits(torch.tensor([1, 2], dtype=torch.int32), 7, 2)


def test_append_fused_rule():
    """b2_paged_decode_append rides inside the decode grid only where one CTA reads a (row, kv head): the single-pass grid
    (num_splits = 1) with all q heads of a kv head in one CTA (G <= 16).  Pure host code."""
    import b200attn
    assert b200attn.append_is_fused(32, 8, 1) and b200attn.append_is_fused(64, 8, 1) and b200attn.append_is_fused(16, 1, 1)
    assert not b200attn.append_is_fused(32, 8, 0)        # dynamic persistent: rows are cut into chunks over many CTAs
    assert not b200attn.append_is_fused(32, 8, 4)        # split-K
    assert not b200attn.append_is_fused(40, 2, 1)        # G = 20: two head-group CTAs share a kv head
    with pytest.raises(b200attn.B2Error):
        b200attn.append_is_fused(30, 8, 1)               # Hq not a multiple of Hkv
