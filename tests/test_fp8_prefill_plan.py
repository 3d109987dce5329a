"""Host logic of the fp8 prefill path (b200attn.ops.fp8_prefill_plan): which cache blocks get dequantised and the block
table of the scratch cache.  Pure index arithmetic, no GPU."""
import pytest
import torch

from b200attn import ops


def test_plan_gathers_visible_blocks_in_chunk_order():
    bt = torch.tensor([[7, 3, 9, 0], [5, 1, 0, 0], [2, 8, 6, 4]], dtype=torch.int32)
    src, sbt = ops.fp8_prefill_plan(bt, kv_lens_host=[33, 16, 64], bs=16)     # 3, 1 and 4 blocks visible
    assert src.dtype == torch.int32 and src.tolist() == [7, 3, 9, 5, 2, 8, 6, 4]
    assert sbt.dtype == torch.int32 and sbt.shape == (3, 4)
    assert sbt[0, :3].tolist() == [0, 1, 2] and sbt[1, :1].tolist() == [3] and sbt[2].tolist() == [4, 5, 6, 7]
    # a chunk's scratch table points at the copies of its own blocks, position by position
    for c, n in enumerate([3, 1, 4]):
        assert src[sbt[c, :n].long()].tolist() == bt[c, :n].tolist()


def test_plan_handles_empty_chunks_and_rejects_mismatches():
    bt = torch.tensor([[4, 2], [1, 0]], dtype=torch.int32)
    src, sbt = ops.fp8_prefill_plan(bt, [0, 17], bs=16)
    assert src.tolist() == [1, 0] and sbt[1].tolist() == [0, 1]
    with pytest.raises(ValueError):
        ops.fp8_prefill_plan(bt, [5], bs=16)                 # one length per chunk
    with pytest.raises(ValueError):
        ops.fp8_prefill_plan(bt, [5, 100], bs=16)            # more blocks than the table holds
