"""Host logic: the C++ block manager behind the C ABI (b2_bm_*), checked against a pure-Python
model of the same allocator semantics."""
import numpy as np
import pytest

import b200attn
from b200attn import _lib


def test_alloc_append_free_roundtrip():
    bm = b200attn.BlockManager(10, 16)
    assert bm.num_free == 10
    bm.alloc_seq(1, 0)
    s = bm.append(1, 17)                      # 2 blocks
    assert bm.num_free == 8 and bm.seq_len(1) == 17
    blocks = s // 16
    assert (s % 16).tolist() == list(range(16)) + [0]
    assert len(set(blocks.tolist())) == 2
    s2 = bm.append(1, 15)                     # fills block 2 exactly, no new block
    assert bm.num_free == 8 and (s2 // 16 == blocks[-1]).all()
    bm.free_seq(1)
    assert bm.num_free == 10 and bm.num_seqs == 0


def test_tables_and_plan_decode():
    bm = b200attn.BlockManager(64, 16)
    for sid, n in ((10, 5), (11, 16), (12, 33)):
        bm.alloc_seq(sid)
        bm.append(sid, n)
    bt, sl = bm.fill_tables([12, 10, 11], max_nb=4)
    assert sl.tolist() == [33, 5, 16]
    assert (bt[1, 1:] == 0).all() and (bt[2, 1:] == 0).all() and bt[0, 3] == 0
    bt2, sl2, slots = bm.plan_decode([12, 10, 11], max_nb=4)
    assert sl2.tolist() == [34, 6, 17]
    # the new token of each sequence lands at position seq_len-1 of its table
    for b in range(3):
        pos = sl2[b] - 1
        assert slots[b] == bt2[b, pos // 16] * 16 + pos % 16
    assert bt2[2, 1] != 0 or bt2[2, 0] != bt2[2, 1]   # seq 11 crossed into a second block


def test_slots_never_alias_across_sequences():
    rng = np.random.default_rng(0)
    bm = b200attn.BlockManager(256, 8)
    owned = {}
    for sid in range(20):
        bm.alloc_seq(sid)
        owned[sid] = []
    for _ in range(300):
        sid = int(rng.integers(0, 20))
        n = int(rng.integers(1, 12))
        if sid in owned and bm.can_append(sid, n):
            owned[sid] += bm.append(sid, n).tolist()
        if rng.random() < 0.05 and sid in owned:
            bm.free_seq(sid)
            del owned[sid]
    allslots = [s for v in owned.values() for s in v]
    assert len(allslots) == len(set(allslots))
    used_blocks = {s // 8 for s in allslots}
    assert bm.num_free == 256 - len(used_blocks)


def test_errors():
    bm = b200attn.BlockManager(2, 16)
    bm.alloc_seq(1)
    with pytest.raises(b200attn.B2Error) as e:
        bm.alloc_seq(1)
    assert e.value.code == _lib.B2_EEXISTS
    with pytest.raises(b200attn.B2Error) as e:
        bm.append(2, 1)
    assert e.value.code == _lib.B2_ENOTFOUND
    bm.append(1, 32)
    assert not bm.can_append(1, 1)
    with pytest.raises(b200attn.B2Error) as e:
        bm.append(1, 1)
    assert e.value.code == _lib.B2_ENOMEM
    assert bm.seq_len(1) == 32                # failed append changed nothing
    with pytest.raises(b200attn.B2Error) as e:
        bm.plan_decode([1], max_nb=2)         # would exceed max_nb
    assert e.value.code == _lib.B2_EINVAL
    with pytest.raises(b200attn.B2Error):
        b200attn.BlockManager(0, 16)
