"""Host-side model of the dynamic persistent schedule in csrc/paged_decode.cu (schedule C).

The kernels derive everything from integer arithmetic on the flattened chunk space: the static share
S = floor(pct% * Ctot / NC) per CTA, the run-start rule, the slot rule and the workspace bound computed on the
host.  This test restates that arithmetic and checks, on random ragged batches and random claim orders, the
invariants the decode and merge kernels rely on:
  * static claims and dynamic claims tile the chunk space exactly once;
  * a run never spans two items; a run that covers its item is written directly, every other run writes the slot
    dyn_slot(first chunk), all slots distinct and below the bound the host sizes the workspace with;
  * merge_chunks_kernel's enumeration (dyn_is_start / dyn_slot over the item's chunks) visits exactly the slots
    that were written for that item, and skips exactly the items that were written directly.
"""
import random

TOK = 64


def chunks_of(L, ct):
    return (L + ct - 1) // ct if L > 0 else 0


def static_chunks(Ctot, NC, pct):
    return (Ctot * pct // 100) // NC


def is_start(c, k, S, stat_total):
    return k == 0 or c >= stat_total or c % S == 0


def slot_of(c, S, stat_total, NC):
    if c >= stat_total:
        return 2 * NC + (c - stat_total)
    x = c // S
    return 2 * x if c % S == 0 else 2 * x + 1


def host_slots(B, ipb, max_ctx, ct, pct, NC):
    ctot_max = B * ipb * ((max_ctx + ct - 1) // ct)
    return 2 * NC + ctot_max * (100 - pct) // 100 + NC + 2


def simulate(seq_lens, ipb, ct, pct, NC, rng):
    """-> (direct: set(item), written: {slot: item}, items: {item: (ci0, nck)})"""
    items, pos, c = {}, [], 0
    for b, L in enumerate(seq_lens):
        nck = chunks_of(L, ct)
        for j in range(ipb):
            if nck:
                items[(b, j)] = (c, nck)
            for k in range(nck):
                pos.append((b, j, k, nck))
            c += nck
    Ctot = c
    S = static_chunks(Ctot, NC, pct)
    stat_total = S * NC
    claimed = [0] * Ctot
    direct, written = set(), {}

    def finish_run(c0, k0, c_last):
        b, j, k, nck = pos[c_last]
        if k0 == 0 and k == nck - 1:
            assert (b, j) not in direct
            direct.add((b, j))
        else:
            s = slot_of(c0, S, stat_total, NC)
            assert s not in written, "slot written twice"
            written[s] = (b, j)

    # static claims: consecutive chunks of one item accumulate into one run
    for x in range(NC):
        c0 = None
        for cid in range(x * S, (x + 1) * S):
            claimed[cid] += 1
            b, j, k, nck = pos[cid]
            if c0 is None:
                c0, k0 = cid, k
            nxt_static = cid + 1 < (x + 1) * S
            cont = nxt_static and k + 1 < nck
            if not cont:
                finish_run(c0, k0, cid)
                c0 = None
    # dynamic claims: any CTA, any order; each chunk is its own run
    dyn = list(range(stat_total, Ctot))
    rng.shuffle(dyn)
    for cid in dyn:
        claimed[cid] += 1
        finish_run(cid, pos[cid][2], cid)
    assert all(v == 1 for v in claimed)
    return direct, written, items, Ctot, S, stat_total


def merge(items, S, stat_total, NC):
    """-> {item: [slots]} as merge_chunks_kernel reads them; items it skips are absent"""
    reads = {}
    for item, (ci0, nck) in items.items():
        if nck <= 1:
            continue
        starts = [c for c in range(ci0, ci0 + nck) if is_start(c, c - ci0, S, stat_total)]
        if len(starts) == 1:
            continue
        reads[item] = [slot_of(c, S, stat_total, NC) for c in starts]
    return reads


def check(seq_lens, ipb, ct, pct, NC, seed=0):
    rng = random.Random(seed)
    direct, written, items, Ctot, S, stat_total = simulate(seq_lens, ipb, ct, pct, NC, rng)
    reads = merge(items, S, stat_total, NC)
    bound = host_slots(len(seq_lens), ipb, max(seq_lens + [1]), ct, pct, NC)
    assert all(0 <= s < bound for s in written), (max(written), bound)
    by_item = {}
    for s, item in written.items():
        by_item.setdefault(item, []).append(s)
    for item in items:
        assert (item in direct) != (item in reads), item
    assert set(by_item) == set(reads)
    for item, slots in reads.items():
        assert sorted(slots) == sorted(by_item[item]), (item, slots, by_item[item])


def test_headline_config():
    for pct in (0, 50, 85, 100):
        for ct in (128, 256, 512):
            check([4096] * 64, ipb=8, ct=ct, pct=pct, NC=296)


def test_tiny_and_empty():
    check([512] * 4, ipb=2, ct=64, pct=85, NC=296)
    check([0, 0, 5, 0], ipb=2, ct=64, pct=85, NC=296)
    check([0, 0], ipb=1, ct=64, pct=85, NC=296)
    check([1], ipb=1, ct=512, pct=0, NC=4)


def test_random_ragged_batches():
    rng = random.Random(1)
    for trial in range(60):
        B = rng.randint(1, 90)
        ctx = rng.choice([70, 600, 4096, 9000])
        lens = [rng.choice([0, 1, rng.randint(1, ctx)]) if rng.random() < 0.2 else rng.randint(1, ctx) for _ in range(B)]
        check(lens, ipb=rng.choice([1, 2, 8]), ct=rng.choice([64, 128, 256, 512]), pct=rng.choice([0, 30, 70, 85, 100]),
              NC=rng.choice([4, 37, 296]), seed=trial)
