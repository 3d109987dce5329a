"""Host-side model of the persistent (stream-K) schedule in csrc/paged_decode.cu.

The device code derives everything from integer arithmetic on the flattened unit space
(range_begin, effective_ctas, the slot rule, the boundary-merge ownership rule).  This test restates
that arithmetic in Python and checks, on random ragged batches, the invariants the kernels rely on:
  * the CTA ranges tile the unit space and every segment lies inside one item;
  * a CTA writes at most two partial segments: slot 0 if the segment is the first it processes, else slot 1;
  * merge_boundary_kernel's rules (owner = first boundary an item crosses, slot = range_begin(cc) >= s_i ? 0 : 1,
    c_last = ((e_i*NC) - 1) // U) visit exactly the slots that were written for that item, once."""
import random

MIN_TOK_PER_CTA = 128


def range_begin(c, U, NC):
    return (c * U) // NC


def effective_ctas(U, grid, min_units):
    want = (U + min_units - 1) // min_units
    return min(grid, max(want, 1))


def main_kernel(seq_lens, IPB, bs, grid):
    """-> (written: {(cta, slot): item}, whole: set(item), NC, U, item_range: {item: (s, e)})"""
    nbs = [(L + bs - 1) // bs for L in seq_lens]
    U = sum(nbs) * IPB
    written, whole, item_range = {}, set(), {}
    acc = 0
    for b, nb in enumerate(nbs):
        for j in range(IPB):
            if nb:
                item_range[(b, j)] = (acc, acc + nb)
            acc += nb
    if U == 0:
        return written, whole, 0, 0, item_range
    NC = effective_ctas(U, grid, MIN_TOK_PER_CTA // bs)
    covered = 0
    for c in range(NC):
        u0, u1 = range_begin(c, U, NC), range_begin(c + 1, U, NC)
        assert u0 == covered
        covered = u1
        first = True
        u = u0
        while u < u1:
            # locate (b, j, blk) of unit u the slow way
            item = next(k for k, (s, e) in item_range.items() if s <= u < e)
            s, e = item_range[item]
            take = min(e - u, u1 - u)
            if u == s and take == e - s:
                assert item not in whole
                whole.add(item)
            else:
                slot = 0 if first else 1
                assert (c, slot) not in written, "a CTA wrote the same slot twice"
                written[(c, slot)] = item
            first = False
            u += take
    assert covered == U
    return written, whole, NC, U, item_range


def merge_kernel(seq_lens, IPB, bs, grid, NC, U, item_range):
    """-> {item: [(cta, slot), ...]} as merge_boundary_kernel would read them"""
    reads = {}
    for c in range(grid):
        if U == 0 or c + 1 >= NC:
            continue
        ub = range_begin(c + 1, U, NC)
        if ub == 0 or ub >= U:
            continue
        item = next(k for k, (s, e) in item_range.items() if s <= ub < e)
        s_i, e_i = item_range[item]
        if ub == s_i:                      # blk == 0: boundary between two items
            continue
        if s_i < range_begin(c, U, NC):    # crossed the previous boundary too
            continue
        c_last = (e_i * NC - 1) // U
        assert item not in reads, "two blocks merged the same item"
        reads[item] = [(cc, 0 if range_begin(cc, U, NC) >= s_i else 1) for cc in range(c, c_last + 1)]
    return reads


def check(seq_lens, IPB, bs, grid):
    written, whole, NC, U, item_range = main_kernel(seq_lens, IPB, bs, grid)
    reads = merge_kernel(seq_lens, IPB, bs, grid, NC, U, item_range)
    # every item is produced exactly one way
    for item in item_range:
        assert (item in whole) != (item in reads), item
    by_item = {}
    for key, item in written.items():
        by_item.setdefault(item, []).append(key)
    for item, slots in reads.items():
        assert sorted(slots) == sorted(by_item[item]), (item, slots, by_item[item])
    assert set(by_item) == set(reads)


def test_uniform_headline_config():
    check([4096] * 64, IPB=8, bs=16, grid=296)


def test_tiny_and_empty():
    check([0, 0, 0], 2, 16, 296)
    check([1], 1, 16, 296)
    check([0, 5, 0, 17, 0], 2, 16, 296)
    check([512] * 4, 2, 16, 296)


def test_random_ragged_batches():
    rng = random.Random(0)
    for _ in range(150):
        B = rng.randint(1, 40)
        bs = rng.choice([8, 16, 32])
        IPB = rng.choice([1, 2, 3, 8])
        grid = rng.choice([1, 2, 7, 296])
        lens = [rng.choice([0, 1, rng.randint(1, 40), rng.randint(1, 3000)]) for _ in range(B)]
        check(lens, IPB, bs, grid)
