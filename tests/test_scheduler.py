"""Continuous-batching scheduler (C++ behind the C ABI, b2_sched_*): invariants on simulated request streams.
Host only; the limits are the reference's launch parameters (benchmark-vllm/models.py:8-10,15)."""
import random

import numpy as np
import pytest

import b200attn


def drive(sched, reqs, arrivals=None, eos=None, max_steps=100000):
    """reqs: {id: (prompt_len, max_new)}; arrivals: {step: [ids]} (default: all at step 0); eos: {id: stop after n tokens}.
    Mirrors what an engine does with each plan and checks the invariants every step.  Returns per-request stats."""
    arrivals = arrivals or {0: list(reqs)}
    eos = eos or {}
    bs, max_nb = sched.block_size, sched.max_nb
    kv = {}            # id -> tokens the engine believes are in the cache
    sampled = {i: 0 for i in reqs}
    done, preemptions, steps = set(), 0, 0
    pending = sum(len(v) for v in arrivals.values())
    while True:
        for i in arrivals.get(steps, []):
            sched.add_request(i, *reqs[i])
            pending -= 1
        plan = sched.step()
        for i in plan.finished_ids:
            assert sampled[i] >= 1 and i not in done
            done.add(int(i))
            kv.pop(int(i), None)
        for i in plan.preempted_ids:
            kv.pop(int(i), None)
            preemptions += 1
        assert len(plan.decode_ids) + int(plan.prefill_len.sum()) <= sched_budget(sched)
        assert sched.num_running <= sched.max_num_seqs
        live_slots = set()
        # decode rows
        for r, i in enumerate(plan.decode_ids):
            i = int(i)
            assert kv[i] + 1 == plan.seq_lens[r], (i, kv[i], plan.seq_lens[r])
            pos = kv[i]
            blk = plan.block_tables[r, pos // bs]
            assert plan.slot_mapping[r] == blk * bs + pos % bs
            kv[i] += 1
            sampled[i] += 1
        # prefill chunks
        off = 0
        for c, i in enumerate(plan.prefill_ids):
            i = int(i)
            n, st = int(plan.prefill_len[c]), int(plan.prefill_start[c])
            assert st == kv.get(i, 0) and n > 0
            slots = plan.prefill_slots[off:off + n]
            off += n
            pos = np.arange(st, st + n)
            assert np.array_equal(slots, plan.prefill_block_tables[c, pos // bs].astype(np.int64) * bs + pos % bs)
            kv[i] = st + n
            if plan.prefill_last[c] == 1:
                assert sampled[i] == 0 and kv[i] == reqs[i][0]
                sampled[i] = 1
            elif plan.prefill_last[c] == 2:
                assert sampled[i] >= 1 and kv[i] == reqs[i][0] + sampled[i] - 1   # recomputed prompt + g - 1 tokens
        assert off == len(plan.prefill_slots)
        # no block is owned by two live sequences
        owner = {}
        tables = [(int(i), plan.block_tables[r]) for r, i in enumerate(plan.decode_ids)] + \
                 [(int(i), plan.prefill_block_tables[c]) for c, i in enumerate(plan.prefill_ids)]
        for i, row in tables:
            for b in row[: (kv[i] + bs - 1) // bs]:
                assert owner.setdefault(int(b), i) == i, f"block {b} shared by {owner[int(b)]} and {i}"
        del live_slots
        # EOS decided by the engine
        for i, n in eos.items():
            if i not in done and sampled[i] >= n and i in kv:
                sched.finish(i)
                done.add(i)
                kv.pop(i)
        steps += 1
        if plan.empty and pending == 0 and sched.num_waiting == 0 and sched.num_running == 0:
            break
        assert steps < max_steps, "scheduler does not drain"
    assert sched.num_free_blocks == sched.num_blocks
    return dict(sampled=sampled, done=done, preemptions=preemptions, steps=steps)


def sched_budget(s):
    return s._budget


def make(num_blocks, bs, max_seqs, max_len, budget):
    s = b200attn.Scheduler(num_blocks, bs, max_seqs, max_len, budget)
    s._budget = budget
    return s


def test_all_requests_finish_with_their_token_budget():
    s = make(4096, 16, 8, 2048, 512)
    reqs = {i: (50 + 37 * i, 5 + i) for i in range(20)}
    st = drive(s, reqs)
    assert st["done"] == set(reqs) and st["preemptions"] == 0
    assert all(st["sampled"][i] == reqs[i][1] for i in reqs)


def test_chunked_prefill_of_a_long_prompt():
    s = make(4096, 16, 4, 40000, 2048)
    st = drive(s, {1: (32768, 3), 2: (128, 8)})
    assert st["done"] == {1, 2} and st["sampled"] == {1: 3, 2: 8}
    assert st["steps"] >= 32768 // 2048


def test_preemption_under_a_tight_pool_recomputes_and_completes():
    # 24 blocks of 16 = 384 token slots for 6 sequences that each want 40 + 60 tokens: must preempt
    s = make(24, 16, 6, 256, 128)
    reqs = {i: (40, 60) for i in range(6)}
    st = drive(s, reqs)
    assert st["done"] == set(reqs) and st["preemptions"] > 0
    assert all(st["sampled"][i] == 60 for i in reqs)


def test_eos_and_late_arrivals_and_max_model_len():
    s = make(512, 16, 4, 96, 64)
    reqs = {1: (30, 1000), 2: (10, 1000), 3: (20, 4), 4: (64, 50)}
    st = drive(s, reqs, arrivals={0: [1, 2], 5: [3], 9: [4]}, eos={2: 7})
    assert st["sampled"][2] == 7 and st["sampled"][3] == 4
    assert st["sampled"][1] == 96 - 30 and st["sampled"][4] == 96 - 64      # stopped by max_model_len
    assert st["done"] == {1, 2, 3, 4}


def test_random_streams():
    rng = random.Random(0)
    for trial in range(25):
        bs = rng.choice([8, 16, 32])
        max_len = rng.choice([128, 600, 3000])
        nblk = rng.randint(max_len // bs + 2, 6 * max_len // bs)
        s = make(nblk, bs, rng.randint(1, 12), max_len, rng.choice([32, 256, 2048]))
        reqs, arrivals = {}, {}
        for i in range(rng.randint(1, 30)):
            reqs[i] = (rng.randint(1, max_len - 1), rng.randint(1, 80))
            arrivals.setdefault(rng.randint(0, 40), []).append(i)
        st = drive(s, reqs, arrivals=arrivals, eos={i: rng.randint(1, 20) for i in reqs if rng.random() < 0.3})
        assert st["done"] == set(reqs)


def test_errors():
    s = make(8, 16, 2, 64, 64)
    s.add_request(1, 10, 5)
    with pytest.raises(b200attn.B2Error):
        s.add_request(1, 10, 5)
    with pytest.raises(b200attn.B2Error):
        s.add_request(2, 64, 5)          # no room under max_model_len
    with pytest.raises(b200attn.B2Error):
        s.add_request(3, 0, 5)
    with pytest.raises(b200attn.B2Error):
        s.finish(99)
    big = make(2, 16, 2, 4096, 64)
    with pytest.raises(b200attn.B2Error):
        big.add_request(1, 100, 5)       # can never fit a pool of 2 blocks


def test_repeated_preemption_of_the_same_request():
    # a pool so tight that requests are preempted again after they resumed decoding
    s = make(10, 8, 4, 200, 64)
    reqs = {i: (9, 50) for i in range(4)}
    st = drive(s, reqs)
    assert st["done"] == set(reqs) and st["preemptions"] >= 4
    assert all(st["sampled"][i] == 50 for i in reqs)


def test_request_that_outgrows_the_whole_pool_ends_by_length_instead_of_livelocking():
    """4 blocks of 16, prompt 40, 100 new tokens -- alone in the running set the request used to be
    preempted, recomputed and preempted again for ever (175 preemptions in 200 steps, generated stuck at 25).
    Its budget is now cut to what the pool can hold (64 cached tokens = 40 + 25 - 1) and it finishes by length."""
    s = make(4, 16, 2, 4096, 64)
    reqs = {1: (40, 100)}
    kv, sampled, finished, pre = 0, 0, False, 0
    s.add_request(1, 40, 100)
    for _ in range(200):
        plan = s.step()
        pre += len(plan.preempted_ids)
        if 1 in plan.finished_ids:
            finished = True
            break
        sampled += len(plan.decode_ids) + int((plan.prefill_last == 1).sum())
    assert finished and pre == 0
    assert sampled == 64 - 40 + 1                      # the pool holds 64 tokens: 40 prompt + 24 generated cached, 25 sampled
    assert s.num_free_blocks == 4 and s.num_running == 0 and s.num_waiting == 0
    # and a second request behind it is served
    st = drive(s, {2: (10, 5)})
    assert st["done"] == {2}


def test_memory_pressure_does_not_thrash_preempted_requests():
    """8 requests (prompts ~100 tokens, 400 new tokens) on a 64-block pool.  A victim used to be re-admitted
    in the very step that preempted it, on the blocks it had just freed, and be evicted again at the next block
    boundary: 171 preemptions and 26 176 prefill tokens scheduled for 908 prompt tokens.  Now a step that preempts
    admits nothing, and a preempted request returns only when its whole recompute fits."""
    s = make(64, 16, 8, 1024, 2048)
    rng = random.Random(3)
    reqs = {i: (rng.randint(90, 130), 400) for i in range(8)}
    prompt_tokens = sum(p for p, _ in reqs.values())
    prefill_tokens = [0]
    orig_step = s.step

    def counting_step():
        plan = orig_step()
        prefill_tokens[0] += int(plan.prefill_len.sum())
        assert not (len(plan.preempted_ids) and set(plan.preempted_ids.tolist()) & set(plan.prefill_ids.tolist())), \
            "a request preempted in this step was re-admitted in the same step"
        return plan

    s.step = counting_step
    st = drive(s, reqs)
    assert st["done"] == set(reqs) and all(st["sampled"][i] == 400 for i in reqs)
    # 64 blocks = 1024 token slots for 8 x ~500 tokens: preemption is unavoidable, thrashing is not
    assert st["preemptions"] <= 40, st["preemptions"]
    assert prefill_tokens[0] <= 12 * prompt_tokens, (prefill_tokens[0], prompt_tokens)
