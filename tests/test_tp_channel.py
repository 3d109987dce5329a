"""CPU: the tensor-parallel step channel (b200attn.tp.StepChannel) and the step message codec -- how rank 0 tells the
other ranks what it admitted before each step (`--tensor-parallel-size N`, reference benchmark-vllm/models.py:11)."""
import multiprocessing as mp
import os
import uuid

import pytest

from b200attn import tp
from b200attn.engine import Request, SamplingParams


def _reader(path, world, rank, n_msgs, q):
    ch = tp.StepChannel(path, world, rank, create=False, capacity=1 << 12)
    got = [ch.recv(timeout=20.0) for _ in range(n_msgs)]
    q.put((rank, got))


@pytest.mark.parametrize("world", [2, 4])
def test_step_channel_delivers_in_order_to_every_reader(world):
    path = f"/dev/shm/b200attn-test-{os.getpid()}-{uuid.uuid4().hex[:8]}"
    ch = tp.StepChannel(path, world, 0, create=True, capacity=1 << 12)
    try:
        msgs = [b"", b"a", bytes(range(256)) * 3, b"", os.urandom(4096), os.urandom(3 * 4096 + 17), b"x" * 4095, b""]
        ctx = mp.get_context("fork")
        q = ctx.Queue()
        procs = [ctx.Process(target=_reader, args=(path, world, r, len(msgs), q)) for r in range(1, world)]
        for p in procs:
            p.start()
        for m in msgs:
            ch.send(m, timeout=20.0)
        res = dict(q.get(timeout=30) for _ in procs)
        for p in procs:
            p.join(timeout=10)
        assert sorted(res) == list(range(1, world))
        for r, got in res.items():
            assert got == msgs, f"rank {r} saw a different message stream"
    finally:
        ch.unlink()


def test_step_channel_times_out_without_readers():
    path = f"/dev/shm/b200attn-test-{os.getpid()}-{uuid.uuid4().hex[:8]}"
    ch = tp.StepChannel(path, 2, 0, create=True, capacity=1 << 12)
    try:
        ch.send(b"first")                      # the slot is free
        with pytest.raises(TimeoutError):
            ch.send(b"second", timeout=0.05)   # nobody acknowledged the first
    finally:
        ch.unlink()


def test_step_message_codec():
    assert tp.encode_step([], []) == b""
    assert tp.decode_step(b"") == (tp.MSG_STEP, [], [])
    r = Request(7, [1, 2, 3], SamplingParams(max_tokens=5, temperature=0.5, seed=3, stop_token_ids=(9,)), seed=42)
    cmd, reqs, aborts = tp.decode_step(tp.encode_step([r], [4, 5]))
    assert cmd == tp.MSG_STEP and aborts == [4, 5]
    (rid, prompt, params, seed), = reqs
    assert (rid, prompt, seed) == (7, [1, 2, 3], 42) and params == r.params
    assert tp.decode_step(tp.encode_step([], [], tp.MSG_SHUTDOWN))[0] == tp.MSG_SHUTDOWN
