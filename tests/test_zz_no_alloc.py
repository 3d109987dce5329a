"""b2_paged_decode_static: the decode entry point that allocates nothing (its split-K scratch is a static device
array inside the library, include/b200attn.h).  Called through the C ABI and checked against b2_paged_decode on the
same inputs, two calls in a row on the shared array, plus the over-the-limit refusal.  Collected last on purpose
(the file name)."""
import ctypes as C

import pytest
import torch

import b200attn
from b200attn import _lib
import paged_attention_ref as R

pytestmark = pytest.mark.gpu

ATOL, RTOL = 2e-3, 1e-2            # the tolerance of tests/test_gpu_parity.py
STATIC_BYTES = 32 << 20            # B2_STATIC_SCRATCH_BYTES


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _caches(inp, dev, fp8):
    shape = (inp["num_blocks"], inp["Hkv"], inp["bs"], inp["d"])
    k, v, slots = inp["k"].to(dev), inp["v"].to(dev), inp["slot_mapping"].to(dev)
    if not fp8:
        kc = torch.zeros(shape, dtype=torch.bfloat16, device=dev)
        vc = torch.zeros(shape, dtype=torch.bfloat16, device=dev)
        b200attn.reshape_and_cache(k, v, kc, vc, slots)
        return kc, vc, None, None, 0
    ks = torch.tensor([float(inp["k"].float().abs().max()) / 448.0], dtype=torch.float32, device=dev)
    vs = torch.tensor([float(inp["v"].float().abs().max()) / 448.0], dtype=torch.float32, device=dev)
    kc = torch.zeros(shape, dtype=torch.uint8, device=dev)
    vc = torch.zeros(shape, dtype=torch.uint8, device=dev)
    b200attn.reshape_and_cache(k, v, kc, vc, slots, ks, vs)
    return kc, vc, ks, vs, 1


def _both(lib, inp, dev, fp8, splits):
    """(output of b2_paged_decode, a closure running b2_paged_decode_static into a fresh NaN-filled buffer)."""
    kc, vc, ks, vs, kvd = _caches(inp, dev, fp8)
    q = inp["q"].to(dev)
    bt, sl = inp["block_tables"].to(dev), inp["seq_lens"].to(dev)
    B, Hq, d = q.shape
    nb, Hkv, bs, _ = kc.shape
    max_nb = bt.shape[1]
    geom = (B, Hq, Hkv, d, bs, max_nb, nb, Hq * d, Hq * d, kc.stride(0), kc.stride(1), 1.0 / d ** 0.5, kvd)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    need = lib.b2_paged_decode_workspace_bytes(B, Hq, Hkv, d, bs, max_nb, splits)
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device=dev)
    ref = torch.full((B, Hq, d), float("nan"), dtype=torch.bfloat16, device=dev)
    _lib.check(lib.b2_paged_decode(_ptr(ref), _ptr(q), _ptr(kc), _ptr(vc), _ptr(bt), _ptr(sl), *geom,
                                   _ptr(ks), _ptr(vs), _ptr(ws), need, splits, stream))
    keep = (kc, vc, ks, vs, q, bt, sl)

    def static_call():
        out = torch.full((B, Hq, d), float("nan"), dtype=torch.bfloat16, device=dev)
        _lib.check(lib.b2_paged_decode_static(_ptr(out), _ptr(q), _ptr(keep[0]), _ptr(keep[1]), _ptr(bt), _ptr(sl),
                                              *geom, _ptr(ks), _ptr(vs), splits, stream))
        torch.cuda.synchronize(dev)
        return out

    torch.cuda.synchronize(dev)
    return ref, static_call


def _check(got, ref, what):
    got, ref = got.float().cpu(), ref.float().cpu()
    assert torch.isfinite(got).all(), f"{what}: non-finite output"
    err = (got - ref).abs()
    bad = err > ATOL + RTOL * ref.abs()
    assert not bad.any(), f"{what}: {int(bad.sum())}/{bad.numel()} outside {ATOL}+{RTOL}|ref|, max err {float(err.max()):.3e}"


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("splits", [0, 4, 1])
def test_static_scratch_decode_matches_workspace_decode_twice(cuda, fp8, splits):
    lib = b200attn.load()
    # two different problems back to back on the one shared array: ragged 24 x up to 3000 tokens, then config 1
    # (4 sequences) -- the second call finds the first call's partials in the scratch and must not read them
    first = R.make_inputs(24, 32, 8, 128, 16, 3000, seed=11, ragged=True)
    second = R.make_inputs(4, 8, 2, 128, 16, 512, seed=0)
    ref1, call1 = _both(lib, first, cuda, fp8, splits)
    ref2, call2 = _both(lib, second, cuda, fp8, splits)
    _check(call1(), ref1, f"first call (fp8={fp8}, splits={splits})")
    _check(call2(), ref2, f"second call (fp8={fp8}, splits={splits})")
    _check(call1(), ref1, f"third call, first problem again (fp8={fp8}, splits={splits})")


def test_static_scratch_decode_allocates_nothing(cuda):
    lib = b200attn.load()
    inp = R.make_inputs(16, 32, 8, 128, 16, 2048, seed=3, ragged=True)
    ref, call = _both(lib, inp, cuda, False, 0)
    call()                                            # set-up (symbol address) done
    free0, _ = torch.cuda.mem_get_info(cuda)
    out = call()
    free1, _ = torch.cuda.mem_get_info(cuda)
    _check(out, ref, "allocation check call")
    # `out` itself comes from torch's caching allocator (a 128 KiB block already cached after the first call)
    assert free1 == free0, f"device free memory changed across the call: {free0} -> {free1}"


def test_static_scratch_decode_rejects_sizes_beyond_the_array(cuda):
    lib = b200attn.load()
    B, Hq, Hkv, d, bs, max_nb = 256, 32, 8, 128, 16, 2048
    assert lib.b2_paged_decode_workspace_bytes(B, Hq, Hkv, d, bs, max_nb, 0) > STATIC_BYTES
    nb = 64
    kc = torch.zeros(nb, Hkv, bs, d, dtype=torch.bfloat16, device=cuda)
    q = torch.zeros(B, Hq, d, dtype=torch.bfloat16, device=cuda)
    out = torch.zeros_like(q)
    bt = torch.zeros(B, max_nb, dtype=torch.int32, device=cuda)
    sl = torch.ones(B, dtype=torch.int32, device=cuda)
    rc = lib.b2_paged_decode_static(_ptr(out), _ptr(q), _ptr(kc), _ptr(kc), _ptr(bt), _ptr(sl), B, Hq, Hkv, d, bs,
                                    max_nb, nb, Hq * d, Hq * d, kc.stride(0), kc.stride(1), 0.1, 0, None, None, 0,
                                    C.c_void_p(torch.cuda.current_stream(cuda).cuda_stream))
    assert rc == -4                                    # B2_ENOMEM
    msg = lib.b2_last_error().decode()
    assert "B2_STATIC_SCRATCH_BYTES" in msg and "b2_paged_decode" in msg, msg
