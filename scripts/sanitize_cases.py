#!/usr/bin/env python
"""Small-shape pass over every kernel family for compute-sanitizer (memcheck / racecheck /
initcheck / synccheck on the parity cases):

    compute-sanitizer --tool memcheck python scripts/sanitize_cases.py

All three decode schedules (single pass, grid split + merge, dynamic persistent + merge), bf16 and fp8 caches, ragged
lengths incl. 0 and 1, KV write (both encodings, skipped slots), the decode-step write over resident block tables,
tcgen05 and mma.sync prompt-chunk attention, the element-wise layer ops, sampling, and the tensor-parallel exchange
kernels with two emulated ranks on one device.  Every case is also checked against the CPU oracle, so a sanitizer run
that reports 0 errors ran the real code paths."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT / "vllm-backend-public_b200", ROOT / "oracle", ROOT / "tests"):
    sys.path.insert(0, str(p))

import torch  # noqa: E402

import b200attn  # noqa: E402
import b200attn.model  # noqa: E402
from b200attn import ops  # noqa: E402
import paged_attention_ref as R  # noqa: E402
import layer_ref as L  # noqa: E402

dev = torch.device("cuda:0")
only = set(sys.argv[1:])


def want(name):
    return not only or name in only


class Guard:
    """Stand-in for memcheck / initcheck where compute-sanitizer is not available: every buffer a kernel may write sits
    between two 4 KiB red zones that must still hold their pattern afterwards (out-of-bounds writes), and scratch
    buffers start as all-ones bytes = NaN (a read of scratch the kernel did not write first shows up in the result)."""
    PAD = 4096

    def __init__(self):
        self.bufs = []

    def alloc(self, shape, dtype, fill=None):
        n = int(torch.tensor(shape).prod()) * torch.empty(0, dtype=dtype).element_size()
        buf = torch.full((n + 2 * self.PAD,), 0xA5, dtype=torch.uint8, device=dev)
        inner = buf[self.PAD:self.PAD + n]
        inner.fill_(0xFF if fill == "nan" else 0)
        self.bufs.append((buf, n))
        return inner.view(dtype).view(shape)

    def check(self, what):
        torch.cuda.synchronize()
        for buf, n in self.bufs:
            assert bool((buf[:self.PAD] == 0xA5).all()) and bool((buf[self.PAD + n:] == 0xA5).all()), \
                f"{what}: a red zone was overwritten (out-of-bounds write)"
        self.bufs = []


def close(out, ref, what, atol=2e-3, rtol=1e-2):
    err = (out.float().cpu() - ref).abs()
    assert bool((err <= atol + rtol * ref.abs()).all()), f"{what}: max err {float(err.max())}"


def decode_cases():
    for fp8 in (False, True):
        for bs in (16, 8):
            inp = R.make_inputs(5, 8, 2, 128, bs, 300, seed=3)
            lens = torch.tensor([300, 1, 0, 77, 129], dtype=torch.int32)
            inp["seq_lens"] = lens
            shape = (inp["num_blocks"], 2, bs, 128)
            ks = vs = None
            G = Guard()
            if fp8:
                kc = G.alloc(shape, torch.uint8)
                vc = G.alloc(shape, torch.uint8)
                ks = torch.tensor([0.05], device=dev)
                vs = torch.tensor([0.04], device=dev)
            else:
                kc = G.alloc(shape, torch.bfloat16)
                vc = G.alloc(shape, torch.bfloat16)
            slots = inp["slot_mapping"].clone()
            slots[::7] = -1                                     # skipped tokens
            ops.reshape_and_cache(inp["k"].to(dev), inp["v"].to(dev), kc, vc, slots.to(dev), ks, vs)
            kc_ref = torch.zeros(shape, dtype=torch.uint8 if fp8 else torch.bfloat16)
            vc_ref = torch.zeros_like(kc_ref)
            R.reshape_and_cache_ref(inp["k"], inp["v"], kc_ref, vc_ref, slots, kv_dtype=1 if fp8 else 0,
                                    k_scale=0.05 if fp8 else 1.0, v_scale=0.04 if fp8 else 1.0)
            assert torch.equal(kc.cpu().view(torch.uint8), kc_ref.view(torch.uint8)), "KV write not bit-exact"
            ref = R.paged_decode_ref(inp["q"], kc_ref, vc_ref, inp["block_tables"], lens, kv_dtype=1 if fp8 else 0,
                                     k_scale=0.05 if fp8 else 1.0, v_scale=0.04 if fp8 else 1.0, return_f32=True)
            G.check(f"KV write fp8={fp8} bs={bs}")
            for splits in (1, 3, 0):
                runs = []
                for rep in range(3):                       # same launch three times: a race would show as differing bits
                    out = G.alloc((5, 8, 128), torch.bfloat16, fill="nan")
                    need = max(ops.decode_workspace_bytes(5, 8, 2, 128, bs, inp["block_tables"].shape[1], splits), 256)
                    ws = G.alloc((need,), torch.uint8, fill="nan")          # scratch starts as NaN
                    ops.paged_decode(inp["q"].to(dev), kc, vc, inp["block_tables"].to(dev), lens.to(dev),
                                     k_scale=ks, v_scale=vs, num_splits=splits, out=out, workspace=ws)
                    torch.cuda.synchronize()
                    assert bool(torch.isfinite(out.float()).all()), "NaN in the result: uninitialised scratch was read"
                    close(out, ref, f"decode fp8={fp8} bs={bs} splits={splits}")
                    runs.append(out.clone())
                    G.check(f"decode fp8={fp8} bs={bs} splits={splits}")
                assert all(torch.equal(runs[0].view(torch.int16), r.view(torch.int16)) for r in runs[1:]), \
                    f"decode fp8={fp8} bs={bs} splits={splits}: results differ between identical launches"
    print("decode cases ok")


def step_cases():
    from b200attn.step import DecodeStep
    B, Hq, Hkv, bs, ctx = 4, 8, 2, 16, 200
    inp = R.make_inputs(B, Hq, Hkv, 128, bs, ctx, seed=5)
    shape = (inp["num_blocks"], Hkv, bs, 128)
    kc = torch.zeros(shape, dtype=torch.bfloat16, device=dev)
    vc = torch.zeros_like(kc)
    ops.reshape_and_cache(inp["k"].to(dev), inp["v"].to(dev), kc, vc, inp["slot_mapping"].to(dev))
    step = DecodeStep(kc, vc, max_batch=B, Hq=Hq, max_nb=inp["block_tables"].shape[1])
    step.h_qkv.normal_()
    step.h_block_tables.copy_(inp["block_tables"])
    step.upload_block_tables()
    step.h_seq_lens.copy_(inp["seq_lens"])
    last = (inp["seq_lens"] - 1).long()
    step.h_slots.copy_(inp["block_tables"][torch.arange(B), last // bs].long() * bs + last % bs)
    for splits in (1, 0):
        step.run(B, splits)
    torch.cuda.synchronize()
    print("decode step ok")


def layer_cases():
    T, Hq, Hkv, d, bs = 150, 4, 2, 128, 16
    g = torch.Generator().manual_seed(0)
    x = torch.randn(T, 1024, generator=g).to(torch.bfloat16).to(dev)
    w = torch.randn(1024, generator=g).to(torch.bfloat16).to(dev)
    res = torch.randn(T, 1024, generator=g).to(torch.bfloat16).to(dev)
    ops.rms_norm(x, w, 1e-5, residual=res)
    ops.silu_and_mul(torch.randn(T, 2 * 704, generator=g).to(torch.bfloat16).to(dev))
    logits = torch.randn(6, 1000, generator=g).to(torch.bfloat16).to(dev)
    ops.sample(logits, torch.tensor([0, 1.0, 0.5, 0, 2.0, 1.0], device=dev), torch.arange(6, device=dev))
    # rotary + cache write + prompt-chunk attention (both kernels) over two chunks with a prefix
    nb = 40
    kc = torch.zeros(nb, Hkv, bs, d, dtype=torch.bfloat16, device=dev)
    vc = torch.zeros_like(kc)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * d, generator=g).to(torch.bfloat16).to(dev)
    cos_sin = b200attn.model.rope_table(512, d, 10000.0, None, dev)
    lens = [100, 50]
    bt = torch.arange(nb, dtype=torch.int32).view(2, 20)
    pos = torch.cat([torch.arange(n, dtype=torch.int32) for n in lens])
    slots = torch.cat([bt[c, torch.arange(n) // bs].long() * bs + torch.arange(n) % bs for c, n in enumerate(lens)])
    ops.rope_and_cache(qkv, pos.to(dev), cos_sin, kc, vc, slots.to(dev), Hq)
    q = qkv.view(T, Hq + 2 * Hkv, d)[:, :Hq]
    cu = torch.tensor([0, 100, 150], dtype=torch.int32, device=dev)
    cs = torch.zeros(2, dtype=torch.int32, device=dev)
    outs = []
    for impl in ("tc5", "mma"):
        os.environ["B2_PREFILL_IMPL"] = impl
        outs.append(ops.prefill_attention(q, kc, vc, bt.to(dev), cu, cs, total_tokens=T))
    os.environ.pop("B2_PREFILL_IMPL", None)
    torch.cuda.synchronize()
    assert float((outs[0].float() - outs[1].float()).abs().max()) < 2e-2, "prefill kernels disagree"
    print("layer cases ok")


def comm_cases():
    import test_gpu_comm as TC
    TC._two_rank_protocol()


if __name__ == "__main__":
    if want("decode"):
        decode_cases()
    if want("step"):
        step_cases()
    if want("layer"):
        layer_cases()
    if want("comm"):
        comm_cases()
    torch.cuda.synchronize()
    print("sanitize cases done")
