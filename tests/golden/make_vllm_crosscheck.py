"""Pin the oracle against vLLM's own CUDA kernels (run on the GPU box; the wheel in the image is vllm 0.22).

The reference launches vLLM as an external server (`vllm/vllm-openai:v0.9.1`,
reference benchmark-vllm/vllm_manager.py:371-374) and holds no vectors for this path.  vLLM is the dependency whose
`reshape_and_cache` / `paged_attention_v1/v2` semantics the oracle restates, so its kernels, run here on the same
seeded inputs, are the closest thing to reference outputs.  This script writes
    gpurun_out/vllm_crosscheck.pt      {case: {"out_v1", "out_v2", "k_cache_sha", "v_cache_sha", ...}}
    gpurun_out/vllm_crosscheck.json    what ran, versions, or why it could not
which are then committed under tests/golden/ and checked by tests/test_vllm_crosscheck.py (oracle, CPU) and
tests/test_gpu_parity.py (our kernels).  Nothing here is imported by the product.

    python tests/golden/make_vllm_crosscheck.py
"""
import hashlib
import json
import sys
import traceback
from pathlib import Path

import torch

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT / "oracle"))
import paged_attention_ref as R  # noqa: E402

OUT = ROOT / "gpurun_out"
OUT.mkdir(exist_ok=True)

CASES = {
    # name: make_inputs kwargs, fp8
    "config1_b4_hq8_hkv2_ctx512": (dict(B=4, Hq=8, Hkv=2, d=128, bs=16, ctx=512, seed=0, ragged=False), False),
    "ragged_b16_hq32_hkv8_ctx700": (dict(B=16, Hq=32, Hkv=8, d=128, bs=16, ctx=700, seed=3, ragged=True), False),
    "ragged_b8_hq64_hkv8_ctx1100": (dict(B=8, Hq=64, Hkv=8, d=128, bs=16, ctx=1100, seed=5, ragged=True), False),
    "fp8_b8_hq32_hkv8_ctx600": (dict(B=8, Hq=32, Hkv=8, d=128, bs=16, ctx=600, seed=4, ragged=True), True),
    # r2 (VERDICT r1 "the vLLM pin is thin"): other block sizes, contexts long enough that v2 really partitions
    # (512-token partitions), G = 1 and G = 16, single-token sequences, a long fp8 case
    "bs8_b6_hq8_hkv2_ctx300": (dict(B=6, Hq=8, Hkv=2, d=128, bs=8, ctx=300, seed=6, ragged=True), False),
    "bs32_b6_hq32_hkv8_ctx900": (dict(B=6, Hq=32, Hkv=8, d=128, bs=32, ctx=900, seed=7, ragged=True), False),
    "long_b4_hq32_hkv8_ctx4500": (dict(B=4, Hq=32, Hkv=8, d=128, bs=16, ctx=4500, seed=8, ragged=False), False),
    "longragged_b6_hq32_hkv8_ctx5000": (dict(B=6, Hq=32, Hkv=8, d=128, bs=16, ctx=5000, seed=9, ragged=True), False),
    "g1_b5_hq4_hkv4_ctx400": (dict(B=5, Hq=4, Hkv=4, d=128, bs=16, ctx=400, seed=10, ragged=True), False),
    "g16_b5_hq16_hkv1_ctx600": (dict(B=5, Hq=16, Hkv=1, d=128, bs=16, ctx=600, seed=11, ragged=True), False),
    "len1_b3_hq8_hkv2": (dict(B=3, Hq=8, Hkv=2, d=128, bs=16, ctx=1, seed=12, ragged=False), False),
    "fp8_bs32_b4_hq32_hkv8_ctx4200": (dict(B=4, Hq=32, Hkv=8, d=128, bs=32, ctx=4200, seed=13, ragged=True), True),
    "fp8_bs8_g16_b4_hq16_hkv1_ctx700": (dict(B=4, Hq=16, Hkv=1, d=128, bs=8, ctx=700, seed=14, ragged=True), True),
}


def sha(t):
    return hashlib.sha256(t.contiguous().view(torch.uint8).cpu().numpy().tobytes()).hexdigest()


def main():
    report = {"torch": torch.__version__, "cases": {}}
    try:
        import vllm
        from vllm import _custom_ops as vops
        report["vllm"] = vllm.__version__
        assert hasattr(torch.ops._C, "paged_attention_v1"), "torch.ops._C.paged_attention_v1 not registered"
    except Exception as e:  # noqa: BLE001
        report["unavailable"] = f"{type(e).__name__}: {e}"
        (OUT / "vllm_crosscheck.json").write_text(json.dumps(report, indent=1))
        print(json.dumps(report, indent=1))
        return
    dev = torch.device("cuda:0")
    report["gpu"] = torch.cuda.get_device_name(0)
    results = {}
    for name, (kw, fp8) in CASES.items():
        try:
            inp = R.make_inputs(**kw)
            B, Hq, Hkv, d, bs = kw["B"], kw["Hq"], kw["Hkv"], kw["d"], kw["bs"]
            nb = inp["num_blocks"]
            x = 16 if fp8 else 8                      # vLLM packs 16 bytes of one token along the last K-cache axis
            cdt = torch.uint8 if fp8 else torch.bfloat16
            kc = torch.zeros(nb, Hkv, d // x, bs, x, dtype=cdt, device=dev)
            vc = torch.zeros(nb, Hkv, d, bs, dtype=cdt, device=dev)
            if fp8:
                ks = float(inp["k"].float().abs().max()) / 448.0
                vs = float(inp["v"].float().abs().max()) / 448.0
            else:
                ks = vs = 1.0
            ksd = torch.tensor([ks], dtype=torch.float32, device=dev)
            vsd = torch.tensor([vs], dtype=torch.float32, device=dev)
            kvd = "fp8" if fp8 else "auto"
            vops.reshape_and_cache(inp["k"].to(dev), inp["v"].to(dev), kc, vc, inp["slot_mapping"].to(dev), kvd, ksd, vsd)
            # back to [num_blocks, Hkv, bs, d] (the layout of this repo and of the oracle)
            k_ours = kc.permute(0, 1, 3, 2, 4).reshape(nb, Hkv, bs, d).contiguous()
            v_ours = vc.permute(0, 1, 3, 2).contiguous()
            q = inp["q"].to(dev)
            bt, sl = inp["block_tables"].to(dev), inp["seq_lens"].to(dev)
            max_len = int(inp["seq_lens"].max())
            scale = d ** -0.5
            out1 = torch.zeros(B, Hq, d, dtype=torch.bfloat16, device=dev)
            vops.paged_attention_v1(out1, q, kc, vc, Hkv, scale, bt, sl, bs, max_len, None, kvd, ksd, vsd)
            part = 512
            nparts = (max_len + part - 1) // part
            out2 = torch.zeros_like(out1)
            exp_sums = torch.empty(B, Hq, nparts, dtype=torch.float32, device=dev)
            max_logits = torch.empty_like(exp_sums)
            tmp = torch.empty(B, Hq, nparts, d, dtype=torch.bfloat16, device=dev)
            vops.paged_attention_v2(out2, exp_sums, max_logits, tmp, q, kc, vc, Hkv, scale, bt, sl, bs, max_len, None,
                                    kvd, ksd, vsd)
            torch.cuda.synchronize()
            results[name] = dict(out_v1=out1.cpu(), out_v2=out2.cpu(), k_cache_sha=sha(k_ours), v_cache_sha=sha(v_ours),
                                 k_scale=ks, v_scale=vs, make_inputs=kw, fp8=fp8)
            report["cases"][name] = dict(ok=True, k_cache_sha=sha(k_ours), v_cache_sha=sha(v_ours),
                                         out_v1_sha=sha(out1), out_v2_sha=sha(out2))
        except Exception as e:  # noqa: BLE001
            report["cases"][name] = dict(ok=False, error=f"{type(e).__name__}: {e}", tb=traceback.format_exc()[-1500:])
    if results:
        torch.save(results, OUT / "vllm_crosscheck.pt")
    (OUT / "vllm_crosscheck.json").write_text(json.dumps(report, indent=1))
    print(json.dumps(report, indent=1)[:4000])


if __name__ == "__main__":
    main()
