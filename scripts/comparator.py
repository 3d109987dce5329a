"""Same-box comparator: the paged-decode kernels of the wheels that happen to be in this image,
timed on the same seeded problems as bench.py, beside this repo's kernel.  Not the reference's pinned vLLM v0.9.1 image
(reference benchmark-vllm/vllm_manager.py:371-374) -- the closest thing to "the kernel the launched engine would run"
that can execute here.  Run by bench_legs.comparator_leg() in a subprocess; writes one JSON file, rewritten after every
comparator so a time-out keeps what finished.

    python scripts/comparator.py --out gpurun_out/comparator.json [--launches 100]

  vllm         torch.ops._C.paged_attention_v2 (partition 512), its own cache layout, filled by its reshape_and_cache
  flash_attn   flash_attn_with_kvcache over a paged cache (FA2 needs page_block_size % 256 == 0, so pages of 256 tokens)
  flashinfer   BatchDecodeWithPagedKVCacheWrapper, HND layout, page 16 -- reads this repo's cache tensors as they are
Every comparator: CUDA events around `launches` back-to-back calls after 10 warm-ups (working set >> L2), the result
compared with this repo's output on the same inputs (max abs difference, reported, not asserted).
"""
import argparse
import json
import sys
import time
import traceback
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "vllm-backend-public_b200")]
import torch  # noqa: E402

import bench  # noqa: E402  (problem generator and byte counts only)

D, BS = bench.D, bench.BS


def timed(fn, launches):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(launches):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / launches


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--launches", type=int, default=100)
    ap.add_argument("--workloads", nargs="*", default=["llama3-8b_b64_ctx4k_bf16kv", "llama3-8b_b128_ctx8k_fp8kv"])
    a = ap.parse_args()
    out_path = Path(a.out)
    dev = torch.device("cuda:0")
    import b200attn
    report = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__, "launches": a.launches,
              "timing": "CUDA events around back-to-back launches after 10 warm-ups; KV working set >> L2", "results": []}

    def flush():
        out_path.write_text(json.dumps(report))

    for wl_name in a.workloads:
        wl = bench.WORKLOADS[wl_name]
        B, Hq, Hkv, ctx, fp8 = wl["B"], wl["Hq"], wl["Hkv"], wl["ctx"], wl["fp8"]
        nbytes = bench.algorithmic_bytes(B, Hq, Hkv, ctx, 1 if fp8 else 2)
        prob = bench.build_problem(dev, B, Hq, Hkv, ctx, fp8, seed=0)
        ks, vs = prob["scales"] or (None, None)
        bt, sl = prob["block_tables"].to(dev), prob["seq_lens"].to(dev)
        scale = D ** -0.5
        ns, _ = bench.pick_schedule(prob, -1)
        ws = torch.empty(max(b200attn.decode_workspace_bytes(B, Hq, Hkv, D, BS, prob["max_nb"], ns), 16), dtype=torch.uint8, device=dev)
        ours = torch.empty(B, Hq, D, dtype=torch.bfloat16, device=dev)
        run_ours = lambda: b200attn.paged_decode(prob["q"], prob["k_cache"], prob["v_cache"], bt, sl, k_scale=ks,  # noqa: E731
                                                 v_scale=vs, out=ours, workspace=ws, num_splits=ns)
        ms = timed(run_ours, a.launches)
        report["results"].append({"name": "b200attn (this repo)", "workload": wl_name, "kernel_ms": round(ms, 5),
                                  "GBps": round(nbytes / ms / 1e6, 1), "schedule": bench.schedule_name(ns)})
        flush()
        ours_f = ours.float()

        def record(name, version, fn, result, note=""):
            try:
                ms = timed(fn, a.launches)
                diff = float((result().float() - ours_f).abs().max())
                report["results"].append({"name": name, "version": version, "workload": wl_name, "kernel_ms": round(ms, 5),
                                          "GBps": round(nbytes / ms / 1e6, 1), "max_abs_diff_vs_ours": diff,
                                          "ours_over_theirs_time": None, "note": note})
            except Exception as e:                                         # noqa: BLE001
                report["results"].append({"name": name, "version": version, "workload": wl_name,
                                          "unavailable": f"{type(e).__name__}: {e}"[:300], "tb": traceback.format_exc()[-600:]})
            flush()

        # ------------------------------------------------------------------ vLLM paged_attention_v2 / v1
        try:
            import vllm
            from vllm import _custom_ops as vops
            assert hasattr(torch.ops._C, "paged_attention_v2"), "torch.ops._C.paged_attention_v2 not registered"
            nb = prob["num_blocks"]
            x = 16 if fp8 else 8
            cdt = torch.uint8 if fp8 else torch.bfloat16
            kc = torch.zeros(nb, Hkv, D // x, BS, x, dtype=cdt, device=dev)
            vc = torch.zeros(nb, Hkv, D, BS, dtype=cdt, device=dev)
            kvd = "fp8" if fp8 else "auto"
            ksd = ks if ks is not None else torch.ones(1, dtype=torch.float32, device=dev)
            vsd = vs if vs is not None else torch.ones(1, dtype=torch.float32, device=dev)
            pos = torch.arange(ctx, device=dev)
            for b in range(B):
                slots = bt[b, pos // BS].to(torch.int64) * BS + (pos % BS)
                k, v = bench.gen_sequence_kv(dev, 0, b, ctx, Hkv)
                vops.reshape_and_cache(k, v, kc, vc, slots, kvd, ksd, vsd)
            vops.reshape_and_cache(prob["k_new"], prob["v_new"], kc, vc, prob["slot_new"].to(dev), kvd, ksd, vsd)
            part = 512
            nparts = (ctx + part - 1) // part
            o2 = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device=dev)
            exp_sums = torch.empty(B, Hq, nparts, dtype=torch.float32, device=dev)
            max_logits = torch.empty_like(exp_sums)
            tmp = torch.empty(B, Hq, nparts, D, dtype=torch.bfloat16, device=dev)
            record("vllm paged_attention_v2", vllm.__version__,
                   lambda: vops.paged_attention_v2(o2, exp_sums, max_logits, tmp, prob["q"], kc, vc, Hkv, scale, bt, sl, BS, ctx,
                                                   None, kvd, ksd, vsd),
                   lambda: o2, "partition 512; vLLM cache layout [nb, Hkv, d/x, bs, x] / [nb, Hkv, d, bs]")
            del kc, vc, tmp
        except Exception as e:                                             # noqa: BLE001
            report["results"].append({"name": "vllm paged_attention_v2", "workload": wl_name,
                                      "unavailable": f"{type(e).__name__}: {e}"[:300]})
            flush()
        torch.cuda.empty_cache()

        # ------------------------------------------------------------------ flashinfer batch decode (HND, page 16)
        try:
            import flashinfer
            fws = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
            w = flashinfer.BatchDecodeWithPagedKVCacheWrapper(fws, kv_layout="HND", use_tensor_cores=True)
            nbs = (prob["seq_lens"].long() + BS - 1) // BS
            indptr = torch.zeros(B + 1, dtype=torch.int32)
            indptr[1:] = torch.cumsum(nbs, 0)
            indices = torch.cat([prob["block_tables"][b, :int(nbs[b])] for b in range(B)]).to(torch.int32)
            last = ((prob["seq_lens"] - 1) % BS + 1).to(torch.int32)
            kvt = torch.float8_e4m3fn if fp8 else torch.bfloat16
            w.plan(indptr.to(dev), indices.to(dev), last.to(dev), Hq, Hkv, D, BS, q_data_type=torch.bfloat16,
                   kv_data_type=kvt)
            kcf = prob["k_cache"].view(kvt) if fp8 else prob["k_cache"]
            vcf = prob["v_cache"].view(kvt) if fp8 else prob["v_cache"]
            kw = dict(k_scale=6.0 / 448.0, v_scale=6.0 / 448.0) if fp8 else {}
            fo = {}
            def fi():                                                       # noqa: E306
                fo["o"] = w.run(prob["q"], (kcf, vcf), **kw)
            record("flashinfer BatchDecodeWithPagedKVCacheWrapper", getattr(flashinfer, "__version__", "?"), fi,
                   lambda: fo["o"], "HND layout, page 16: reads this repo's cache tensors as they are; use_tensor_cores=True")
            del fws
        except Exception as e:                                             # noqa: BLE001
            report["results"].append({"name": "flashinfer BatchDecodeWithPagedKVCacheWrapper", "workload": wl_name,
                                      "unavailable": f"{type(e).__name__}: {e}"[:300]})
            flush()
        torch.cuda.empty_cache()

        # ------------------------------------------------------------------ flash_attn_with_kvcache (bf16 only, page 256)
        if not fp8:
            try:
                import flash_attn
                from flash_attn import flash_attn_with_kvcache
                page = 256
                npg = ctx // page
                g = torch.Generator().manual_seed(5)
                perm = torch.randperm(B * npg + 8, generator=g).to(torch.int32)
                fbt = perm[: B * npg].reshape(B, npg).contiguous().to(dev)
                kf = torch.zeros(B * npg + 8, page, Hkv, D, dtype=torch.bfloat16, device=dev)
                vf = torch.zeros_like(kf)
                for b in range(B):
                    k, v = bench.gen_sequence_kv(dev, 0, b, ctx, Hkv)
                    k[ctx - 1], v[ctx - 1] = prob["k_new"][b], prob["v_new"][b]
                    kf[fbt[b].long()] = k.view(npg, page, Hkv, D)
                    vf[fbt[b].long()] = v.view(npg, page, Hkv, D)
                q4 = prob["q"].unsqueeze(1).contiguous()
                fo2 = {}
                def fa():                                                   # noqa: E306
                    fo2["o"] = flash_attn_with_kvcache(q4, kf, vf, cache_seqlens=sl, block_table=fbt, softmax_scale=scale, causal=True)
                record("flash_attn_with_kvcache", flash_attn.__version__, fa, lambda: fo2["o"].squeeze(1),
                       "FA2 paged KV needs page_block_size % 256 == 0: pages of 256 tokens, NHD layout")
                del kf, vf
            except Exception as e:                                         # noqa: BLE001
                report["results"].append({"name": "flash_attn_with_kvcache", "workload": wl_name,
                                          "unavailable": f"{type(e).__name__}: {e}"[:300]})
                flush()
        del prob
        torch.cuda.empty_cache()

    # ratios against this repo's kernel on the same workload
    ours_ms = {r["workload"]: r["kernel_ms"] for r in report["results"] if r["name"].startswith("b200attn")}
    for r in report["results"]:
        if "kernel_ms" in r and not r["name"].startswith("b200attn"):
            r["ours_over_theirs_time"] = round(ours_ms[r["workload"]] / r["kernel_ms"], 4)
    flush()


if __name__ == "__main__":
    t0 = time.time()
    main()
