"""Secondary legs of bench.py (and scripts/bench_model.py): measurements that ride along in the one JSON line.

  decode_tok_s_leg   the model-level half of BASELINE.json's metric: whole-model decode tok/s, Llama-3-8B geometry,
                     batch 64, ctx 4096, random-init weights, through b200attn.engine.LLMEngine (C++ scheduler,
                     CUDA-graph decode steps), with its weights + KV HBM roofline.
  comparator_leg     the same-box comparator BASELINE.md section 4 plans: the paged-decode kernels of the vLLM /
                     flash-attn / flashinfer wheels that happen to be in this image, timed on the same seeded problem.
                     Runs in a SUBPROCESS (scripts/comparator.py) with a time limit: those imports log to stdout, may
                     JIT-compile, and must never be able to break or pollute the bench line.  Absent wheel, import
                     failure or time-out -> "unavailable"; never an error.
Not product code: nothing under vllm-backend-public_b200/ imports this file.
"""
from __future__ import annotations

import json
import os
import random
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
for _p in (ROOT / "vllm-backend-public_b200",):
    if str(_p) not in sys.path:
        sys.path.insert(0, str(_p))

import torch  # noqa: E402


def _hbm_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    try:
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def _event_time(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def decode_tok_s_leg(dev, model: str = "llama-3-8b", B: int = 64, ctx: int = 4096, steps: int = 100, kv: str = "auto",
                     tp_group=None) -> dict:
    """Whole-model decode at batch B, context ctx: e2e = engine.step() (scheduler, H2D of the step rows, CUDA-graph
    replay of 32 layers + lm_head + sample, D2H of the sampled ids, sync); device = the graph replay alone.
    Roofline: (weights streamed per step + KV bytes) / measured HBM bandwidth."""
    try:
        from b200attn.engine import LLMEngine, SamplingParams
        from b200attn.model import resolve_config
        cfg = resolve_config(model)
        HBM, hbm_src = _hbm_peak()
        # sequences decode while later prompts are still being prefilled (decode-first schedule): give every request
        # enough new tokens to still be running when the timed steps start
        max_new = steps + 40 + B * ((ctx + 8191) // 8192 + 1)
        eng = LLMEngine(cfg, dev, max_num_seqs=B, max_model_len=ctx + max_new + 64, max_num_batched_tokens=8192,
                        kv_cache_dtype=kv)
        if ctx + max_new > eng.max_model_len:
            return {"unavailable": f"ctx {ctx} + {max_new} new tokens exceed max_model_len {eng.max_model_len} of {cfg.name}"}
        kv_elem = 1 if kv == "fp8" else 2
        rng = random.Random(0)
        t0 = time.perf_counter()
        for _ in range(B):
            eng.add_request([rng.randrange(cfg.vocab) for _ in range(ctx - 1)], SamplingParams(max_tokens=max_new, temperature=0.0))
        while eng.sched.num_waiting or eng.stats.decode_steps < 5:           # prefill + 5 warm-up decode steps
            eng.step()
        t_prefill = time.perf_counter() - t0
        torch.cuda.synchronize()
        n0 = eng.stats.generated_tokens
        t0 = time.perf_counter()
        for _ in range(steps):
            eng.step()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        toks = eng.stats.generated_tokens - n0
        replay = next((g for (b, _), g in eng._graphs.items() if b == B), None) or list(eng._graphs.values())[-1]
        with torch.cuda.stream(eng._stream):
            med, best = _event_time(replay, iters=20)
        avg_ctx = ctx + max(eng.stats.generated_tokens // max(B, 1) - steps // 2, 0)   # sequences grew while others prefilled
        kv_bytes = B * avg_ctx * cfg.kv_bytes_per_token(kv_elem)
        w_bytes = eng.model.weight_bytes() - cfg.vocab * cfg.hidden * 2        # the embedding table is gathered, not streamed
        ceiling = B / ((kv_bytes + w_bytes) / (HBM * 1e9))
        rec = dict(model=cfg.name, kv_cache=kv, B=B, ctx=ctx, avg_ctx_during_timing=avg_ctx, running=int(eng.sched.num_running),
                   steps=steps, value=round(toks / wall, 1), unit="tok/s", e2e_tok_s=round(toks / wall, 1),
                   e2e_ms_per_step=round(wall / steps * 1e3, 4), device_ms_per_step=round(med, 4),
                   device_tok_s=round(B / med * 1e3, 1), hbm_bytes_per_step=int(kv_bytes + w_bytes),
                   roofline_tok_s_at_measured_hbm=round(ceiling, 1), peak_source=hbm_src,
                   frac_of_roofline=round((B / med * 1e3) / ceiling, 4),
                   frac_of_roofline_e2e=round((toks / wall) / ceiling, 4),
                   achieved_GBps=round((kv_bytes + w_bytes) / med / 1e6, 1),
                   target_tok_s_at_70pct_of_8TBps=7108, prefill_s=round(t_prefill, 3),
                   prefill_tok_s=round(B * (ctx - 1) / t_prefill, 1), weights="random-init", data="synthetic token ids",
                   note="e2e = engine.step() incl. scheduler, H2D of the step rows, graph replay, D2H of the sampled ids, "
                        "sync; device = CUDA-graph replay alone (32 layers + lm_head + sample); GEMMs are cuBLAS")
        del eng
        torch.cuda.empty_cache()
        return rec
    except Exception as e:                                                    # noqa: BLE001
        torch.cuda.empty_cache()
        return {"error": f"{type(e).__name__}: {e}"[:400]}


def comparator_leg(timeout_s: int = 420) -> dict:
    """vLLM paged_attention_v2 / flash_attn_with_kvcache / flashinfer decode on BASELINE configs 2 and 3, same box.
    See scripts/comparator.py; this function only runs it under a time limit and reads its JSON."""
    script = ROOT / "scripts" / "comparator.py"
    if os.environ.get("B2_BENCH_COMPARATOR", "1") != "1":
        return {"unavailable": "switched off (B2_BENCH_COMPARATOR=0)"}
    with tempfile.TemporaryDirectory() as td:
        out = Path(td) / "comparator.json"
        try:
            r = subprocess.run([sys.executable, str(script), "--out", str(out)], stdout=subprocess.DEVNULL,
                               stderr=subprocess.PIPE, timeout=timeout_s, text=True)
        except subprocess.TimeoutExpired:
            if out.exists():            # the script rewrites its output after every comparator: keep what finished
                try:
                    rec = json.loads(out.read_text())
                    rec["truncated"] = f"subprocess exceeded {timeout_s} s; later comparators did not run"
                    return rec
                except Exception:
                    pass
            return {"unavailable": f"comparator subprocess exceeded {timeout_s} s"}
        except Exception as e:                                                # noqa: BLE001
            return {"unavailable": f"{type(e).__name__}: {e}"[:300]}
        if out.exists():
            try:
                return json.loads(out.read_text())
            except Exception as e:                                            # noqa: BLE001
                return {"unavailable": f"unreadable comparator output: {e}"[:300]}
        return {"unavailable": f"comparator exited {r.returncode}: {r.stderr[-300:]}"}


# ------------------------------------------------------------------------------------------------------------------
# tensor-parallel serving (N > 1): every rank calls these; rank 0 drives the engine, the others mirror it
# ------------------------------------------------------------------------------------------------------------------
def tp_parity_leg(dev, group, model: str = "tiny8", seed: int = 1, gap: float = 0.06) -> dict:
    """The served engine at --tensor-parallel-size N against the dense CPU oracle (teacher forcing, as
    tests/test_gpu_engine.py): every generated token must be within `gap` logits of the oracle's maximum at its
    position.  Rank 0 also serves the same prompts at tp = 1 and reports how many tokens agree exactly."""
    from b200attn.engine import LLMEngine, SamplingParams
    from b200attn.model import REGISTRY, LlamaModel
    cfg = REGISTRY[model]
    kw = dict(max_num_seqs=8, max_model_len=1024, block_size=16, max_num_batched_tokens=256, num_blocks=600, seed=seed)
    rng = random.Random(0)
    prompts = [[rng.randrange(cfg.vocab) for _ in range(n)] for n in (1, 5, 16, 17, 100, 300, 520, 33, 64, 257, 2, 129)]
    rec = None
    try:
        eng = LLMEngine(cfg, dev, tp=group, **kw)
        if group.rank == 0:
            reqs = [eng.add_request(p, SamplingParams(max_tokens=24, temperature=0.0)) for p in prompts]
            eng.run_until_idle()
            eng.shutdown_workers()
            stats = eng.stats
        else:
            eng.worker_loop()
        eng.close()
        del eng
        torch.cuda.empty_cache()
        if group.rank == 0:
            if str(ROOT / "oracle") not in sys.path:
                sys.path.insert(0, str(ROOT / "oracle"))
            import layer_ref as L                                       # the checker (oracle/), never the product path
            full = LlamaModel(cfg, dev, num_blocks=1, seed=seed)        # same seed -> the unsharded weights
            w = {k: v.cpu() for k, v in full.w.items()}
            table = full.cos_sin.cpu()
            del full
            worst, n_tok = 0.0, 0
            for r in reqs:
                ids = r.prompt_ids + r.output_ids
                logits = L.llama_forward_ref(w, ids[:-1], layers=cfg.layers, Hq=cfg.Hq, Hkv=cfg.Hkv, d=cfg.d, eps=cfg.eps,
                                             cos_sin=table)
                rows = logits[len(r.prompt_ids) - 1:]
                picked = rows.gather(1, torch.tensor(r.output_ids).unsqueeze(1)).squeeze(1)
                worst = max(worst, float((rows.max(dim=1).values - picked).max()))
                n_tok += len(r.output_ids)
            e1 = LLMEngine(cfg, dev, **kw)
            out1 = e1.generate(prompts, SamplingParams(max_tokens=24, temperature=0.0))
            del e1
            torch.cuda.empty_cache()
            same = sum(int(a == b) for r, o in zip(reqs, out1) for a, b in zip(r.output_ids, o))
            rec = {"model": model, "tp": group.size, "requests": len(reqs), "tokens_checked": n_tok,
                   "all_finished": all(r.finished and len(r.output_ids) == 24 for r in reqs),
                   "max_logit_gap_vs_oracle": round(worst, 5), "gap_allowed": gap, "ok": bool(worst <= gap),
                   "tokens_equal_to_tp1_engine": same, "decode_steps": stats.decode_steps, "mixed_steps": stats.mixed_steps,
                   "check": "teacher forcing against oracle/layer_ref.py::llama_forward_ref on the unsharded weights; "
                            "prompts 1..520 tokens, chunked prefill (NCCL all-reduce for > 256-row steps), CUDA-graph "
                            "decode with the fused NVLink all-reduce + RMSNorm"}
    except Exception as e:                                                    # noqa: BLE001
        rec = {"error": f"{type(e).__name__}: {e}"[:400]}
    group.barrier()
    return rec


_NVLINK_COUNTERS = [True]        # flips to False the first time NVML says the counters are not there (this pool: N/A)


def _nvlink_kib(index: int):
    """(tx, rx) data KiB over all NVLink links of GPU `index` (NVML field counters), or None."""
    if not _NVLINK_COUNTERS[0]:
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        v = pynvml.nvmlDeviceGetFieldValues(h, [(pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, 0xFFFFFFFF),
                                                (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, 0xFFFFFFFF)])
        if any(x.nvmlReturn != 0 for x in v):
            _NVLINK_COUNTERS[0] = False
            return None
        return int(v[0].value.ullVal), int(v[1].value.ullVal)
    except Exception:                                                         # noqa: BLE001
        _NVLINK_COUNTERS[0] = False
        return None


def tp_comm_leg(dev, group, T: int = 64, H: int = 4096, iters: int = 200) -> dict:
    """The exchange after o_proj / down_proj at decode size: the fused one-shot NVLink kernel against NCCL all-reduce +
    the plain fused-add RMSNorm, back-to-back launches, CUDA events on rank 0 (every rank runs the same loop)."""
    from b200attn import ops
    try:
        group.setup_comm(256, H, 0, 0)
        w = torch.ones(H, dtype=torch.bfloat16, device=dev)
        resid = torch.zeros(T, H, dtype=torch.bfloat16, device=dev)
        out = torch.empty(T, H, dtype=torch.bfloat16, device=dev)
        for s in (0, 1):
            group.inputs[s][:T].normal_()
        x = torch.randn(T, H, device=dev).to(torch.bfloat16)

        def fused():
            for i in range(iters):
                group.allreduce_rms_norm(i & 1, T, w, resid, 1e-5, out=out)

        def push():
            for i in range(iters):
                group.push_allreduce_rms_norm(x, w, resid, 1e-5, out=out)

        def nccl():
            for _ in range(iters):
                group.all_reduce_nccl(x)
                ops.rms_norm(x, w, 1e-5, residual=resid, out=out)

        res = {}
        gpu_index = int(os.environ.get("CUDA_VISIBLE_DEVICES", "0,1,2,3,4,5,6,7").split(",")[dev.index or 0])
        for name, fn in (("push_allreduce_rmsnorm_us", push), ("pull_allreduce_rmsnorm_us", fused),
                         ("nccl_allreduce_plus_rmsnorm_us", nccl)):
            fn()
            group.barrier()
            k0 = _nvlink_kib(gpu_index)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            res[name] = round(e0.elapsed_time(e1) / iters * 1e3, 2)
            group.barrier()
            k1 = _nvlink_kib(gpu_index)
            if k0 is not None and k1 is not None:                 # NVML link counters of this GPU around the loop
                res[name.replace("_us", "_nvlink_bytes_per_call")] = {"tx": round((k1[0] - k0[0]) * 1024 / iters),
                                                                      "rx": round((k1[1] - k0[1]) * 1024 / iters)}
            x.normal_()
        # the same loops replayed as ONE CUDA graph: device time per exchange without the host launch cost
        # (how the decode step runs them)
        st = torch.cuda.Stream(dev)
        for name, fn in (("push_in_graph_us", push), ("pull_in_graph_us", fused), ("nccl_in_graph_us", nccl)):
            try:
                group.barrier()
                with torch.cuda.stream(st):
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=st):
                        fn()
                    g.replay()
                    st.synchronize()
                    group.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(st)
                    g.replay()
                    e1.record(st)
                    st.synchronize()
                res[name] = round(e0.elapsed_time(e1) / iters * 1e3, 2)
                del g
            except Exception as e:                                            # noqa: BLE001
                res[name] = f"{type(e).__name__}: {e}"[:200]
        if os.environ.get("B2_COMM_DEBUG") == "1":
            import ctypes
            torch.cuda.synchronize()
            st4 = (ctypes.c_uint64 * 4)()
            group._lib.b2_comm_debug_stamps(group._comm, st4)
            res["push_stamps_ns"] = {"pushed": st4[1] - st4[0], "all_arrived": st4[2] - st4[0], "row_done": st4[3] - st4[0]}
        group.barrier()
        group.close()
        res.update(T=T, H=H, tp=group.size, message_bytes=T * H * 2, iters=iters,
                   note="*_us: per call, back-to-back EAGER launches from Python on every rank -- host launch pacing and jitter of the "
                        "slowest rank included (every exchange waits for all ranks' launches), so they vary run to run; "
                        "*_in_graph_us: the same loop replayed as one CUDA graph = the device cost, and how the decode "
                        "step runs it; the fused kernel reads tp x message bytes over NVLink per rank")
        return res
    except Exception as e:                                                    # noqa: BLE001
        return {"error": f"{type(e).__name__}: {e}"[:400]}


def tp_decode_leg(dev, group, model: str = "llama-3-8b", B: int = 64, ctx: int = 4096, steps: int = 50) -> dict:
    """Whole-model decode tok/s served at --tensor-parallel-size N (strong scaling of decode_tok_s_leg's workload)."""
    try:
        from b200attn.engine import LLMEngine, SamplingParams
        from b200attn.model import resolve_config
        cfg = resolve_config(model)
        HBM, hbm_src = _hbm_peak()
        max_new = steps + 40 + B * ((ctx + 8191) // 8192 + 1)
        eng = LLMEngine(cfg, dev, max_num_seqs=B, max_model_len=ctx + max_new + 64, max_num_batched_tokens=8192, tp=group)
        rec = None
        if group.rank != 0:
            eng.worker_loop()
        else:
            rng = random.Random(0)
            t0 = time.perf_counter()
            for _ in range(B):
                eng.add_request([rng.randrange(cfg.vocab) for _ in range(ctx - 1)],
                                SamplingParams(max_tokens=max_new, temperature=0.0))
            while eng.sched.num_waiting or eng.stats.decode_steps < 5:
                eng.step()
            t_prefill = time.perf_counter() - t0
            torch.cuda.synchronize()
            eng.time_device = True
            n0 = eng.stats.generated_tokens
            t0 = time.perf_counter()
            for _ in range(steps):
                eng.step()
            torch.cuda.synchronize()
            wall = time.perf_counter() - t0
            toks = eng.stats.generated_tokens - n0
            dev_ms = eng.stats.device_ms / max(eng.stats.device_steps, 1)
            avg_ctx = ctx + max(eng.stats.generated_tokens // max(B, 1) - steps // 2, 0)
            kv_bytes = B * avg_ctx * cfg.kv_bytes_per_token(2, group.size)
            w_bytes = eng.model.weight_bytes() - cfg.vocab * cfg.hidden * 2
            ceiling = B / ((kv_bytes + w_bytes) / (HBM * 1e9))
            rec = dict(model=cfg.name, tp=group.size, B=B, ctx=ctx, avg_ctx_during_timing=avg_ctx, steps=steps,
                       running=int(eng.sched.num_running), value=round(toks / wall, 1), unit="tok/s",
                       e2e_ms_per_step=round(wall / steps * 1e3, 4), device_ms_per_step=round(dev_ms, 4),
                       device_tok_s=round(B / dev_ms * 1e3, 1), hbm_bytes_per_step_per_gpu=int(kv_bytes + w_bytes),
                       roofline_tok_s_at_measured_hbm=round(ceiling, 1), peak_source=hbm_src,
                       frac_of_roofline=round((B / dev_ms * 1e3) / ceiling, 4),
                       exchanges_per_step=2 * cfg.layers + (1 if eng.model.vocab_parallel else 0),
                       prefill_s=round(t_prefill, 3), prefill_tok_s=round(B * (ctx - 1) / t_prefill, 1),
                       weights="random-init", data="synthetic token ids",
                       note="rank 0: engine.step() incl. the step message to the other ranks, scheduler, H2D, graph replay "
                            "(per layer: 2 fused NVLink all-reduce + RMSNorm kernels; vocab-parallel lm_head + peer-store "
                            "all-gather), D2H of the sampled ids; device = CUDA events around the replay on rank 0")
            eng.shutdown_workers()
        eng.close()
        del eng
        torch.cuda.empty_cache()
        group.barrier()
        return rec
    except Exception as e:                                                    # noqa: BLE001
        torch.cuda.empty_cache()
        return {"error": f"{type(e).__name__}: {e}"[:400]}
