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
