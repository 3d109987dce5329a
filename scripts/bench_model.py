"""Measurements for the rows next to the hot path, one JSON line each:
  prefill attention TFLOP/s (Llama-3-8B heads, 8k-token chunks at several prefix lengths), element-wise ops GB/s,
  whole-model decode tok/s at Llama-3-8B b64 ctx4k (BASELINE.json's second headline figure; roofline = weights + KV bytes).
    python scripts/bench_model.py [--skip-model] > gpurun_out/bench_model.jsonl
CUDA events on the launching stream, >= 3 warm-ups, inputs larger than L2 or L2 flushed (said per line)."""
import argparse
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT / "vllm-backend-public_b200")]
import torch  # noqa: E402
from b200attn import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--skip-model", action="store_true")
ap.add_argument("--model", default="llama-3-8b")
ap.add_argument("--steps", type=int, default=100)
ap.add_argument("--only", choices=["all", "prefill", "ops", "decode"], default="all")
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--ctx", type=int, default=4096)
ap.add_argument("--kv", choices=["auto", "fp8"], default="auto", help="KV cache dtype of the whole-model decode run")
a = ap.parse_args()
dev = torch.device("cuda:0")
peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
HBM = float(peaks.get("hbm_gbs", 6650.0))
TF = float(peaks.get("bf16_tflops", 1650.0))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=20, warm=3, flush_l2=False, reps=1):
    """ms per call.  reps > 1: that many back-to-back launches inside one event pair (streaming kernels of tens of
    microseconds whose working set exceeds L2: a lone launch between two events also times the launch latency --
    ncu shows rms_norm at 38 us where the lone-launch timing reads 59)."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush_l2:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / reps)
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def emit(**kw):
    print(json.dumps(kw), flush=True)


# ---------------------------------------------------------------- prefill attention
Hq, Hkv, d, bs = 32, 8, 128, 16
g = torch.Generator(device=dev).manual_seed(0)
for chunk, start in ((8192, 0), (8192, 8192), (8192, 24576), (2048, 0), (512, 3584)) if a.only in ("all", "prefill") else ():
    total = start + chunk
    nb = (total + bs - 1) // bs
    num_blocks = nb + 8
    kc = torch.randn(num_blocks, Hkv, bs, d, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    vc = torch.randn(num_blocks, Hkv, bs, d, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    q = torch.randn(chunk, Hq, d, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    bt = torch.randperm(num_blocks, device=dev)[:nb].to(torch.int32).unsqueeze(0).contiguous()
    cu = torch.tensor([0, chunk], dtype=torch.int32, device=dev)
    st = torch.tensor([start], dtype=torch.int32, device=dev)
    out = torch.empty_like(q)
    med, best = timeit(lambda: ops.prefill_attention(q, kc, vc, bt, cu, st, out=out), iters=10)
    pairs = chunk * start + chunk * (chunk + 1) // 2
    flops = 4 * d * Hq * pairs
    emit(op="prefill_attention", chunk=chunk, prefix=start, Hq=Hq, Hkv=Hkv, ms=med, ms_min=best,
         tflops=flops / med / 1e9, frac_of_measured_bf16=flops / med / 1e9 / TF, peak_tflops=TF,
         note="causal flops only (4*d per visible (q,k) pair)")

# ---------------------------------------------------------------- element-wise ops
for T in (64, 8192) if a.only in ("all", "ops") else ():
    H, I = 4096, 14336
    x = torch.randn(T, H, device=dev).to(torch.bfloat16)
    r = torch.randn(T, H, device=dev).to(torch.bfloat16)
    w = torch.ones(H, device=dev, dtype=torch.bfloat16)
    o = torch.empty_like(x)
    small = T == 64
    reps = 1 if small else 20
    med, best = timeit(lambda: ops.rms_norm(x, w, 1e-5, residual=r, out=o), flush_l2=small, reps=reps)
    emit(op="rms_norm+residual", T=T, H=H, us=med * 1e3, GBps=4 * T * H * 2 / med / 1e6, frac_hbm=4 * T * H * 2 / med / 1e6 / HBM,
         l2="flushed" if small else "working set > L2" if 4 * T * H * 2 > 126e6 else "fits L2 (reported as is)")
    gu = torch.randn(T, 2 * I, device=dev).to(torch.bfloat16)
    oo = torch.empty(T, I, device=dev, dtype=torch.bfloat16)
    med, best = timeit(lambda: ops.silu_and_mul(gu, out=oo), flush_l2=small, reps=reps)
    emit(op="silu_and_mul", T=T, I=I, us=med * 1e3, GBps=3 * T * I * 2 / med / 1e6, frac_hbm=3 * T * I * 2 / med / 1e6 / HBM,
         l2="flushed" if small else "working set > L2")
    qkv = torch.randn(T, (Hq + 2 * Hkv) * d, device=dev).to(torch.bfloat16)
    nblk = (T + bs - 1) // bs + 4
    kc = torch.zeros(nblk, Hkv, bs, d, device=dev, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    pos = torch.arange(T, device=dev, dtype=torch.int32)
    slots = torch.randperm(nblk * bs, device=dev)[:T].to(torch.int64)
    table = torch.randn(max(T, 64), d, device=dev)
    med, best = timeit(lambda: ops.rope_and_cache(qkv, pos, table, kc, vc, slots, Hq), flush_l2=small, reps=reps)
    byt = T * ((Hq + Hkv) * d * 2 * 2 + Hkv * d * 2 + 2 * Hkv * d * 2 + d * 4)
    emit(op="rope_and_cache", T=T, us=med * 1e3, GBps=byt / med / 1e6, frac_hbm=byt / med / 1e6 / HBM,
         l2="flushed" if small else "working set > L2" if byt > 126e6 else "fits L2 (reported as is)")
logits = torch.randn(64, 128256, device=dev).to(torch.bfloat16)
med, best = timeit(lambda: ops.sample(logits), flush_l2=True)
emit(op="sample(greedy)", B=64, V=128256, us=med * 1e3, GBps=64 * 128256 * 2 / med / 1e6, frac_hbm=64 * 128256 * 2 / med / 1e6 / HBM,
     l2="flushed")

if a.skip_model:
    sys.exit(0)
# ---------------------------------------------------------------- whole-model decode tok/s, b64 ctx4k
if a.only not in ("all", "decode"):
    sys.exit(0)
kc = vc = q = out = gu = oo = logits = None
torch.cuda.empty_cache()
from b200attn.engine import LLMEngine, SamplingParams  # noqa: E402
from b200attn.model import resolve_config  # noqa: E402
cfg = resolve_config(a.model)
B, ctx = a.batch, a.ctx
# sequences decode while later prompts are still being prefilled (decode-first schedule): give every request enough
# new tokens to still be running when the timed steps start
max_new = a.steps + 40 + B * ((ctx + 8191) // 8192 + 1)
eng = LLMEngine(cfg, dev, max_num_seqs=B, max_model_len=ctx + max_new + 64, max_num_batched_tokens=8192,
                kv_cache_dtype=a.kv)
kv_elem = 1 if a.kv == "fp8" else 2
import random  # noqa: E402
rng = random.Random(0)
t0 = time.perf_counter()
for i in range(B):
    eng.add_request([rng.randrange(cfg.vocab) for _ in range(ctx - 1)], SamplingParams(max_tokens=max_new, temperature=0.0))
if ctx + max_new > eng.max_model_len:
    sys.exit(f"ctx {ctx} + {max_new} new tokens exceed max_model_len {eng.max_model_len} of {cfg.name} "
             "(use --model llama-3.1-8b for contexts past 8k)")
while eng.sched.num_waiting or eng.stats.decode_steps < 5:           # prefill + 5 warm-up decode steps
    eng.step()
t_prefill = time.perf_counter() - t0
torch.cuda.synchronize()
n0 = eng.stats.generated_tokens
t0 = time.perf_counter()
for _ in range(a.steps):
    eng.step()
torch.cuda.synchronize()
wall = time.perf_counter() - t0
toks = eng.stats.generated_tokens - n0
# device time of the graph replay alone (inputs resident)
replay = next((g for (b, _), g in eng._graphs.items() if b == B), None) or list(eng._graphs.values())[-1]
with torch.cuda.stream(eng._stream):
    med, best = timeit(replay, iters=20)
n_run = eng.sched.num_running if hasattr(eng.sched, "num_running") else B
avg_ctx = ctx + max(eng.stats.generated_tokens // max(B, 1) - a.steps // 2, 0)   # sequences grew while others prefilled
kv_bytes = B * avg_ctx * cfg.kv_bytes_per_token(kv_elem)
w_bytes = eng.model.weight_bytes() - cfg.vocab * cfg.hidden * 2        # the embedding table is gathered, not streamed
ceiling = B / ((kv_bytes + w_bytes) / (HBM * 1e9))
emit(op="decode_tok_s", model=cfg.name, kv_cache=a.kv, B=B, ctx=ctx, avg_ctx_during_timing=avg_ctx, running=n_run, steps=a.steps, e2e_tok_s=toks / wall, e2e_ms_per_step=wall / a.steps * 1e3,
     device_ms_per_step=med, device_tok_s=B / med * 1e3, hbm_bytes_per_step=kv_bytes + w_bytes,
     roofline_tok_s_at_measured_hbm=ceiling, frac_of_roofline=(B / med * 1e3) / ceiling,
     achieved_GBps=(kv_bytes + w_bytes) / med / 1e6, prefill_s=t_prefill, prefill_tok_s=B * (ctx - 1) / t_prefill,
     note="e2e = engine.step() incl. scheduler, H2D of the step rows, graph replay, D2H of the sampled ids, sync; "
          "device = CUDA-graph replay alone (32 layers + lm_head + sample)")
