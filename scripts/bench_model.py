"""Measurements for the rows next to the hot path, one JSON line each:
  prefill attention TFLOP/s (Llama-3-8B heads, 8k-token chunks at several prefix lengths), element-wise ops GB/s,
  whole-model decode tok/s at Llama-3-8B b64 ctx4k (BASELINE.json's second headline figure; roofline = weights + KV bytes).
    python scripts/bench_model.py [--skip-model] > gpurun_out/bench_model.jsonl
CUDA events on the launching stream, >= 3 warm-ups, inputs larger than L2 or L2 flushed (said per line)."""
import argparse
import json
import os
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
ap.add_argument("--only", choices=["all", "prefill", "ops", "decode", "gemm"], default="all")
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--ctx", type=int, default=4096)
ap.add_argument("--prefill-bs", type=int, default=16, help="KV block size of the prefill attention runs (one TMA box = bs rows)")
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
Hq, Hkv, d, bs = 32, 8, 128, a.prefill_bs
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
    # small chunks: 8 launches captured in a CUDA graph and replayed inside one event pair.  A lone 2k-chunk launch
    # (~30 us of kernel) between two events -- or eight eager ones -- times the host side of the call (python argument
    # checks, three tensor-map encodes, the launch: ~50 us), which the serving step overlaps with the kernels in front
    reps = 1 if chunk >= 8192 else 8
    call = lambda: ops.prefill_attention(q, kc, vc, bt, cu, st, out=out)
    if reps > 1:
        call()
        torch.cuda.synchronize()
        side = torch.cuda.Stream()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            call()
            side.synchronize()
            with torch.cuda.graph(gr, stream=side):
                for _ in range(reps):
                    call()
        med, best = timeit(gr.replay, iters=10)
        med, best = med / reps, best / reps
    else:
        med, best = timeit(call, iters=10)
    pairs = chunk * start + chunk * (chunk + 1) // 2
    flops = 4 * d * Hq * pairs
    emit(op="prefill_attention", chunk=chunk, prefix=start, block_size=bs, Hq=Hq, Hkv=Hkv, launches_per_timing=reps, timed="cuda graph of the launches" if reps > 1 else "eager launch", ms=med, ms_min=best,
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
    if small:     # a decode step: every token in another block
        slots = torch.randperm(nblk * bs, device=dev)[:T].to(torch.int64)
    else:         # a prompt chunk: tokens fill 16-token blocks in order, the blocks are scattered over the pool
        blocks = torch.randperm(nblk, device=dev)
        ar = torch.arange(T, device=dev)
        slots = (blocks[ar // bs] * bs + ar % bs).to(torch.int64)
    table = torch.randn(max(T, 64), d, device=dev)
    med, best = timeit(lambda: ops.rope_and_cache(qkv, pos, table, kc, vc, slots, Hq), flush_l2=small, reps=reps)
    byt = T * ((Hq + Hkv) * d * 2 * 2 + Hkv * d * 2 + 2 * Hkv * d * 2 + d * 4)
    emit(op="rope_and_cache", T=T, us=med * 1e3, GBps=byt / med / 1e6, frac_hbm=byt / med / 1e6 / HBM,
         l2="flushed" if small else "working set > L2" if byt > 126e6 else "fits L2 (reported as is)",
         kernel="streaming (persistent grid, shared-memory ring fed by bulk copies)" if T >= 2 * 4 * 148 and os.environ.get("B2_ROPE_IMPL", "")[:1] != "t" else "one token per CTA")
    med, best = timeit(lambda: ops.rope_and_cache(qkv, pos, table, kc, vc, slots, Hq, write_k=False), flush_l2=small, reps=reps)
    byt2 = byt - T * Hkv * d * 2
    emit(op="rope_and_cache(write_k=0)", T=T, us=med * 1e3, GBps=byt2 / med / 1e6, frac_hbm=byt2 / med / 1e6 / HBM,
         note="rotated K goes to the cache only (what the engine runs)")
logits = torch.randn(64, 128256, device=dev).to(torch.bfloat16)
med, best = timeit(lambda: ops.sample(logits), flush_l2=True)
emit(op="sample(greedy)", B=64, V=128256, us=med * 1e3, GBps=64 * 128256 * 2 / med / 1e6, frac_hbm=64 * 128256 * 2 / med / 1e6 / HBM,
     l2="flushed")

# ---------------------------------------------------------------- the decode-step GEMMs (cuBLAS), per rank at tp 1 / 8
if a.only in ("all", "gemm"):
    M, Hd, I, QKV = 64, 4096, 14336, 6144
    for tp in (1, 2, 8):
        for name, K, N in (("wqkv", Hd, QKV // tp), ("wo", Hd // tp, Hd), ("wgu", Hd, 2 * I // tp), ("wdown", I // tp, Hd)):
            xg = torch.randn(M, K, device=dev).to(torch.bfloat16)
            # 16 different weight matrices back to back (as 16 layers): the weights stream from HBM, not from L2
            ws = [torch.randn(N, K, device=dev).to(torch.bfloat16) for _ in range(16)]
            og = torch.empty(M, N, device=dev, dtype=torch.bfloat16)

            def graphed(fn):
                """the loop as ONE CUDA graph: device time per GEMM without the host launch cost (how a decode step runs)"""
                st_ = torch.cuda.Stream()
                with torch.cuda.stream(st_):
                    fn()
                    st_.synchronize()
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr, stream=st_):
                        fn()
                st_.synchronize()
                return gr.replay

            def run():
                for wgt in ws:
                    torch.mm(xg, wgt.t(), out=og)
            run = graphed(run)
            med, best = timeit(run, iters=20)
            byt = N * K * 2 + M * K * 2 + M * N * 2
            emit(op="decode_gemm", name=name, tp=tp, M=M, K=K, N=N, us=med * 1e3 / 16, GBps=byt * 16 / med / 1e6,
                 frac_hbm=byt * 16 / med / 1e6 / HBM, note="cuBLAS via torch.mm, 16 distinct weights back to back, replayed as one CUDA graph")

            def run2():
                for wgt in ws:
                    ops.skinny_gemm(xg, wgt, out=og)
            run2 = graphed(run2)
            med2, best2 = timeit(run2, iters=20)
            emit(op="decode_gemm_tc5", name=name, tp=tp, M=M, K=K, N=N, us=med2 * 1e3 / 16, GBps=byt * 16 / med2 / 1e6,
                 frac_hbm=byt * 16 / med2 / 1e6 / HBM, vs_cublas=round(med / med2, 3),
                 note="b2_skinny_gemm (tcgen05, swap-AB, cluster split-K), same 16 weights")
            del ws

if a.skip_model:
    sys.exit(0)
# ---------------------------------------------------------------- whole-model decode tok/s, b64 ctx4k
if a.only not in ("all", "decode"):
    sys.exit(0)
kc = vc = q = out = gu = oo = logits = None
torch.cuda.empty_cache()
sys.path.insert(0, str(ROOT))
import bench_legs  # noqa: E402
rec = bench_legs.decode_tok_s_leg(dev, model=a.model, B=a.batch, ctx=a.ctx, steps=a.steps, kv=a.kv)
emit(op="decode_tok_s", **rec)
