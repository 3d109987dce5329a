"""Per-CTA timeline of the decode kernels (diagnostic).  Needs `make -C vllm-backend-public_b200/csrc trace`.

    B200ATTN_LIB=vllm-backend-public_b200/lib/libb200attn_trace.so python scripts/trace_schedule.py [workload]

For each schedule (grid split 1, dynamic with several chunk sizes / static shares) it runs one traced launch on the headline config
and prints: kernel span, distribution of CTA start / first-data / end times, per-CTA busy time, idle tail.
"""
import ctypes as C
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
os.environ.setdefault("B200ATTN_LIB", str(ROOT / "vllm-backend-public_b200/lib/libb200attn_trace.so"))
sys.path.insert(0, str(ROOT / "vllm-backend-public_b200"))
sys.path.insert(0, str(ROOT))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import b200attn  # noqa: E402
import bench  # noqa: E402

wl_name = sys.argv[1] if len(sys.argv) > 1 else "llama3-8b_b64_ctx4k_bf16kv"
wl = bench.WORKLOADS[wl_name]
dev = torch.device("cuda:0")
lib = b200attn.load()
lib.b2_debug_read_trace.argtypes = [C.c_void_p, C.c_int]
prob = bench.build_problem(dev, wl["B"], wl["Hq"], wl["Hkv"], wl["ctx"], wl["fp8"], seed=0)
if len(sys.argv) > 2 and sys.argv[2] == "ragged":        # the serving shape: lengths log-uniform in [128, ctx] (as bench.py --ragged)
    import math
    g = torch.Generator().manual_seed(1234)
    lens = torch.exp(torch.empty(wl["B"]).uniform_(math.log(128.0), math.log(float(wl["ctx"])), generator=g)).to(torch.int32)
    prob["seq_lens"] = lens.clamp_(128, wl["ctx"])
    print("ragged lengths: sum", int(prob["seq_lens"].sum()), "max", int(prob["seq_lens"].max()))
bt, sl = prob["block_tables"].to(dev), prob["seq_lens"].to(dev)
ks, vs = prob["scales"] or (None, None)


def pct(a, q):
    return float(np.percentile(a, q))


def run(label, splits, env):
    for k in ("B2_CHUNK_TOKENS", "B2_STATIC_PCT"):
        os.environ.pop(k, None)
    os.environ.update(env)
    ws = torch.zeros(max(16, b200attn.decode_workspace_bytes(wl["B"], wl["Hq"], wl["Hkv"], 128, 16, prob["max_nb"], splits)),
                     dtype=torch.uint8, device=dev)
    out = torch.empty(wl["B"], wl["Hq"], 128, dtype=torch.bfloat16, device=dev)
    call = lambda: b200attn.paged_decode(prob["q"], prob["k_cache"], prob["v_cache"], bt, sl, k_scale=ks, v_scale=vs,  # noqa: E731
                                         out=out, workspace=ws, num_splits=splits)
    for _ in range(5):
        call()
    torch.cuda.synchronize()
    lib.b2_debug_clear_trace()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    call()
    e1.record()
    torch.cuda.synchronize()
    buf = np.zeros(2048 * 8, dtype=np.uint64)
    lib.b2_debug_read_trace(buf.ctypes.data, buf.size)
    t = buf.reshape(2048, 8).astype(np.int64)
    used = t[:, 0] > 0
    t = t[used]
    t0 = t[:, 0].min()
    start = (t[:, 0] - t0) / 1e3
    ready = (t[:, 1] - t0) / 1e3
    end = (t[:, 3] - t0) / 1e3
    ok = t[:, 3] > 0
    span = end[ok].max()
    print(f"== {label}: event {e0.elapsed_time(e1) * 1e3:.1f} us, traced CTAs {used.sum()}, span(first start -> last consumer end) {span:.1f} us")
    print(f"   start   us: min {start.min():.1f} p50 {pct(start, 50):.1f} p90 {pct(start, 90):.1f} max {start.max():.1f}")
    print(f"   ready   us (prologue done): p50 {pct(ready[ok] - start[ok], 50):.2f} max {(ready[ok] - start[ok]).max():.2f} after start")
    print(f"   end     us: min {end[ok].min():.1f} p10 {pct(end[ok], 10):.1f} p50 {pct(end[ok], 50):.1f} p90 {pct(end[ok], 90):.1f} max {end[ok].max():.1f}")
    busy = end[ok] - start[ok]
    print(f"   busy    us: min {busy.min():.1f} p50 {pct(busy, 50):.1f} max {busy.max():.1f}   sum/({used.sum()} x span) = {busy.sum() / (ok.sum() * span):.3f}")
    if t[:, 5].max() > 0:
        st = t[:, 5][ok]
        print(f"   stages per CTA: min {st.min()} p50 {int(pct(st, 50))} max {st.max()}  (x64 tokens)")
    # slots busy over time (how many CTAs are alive), sampled every 10 us
    grid = np.arange(0, span, 10.0)
    alive = [(int(((start[ok] <= g) & (end[ok] > g)).sum())) for g in grid]
    print("   alive CTAs every 10 us:", alive)


run("grid split, num_splits=1", 1, {})
run("dynamic pct85 ct512", 0, {"B2_STATIC_PCT": "85", "B2_CHUNK_TOKENS": "512"})
run("dynamic pct85 ct256", 0, {"B2_STATIC_PCT": "85", "B2_CHUNK_TOKENS": "256"})
run("dynamic default", 0, {})
run("dynamic pct0 ct128", 0, {"B2_STATIC_PCT": "0", "B2_CHUNK_TOKENS": "128"})
