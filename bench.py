#!/usr/bin/env python
"""bench.py -- paged-attention decode on B200: achieved HBM GB/s on BASELINE.json's headline config.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One "step" = one pass of the hot path over one decode batch of one layer:
    reshape_and_cache (the step's new K/V token of every sequence)  ->  paged_decode.
Workload at N = 1 (BASELINE.json configs[1], the configuration the metric is quoted on):
    Llama-3-8B attention geometry, 32 q / 8 kv heads, d = 128, block_size 16, batch 64, ctx 4096, bf16 KV.
At N > 1 (torchrun, one rank per GPU, NCCL only for the timing barrier) the KV heads are sharded over the
ranks and the batch grows with N (global batch 64*N), so per-GPU work is fixed: "scaling": "weak".  There is
no collective on the data path.

Printed on rank 0: ONE JSON line (see the keys below).  `value` is whole-job GB/s with every input resident
in HBM; `e2e` is the same metric through the public API with HOST (pinned) buffers, H2D/D2H copies inside
the timed region; `roofline` is for the dominant kernel (b2_paged_decode) from CUDA events recorded on the
launching stream inside the timed region; `cpu_baseline` is the CPU oracle on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
for _p in (ROOT / "vllm-backend-public_b200", ROOT / "oracle"):
    sys.path.insert(0, str(_p))

import torch  # noqa: E402

METRIC = "paged_decode_hbm_GBps"
UNIT = "GB/s"
NOMINAL_HBM_GBS = 8000.0

WORKLOADS = {
    # name: (B per GPU-equivalent, Hq, Hkv, ctx, kv dtype)
    "llama3-8b_b64_ctx4k_bf16kv": dict(B=64, Hq=32, Hkv=8, ctx=4096, fp8=False),     # BASELINE configs[1]
    "llama3-8b_b128_ctx8k_fp8kv": dict(B=128, Hq=32, Hkv=8, ctx=8192, fp8=True),     # BASELINE configs[2]
    "llama3-70b_b256_ctx8k_bf16kv": dict(B=256, Hq=64, Hkv=8, ctx=8192, fp8=False),  # BASELINE configs[3]
    "oracle_b4_ctx512": dict(B=4, Hq=8, Hkv=2, ctx=512, fp8=False),                 # BASELINE configs[0]
    "llama3-70b_tp8_shard": dict(B=256, Hq=8, Hkv=1, ctx=8192, fp8=False),           # one rank's share of configs[3] at tp = 8
}
D = 128
BS = int(os.environ.get("B2_BENCH_BLOCK_SIZE", "16"))   # BASELINE fixes 16; other sizes are experiments only


def algorithmic_bytes(B, Hq, Hkv, ctx, kv_bytes):
    """K+V read once, q read, out written, block table, seq_lens."""
    nb = (ctx + BS - 1) // BS
    return B * ctx * 2 * Hkv * D * kv_bytes + 2 * B * Hq * D * 2 + B * nb * 4 + B * 4


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# --------------------------------------------------------------------------------------------
# clocks during the timed region (NVML poll thread)
# --------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def __enter__(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t is not None:
            self._t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": int(statistics.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def physical_gpu_index(local_rank: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# --------------------------------------------------------------------------------------------
# synthetic problem (device-side generation)
# --------------------------------------------------------------------------------------------
def _seq_gen(dev, seed, b):
    return torch.Generator(device=dev).manual_seed(seed * 1000003 + b)


def gen_sequence_kv(dev, seed, b, ctx, Hkv):
    """K and V of sequence b, all kv heads: the same numbers on every rank (each keeps its head slice)."""
    g = _seq_gen(dev, seed, b)
    k = torch.randn(ctx, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    v = torch.randn(ctx, Hkv, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    return k, v


def build_problem(dev, B, Hq, Hkv, ctx, fp8, seed, shard=None, rows=None):
    """Random-permuted block tables; cache filled with N(0,1) K/V through b2_reshape_and_cache.

    The problem is a function of (seed, shape) only: every rank generates the full-head tensors of a sequence from a
    per-sequence seed and keeps the heads of `shard` (b200attn.parallel.HeadShard), so the sharded and the unsharded
    run see the same numbers.  rows: only these sequences of the B-sequence problem (same data, compact batch)."""
    import b200attn
    g = torch.Generator().manual_seed(seed)
    max_nb = (ctx + BS - 1) // BS
    seqs = list(range(B)) if rows is None else list(rows)
    n = len(seqs)
    num_blocks = n * max_nb + 64
    perm = torch.randperm(num_blocks, generator=g).to(torch.int32)
    block_tables = perm[: n * max_nb].reshape(n, max_nb).contiguous()
    seq_lens = torch.full((n,), ctx, dtype=torch.int32)
    kvh = slice(0, Hkv) if shard is None else shard.kv_heads
    qh = slice(0, Hq) if shard is None else shard.q_heads
    Hkv_l, Hq_l = kvh.stop - kvh.start, qh.stop - qh.start
    dt = torch.uint8 if fp8 else torch.bfloat16
    k_cache = torch.zeros(num_blocks, Hkv_l, BS, D, dtype=dt, device=dev)
    v_cache = torch.zeros_like(k_cache)
    scales = None
    if fp8:
        # N(0,1) inputs: |x| < ~6; per-tensor scale amax/448 with amax fixed at 6 (recorded in config)
        scales = (torch.tensor([6.0 / 448.0], device=dev), torch.tensor([6.0 / 448.0], device=dev))
    bt_dev = block_tables.to(dev)
    pos = torch.arange(ctx, device=dev)
    for i, b in enumerate(seqs):      # one sequence per launch keeps the staging buffers small
        slots = bt_dev[i, pos // BS].to(torch.int64) * BS + (pos % BS)
        k, v = gen_sequence_kv(dev, seed, b, ctx, Hkv)
        b200attn.reshape_and_cache(k[:, kvh], v[:, kvh], k_cache, v_cache, slots, *(scales or (None, None)))
    gq = torch.Generator(device=dev).manual_seed(seed + 77)
    q_all = torch.randn(B, Hq, D, generator=gq, device=dev, dtype=torch.float32).to(torch.bfloat16)
    # the step's new token (position ctx-1 of every sequence is rewritten each step)
    k_all = torch.randn(B, Hkv, D, generator=gq, device=dev, dtype=torch.float32).to(torch.bfloat16)
    v_all = torch.randn(B, Hkv, D, generator=gq, device=dev, dtype=torch.float32).to(torch.bfloat16)
    idx = torch.tensor(seqs, device=dev)
    q = q_all[idx][:, qh].contiguous()
    k_new = k_all[idx][:, kvh].contiguous()
    v_new = v_all[idx][:, kvh].contiguous()
    last = ctx - 1
    slot_new = (block_tables[:, last // BS].to(torch.int64) * BS + last % BS)
    torch.cuda.synchronize(dev)
    return dict(q=q, k_new=k_new, v_new=v_new, k_cache=k_cache, v_cache=v_cache, block_tables=block_tables,
                seq_lens=seq_lens, slot_new=slot_new, scales=scales, max_nb=max_nb, num_blocks=num_blocks,
                B=n, Hq=Hq_l, Hkv=Hkv_l, ctx=ctx, fp8=fp8)


def pick_schedule(prob, splits_arg):
    import b200attn
    if splits_arg < 0:
        return b200attn.plan_num_splits(prob["seq_lens"], prob["Hq"], prob["Hkv"], D, BS,
                                        b200attn.KV_FP8_E4M3 if prob["fp8"] else b200attn.KV_BF16), \
            "host planner (b2_paged_decode_plan_splits on host seq_lens)"
    return splits_arg, "--splits"


def schedule_name(num_splits):
    return {0: "dynamic persistent (zero + decode + merge kernels)", 1: "single-pass grid"}.get(
        num_splits, f"grid split x{num_splits} + merge")


def step_is_fused(prob, num_splits):
    """The step runs as ONE launch (b2_paged_decode_append: the CTA that owns (row, kv head) writes the row's new K/V
    itself) when the schedule is the single-pass grid and G <= 16; B2_BENCH_APPEND=0 keeps the two-call step (A/B)."""
    import b200attn
    return os.environ.get("B2_BENCH_APPEND", "1") != "0" and b200attn.append_is_fused(prob["Hq"], prob["Hkv"], num_splits)


def step_description(fused):
    if fused:
        return ("b2_paged_decode_append, one layer: KV write of 1 new token/seq fused into the single-pass decode grid "
                "(one launch; same bits as reshape_and_cache + paged_decode)")
    return "reshape_and_cache(1 new token/seq) + paged_decode, one layer"


def time_resident(dev, prob, num_splits, K, W, barrier, clock_index=None):
    """W warm-up steps, then K timed steps of [KV write of 1 new token / sequence + paged decode] with every input
    resident in HBM; CUDA events on the launching stream.  The step is b2_paged_decode_append (one launch) when
    step_is_fused, else b2_reshape_and_cache + b2_paged_decode.  -> (ms per step, mean / min ms of the decode call --
    the fused call includes the write --, out tensor, clocks summary)."""
    import b200attn
    fused = step_is_fused(prob, num_splits)
    B = prob["B"]
    bt_d, sl_d, slot_d = prob["block_tables"].to(dev), prob["seq_lens"].to(dev), prob["slot_new"].to(dev)
    ks, vs = prob["scales"] or (None, None)
    ws_bytes = b200attn.decode_workspace_bytes(B, prob["Hq"], prob["Hkv"], D, BS, prob["max_nb"], num_splits)
    ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=dev)
    out = torch.empty(B, prob["Hq"], D, dtype=torch.bfloat16, device=dev)
    stream = torch.cuda.current_stream(dev)

    def kv_write():
        if not fused:
            b200attn.reshape_and_cache(prob["k_new"], prob["v_new"], prob["k_cache"], prob["v_cache"], slot_d, ks, vs)

    def decode():
        if fused:
            b200attn.paged_decode_append(prob["q"], prob["k_new"], prob["v_new"], slot_d, prob["k_cache"],
                                         prob["v_cache"], bt_d, sl_d, k_scale=ks, v_scale=vs, out=out, workspace=ws,
                                         num_splits=num_splits)
        else:
            b200attn.paged_decode(prob["q"], prob["k_cache"], prob["v_cache"], bt_d, sl_d, k_scale=ks, v_scale=vs,
                                  out=out, workspace=ws, num_splits=num_splits)

    for _ in range(W):
        kv_write()
        decode()
    ev_all = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
    if fused:
        # the step IS one launch: one event between consecutive steps times each launch (interval i = launch i plus its
        # launch gap) -- two records per step would only add instrumentation to the timed region
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
        barrier()
        with ClockSampler(clock_index if clock_index is not None else 0) as clk:
            ev_all[0].record(stream)
            marks[0].record(stream)
            for i in range(K):
                decode()
                marks[i + 1].record(stream)
            ev_all[1].record(stream)
            barrier()
        dec_ms = [marks[i].elapsed_time(marks[i + 1]) for i in range(K)]
        return ev_all[0].elapsed_time(ev_all[1]) / K, statistics.mean(dec_ms), min(dec_ms), out, clk.summary()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    with ClockSampler(clock_index if clock_index is not None else 0) as clk:
        ev_all[0].record(stream)
        for i in range(K):
            kv_write()
            ev[i][0].record(stream)
            decode()
            ev[i][1].record(stream)
        ev_all[1].record(stream)
        barrier()
    dec_ms = [a.elapsed_time(b) for a, b in ev]
    return ev_all[0].elapsed_time(ev_all[1]) / K, statistics.mean(dec_ms), min(dec_ms), out, clk.summary()


# --------------------------------------------------------------------------------------------
# CPU arm: the oracle on a bounded sample of the same workload
# --------------------------------------------------------------------------------------------
def cpu_sample_problem(Hq, Hkv, ctx, fp8, n_seq, seed=0):
    import paged_attention_ref as R
    inp = R.make_inputs(n_seq, Hq, Hkv, D, BS, ctx, seed=seed, slack_blocks=4)
    dt = torch.uint8 if fp8 else torch.bfloat16
    kc = torch.zeros(inp["num_blocks"], Hkv, BS, D, dtype=dt)
    vc = torch.zeros_like(kc)
    kw = {}
    if fp8:
        kw = dict(kv_dtype=R.KV_FP8_E4M3, k_scale=6.0 / 448.0, v_scale=6.0 / 448.0)
        R.reshape_and_cache_ref(inp["k"], inp["v"], kc, vc, inp["slot_mapping"], R.KV_FP8_E4M3, 6.0 / 448.0, 6.0 / 448.0)
    else:
        R.reshape_and_cache_ref(inp["k"], inp["v"], kc, vc, inp["slot_mapping"])
    return inp, kc, vc, kw


def time_cpu_oracle(Hq, Hkv, ctx, fp8, n_seq, steps, warmup):
    """Each step: the oracle's KV write of one new token per sequence + paged decode, on n_seq sequences."""
    import paged_attention_ref as R
    torch.set_num_threads(os.cpu_count() or 1)
    inp, kc, vc, kw = cpu_sample_problem(Hq, Hkv, ctx, fp8, n_seq)
    last = ctx - 1
    slot_new = inp["block_tables"][:, last // BS].to(torch.int64) * BS + last % BS
    k_new = inp["k"][:n_seq].clone()
    v_new = inp["v"][:n_seq].clone()

    def step():
        if fp8:
            R.reshape_and_cache_ref(k_new, v_new, kc, vc, slot_new, R.KV_FP8_E4M3, kw["k_scale"], kw["v_scale"])
        else:
            R.reshape_and_cache_ref(k_new, v_new, kc, vc, slot_new)
        return R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"], **kw)

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    nbytes = algorithmic_bytes(n_seq, Hq, Hkv, ctx, 1 if fp8 else 2)
    return nbytes / dt / 1e9, dt, torch.get_num_threads()


def workload_config(wl_name, wl, B, tp):
    """The keys that define the workload -- identical in the GPU arm and the reference (CPU) arm."""
    return {"workload": wl_name, "global_batch": B, "ctx": wl["ctx"], "Hq": wl["Hq"], "Hkv": wl["Hkv"], "d": D,
            "block_size": BS, "kv_dtype": "fp8_e4m3" if wl["fp8"] else "bf16",
            "step": "reshape_and_cache(1 new token/seq) + paged_decode, one layer",
            "parallelism": f"kv-head-shard tp{tp}" if tp > 1 else "1 GPU"}


def run_reference_arm(args, wl_name, wl):
    """--impl reference: the reference repository has no implementation of this path (it launches an
    external vLLM container), so the CPU arm is the oracle port, on all host threads, on a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_seq = 8
    steps = max(1, min(args.steps, 20))
    warmup = max(1, min(args.warmup, 5))
    gbs, dt, cores = time_cpu_oracle(wl["Hq"], wl["Hkv"], wl["ctx"], wl["fp8"], n_seq, steps, warmup)
    sample = (f"{n_seq} of {wl['B']} sequences of {wl_name} per step (ctx {wl['ctx']}, all {wl['Hkv']} kv heads); "
              f"{steps} steps after {warmup} warm-up; fp32 PyTorch oracle; GB/s = the sample's algorithmic bytes / time")
    cfg = workload_config(wl_name, wl, wl["B"] * args.gpus, args.gpus)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 4), "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": round(gbs, 4), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(gbs, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
def roofline_record(bytes_rank, dec_mean_ms, dec_min_ms, traffic=None, fused=False):
    peak, peak_src = measured_peak()
    achieved = bytes_rank / (dec_mean_ms * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
            "kernel": ("b2_paged_decode_append (one launch: the step's KV write + the decode grid)" if fused
                       else "b2_paged_decode (every kernel the call launches)"),
            "kernel_ms": round(dec_mean_ms, 5), "kernel_ms_min": round(dec_min_ms, 5),
            "frac_of_8TBps": round(achieved / NOMINAL_HBM_GBS, 4)}


def committed_traffic(wl_name, B, Hq_l, Hkv_l):
    """dram bytes per launch from the committed `ncu --set full` capture -- only when it was taken on this very
    per-GPU shape (profiles/traffic.json records the shape next to the bytes); otherwise null."""
    tp_path = ROOT / "profiles" / "traffic.json"
    try:
        rec = json.loads(tp_path.read_text()).get(wl_name)
    except Exception:
        return None
    if isinstance(rec, dict):
        shape = rec.get("per_gpu", {})
        if (shape.get("B"), shape.get("Hq"), shape.get("Hkv")) == (B, Hq_l, Hkv_l):
            return rec.get("dram_bytes_per_launch")
        return None
    return None


def sub_config_record(dev, name, K, W, splits_arg):
    """One more BASELINE config measured the same way as the headline (device-resident steps, CUDA events): the
    record that goes under `configs` in the JSON line."""
    wl = WORKLOADS[name]
    prob = build_problem(dev, wl["B"], wl["Hq"], wl["Hkv"], wl["ctx"], wl["fp8"], seed=0)
    num_splits, src = pick_schedule(prob, splits_arg)
    ms_step, dec_mean, dec_min, _, clk = time_resident(dev, prob, num_splits, K, W, lambda: torch.cuda.synchronize(dev),
                                                       physical_gpu_index(dev.index or 0))
    nbytes = algorithmic_bytes(wl["B"], wl["Hq"], wl["Hkv"], wl["ctx"], 1 if wl["fp8"] else 2)
    peak, _ = measured_peak()
    rec = {"workload": name, "B": wl["B"], "Hq": wl["Hq"], "Hkv": wl["Hkv"], "ctx": wl["ctx"],
           "kv_dtype": "fp8_e4m3" if wl["fp8"] else "bf16", "algorithmic_bytes": nbytes,
           "value": round(nbytes / (ms_step * 1e-3) / 1e9, 2), "ms_per_step": round(ms_step, 5),
           "kernel_ms": round(dec_mean, 5), "kernel_ms_min": round(dec_min, 5),
           "achieved": round(nbytes / (dec_mean * 1e-3) / 1e9, 2),
           "frac": round(nbytes / (dec_mean * 1e-3) / 1e9 / peak, 4),
           "frac_of_8TBps": round(nbytes / (dec_mean * 1e-3) / 1e9 / NOMINAL_HBM_GBS, 4),
           "num_splits": num_splits, "schedule": schedule_name(num_splits), "schedule_chosen_by": src,
           "step_impl": step_description(step_is_fused(prob, num_splits)),
           "steps": K, "warmup": W, "clocks": clk}
    del prob
    torch.cuda.empty_cache()
    return rec


def sharded_parity_check(dev, dist, rank, wl, shard, out_local, seed, n_rows=8, n_oracle=4):
    """The head-sharded result, gathered once (outside every timed region), against (1) rank 0 running the UNSHARDED
    problem (all heads, b2_paged_decode) on a slice of the batch and (2) the CPU oracle on some of those rows.
    Rows: first, last and random ones.  Bound: |d| <= 2e-3 + 1e-2 |ref| (north_star, bf16 KV)."""
    import b200attn
    from b200attn import parallel
    B, Hq, Hkv, ctx = wl["B"], wl["Hq"], wl["Hkv"], wl["ctx"]
    full = parallel.gather_heads(out_local, shard)            # [B, Hq, d] on every rank
    rec = None
    if rank == 0:
        g = torch.Generator().manual_seed(99)
        rows = sorted({0, B - 1} | set(torch.randperm(B, generator=g)[: max(n_rows - 2, 0)].tolist()))
        sub = build_problem(dev, B, Hq, Hkv, ctx, wl["fp8"], seed, shard=None, rows=rows)
        ks, vs = sub["scales"] or (None, None)
        # the timed steps rewrote position ctx-1 with k_new / v_new: do the same here
        b200attn.reshape_and_cache(sub["k_new"], sub["v_new"], sub["k_cache"], sub["v_cache"], sub["slot_new"].to(dev), ks, vs)
        un = b200attn.paged_decode(sub["q"], sub["k_cache"], sub["v_cache"], sub["block_tables"].to(dev),
                                   sub["seq_lens"].to(dev), k_scale=ks, v_scale=vs, num_splits=1)
        got = full[torch.tensor(rows, device=dev)].float()
        err_un = (got - un.float()).abs()
        ok_un = bool((err_un <= 2e-3 + 1e-2 * un.float().abs()).all())
        # oracle on the first n_oracle of those rows, from the unsharded device cache copied back
        import paged_attention_ref as R
        orows = list(range(min(n_oracle, len(rows))))
        need = torch.unique(sub["block_tables"][orows].flatten()).long()
        kc_ref = torch.zeros(sub["k_cache"].shape, dtype=sub["k_cache"].dtype)
        vc_ref = torch.zeros_like(kc_ref)
        kc_ref[need] = sub["k_cache"][need.to(dev)].cpu()
        vc_ref[need] = sub["v_cache"][need.to(dev)].cpu()
        okw = dict(kv_dtype=R.KV_FP8_E4M3, k_scale=6.0 / 448.0, v_scale=6.0 / 448.0) if wl["fp8"] else {}
        ref = R.paged_decode_ref(sub["q"][orows].cpu(), kc_ref, vc_ref, sub["block_tables"][orows], sub["seq_lens"][orows],
                                 return_f32=True, **okw)
        err_or = (got[orows].cpu() - ref).abs()
        ok_or = bool((err_or <= 2e-3 + 1e-2 * ref.abs()).all()) and bool(torch.isfinite(got).all())
        rec = {"rows_checked": len(rows), "rows": rows, "heads_checked": Hq,
               "max_abs_err_vs_unsharded_kernel": float(err_un.max()), "within_bound_vs_unsharded": ok_un,
               "oracle_rows": len(orows), "max_abs_err_vs_oracle": float(err_or.max()), "within_bound_vs_oracle": ok_or,
               "max_abs_err": float(max(err_un.max(), err_or.max())), "bound": "2e-3 + 1e-2*|ref|",
               "ok": ok_un and ok_or,
               "how": "out [B, Hq/tp, d] of every rank all-gathered once (NCCL, outside the timed regions) and compared on "
                      "rank 0 with b2_paged_decode over ALL heads of the same sequences, and with the CPU oracle"}
        del sub
        torch.cuda.empty_cache()
    return rec


def run_ours(args, wl_name, wl):
    import b200attn
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit(f"--gpus {args.gpus} needs torchrun --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    dist = None
    if world > 1:
        import torch.distributed as dist
        from datetime import timedelta
        # a rank that fails alone must not leave the others waiting for NCCL's default 10 minutes
        dist.init_process_group("nccl", device_id=dev, timeout=timedelta(minutes=4))

    tp = world
    if wl["Hkv"] % tp:
        raise SystemExit(f"Hkv={wl['Hkv']} is not divisible by tp={tp}")
    # KV-head sharding, weak scaling: rank r owns kv heads [r*Hkv/tp, (r+1)*Hkv/tp) of a batch of B*tp sequences
    # (--strong keeps the batch fixed instead: BASELINE configs[3], B = 256 sharded over tp = 2/4/8)
    from b200attn import parallel
    shard = parallel.HeadShard(tp, rank, wl["Hq"], wl["Hkv"])
    B = wl["B"] if args.strong else wl["B"] * tp
    Hq_l, Hkv_l = shard.Hq_local, shard.Hkv_local
    ctx, fp8 = wl["ctx"], wl["fp8"]
    kv_bytes = 1 if fp8 else 2
    prob = build_problem(dev, B, wl["Hq"], wl["Hkv"], ctx, fp8, seed=0, shard=shard)
    bytes_rank = algorithmic_bytes(B, Hq_l, Hkv_l, ctx, kv_bytes)
    if args.ragged:
        # experiment (not a BASELINE config): sequence lengths log-uniform in [128, ctx], the shape of a serving batch
        g = torch.Generator().manual_seed(1234)
        lens = torch.exp(torch.empty(B).uniform_(math.log(128.0), math.log(float(ctx)), generator=g)).to(torch.int32)
        lens = lens.clamp_(128, ctx)
        prob["seq_lens"] = lens
        last = (lens - 1).long()
        prob["slot_new"] = prob["block_tables"][torch.arange(B), last // BS].to(torch.int64) * BS + last % BS
        nbs = (lens.long() + BS - 1) // BS
        bytes_rank = int(lens.long().sum()) * 2 * Hkv_l * D * kv_bytes + 2 * B * Hq_l * D * 2 + int(nbs.sum()) * 4 + B * 4

    ks, vs = prob["scales"] or (None, None)
    # schedule: the engine's scheduler holds seq_lens on the host and picks the schedule from them (--splits -1)
    num_splits, schedule_src = pick_schedule(prob, args.splits)
    schedule = schedule_name(num_splits)
    if step_is_fused(prob, num_splits):
        launches_per_step = 1                                  # b2_paged_decode_append: the write rides in the decode grid
    elif num_splits == 1:
        launches_per_step = 2                                  # reshape + decode
    elif num_splits == 0:
        launches_per_step = 4                                  # reshape + zero-counter + decode + merge
    else:
        launches_per_step = 3                                  # reshape + decode + merge

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(*vals):
        t = torch.tensor(list(vals), dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t]

    K = args.steps
    # ---------------- device-resident timing: `value`, `roofline` ----------------
    ms_per_step, dec_mean_ms, dec_min_ms, out, clocks = time_resident(dev, prob, num_splits, K, args.warmup, barrier,
                                                                      physical_gpu_index(local_rank))
    ms_per_step, dec_mean_ms = max_over_ranks(ms_per_step, dec_mean_ms)
    value = bytes_rank * world / (ms_per_step * 1e-3) / 1e9

    # ---------------- end to end through the public API with HOST buffers: `e2e` ----------------
    # b200attn.DecodeStep: the step's q|k|v rows and seq_lens|slots in pinned host memory, block tables RESIDENT on
    # the device (kept current by the decode-step KV write), both kernels, result rows stored by the kernels straight
    # into pinned host memory, stream sync; replayed as one CUDA graph per step.
    step = b200attn.DecodeStep(prob["k_cache"], prob["v_cache"], max_batch=B, Hq=Hq_l, max_nb=prob["max_nb"],
                               k_scale=ks, v_scale=vs,
                               zero_copy_inputs=os.environ.get("B2_E2E_ZERO_COPY", "1") == "1")
    step.h_qkv[:, :Hq_l].copy_(prob["q"].cpu())
    step.h_qkv[:, Hq_l:Hq_l + Hkv_l].copy_(prob["k_new"].cpu())
    step.h_qkv[:, Hq_l + Hkv_l:].copy_(prob["v_new"].cpu())
    step.h_block_tables.copy_(prob["block_tables"])
    step.upload_block_tables()          # resident tables: uploaded once, kept current by the decode-step KV write
    step.h_seq_lens.copy_(prob["seq_lens"])
    step.h_slots.copy_(prob["slot_new"])
    h2d, d2h = step.h2d_bytes, step.d2h_bytes
    e2e_splits = None if args.splits < 0 else args.splits      # None: planned on the host inside every step
    for _ in range(max(3, args.warmup // 4)):
        h_res = step.run(B, e2e_splits)
    if not torch.equal(h_res.view(torch.int16), out.cpu().view(torch.int16)) and rank == 0:
        diff = (h_res.float() - out.cpu().float()).abs().max().item()
        if not diff < 2e-2:
            raise SystemExit(f"e2e result differs from the device-resident result (max abs diff {diff})")
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        step.run(B, e2e_splits)
    barrier()
    e2e_ms, = max_over_ranks((time.perf_counter() - t0) * 1e3 / K)
    e2e_value = bytes_rank * world / (e2e_ms * 1e-3) / 1e9
    e2e_zero_copy = step.zero_copy_inputs and num_splits != 0
    del step

    # optional consumer-side all-gather of the head-sharded output (off the attention path; reported, not in `value`)
    allgather_us = None
    if dist is not None:
        stream = torch.cuda.current_stream(dev)
        for _ in range(3):
            parallel.gather_heads(out, shard)
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        g0.record(stream)
        for _ in range(20):
            parallel.gather_heads(out, shard)
        g1.record(stream)
        barrier()
        allgather_us = round(max_over_ranks(g0.elapsed_time(g1) / 20 * 1e3)[0], 2)

    line = None
    if rank == 0:
        cfg = workload_config(wl_name + ("+ragged(log-uniform 128..ctx)" if args.ragged else ""), wl, B, tp)
        cfg.update({
            "collective": "none on the attention path" + (f"; optional NCCL all-gather of out [B, Hq, d]: {allgather_us} us"
                                                           if allgather_us is not None else ""),
            "per_gpu": {"B": B, "Hq": Hq_l, "Hkv": Hkv_l}, "num_splits": num_splits,
            "step_impl": step_description(step_is_fused(prob, num_splits)),
            "schedule": schedule, "schedule_chosen_by": schedule_src,
            "algorithmic_bytes_per_step_per_gpu": bytes_rank,
            "l2": "KV working set per GPU >> 126 MB L2, every step streams it from HBM; no flush needed"})
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
            "ms_per_step": round(ms_per_step, 5), "higher_is_better": True, "scaling": "strong" if args.strong else "weak", "vs_baseline": None,
            "dtype": "bf16" if not fp8 else "fp8_e4m3 KV / bf16 q,out (fp32 accumulate)", "data": "synthetic",
            "config": cfg,
            "pct_of_8TBps": round(100.0 * value / world / NOMINAL_HBM_GBS, 2),
            "roofline": roofline_record(bytes_rank, dec_mean_ms, dec_min_ms, committed_traffic(wl_name, B, Hq_l, Hkv_l),
                                        fused=step_is_fused(prob, num_splits)),
            "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": round(e2e_ms, 5),
                    "path": ("b200attn.DecodeStep: block tables resident on the device (uploaded when a sequence takes a "
                             "batch row; the decode-step KV write keeps them current), per step an H2D copy of "
                             "seq_lens|slots from pinned memory, "
                             + ("q|k|v rows read by the kernels straight from pinned host memory (the H2D leg, no "
                                "staging copy), " if e2e_zero_copy else "H2D copy of the q|k|v rows, ")
                             + "reshape_and_cache_decode + decode-plan run, result rows stored by the kernels into pinned "
                               "host memory as 16-byte vectors (the D2H leg), stream sync; one CUDA-graph replay per "
                               "step, schedule re-planned on the host every 8 steps (lengths grow by one token per step)")},
            "gpu_launches": launches_per_step * K,
            "clocks": clocks,
        }
    del prob, out
    torch.cuda.empty_cache()

    # ---------------- the other BASELINE configs, the model-level half of the metric, the comparator (N = 1) ----------
    extras = not args.lean and not args.ragged and wl_name == "llama3-8b_b64_ctx4k_bf16kv"
    if extras and world == 1:
        import bench_legs
        Kx = max(3, min(K, 200))
        line["configs"] = {}
        for name, key in (("llama3-8b_b128_ctx8k_fp8kv", "fp8_b128_ctx8k"), ("llama3-70b_b256_ctx8k_bf16kv", "llama70b_b256_ctx8k_tp1")):
            try:
                line["configs"][key] = sub_config_record(dev, name, Kx, max(3, min(args.warmup, 20)), args.splits)
            except Exception as e:                                        # noqa: BLE001
                line["configs"][key] = {"error": f"{type(e).__name__}: {e}"[:300]}
        line["comparator"] = bench_legs.comparator_leg()
        line["decode_tok_s"] = bench_legs.decode_tok_s_leg(dev, steps=max(20, min(K, 100)))

    # ---------------- BASELINE configs[3], strong scaling over the ranks + sharded-vs-unsharded parity (N > 1) --------
    if extras and world > 1:
        wl4 = WORKLOADS["llama3-70b_b256_ctx8k_bf16kv"]
        sh4 = parallel.HeadShard(tp, rank, wl4["Hq"], wl4["Hkv"])
        Kx, Wx = max(3, min(K, 100)), max(3, min(args.warmup, 10))
        p4 = build_problem(dev, wl4["B"], wl4["Hq"], wl4["Hkv"], wl4["ctx"], False, seed=4, shard=sh4)
        ns4, src4 = pick_schedule(p4, args.splits)
        ms4, dec4, dec4_min, out4, clk4 = time_resident(dev, p4, ns4, Kx, Wx, barrier, physical_gpu_index(local_rank))
        ms4, dec4 = max_over_ranks(ms4, dec4)
        parity = sharded_parity_check(dev, dist, rank, wl4, sh4, out4, seed=4)
        bytes4 = algorithmic_bytes(wl4["B"], sh4.Hq_local, sh4.Hkv_local, wl4["ctx"], 2)
        del p4, out4
        torch.cuda.empty_cache()
        barrier()
        tp1 = None
        if rank == 0:          # the tp = 1 record of the same config, on this box, for the efficiency figure
            tp1 = sub_config_record(dev, "llama3-70b_b256_ctx8k_bf16kv", Kx, Wx, args.splits)
        barrier()
        if rank == 0:
            peak, _ = measured_peak()
            val4 = bytes4 * world / (ms4 * 1e-3) / 1e9
            line["strong_config4"] = {
                "workload": "llama3-70b_b256_ctx8k_bf16kv", "scaling": "strong", "global_batch": wl4["B"], "tp": tp,
                "per_gpu": {"B": wl4["B"], "Hq": sh4.Hq_local, "Hkv": sh4.Hkv_local, "algorithmic_bytes": bytes4},
                "value": round(val4, 2), "unit": UNIT, "ms_per_step": round(ms4, 5),
                "per_gpu_kernel_ms": round(dec4, 5), "per_gpu_achieved": round(bytes4 / (dec4 * 1e-3) / 1e9, 2),
                "per_gpu_frac": round(bytes4 / (dec4 * 1e-3) / 1e9 / peak, 4),
                "num_splits": ns4, "schedule": schedule_name(ns4), "schedule_chosen_by": src4, "steps": Kx, "warmup": Wx,
                "tp1": {k: tp1[k] for k in ("value", "ms_per_step", "kernel_ms", "achieved", "frac", "num_splits")},
                # the same sequences on tp GPUs vs on one: time ratio per GPU (block tables are replicated, so the
                # per-GPU bytes are not exactly 1/tp)
                "efficiency_vs_tp1": round(tp1["ms_per_step"] / (ms4 * world), 4),
                "speedup_vs_tp1": round(tp1["ms_per_step"] / ms4, 3),
                "collective": "none on the attention path", "clocks": clk4}
            line["sharded_parity"] = parity

    # ---------------- the served engine at --tensor-parallel-size N (rows a4 / n1): parity, exchange cost, tok/s ------
    if extras and world > 1:
        import bench_legs
        from b200attn.tp import TPGroup
        group = TPGroup(dev)
        tp_rec = {"parity": bench_legs.tp_parity_leg(dev, group),
                  "exchange": bench_legs.tp_comm_leg(dev, group),
                  "decode_tok_s": bench_legs.tp_decode_leg(dev, group, steps=max(20, min(K, 50)))}
        if rank == 0:
            line["tp_serving"] = tp_rec

    if rank == 0:
        if world == 1 and not args.no_cpu:
            n_seq = 8
            gbs, dt, cores = time_cpu_oracle(wl["Hq"], wl["Hkv"], ctx, fp8, n_seq, steps=10, warmup=1)
            line["cpu_baseline"] = {
                "value": round(gbs, 4), "unit": UNIT, "cores": cores, "kind": "port",
                "sample": f"{n_seq} of {wl['B']} sequences per step, 10 steps after 1 warm-up, fp32 PyTorch oracle "
                          f"({dt * 1e3:.1f} ms/step)"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="llama3-8b_b64_ctx4k_bf16kv")
    ap.add_argument("--splits", type=int, default=-1,
                    help="num_splits for b2_paged_decode: -1 = host planner (default), 0 = dynamic persistent, 1 = single pass")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--lean", action="store_true",
                    help="headline leg only: skip the sub-records (other BASELINE configs, decode tok/s, comparator, "
                         "strong scaling + sharded parity)")
    ap.add_argument("--strong", action="store_true", help="N > 1: keep the global batch fixed (strong scaling)")
    ap.add_argument("--batch", type=int, default=0, help="override the workload batch size (experiments only)")
    ap.add_argument("--ragged", action="store_true",
                    help="experiment: sequence lengths log-uniform in [128, ctx] instead of all = ctx")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.batch:
        wl["B"] = args.batch
    if args.impl == "reference":
        run_reference_arm(args, args.workload, wl)
    else:
        run_ours(args, args.workload, wl)


if __name__ == "__main__":
    main()
