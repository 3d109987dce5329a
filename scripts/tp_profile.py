#!/usr/bin/env python
"""Where does a mixed (prompt-chunk) step go at tensor_parallel_size N?  torchrun --nproc-per-node N scripts/tp_profile.py
Rank 0 profiles two mixed steps with torch.profiler and prints the top CUDA kernels / CPU ops and the wall time per step."""
import os
import random
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "vllm-backend-public_b200"):
    sys.path.insert(0, str(p))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", "0"))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
from b200attn.engine import LLMEngine, SamplingParams  # noqa: E402
from b200attn.model import resolve_config  # noqa: E402
from b200attn.tp import TPGroup  # noqa: E402

group = TPGroup(dev)
cfg = resolve_config(os.environ.get("MODEL", "llama-3.1-8b"))
eng = LLMEngine(cfg, dev, max_num_seqs=64, max_model_len=16384, max_num_batched_tokens=8192, tp=group if group.size > 1 else None)
if rank != 0:
    eng.worker_loop()
else:
    rng = random.Random(0)
    for n in (8000, 8000, 8000, 8000, 8000, 8000, 12000, 12000):
        eng.add_request([rng.randrange(cfg.vocab) for _ in range(n)], SamplingParams(max_tokens=64, temperature=0.0))
    for _ in range(2):
        eng.step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n0 = eng.stats.mixed_steps
    for _ in range(3):
        eng.step()
    torch.cuda.synchronize()
    print(f"tp={group.size}: {(time.perf_counter() - t0) / 3 * 1e3:.1f} ms per mixed step ({eng.stats.mixed_steps - n0} mixed)", flush=True)
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            eng.step()
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=70))
    print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=12, max_name_column_width=70))
    eng.run_until_idle()
    eng.shutdown_workers()
eng.close()
dist.barrier()
dist.destroy_process_group()
