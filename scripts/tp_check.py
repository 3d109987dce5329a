#!/usr/bin/env python
"""Tensor-parallel serving check, run under torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/tp_check.py [--legs parity,exchange,decode] [--model llama-3-8b] [--steps 50]

Rank 0 prints one JSON line: teacher-forcing parity of the served engine against the CPU oracle, the fused NVLink
all-reduce + RMSNorm kernel against NCCL + plain norm, whole-model decode tok/s at this tp."""
import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "vllm-backend-public_b200", ROOT / "oracle"):
    sys.path.insert(0, str(p))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--legs", default="parity,exchange,decode")
    ap.add_argument("--model", default="llama-3-8b")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=50)
    args = ap.parse_args()
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    from datetime import timedelta
    dist.init_process_group("nccl", device_id=dev, timeout=timedelta(minutes=4))
    import bench_legs
    from b200attn.tp import TPGroup
    group = TPGroup(dev)
    legs = args.legs.split(",")
    rec = {"tp": group.size}
    if "parity" in legs:
        rec["parity"] = bench_legs.tp_parity_leg(dev, group)
    if "exchange" in legs:
        rec["exchange"] = bench_legs.tp_comm_leg(dev, group)
        rec["exchange_T8"] = bench_legs.tp_comm_leg(dev, group, T=8)
        rec["exchange_T256"] = bench_legs.tp_comm_leg(dev, group, T=256)
    if "decode" in legs:
        rec["decode_tok_s"] = bench_legs.tp_decode_leg(dev, group, model=args.model, B=args.batch, ctx=args.ctx,
                                                       steps=args.steps)
    if rank == 0:
        print(json.dumps(rec), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
