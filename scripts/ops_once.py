"""One call each of the element-wise layer ops at T = 8192 (Llama-3-8B widths) for ncu:
    ncu --set full --clock-control none --import-source on -k regex:'rope_and_cache|rms_norm' -o gpurun_out/prof_ops python scripts/ops_once.py"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "vllm-backend-public_b200"))
import torch
from b200attn import ops
from b200attn.model import rope_table
dev = torch.device("cuda:0")
T, H, Hq, Hkv, d, bs = 8192, 4096, 32, 8, 128, 16
x = torch.randn(T, H, device=dev).to(torch.bfloat16)
res = torch.randn(T, H, device=dev).to(torch.bfloat16)
w = torch.ones(H, device=dev, dtype=torch.bfloat16)
qkv = torch.randn(T, (Hq + 2 * Hkv) * d, device=dev).to(torch.bfloat16)
pos = torch.arange(T, dtype=torch.int32, device=dev)
cs = rope_table(T, d, 500000.0, None, dev)
nb = T // bs
kc = torch.zeros(nb, Hkv, bs, d, dtype=torch.bfloat16, device=dev)
vc = torch.zeros_like(kc)
import os
if os.environ.get("SLOTS", "prefill") == "random":
    slots = torch.randperm(T, device=dev).to(torch.int64)                      # every token in another block (worst case)
else:   # a prompt chunk: tokens fill 16-token blocks in order, the blocks are scattered over the pool
    blocks = torch.randperm(nb, device=dev)
    slots = (blocks[torch.arange(T, device=dev) // bs] * bs + torch.arange(T, device=dev) % bs).to(torch.int64)
for _ in range(3):
    ops.rms_norm(x, w, 1e-5, residual=res)
    ops.rope_and_cache(qkv, pos, cs, kc, vc, slots, Hq, write_k=os.environ.get("WRITE_K", "1") == "1")
torch.cuda.synchronize()
print("ok")
