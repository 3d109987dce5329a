"""One prompt-chunk attention call at Llama-3-8B heads (8k chunk after an 8k prefix) for ncu:
    ncu --set full --clock-control none --import-source on -k regex:prefill -c 1 -o gpurun_out/prof_prefill python scripts/prefill_once.py"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "vllm-backend-public_b200"))
import torch
from b200attn import ops
dev = torch.device("cuda:0")
Hq, Hkv, d, bs, chunk, start = 32, 8, 128, 16, 8192, 8192
nb = (start + chunk) // bs
kc = torch.randn(nb + 8, Hkv, bs, d, device=dev).to(torch.bfloat16)
vc = torch.randn(nb + 8, Hkv, bs, d, device=dev).to(torch.bfloat16)
q = torch.randn(chunk, Hq, d, device=dev).to(torch.bfloat16)
bt = torch.randperm(nb + 8, device=dev)[:nb].to(torch.int32).unsqueeze(0).contiguous()
cu = torch.tensor([0, chunk], dtype=torch.int32, device=dev)
st = torch.tensor([start], dtype=torch.int32, device=dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    out = ops.prefill_attention(q, kc, vc, bt, cu, st)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
