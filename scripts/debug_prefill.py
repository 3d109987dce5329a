"""Debug aid: one small prompt-chunk attention case against the oracle, per-row error report."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "vllm-backend-public_b200")); sys.path.insert(0, str(ROOT))
import torch
from b200attn import ops
dev = torch.device("cuda:0")
length, start = int(sys.argv[1]), int(sys.argv[2])
Hq, Hkv, d, bs = 8, 2, 128, 16
torch.manual_seed(0)
total = start + length
nb = (total + bs - 1) // bs
kc = torch.randn(nb + 4, Hkv, bs, d).to(torch.bfloat16)
vc = torch.randn(nb + 4, Hkv, bs, d).to(torch.bfloat16)
q = torch.randn(length, Hq, d).to(torch.bfloat16)
bt = torch.randperm(nb + 4)[:nb].to(torch.int32).unsqueeze(0).contiguous()
cu = torch.tensor([0, length], dtype=torch.int32)
st = torch.tensor([start], dtype=torch.int32)
out = ops.prefill_attention(q.to(dev), kc.to(dev), vc.to(dev), bt.to(dev), cu.to(dev), st.to(dev)).float().cpu()
# dense reference
K = kc[bt[0].long()].permute(1, 0, 2, 3).reshape(Hkv, nb * bs, d)[:, :total].float()
V = vc[bt[0].long()].permute(1, 0, 2, 3).reshape(Hkv, nb * bs, d)[:, :total].float()
G = Hq // Hkv
ref = torch.empty(length, Hq, d)
for h in range(Hq):
    s = (q[:, h].float() @ K[h // G].T) * d ** -0.5
    pos = start + torch.arange(length)[:, None]
    s = s.masked_fill(torch.arange(total)[None, :] > pos, float("-inf"))
    ref[:, h] = torch.softmax(s, -1) @ V[h // G]
err = (out - ref).abs().amax(dim=(1, 2))
bad = (err > 1e-2).nonzero().flatten().tolist()
print("max err", float(err.max()), "bad rows", len(bad), bad[:10], "...", bad[-5:])
