"""Timeline of the prefill kernel's softmax warps (diagnostic; needs `make -C vllm-backend-public_b200/csrc trace`):
clock64 stamps of tile A / tile B of CTA (0, 0) for the first 32 key tiles of an 8k chunk over an 8k prefix."""
import ctypes as C
import os
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
os.environ.setdefault("B200ATTN_LIB", str(ROOT / "vllm-backend-public_b200/lib/libb200attn_trace.so"))
sys.path.insert(0, str(ROOT / "vllm-backend-public_b200"))
import numpy as np
import torch
import b200attn
from b200attn import ops
dev = torch.device("cuda:0")
Hq, Hkv, d, bs = int(os.environ.get("HQ", 32)), int(os.environ.get("HKV", 8)), 128, int(os.environ.get("BS", 16))
chunk, start = int(os.environ.get("CHUNK", 8192)), int(os.environ.get("START", 8192))
print(f"Hq {Hq} Hkv {Hkv} block_size {bs} chunk {chunk} prefix {start}")
nb = (start + chunk) // bs
kc = torch.randn(nb + 8, Hkv, bs, d, device=dev).to(torch.bfloat16)
vc = torch.randn(nb + 8, Hkv, bs, d, device=dev).to(torch.bfloat16)
q = torch.randn(chunk, Hq, d, device=dev).to(torch.bfloat16)
bt = torch.randperm(nb + 8, device=dev)[:nb].to(torch.int32).unsqueeze(0).contiguous()
cu = torch.tensor([0, chunk], dtype=torch.int32, device=dev)
st = torch.tensor([start], dtype=torch.int32, device=dev)
for _ in range(3):
    ops.prefill_attention(q, kc, vc, bt, cu, st)
torch.cuda.synchronize()
lib = b200attn.load()
buf = np.zeros(2 * 32 * 16, dtype=np.uint64)
lib.b2_debug_read_prefill_trace.argtypes = [C.c_void_p]
lib.b2_debug_read_prefill_trace(buf.ctypes.data)
t = buf.reshape(2, 32, 16).astype(np.int64)
t0 = t[:, :, 0][t[:, :, 0] > 0].min()
print("cycles relative to the first stamp; columns: S ready | S in regs | max known | P computed | P handed over")
lo, hi = (int(v) for v in os.environ.get("TRACE_STEPS", "8:24").split(":"))
for j in range(lo, hi):
    for x in (0, 1):
        r = t[x, j, :5] - t0
        i = t[x, j, 5:] - t0
        w = t[x, j, 9:12] - t[x, j, 4]
        print(f"step {j:2d} tile {'AB'[x]}: " + " ".join(f"{v:8d}" for v in r) +
              f"   load {r[1]-r[0]:5d} max {r[2]-r[1]:5d} exp {r[3]-r[2]:5d} hand {r[4]-r[3]:5d}  wait-for-next-S {t[x, j+1, 0]-t[x, j, 4]:6d}"
              f" | issuer: P seen +{i[0]-r[4]:5d} PV issued +{i[1]-i[0]:5d} QK issued +{i[2]-i[1]:5d} -> S ready +{t[x, j+1, 0]-t0-i[2]:5d}"
              f" | warps 1-3 hand over {w[0]:+5d} {w[1]:+5d} {w[2]:+5d} after warp 0" + (f" | V,K waits done {i[3]-r[4]:+6d} vs P handed" if x == 0 else ""))
print("TMA: tile | stage free -> boxes issued -> seen full by the MMA warp (cycles)")
for j in range(lo, hi):
    print(f"  K({j}): free {t[0, j, 12]-t0:7d} issue +{t[0, j, 13]-t[0, j, 12]:4d} landed +{t[0, j, 15]-t[0, j, 12]:5d} full-seen +{t[0, j, 14]-t[0, j, 12]:6d}"
          f"   V({j}): free {t[1, j, 12]-t0:7d} issue +{t[1, j, 13]-t[1, j, 12]:4d} landed +{t[1, j, 15]-t[1, j, 12]:5d} full-seen +{t[1, j, 14]-t[1, j, 12]:6d}")
a = t[0, lo:hi]
print("period (tile A, S ready to next S ready): mean", float(np.diff(a[:, 0]).mean()))
