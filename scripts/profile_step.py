"""Where the host-side time of DecodeStep.run goes (config 2): plan / stream context / graph replay / sync."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "vllm-backend-public_b200")); sys.path.insert(0, str(ROOT))
import torch
import b200attn
dev = torch.device("cuda:0")
B, Hq, Hkv, d, bs, ctx = 64, 32, 8, 128, 16, 4096
nb = ctx // bs
kc = torch.zeros(B * nb + 64, Hkv, bs, d, dtype=torch.bfloat16, device=dev)
vc = torch.zeros_like(kc)
step = b200attn.DecodeStep(kc, vc, max_batch=B, Hq=Hq, max_nb=nb)
step.h_block_tables.copy_(torch.randperm(B * nb)[: B * nb].view(B, nb).to(torch.int32))
step.h_seq_lens.fill_(ctx)
step.h_slots.copy_((step.h_block_tables[:, -1].long() * bs + bs - 1))
for _ in range(20):
    step.run(B)
N = 2000
def T(fn):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(N): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / N * 1e6
print("run() total us", T(lambda: step.run(B)))
print("run(splits=1) us", T(lambda: step.run(B, 1)))
print("plan us", T(lambda: step.plan(B)))
g = step._graphs[(B, 1)]
def ctx_only():
    with torch.cuda.stream(step._stream): pass
print("stream ctx us", T(ctx_only))
def replay_sync():
    with torch.cuda.stream(step._stream): g.replay()
    step._stream.synchronize()
print("ctx+replay+sync us", T(replay_sync))
def replay_nosync():
    g.replay()
print("replay on current stream, sync at end: us/iter", T(replay_nosync))
def replay_cur_sync():
    g.replay(); torch.cuda.current_stream().synchronize()
print("replay current + sync us", T(replay_cur_sync))
