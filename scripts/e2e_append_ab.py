"""DecodeStep (the e2e path of bench.py, config 2, zero-copy pinned inputs) with the fused plan call against the two-call
step: wall time of run() and the GPU time of back-to-back graph replays, alternating on one box."""
import os, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "vllm-backend-public_b200")); sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "oracle"))
import torch
import b200attn
import bench

dev = torch.device("cuda:0")
wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "llama3-8b_b64_ctx4k_bf16kv"]
zero_copy = (sys.argv[2] if len(sys.argv) > 2 else "1") == "1"
prob = bench.build_problem(dev, wl["B"], wl["Hq"], wl["Hkv"], wl["ctx"], wl["fp8"], seed=0)
B, Hq, Hkv = wl["B"], wl["Hq"], wl["Hkv"]
ks, vs = prob["scales"] or (None, None)


def make(fuse):
    os.environ["B2_STEP_FUSE_APPEND"] = "1" if fuse else "0"
    st = b200attn.DecodeStep(prob["k_cache"], prob["v_cache"], max_batch=B, Hq=Hq, max_nb=prob["max_nb"], k_scale=ks,
                             v_scale=vs, zero_copy_inputs=zero_copy)
    st.h_qkv[:, :Hq].copy_(prob["q"].cpu()); st.h_qkv[:, Hq:Hq + Hkv].copy_(prob["k_new"].cpu())
    st.h_qkv[:, Hq + Hkv:].copy_(prob["v_new"].cpu())
    st.h_block_tables.copy_(prob["block_tables"]); st.upload_block_tables()
    st.h_seq_lens.copy_(prob["seq_lens"]); st.h_slots.copy_(prob["slot_new"])
    for _ in range(10):
        st.run(B)
    return st


steps = {1: make(True), 0: make(False)}
assert torch.equal(steps[1].h_out.view(torch.int16), steps[0].h_out.view(torch.int16)), "fused e2e result differs"
N = 1000
for rep in range(3):
    for f in (1, 0):
        st = steps[f]
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(N):
            st.run(B)
        wall = (time.perf_counter() - t0) / N * 1e6
        (key, g), = [(k, v) for k, v in st._graphs.items()][:1]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(N):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        print(f"fuse_append={f} splits={key[1]} run() wall {wall:.2f} us/step, back-to-back graph replays {e0.elapsed_time(e1) / N * 1e3:.2f} us/step", flush=True)
