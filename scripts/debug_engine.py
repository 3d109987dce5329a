import random, sys, torch
sys.path[:0] = ["vllm-backend-public_b200", "oracle", "tests"]
from b200attn.engine import LLMEngine, SamplingParams
from b200attn.model import REGISTRY
import layer_ref as L

def gaps(eng, r):
    c = eng.cfg
    w = {k: v.cpu() for k, v in eng.model.w.items()}
    ids = r.prompt_ids + r.output_ids
    logits = L.llama_forward_ref(w, ids[:-1], layers=c.layers, Hq=c.Hq, Hkv=c.Hkv, d=c.d, eps=c.eps, cos_sin=eng.model.cos_sin.cpu())
    rows = logits[len(r.prompt_ids) - 1:]
    picked = rows.gather(1, torch.tensor(r.output_ids).unsqueeze(1)).squeeze(1)
    return [round(float(x), 3) for x in (rows.max(dim=1).values - picked)]

def run(lens, eager, **kw):
    args = dict(max_num_seqs=8, max_model_len=1024, block_size=16, max_num_batched_tokens=256, seed=1, num_blocks=600, enforce_eager=eager)
    args.update(kw)
    eng = LLMEngine(REGISTRY["tiny"], torch.device("cuda:0"), **args)
    rng = random.Random(0)
    prompts = [[rng.randrange(1024) for _ in range(n)] for n in lens]
    reqs = [eng.add_request(p, SamplingParams(max_tokens=24, temperature=0.0)) for p in prompts]
    eng.run_until_idle()
    print("lens", lens, "eager", eager, kw, "steps d/m", eng.stats.decode_steps, eng.stats.mixed_steps)
    for r in reqs:
        g = gaps(eng, r)
        print("  P=%d max gap %.3f" % (len(r.prompt_ids), max(g)), g if max(g) > 0.03 else "")

run([1], True)
run([1], False)
run([40], True)
run([1, 5], True)
run([1, 5, 16, 17], True)
run([300], True)
run([1, 5, 16, 17, 100, 300], True)
