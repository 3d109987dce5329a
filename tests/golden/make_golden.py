"""Regenerates tests/golden/config1.json + config1_out.pt from the oracle (BASELINE.json configs[0]:
8 q / 2 kv heads, head_dim 128, block_size 16, batch 4, ctx 512).  The reference repository has no
implementation or vectors for this path, so the fixture is the oracle's own frozen output; it is
cross-checked on the GPU box against the vLLM wheel's paged_attention_v1 by
tests/golden/make_vllm_crosscheck.py when that wheel loads there."""
import hashlib
import json
import sys
from pathlib import Path

import torch

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent / "oracle"))
import paged_attention_ref as R  # noqa: E402

shape = dict(B=4, Hq=8, Hkv=2, d=128, bs=16, ctx=512)
seed = 0
inp = R.make_inputs(**shape, seed=seed)
kc = torch.zeros(inp["num_blocks"], 2, 16, 128, dtype=torch.bfloat16)
vc = torch.zeros_like(kc)
R.reshape_and_cache_ref(inp["k"], inp["v"], kc, vc, inp["slot_mapping"])
out = R.paged_decode_ref(inp["q"], kc, vc, inp["block_tables"], inp["seq_lens"])


def sha(t):
    return hashlib.sha256(t.contiguous().view(torch.uint8).numpy().tobytes()).hexdigest()


meta = dict(shape=shape, seed=seed, torch=torch.__version__,
            sha256=dict(q=sha(inp["q"]), k_cache=sha(kc), v_cache=sha(vc), block_tables=sha(inp["block_tables"]),
                        out=sha(out)))
(HERE / "config1.json").write_text(json.dumps(meta, indent=1))
torch.save(out, HERE / "config1_out.pt")
print(json.dumps(meta, indent=1))
