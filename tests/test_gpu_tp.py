"""GPU (>= 2 devices): the served engine at --tensor-parallel-size 2 (reference benchmark-vllm/models.py:11,
vllm_templates.py:55) -- one process per GPU under torchrun, teacher-forcing parity against the dense CPU oracle.
Skipped on a single-GPU box; bench.py --gpus N runs the same check (`tp_serving.parity`) where the driver sees it."""
import json
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_tp2_engine_matches_oracle(cuda):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29517", str(ROOT / "scripts" / "tp_check.py"), "--legs", "parity"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-3000:])
    rec = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    p = rec["parity"]
    assert p.get("ok") and p["all_finished"], p
    assert p["decode_steps"] > 0 and p["mixed_steps"] > 0
