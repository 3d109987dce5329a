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


def test_server_tensor_parallel_size_2(cuda):
    """The launch vector the reference renders with tensor_parallel_size: 2 -> rank 0 serves HTTP and spawns rank 1;
    greedy text equals what the same server produces at tensor_parallel_size 1 would only hold up to bf16 near-ties,
    so the check here is the contract: health, a completion of the asked length, repeatability, clean shutdown."""
    import os
    import socket
    import time
    import requests
    from b200attn.engine_args import EngineArgs
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    args = EngineArgs(model_name="tiny8", gpu_memory_utilization=0.0, max_num_seqs=8, block_size=16, max_model_len=1024,
                      tensor_parallel_size=2, gpu_resource_count=2, served_model_name="b200-tiny8", port=port,
                      host="127.0.0.1", additional_args={"max-num-batched-tokens": 256, "num-gpu-blocks-override": 600})
    env = dict(os.environ, PYTHONPATH=str(ROOT / "vllm-backend-public_b200"))
    proc = subprocess.Popen([sys.executable, "-m", "b200attn.server"] + args.to_cli_args(), env=env,
                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    base = f"http://127.0.0.1:{port}"
    try:
        for _ in range(900):
            if proc.poll() is not None:
                raise RuntimeError("server exited: " + proc.stdout.read().decode()[-3000:])
            try:
                if requests.get(base + "/health", timeout=1).status_code == 200:
                    break
            except requests.RequestException:
                pass
            time.sleep(0.2)
        else:
            raise RuntimeError("server did not become healthy")
        body = {"prompt": list(range(1, 300)), "max_tokens": 20, "temperature": 0}
        a = requests.post(base + "/v1/completions", json=body, timeout=120).json()
        b = requests.post(base + "/v1/completions", json=body, timeout=120).json()
        assert a["usage"]["completion_tokens"] == 20 and a["choices"][0]["text"] == b["choices"][0]["text"]
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(8) as ex:
            outs = list(ex.map(lambda i: requests.post(base + "/v1/completions", json={
                "prompt": [(5 * i + j) % 2048 for j in range(10 + 60 * i)], "max_tokens": 12, "temperature": 0.8,
                "seed": i}, timeout=120).json(), range(8)))
        assert all(o["usage"]["completion_tokens"] == 12 for o in outs)
        # SIGTERM (what Kubernetes sends the pod): rank 0 must run its orderly shutdown -- tell rank 1 to stop, close
        # the peer-mapped buffers together, exit 0 -- and no rank may survive it.  (uvicorn re-raises the signal after
        # draining; without a handler of our own the process died there and rank 1 stayed behind with its GPU spinning
        # inside a barrier.)
        import psutil
        kids = psutil.Process(proc.pid).children(recursive=True)
        assert any("b200attn.tp" in " ".join(k.cmdline()) for k in kids), "rank 1 is not a child of the server"
        proc.terminate()
        assert proc.wait(40) == 0, "server did not exit 0 on SIGTERM"
        gone, alive = psutil.wait_procs(kids, timeout=10)
        assert not alive, f"tensor-parallel rank processes survived the server: {[k.pid for k in alive]}"
    finally:
        if proc.poll() is None:
            proc.kill()
            proc.wait(10)


def test_tp_worker_dies_with_a_killed_server(cuda):
    """SIGKILL of rank 0 (no shutdown code runs at all): rank 1 must not stay behind holding its GPU -- the
    watcher thread of b200attn.tp.worker_main sees the launcher gone and leaves."""
    import os
    import socket
    import time
    import psutil
    import requests
    from b200attn.engine_args import EngineArgs
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    args = EngineArgs(model_name="tiny8", gpu_memory_utilization=0.0, max_num_seqs=8, block_size=16, max_model_len=1024,
                      tensor_parallel_size=2, gpu_resource_count=2, served_model_name="b200-tiny8", port=port,
                      host="127.0.0.1", additional_args={"max-num-batched-tokens": 256, "num-gpu-blocks-override": 600})
    env = dict(os.environ, PYTHONPATH=str(ROOT / "vllm-backend-public_b200"))
    proc = subprocess.Popen([sys.executable, "-m", "b200attn.server"] + args.to_cli_args(), env=env,
                            stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    kids = []
    try:
        for _ in range(900):
            assert proc.poll() is None, "server exited during start-up"
            try:
                if requests.get(f"http://127.0.0.1:{port}/health", timeout=1).status_code == 200:
                    break
            except requests.RequestException:
                pass
            time.sleep(0.2)
        else:
            raise RuntimeError("server did not become healthy")
        kids = psutil.Process(proc.pid).children(recursive=True)
        assert kids
        proc.kill()
        proc.wait(10)
        gone, alive = psutil.wait_procs(kids, timeout=15)
        assert not alive, f"rank processes survived a killed server: {[k.pid for k in alive]}"
    finally:
        if proc.poll() is None:
            proc.kill()
        for k in kids:
            if k.is_running():
                k.kill()
