"""BASELINE.json configs[4]: random-init Llama-3-8B behind the OpenAI endpoint on port 8005, driven by the load generator
(mixed prompt lengths 128-32k, log-uniform, seed 0).  Usage (on the GPU box):
    python scripts/run_e2e.py [--num-requests 256] [--concurrency 64] [--max-tokens 128] [--model llama-3.1-8b] [--tag x]
The client is the evaluation Job of the reference's loop: b200attn/templates/evaluate.yaml rendered by the
process_template rule (reference benchmark-eval/main.py:55-58,113) for this model and engine URL, and the Job's
container command run locally (b200attn.evaljob) -- no Kubernetes on the box.  --tensor-parallel-size N serves on N GPUs.
Writes gpurun_out/e2e_<tag>.json (raw + standardized EvaluationPayload) and prints the standardized metrics."""
import argparse
import os
import subprocess
import sys
import time
from pathlib import Path

import requests

ROOT = Path(__file__).resolve().parent.parent
ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-3.1-8b")
ap.add_argument("--num-requests", type=int, default=256)
ap.add_argument("--concurrency", type=int, default=64)
ap.add_argument("--min-prompt", type=int, default=128)
ap.add_argument("--max-prompt", type=int, default=32768)
ap.add_argument("--max-tokens", type=int, default=128)
ap.add_argument("--max-num-batched-tokens", type=int, default=8192)
ap.add_argument("--port", type=int, default=8005)
ap.add_argument("--tensor-parallel-size", type=int, default=1)
ap.add_argument("--tag", default="r2")
a = ap.parse_args()
sys.path.insert(0, str(ROOT / "vllm-backend-public_b200"))
from b200attn import evaljob  # noqa: E402

env = dict(os.environ, PYTHONPATH=str(ROOT / "vllm-backend-public_b200"))
max_len = a.max_prompt + a.max_tokens + 16
srv = subprocess.Popen([sys.executable, "-m", "b200attn.server", "--model", a.model, "--port", str(a.port), "--host", "127.0.0.1",
                        "--max-num-seqs", str(a.concurrency), "--max-model-len", str(max_len), "--block-size", "16",
                        "--gpu-memory-utilization", "0.9", "--dtype", "bfloat16", "--served-model-name", a.model.replace("_", "-").replace(".", "-"),   # the name the rendered Job asks for
                        "--max-num-batched-tokens", str(a.max_num_batched_tokens),
                        "--tensor-parallel-size", str(a.tensor_parallel_size)], env=env)
base = f"http://127.0.0.1:{a.port}"
try:
    t0 = time.time()
    while True:
        if srv.poll() is not None:
            sys.exit("server exited early")
        try:
            if requests.get(base + "/health", timeout=1).status_code == 200:
                break
        except requests.RequestException:
            pass
        if time.time() - t0 > 600:
            sys.exit("server not healthy after 600 s")
        time.sleep(0.5)
    print(f"server healthy after {time.time() - t0:.1f}s", flush=True)
    out = ROOT / "gpurun_out" / f"e2e_{a.tag}.json"
    out.parent.mkdir(exist_ok=True)
    # the evaluation Job: template -> process_template -> YAML check -> run its container command here
    text, job = evaljob.render_for_model(a.model, base)
    (out.parent / f"e2e_{a.tag}_job.yaml").write_text(text)
    print("job", job["metadata"]["name"], "->", evaljob.job_command(job)[1]["VLLM_MODEL_ENDPOINT"], flush=True)
    rc = evaljob.run_job_locally(job, extra_args=["--num-requests", str(a.num_requests), "--concurrency", str(a.concurrency),
                                                  "--min-prompt", str(a.min_prompt), "--max-prompt", str(a.max_prompt),
                                                  "--max-tokens", str(a.max_tokens), "--out", str(out),
                                                  "--config", "/nonexistent"])
    print(requests.get(base + "/metrics").text)
    sys.exit(rc)
finally:
    srv.terminate()
    try:
        srv.wait(15)
    except subprocess.TimeoutExpired:
        srv.kill()
