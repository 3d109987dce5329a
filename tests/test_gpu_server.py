"""GPU: the real engine process behind the reference's launch vector, driven over HTTP (boundary b1 end to end):
`python -m b200attn.server <argv the reference renders>` -> /health -> /v1/models -> /v1/completions -> load generator
-> EvaluationPayload."""
import json
import os
import socket
import subprocess
import sys
import time
from pathlib import Path

import pytest
import requests

from b200attn import loadgen
from b200attn.engine_args import EngineArgs

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.fixture(scope="module")
def server(cuda):
    port = _free_port()
    # the vector the reference's template would render for this config (CPU-era flags included)
    args = EngineArgs(model_name="tiny", gpu_memory_utilization=0.0, max_num_seqs=8, block_size=16, dtype="float32",
                      max_model_len=1024, served_model_name="b200-tiny", port=port, host="127.0.0.1",
                      additional_args={"disable-log-stats": True, "enforce-eager": True, "max-num-batched-tokens": 256,
                                       "num-gpu-blocks-override": 600})
    env = dict(os.environ, PYTHONPATH=str(ROOT / "vllm-backend-public_b200"))
    proc = subprocess.Popen([sys.executable, "-m", "b200attn.server"] + args.to_cli_args(), env=env,
                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    base = f"http://127.0.0.1:{port}"
    try:
        for _ in range(600):
            if proc.poll() is not None:
                raise RuntimeError("server exited: " + proc.stdout.read().decode()[-2000:])
            try:
                if requests.get(base + "/health", timeout=1).status_code == 200:
                    break
            except requests.RequestException:
                pass
            time.sleep(0.2)
        else:
            raise RuntimeError("server did not become healthy")
        yield base
    finally:
        proc.terminate()
        try:
            rc = proc.wait(20)
        except subprocess.TimeoutExpired:
            proc.kill()
            rc = None
    # SIGTERM must end in the server's own shutdown path (exit code 0), not in the re-raised signal (-15)
    assert rc == 0, f"server exit code {rc} after SIGTERM"


def test_models_and_completion(server):
    d = requests.get(server + "/v1/models").json()
    assert d["data"][0]["id"] == "b200-tiny" and d["data"][0]["max_model_len"] == 1024
    body = {"model": "b200-tiny", "prompt": "The quick brown fox", "max_tokens": 12, "temperature": 0}
    a = requests.post(server + "/v1/completions", json=body).json()
    b = requests.post(server + "/v1/completions", json=body).json()
    assert a["usage"] == {"prompt_tokens": 19, "completion_tokens": 12, "total_tokens": 31}
    assert a["choices"][0]["text"] == b["choices"][0]["text"] and a["choices"][0]["finish_reason"] == "length"
    # streaming returns the same greedy text
    with requests.post(server + "/v1/completions", json=dict(body, stream=True), stream=True) as r:
        events = [l[6:] for l in r.iter_lines(decode_unicode=True) if l and l.startswith("data: ")]
    assert events[-1] == "[DONE]"
    text = "".join(c["text"] for e in events[:-1] for c in json.loads(e)["choices"])
    assert text == a["choices"][0]["text"]


def test_token_prompts_concurrent(server):
    from concurrent.futures import ThreadPoolExecutor

    def one(i):
        body = {"prompt": [(7 * i + j) % 1024 for j in range(20 + 37 * i)], "max_tokens": 16, "temperature": 0.7, "seed": i}
        return requests.post(server + "/v1/completions", json=body).json()

    with ThreadPoolExecutor(16) as ex:
        outs = list(ex.map(one, range(16)))
    assert all(o["usage"]["completion_tokens"] == 16 for o in outs)
    again = one(3)
    assert again["choices"][0]["text"] == outs[3]["choices"][0]["text"]          # seeded sampling repeats


def test_context_limit_and_errors(server):
    r = requests.post(server + "/v1/completions", json={"prompt": [1] * 1024, "max_tokens": 1})
    assert r.status_code == 400 and r.json()["error"]["code"] == "context_length_exceeded"
    r = requests.post(server + "/v1/completions", json={"prompt": [1] * 1000, "max_tokens": 100}).json()
    assert r["usage"]["completion_tokens"] == 24 and r["choices"][0]["finish_reason"] == "length"
    assert requests.post(server + "/v1/completions", json={"prompt": "x", "model": "nope"}).status_code == 404


def test_loadgen_against_engine(server, tmp_path):
    out = tmp_path / "res.json"
    rc = loadgen.main(["--endpoint", server + "/v1/completions", "--num-requests", "24", "--concurrency", "8",
                       "--min-prompt", "16", "--max-prompt", "900", "--max-tokens", "16", "--out", str(out),
                       "--config", "/nonexistent", "--ready-delay", "0.1"])
    assert rc == 0
    res = json.loads(out.read_text())
    m = res["standardized"]["data"]["metrics"]
    assert m["num_ok"] == 24 and m["completion_tokens"] == 24 * 16 and m["output_tok_per_s"] > 0
    for f in loadgen.REQUIRED_PAYLOAD_FIELDS:
        assert f in res["raw"] and f in res["standardized"]
    met = requests.get(server + "/metrics").text
    assert "b200attn_generation_tokens_total" in met
