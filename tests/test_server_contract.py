"""HTTP contract of the engine process (boundary b1) on CPU: the FastAPI app in front of a STUB engine (test double with
the LLMEngine interface: the real engine needs a GPU and is exercised by tests/test_gpu_server.py), the load generator
against that app over real sockets, and its EvaluationPayload against the reference's schema
(tests/golden/reference_contract.json, generated from reference benchmark-results/models.py)."""
import asyncio
import json
import socket
import threading
import time
from pathlib import Path
from types import SimpleNamespace

import pytest

from b200attn import loadgen
from b200attn.engine_args import EngineArgs
from b200attn.server import EngineRunner, create_app
from b200attn.tokenizer import ByteTokenizer

FIX = json.loads((Path(__file__).parent / "golden" / "reference_contract.json").read_text())


class StubEngine:
    """Emits the alphabet, one token per request per step, through the engine's callback interface."""

    def __init__(self, max_model_len=64, vocab=1024):
        from b200attn.engine import Request, EngineStats                    # dataclasses only (no CUDA at import)
        self._Request = Request
        self.max_model_len, self.cfg = max_model_len, SimpleNamespace(vocab=vocab)
        self.block_size, self.num_blocks = 16, 100
        self.requests, self._lock, self._next = {}, threading.Lock(), 1
        self.stats = EngineStats()
        self.sched = SimpleNamespace(num_running=0, num_waiting=0, num_free_blocks=100)

    def add_request(self, prompt_ids, params, on_token=None):
        with self._lock:
            r = self._Request(self._next, list(prompt_ids), params, on_token=on_token)
            self._next += 1
            self.requests[r.id] = r
        return r

    def abort(self, rid):
        with self._lock:
            r = self.requests.pop(rid, None)
        if r is not None and not r.finished:
            r.finished, r.finish_reason = True, "abort"
            r.on_token(r, None)

    def has_work(self):
        return bool(self.requests)

    def step(self):
        with self._lock:
            live = list(self.requests.values())
        for r in live:
            tok = ord("a") + (len(r.output_ids) + len(r.prompt_ids)) % 26
            r.output_ids.append(tok)
            self.stats.generated_tokens += 1
            r.on_token(r, tok)
            if len(r.output_ids) >= r.params.max_tokens or len(r.prompt_ids) + len(r.output_ids) >= self.max_model_len:
                r.finished, r.finish_reason = True, "length"
                with self._lock:
                    self.requests.pop(r.id, None)
                r.on_token(r, None)
        time.sleep(0.001)


class StubRunner(EngineRunner):
    def _main(self):
        self.engine = StubEngine()
        self.tokenizer = ByteTokenizer(1024)
        self.ready.set()
        while not self._stop:
            if self.engine.has_work():
                self.engine.step()
                self._flush_events()            # the real engine calls this through its after_step hook
            else:
                self._wake.wait(0.01)
                self._wake.clear()


@pytest.fixture(scope="module")
def client():
    from fastapi.testclient import TestClient
    args = EngineArgs.from_cli(["--model", "tiny", "--served-model-name", "stub-model", "--port", "8005"])
    runner = StubRunner(args)
    runner.start()
    runner.ready.wait(5)
    with TestClient(create_app(runner)) as c:
        yield c
    runner.stop()


def test_health_and_models(client):
    assert client.get("/health").status_code == 200
    r = client.get("/v1/models")
    assert r.status_code == 200
    d = r.json()
    assert d["object"] == "list" and d["data"][0]["id"] == "stub-model" and d["data"][0]["root"] == "tiny"


def test_health_is_503_until_ready():
    from fastapi.testclient import TestClient

    class Never(EngineRunner):
        def _main(self):
            time.sleep(0.2)

    runner = Never(EngineArgs.from_cli(["--model", "tiny"]))
    with TestClient(create_app(runner)) as c:
        assert c.get("/health").status_code == 503
        assert c.get("/v1/models").status_code == 503
        assert c.post("/v1/completions", json={"prompt": "x"}).status_code == 503


def test_completion_text_and_usage(client):
    r = client.post("/v1/completions", json={"model": "stub-model", "prompt": "hello", "max_tokens": 5, "temperature": 0})
    assert r.status_code == 200
    d = r.json()
    assert d["object"] == "text_completion" and d["model"] == "stub-model" and d["id"].startswith("cmpl-")
    ch = d["choices"][0]
    assert ch["index"] == 0 and ch["finish_reason"] == "length" and len(ch["text"]) == 5
    assert d["usage"] == {"prompt_tokens": 5, "completion_tokens": 5, "total_tokens": 10}


def test_prompt_forms_n_echo_and_stop(client):
    r = client.post("/v1/completions", json={"prompt": [1, 2, 3], "max_tokens": 3}).json()
    assert r["usage"]["prompt_tokens"] == 3 and len(r["choices"]) == 1
    r = client.post("/v1/completions", json={"prompt": ["ab", "cde"], "max_tokens": 2, "n": 2}).json()
    assert [c["index"] for c in r["choices"]] == [0, 1, 2, 3] and r["usage"]["prompt_tokens"] == 5
    r = client.post("/v1/completions", json={"prompt": [[1, 2], [3]], "max_tokens": 2}).json()
    assert len(r["choices"]) == 2
    r = client.post("/v1/completions", json={"prompt": "xy", "max_tokens": 2, "echo": True}).json()
    assert r["choices"][0]["text"].startswith("xy") and len(r["choices"][0]["text"]) == 4
    # the stub emits c d e f ... after a 2-token prompt: stop at "ef"
    r = client.post("/v1/completions", json={"prompt": "xy", "max_tokens": 20, "stop": ["ef"]}).json()
    assert r["choices"][0]["text"] == "cd" and r["choices"][0]["finish_reason"] == "stop"
    # context limit of the stub is 64 tokens: generation stops by length there
    r = client.post("/v1/completions", json={"prompt": "x" * 60, "max_tokens": 100}).json()
    assert r["usage"]["completion_tokens"] == 4 and r["choices"][0]["finish_reason"] == "length"


def test_streaming_sse(client):
    with client.stream("POST", "/v1/completions", json={"prompt": "hi", "max_tokens": 6, "stream": True,
                                                        "stream_options": {"include_usage": True}}) as r:
        assert r.status_code == 200 and r.headers["content-type"].startswith("text/event-stream")
        events = [l[6:] for l in r.iter_lines() if l.startswith("data: ")]
    assert events[-1] == "[DONE]"
    objs = [json.loads(e) for e in events[:-1]]
    text = "".join(c["text"] for o in objs for c in o["choices"])
    assert len(text) == 6
    assert [c["finish_reason"] for o in objs for c in o["choices"]][-1] == "length"
    assert objs[-1]["usage"]["completion_tokens"] == 6 and objs[-1]["choices"] == []


def test_chat_completions(client):
    r = client.post("/v1/chat/completions", json={"messages": [{"role": "user", "content": "hi"}], "max_tokens": 4})
    d = r.json()
    assert r.status_code == 200 and d["object"] == "chat.completion"
    assert d["choices"][0]["message"]["role"] == "assistant" and len(d["choices"][0]["message"]["content"]) == 4


@pytest.mark.parametrize("body,status,param", [
    ({}, 400, "prompt"),
    ({"prompt": ""}, 400, "prompt"),
    ({"prompt": []}, 400, "prompt"),
    ({"prompt": [1, "a"]}, 400, "prompt"),
    ({"prompt": "x", "max_tokens": 0}, 400, "max_tokens"),
    ({"prompt": "x", "temperature": -1}, 400, "temperature"),
    ({"prompt": "x", "top_p": 0.0}, 400, "top_p"),
    ({"prompt": "x", "top_k": "many"}, 400, "top_k"),
    ({"prompt": "x", "n": 0}, 400, "n"),
    ({"prompt": "x", "stop": 5}, 400, "stop"),
    ({"prompt": "x" * 64}, 400, "prompt"),
    ({"prompt": [5000]}, 400, "prompt"),
    ({"prompt": "x", "model": "other"}, 404, "model"),
])
def test_errors_use_openai_envelope(client, body, status, param):
    r = client.post("/v1/completions", json=body)
    assert r.status_code == status
    e = r.json()["error"]
    assert set(e) == {"message", "type", "param", "code"} and e["param"] == param


def test_invalid_json(client):
    r = client.post("/v1/completions", content=b"{nope", headers={"content-type": "application/json"})
    assert r.status_code == 400 and "error" in r.json()


# ---------------------------------------------------------------------------------------------
# load generator over real sockets + EvaluationPayload schema
# ---------------------------------------------------------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _serve(app, port):
    import uvicorn
    server = uvicorn.Server(uvicorn.Config(app, host="127.0.0.1", port=port, log_level="error"))
    t = threading.Thread(target=server.run, daemon=True)
    t.start()
    for _ in range(100):
        if server.started:
            break
        time.sleep(0.05)
    return server, t


def _check_payload(p):
    schema = FIX["evaluation_payload_schema"]
    assert set(schema["required"]) == set(loadgen.REQUIRED_PAYLOAD_FIELDS)
    for f in schema["required"]:
        assert f in p
    assert set(p) <= set(schema["properties"])
    for f in ("benchmark_name", "timestamp", "model_id", "tokenizer_id", "source"):
        assert isinstance(p[f], str)


def test_loadgen_end_to_end_with_results_service(tmp_path):
    from fastapi import FastAPI, Request
    runner = StubRunner(EngineArgs.from_cli(["--model", "tiny", "--served-model-name", "stub-model"]))
    runner.start()
    runner.ready.wait(5)
    eport, rport = _free_port(), _free_port()
    eserver, _ = _serve(create_app(runner), eport)
    stored = {}
    results = FastAPI()

    @results.post("/raw_input")
    async def raw(request: Request):                       # the storage rule of reference results_manager.py:18-52
        p = await request.json()
        _check_payload(p)
        stored["raw:" + loadgen.payload_pk(p)] = p
        return {"status": "success", "run_id": p["run_id"], "saved_as": loadgen.payload_pk(p)}

    @results.post("/standardized_output")
    async def std(request: Request):
        p = await request.json()
        _check_payload(p)
        stored["std:" + loadgen.payload_pk(p)] = p
        return {"status": "success", "run_id": p["run_id"], "saved_as": loadgen.payload_pk(p)}

    rserver, _ = _serve(results, rport)
    out = tmp_path / "res.json"
    rc = loadgen.main(["--endpoint", f"http://127.0.0.1:{eport}/v1/completions", "--num-requests", "12", "--concurrency", "4",
                       "--min-prompt", "4", "--max-prompt", "40", "--max-tokens", "8", "--results-url",
                       f"http://127.0.0.1:{rport}", "--out", str(out), "--config", "/nonexistent", "--ready-delay", "0.1"])
    eserver.should_exit = rserver.should_exit = True
    runner.stop()
    assert rc == 0
    res = json.loads(out.read_text())
    m = res["standardized"]["data"]["metrics"]
    assert m["num_ok"] == 12 and m["num_failed"] == 0 and m["completion_tokens"] == 12 * 8
    assert m["ttft_s"]["p50"] > 0 and m["output_tok_per_s"] > 0
    assert res["posted"]["/raw_input"]["status"] == 200 and res["posted"]["/standardized_output"]["status"] == 200
    assert len(stored) == 2
    _check_payload(res["raw"])
    assert len(res["raw"]["data"]["requests"]) == 12


def test_payload_pk_matches_reference_rule():
    ex = FIX["evaluation_payload_pk_example"]
    assert loadgen.payload_pk(ex["payload"]) == ex["pk"]


def test_prompt_lengths_are_log_uniform_and_seeded():
    a = loadgen.prompt_lengths(2000, 128, 32768, seed=0)
    assert a == loadgen.prompt_lengths(2000, 128, 32768, seed=0) and min(a) >= 128 and max(a) <= 32768
    import math
    below = sum(1 for x in a if x < math.sqrt(128 * 32768)) / len(a)
    assert 0.45 < below < 0.55                              # the geometric midpoint splits a log-uniform draw in half


def test_process_template_placeholders():
    """The eval template placeholders benchmark-eval fills (reference benchmark-eval/main.py:55-58) end up as the
    endpoint the load generator reads from VLLM_MODEL_ENDPOINT."""
    pt = FIX["process_template"]
    for model, url, rendered in pt["cases"]:
        ours = pt["template"].replace("{model_name}", model).replace("{inference_engine_url}", url) \
            .replace("{model_name_lower}", model.lower())
        assert ours == rendered
        import yaml
        env = yaml.safe_load(rendered)["spec"]["template"]["spec"]["containers"][0]["env"]
        ep = next(e["value"] for e in env if e["name"] == "VLLM_MODEL_ENDPOINT")
        assert ep == url + "/v1/completions" and ep.split("/v1/")[0] == url


def test_chat_uses_the_models_own_template_when_the_tokenizer_ships_one(tmp_path):
    """/v1/chat/completions must prompt instruct weights with THEIR chat template (tokenizer_config.json),
    not a hard-coded format; the byte tokenizer (random-init weights) has none and keeps the plain fallback."""
    import json
    from tokenizers import Tokenizer, models, pre_tokenizers
    from b200attn.tokenizer import ByteTokenizer, HFTokenizer
    vocab = {"<s>": 0, "</s>": 1, "[INST]": 2, "[/INST]": 3, "hi": 4, "there": 5, "<unk>": 6}
    tok = Tokenizer(models.WordLevel(vocab, unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.WhitespaceSplit()
    tok.save(str(tmp_path / "tokenizer.json"))
    (tmp_path / "tokenizer_config.json").write_text(json.dumps({
        "bos_token": {"content": "<s>"}, "eos_token": "</s>",
        "chat_template": "{{ bos_token }} {% for m in messages %}{% if m['role'] == 'user' %}[INST] {{ m['content'] }} [/INST] "
                         "{% else %}{{ m['content'] }} {{ eos_token }} {% endif %}{% endfor %}"}))
    t = HFTokenizer(tmp_path)
    text = t.apply_chat_template([{"role": "user", "content": "hi"}, {"role": "assistant", "content": "there"},
                                  {"role": "user", "content": "hi there"}])
    assert t.encode_rendered(text) == [0, 2, 4, 3, 5, 1, 2, 4, 5, 3]
    assert not hasattr(ByteTokenizer(300), "apply_chat_template")


def test_evaljob_renders_like_the_reference_and_describes_a_runnable_job():
    """b200attn.evaljob.process_template against the outputs of the reference's own function
    (tests/golden/reference_contract.json, generated from benchmark-eval/main.py:55-58), and the committed
    evaluate.yaml renders to a valid Job whose container is the load generator pointed at the engine."""
    import json
    from pathlib import Path
    from b200attn import evaljob
    fx = json.loads((Path(__file__).parent / "golden" / "reference_contract.json").read_text())["process_template"]
    for model, url, rendered in fx["cases"]:
        assert evaljob.process_template(fx["template"], model, url) == rendered
    text, job = evaljob.render_for_model("B200_Llama-3.1-8B", "http://127.0.0.1:8005")
    assert "{" not in text.split("spec:", 1)[1].replace('["python", "-m", "b200attn.loadgen"]', "")   # no placeholder left
    assert job["metadata"]["name"] == "eval-b200-llama-3-1-8b"
    argv, env = evaljob.job_command(job)
    assert argv[:3] == ["python", "-m", "b200attn.loadgen"] and argv[argv.index("--model") + 1] == "B200-Llama-3-1-8B"
    assert env["VLLM_MODEL_ENDPOINT"] == "http://127.0.0.1:8005/v1/completions"
    import pytest
    with pytest.raises(ValueError):
        evaljob.render_for_model("m", "u", template="kind: [unclosed")


_SIGTERM_CHILD = r'''
import sys, threading
from b200attn import server
from b200attn.tokenizer import ByteTokenizer

class Runner(server.EngineRunner):
    """No engine: ready at once, and stop() leaves a mark -- what rank 0 must reach to release its worker ranks."""
    def _main(self):
        self.tokenizer = ByteTokenizer(1024)
        self.ready.set()
        while not self._stop:
            self._wake.wait(0.05)
    def stop(self):
        super().stop()
        open(sys.argv[2], "w").write("stopped")

server.EngineRunner = Runner
sys.exit(server.main(["--model", "tiny", "--host", "127.0.0.1", "--port", sys.argv[1]]))
'''


def test_sigterm_runs_the_shutdown_path(tmp_path):
    """The pod's SIGTERM must reach EngineRunner.stop() and exit 0.  uvicorn re-raises the signal once its loop has
    drained; with the default handler the process died inside server.run() and a tensor-parallel rank 0 never told
    ranks 1..N-1 to stop (they stayed behind spinning inside a barrier)."""
    import os
    import signal
    import subprocess
    import sys
    import requests
    port = _free_port()
    mark = tmp_path / "stopped"
    env = dict(os.environ, PYTHONPATH=str(Path(__file__).resolve().parent.parent / "vllm-backend-public_b200"))
    proc = subprocess.Popen([sys.executable, "-c", _SIGTERM_CHILD, str(port), str(mark)], env=env,
                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    try:
        for _ in range(300):
            assert proc.poll() is None, proc.stdout.read().decode()[-2000:]
            try:
                if requests.get(f"http://127.0.0.1:{port}/health", timeout=1).status_code == 200:
                    break
            except requests.RequestException:
                pass
            time.sleep(0.1)
        else:
            raise AssertionError("server did not become healthy")
        proc.send_signal(signal.SIGTERM)
        rc = proc.wait(20)
    finally:
        if proc.poll() is None:
            proc.kill()
    assert rc == 0, f"exit code {rc}: {proc.stdout.read().decode()[-2000:]}"
    assert mark.read_text() == "stopped"
