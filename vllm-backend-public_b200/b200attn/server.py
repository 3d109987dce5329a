"""OpenAI-compatible HTTP front of the engine: the endpoints the reference's services and Jobs call.

    python -m b200attn.server --model llama-3-8b --port 8005 --max-num-seqs 64 --max-model-len 33024 ...

  GET  /health           K8s liveness/readiness probes (reference benchmark-vllm/vllm_templates.py:182-201,352-369) and
                         benchmark-deployer's pod check (reference benchmark-deployer/deployer_manager.py:1597-1606):
                         200 once the weights, the KV pool and the decode graphs are ready, 503 before.
  GET  /v1/models        benchmark-eval's readiness poll (reference benchmark-eval/main.py:101-110).
  POST /v1/completions   the URL injected into benchmark Jobs as VLLM_MODEL_ENDPOINT
                         (reference benchmark-deployer/deployer_manager.py:2195-2199, docs/deployment/post_run.md:46).
  POST /v1/chat/completions, GET /version, GET /metrics: conveniences vLLM's server also has.

Flags are the reference's launch vector (b200attn.engine_args).  The engine runs in one thread (single writer of the
scheduler); HTTP handlers hand requests over and receive tokens through per-request asyncio queues.
Errors use OpenAI's envelope: {"error": {"message", "type", "param", "code"}} with 4xx/5xx.
"""
from __future__ import annotations

import asyncio
import os
import json
import sys
import threading
import time
import uuid
from typing import Any, Dict, List, Optional, Tuple

from fastapi import FastAPI, Request
from fastapi.responses import JSONResponse, PlainTextResponse, StreamingResponse

from .engine_args import EngineArgs


class EngineRunner:
    """Owns the engine thread.  `submit` is called from the event loop; tokens come back through asyncio queues."""

    def __init__(self, args: EngineArgs, tokenizer_path: Optional[str] = None):
        self.args = args
        self.tokenizer_path = tokenizer_path
        self.engine = None
        self.tokenizer = None
        self.ready = threading.Event()
        self.failed: Optional[str] = None
        self._wake = threading.Event()
        self._stop = False
        self._thread = threading.Thread(target=self._thread_main, name="b200attn-engine", daemon=True)
        self._engine_tid = None
        self.t_start = time.time()
        self.workers: list = []                  # child processes of ranks 1..tp-1
        self._events: list = []                  # (queue, token) pairs of the running step (engine thread only)
        self._loop = None

    def start(self) -> None:
        self._thread.start()

    def _thread_main(self) -> None:
        self._engine_tid = threading.get_ident()
        self._main()

    def stop(self) -> None:
        self._stop = True
        self._wake.set()
        self._thread.join(timeout=10)
        if self.workers:
            try:
                if self.engine is not None and self.failed is None:
                    self.engine.shutdown_workers()
                    self.engine.close()              # the workers close too: the barrier inside must be matched
            except Exception:
                pass
            for p in self.workers:
                try:
                    p.wait(timeout=10)
                except Exception:
                    p.kill()
            try:                                     # leave no NCCL communicator for interpreter shutdown to wait on
                import torch.distributed as dist
                if dist.is_initialized() and all(p.poll() is not None for p in self.workers):
                    dist.destroy_process_group()
            except Exception:
                pass

    def _main(self) -> None:
        try:
            import torch
            from .engine import LLMEngine
            from .model import resolve_config
            from .tokenizer import load_tokenizer
            from pathlib import Path
            cfg = resolve_config(self.args.model_name)
            kw = self.args.engine_kwargs()
            wdir = Path(self.args.model_name) if Path(self.args.model_name).is_dir() else None
            self.tokenizer = load_tokenizer(self.args.model_name, cfg.vocab, self.tokenizer_path)
            group = None
            tpn = self.args.tensor_parallel_size
            if tpn > 1:
                # --tensor-parallel-size N: this process is rank 0 (GPU 0, the HTTP front end, the request queue);
                # ranks 1..N-1 are child processes on GPUs 1..N-1 mirroring every step (b200attn.tp)
                from . import tp
                if torch.cuda.device_count() < tpn:
                    raise RuntimeError(f"--tensor-parallel-size {tpn} needs {tpn} GPUs, {torch.cuda.device_count()} visible")
                port = tp.free_port()
                spec = {"model": self.args.model_name, "engine": dict(kw, eos_token_id=self.tokenizer.eos_token_id)}
                self.workers = tp.launch_workers(tpn, spec, port)
                torch.cuda.set_device(0)
                tp.init_distributed(0, tpn, torch.device("cuda", 0), master_port=port)
                group = tp.TPGroup(torch.device("cuda", 0))
            self.engine = LLMEngine(cfg, torch.device("cuda", 0), weights_dir=wdir,
                                    eos_token_id=self.tokenizer.eos_token_id, tp=group, **kw)
            n = self.engine.capture_all()
            self.engine.after_step = self._flush_events
            print(f"[b200attn] {cfg.name}: {self.engine.model.weight_bytes() / 2**30:.1f} GiB weights, "
                  f"{self.engine.num_blocks} KV blocks x {self.engine.block_size} tokens, max_model_len "
                  f"{self.engine.max_model_len}, {n} decode graphs, ready after {time.time() - self.t_start:.1f}s",
                  file=sys.stderr, flush=True)
            self.ready.set()
        except Exception as e:                                    # start-up failure: /health stays 503, process exits
            import traceback
            traceback.print_exc()
            self.failed = f"{type(e).__name__}: {e}"
            self.ready.set()
            return
        eng = self.engine
        while not self._stop:
            dead = [p.pid for p in self.workers if p.poll() is not None]
            if dead:                                              # a tensor-parallel rank died: the engine cannot step
                self.failed = f"tensor-parallel worker process(es) {dead} exited"
                for r in list(eng.requests.values()):
                    r.finished, r.finish_reason = True, "error"
                    if r.on_token:
                        r.on_token(r, None)
                self._flush_events()
                return
            if eng.has_work():
                try:
                    eng.step()
                except Exception as e:                            # CUDA fatal: fail loudly so the orchestrator restarts us
                    import traceback
                    traceback.print_exc()
                    self.failed = f"{type(e).__name__}: {e}"
                    with eng._lock:                               # requests queued but never admitted must end too
                        queued, eng._incoming = eng._incoming, []
                    for r in list(eng.requests.values()) + queued:
                        r.finished, r.finish_reason = True, "error"
                        if r.on_token:
                            r.on_token(r, None)
                    self._flush_events()
                    return
            else:
                self._wake.wait(timeout=0.05)
                self._wake.clear()

    def submit(self, prompt_ids: List[int], params, loop: asyncio.AbstractEventLoop) -> Tuple[Any, asyncio.Queue]:
        q: asyncio.Queue = asyncio.Queue()

        def on_token(req, tok):
            if threading.get_ident() == self._engine_tid:        # engine thread, inside step(): only collect
                self._events.append((q, tok))
            else:                                                # any other caller: deliver now
                loop.call_soon_threadsafe(q.put_nowait, tok)

        self._loop = loop
        req = self.engine.add_request(prompt_ids, params, on_token)
        self._wake.set()
        return req, q

    def _flush_events(self) -> None:
        """Engine thread, once per step: hand every token of the step to the event loop with ONE wake-up (a
        call_soon_threadsafe per token costs a self-pipe write and a loop iteration each)."""
        if self._events and self._loop is not None:
            evs, self._events = self._events, []
            try:
                self._loop.call_soon_threadsafe(self._deliver, evs)
            except RuntimeError:                     # loop already closed (shutdown)
                pass

    @staticmethod
    def _deliver(evs) -> None:
        for q, tok in evs:
            q.put_nowait(tok)


# ------------------------------------------------------------------------------------------------
# request parsing
# ------------------------------------------------------------------------------------------------
class ApiError(Exception):
    def __init__(self, status: int, message: str, type_: str = "invalid_request_error", param: Optional[str] = None,
                 code: Optional[str] = None):
        super().__init__(message)
        self.status, self.message, self.type, self.param, self.code = status, message, type_, param, code

    def body(self) -> Dict[str, Any]:
        return {"error": {"message": self.message, "type": self.type, "param": self.param, "code": self.code}}


def parse_prompts(prompt: Any, tokenizer) -> List[List[int]]:
    """OpenAI `prompt`: a string, a list of strings, a list of token ids, or a list of lists of token ids."""
    if isinstance(prompt, str):
        return [tokenizer.encode(prompt)]
    if isinstance(prompt, list) and prompt:
        if all(isinstance(x, int) and not isinstance(x, bool) for x in prompt):
            return [list(prompt)]
        if all(isinstance(x, str) for x in prompt):
            return [tokenizer.encode(x) for x in prompt]
        if all(isinstance(x, list) and x and all(isinstance(t, int) and not isinstance(t, bool) for t in x) for x in prompt):
            return [list(x) for x in prompt]
    raise ApiError(400, "prompt must be a string, a list of strings, a list of token ids or a list of token-id lists",
                   param="prompt")


def sampling_from_body(body: Dict[str, Any], default_max_tokens: int = 16):
    from .engine import SamplingParams

    def num(key, default, cast, lo=None, hi=None):
        v = body.get(key, default)
        if v is None:
            v = default
        if isinstance(v, bool) or not isinstance(v, (int, float)):
            raise ApiError(400, f"{key} must be a number", param=key)
        v = cast(v)
        if (lo is not None and v < lo) or (hi is not None and v > hi):
            raise ApiError(400, f"{key}={v} outside [{lo}, {hi}]", param=key)
        return v

    max_tokens = num("max_tokens", body.get("max_completion_tokens", default_max_tokens), int, 1)
    temperature = num("temperature", 1.0, float, 0.0, 100.0)
    top_p = num("top_p", 1.0, float, 0.0, 1.0)
    top_k = num("top_k", 0, int, -1)                      # vLLM extension of the OpenAI body: -1 / 0 = no limit
    n = num("n", 1, int, 1, 64)
    if top_p <= 0.0:
        raise ApiError(400, "top_p must be in (0, 1]", param="top_p")
    seed = body.get("seed")
    if seed is not None and (isinstance(seed, bool) or not isinstance(seed, int)):
        raise ApiError(400, "seed must be an integer", param="seed")
    stop = body.get("stop")
    if stop is None:
        stop = []
    elif isinstance(stop, str):
        stop = [stop]
    elif not (isinstance(stop, list) and all(isinstance(s, str) and s for s in stop) and len(stop) <= 4):
        raise ApiError(400, "stop must be a string or a list of up to 4 non-empty strings", param="stop")
    stop_ids = body.get("stop_token_ids") or []
    if not (isinstance(stop_ids, list) and all(isinstance(t, int) for t in stop_ids)):
        raise ApiError(400, "stop_token_ids must be a list of integers", param="stop_token_ids")
    sp = SamplingParams(max_tokens=max_tokens, temperature=temperature, top_k=max(top_k, 0), top_p=top_p, seed=seed,
                        ignore_eos=bool(body.get("ignore_eos", False)), stop_token_ids=tuple(stop_ids))
    return sp, n, stop


class _Choice:
    """One generated sequence being collected by an HTTP handler."""

    def __init__(self, index: int, req, queue: asyncio.Queue, stop: List[str], tokenizer, echo_text: str = ""):
        self.index, self.req, self.queue, self.stop, self.tok = index, req, queue, stop, tokenizer
        self.ids: List[int] = []
        self.text = ""
        self.sent = 0                       # characters already streamed
        self.finish_reason: Optional[str] = None
        self.echo_text = echo_text

    def _hold_back(self) -> int:
        """Characters that may still turn into a stop string: do not stream them yet."""
        return max((len(s) - 1 for s in self.stop), default=0)

    async def next_delta(self, runner: "EngineRunner") -> Optional[str]:
        """Wait for the next token; return newly releasable text ('' if none), or None when the choice is finished."""
        if self.finish_reason is not None:
            return None
        tok = await self.queue.get()
        if tok is None:
            reason = self.req.finish_reason or "stop"
            if reason.startswith("error:"):                 # refused at admission (it can never fit): the request's fault
                raise ApiError(400, reason[6:].strip(), param="prompt")
            if reason.startswith("error"):
                raise ApiError(500, reason, "server_error")
            self.finish_reason = "length" if reason == "length" else "stop"
            delta = self.text[self.sent:]
            self.sent = len(self.text)
            return delta
        self.ids.append(tok)
        self.text = self.tok.decode(self.ids)
        for s in self.stop:
            k = self.text.find(s, max(0, self.sent - len(s)))
            if k >= 0:
                self.text = self.text[:k]
                self.finish_reason = "stop"
                runner.engine.abort(self.req.id)
                delta = self.text[self.sent:]
                self.sent = len(self.text)
                return delta
        upto = max(self.sent, len(self.text) - self._hold_back())
        if self.text.endswith("�"):     # incomplete UTF-8 sequence: wait for the next byte
            upto = max(self.sent, min(upto, len(self.text) - 1))
        delta = self.text[self.sent:upto]
        self.sent = upto
        return delta


# ------------------------------------------------------------------------------------------------
# the app
# ------------------------------------------------------------------------------------------------
def create_app(runner: EngineRunner):
    app = FastAPI(title="b200attn", docs_url=None, redoc_url=None)
    args = runner.args
    counters = {"requests": 0, "errors": 0}

    def err(e: ApiError):
        counters["errors"] += 1
        return JSONResponse(e.body(), status_code=e.status)

    def need_ready():
        if runner.failed:
            raise ApiError(503, f"engine failed: {runner.failed}", "server_error")
        if not runner.ready.is_set():
            raise ApiError(503, "engine is starting", "server_error")

    @app.get("/health")
    async def health():
        if runner.failed or not runner.ready.is_set():
            return PlainTextResponse("", status_code=503)
        return PlainTextResponse("", status_code=200)

    @app.get("/ping")
    async def ping():
        return await health()

    @app.get("/version")
    async def version():
        from . import _lib
        return {"version": "b200attn-abi%d" % _lib.ABI_VERSION}

    @app.get("/v1/models")
    async def models():
        try:
            need_ready()
        except ApiError as e:
            return err(e)
        eng = runner.engine
        return {"object": "list", "data": [{
            "id": args.public_name, "object": "model", "created": int(runner.t_start), "owned_by": "b200attn",
            "root": args.model_name, "parent": None, "max_model_len": eng.max_model_len}]}

    @app.get("/metrics")
    async def metrics():
        lines = [f"b200attn_http_requests_total {counters['requests']}", f"b200attn_http_errors_total {counters['errors']}"]
        eng = runner.engine
        if eng is not None:
            # a snapshot the engine thread publishes after every step: the C++ scheduler is single-writer and is never
            # read from the event loop
            m = eng.metrics
            lines += [f"b200attn_num_requests_running {m['running']}",
                      f"b200attn_num_requests_waiting {m['waiting']}",
                      f"b200attn_kv_blocks_free {m['free_blocks']}", f"b200attn_kv_blocks_total {eng.num_blocks}",
                      f"b200attn_prompt_tokens_total {m['prompt_tokens']}", f"b200attn_generation_tokens_total {m['generated_tokens']}",
                      f"b200attn_engine_steps_total {m['steps']}", f"b200attn_preemptions_total {m['preemptions']}",
                      f"b200attn_engine_mixed_steps_total {m.get('mixed_steps', 0)}",
                      f"b200attn_engine_busy_seconds_total {m.get('busy_s', 0)}",
                      f"b200attn_engine_gpu_wait_seconds_total {m.get('sync_wait_s', 0)}",
                      f"b200attn_engine_mixed_step_seconds_total {m.get('mixed_busy_s', 0)}"]
        return PlainTextResponse("\n".join(lines) + "\n")

    def check_model(body):
        m = body.get("model")
        if m is not None and m not in (args.public_name, args.model_name):
            raise ApiError(404, f"The model `{m}` does not exist.", "NotFoundError", "model", "model_not_found")

    async def run_completion(body: Dict[str, Any], prompts: List[List[int]], chat: bool):
        """Shared by /v1/completions and /v1/chat/completions."""
        eng, tok = runner.engine, runner.tokenizer
        sp, n, stop = sampling_from_body(body)
        for p in prompts:
            if not p:
                raise ApiError(400, "empty prompt", param="prompt")
            if len(p) + 1 > eng.max_model_len:
                raise ApiError(400, f"This model's maximum context length is {eng.max_model_len} tokens; the prompt has "
                               f"{len(p)}", param="prompt", code="context_length_exceeded")
            if max(p) >= eng.cfg.vocab or min(p) < 0:
                raise ApiError(400, f"token id outside the vocabulary [0, {eng.cfg.vocab})", param="prompt")
            if (len(p) + eng.block_size) // eng.block_size > eng.num_blocks:
                raise ApiError(400, f"a prompt of {len(p)} tokens cannot fit the KV pool of {eng.num_blocks} blocks",
                               param="prompt", code="context_length_exceeded")
        loop = asyncio.get_running_loop()
        echo = bool(body.get("echo", False)) and not chat
        choices: List[_Choice] = []
        from dataclasses import replace
        for pi, p in enumerate(prompts):
            for k in range(n):
                spk = sp if (n == 1 or sp.seed is None) else replace(sp, seed=sp.seed + k)
                req, q = runner.submit(p, spk, loop)
                choices.append(_Choice(pi * n + k, req, q, stop, tok, tok.decode(p) if echo else ""))
        rid = ("chatcmpl-" if chat else "cmpl-") + uuid.uuid4().hex
        created = int(time.time())
        obj = "chat.completion" if chat else "text_completion"
        prompt_tokens = sum(len(p) for p in prompts)

        def usage():
            c = sum(len(ch.ids) for ch in choices)
            return {"prompt_tokens": prompt_tokens, "completion_tokens": c, "total_tokens": prompt_tokens + c}

        def abort_all():
            for ch in choices:
                if ch.finish_reason is None:
                    eng.abort(ch.req.id)

        if not body.get("stream", False):
            try:
                for ch in choices:
                    while await ch.next_delta(runner) is not None:
                        pass
            except BaseException:
                abort_all()
                raise
            out = []
            for ch in choices:
                if chat:
                    out.append({"index": ch.index, "message": {"role": "assistant", "content": ch.text},
                                "finish_reason": ch.finish_reason, "logprobs": None})
                else:
                    out.append({"index": ch.index, "text": ch.echo_text + ch.text, "logprobs": None,
                                "finish_reason": ch.finish_reason})
            return JSONResponse({"id": rid, "object": obj, "created": created, "model": args.public_name,
                                 "choices": out, "usage": usage()})

        include_usage = bool((body.get("stream_options") or {}).get("include_usage", False))

        async def sse():
            def chunk(ch, delta, finish):
                if chat:
                    c = {"index": ch.index, "delta": ({"content": delta} if delta else {}), "finish_reason": finish}
                else:
                    c = {"index": ch.index, "text": delta, "logprobs": None, "finish_reason": finish}
                d = {"id": rid, "object": obj + (".chunk" if chat else ""), "created": created,
                     "model": args.public_name, "choices": [c]}
                return "data: " + json.dumps(d) + "\n\n"
            try:
                if len(choices) == 1 and not chat and not choices[0].echo_text:
                    # the common case (one prompt, n = 1): no task per token, and only the delta goes through the JSON
                    # encoder -- at tensor-parallel speeds the event loop, not the GPU, bounds the tokens per second
                    ch = choices[0]
                    head = ('data: {"id": ' + json.dumps(rid) + ', "object": ' + json.dumps(obj) + ', "created": '
                            + str(created) + ', "model": ' + json.dumps(args.public_name)
                            + ', "choices": [{"index": ' + str(ch.index) + ', "text": ')
                    tail_none = ', "logprobs": null, "finish_reason": null}]}\n\n'
                    while True:
                        delta = await ch.next_delta(runner)
                        if delta is None:
                            break
                        if ch.finish_reason is not None:
                            yield (head + json.dumps(delta) + ', "logprobs": null, "finish_reason": '
                                   + json.dumps(ch.finish_reason) + '}]}\n\n')
                            break
                        if delta:
                            yield head + json.dumps(delta) + tail_none
                    if include_usage:
                        yield "data: " + json.dumps({"id": rid, "object": obj, "created": created,
                                                     "model": args.public_name, "choices": [], "usage": usage()}) + "\n\n"
                    yield "data: [DONE]\n\n"
                    return
                pending = {asyncio.ensure_future(ch.next_delta(runner)): ch for ch in choices}
                for ch in choices:
                    if chat:
                        yield "data: " + json.dumps({"id": rid, "object": obj + ".chunk", "created": created,
                                                     "model": args.public_name, "choices": [
                                                         {"index": ch.index, "delta": {"role": "assistant"},
                                                          "finish_reason": None}]}) + "\n\n"
                    elif ch.echo_text:
                        yield chunk(ch, ch.echo_text, None)
                while pending:
                    done, _ = await asyncio.wait(pending.keys(), return_when=asyncio.FIRST_COMPLETED)
                    for fut in done:
                        ch = pending.pop(fut)
                        delta = fut.result()
                        if delta is None:
                            continue
                        if ch.finish_reason is not None:
                            yield chunk(ch, delta, ch.finish_reason)
                        else:
                            if delta:
                                yield chunk(ch, delta, None)
                            pending[asyncio.ensure_future(ch.next_delta(runner))] = ch
                if include_usage:
                    yield "data: " + json.dumps({"id": rid, "object": obj + (".chunk" if chat else ""), "created": created,
                                                 "model": args.public_name, "choices": [], "usage": usage()}) + "\n\n"
                yield "data: [DONE]\n\n"
            finally:
                abort_all()                 # client went away: free the sequences

        return StreamingResponse(sse(), media_type="text/event-stream")

    async def read_body(request: Request) -> Dict[str, Any]:
        try:
            body = await request.json()
        except Exception:
            raise ApiError(400, "request body is not valid JSON")
        if not isinstance(body, dict):
            raise ApiError(400, "request body must be a JSON object")
        return body

    @app.post("/v1/completions")
    async def completions(request: Request):
        counters["requests"] += 1
        try:
            need_ready()
            body = await read_body(request)
            check_model(body)
            if "prompt" not in body:
                raise ApiError(400, "prompt is required", param="prompt")
            prompts = parse_prompts(body["prompt"], runner.tokenizer)
            return await run_completion(body, prompts, chat=False)
        except ApiError as e:
            return err(e)
        except ValueError as e:
            return err(ApiError(400, str(e)))

    @app.post("/v1/chat/completions")
    async def chat_completions(request: Request):
        counters["requests"] += 1
        try:
            need_ready()
            body = await read_body(request)
            check_model(body)
            msgs = body.get("messages")
            if not (isinstance(msgs, list) and msgs and all(isinstance(m, dict) and "role" in m for m in msgs)):
                raise ApiError(400, "messages must be a non-empty list of {role, content}", param="messages")
            tok = runner.tokenizer
            text = tok.apply_chat_template(msgs) if hasattr(tok, "apply_chat_template") else None
            if text is not None:                          # the model's own template (instruct weights from an HF directory)
                ids = tok.encode_rendered(text)
            else:                                         # no template shipped (byte tokenizer, random-init weights)
                text = "".join(f"<|{m['role']}|>\n{m.get('content') or ''}\n" for m in msgs) + "<|assistant|>\n"
                ids = tok.encode(text)
            return await run_completion(body, [ids], chat=True)
        except ApiError as e:
            return err(e)
        except ValueError as e:
            return err(ApiError(400, str(e)))

    return app


def main(argv: Optional[List[str]] = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    tokenizer = None
    if "--tokenizer" in argv:
        i = argv.index("--tokenizer")
        tokenizer = argv[i + 1]
        del argv[i:i + 2]
    if "--config" in argv:                        # a VLLMConfig YAML of the reference instead of flags
        i = argv.index("--config")
        args = EngineArgs.from_yaml(argv[i + 1])
    else:
        args = EngineArgs.from_cli(argv)
    for note, on in (("--device %s ignored: runs on cuda:0" % args.device, args.device not in (None, "cuda", "auto")),
                     ("--dtype %s served as bfloat16" % args.dtype, args.dtype not in ("auto", "bfloat16")),
                     ("--quantization %s ignored" % args.quantization, bool(args.quantization))):
        if on:
            print("[b200attn] " + note, file=sys.stderr)
    import uvicorn
    # the engine thread and the HTTP event loop share this interpreter: a 5 ms GIL switch interval would cost the engine
    # up to 5 ms every time it gives the lock away at a launch; and no CPU math here is worth an OpenMP pool
    sys.setswitchinterval(1e-4)
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    runner = EngineRunner(args, tokenizer)
    runner.start()
    app = create_app(runner)
    config = uvicorn.Config(app, host=args.host, port=args.port, log_level="warning", access_log=False)
    server = uvicorn.Server(config)

    def watchdog():                               # exit non-zero when the engine dies so the orchestrator restarts the pod
        runner.ready.wait()
        while runner.failed is None and not server.should_exit:
            time.sleep(0.5)
        if runner.failed:
            print(f"[b200attn] fatal: {runner.failed}", file=sys.stderr, flush=True)
            server.should_exit = True

    threading.Thread(target=watchdog, daemon=True).start()

    # uvicorn handles SIGTERM / SIGINT itself, but when its loop has drained it restores the previous handlers and
    # RE-RAISES the signal: with the default handler the process would die inside server.run(), runner.stop() would
    # never run and the tensor-parallel ranks would be orphaned.  A handler of our own makes the re-raise harmless.
    def on_signal(signum, frame):
        server.should_exit = True

    try:
        import signal
        for sig in (signal.SIGTERM, signal.SIGINT):
            signal.signal(sig, on_signal)
    except ValueError:                            # not the main thread (embedded use): the embedder owns the signals
        pass
    server.run()
    runner.stop()
    return 1 if runner.failed else 0


if __name__ == "__main__":
    sys.exit(main())
