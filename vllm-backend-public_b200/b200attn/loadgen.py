"""Benchmark client: what a reference benchmark Job does to an engine.

It is the container side of `benchmark-deployer /run`: the engine URL arrives as env `VLLM_MODEL_ENDPOINT`
(reference benchmark-deployer/deployer_manager.py:2195-2199, e.g. http://host:8000/v1/completions), an optional JSON
config is mounted at /app/configs/eval_config.json (reference deployer_manager.py:2145-2189), readiness is polled on
`GET {engine}/v1/models` like benchmark-eval does (reference benchmark-eval/main.py:101-110), and the results are
POSTed as `EvaluationPayload` (reference benchmark-results/models.py:6-19) to `{results}/raw_input` and
`{results}/standardized_output` (reference benchmark-results/main.py:55-63), whose primary key is
f"{timestamp}-{benchmark_name}-{run_id}" (reference benchmark-results/results_manager.py:22).

Workload (BASELINE.json configs[4]): `num_requests` prompts with lengths drawn log-uniformly from
[min_prompt, max_prompt] tokens (seed 0), sent as token-id prompts, `max_tokens` each, at most `concurrency` in
flight, streamed so TTFT and inter-token latency are measured at the client.

    python -m b200attn.loadgen --endpoint http://127.0.0.1:8005/v1/completions --model llama-3-8b \
        --num-requests 256 --concurrency 64 --min-prompt 128 --max-prompt 32768 --max-tokens 128 --out result.json
"""
from __future__ import annotations

import argparse
import asyncio
import json
import math
import os
import random
import statistics
import sys
import time
import uuid
from datetime import datetime, timezone
from typing import Any, Dict, List, Optional

REQUIRED_PAYLOAD_FIELDS = ("benchmark_name", "data", "timestamp", "model_id", "tokenizer_id", "source")


def make_payload(benchmark_name: str, data: Any, model_id: str, tokenizer_id: str, source: str,
                 timestamp: Optional[str] = None, run_id: Optional[str] = None) -> Dict[str, Any]:
    """The EvaluationPayload JSON object (reference benchmark-results/models.py:6-19)."""
    return {"run_id": run_id, "benchmark_name": benchmark_name, "data": data,
            "timestamp": timestamp or datetime.now(timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ"),
            "model_id": model_id, "tokenizer_id": tokenizer_id, "source": source}


def payload_pk(p: Dict[str, Any]) -> str:
    """Primary key benchmark-results stores the payload under (reference results_manager.py:22,58)."""
    return f"{p['timestamp']}-{p['benchmark_name']}-{p.get('run_id') or None}"


def prompt_lengths(n: int, lo: int, hi: int, seed: int = 0) -> List[int]:
    """Log-uniform integer lengths in [lo, hi]."""
    rng = random.Random(seed)
    return [min(hi, max(lo, int(round(math.exp(rng.uniform(math.log(lo), math.log(hi))))))) for _ in range(n)]


def make_prompt(length: int, vocab: int, seed: int) -> List[int]:
    rng = random.Random(seed)
    return [rng.randrange(vocab) for _ in range(length)]


def pct(xs: List[float], q: float) -> Optional[float]:
    if not xs:
        return None
    s = sorted(xs)
    k = (len(s) - 1) * q
    f, c = math.floor(k), math.ceil(k)
    return s[f] + (s[c] - s[f]) * (k - f)


def summarise(records: List[Dict[str, Any]], wall_s: float) -> Dict[str, Any]:
    ok = [r for r in records if r.get("error") is None]
    ttft = [r["ttft_s"] for r in ok if r.get("ttft_s") is not None]
    itl = [x for r in ok for x in r.get("itl_s", [])]
    lat = [r["latency_s"] for r in ok]
    out_tok = sum(r["completion_tokens"] for r in ok)
    in_tok = sum(r["prompt_tokens"] for r in ok)

    def stats(xs):
        return None if not xs else {"mean": statistics.fmean(xs), "p50": pct(xs, 0.5), "p90": pct(xs, 0.9),
                                    "p99": pct(xs, 0.99), "max": max(xs)}
    return {"num_requests": len(records), "num_ok": len(ok), "num_failed": len(records) - len(ok), "wall_s": wall_s,
            "prompt_tokens": in_tok, "completion_tokens": out_tok,
            "output_tok_per_s": out_tok / wall_s if wall_s > 0 else None,
            "total_tok_per_s": (in_tok + out_tok) / wall_s if wall_s > 0 else None,
            "requests_per_s": len(ok) / wall_s if wall_s > 0 else None,
            "ttft_s": stats(ttft), "itl_s": stats(itl), "latency_s": stats(lat)}


async def wait_ready(base: str, tries: int, delay_s: float) -> None:
    """benchmark-eval's readiness loop (reference benchmark-eval/main.py:101-110): GET /v1/models until 200."""
    import aiohttp
    last = None
    async with aiohttp.ClientSession() as s:
        for _ in range(tries):
            try:
                async with s.get(base + "/v1/models", timeout=aiohttp.ClientTimeout(total=10)) as r:
                    if r.status == 200:
                        return
                    last = f"HTTP {r.status}"
            except Exception as e:
                last = repr(e)
            await asyncio.sleep(delay_s)
    raise RuntimeError(f"engine at {base} not ready after {tries} tries: {last}")


async def one_request(session, endpoint: str, model: str, prompt: List[int], max_tokens: int, temperature: float,
                      seed: int) -> Dict[str, Any]:
    body = {"model": model, "prompt": prompt, "max_tokens": max_tokens, "temperature": temperature, "seed": seed,
            "stream": True, "ignore_eos": True, "stream_options": {"include_usage": True}}
    rec: Dict[str, Any] = {"prompt_tokens": len(prompt), "completion_tokens": 0, "ttft_s": None, "itl_s": [],
                           "latency_s": None, "error": None, "finish_reason": None}
    t0 = time.perf_counter()
    last = None
    try:
        async with session.post(endpoint, json=body) as resp:
            if resp.status != 200:
                rec["error"] = f"HTTP {resp.status}: {(await resp.text())[:200]}"
                return rec
            buf = b""
            async for chunk in resp.content.iter_any():
                now = time.perf_counter()
                buf += chunk
                while b"\n\n" in buf:
                    line, buf = buf.split(b"\n\n", 1)
                    if not line.startswith(b"data: "):
                        continue
                    data = line[6:]
                    if data == b"[DONE]":
                        continue
                    ev = json.loads(data)
                    if ev.get("usage"):
                        rec["completion_tokens"] = ev["usage"]["completion_tokens"]
                    for ch in ev.get("choices", []):
                        if ch.get("finish_reason"):
                            rec["finish_reason"] = ch["finish_reason"]
                        if rec["ttft_s"] is None:
                            rec["ttft_s"] = now - t0
                        elif last is not None and now > last:
                            rec["itl_s"].append(now - last)
                        last = now
    except Exception as e:
        rec["error"] = repr(e)
    rec["latency_s"] = time.perf_counter() - t0
    return rec


async def run_load(endpoint: str, model: str, lengths: List[int], vocab: int, max_tokens: int, concurrency: int,
                   temperature: float = 0.0, seed: int = 0) -> Dict[str, Any]:
    import aiohttp
    sem = asyncio.Semaphore(concurrency)
    records: List[Optional[Dict[str, Any]]] = [None] * len(lengths)
    conn = aiohttp.TCPConnector(limit=0)
    timeout = aiohttp.ClientTimeout(total=None, sock_read=600)

    async with aiohttp.ClientSession(connector=conn, timeout=timeout) as session:
        async def worker(i: int, n: int):
            prompt = make_prompt(n, vocab, seed * 1000003 + i)
            async with sem:
                records[i] = await one_request(session, endpoint, model, prompt, max_tokens, temperature, seed + i)

        t0 = time.perf_counter()
        await asyncio.gather(*(worker(i, n) for i, n in enumerate(lengths)))
        wall = time.perf_counter() - t0
    return {"records": records, "summary": summarise(records, wall)}


async def post_results(results_url: str, raw: Dict[str, Any], std: Dict[str, Any]) -> Dict[str, Any]:
    import aiohttp
    out = {}
    async with aiohttp.ClientSession() as s:
        for path, payload in (("/raw_input", raw), ("/standardized_output", std)):
            try:
                async with s.post(results_url.rstrip("/") + path, json=payload) as r:
                    out[path] = {"status": r.status, "body": (await r.text())[:300]}
            except Exception as e:
                out[path] = {"status": None, "body": repr(e)}
    return out


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--endpoint", default=os.environ.get("VLLM_MODEL_ENDPOINT"))
    ap.add_argument("--config", default="/app/configs/eval_config.json", help="optional JSON overriding the flags")
    ap.add_argument("--model", default=None)
    ap.add_argument("--num-requests", type=int, default=256)
    ap.add_argument("--concurrency", type=int, default=64)
    ap.add_argument("--min-prompt", type=int, default=128)
    ap.add_argument("--max-prompt", type=int, default=32768)
    ap.add_argument("--max-tokens", type=int, default=128)
    ap.add_argument("--temperature", type=float, default=0.0)
    ap.add_argument("--vocab", type=int, default=None, help="token ids are drawn from [0, vocab); default from the model list")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--benchmark-name", default="b200attn-decode-mixed-prompts")
    ap.add_argument("--tokenizer-id", default=None)
    ap.add_argument("--source", default="b200attn.loadgen")
    ap.add_argument("--results-url", default=os.environ.get("BENCHMARK_RESULTS_URL"))
    ap.add_argument("--ready-tries", type=int, default=60)
    ap.add_argument("--ready-delay", type=float, default=1.0)
    ap.add_argument("--out", default=None, help="write {raw, standardized} payloads to this JSON file")
    a = ap.parse_args(argv)
    if a.config and os.path.exists(a.config):
        for k, v in json.loads(open(a.config).read()).items():
            if hasattr(a, k.replace("-", "_")):
                setattr(a, k.replace("-", "_"), v)
    if not a.endpoint:
        ap.error("--endpoint or env VLLM_MODEL_ENDPOINT is required")
    base = a.endpoint.split("/v1/")[0].rstrip("/")

    async def go():
        import aiohttp
        await wait_ready(base, a.ready_tries, a.ready_delay)
        async with aiohttp.ClientSession() as s:
            async with s.get(base + "/v1/models") as r:
                info = (await r.json())["data"][0]
        model = a.model or info["id"]
        vocab = a.vocab or 256
        max_len = info.get("max_model_len")
        hi = a.max_prompt if not max_len else min(a.max_prompt, max_len - a.max_tokens - 1)
        lengths = prompt_lengths(a.num_requests, min(a.min_prompt, hi), hi, a.seed)
        res = await run_load(a.endpoint, model, lengths, vocab, a.max_tokens, a.concurrency, a.temperature, a.seed)
        ts = datetime.now(timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")
        run_id = "run-" + uuid.uuid4().hex[:12]
        workload = {"num_requests": a.num_requests, "concurrency": a.concurrency, "min_prompt": a.min_prompt,
                    "max_prompt": hi, "max_tokens": a.max_tokens, "temperature": a.temperature, "seed": a.seed,
                    "prompt_length_distribution": "log-uniform", "endpoint": a.endpoint, "vocab": vocab}
        raw = make_payload(a.benchmark_name, {"workload": workload, "requests": res["records"]}, model,
                           a.tokenizer_id or info.get("root", model), a.source, ts, run_id)
        std = make_payload(a.benchmark_name, {"workload": workload, "metrics": res["summary"]}, model,
                           a.tokenizer_id or info.get("root", model), a.source, ts, run_id)
        posted = await post_results(a.results_url, raw, std) if a.results_url else None
        return raw, std, posted

    raw, std, posted = asyncio.run(go())
    if a.out:
        with open(a.out, "w") as f:
            json.dump({"raw": raw, "standardized": std, "posted": posted}, f)
    print(json.dumps({"pk": payload_pk(std), "metrics": std["data"]["metrics"], "posted": posted}))
    return 0 if std["data"]["metrics"]["num_failed"] == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
