"""Closed-loop load generator for /v1/completions: N concurrent users, each sending
the next request as soon as the previous one finishes, fixed prompt / output lengths, streaming.
Reports request/s, output tokens/s, TTFT p50/p95/p99 (time from request send to the first streamed
chunk — the definition used by vLLM's benchmark client, SP:vllm/benchmarks/lib/endpoint_request_func.py
:193-238), inter-token latency and end-to-end latency. Locust (reference requirements.txt:35) is not on
the image; aiohttp is."""
from __future__ import annotations

import asyncio
import json
import random
import statistics
import time


def _pct(xs, p):
    if not xs:
        return None
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(round(p / 100 * (len(xs) - 1))))]


async def _one(session, url, model, prompt_ids, max_tokens, rec, extra=None):
    body = {"model": model, "prompt": prompt_ids, "max_tokens": max_tokens, "temperature": 0.0,
            "ignore_eos": True, "stream": True}
    if extra:
        body.update(extra)
    t0 = time.perf_counter()
    first = None
    n = 0
    last = t0
    itl = []
    try:
        async with session.post(url, json=body) as resp:
            if resp.status != 200:
                rec["errors"] += 1
                return
            async for raw in resp.content:
                line = raw.decode().strip()
                if not line.startswith("data:"):
                    continue
                data = line[5:].strip()
                if data == "[DONE]":
                    break
                now = time.perf_counter()
                if first is None:
                    first = now - t0
                else:
                    itl.append(now - last)
                last = now
                ch = json.loads(data).get("choices") or []
                if ch:
                    n += 1
    except Exception:
        rec["errors"] += 1
        return
    rec["ttft"].append(first if first is not None else float("nan"))
    rec["e2e"].append(time.perf_counter() - t0)
    rec["itl"] += itl
    rec["tokens"] += n
    rec["requests"] += 1


async def run_load(base_url: str, model: str, concurrency: int, num_requests: int, input_len: int = 512,
                   output_len: int = 128, vocab: int = 32000, seed: int = 0, extra: dict | None = None,
                   think_time: float = 0.0, shared_prefix_len: int = 0, request_rate: float = 0.0) -> dict:
    """One load level. Closed loop (default): `concurrency` users, each sending its next request when the previous
    one has finished (after `think_time` seconds, Locust's wait_time). Open loop (`request_rate` > 0): requests
    arrive as a Poisson process of that rate whatever the server does, at most `concurrency` in flight.
    shared_prefix_len: the first that many prompt tokens are the same for every request (a system prompt; what a
    server-side prefix cache can exploit) — otherwise no prompt token sequence is ever sent twice."""
    import aiohttp

    # every (seed, concurrency level) gets its own prompts: no prompt is ever sent twice, so a server-side prefix
    # cache cannot turn prefill into a lookup
    rng = random.Random(seed * 1000003 + concurrency * 7919 + num_requests)
    url = base_url.rstrip("/") + "/v1/completions"
    rec = {"ttft": [], "e2e": [], "itl": [], "tokens": 0, "requests": 0, "errors": 0}
    k = max(0, min(shared_prefix_len, input_len - 1))
    prefix = [1] + [rng.randrange(3, vocab) for _ in range(k)]
    pending = [prefix + [rng.randrange(3, vocab) for _ in range(input_len - 1 - k)] for _ in range(num_requests)]
    lock = asyncio.Lock()

    async def user(session):
        while True:
            async with lock:
                if not pending:
                    return
                p = pending.pop()
            await _one(session, url, model, p, output_len, rec, extra)
            if think_time > 0:
                await asyncio.sleep(think_time)

    async def open_loop(session):
        sem = asyncio.Semaphore(concurrency)
        tasks = []

        async def fire(p):
            async with sem:
                await _one(session, url, model, p, output_len, rec, extra)

        while pending:
            tasks.append(asyncio.create_task(fire(pending.pop())))
            await asyncio.sleep(rng.expovariate(request_rate))
        await asyncio.gather(*tasks)

    timeout = aiohttp.ClientTimeout(total=3600)
    t0 = time.perf_counter()
    async with aiohttp.ClientSession(timeout=timeout, connector=aiohttp.TCPConnector(limit=0)) as s:
        if request_rate > 0:
            await open_loop(s)
        else:
            await asyncio.gather(*[user(s) for _ in range(concurrency)])
    wall = time.perf_counter() - t0
    return {"concurrency": concurrency, "requests": rec["requests"], "errors": rec["errors"], "wall_s": wall,
            "input_len": input_len, "output_len": output_len, "request_rate": request_rate, "think_time_s": think_time,
            "shared_prefix_len": k,
            "requests_per_s": rec["requests"] / wall, "output_tokens_per_s": rec["tokens"] / wall,
            "total_tokens_per_s": (rec["tokens"] + rec["requests"] * input_len) / wall,
            "ttft_p50_ms": 1e3 * (_pct(rec["ttft"], 50) or 0), "ttft_p95_ms": 1e3 * (_pct(rec["ttft"], 95) or 0),
            "ttft_p99_ms": 1e3 * (_pct(rec["ttft"], 99) or 0),
            "itl_mean_ms": 1e3 * (statistics.mean(rec["itl"]) if rec["itl"] else 0),
            "e2e_p50_ms": 1e3 * (_pct(rec["e2e"], 50) or 0), "e2e_p99_ms": 1e3 * (_pct(rec["e2e"], 99) or 0)}


async def sweep(base_url, model, concurrencies, requests_per_level=None, **kw):
    out = []
    for c in concurrencies:
        n = requests_per_level or max(4 * c, 16)
        out.append(await run_load(base_url, model, c, n, **kw))
        print(json.dumps(out[-1]), flush=True)
    return out


def main(argv=None) -> int:
    """`python -m dlti_b200.serve.loadgen --url http://127.0.0.1:8000 --model llama2-7b --concurrency 1,16,64`:
    the load generator as its own process (a client that shares the server's interpreter would compete with the
    engine thread for the GIL and distort the latencies it measures)."""
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--url", required=True)
    ap.add_argument("--model", required=True)
    ap.add_argument("--concurrency", default="1,8,32,64")
    ap.add_argument("--requests_per_level", type=int, default=None)
    ap.add_argument("--input_len", type=int, default=512)
    ap.add_argument("--output_len", type=int, default=128)
    ap.add_argument("--vocab", type=int, default=32000)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--think_time", type=float, default=0.0, help="seconds a user waits between two requests")
    ap.add_argument("--shared_prefix_len", type=int, default=0, help="prompt tokens common to all requests")
    ap.add_argument("--request_rate", type=float, default=0.0, help="> 0: open loop, Poisson arrivals per second")
    ap.add_argument("--extra", type=str, default=None, help="JSON object merged into every request body")
    ap.add_argument("--out", type=str, default=None)
    a = ap.parse_args(argv)
    res = asyncio.run(sweep(a.url, a.model, [int(x) for x in a.concurrency.split(",")], a.requests_per_level,
                            input_len=a.input_len, output_len=a.output_len, vocab=a.vocab, seed=a.seed,
                            extra=json.loads(a.extra) if a.extra else None, think_time=a.think_time,
                            shared_prefix_len=a.shared_prefix_len, request_rate=a.request_rate))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
