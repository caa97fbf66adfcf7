"""OpenAI-compatible HTTP front-end: POST /v1/completions (JSON or SSE
stream), GET /v1/models, GET /health, GET /metrics (Prometheus text). Schema follows what vLLM serves
(SP:vllm/entrypoints/openai/completion/protocol.py:42-90,466-545); SSE framing is
`data: {json}\\n\\n` ... `data: [DONE]\\n\\n`. The reference only declares this capability
(reference README.md:10,16; requirements.txt:18).

The engine loop runs in its own thread; the asyncio side hands requests over a thread-safe queue and
receives tokens through per-request asyncio queues.
"""
from __future__ import annotations

import asyncio
import json
import queue
import threading
import time
import uuid

try:  # module-level so FastAPI can resolve the (stringified) annotations of the route handlers
    from fastapi import FastAPI, Request
    from fastapi.responses import JSONResponse, PlainTextResponse, StreamingResponse
except Exception:  # pragma: no cover - fastapi is an optional dependency of the package
    FastAPI = Request = JSONResponse = PlainTextResponse = StreamingResponse = None

from .engine import LLMEngine
from .params import SamplingParams


class AsyncEngine:
    def __init__(self, engine: LLMEngine):
        self.engine = engine
        self.inbox: queue.Queue = queue.Queue()
        self.streams: dict[str, tuple[asyncio.AbstractEventLoop, asyncio.Queue]] = {}
        self.stop = False
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.metrics = {"requests_total": 0, "tokens_total": 0, "ttft_sum": 0.0, "ttft_count": 0}

    def start(self):
        self.thread.start()

    def _loop(self):
        import torch
        if self.engine.device.type == "cuda":
            torch.cuda.set_device(self.engine.device)
        eng = self.engine
        while not self.stop:
            try:
                while True:
                    kind, payload = self.inbox.get_nowait() if eng.has_unfinished() else self.inbox.get(timeout=0.05)
                    if kind == "add":
                        rid, prompt, params = payload
                        try:
                            eng.add_request(prompt, params, rid)
                        except Exception as e:  # bad request (too long, ...)
                            self._emit(rid, {"error": str(e)})
                    elif kind == "abort":
                        eng.abort(payload)
            except queue.Empty:
                pass
            if not eng.has_unfinished():
                continue
            for out in eng.step():
                rid = out.request_id
                r = eng.requests.get(rid)
                if r is None:
                    continue
                if len(r.output_ids) == 1:
                    self.metrics["ttft_sum"] += time.time() - r.arrival
                    self.metrics["ttft_count"] += 1
                self.metrics["tokens_total"] += len(out.new_token_ids)
                delta = eng.text_delta(rid, final=out.finished)
                self._emit(rid, {"text": delta, "token_ids": out.new_token_ids, "finished": out.finished,
                                 "finish_reason": out.finish_reason, "logprobs": out.logprobs,
                                 "prompt_tokens": len(r.prompt_ids), "completion_tokens": len(r.output_ids)})
                if out.finished:
                    eng.release(rid)
        eng.shutdown_workers()

    def _emit(self, rid, item):
        s = self.streams.get(rid)
        if s:
            loop, q = s
            loop.call_soon_threadsafe(q.put_nowait, item)

    async def generate(self, prompt, params: SamplingParams, rid: str):
        q: asyncio.Queue = asyncio.Queue()
        self.streams[rid] = (asyncio.get_running_loop(), q)
        self.metrics["requests_total"] += 1
        self.inbox.put(("add", (rid, prompt, params)))
        try:
            while True:
                item = await q.get()
                yield item
                if item.get("finished") or "error" in item:
                    return
        finally:
            self.streams.pop(rid, None)

    def abort(self, rid: str):
        self.inbox.put(("abort", rid))


def _params_from_body(b: dict) -> SamplingParams:
    return SamplingParams(
        max_tokens=b.get("max_tokens", 16) or 16, temperature=float(b.get("temperature", 1.0) if b.get("temperature") is not None else 1.0),
        top_p=float(b.get("top_p", 1.0) or 1.0), top_k=int(b.get("top_k", 0) or 0), n=int(b.get("n", 1) or 1),
        seed=b.get("seed"), stop=b.get("stop") or [], stop_token_ids=b.get("stop_token_ids") or [],
        ignore_eos=bool(b.get("ignore_eos", False)), min_tokens=int(b.get("min_tokens", 0) or 0),
        logprobs=b.get("logprobs"), echo=bool(b.get("echo", False)),
        skip_special_tokens=bool(b.get("skip_special_tokens", True)))


def _error(message: str, typ: str = "invalid_request_error", code: int = 400):
    return JSONResponse({"error": {"message": message, "type": typ, "param": None, "code": code}}, status_code=code)


def build_app(aeng: AsyncEngine, model_name: str):
    if FastAPI is None:
        raise RuntimeError("fastapi is required for the HTTP server")
    app = FastAPI(title="dlti_b200 OpenAI-compatible server")
    created = int(time.time())

    @app.get("/health")
    async def health():
        return PlainTextResponse("")

    @app.get("/v1/models")
    async def models():
        return {"object": "list", "data": [{"id": model_name, "object": "model", "created": created,
                                             "owned_by": "dlti_b200", "max_model_len": aeng.engine.max_model_len}]}

    @app.get("/metrics")
    async def metrics():
        m, e = aeng.metrics, aeng.engine
        lines = [f"dlti_requests_total {m['requests_total']}", f"dlti_generation_tokens_total {m['tokens_total']}",
                 f"dlti_time_to_first_token_seconds_sum {m['ttft_sum']}", f"dlti_time_to_first_token_seconds_count {m['ttft_count']}",
                 f"dlti_num_requests_running {len(e.sched.running)}", f"dlti_num_requests_waiting {len(e.sched.waiting)}",
                 f"dlti_kv_cache_usage {e.blocks.usage():.4f}", f"dlti_num_preemptions_total {e.sched.num_preemptions}"]
        return PlainTextResponse("\n".join(lines) + "\n")

    @app.post("/v1/completions")
    async def completions(request: Request):
        try:
            body = await request.json()
        except Exception:
            return _error("request body is not valid JSON")
        if "prompt" not in body:
            return _error("'prompt' is required")
        if body.get("model") not in (None, model_name):
            return _error(f"The model `{body.get('model')}` does not exist.", "NotFoundError", 404)
        try:
            sp = _params_from_body(body)
        except Exception as e:
            return _error(str(e))
        prompt = body["prompt"]
        if isinstance(prompt, list) and prompt and isinstance(prompt[0], int):
            prompts = [prompt]                      # token ids
        elif isinstance(prompt, list) and prompt and isinstance(prompt[0], list):
            prompts = prompt
        elif isinstance(prompt, list):
            prompts = prompt
        else:
            prompts = [prompt]
        if not prompts:
            return _error("'prompt' is empty")
        base_id = f"cmpl-{uuid.uuid4().hex}"
        stream = bool(body.get("stream", False))
        include_usage = bool((body.get("stream_options") or {}).get("include_usage", False))
        jobs = [(i * sp.n + j, p) for i, p in enumerate(prompts) for j in range(sp.n)]
        tok = aeng.engine.tokenizer

        def chunk(idx, text, finish, lp=None):
            return {"id": base_id, "object": "text_completion", "created": int(time.time()), "model": model_name,
                    "choices": [{"index": idx, "text": text, "logprobs": lp, "finish_reason": finish}]}

        if stream:
            async def sse():
                q: asyncio.Queue = asyncio.Queue()
                usage = {"prompt_tokens": 0, "completion_tokens": 0}

                async def pump(idx, p):
                    rid = f"{base_id}-{idx}"
                    if sp.echo and isinstance(p, str):
                        await q.put(chunk(idx, p, None))
                    async for item in aeng.generate(p, SamplingParams(**vars(sp)), rid):
                        if "error" in item:
                            await q.put({"error": {"message": item["error"], "type": "invalid_request_error", "param": None, "code": 400}})
                            break
                        lp = {"token_logprobs": item["logprobs"]} if item.get("logprobs") else None
                        await q.put(chunk(idx, item["text"], item["finish_reason"], lp))
                        if item["finished"]:
                            usage["prompt_tokens"] += item["prompt_tokens"]
                            usage["completion_tokens"] += item["completion_tokens"]
                    await q.put(None)

                tasks = [asyncio.create_task(pump(i, p)) for i, p in jobs]
                done = 0
                try:
                    while done < len(tasks):
                        item = await q.get()
                        if item is None:
                            done += 1
                            continue
                        yield f"data: {json.dumps(item)}\n\n"
                    if include_usage:
                        u = dict(usage, total_tokens=usage["prompt_tokens"] + usage["completion_tokens"])
                        yield f"data: {json.dumps({'id': base_id, 'object': 'text_completion', 'created': int(time.time()), 'model': model_name, 'choices': [], 'usage': u})}\n\n"
                    yield "data: [DONE]\n\n"
                except asyncio.CancelledError:           # client went away: cancel the requests
                    for i, _ in jobs:
                        aeng.abort(f"{base_id}-{i}")
                    raise

            return StreamingResponse(sse(), media_type="text/event-stream")

        async def run_one(idx, p):
            rid = f"{base_id}-{idx}"
            text, lps, fin, pt, ct = (p if (sp.echo and isinstance(p, str)) else ""), [], None, 0, 0
            async for item in aeng.generate(p, SamplingParams(**vars(sp)), rid):
                if "error" in item:
                    raise ValueError(item["error"])
                text += item["text"]
                if item.get("logprobs"):
                    lps += item["logprobs"]
                if item["finished"]:
                    fin, pt, ct = item["finish_reason"], item["prompt_tokens"], item["completion_tokens"]
            return idx, text, fin, lps, pt, ct

        try:
            results = await asyncio.gather(*[run_one(i, p) for i, p in jobs])
        except ValueError as e:
            return _error(str(e))
        choices = [{"index": i, "text": t, "logprobs": ({"token_logprobs": l} if l else None), "finish_reason": f}
                   for i, t, f, l, _, _ in sorted(results)]
        pt = sum(r[4] for r in results)
        ct = sum(r[5] for r in results)
        return JSONResponse({"id": base_id, "object": "text_completion", "created": int(time.time()), "model": model_name,
                             "choices": choices, "usage": {"prompt_tokens": pt, "completion_tokens": ct, "total_tokens": pt + ct}})

    return app
