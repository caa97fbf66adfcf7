"""OpenAI-compatible HTTP front-end: POST /v1/completions and POST
/v1/chat/completions (JSON or SSE stream), GET /v1/models, GET /health, GET /metrics (Prometheus text). Schema follows what vLLM serves
(SP:vllm/entrypoints/openai/completion/protocol.py:42-90,466-545); SSE framing is
`data: {json}\\n\\n` ... `data: [DONE]\\n\\n`. The reference only declares this capability
(reference README.md:10,16; requirements.txt:18).

Two deployments of the same app: AsyncEngine runs the engine loop in a thread of the HTTP process (requests
over a thread-safe queue, tokens through per-request asyncio queues); EngineProxy + engine_process_loop put the
HTTP server in a process of its own, connected to the engine process by a pipe (one message per engine step) —
the layout vLLM uses for its API server / engine core (SP:vllm/v1/engine/core_client.py), so the engine never
competes with request handling for the interpreter lock.
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


# Prometheus histograms (vLLM exposes the same three: time to first token, time per output token, end-to-end
# request latency). Plain dicts so the two-process front-end can receive them through the pipe.
_HIST = {"ttft": (0.005, 0.01, 0.02, 0.04, 0.08, 0.16, 0.32, 0.64, 1.28, 2.56, 5.12, 10.0),
         "tpot": (0.001, 0.002, 0.003, 0.004, 0.006, 0.008, 0.012, 0.016, 0.024, 0.032, 0.064, 0.128),
         "e2e": (0.05, 0.1, 0.2, 0.4, 0.8, 1.6, 3.2, 6.4, 12.8, 25.6, 51.2, 102.4)}
_HIST_NAMES = {"ttft": "dlti_time_to_first_token_seconds", "tpot": "dlti_time_per_output_token_seconds",
               "e2e": "dlti_e2e_request_latency_seconds"}


def _new_metrics() -> dict:
    m = {"requests_total": 0, "tokens_total": 0, "ttft_sum": 0.0, "ttft_count": 0}
    for k, edges in _HIST.items():
        m["hist_" + k] = {"counts": [0] * (len(edges) + 1), "sum": 0.0}
    return m


def _observe(metrics: dict, key: str, value: float) -> None:
    h = metrics["hist_" + key]
    edges = _HIST[key]
    i = 0
    while i < len(edges) and value > edges[i]:
        i += 1
    h["counts"][i] += 1
    h["sum"] += value


def _render_histograms(m: dict) -> list[str]:
    lines = []
    for key, edges in _HIST.items():
        h = m.get("hist_" + key)
        if not h:
            continue
        name = _HIST_NAMES[key]
        lines.append(f"# TYPE {name} histogram")
        cum = 0
        for e, c in zip(edges, h["counts"]):
            cum += c
            lines.append(f'{name}_bucket{{le="{e}"}} {cum}')
        cum += h["counts"][-1]
        lines.append(f'{name}_bucket{{le="+Inf"}} {cum}')
        lines.append(f"{name}_sum {h['sum']}")
        lines.append(f"{name}_count {cum}")
    return lines


def _engine_stats(eng: LLMEngine, metrics: dict) -> dict:
    return dict(metrics, running=len(eng.sched.running), waiting=len(eng.sched.waiting),
                kv_usage=eng.blocks.usage(), preemptions=eng.sched.num_preemptions,
                max_model_len=eng.max_model_len, prefix_hits=eng.blocks.hits, prefix_queries=eng.blocks.queries)


def _handle_command(eng: LLMEngine, kind, payload, emit_error) -> None:
    if kind == "add":
        rid, prompt, params = payload
        try:
            eng.add_request(prompt, params, rid)
        except Exception as e:  # bad request (too long, ...)
            emit_error(rid, str(e))
    elif kind == "abort":
        eng.abort(payload)


def _step_items(eng: LLMEngine, metrics: dict) -> list:
    """One engine step -> [(request id, stream item)] for every request that produced a token."""
    batch = []
    for out in eng.step():
        rid = out.request_id
        r = eng.requests.get(rid)
        if r is None:
            continue
        if len(r.output_ids) == 1:
            metrics["ttft_sum"] += time.time() - r.arrival
            metrics["ttft_count"] += 1
            _observe(metrics, "ttft", time.time() - r.arrival)
        if out.finished and r.finish_time is not None:
            _observe(metrics, "e2e", r.finish_time - r.arrival)
            if len(r.output_ids) > 1 and r.first_token_time is not None:
                _observe(metrics, "tpot", (r.finish_time - r.first_token_time) / (len(r.output_ids) - 1))
        metrics["tokens_total"] += len(out.new_token_ids)
        delta = eng.text_delta(rid, final=out.finished)
        batch.append((rid, {"text": delta, "token_ids": out.new_token_ids, "finished": out.finished,
                            "finish_reason": out.finish_reason, "logprobs": out.logprobs,
                            "prompt_tokens": len(r.prompt_ids), "completion_tokens": len(r.output_ids)}))
        if out.finished:
            eng.release(rid)
    return batch


class _Streams:
    """asyncio side shared by the in-process and the two-process front-ends: per-request queues fed from
    another thread, one hand-over to the event loop per engine step."""

    def __init__(self):
        self.streams: dict[str, tuple[asyncio.AbstractEventLoop, asyncio.Queue]] = {}

    def _emit_many(self, items):
        by_loop: dict = {}
        for rid, item in items:
            s = self.streams.get(rid)
            if s:
                by_loop.setdefault(s[0], []).append((s[1], item))
        for loop, pairs in by_loop.items():
            loop.call_soon_threadsafe(self._deliver, pairs)

    @staticmethod
    def _deliver(pairs):
        for q, item in pairs:
            q.put_nowait(item)

    def _submit(self, kind, payload):
        raise NotImplementedError

    async def generate(self, prompt, params: SamplingParams, rid: str):
        q: asyncio.Queue = asyncio.Queue()
        self.streams[rid] = (asyncio.get_running_loop(), q)
        self._submit("add", (rid, prompt, params))
        try:
            while True:
                item = await q.get()
                yield item
                if item.get("finished") or "error" in item:
                    return
        finally:
            self.streams.pop(rid, None)

    def abort(self, rid: str):
        self._submit("abort", rid)


class AsyncEngine(_Streams):
    """Engine loop in a thread of the HTTP server's process (tests, notebooks, small deployments)."""

    def __init__(self, engine: LLMEngine):
        super().__init__()
        self.engine = engine
        self.inbox: queue.Queue = queue.Queue()
        self.stop = False
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.metrics = _new_metrics()

    def start(self):
        # the engine thread re-acquires the GIL after every device sync; with the default 5 ms switch interval a
        # busy HTTP thread can hold it that long, which is most of a decode step
        import sys
        sys.setswitchinterval(min(sys.getswitchinterval(), 2e-4))
        self.thread.start()

    def stats(self) -> dict:
        return _engine_stats(self.engine, self.metrics)

    def _submit(self, kind, payload):
        if kind == "add":
            self.metrics["requests_total"] += 1
        self.inbox.put((kind, payload))

    def _loop(self):
        import torch
        if self.engine.device.type == "cuda":
            torch.cuda.set_device(self.engine.device)
        eng = self.engine
        while not self.stop:
            try:
                while True:
                    kind, payload = self.inbox.get_nowait() if eng.has_unfinished() else self.inbox.get(timeout=0.05)
                    _handle_command(eng, kind, payload, lambda rid, msg: self._emit_many([(rid, {"error": msg})]))
            except queue.Empty:
                pass
            if not eng.has_unfinished():
                continue
            self._emit_many(_step_items(eng, self.metrics))
        eng.shutdown_workers()


class EngineProxy(_Streams):
    """Front-end half of the two-process server: lives in the HTTP process, talks to the engine process over a
    pipe. The engine's per-step host work then never waits on the HTTP threads for the interpreter lock, and the
    SSE encoding of step N runs while the GPU computes step N+1."""

    def __init__(self, conn, max_model_len: int):
        super().__init__()
        self.conn = conn
        self._stats = {"requests_total": 0, "tokens_total": 0, "ttft_sum": 0.0, "ttft_count": 0, "running": 0,
                       "waiting": 0, "kv_usage": 0.0, "preemptions": 0, "max_model_len": max_model_len}
        self._lock = threading.Lock()
        self.dead = False
        self.thread = threading.Thread(target=self._reader, daemon=True)

    def start(self):
        self.thread.start()

    def stats(self) -> dict:
        return dict(self._stats)

    def _submit(self, kind, payload):
        rid = payload[0] if kind == "add" else payload
        if kind == "add":
            payload = (rid, payload[1], vars(payload[2]))
        try:
            if self.dead:
                raise OSError("engine process exited")
            with self._lock:
                self.conn.send((kind, payload))
        except (OSError, ValueError, BrokenPipeError):
            self.dead = True
            if kind == "add":
                self._emit_many([(rid, {"error": "engine process exited"})])

    def _reader(self):
        while True:
            try:
                items, stats = self.conn.recv()
            except (EOFError, OSError):
                # the engine process is gone: fail the open requests instead of leaving them waiting, and
                # report unhealthy so a supervisor restarts the deployment
                self.dead = True
                self._emit_many([(rid, {"error": "engine process exited"}) for rid in list(self.streams)])
                return
            self._stats.update(stats)
            self._emit_many(items)


def engine_process_loop(eng: LLMEngine, conn, should_stop=lambda: False) -> None:
    """Engine half of the two-process server (runs in the process that owns the GPU): commands come in and
    per-step outputs go out over `conn`, one message per engine step."""
    metrics = _new_metrics()
    errors: list = []
    try:
        while not should_stop():
            ready = conn.poll(0 if eng.has_unfinished() else 0.05)
            while ready:
                kind, payload = conn.recv()
                if kind == "add":
                    metrics["requests_total"] += 1
                    payload = (payload[0], payload[1], SamplingParams(**payload[2]))
                _handle_command(eng, kind, payload, lambda rid, msg: errors.append((rid, {"error": msg})))
                ready = conn.poll(0)
            items = errors[:]
            errors.clear()
            if eng.has_unfinished():
                items += _step_items(eng, metrics)
            if items:
                conn.send((items, _engine_stats(eng, metrics)))
    except (EOFError, OSError, BrokenPipeError):
        pass
    eng.shutdown_workers()


def frontend_main(conn, host: str, port: int, model_name: str, max_model_len: int) -> None:
    """Entry point of the HTTP process (spawned by serve_two_process)."""
    import uvicorn
    proxy = EngineProxy(conn, max_model_len)
    proxy.start()
    uvicorn.run(build_app(proxy, model_name), host=host, port=port, log_level="warning")


def start_frontend_process(eng: LLMEngine, host: str, port: int, model_name: str):
    """Spawn the HTTP front-end as its own process; returns (process, engine-side connection). The caller then
    runs engine_process_loop(eng, conn) in the thread that owns the CUDA context."""
    import multiprocessing as mp
    import os
    ctx = mp.get_context("spawn")
    parent, child = ctx.Pipe(duplex=True)
    saved = os.environ.get("CUDA_VISIBLE_DEVICES")
    os.environ["CUDA_VISIBLE_DEVICES"] = ""              # the front-end never touches a GPU
    try:
        proc = ctx.Process(target=frontend_main, args=(child, host, port, model_name, eng.max_model_len), daemon=True)
        proc.start()
    finally:
        if saved is None:
            os.environ.pop("CUDA_VISIBLE_DEVICES", None)
        else:
            os.environ["CUDA_VISIBLE_DEVICES"] = saved
    child.close()
    return proc, parent


def _params_from_body(b: dict) -> SamplingParams:
    return SamplingParams(
        max_tokens=b.get("max_tokens", 16) or 16, temperature=float(b.get("temperature", 1.0) if b.get("temperature") is not None else 1.0),
        top_p=float(b.get("top_p", 1.0) or 1.0), top_k=int(b.get("top_k", 0) or 0), n=int(b.get("n", 1) or 1),
        seed=b.get("seed"), stop=b.get("stop") or [], stop_token_ids=b.get("stop_token_ids") or [],
        ignore_eos=bool(b.get("ignore_eos", False)), min_tokens=int(b.get("min_tokens", 0) or 0),
        logprobs=b.get("logprobs"), echo=bool(b.get("echo", False)),
        skip_special_tokens=bool(b.get("skip_special_tokens", True)),
        detokenize=bool(b.get("detokenize", True)))


def _error(message: str, typ: str = "invalid_request_error", code: int = 400):
    return JSONResponse({"error": {"message": message, "type": typ, "param": None, "code": code}}, status_code=code)


def llama2_chat_prompt(messages: list[dict]) -> str:
    """OpenAI `messages` -> Llama-2 chat prompt text (the format the reference trains on:
    `<s>[INST] question [/INST] answer</s>`, reference scripts/prepare_dataset.py:24-32, with Meta's <<SYS>>
    block for a system message). The tokenizer adds the leading BOS, so the first `<s>` is not written here;
    the text ends after the last `[/INST]` so the model continues with the assistant turn."""
    if not messages or not isinstance(messages, list):
        raise ValueError("'messages' must be a non-empty list")
    system = ""
    turns: list[tuple[str, str]] = []
    for m_ in messages:
        role, content = m_.get("role"), m_.get("content")
        if not isinstance(content, str) or role not in ("system", "user", "assistant"):
            raise ValueError("each message needs a role of system|user|assistant and string content")
        if role == "system":
            system = content
        else:
            turns.append((role, content))
    if not turns or turns[-1][0] != "user":
        raise ValueError("the last message must come from the user")
    out, first, i = "", True, 0
    while i < len(turns):
        role, content = turns[i]
        if role != "user":
            raise ValueError("user and assistant messages must alternate, starting with the user")
        q = content.strip()
        if first and system:
            q = f"<<SYS>>\n{system.strip()}\n<</SYS>>\n\n{q}"
        out += ("" if first else "<s>") + f"[INST] {q} [/INST]"
        first = False
        if i + 1 < len(turns):
            if turns[i + 1][0] != "assistant":
                raise ValueError("user and assistant messages must alternate, starting with the user")
            out += f" {turns[i + 1][1].strip()} </s>"
        i += 2
    return out


def build_app(aeng: AsyncEngine, model_name: str):
    if FastAPI is None:
        raise RuntimeError("fastapi is required for the HTTP server")
    app = FastAPI(title="dlti_b200 OpenAI-compatible server")
    created = int(time.time())

    @app.get("/health")
    async def health():
        if getattr(aeng, "dead", False):          # two-process layout: the engine side of the pipe is closed
            return PlainTextResponse("engine process exited", status_code=503)
        return PlainTextResponse("")

    @app.get("/v1/models")
    async def models():
        return {"object": "list", "data": [{"id": model_name, "object": "model", "created": created,
                                             "owned_by": "dlti_b200", "max_model_len": aeng.stats()["max_model_len"]}]}

    @app.get("/metrics")
    async def metrics():
        m = aeng.stats()
        lines = [f"dlti_requests_total {m['requests_total']}", f"dlti_generation_tokens_total {m['tokens_total']}",
                 f"dlti_time_to_first_token_seconds_sum {m['ttft_sum']}", f"dlti_time_to_first_token_seconds_count {m['ttft_count']}",
                 f"dlti_num_requests_running {m['running']}", f"dlti_num_requests_waiting {m['waiting']}",
                 f"dlti_kv_cache_usage {m['kv_usage']:.4f}", f"dlti_num_preemptions_total {m['preemptions']}"]
        lines += [f"dlti_prefix_cache_hits_tokens_total {m.get('prefix_hits', 0)}",
                  f"dlti_prefix_cache_queries_tokens_total {m.get('prefix_queries', 0)}"]
        lines += _render_histograms(m)
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

    @app.post("/v1/chat/completions")
    async def chat_completions(request: Request):
        """Chat flavour of the same engine path: messages -> Llama-2 chat prompt -> completion. Response objects
        follow the OpenAI chat schema (`chat.completion` / `chat.completion.chunk` with `delta`)."""
        try:
            body = await request.json()
        except Exception:
            return _error("request body is not valid JSON")
        if body.get("model") not in (None, model_name):
            return _error(f"The model `{body.get('model')}` does not exist.", "NotFoundError", 404)
        try:
            prompt = llama2_chat_prompt(body.get("messages"))
            b2 = dict(body)
            if b2.get("max_tokens") is None:
                b2["max_tokens"] = body.get("max_completion_tokens") or 256
            b2["logprobs"] = None
            sp = _params_from_body(b2)
        except Exception as e:
            return _error(str(e))
        base_id = f"chatcmpl-{uuid.uuid4().hex}"
        stream = bool(body.get("stream", False))
        include_usage = bool((body.get("stream_options") or {}).get("include_usage", False))
        created_at = int(time.time())

        def obj(kind, choices, usage=None):
            o = {"id": base_id, "object": kind, "created": created_at, "model": model_name, "choices": choices}
            if usage is not None:
                o["usage"] = usage
            return o

        if stream:
            async def sse():
                q: asyncio.Queue = asyncio.Queue()
                usage = {"prompt_tokens": 0, "completion_tokens": 0}

                async def pump(idx):
                    rid = f"{base_id}-{idx}"
                    await q.put(obj("chat.completion.chunk", [{"index": idx, "delta": {"role": "assistant", "content": ""},
                                                               "finish_reason": None}]))
                    async for item in aeng.generate(prompt, SamplingParams(**vars(sp)), rid):
                        if "error" in item:
                            await q.put({"error": {"message": item["error"], "type": "invalid_request_error", "param": None, "code": 400}})
                            break
                        await q.put(obj("chat.completion.chunk", [{"index": idx, "delta": {"content": item["text"]},
                                                                   "finish_reason": item["finish_reason"]}]))
                        if item["finished"]:
                            usage["prompt_tokens"] += item["prompt_tokens"]
                            usage["completion_tokens"] += item["completion_tokens"]
                    await q.put(None)

                tasks = [asyncio.create_task(pump(i)) for i in range(sp.n)]
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
                        yield f"data: {json.dumps(obj('chat.completion.chunk', [], u))}\n\n"
                    yield "data: [DONE]\n\n"
                except asyncio.CancelledError:
                    for i in range(sp.n):
                        aeng.abort(f"{base_id}-{i}")
                    raise

            return StreamingResponse(sse(), media_type="text/event-stream")

        async def run_one(idx):
            text, fin, pt, ct = "", None, 0, 0
            async for item in aeng.generate(prompt, SamplingParams(**vars(sp)), f"{base_id}-{idx}"):
                if "error" in item:
                    raise ValueError(item["error"])
                text += item["text"]
                if item["finished"]:
                    fin, pt, ct = item["finish_reason"], item["prompt_tokens"], item["completion_tokens"]
            return idx, text, fin, pt, ct

        try:
            results = await asyncio.gather(*[run_one(i) for i in range(sp.n)])
        except ValueError as e:
            return _error(str(e))
        choices = [{"index": i, "message": {"role": "assistant", "content": t.strip()}, "finish_reason": f}
                   for i, t, f, _, _ in sorted(results)]
        pt, ct = sum(r[3] for r in results), sum(r[4] for r in results)
        return JSONResponse(obj("chat.completion", choices, {"prompt_tokens": pt, "completion_tokens": ct,
                                                             "total_tokens": pt + ct}))

    return app
