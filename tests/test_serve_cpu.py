"""Serving tier on CPU: block manager / scheduler invariants, engine vs naive re-forward (greedy
token identity), preemption, OpenAI HTTP schema + SSE framing."""
import json

import pytest
import torch

from dlti_b200.models.config import PRESETS
from dlti_b200.models.llama import LoraLlama, ResidentWeights, pack_batch
from dlti_b200.serve.engine import LLMEngine
from dlti_b200.serve.kv_cache import BlockManager
from dlti_b200.serve.model_runner import TPContext, TPWeights
from dlti_b200.serve.params import Request, SamplingParams
from dlti_b200.serve.scheduler import Scheduler
from dlti_b200.serve.tokenizer import ByteTokenizer, IncrementalDetokenizer


def make_engine(num_blocks=64, **kw):
    cfg = PRESETS["tiny"]
    full = ResidentWeights.random_init(cfg, "cpu", torch.float32, seed=3)
    for t in full.all_matrices():
        t.mul_(4.0)
    w = TPWeights(cfg, 0, 1, "cpu", torch.float32).load_from_full(full)
    eng = LLMEngine(cfg, w, TPContext(), ByteTokenizer(cfg.vocab_size), num_blocks=num_blocks, max_model_len=256,
                    use_cuda_graphs=False, **kw)
    return cfg, full, eng


def naive_greedy(cfg, full, prompt, n):
    m = LoraLlama(cfg, full, None)
    m.training = False
    ids = list(prompt)
    for _ in range(n):
        t = torch.tensor([ids])
        logits = m.forward_logits(pack_batch(t, t.clone(), None))
        ids.append(int(logits[-1].argmax()))
    return ids[len(prompt):]


def test_block_manager_and_scheduler_invariants():
    bm = BlockManager(8, 4)
    blocks = []
    assert bm.ensure(blocks, 9) and len(blocks) == 3 and bm.num_free == 5
    assert bm.ensure(blocks, 10) and len(blocks) == 3
    assert not bm.ensure([], 100)
    bm.release(blocks)
    assert bm.num_free == 8 and blocks == []
    sch = Scheduler(BlockManager(6, 4), max_num_seqs=4, max_num_batched_tokens=32, max_model_len=64)
    rs = [Request(f"r{i}", list(range(3, 10)), SamplingParams(max_tokens=8)) for i in range(4)]
    for r in rs:
        sch.add(r)
    b = sch.schedule()
    assert b.kind == "prefill" and [r.request_id for r in b.requests] == ["r0", "r1", "r2"]   # 2 blocks each
    for r in b.requests:
        r.output_ids.append(5)
    b2 = sch.schedule()        # no block left for r3, and the running three each need a 3rd block: the
    assert b2.kind == "decode" and sch.num_preemptions >= 1 and 1 <= len(b2.requests) < 3   # youngest is preempted
    assert sch.waiting[0].request_id in ("r2", "r1") and sch.waiting[0].blocks == []
    used = sum(len(r.blocks) for r in sch.running)
    assert used + sch.blocks.num_free == 6                                                    # no leaks
    sch.finish(sch.running[0], "length")
    assert sum(len(r.blocks) for r in sch.running) + sch.blocks.num_free == 6


def test_engine_greedy_matches_naive_reforward():
    cfg, full, eng = make_engine()
    prompts = [[1, 40, 41, 42, 43, 44], [1, 7, 8], [1] + list(range(100, 130))]
    outs = eng.generate(prompts, SamplingParams(max_tokens=6, temperature=0.0, ignore_eos=True))
    for p, o in zip(prompts, outs):
        assert o["token_ids"] == naive_greedy(cfg, full, p, 6), (p, o["token_ids"])
        assert o["finish_reason"] == "length"
    assert eng.blocks.num_free == eng.blocks.num_blocks and not eng.requests


def test_engine_preemption_is_transparent():
    cfg, full, eng = make_engine(num_blocks=6)            # 96 token slots for 3 sequences that need > 32 each
    prompts = [[1] + list(range(10 + i, 36 + i)) for i in range(3)]
    outs = eng.generate(prompts, SamplingParams(max_tokens=12, temperature=0.0, ignore_eos=True))
    assert eng.sched.num_preemptions > 0
    for p, o in zip(prompts, outs):
        assert o["token_ids"] == naive_greedy(cfg, full, p, 12)


def test_sampling_params_and_stop():
    cfg, full, eng = make_engine()
    o = eng.generate([[1, 50, 51]], SamplingParams(max_tokens=8, temperature=0.9, top_k=5, seed=7, ignore_eos=True))[0]
    o2 = eng.generate([[1, 50, 51]], SamplingParams(max_tokens=8, temperature=0.9, top_k=5, seed=7, ignore_eos=True))[0]
    assert len(o["token_ids"]) == 8 and o["token_ids"] == o2["token_ids"]
    g = naive_greedy(cfg, full, [1, 50, 51], 4)
    o3 = eng.generate([[1, 50, 51]], SamplingParams(max_tokens=8, temperature=0.0, stop_token_ids=[g[2]]))[0]
    assert o3["token_ids"] == g[:3] and o3["finish_reason"] == "stop"
    with pytest.raises(ValueError):
        SamplingParams(top_p=0.0)
    d = IncrementalDetokenizer(ByteTokenizer())
    euro = [b + 3 for b in "€".encode()]
    assert d.step(euro[:1]) == "" and d.step(euro[:2]) == "" and d.step(euro) == "€"


def test_openai_http_surface():
    pytest.importorskip("fastapi")
    from fastapi.testclient import TestClient

    from dlti_b200.serve.api_server import AsyncEngine, build_app

    cfg, full, eng = make_engine()
    aeng = AsyncEngine(eng)
    aeng.start()
    try:
        with TestClient(build_app(aeng, "tiny")) as c:
            assert c.get("/health").status_code == 200
            assert c.get("/v1/models").json()["data"][0]["id"] == "tiny"
            r = c.post("/v1/completions", json={"model": "tiny", "prompt": "hi", "max_tokens": 5, "temperature": 0,
                                                "ignore_eos": True})
            j = r.json()
            assert r.status_code == 200 and j["object"] == "text_completion" and j["id"].startswith("cmpl-")
            assert j["choices"][0]["finish_reason"] == "length" and j["choices"][0]["index"] == 0
            assert j["usage"] == {"prompt_tokens": 3, "completion_tokens": 5, "total_tokens": 8}
            # token-id prompt, n=2, batch of prompts
            r = c.post("/v1/completions", json={"prompt": [[1, 5, 6], [1, 9]], "max_tokens": 3, "n": 2, "temperature": 0,
                                                "ignore_eos": True})
            assert len(r.json()["choices"]) == 4 and r.json()["usage"]["completion_tokens"] == 12
            # SSE framing
            with c.stream("POST", "/v1/completions", json={"prompt": [1, 5, 6], "max_tokens": 4, "stream": True,
                                                           "temperature": 0, "ignore_eos": True,
                                                           "stream_options": {"include_usage": True}}) as s:
                lines = [l for l in s.iter_lines() if l]
            assert all(l.startswith("data: ") for l in lines) and lines[-1] == "data: [DONE]"
            chunks = [json.loads(l[6:]) for l in lines[:-1]]
            assert len([ch for ch in chunks if ch["choices"]]) == 4
            assert chunks[-1]["usage"]["completion_tokens"] == 4
            assert [ch["choices"][0]["finish_reason"] for ch in chunks if ch["choices"]][-1] == "length"
            # errors
            e = c.post("/v1/completions", json={"max_tokens": 3})
            assert e.status_code == 400 and e.json()["error"]["type"] == "invalid_request_error"
            assert c.post("/v1/completions", json={"model": "nope", "prompt": "x"}).status_code == 404
            assert "dlti_generation_tokens_total" in c.get("/metrics").text
    finally:
        aeng.stop = True
        aeng.thread.join(timeout=5)
