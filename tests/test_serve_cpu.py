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


def test_scheduler_budget_and_sequence_limits():
    """Admission respects the token budget (decode rows count one token each), the sequence limit and FCFS order; a
    prompt larger than the whole budget is still served, alone; prefill-first mode never mixes."""
    def reqs(lens):
        return [Request(f"r{i}", list(range(3, 3 + n)), SamplingParams(max_tokens=4)) for i, n in enumerate(lens)]

    sch = Scheduler(BlockManager(64, 4), max_num_seqs=3, max_num_batched_tokens=20, max_model_len=64)
    for r in reqs([8, 8, 8, 8]):
        sch.add(r)
    b = sch.schedule()
    assert b.kind == "prefill" and [r.request_id for r in b.requests] == ["r0", "r1"]       # 8 + 8 <= 20 < 24
    for r in b.requests:
        r.output_ids.append(5)
    b = sch.schedule()                                                                     # 2 decode rows + r2 (18 <= 20)
    assert b.kind == "mixed" and b.num_decode == 2 and [r.request_id for r in b.requests] == ["r0", "r1", "r2"]
    for r in b.requests:
        r.output_ids.append(5)
    b = sch.schedule()                                                                     # max_num_seqs = 3: r3 waits
    assert b.kind == "decode" and len(b.requests) == 3 and sch.waiting[0].request_id == "r3"
    sch.finish(sch.running[0], "length")
    b = sch.schedule()
    assert b.kind == "mixed" and b.requests[-1].request_id == "r3"
    # a prompt over the budget runs alone, and only when nothing else is in the step
    sch = Scheduler(BlockManager(64, 4), max_num_seqs=4, max_num_batched_tokens=16, max_model_len=64)
    small, big = reqs([4, 30])
    sch.add(small)
    sch.add(big)
    b = sch.schedule()
    assert [r.request_id for r in b.requests] == ["r0"]
    small.output_ids.append(5)
    assert sch.schedule().kind == "decode"                                                 # big does not fit next to a decode row
    sch.finish(small, "length")
    b = sch.schedule()
    assert b.kind == "prefill" and b.requests == [big]
    # prefill-first policy
    sch = Scheduler(BlockManager(64, 4), max_num_seqs=4, max_num_batched_tokens=64, max_model_len=64, mixed=False)
    a, c = reqs([6, 6])
    sch.add(a)
    assert sch.schedule().kind == "prefill"
    a.output_ids.append(5)
    sch.add(c)
    b = sch.schedule()
    assert b.kind == "prefill" and b.requests == [c] and b.num_decode == 0
    c.output_ids.append(5)
    b = sch.schedule()
    assert b.kind == "decode" and len(b.requests) == 2
    with pytest.raises(ValueError):
        sch.add(Request("long", list(range(80)), SamplingParams()))


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


@pytest.mark.parametrize("mixed", [True, False])
def test_staggered_arrivals_mixed_batches(mixed):
    """Prompts that arrive while other sequences are generating join the running step (mixed=True: decode rows
    and prompt rows in one forward pass; False: prefill-first steps). Tokens match the naive oracle either way."""
    cfg, full, eng = make_engine(mixed_batches=mixed)
    sp = SamplingParams(max_tokens=10, temperature=0.0, ignore_eos=True)
    prompts = [[1, 40, 41, 42, 43, 44], [1] + list(range(100, 130)), [1, 7, 8], [1] + list(range(60, 81)), [1, 9]]
    arrive = {0: [0], 2: [1, 2], 3: [3], 7: [4]}                 # engine step -> prompts that arrive before it
    rids, kinds, step = {}, [], 0
    while step < 4 or eng.has_unfinished():
        for i in arrive.get(step, []):
            rids[i] = eng.add_request(prompts[i], SamplingParams(**vars(sp)))
        before = dict(eng.stats)
        eng.step()
        kinds += [k for k in ("prefill", "decode", "mixed") if eng.stats[k + "_steps"] > before[k + "_steps"]]
        step += 1
    assert ("mixed" in kinds) == mixed and kinds[0] == "prefill"
    if mixed:
        assert kinds.count("mixed") == 3 and "prefill" not in kinds[1:]
    for i, p in enumerate(prompts):
        assert eng.requests[rids[i]].output_ids == naive_greedy(cfg, full, p, 10), i
    assert eng.blocks.num_free == eng.blocks.num_blocks


@pytest.mark.parametrize("num_blocks", [64, 4])
def test_async_scheduling_matches_synchronous(num_blocks):
    """async_scheduling launches step N+1 before the host has read step N (decode rows are fed from the
    device-side samples). Text, finish reasons and log-probs must not depend on it: greedy and seeded sampling,
    ends by stop token (discovered one step late: the extra row is discarded), by length, under preemption
    (4 blocks) and with requests arriving mid-flight."""
    def run(async_mode):
        cfg, full, eng = make_engine(num_blocks=num_blocks, async_scheduling=async_mode)
        g = naive_greedy(cfg, full, [1, 50, 51], 6)
        jobs = [([1, 50, 51], SamplingParams(max_tokens=12, temperature=0.0, stop_token_ids=[g[3]])),
                ([1] + list(range(10, 36)), SamplingParams(max_tokens=14, temperature=0.0, ignore_eos=True, logprobs=1)),
                ([1] + list(range(7, 20)), SamplingParams(max_tokens=9, temperature=0.8, top_k=6, seed=11, ignore_eos=True)),
                ([1] + list(range(60, 95)), SamplingParams(max_tokens=10, temperature=0.0, ignore_eos=True)),
                ([1, 5], SamplingParams(max_tokens=1, temperature=0.0, ignore_eos=True))]
        arrive = {0: [0, 1, 2], 4: [3, 4]}
        rids, step, n_out = {}, 0, 0
        while step < 5 or eng.has_unfinished():
            for i in arrive.get(step, []):
                rids[i] = eng.add_request(*jobs[i])
            n_out += len(eng.step())
            step += 1
        res = [(eng.requests[rids[i]].output_ids, eng.requests[rids[i]].finish_reason,
                [round(x, 4) for x in eng.requests[rids[i]].logprobs]) for i in range(len(jobs))]
        assert n_out == sum(len(r[0]) for r in res)
        assert eng.blocks.num_free == eng.blocks.num_blocks and eng._inflight is None
        assert all(r.pending == 0 for r in eng.requests.values())
        return res, eng.sched.num_preemptions, g

    sync, _, g = run(False)
    asyn, preempt, _ = run(True)
    assert sync == asyn
    assert sync[0][0] == g[:4] and sync[0][1] == "stop" and sync[4][1] == "length" and len(sync[4][0]) == 1
    assert len(sync[1][2]) == 14 and (preempt > 0) == (num_blocks == 4)


@pytest.mark.parametrize("async_mode", [True, False])
def test_abort_while_a_step_is_in_flight(async_mode):
    """A request cancelled between two steps (client went away) may still have a row in the step that is in flight:
    its sample is dropped, its blocks are returned once, the other sequences are unaffected."""
    cfg, full, eng = make_engine(async_scheduling=async_mode)
    sp = SamplingParams(max_tokens=8, temperature=0.0, ignore_eos=True)
    keep = eng.add_request([1, 40, 41, 42], SamplingParams(**vars(sp)))
    drop = eng.add_request([1] + list(range(50, 70)), SamplingParams(**vars(sp)))
    for _ in range(3):
        eng.step()
    n_before = len(eng.requests[drop].output_ids)
    dropped = eng.requests[drop]
    eng.abort(drop)
    assert drop not in eng.requests and dropped.finish_reason == "abort"
    late = eng.add_request([1, 7, 8], SamplingParams(**vars(sp)))          # reuses the freed blocks
    while eng.has_unfinished():
        eng.step()
    assert len(dropped.output_ids) == n_before and dropped.blocks == []
    assert eng.requests[keep].output_ids == naive_greedy(cfg, full, [1, 40, 41, 42], 8)
    assert eng.requests[late].output_ids == naive_greedy(cfg, full, [1, 7, 8], 8)
    assert eng.blocks.num_free == eng.blocks.num_blocks and eng._inflight is None
    assert len(set(eng.blocks.free)) == eng.blocks.num_blocks             # no block released twice


_ORACLE_CACHE: dict = {}


@pytest.mark.parametrize("seed", list(range(24)) + [121, 123, 140, 146, 160, 162, 180, 203, 221, 255, 266, 284])
def test_engine_randomised_schedules_match_oracle(seed):
    """Randomised end-to-end check of the scheduling machinery: random block pool / sequence limit / token budget,
    sync or async stepping, mixed or prefill-first steps, chunked prefill on or off, random prompt lengths around
    the block size, staggered arrivals, stop tokens, one-token requests, a mid-flight abort. Every surviving
    request must produce exactly the oracle's greedy tokens and finish reason; no block may leak or be freed
    twice. (The listed seeds above 24 are ones that preempt.)"""
    import random
    rng = random.Random(seed)
    nb = rng.choice([5, 6, 8, 12, 40])
    mode, mixed = rng.random() < 0.7, rng.random() < 0.7
    chunk = rng.choice([0, 0, 8, 16, 24])
    budget = rng.choice([32, 48, 64, 8192]) if chunk else rng.choice([64, 8192])
    apc = rng.random() < 0.5
    cfg, full, eng = make_engine(num_blocks=nb, async_scheduling=mode, mixed_batches=mixed, prefill_chunk=chunk,
                                 max_num_batched_tokens=budget, max_num_seqs=rng.choice([2, 3, 8]),
                                 enable_prefix_caching=apc)
    shared = [1] + [rng.randrange(3, 200) for _ in range(40)]          # prompts often start with the same tokens

    def oracle(p, n):
        key = (tuple(p), n)
        if key not in _ORACLE_CACHE:
            _ORACLE_CACHE[key] = naive_greedy(cfg, full, p, n)
        return _ORACLE_CACHE[key]

    jobs = []
    for _ in range(rng.randint(2, 6)):
        n = min(rng.choice([2, 3, 7, 15, 16, 17, 30, 45]), nb * 16 - 20)
        p = [1] + [rng.randrange(3, 200) for _ in range(n - 1)]
        if rng.random() < 0.5:
            k = rng.choice([16, 17, 32, 33])
            p = (shared[:k] + p[1:])[:n] if n > 4 else p
        mt = rng.choice([1, 2, 5, 9, 14])
        jobs.append((p, 1 if len(p) + mt + 1 > nb * 16 else mt, rng.randint(0, 8)))
    rids, aborted, step = {}, set(), 0
    while step < 10 or eng.has_unfinished():
        for i, (p, mt, at) in enumerate(jobs):
            if at == step:
                stop = [oracle(p, mt)[mt // 2]] if (i % 3 == 0 and mt > 2) else []
                rids[i] = eng.add_request(p, SamplingParams(max_tokens=mt, temperature=0.0, ignore_eos=True,
                                                            stop_token_ids=stop))
        if step == 6 and rids and rng.random() < 0.3:
            i = rng.choice(list(rids))
            r = eng.requests.get(rids[i])
            if r is not None and r.status != "finished":
                eng.abort(rids[i])
                aborted.add(i)
        eng.step()
        step += 1
        assert step < 2000, "no progress"
    for i, (p, mt, _) in enumerate(jobs):
        if i in aborted:
            continue
        r, g = eng.requests[rids[i]], oracle(p, mt)
        if r.params.stop_token_ids:
            k = g.index(r.params.stop_token_ids[0]) + 1
            assert r.output_ids == g[:k] and r.finish_reason == "stop", (i, r.output_ids, g)
        else:
            assert r.output_ids == g and r.finish_reason == "length", (i, r.output_ids, g)
    assert eng.blocks.num_free == nb and len(set(eng.blocks.free)) == nb and not any(eng.blocks.ref)
    assert not eng.sched.waiting and not eng.sched.running and eng._inflight is None


def test_prefix_caching_shares_blocks_and_stays_exact():
    """enable_prefix_caching: a prompt that starts like an earlier one takes references on the earlier prompt's full
    blocks and only computes the rest (its attention reads the shared prefix from the cache). Outputs stay equal to
    the oracle, shared blocks are neither freed early nor leaked, and cached blocks are recycled when memory runs
    out."""
    cfg, full, eng = make_engine(num_blocks=12, enable_prefix_caching=True)
    sp = SamplingParams(max_tokens=6, temperature=0.0, ignore_eos=True)
    sys_prompt = [1] + list(range(30, 30 + 49))                      # 50 tokens = 3 full blocks + 2
    a, b, c = sys_prompt + [200, 201], sys_prompt + [210], sys_prompt[:35] + [220, 221, 222]
    out_a = eng.generate([a], sp)[0]["token_ids"]
    assert out_a == naive_greedy(cfg, full, a, 6) and eng.blocks.hits == 0
    assert len(eng.blocks.cached) == 3 and eng.blocks.num_free == 12   # published, free, still findable
    before = eng.stats["prompt_tokens"]
    rb, rc = eng.add_request(b, SamplingParams(**vars(sp))), eng.add_request(c, SamplingParams(**vars(sp)))
    eng.step()
    assert eng.requests[rb].blocks[:3] == eng.requests[rc].blocks[:2] + [eng.requests[rb].blocks[2]]
    assert [eng.blocks.ref[x] for x in eng.requests[rb].blocks[:3]] == [2, 2, 1]
    assert eng.stats["prompt_tokens"] - before == (len(b) - 48) + (len(c) - 32) and eng.blocks.hits == 80
    while eng.has_unfinished():
        eng.step()
    assert eng.requests[rb].output_ids == naive_greedy(cfg, full, b, 6)
    assert eng.requests[rc].output_ids == naive_greedy(cfg, full, c, 6)
    assert eng.blocks.num_free == 12 and not any(eng.blocks.ref)
    # memory pressure: unrelated prompts recycle the cached blocks; a later hit on a recycled prefix is a miss
    fill = [[1] + list(range(100 + 7 * i, 100 + 7 * i + 60)) for i in range(3)]
    outs = eng.generate(fill, sp)
    for p, o in zip(fill, outs):
        assert o["token_ids"] == naive_greedy(cfg, full, p, 6)
    hits = eng.blocks.hits
    assert eng.generate([a], sp)[0]["token_ids"] == out_a
    assert eng.blocks.hits - hits < 48 and eng.blocks.num_free == 12 and not any(eng.blocks.ref)
    assert len(eng.blocks.cached) == len(eng.blocks.hash_of) <= 12


def test_step_layout_roundtrip():
    import numpy as np
    from dlti_b200.serve.engine import StepLayout
    for kw, want in [(dict(), dict(ids=11, slots=11, last=5, seeds=5, pslots=0, pdst=0, csrc=0, pos=11, src=11, ctx=3,
                                   cu=3, ecu=0, topk=5, temp=5, topp=5)),
                     (dict(ns=1, P=6), dict(ids=11, slots=11, last=4, seeds=4, pslots=6, pdst=6, csrc=8, pos=11, src=11,
                                            ctx=3, cu=3, ecu=3, topk=4, temp=4, topp=4))]:
        lay = StepLayout(T=11, nd=3, npf=2, mb=8, **kw)
        buf = np.zeros(lay.words, dtype=np.int64)
        v = lay.views(buf, np)
        assert {k: len(a) for k, a in v.items() if k != "tables"} == want
        assert v["tables"].shape == (3, 8) and v["temp"].dtype == np.float32 and v["pos"].dtype == np.int32
        for i, (k, a) in enumerate(v.items()):                       # sections do not overlap
            a[...] = i + 1
        for i, (k, a) in enumerate(v.items()):
            assert (a == i + 1).all(), k
        t = lay.views(torch.from_numpy(buf), torch)
        assert t["tables"].shape == (3, 8) and int(t["cu"][0]) == v["cu"][0] and t["topp"].dtype == torch.float32


@pytest.mark.parametrize("num_blocks", [10, 7])
@pytest.mark.parametrize("async_mode,mixed", [(True, True), (False, True), (True, False)])
def test_chunked_prefill_matches_whole_prompt_prefill(async_mode, mixed, num_blocks):
    """prefill_chunk: prompts enter the cache a few tokens per step (their later chunks attend to the cached
    prefix through the extended [prefix | chunk] attention layout) while other sequences keep decoding. Same
    tokens as the naive oracle, including a prompt longer than the step's whole token budget and (7 blocks)
    preemption — also of a sequence that is preempted while its final token is still in flight: it must end when
    that token is read, not be prefilled again."""
    cfg, full, eng = make_engine(num_blocks=num_blocks, async_scheduling=async_mode, mixed_batches=mixed,
                                 prefill_chunk=16, max_num_batched_tokens=40)
    sp = SamplingParams(max_tokens=8, temperature=0.0, ignore_eos=True)
    prompts = [[1, 40, 41, 42], [1] + list(range(100, 170)), [1] + list(range(60, 93)), [1, 9, 8]]
    arrive = {0: [0], 1: [1, 2], 6: [3]} if num_blocks == 10 else {0: [0, 1, 2, 3]}
    rids, step = {}, 0
    while step < 7 or eng.has_unfinished():
        for i in arrive.get(step, []):
            rids[i] = eng.add_request(prompts[i], SamplingParams(**vars(sp)))
        before = eng.stats["generated_tokens"], eng.stats["prompt_tokens"]
        eng.step()
        step += 1
        assert eng.stats["prompt_tokens"] - before[1] <= 40
    for i, p in enumerate(prompts):
        assert eng.requests[rids[i]].output_ids == naive_greedy(cfg, full, p, 8), i
    assert eng.stats["prompt_tokens"] >= sum(len(p) for p in prompts)          # 71 tokens took >= 5 chunks
    if num_blocks == 7 and async_mode:
        assert eng.sched.num_preemptions > 0
    assert eng.blocks.num_free == eng.blocks.num_blocks and eng._inflight is None and not eng.sched.waiting
    assert all(r.pending == 0 for r in eng.requests.values())


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
            # chat flavour: messages -> Llama-2 chat prompt -> the same engine path
            from dlti_b200.serve.api_server import llama2_chat_prompt
            msgs = [{"role": "system", "content": "be brief"}, {"role": "user", "content": "hi"},
                    {"role": "assistant", "content": "hello"}, {"role": "user", "content": "bye"}]
            ptxt = llama2_chat_prompt(msgs)
            assert ptxt == "[INST] <<SYS>>\nbe brief\n<</SYS>>\n\nhi [/INST] hello </s><s>[INST] bye [/INST]"
            r = c.post("/v1/chat/completions", json={"messages": msgs, "max_tokens": 4, "temperature": 0, "ignore_eos": True})
            j = r.json()
            assert r.status_code == 200 and j["object"] == "chat.completion" and j["id"].startswith("chatcmpl-")
            assert j["choices"][0]["message"]["role"] == "assistant" and j["usage"]["completion_tokens"] == 4
            r2 = c.post("/v1/completions", json={"prompt": ptxt, "max_tokens": 4, "temperature": 0, "ignore_eos": True})
            assert r2.json()["choices"][0]["text"].strip() == j["choices"][0]["message"]["content"]
            with c.stream("POST", "/v1/chat/completions", json={"messages": msgs[1:2], "max_tokens": 3, "stream": True,
                                                                 "temperature": 0, "ignore_eos": True}) as s:
                cl = [l for l in s.iter_lines() if l]
            cc = [json.loads(l[6:]) for l in cl[:-1]]
            assert cl[-1] == "data: [DONE]" and cc[0]["choices"][0]["delta"]["role"] == "assistant"
            assert all(x["object"] == "chat.completion.chunk" for x in cc) and cc[-1]["choices"][0]["finish_reason"] == "length"
            assert c.post("/v1/chat/completions", json={"messages": [{"role": "assistant", "content": "x"}]}).status_code == 400
            # errors
            e = c.post("/v1/completions", json={"max_tokens": 3})
            assert e.status_code == 400 and e.json()["error"]["type"] == "invalid_request_error"
            assert c.post("/v1/completions", json={"model": "nope", "prompt": "x"}).status_code == 404
            mt = c.get("/metrics").text
            assert "dlti_generation_tokens_total" in mt
            # histograms: every finished request is in the +Inf bucket of the latency histograms
            import re
            n_req = int(re.search(r"dlti_requests_total (\d+)", mt).group(1))
            assert int(re.search(r'dlti_e2e_request_latency_seconds_bucket\{le="\+Inf"\} (\d+)', mt).group(1)) == n_req - 0
            assert 'dlti_time_to_first_token_seconds_bucket{le="0.005"}' in mt and "dlti_time_per_output_token_seconds_sum" in mt
    finally:
        aeng.stop = True
        aeng.thread.join(timeout=5)


def test_two_process_frontend():
    """HTTP front-end in its own process, engine loop here, one pipe message per engine step: same answers as the
    in-process server, streaming included; the front-end's /metrics reflects the engine's counters."""
    pytest.importorskip("fastapi")
    import socket
    import threading
    import time
    import urllib.request

    from dlti_b200.serve.api_server import engine_process_loop, start_frontend_process

    cfg, full, eng = make_engine()
    with socket.socket() as so:
        so.bind(("127.0.0.1", 0))
        port = so.getsockname()[1]
    proc, conn = start_frontend_process(eng, "127.0.0.1", port, "tiny")
    stop = threading.Event()
    th = threading.Thread(target=engine_process_loop, args=(eng, conn, stop.is_set), daemon=True)
    th.start()
    url = f"http://127.0.0.1:{port}"
    try:
        for _ in range(600):
            try:
                urllib.request.urlopen(url + "/health", timeout=1)
                break
            except Exception:
                assert proc.is_alive()
                time.sleep(0.1)

        def post(body):
            req = urllib.request.Request(url + "/v1/completions", data=json.dumps(body).encode(),
                                         headers={"Content-Type": "application/json"})
            return urllib.request.urlopen(req, timeout=60)

        prompt = [1, 40, 41, 42, 43, 44]
        j = json.loads(post({"prompt": prompt, "max_tokens": 6, "temperature": 0, "ignore_eos": True}).read())
        assert j["usage"] == {"prompt_tokens": 6, "completion_tokens": 6, "total_tokens": 12}
        want = ByteTokenizer(cfg.vocab_size).decode(naive_greedy(cfg, full, prompt, 6))
        assert j["choices"][0]["text"] == want and j["choices"][0]["finish_reason"] == "length"
        lines = [l.decode().strip() for l in post({"prompt": prompt, "max_tokens": 6, "temperature": 0, "ignore_eos": True,
                                                   "stream": True}) if l.strip()]
        assert lines[-1] == "data: [DONE]"
        chunks = [json.loads(l[6:]) for l in lines[:-1]]
        assert len(chunks) == 6 and "".join(c["choices"][0]["text"] for c in chunks) == want
        try:
            post({"prompt": list(range(3, 400)), "max_tokens": 4})              # longer than max_model_len
            raise AssertionError("expected HTTP 400")
        except urllib.error.HTTPError as e:
            assert e.code == 400
        m = urllib.request.urlopen(url + "/metrics", timeout=5).read().decode()
        assert "dlti_generation_tokens_total 12" in m and "dlti_num_requests_running 0" in m
        assert 'dlti_e2e_request_latency_seconds_bucket{le="+Inf"} 2' in m            # histograms cross the pipe too
        # engine side goes away: the front-end turns unhealthy and fails new requests instead of hanging
        stop.set()
        th.join(timeout=5)
        conn.close()
        for _ in range(100):
            try:
                urllib.request.urlopen(url + "/health", timeout=2)
                time.sleep(0.05)
            except urllib.error.HTTPError as e:
                assert e.code == 503
                break
        else:
            raise AssertionError("/health stayed 200 after the engine side closed")
        try:
            post({"prompt": prompt, "max_tokens": 2})
            raise AssertionError("expected an error response")
        except urllib.error.HTTPError as e:
            assert e.code == 400 and "engine process exited" in e.read().decode()
    finally:
        stop.set()
        th.join(timeout=5)
        proc.terminate()
        proc.join(timeout=5)
    assert eng.blocks.num_free == eng.blocks.num_blocks


TP_WORKER = r"""
import os, sys, json, torch
sys.path.insert(0, {root!r})
from dlti_b200.models.config import PRESETS
from dlti_b200.parallel import comm
from dlti_b200.serve.engine import LLMEngine
from dlti_b200.serve.model_runner import TPContext, TPWeights
from dlti_b200.serve.params import SamplingParams
from dlti_b200.serve.tokenizer import ByteTokenizer
info = comm.init_distributed("cpu")
cfg = PRESETS["tiny"]
w = TPWeights(cfg, info.rank, info.world_size, "cpu", torch.float32).init_random(5)
for l in w.layers:
    for k in ("w_qkv", "w_o", "w_gu", "w_down"):
        l[k].mul_(4.0)
w.embed.mul_(4.0); w.lm_head.mul_(4.0)
eng = LLMEngine(cfg, w, TPContext(info, "nccl", None, cfg), ByteTokenizer(cfg.vocab_size), info=info, num_blocks=6,
                max_model_len=256, use_cuda_graphs=False, async_scheduling=(sys.argv[2] == "1"),
                prefill_chunk=int(sys.argv[3]), max_num_batched_tokens=48)
if info.rank != 0:
    eng.worker_loop()
else:
    prompts = [[1] + list(range(40, 40 + n)) for n in (5, 33, 20)]
    sps = [SamplingParams(max_tokens=12, temperature=0.0, ignore_eos=True),
           SamplingParams(max_tokens=9, temperature=0.0, ignore_eos=True, logprobs=1),
           SamplingParams(max_tokens=10, temperature=0.7, top_k=8, seed=3, ignore_eos=True)]
    out = eng.generate(prompts, sps)
    eng.shutdown_workers()
    json.dump({{"tokens": [o["token_ids"] for o in out], "preempt": eng.sched.num_preemptions}}, open(sys.argv[1], "w"))
comm.shutdown()
"""


@pytest.mark.parametrize("async_mode,chunk", [("1", 0), ("0", 0), ("1", 16)])
def test_tensor_parallel_engine_over_gloo(tmp_path, async_mode, chunk):
    """Host-side TP plumbing on CPU (gloo, world size 2): weight slicing, the packed-step broadcast, the worker loop
    with its own sampling + device-fed tokens, log-probs, preemption (6 blocks) — TP=2 must generate what one
    process generates from the same seeded weights."""
    import os
    import socket
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "tpw.py"
    script.write_text(TP_WORKER.format(root=root))
    env = dict(os.environ, PYTHONPATH=root, OMP_NUM_THREADS="2")
    outs = {}
    for world in (1, 2):
        out = tmp_path / f"w{world}.json"
        if world == 1:
            cmd = [sys.executable, str(script), str(out), async_mode, str(chunk)]
            e = dict(env, WORLD_SIZE="1", RANK="0", LOCAL_RANK="0")
        else:
            with socket.socket() as so:
                so.bind(("127.0.0.1", 0))
                port = so.getsockname()[1]
            cmd = [sys.executable, "-m", "dlti_b200.launch", "--num_gpus=2", f"--master_port={port}", "--no_local_rank",
                   str(script), str(out), async_mode, str(chunk)]
            e = env
        r = subprocess.run(cmd, env=e, capture_output=True, text=True, timeout=300, cwd=root)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        outs[world] = json.load(open(out))
    assert outs[1]["tokens"] == outs[2]["tokens"]
    assert [len(t) for t in outs[2]["tokens"]] == [12, 9, 10]


def test_load_generator_against_live_server(tmp_path):
    """The load generator as its own process against the two-process server (with prefix caching on): closed loop
    with unique prompts, closed loop with a shared system prompt (the server reports prefix-cache hits), and
    open-loop Poisson arrivals; every level reports the full metric set and no errors."""
    pytest.importorskip("fastapi")
    pytest.importorskip("aiohttp")
    import os
    import socket
    import subprocess
    import sys
    import threading
    import time
    import urllib.request

    from dlti_b200.serve.api_server import engine_process_loop, start_frontend_process

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cfg, full, eng = make_engine(enable_prefix_caching=True)
    with socket.socket() as so:
        so.bind(("127.0.0.1", 0))
        port = so.getsockname()[1]
    proc, conn = start_frontend_process(eng, "127.0.0.1", port, "tiny")
    stop = threading.Event()
    th = threading.Thread(target=engine_process_loop, args=(eng, conn, stop.is_set), daemon=True)
    th.start()
    url = f"http://127.0.0.1:{port}"
    try:
        for _ in range(600):
            try:
                urllib.request.urlopen(url + "/health", timeout=1)
                break
            except Exception:
                time.sleep(0.1)

        def load(*extra):
            out = tmp_path / "lg.json"
            r = subprocess.run([sys.executable, "-m", "dlti_b200.serve.loadgen", "--url", url, "--model", "tiny",
                                "--input_len", "40", "--output_len", "5", "--vocab", str(cfg.vocab_size), "--out", str(out),
                                *extra], capture_output=True, text=True, timeout=120, cwd=root,
                               env=dict(os.environ, PYTHONPATH=root))
            assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
            return json.load(open(out))

        res = load("--concurrency", "1,3", "--requests_per_level", "6")
        assert [l["concurrency"] for l in res] == [1, 3]
        for l in res:
            assert l["requests"] == 6 and l["errors"] == 0 and l["output_tokens_per_s"] > 0
            assert 0 < l["ttft_p50_ms"] <= l["ttft_p99_ms"] <= l["e2e_p99_ms"] and l["itl_mean_ms"] > 0
        assert eng.blocks.hits == 0                                   # unique prompts: nothing to share
        res = load("--concurrency", "2", "--requests_per_level", "6", "--shared_prefix_len", "33", "--think_time", "0.01")
        assert res[0]["errors"] == 0 and res[0]["shared_prefix_len"] == 33 and eng.blocks.hits >= 32 * 4
        res = load("--concurrency", "4", "--requests_per_level", "8", "--request_rate", "200")
        assert res[0]["requests"] == 8 and res[0]["errors"] == 0 and res[0]["request_rate"] == 200
        m = urllib.request.urlopen(url + "/metrics", timeout=5).read().decode()
        assert "dlti_prefix_cache_hits_tokens_total" in m
    finally:
        stop.set()
        th.join(timeout=5)
        proc.terminate()
        proc.join(timeout=5)
    assert eng.blocks.num_free == eng.blocks.num_blocks and not any(eng.blocks.ref)


class _LetterTokenizer(ByteTokenizer):
    """One lowercase letter per token: a stop string of n characters always spans n tokens."""

    def decode(self, ids, skip_special_tokens=True):
        return "".join(chr(97 + i % 26) for i in ids)


@pytest.mark.parametrize("async_sched", [False, True])
def test_stop_strings_end_generation_and_are_not_streamed(async_sched):
    """OpenAI `stop` strings: a match that spans several tokens ends the request with finish_reason "stop", the
    text is cut before the match, and no later text reaches a streaming client."""
    cfg, full, eng = make_engine(async_scheduling=async_sched)
    eng.tokenizer = _LetterTokenizer(cfg.vocab_size)
    prompt = [1, 50, 51]
    ref = eng.generate([prompt], SamplingParams(max_tokens=24, temperature=0.0, ignore_eos=True))[0]
    text = ref["text"]
    assert len(text) == 24 and ref["finish_reason"] == "length"
    # a 3-token stop string whose first occurrence starts at character k
    k = next(i for i in range(4, 16) if text.find(text[i:i + 3]) == i)
    stop = text[k:k + 3]
    # non-streaming
    o = eng.generate([prompt], SamplingParams(max_tokens=24, temperature=0.0, ignore_eos=True, stop=[stop, "~~~~~~"]))[0]
    assert o["finish_reason"] == "stop" and o["text"] == text[:k]
    assert k + 3 <= len(o["token_ids"]) <= k + 4          # ended at the match (async: at most one token late)
    # streaming: pieces as the server hands them out, one per engine step
    rid = eng.add_request(prompt, SamplingParams(max_tokens=24, temperature=0.0, ignore_eos=True, stop=[stop, "~~~~~~"]))
    pieces, reason = [], None
    while eng.has_unfinished():
        for out in eng.step():
            pieces.append(eng.text_delta(out.request_id, final=out.finished))
            if out.finished:
                reason = out.finish_reason
    assert reason == "stop" and "".join(pieces) == text[:k]
    assert all(stop not in "".join(pieces[:i + 1]) for i in range(len(pieces)))
    eng.release(rid)
    # a stop string that never occurs changes nothing, and the held-back tail is flushed at the end
    o = eng.generate([prompt], SamplingParams(max_tokens=24, temperature=0.0, ignore_eos=True, stop=["~~~~~~"]))[0]
    assert o["finish_reason"] == "length" and o["text"] == text


def test_unmerged_adapter_serving_matches_merged_weights():
    """X1: base + PEFT adapter served unmerged (LoRA branch inside the QKV / o_proj GEMMs) gives the logits of the
    merged model W' = W + (alpha/r) B A, for whole prompts and for decode steps through the paged cache."""
    import copy
    from dlti_b200.models.lora import LoraConfig
    from dlti_b200.models.merge import merge_into_weights
    cfg = PRESETS["tiny"]
    full = ResidentWeights.random_init(cfg, "cpu", torch.float32, seed=3)
    for t in full.all_matrices():
        t.mul_(4.0)
    lcfg = LoraConfig(r=8, lora_alpha=16, lora_dropout=0.0)
    g = torch.Generator().manual_seed(11)
    d, q, kv = cfg.hidden_size, cfg.q_size, cfg.kv_size
    state = {}
    for li in range(cfg.num_hidden_layers):
        for mod, (i, o) in {"q_proj": (d, q), "k_proj": (d, kv), "v_proj": (d, kv), "o_proj": (q, d)}.items():
            state[f"base_model.model.model.layers.{li}.self_attn.{mod}.lora_A.weight"] = torch.randn(8, i, generator=g) * 0.2
            state[f"base_model.model.model.layers.{li}.self_attn.{mod}.lora_B.weight"] = torch.randn(o, 8, generator=g) * 0.2
    merged = copy.deepcopy(full)
    merge_into_weights(merged, state, lcfg)

    def engine(weights, adapter):
        w = TPWeights(cfg, 0, 1, "cpu", torch.float32).load_from_full(weights)
        if adapter:
            w.attach_adapter(state, lcfg)
        return LLMEngine(cfg, w, TPContext(), ByteTokenizer(cfg.vocab_size), num_blocks=64, max_model_len=256,
                         use_cuda_graphs=False)

    prompts = [[1] + list(range(40, 40 + n)) for n in (5, 23)]
    sp = SamplingParams(max_tokens=10, temperature=0.0, ignore_eos=True, logprobs=1)
    a = engine(full, True).generate(prompts, sp)
    b = engine(merged, False).generate(prompts, sp)
    c = engine(full, False).generate(prompts, sp)
    assert [o["token_ids"] for o in a] == [o["token_ids"] for o in b]
    assert [o["token_ids"] for o in a] != [o["token_ids"] for o in c]          # the adapter does change the model


def test_detokenize_false_returns_token_ids_only():
    cfg, full, eng = make_engine()
    o = eng.generate([[1, 50, 51]], SamplingParams(max_tokens=6, temperature=0.0, ignore_eos=True, detokenize=False))[0]
    assert len(o["token_ids"]) == 6 and o["text"] == ""
    o2 = eng.generate([[1, 50, 51]], SamplingParams(max_tokens=6, temperature=0.0, ignore_eos=True))[0]
    assert o2["token_ids"] == o["token_ids"] and o2["text"] != ""
