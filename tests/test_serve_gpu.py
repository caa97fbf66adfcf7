"""Serving engine on the sm_100a kernels (GPU tier): paged decode + CUDA-graph replay vs eager, and
engine output vs a naive full re-forward through the training-side model (same weights)."""
import pytest
import torch

from dlti_b200.models.config import PRESETS
from dlti_b200.models.llama import LoraLlama, ResidentWeights, pack_batch
from dlti_b200.serve.engine import LLMEngine
from dlti_b200.serve.model_runner import TPContext, TPWeights
from dlti_b200.serve.params import SamplingParams
from dlti_b200.serve.tokenizer import ByteTokenizer

pytestmark = pytest.mark.gpu


def build(graphs, **kw):
    cfg = PRESETS["tiny-gqa"]
    full = ResidentWeights.random_init(cfg, "cuda", torch.bfloat16, seed=3)
    for t in full.all_matrices():
        t.mul_(4.0)
    w = TPWeights(cfg, 0, 1, "cuda", torch.bfloat16).load_from_full(full)
    eng = LLMEngine(cfg, w, TPContext(), ByteTokenizer(cfg.vocab_size), num_blocks=256, max_model_len=512,
                    use_cuda_graphs=graphs, **kw)
    return cfg, full, eng


def test_engine_graph_vs_eager_vs_naive(monkeypatch):
    import dlti_b200.ops.gemm as G
    prompts = [[1] + list(range(40, 40 + n)) for n in (5, 33, 130)]
    sp = SamplingParams(max_tokens=24, temperature=0.0, ignore_eos=True)
    # exact graph == eager equality needs a reproducible summation order: one K split per output tile (the default
    # stream-K decomposition reduces partial sums with atomics, whose order varies from run to run)
    monkeypatch.setattr(G, "_SKINNY_SPLIT", 1)
    _, _, eng_g1 = build(True)
    out_g1 = eng_g1.generate(prompts, sp)
    _, _, eng_e = build(False)
    out_e = eng_e.generate(prompts, sp)
    assert [o["token_ids"] for o in out_g1] == [o["token_ids"] for o in out_e]     # graph replay == eager
    monkeypatch.setattr(G, "_SKINNY_SPLIT", 0)
    cfg, full, eng_g = build(True)
    out_g = eng_g.generate(prompts, sp)
    assert eng_g.graphs and eng_g.blocks.num_free == eng_g.blocks.num_blocks
    # naive: re-run the whole sequence through the training-side forward each step
    m = LoraLlama(cfg, full, None)
    m.training = False
    agree = total = 0
    for p, o in zip(prompts, out_g):
        ids = list(p)
        for k in range(24):
            t = torch.tensor([ids])
            logits = m.forward_logits(pack_batch(t, t.clone(), None).to("cuda"))
            nxt = int(logits[-1].float().argmax())
            total += 1
            agree += int(nxt == o["token_ids"][k])
            ids.append(o["token_ids"][k])          # teacher-force the engine's token
    assert agree / total > 0.9, (agree, total)


def test_mixed_batches_staggered_arrivals():
    """Prompts arriving while others generate ride along with the decode rows (one forward pass with both the
    paged-decode and the varlen prefill attention kernels); tokens agree with the naive re-forward."""
    cfg, full, eng = build(True)
    assert eng.warmup(4) == 3 and sorted(eng.graphs) == [1, 2, 4]       # graphs captured before the first request
    sp = SamplingParams(max_tokens=20, temperature=0.0, ignore_eos=True)
    prompts = [[1] + list(range(40, 40 + n)) for n in (5, 33, 130, 300, 2)]
    arrive = {0: [0], 2: [1, 2], 5: [3], 9: [4]}
    rids, step = {}, 0
    while step < 10 or eng.has_unfinished():
        for i in arrive.get(step, []):
            rids[i] = eng.add_request(prompts[i], SamplingParams(**vars(sp)))
        eng.step()
        step += 1
    assert eng.stats["mixed_steps"] == 3 and eng.stats["prefill_steps"] == 1 and eng.stats["decode_steps"] > 0
    assert eng.blocks.num_free == eng.blocks.num_blocks
    m = LoraLlama(cfg, full, None)
    m.training = False
    agree = total = 0
    for i, p in enumerate(prompts):
        out = eng.requests[rids[i]].output_ids
        assert len(out) == 20
        ids = list(p)
        for k in range(20):
            t = torch.tensor([ids])
            logits = m.forward_logits(pack_batch(t, t.clone(), None).to("cuda"))
            total += 1
            agree += int(int(logits[-1].float().argmax()) == out[k])
            ids.append(out[k])
    assert agree / total > 0.9, (agree, total)
    # the prefill-first policy produces the same kind of text (not bit-identical: different GEMM paths per step)
    _, _, eng2 = build(True, mixed_batches=False)
    outs = eng2.generate(prompts, sp)
    same = sum(int(a == b) for i, o in enumerate(outs) for a, b in zip(o["token_ids"], eng.requests[rids[i]].output_ids))
    assert same / (20 * len(prompts)) > 0.8 and eng2.stats["mixed_steps"] == 0


CHUNK_SCRIPT = r"""
import sys, torch
sys.path.insert(0, {root!r})
from tests.test_serve_gpu import build
from dlti_b200.serve.params import SamplingParams
prompts = [[1] + list(range(40, 40 + n)) for n in (5, 200, 130, 70)]
sp = SamplingParams(max_tokens=12, temperature=0.0, ignore_eos=True)
_, _, ref = build(True)
want = [o["token_ids"] for o in ref.generate(prompts, sp)]
_, _, eng = build(True, prefill_chunk=64, max_num_batched_tokens=160)
got = [o["token_ids"] for o in eng.generate(prompts, sp)]
same = sum(a == b for x, y in zip(got, want) for a, b in zip(x, y))
print("agreement", same, "of", 12 * len(prompts), eng.stats)
assert eng.stats["prompt_tokens"] == sum(len(p) for p in prompts) and eng.stats["prefill_steps"] + eng.stats["mixed_steps"] >= 4
assert eng.blocks.num_free == eng.blocks.num_blocks
sys.exit(0 if same >= 0.9 * 12 * len(prompts) else 1)
"""


def test_chunked_prefill_gpu(tmp_path):
    """Opt-in chunked prefill (prefill_chunk > 0) on the GPU kernels: later chunks attend to the cached prefix through
    the extended [prefix | chunk] layout. Runs in its own process: the path had only been exercised on the CPU
    oracles when this test was written, and a device fault must not take the remaining tests with it. It has
    passed on B200 since round 1, so a failure (or a timeout) fails the test."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "chunk.py"
    script.write_text(CHUNK_SCRIPT.format(root=root))
    r = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0, "chunked prefill on GPU: " + (r.stdout + r.stderr)[-2500:]


@pytest.mark.parametrize("graphs", [False, True])
def test_unmerged_adapter_on_gpu_kernels(graphs):
    """X1 on the sm_100a kernels: LoRA branch as a second stream-K pass into the decode accumulator (q/k/v), as a
    K-extension of the prefill GEMM, and as extra K columns of o_proj — against the merged model through the naive
    training-side forward (fp32 weights merged, then rounded once)."""
    import copy
    from dlti_b200.models.lora import LoraConfig
    from dlti_b200.models.merge import merge_into_weights
    cfg = PRESETS["tiny-gqa"]
    full = ResidentWeights.random_init(cfg, "cuda", torch.bfloat16, seed=3)
    for t in full.all_matrices():
        t.mul_(4.0)
    lcfg = LoraConfig(r=16, lora_alpha=32, lora_dropout=0.0)
    g = torch.Generator().manual_seed(11)
    d, q, kv = cfg.hidden_size, cfg.q_size, cfg.kv_size
    state = {}
    for li in range(cfg.num_hidden_layers):
        for mod, (i, o) in {"q_proj": (d, q), "k_proj": (d, kv), "v_proj": (d, kv), "o_proj": (q, d)}.items():
            state[f"base_model.model.model.layers.{li}.self_attn.{mod}.lora_A.weight"] = torch.randn(16, i, generator=g) * 0.1
            state[f"base_model.model.model.layers.{li}.self_attn.{mod}.lora_B.weight"] = torch.randn(o, 16, generator=g) * 0.1
    merged = copy.deepcopy(full)
    merge_into_weights(merged, state, lcfg)
    w = TPWeights(cfg, 0, 1, "cuda", torch.bfloat16).load_from_full(full).attach_adapter(state, lcfg)
    eng = LLMEngine(cfg, w, TPContext(), ByteTokenizer(cfg.vocab_size), num_blocks=256, max_model_len=512,
                    use_cuda_graphs=graphs)
    wm = TPWeights(cfg, 0, 1, "cuda", torch.bfloat16).load_from_full(merged)
    engm = LLMEngine(cfg, wm, TPContext(), ByteTokenizer(cfg.vocab_size), num_blocks=256, max_model_len=512,
                     use_cuda_graphs=graphs)
    wb = TPWeights(cfg, 0, 1, "cuda", torch.bfloat16).load_from_full(full)
    engb = LLMEngine(cfg, wb, TPContext(), ByteTokenizer(cfg.vocab_size), num_blocks=256, max_model_len=512,
                     use_cuda_graphs=graphs)
    prompts = [[1] + list(range(40, 40 + n)) for n in (5, 33, 300)]
    sp = SamplingParams(max_tokens=16, temperature=0.0, ignore_eos=True)
    a = [o["token_ids"] for o in eng.generate(prompts, sp)]
    b = [o["token_ids"] for o in engm.generate(prompts, sp)]
    c = [o["token_ids"] for o in engb.generate(prompts, sp)]
    same = sum(x == y for p, r in zip(a, b) for x, y in zip(p, r))
    base_same = sum(x == y for p, r in zip(a, c) for x, y in zip(p, r))
    # the merged weights are rounded to bf16 once more than the separate branch, so a greedy sequence may part ways
    # with its twin at a near-tie (and then stays apart): the first tokens (prefill path) must agree, most sequences
    # must agree in full, and the adapter must not be a no-op
    assert [x[0] for x in a] == [x[0] for x in b], (a, b)
    assert sum(x == y for x, y in zip(a, b)) >= 2 and same >= 32, (same, a, b)
    assert base_same < same
