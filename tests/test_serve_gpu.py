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


def build(graphs):
    cfg = PRESETS["tiny-gqa"]
    full = ResidentWeights.random_init(cfg, "cuda", torch.bfloat16, seed=3)
    for t in full.all_matrices():
        t.mul_(4.0)
    w = TPWeights(cfg, 0, 1, "cuda", torch.bfloat16).load_from_full(full)
    eng = LLMEngine(cfg, w, TPContext(), ByteTokenizer(cfg.vocab_size), num_blocks=256, max_model_len=512,
                    use_cuda_graphs=graphs)
    return cfg, full, eng


def test_engine_graph_vs_eager_vs_naive():
    prompts = [[1] + list(range(40, 40 + n)) for n in (5, 33, 130)]
    sp = SamplingParams(max_tokens=24, temperature=0.0, ignore_eos=True)
    cfg, full, eng_g = build(True)
    out_g = eng_g.generate(prompts, sp)
    _, _, eng_e = build(False)
    out_e = eng_e.generate(prompts, sp)
    assert [o["token_ids"] for o in out_g] == [o["token_ids"] for o in out_e]      # graph replay == eager
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
