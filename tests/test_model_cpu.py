"""Hand-written Llama+LoRA forward/backward vs autograd oracle and vs HF transformers (CPU, fp32)."""
import pytest
import torch

from dlti_b200.models.config import PRESETS, LlamaConfig
from dlti_b200.models.llama import LoraLlama, ResidentWeights, pack_batch
from dlti_b200.models.lora import LoraConfig, LoraParams
from tests.ref_model import ref_loss


def _batch(cfg, B=3, S=24, seed=0, ragged=True):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, cfg.vocab_size, (B, S), generator=g)
    mask = torch.ones(B, S, dtype=torch.long)
    if ragged:
        mask[1, S - 7:] = 0
        mask[2, S - 1:] = 0
    labels = ids.clone()
    labels[mask == 0] = -100
    return pack_batch(ids, labels, mask)


@pytest.mark.parametrize("preset,ckpt,p", [("tiny", True, 0.0), ("tiny", False, 0.05), ("tiny-gqa", True, 0.05)])
def test_forward_backward_matches_autograd(preset, ckpt, p):
    cfg = PRESETS[preset]
    torch.manual_seed(0)
    w = ResidentWeights.random_init(cfg, "cpu", torch.float32, seed=1)
    for t in w.all_matrices():
        t.mul_(3.0)  # make activations non-trivial
    lcfg = LoraConfig(r=8, lora_alpha=16, lora_dropout=p)
    lora = LoraParams(cfg, lcfg, "cpu", torch.float32, seed=3)
    for n in lora.slots:  # non-zero B so every gradient path is exercised
        if n.endswith("lora_B"):
            lora.view(n).normal_(0, 0.05)
    lora.refresh_compute()
    model = LoraLlama(cfg, w, lora, gradient_checkpointing=ckpt, dropout_seed=11)
    b = _batch(cfg)
    gscale = 1.0 / b.n_valid
    model.micro_step = 5
    loss = model.forward_backward(b, gscale)
    rl, leaves, _ = ref_loss(cfg, w, lora, b, p, 11, 5)
    (rl * gscale).backward()
    assert abs(loss.item() - rl.item()) / rl.item() < 1e-4
    for n, leaf in leaves.items():
        got = lora.view(n, lora.grad)
        ref = leaf.grad.float()
        err = (got - ref).abs().max().item() / (ref.abs().max().item() + 1e-12)
        assert err < 2e-3, (n, err)


def test_logits_match_hf_transformers():
    transformers = pytest.importorskip("transformers")
    from dlti_b200.models.hf_io import resident_state_dict

    cfg = PRESETS["tiny-gqa"]
    w = ResidentWeights.random_init(cfg, "cpu", torch.float32, seed=2)
    hf_cfg = transformers.LlamaConfig(**{k: v for k, v in cfg.to_hf_dict().items()
                                         if k not in ("architectures", "model_type", "torch_dtype")})
    hf = transformers.LlamaForCausalLM(hf_cfg).to(torch.float32).eval()
    missing, unexpected = hf.load_state_dict(resident_state_dict(w), strict=False)
    assert not unexpected and all("rotary" in m for m in missing)
    model = LoraLlama(cfg, w, None)
    model.training = False
    b = _batch(cfg, B=2, S=16, ragged=False)
    mine = model.forward_logits(b).view(2, 16, -1)
    with torch.no_grad():
        theirs = hf(input_ids=b.input_ids.view(2, 16)).logits
    torch.testing.assert_close(mine, theirs, rtol=2e-4, atol=2e-4)
