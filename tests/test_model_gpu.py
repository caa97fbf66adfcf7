"""Llama+LoRA training step on the sm_100a kernels (bf16) vs the fp64 autograd oracle (GPU tier)."""
import pytest
import torch

from dlti_b200.models.config import PRESETS
from dlti_b200.models.llama import LoraLlama, ResidentWeights, pack_batch
from dlti_b200.models.lora import LoraConfig, LoraParams
from dlti_b200.ops.gemm import gemm, gemm_ref
from tests.ref_model import ref_loss

pytestmark = pytest.mark.gpu


def _batch(cfg, B, S, seed=0, ragged=True):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, cfg.vocab_size, (B, S), generator=g)
    mask = torch.ones(B, S, dtype=torch.long)
    if ragged:
        mask[1, S - 37:] = 0
    labels = ids.clone()
    labels[mask == 0] = -100
    return pack_batch(ids, labels, mask)


def test_gemm_small_inner_extent():
    """TN GEMMs used for dA/dB: operand inner extent (16) smaller than the 64-wide TMA box."""
    torch.manual_seed(0)
    T, d, r = 1024, 512, 16
    dy = torch.randn(T, d, device="cuda", dtype=torch.bfloat16)
    t = torch.randn(T, 64, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(d, r, device="cuda")
    gemm(dy, t[:, 16:32], a_mn=True, b_mn=True, out=out, accumulate=True, alpha=2.0, split_k=2, block_n=64)
    ref = gemm_ref(dy, t[:, 16:32], a_mn=True, b_mn=True, alpha=2.0, out_dtype=torch.float32)
    assert (out - ref).abs().max() / ref.abs().max() < 1e-2
    x = torch.randn(T, d, device="cuda", dtype=torch.bfloat16)
    out2 = torch.zeros(48, d, device="cuda")
    gemm(t[:, :48], x, a_mn=True, b_mn=True, out=out2, accumulate=True, split_k=2, block_n=128)
    ref2 = gemm_ref(t[:, :48], x, a_mn=True, b_mn=True, out_dtype=torch.float32)
    assert (out2 - ref2).abs().max() / ref2.abs().max() < 1e-2


@pytest.mark.parametrize("preset,ckpt,p", [("tiny", True, 0.05), ("tiny-gqa", False, 0.0), ("tiny-gqa", True, 0.05)])
def test_train_step_matches_oracle(preset, ckpt, p):
    cfg = PRESETS[preset]
    w_cpu = ResidentWeights.random_init(cfg, "cpu", torch.float32, seed=1)
    for t in w_cpu.all_matrices():
        t.mul_(2.0)
    w = ResidentWeights(cfg, "cuda", torch.bfloat16)
    for a, b in zip(list(w.all_matrices()) + list(w.all_norms()),
                    list(w_cpu.all_matrices()) + list(w_cpu.all_norms())):
        a.copy_(b)
        b.copy_(a.float().cpu())  # oracle sees the bf16-rounded weights
    lcfg = LoraConfig(r=16, lora_alpha=32, lora_dropout=p)
    lora = LoraParams(cfg, lcfg, "cuda", torch.bfloat16, seed=3)
    lora_cpu = LoraParams(cfg, lcfg, "cpu", torch.float32, seed=3)
    g = torch.Generator().manual_seed(9)
    for n in lora.slots:
        if n.endswith("lora_B"):
            v = torch.randn(lora.slots[n].shape, generator=g) * 0.05
            lora.view(n).copy_(v)
            lora_cpu.view(n).copy_(v)
    lora.refresh_compute()
    model = LoraLlama(cfg, w, lora, gradient_checkpointing=ckpt, dropout_seed=11)
    b = _batch(cfg, 2, 200)
    gs = torch.tensor([1.0 / b.n_valid], device="cuda")
    model.micro_step = 2
    loss = model.forward_backward(b.to("cuda"), gs)
    torch.cuda.synchronize()
    rl, leaves, _ = ref_loss(cfg, w_cpu, lora_cpu, b, p, 11, 2)
    (rl / b.n_valid).backward()
    assert abs(loss.item() - rl.item()) / rl.item() < 2e-2
    worst = 0.0
    for n, leaf in leaves.items():
        got = lora.view(n, lora.grad).float().cpu()
        ref = leaf.grad.float()
        err = (got - ref).abs().max().item() / (ref.abs().max().item() + 1e-12)
        worst = max(worst, err)
        assert err < 8e-2, (n, err)
    print("worst relative-to-max grad error", worst)
