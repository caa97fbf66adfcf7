"""Helpers shared by the GPU integration tests (not collected as tests themselves)."""
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_short(info, stage: int, backend: str, fused, steps: int = 4, preset: str = "tiny-gqa", seq: int = 256,
              with_norms: bool = False, with_master: bool = False):
    from dlti_b200.config.args import TrainingArgs
    from dlti_b200.engine.trainer import Trainer
    from dlti_b200.models.config import PRESETS
    from dlti_b200.models.llama import pack_batch
    from dlti_b200.models.lora import LoraConfig

    cfg = PRESETS[preset]
    args = TrainingArgs(output_dir="/tmp/dlti_mg", per_device_train_batch_size=2, gradient_accumulation_steps=2,
                        learning_rate=5e-3, precision="bf16", device="cuda", backend=backend,
                        deepspeed=os.path.join(ROOT, "configs", f"ds_config_zero{stage}.json") if stage else None,
                        no_offload=True, max_steps=steps + 2)
    tr = Trainer(cfg, LoraConfig(r=16, lora_alpha=32, lora_dropout=0.05), args, None, None, info=info)
    if fused is not None and tr.fused is None:
        raise RuntimeError("fused backend requested but not active")
    losses, norms = [], []
    g = torch.Generator().manual_seed(7 + info.rank)
    for s in range(steps):
        mbs = []
        for _ in range(2):
            ids = torch.randint(3, cfg.vocab_size, (2, seq), generator=g)
            mbs.append(pack_batch(ids, ids.clone(), None, pin=True))
        losses.append(tr.train_step(mbs))
        norms.append(tr.last_grad_norm)
    torch.cuda.synchronize()
    if tr.fused is not None:
        tr.fused.check_error()
    if with_master:
        return losses, norms, tr.lora.master.detach().clone()
    return (losses, norms) if with_norms else losses
