"""Engine tier on CPU: BASELINE.json config 1 plumbing (tiny Llama, LoRA r=8, fp32,
seq 128: runs to completion and the loss decreases), resume == uninterrupted, fp16 scaler path,
ZeRO stages vs single-process AdamW over gloo (world size 2), consolidated save from all ranks."""
import json
import os
import subprocess
import sys

import pytest
import torch

from dlti_b200.config.args import TrainingArgs
from dlti_b200.data.collate import CausalLMCollator
from dlti_b200.data.synthetic import SyntheticInstructionDataset
from dlti_b200.engine import checkpoint as ckpt
from dlti_b200.engine.trainer import Trainer
from dlti_b200.models.config import PRESETS
from dlti_b200.models.lora import LoraConfig

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def make_trainer(tmp, **kw):
    cfg = PRESETS["tiny"]
    ds = SyntheticInstructionDataset(kw.pop("n", 48), cfg.vocab_size, max_length=kw.pop("max_length", 64), seed=1)
    a = dict(output_dir=str(tmp), num_train_epochs=1, per_device_train_batch_size=2,
             gradient_accumulation_steps=2, learning_rate=5e-3, device="cpu", precision="fp32",
             logging_steps=2, save_strategy="steps", save_steps=4, save_total_limit=2, backend="nccl")
    lcfg = LoraConfig(r=8, lora_alpha=16, lora_dropout=kw.pop("lora_dropout", 0.05))
    a.update(kw)
    return Trainer(cfg, lcfg, TrainingArgs(**a), ds, CausalLMCollator(cfg.eos_token_id))


def test_baseline_config1_loss_decreases(tmp_path):
    tr = make_trainer(tmp_path, n=64, max_length=128, num_train_epochs=3, save_strategy="epoch")
    res = tr.train()
    losses = [e["loss"] for e in tr.log_history]
    assert res["global_step"] == 48 and len(losses) >= 10
    assert sum(losses[-3:]) / 3 < sum(losses[:3]) / 3 - 0.05, losses
    assert set(tr.log_history[0]) >= {"loss", "grad_norm", "learning_rate", "epoch"}
    cks = [p.name for p in ckpt.list_checkpoints(tmp_path)]
    assert cks == ["checkpoint-32", "checkpoint-48"]                 # save per epoch, keep 2
    tr.save_model(tmp_path / "final")
    assert (tmp_path / "final" / "adapter_model.safetensors").exists()
    st = json.load(open(tmp_path / "checkpoint-48" / "trainer_state.json"))
    assert st["global_step"] == 48 and (tmp_path / "checkpoint-48" / "optimizer.pt").exists()


def test_resume_equals_uninterrupted(tmp_path):
    a = make_trainer(tmp_path / "a")
    a.train()
    ref = a.lora.master.clone()
    b = make_trainer(tmp_path / "b")
    b.stop_at_step = 8                                                # "crash" right after the step-8 checkpoint
    b.train()
    assert ckpt.latest_checkpoint(tmp_path / "b").name == "checkpoint-8"
    c = make_trainer(tmp_path / "b")
    c.train(resume_from_checkpoint=True)
    assert c.global_step == a.global_step == 12
    torch.testing.assert_close(c.lora.master, ref, rtol=1e-6, atol=1e-7)
    assert [e["loss"] for e in c.log_history] == [e["loss"] for e in a.log_history]


def test_fp16_scaler_skips_overflow_step(tmp_path):
    z2 = os.path.join(ROOT, "configs/ds_config_zero2.json")
    tr = make_trainer(tmp_path, deepspeed=z2, precision=None, gradient_accumulation_steps=1)
    assert tr.precision == "fp16" and tr.scaler.dynamic and tr.scaler.cur_scale == 65536
    batches = list(tr.loader.iter_epoch(0))
    before = tr.lora.master.clone()
    tr.scaler.cur_scale = float("inf")                                  # force a non-finite gradient
    tr.train_step(batches[:1])
    assert tr.last_skipped and torch.equal(tr.lora.master, before) and tr.optimizer.step_count == 0
    tr.scaler.cur_scale = 65536.0
    tr.train_step(batches[1:2])
    assert not tr.last_skipped and tr.optimizer.step_count == 1
    # DeepSpeed WarmupLR with warmup 0: the first two steps run at lr 0 (scheduler not yet warmed)
    tr.train_step(batches[2:3])
    assert torch.equal(tr.lora.master, before) and tr.scheduler.get_lr() == 5e-3
    tr.train_step(batches[3:4])
    assert not torch.equal(tr.lora.master, before)


WORKER = r'''
import os, sys, json, torch
sys.path.insert(0, {root!r})
from tests.test_trainer_cpu import make_trainer
stage, out = int(sys.argv[1]), sys.argv[2]
cfgp = os.path.join({root!r}, "configs/ds_config_zero%d.json" % stage) if stage else None
tr = make_trainer(out, deepspeed=cfgp, precision="fp32", no_offload=(sys.argv[3] == "0"), max_steps=4,
                  gradient_accumulation_steps=1 if int(os.environ["WORLD_SIZE"]) == 2 else 2,
                  save_strategy="no", lora_dropout=0.0)   # dropout masks depend on the packing
assert tr.stage == stage
res = tr.train()
tr.save_model(os.path.join(out, "final"))
if tr.is_main:
    torch.save({{"master": tr.lora.master[: tr.lora.numel].clone(), "losses": [e["loss"] for e in tr.log_history]}},
               os.path.join(out, "result.pt"))
'''


def _run_world(tmp, stage, world, offload="0"):
    script = tmp / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    out = tmp / f"s{stage}_w{world}_{offload}"
    out.mkdir()
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
    if world == 1:
        env.update(WORLD_SIZE="1", RANK="0", LOCAL_RANK="0")
        cmd = [sys.executable, str(script), str(stage), str(out), offload]
    else:
        port = 29600 + stage * 7 + world + (3 if offload == "1" else 0)
        cmd = [sys.executable, "-m", "dlti_b200.launch", f"--num_gpus={world}", f"--master_port={port}",
               "--no_local_rank", str(script), str(stage), str(out), offload]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return torch.load(out / "result.pt"), out


@pytest.mark.parametrize("stage,offload", [(1, "0"), (2, "0"), (3, "0"), (3, "1")])
def test_zero_stages_match_single_process(tmp_path, stage, offload):
    """ZeRO-N on 2 ranks (micro 2 x accum 1 x world 2) == 1 rank (micro 2 x accum 2): same global
    batch, same fp32 AdamW result (token-normalised loss, SUM reduction)."""
    ref, _ = _run_world(tmp_path, stage, 1)
    got, out = _run_world(tmp_path, stage, 2, offload)
    torch.testing.assert_close(got["master"], ref["master"], rtol=2e-3, atol=1e-4)  # Adam is sign-like for ~0 grads: summation order shows
    for a, b in zip(got["losses"], ref["losses"]):
        assert abs(a - b) < 2e-3
    assert (out / "final" / "adapter_model.safetensors").exists()      # consolidated save, all ranks entered


FAULT_WORKER = r'''
import os, sys, torch
sys.path.insert(0, {root!r})
from tests.test_trainer_cpu import make_trainer
out = sys.argv[1]
cfgp = os.path.join({root!r}, "configs/ds_config_zero2.json")
tr = make_trainer(out, deepspeed=cfgp, precision="fp32", gradient_accumulation_steps=1, save_steps=2,
                  lora_dropout=0.0)
tr.train(resume_from_checkpoint=True if os.environ.get("RESUME") else None)
if tr.is_main:
    torch.save(tr.lora.master[: tr.lora.numel].clone(), os.path.join(out, "final_master.pt"))
'''


def test_fault_injection_fail_fast_and_resume(tmp_path):
    """§5.3: a rank dying (DLTI_FAULT_AT_STEP) makes the launcher tear the job down with a non-zero code;
    restarting with resume continues from the last checkpoint and ends where an uninterrupted run ends."""
    script = tmp_path / "fw.py"
    script.write_text(FAULT_WORKER.format(root=ROOT))
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")

    def launch(out, extra_env):
        import socket
        with socket.socket() as so:               # a fresh port per launch: the killed job's port may linger in TIME_WAIT
            so.bind(("127.0.0.1", 0))
            port = so.getsockname()[1]
        return subprocess.run([sys.executable, "-m", "dlti_b200.launch", "--num_gpus=2", f"--master_port={port}",
                               "--no_local_rank", str(script), str(out)], env=dict(env, **extra_env),
                              capture_output=True, text=True, timeout=600, cwd=ROOT)

    ref = launch(tmp_path / "ref", {})
    assert ref.returncode == 0, ref.stderr[-2000:]
    bad = launch(tmp_path / "run", {"DLTI_FAULT_AT_STEP": "5", "DLTI_FAULT_RANK": "1"})
    assert bad.returncode != 0 and "killing the other ranks" in bad.stderr
    assert ckpt.latest_checkpoint(tmp_path / "run").name == "checkpoint-4"
    ok = launch(tmp_path / "run", {"RESUME": "1"})
    assert ok.returncode == 0, ok.stderr[-2000:]
    a = torch.load(tmp_path / "ref" / "final_master.pt")
    b = torch.load(tmp_path / "run" / "final_master.pt")
    torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)
