"""CPU unit tier: config resolution, naming/CSV, schedules, loss scaler, collator,
sampler, IO round-trips, merge, checkpoint rotation, Philox restatement."""
import json
import math
import os

import pytest
import torch

from dlti_b200.config.ds_config import DSConfig, DSConfigError
from dlti_b200.data.collate import CausalLMCollator
from dlti_b200.data.sampler import DistributedBatchSampler
from dlti_b200.data.synthetic import SyntheticInstructionDataset
from dlti_b200.data.text import format_conversation_for_llama2
from dlti_b200.engine import checkpoint as ckpt
from dlti_b200.models.config import PRESETS, LlamaConfig, resolve_config
from dlti_b200.models.hf_io import load_resident, resident_state_dict, save_hf_checkpoint
from dlti_b200.models.llama import LoraLlama, ResidentWeights, pack_batch
from dlti_b200.models.lora import LoraConfig, LoraParams
from dlti_b200.models.merge import merge_checkpoint
from dlti_b200.models.peft_io import load_adapter_config, load_adapter_state, save_adapter
from dlti_b200.optim.loss_scaler import DynamicLossScaler, build_loss_scaler
from dlti_b200.optim.sched import LinearDecaySchedule, WarmupLR
from dlti_b200.utils import metrics as M

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RES = dict(micro_batch=2, grad_accum=4, world_size=2, lr=2e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
           max_grad_norm=1.0, warmup_steps=0, total_steps=100)


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_reference_configs_resolve(stage):
    ds = DSConfig.load(os.path.join(ROOT, f"configs/ds_config_zero{stage}.json")).resolve(**RES)
    assert ds.zero_stage == stage and ds.precision == "fp16" and ds.gradient_clipping == 1.0
    assert ds.cfg["train_batch_size"] == 16 and ds.cfg["train_micro_batch_size_per_gpu"] == 2
    kind, sp = ds.scheduler
    assert kind == "WarmupLR" and sp["warmup_max_lr"] == 2e-4 and sp["warmup_num_steps"] == 0
    assert ds.has_optimizer == (stage != 2)
    if stage == 3:
        assert ds.offload_optimizer == "cpu" and ds.offload_param == "cpu" and ds.pin_memory
        assert ds.stage3_param_persistence_threshold == 100000 and ds.reduce_bucket_size == 50000000
        ds.override_no_offload()
        assert ds.offload_optimizer == "none" and ds.offload_param == "none"
    if stage == 2:
        assert ds.offload_optimizer == "none" and ds.allgather_bucket_size == 200000000


def test_config_mismatch_and_precision_override():
    raw = json.load(open(os.path.join(ROOT, "configs/ds_config_zero1.json")))
    raw["optimizer"]["params"]["lr"] = 1e-3
    raw["gradient_clipping"] = 0.5
    with pytest.raises(DSConfigError) as e:
        DSConfig(raw).resolve(**RES)
    assert "optimizer.params.lr" in str(e.value) and "gradient_clipping" in str(e.value)
    ds = DSConfig(json.load(open(os.path.join(ROOT, "configs/ds_config_zero3.json")))).resolve(**RES, precision="bf16")
    assert ds.precision == "bf16"
    with pytest.raises(DSConfigError):
        DSConfig({"zero_optimization": {"stage": 7}})


def test_naming_and_csv(tmp_path):
    assert M.create_experiment_name(1, 0) == "baseline"
    assert M.create_experiment_name(2, 2) == "zero2_2gpu" and M.create_experiment_name(8, 3) == "zero3_8gpu"
    assert M.get_zero_stage_from_config(os.path.join(ROOT, "configs/ds_config_zero3.json")) == 3
    p = tmp_path / "results" / "m.csv"
    row = {c: 1 for c in M.CORE_COLUMNS}
    row.update(experiment="baseline", strategy="pytorch_lora", tokens_per_second=5.0)
    M.save_training_metrics(row, str(p))
    M.save_training_metrics(dict(row, experiment="zero2_2gpu"), str(p))
    lines = p.read_text().strip().splitlines()
    assert lines[0].split(",")[:8] == M.CORE_COLUMNS and len(lines) == 3
    import doctest
    assert doctest.testmod(M).failed == 0


def test_schedules():
    s = WarmupLR(0.0, 2e-4, 0)
    assert s.get_lr() == 0.0          # before the first step (DeepSpeed warning path)
    assert s.step() == 0.0            # it 0: log(1)/log(2) = 0 (warmup clamped to 2 steps)
    assert s.step() == pytest.approx(2e-4) and s.step() == pytest.approx(2e-4)
    s = WarmupLR(0.0, 1.0, 100)
    vals = [s.step() for _ in range(120)]
    assert vals[9] == pytest.approx(math.log(10) / math.log(100)) and vals[-1] == 1.0
    lin = LinearDecaySchedule(2e-4, 10)
    assert lin.get_lr() == pytest.approx(2e-4)
    got = [lin.step() for _ in range(10)]
    assert got[0] == pytest.approx(2e-4 * 0.9) and got[-1] == 0.0
    sd = lin.state_dict()
    l2 = LinearDecaySchedule(2e-4, 10)
    l2.load_state_dict(sd)
    assert l2.get_lr() == lin.get_lr()


def test_dynamic_loss_scaler_state_machine():
    sc = build_loss_scaler({"enabled": True, "loss_scale": 0, "loss_scale_window": 4, "initial_scale_power": 16,
                            "hysteresis": 2, "min_loss_scale": 1})
    assert isinstance(sc, DynamicLossScaler) and sc.cur_scale == 65536
    sc.update(True)                      # first overflow absorbed by hysteresis
    assert sc.cur_scale == 65536
    sc.update(True)                      # second halves
    assert sc.cur_scale == 32768
    for _ in range(3):
        sc.update(False)
    assert sc.cur_scale == 32768
    sc.update(False)                     # window of 4 clean iterations since last overflow => x2
    assert sc.cur_scale == 65536
    floor = DynamicLossScaler(initial_scale_power=1, hysteresis=1)
    for _ in range(5):
        floor.update(True)
    assert floor.cur_scale == 1.0
    assert build_loss_scaler(None).cur_scale == 1.0 and build_loss_scaler({"enabled": True, "loss_scale": 128}).cur_scale == 128


def test_collator_and_pack():
    feats = [{"input_ids": [1, 5, 6, 2]}, {"input_ids": [1, 7, 2]}, {"input_ids": [1, 2, 9, 9, 2]}]
    b = CausalLMCollator(pad_token_id=2, mask_by="pad_id")(feats)
    assert b["input_ids"].shape == (3, 5) and b["labels"][1].tolist() == [1, 7, -100, -100, -100]
    assert b["labels"][0].tolist() == [1, 5, 6, -100, -100]        # real EOS masked too (pad == eos)
    b2 = CausalLMCollator(pad_token_id=2, mask_by="attention")(feats)
    assert b2["labels"][0].tolist() == [1, 5, 6, 2, -100]
    pk = pack_batch(b2["input_ids"], b2["labels"], b2["attention_mask"])
    assert pk.cu_seqlens.tolist() == [0, 4, 7, 12] and pk.max_seqlen == 5
    assert pk.input_ids.tolist() == [1, 5, 6, 2, 1, 7, 2, 1, 2, 9, 9, 2]
    assert pk.labels.tolist() == [5, 6, 2, -100, 7, 2, -100, 2, 9, 9, 2, -100]
    assert pk.positions.tolist() == [0, 1, 2, 3, 0, 1, 2, 0, 1, 2, 3, 4] and pk.n_valid == 9


def test_sampler_shards_and_resume():
    s0 = DistributedBatchSampler(10, 2, rank=0, world_size=2, seed=42)
    s1 = DistributedBatchSampler(10, 2, rank=1, world_size=2, seed=42)
    a, b = s0.epoch_indices(0), s1.epoch_indices(0)
    assert len(a) == len(b) == 5 and sorted(a + b) == list(range(10))
    assert s0.epoch_indices(0) == a and s0.epoch_indices(1) != a
    full = list(s0.iter_epoch(0))
    assert list(s0.iter_epoch(0, skip_batches=2)) == full[2:] and len(s0) == 3


def test_synthetic_dataset_shapes():
    ds = SyntheticInstructionDataset(16, 1000, max_length=64)
    x = ds[3]
    assert x == ds[3] and 10 <= len(x["input_ids"]) <= 64 and x["input_ids"][0] == 1
    fx = SyntheticInstructionDataset(4, 1000, max_length=32, fixed_length=True)[0]
    assert len(fx["input_ids"]) == 32
    assert format_conversation_for_llama2({"question": " q ", "answer": "a"})["text"] == "<s>[INST] q [/INST] a</s>"


def test_presets_and_param_counts():
    c = resolve_config("meta-llama/Llama-2-7b-hf")
    assert c.num_params() == 6738415616 and c.params_per_layer() == 202383360
    lp_numel = 32 * (4 * 16 * 4096 + 4 * 4096 * 16)
    assert lp_numel == 16777216
    assert resolve_config("llama2-70b").num_params() == 68976648192
    with pytest.raises(ValueError):
        resolve_config("nope")


def test_hf_and_peft_roundtrip_and_merge(tmp_path):
    cfg = PRESETS["tiny-gqa"]
    w = ResidentWeights.random_init(cfg, "cpu", torch.float32, seed=0)
    files = save_hf_checkpoint(cfg, resident_state_dict(w), tmp_path / "base", max_shard_bytes=3_000_000)
    assert len(files) > 1 and (tmp_path / "base" / "model.safetensors.index.json").exists()
    cfg2, w2 = load_resident(tmp_path / "base", "cpu", torch.float32)
    assert cfg2.num_key_value_heads == 2
    for a, b in zip(w.all_matrices(), w2.all_matrices()):
        assert torch.equal(a, b)
    lcfg = LoraConfig(r=8, lora_alpha=16, base_model_name_or_path="tiny-gqa")
    lora = LoraParams(cfg, lcfg, "cpu", torch.float32, seed=1)
    for n in lora.slots:
        if n.endswith("lora_B"):
            lora.view(n).normal_(0, 0.1)
    save_adapter(lora.state_dict(), lcfg, tmp_path / "ad")
    ac = json.load(open(tmp_path / "ad" / "adapter_config.json"))
    assert ac["peft_type"] == "LORA" and ac["r"] == 8 and ac["lora_alpha"] == 16 and ac["bias"] == "none"
    sd = load_adapter_state(tmp_path / "ad")
    k = "base_model.model.model.layers.1.self_attn.v_proj.lora_B.weight"
    assert k in sd and sd[k].shape == (cfg.kv_size, 8) and sd[k].dtype == torch.float32
    l2 = LoraParams(cfg, load_adapter_config(tmp_path / "ad"), "cpu", torch.float32, seed=5)
    l2.load_state_dict(sd)
    assert torch.equal(l2.master[: l2.numel], lora.master[: lora.numel])
    # merged model == base + adapter
    merge_checkpoint(tmp_path / "base", tmp_path / "ad", tmp_path / "merged", dtype=torch.float32)
    _, wm = load_resident(tmp_path / "merged", "cpu", torch.float32)
    lora.refresh_compute()
    ids = torch.randint(3, cfg.vocab_size, (2, 12))
    b = pack_batch(ids, ids.clone(), None)
    m_ad = LoraLlama(cfg, w, lora); m_ad.training = False
    m_mg = LoraLlama(cfg, wm, None); m_mg.training = False
    torch.testing.assert_close(m_ad.forward_logits(b), m_mg.forward_logits(b), rtol=1e-4, atol=1e-4)


def test_checkpoint_rotation(tmp_path):
    for s in (100, 200, 300, 1000):
        (tmp_path / f"checkpoint-{s}").mkdir()
    (tmp_path / "final").mkdir()
    assert ckpt.latest_checkpoint(tmp_path).name == "checkpoint-1000"
    gone = ckpt.rotate_checkpoints(tmp_path, 2)
    assert [p.name for p in gone] == ["checkpoint-100", "checkpoint-200"]
    assert [p.name for p in ckpt.list_checkpoints(tmp_path)] == ["checkpoint-300", "checkpoint-1000"]


def test_philox_known_answer():
    from dlti_b200.ops.rng import keep_mask, philox4x32
    import numpy as np
    # Random123 known-answer test for philox4x32-10: counter = key = 0
    r = philox4x32(np.array([0]), np.array([0]), np.array([0]), np.array([0]), 0, 0)
    assert [int(x[0]) for x in r] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    m = keep_mask(64, 128, 0.05, 1, 2)
    assert 0.02 < 1 - m.float().mean().item() < 0.09


def test_public_header_declares_every_entry_point():
    """include/dlti_b200.h is the C ABI contract: every DLTI_API definition in csrc/ must be declared."""
    import re
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    header = (root / "include" / "dlti_b200.h").read_text()
    declared = set(re.findall(r"\b(dlti_[a-z0-9_]+)\s*\(", header))
    defined = set()
    for src in list((root / "dlti_b200" / "csrc").glob("*.cu")) + list((root / "dlti_b200" / "csrc").glob("*.cpp")):
        defined |= set(re.findall(r"DLTI_API\s+[a-zA-Z_ ]+\**\s*(dlti_[a-z0-9_]+)\s*\(", src.read_text()))
    assert defined, "no entry points found"
    missing = sorted(defined - declared)
    assert not missing, f"undeclared entry points: {missing}"


def test_debug_index_checks():
    """DLTI_DEBUG_CHECKS: the ops that index memory through caller-supplied integers reject bad indices before
    anything is launched (our stand-in for compute-sanitizer's memcheck on a pool where it cannot run)."""
    from dlti_b200.ops import _debug
    from dlti_b200.ops import attention as A
    from dlti_b200.ops import elementwise as E
    from dlti_b200.ops import serving as S
    H = KV = 2
    hd, bs, nblk = 8, 4, 6
    cos, sin = E.rope_tables(32, hd, 1e4, "cpu")
    kc = torch.zeros(nblk, bs, KV, hd)
    vc = torch.zeros_like(kc)
    qkv = torch.randn(3, (H + 2 * KV) * hd)
    pos = torch.tensor([0, 1, 2], dtype=torch.int32)
    good = torch.tensor([0, 5, -1], dtype=torch.int64)
    bad = torch.tensor([0, nblk * bs, -1], dtype=torch.int64)
    tables = torch.tensor([[1, 2, 0], [9, 0, 0]], dtype=torch.int32)
    table = torch.randn(16, 8)
    S.rope_cache_(qkv.clone(), pos, cos, sin, bad[:1].repeat(3), kc, vc, H, KV, hd)       # checks are off by default
    _debug.enable(True)
    try:
        S.rope_cache_(qkv.clone(), pos, cos, sin, good, kc, vc, H, KV, hd)
        with pytest.raises(_debug.DltiIndexError, match=r"slots\[1\] = 24"):
            S.rope_cache_(qkv.clone(), pos, cos, sin, bad, kc, vc, H, KV, hd)
        with pytest.raises(_debug.DltiIndexError, match="positions"):
            S.rope_cache_(qkv.clone(), torch.tensor([0, 1, 32], dtype=torch.int32), cos, sin, good, kc, vc, H, KV, hd)
        q = torch.randn(2, H * hd)
        S.paged_decode(q, kc, vc, tables, torch.tensor([5, 0], dtype=torch.int32), H, KV, hd)  # row 1 unused: entry 9 is fine
        with pytest.raises(_debug.DltiIndexError, match=r"block_tables\[1, 0\] = 9"):
            S.paged_decode(q, kc, vc, tables, torch.tensor([5, 1], dtype=torch.int32), H, KV, hd)
        with pytest.raises(_debug.DltiIndexError, match="does not fit"):
            S.paged_decode(q, kc, vc, tables, torch.tensor([13, 0], dtype=torch.int32), H, KV, hd)
        with pytest.raises(_debug.DltiIndexError, match="ids"):
            E.embedding(torch.tensor([0, 16]), table)
        with pytest.raises(_debug.DltiIndexError, match="cu_seqlens"):
            A.attention_fwd(qkv, torch.tensor([0, 2, 4], dtype=torch.int32), 2, H, KV, hd)
    finally:
        _debug.enable(False)


def _ring_reader(name, idx, n_readers, n_msgs, q):
    import numpy as np
    from dlti_b200.serve.shm_ring import ShmRing
    r = ShmRing(name, idx, n_readers, n_slots=4, slot_bytes=1 << 16)
    ok = True
    for i in range(n_msgs):
        m = r.read(timeout_ms=20000).view(np.int64)
        ok = ok and m.size == 1 + (i % 97) and int(m[0]) == i and (m.size == 1 or int(m[-1]) == i * 3)
    r.close()
    q.put(ok)


def test_shm_ring_delivers_every_message_in_order_to_every_reader():
    """csrc/shm_ring.cpp (N10): one writer, two readers, a ring much shorter than the message count."""
    import multiprocessing as mp
    import numpy as np
    from dlti_b200.serve.shm_ring import ShmRing
    name = f"/dlti_test_ring_{os.getpid()}"
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    w = ShmRing(name, -1, 2, n_slots=4, slot_bytes=1 << 16)
    procs = [ctx.Process(target=_ring_reader, args=(name, i, 2, 500, q)) for i in range(2)]
    for p in procs:
        p.start()
    for i in range(500):
        m = np.full(1 + (i % 97), i * 3, dtype=np.int64)
        m[0] = i
        w.write(m, timeout_ms=20000)
    res = [q.get(timeout=60) for _ in procs]
    for p in procs:
        p.join(timeout=30)
    w.close()
    assert res == [True, True]
    with pytest.raises(RuntimeError):
        ShmRing(name + "x", -1, 1, n_slots=2, slot_bytes=64).write(np.zeros(100, dtype=np.int64))


def test_baseline_fp16_scaler_follows_torch_gradscaler_and_ds_config_reports_ignored_keys(capsys):
    from dlti_b200.optim.loss_scaler import grad_scaler_like
    s = grad_scaler_like()
    assert s.cur_scale == 65536.0
    s.update(True)                       # GradScaler backs off at once (DeepSpeed's default waits: hysteresis 2)
    assert s.cur_scale == 32768.0
    for _ in range(1999):
        s.update(False)
    assert s.cur_scale == 32768.0
    s.update(False)                      # growth_interval = 2000 clean steps
    assert s.cur_scale == 65536.0
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    c = DSConfig.load(os.path.join(root, "configs", "ds_config_zero3.json"))
    keys = " ".join(c.ignored_keys())
    assert "stage3_prefetch_bucket_size" in keys and "stage3_max_reuse_distance" in keys
    assert "accepted for compatibility" in capsys.readouterr().err


def test_zero3_gather_queue_schedule_dry_run(monkeypatch):
    """Host-side logic of K1 (parallel.fused.GatherQueue + parallel.zero3.Zero3Weights), run without a GPU: the native
    attach call is replaced by a recorder and two optimizer steps of the model's exact call order are replayed. Every
    64-row unit of every gathered matrix must be requested exactly once per fetch; a consumer whose operand is not
    fully requested by EARLIER launches must carry the remainder itself and wait on it; in steady state a step
    fetches 2(L-2) layers + the LM head (two staging slots: the two layers at either turning point are still there)."""
    from dlti_b200.models.config import PRESETS
    from dlti_b200.parallel import fused as F
    from dlti_b200.parallel.comm import DistInfo
    from dlti_b200.parallel.zero3 import Zero3Weights

    calls = []

    def fake_fn(name, argtypes, restype=None, cpu=False):
        def f(*a):
            if name == "dlti_gemm2_attach":
                calls.append(dict(staging=a[1].value if hasattr(a[1], "value") else a[1], j0=a[7], n=a[8], epoch=a[9],
                                  wait=bool(getattr(a[10], "value", a[10]))))
            return 0
        return f

    monkeypatch.setattr(F._lib, "fn", fake_fn)

    class FakeFused:
        rank, world = 1, 4
        info = DistInfo(rank=1, world_size=4, device=torch.device("cpu"))
        err = torch.zeros(1, dtype=torch.int32)

        def alloc_symmetric(self, shape, dtype):
            return torch.zeros(*shape, dtype=dtype)

        def peer_array(self, t):
            return None

    import dataclasses
    cfg = dataclasses.replace(PRESETS["tiny-gqa"], num_hidden_layers=5)
    L = cfg.num_hidden_layers
    ds = DSConfig({"zero_optimization": {"stage": 3}})
    w = Zero3Weights(cfg, ds, FakeFused.info, torch.bfloat16, backend="fused", fused=FakeFused())
    assert w.in_gemm_gather
    by_ptr = {}
    for s in w.stage_slots:
        for m, sm in s.items():
            by_ptr[sm.staging.data_ptr()] = sm
    by_ptr[w.head_stage.staging.data_ptr()] = w.head_stage
    fetches = []                     # (tag, epoch) in bind order
    orig_bind = F.StagedMatrix.bind

    def bind(self, shard, tag):
        orig_bind(self, shard, tag)
        fetches.append((tag, self.epoch, self))
    monkeypatch.setattr(F.StagedMatrix, "bind", bind)
    covered = {}                     # (id(sm), epoch) -> set of positions requested so far

    def launch(use, flops=1e11):
        """One large GEMM launch with `use` as its gather handle; checks the consumer rule."""
        n0 = len(calls)
        sm = use.sm
        before = set(covered.get((id(sm), sm.epoch), ())) if sm is not None else None
        use.attach(flops)
        new = calls[n0:]
        assert len(new) <= 1
        for c in new:
            tgt = by_ptr[c["staging"]]
            got = covered.setdefault((id(tgt), c["epoch"]), set())
            rng = set(range(c["j0"], c["j0"] + c["n"]))
            assert not (rng & got), "a unit was requested twice"
            got |= rng
            if c["wait"]:
                assert tgt is sm and before | rng == set(range(sm.n_units)), "a waiting launch must complete its operand"
        if sm is not None and len(before) < sm.n_units:          # operand incomplete before this launch
            assert new and new[0]["wait"] and by_ptr[new[0]["staging"]] is sm

    nb = 2
    for step in range(2):
        start = len(fetches)
        w.prefetch(0)
        for li in range(L):
            if li + 1 < L:
                w.prefetch(li + 1)
            lw = w.layer(li)
            if li == L - 1:
                w.prefetch_head()
            for _ in range(nb):
                for m in ("w_qkv", "w_o", "w_gu", "w_down"):
                    launch(lw.g(m))
            w.release(li)
        w.get_lm_head()
        for _ in range(nb):
            launch(w.head_gather()); launch(w.head_gather())
        w.release_head()
        w.prefetch(L - 1)
        for li in reversed(range(L)):
            if li > 0:
                w.prefetch(li - 1)
            lw = w.layer(li)
            for _ in range(nb):
                for m in ("w_qkv", "w_o", "w_gu", "w_down", "w_gu", "w_o", "w_qkv"):
                    launch(lw.g(m))
            w.release(li)
        layer_fetches = {t[0][0] for t in fetches[start:] if len(t[0]) == 2}
        n_layer_binds = sum(1 for t in fetches[start:] if len(t[0]) == 2) // 4
        n_head = sum(1 for t in fetches[start:] if t[0] == ("lm_head",))
        assert n_head == 1
        assert n_layer_binds == (2 * L - 2 if step == 0 else 2 * (L - 2)), (step, n_layer_binds)
    w.queue.flush()
    for (tag, epoch, sm) in fetches:                             # every fetch fully requested, exactly once
        assert covered.get((id(sm), epoch)) == set(range(sm.n_units)), (tag, epoch)


def test_tp_prefill_k10_k11_host_orchestration_dry_run(monkeypatch):
    """Host side of the GEMM-fused prefill collectives (serve.model_runner.TPContext, K10 / K11) without a GPU: the
    native calls are recorded. Per row-parallel GEMM: scatter spec with per = ceil(ceil(T / world) / 64) * 64, two flag
    barriers around the owner-side reduce, a fresh zeroed counter slice, and a gather handle whose job covers
    world * per / 64 units starting at this rank's own rows (rot = rank * per / 64, m_rot = rank * per // 256)."""
    from dlti_b200.models.config import PRESETS
    from dlti_b200.parallel import fused as F
    from dlti_b200.parallel.comm import DistInfo
    from dlti_b200.serve import model_runner as MR

    rec = []

    def fake_fn(name, argtypes, restype=None, cpu=False):
        def f(*a):
            rec.append((name, a))
            return 0
        return f

    monkeypatch.setattr(F._lib, "fn", fake_fn)
    monkeypatch.setattr(F._lib, "stream_ptr", lambda: None)

    class FakeFused:
        def __init__(self, rank, world):
            self.rank, self.world = rank, world
            self.info = DistInfo(rank=rank, world_size=world, device=torch.device("cpu"))
            self.err = torch.zeros(1, dtype=torch.int32)
            self.barriers = []

        def alloc_symmetric(self, shape, dtype):
            return torch.zeros(*shape, dtype=dtype)

        def peer_array(self, t):
            return ("peers", t.data_ptr())

        def barrier_graph_safe(self, ch):
            self.barriers.append(ch)

    cfg = PRESETS["tiny-mha8"]
    d = cfg.hidden_size
    for world, rank, T in ((8, 5, 465), (2, 1, 8192), (4, 0, 300)):
        rec.clear()
        fz = FakeFused(rank, world)
        tp = MR.TPContext(fz.info, "fused", fz, cfg, max_tokens=8192, dtype=torch.float32)
        assert tp.fused_large
        residual = torch.zeros(T, d)
        seen = {}

        def partial_fn(out, push, scatter=None):
            seen["scatter"], seen["out_rows"] = scatter, out.shape[0]

        counters = []
        for _ in range(3):
            x = tp.reduce_add_norm(partial_fn, T, residual, torch.ones(d), 1e-5)
            per = -(-(-(-T // world)) // 64) * 64
            assert isinstance(x, MR.ShardedAct) and x.rows().shape == (T, d)
            assert seen["scatter"][1:] == (world, rank, per) and seen["out_rows"] == T
            h = x.handle
            assert h.per == per and h.n_units == world * per // 64 and h.counters.numel() == h.n_units
            assert int(h.counters.abs().sum()) == 0
            counters.append((h.counters.data_ptr(), h.counters.numel()))
            h.attach(1e9)
            att = [a for n, a in rec if n == "dlti_gemm2_attach"][-1]
            assert att[4] == per and att[5] == world and att[6] == rank * (per // 64) and att[7] == 0 and att[8] == h.n_units
            assert att[11] == h.n_units                                   # the launch waits on its A operand
            assert [a for n, a in rec if n == "dlti_gemm2_attach_a_side"][-1][0] == (rank * per) // 256
            n_att = len([1 for n, _ in rec if n == "dlti_gemm2_attach"])
            h.attach(1e9)                                                 # a second consumer: nothing to pull again
            assert len([1 for n, _ in rec if n == "dlti_gemm2_attach"]) == n_att
        assert fz.barriers == [6, 6] * 3
        own = [a for n, a in rec if n == "dlti_tp_own_reduce_norm"]
        lo = min(T, rank * per)
        assert len(own) == 3 and own[0][2] == per and own[0][3] == lo and own[0][4] == min(T, lo + per) - lo
        spans = sorted(counters)
        assert all(a + 4 * n <= b for (a, n), (b, _) in zip(spans, spans[1:])), "counter slices overlap"


@pytest.mark.parametrize("mode", ["fused", "overlap", "tail"])
def test_k9_host_orchestration_dry_run(monkeypatch, mode):
    """parallel.zero.ZeroOptimizer on the fused backend with a recording fake: which peer-memory calls a step makes in
    each DLTI_K9 mode, and that the flat gradient ranges handed to the reduce cover this rank's partition."""
    from dlti_b200.models.config import PRESETS
    from dlti_b200.models.lora import LoraConfig, LoraParams
    from dlti_b200.parallel.comm import DistInfo
    from dlti_b200.parallel.zero import ZeroOptimizer

    monkeypatch.setenv("DLTI_K9", mode)
    cfg = PRESETS["tiny-gqa"]
    world, rank = 4, 2
    info = DistInfo(rank=rank, world_size=world, device=torch.device("cpu"))
    lora = LoraParams(cfg, LoraConfig(r=8, lora_alpha=16), "cpu", torch.float32, pad_to=world * 64)
    calls = []

    class FakeFused:
        def k9_begin_step(self):
            calls.append(("begin",))

        def grad_range_final(self, slot, g_full, grad_part, glo, n, part_lo, part_n, side=True, ctas=16):
            calls.append(("range", slot, glo, n, part_lo, part_n, side))

        def k9_step(self, master, m, v, g, lo, n, **kw):
            calls.append(("step", lo, n, sorted(kw)))
            kw["out2"].copy_(torch.tensor([4.0, 7.0]))
            kw["skipped_out"].zero_()

        def reduce_scatter_accumulate(self, g, part):
            calls.append(("rs",))

        def global_sumsq(self, g, out):
            calls.append(("sumsq",))
            out.fill_(4.0)

        def adamw_push(self, *a, **kw):
            calls.append(("adamw_push",))
            kw["skipped_out"].zero_()

    opt = ZeroOptimizer(lora, DSConfig({"zero_optimization": {"stage": 2}}), info, lr=1e-3, backend="fused", fused=FakeFused())
    assert opt.k9_fused == (mode != "tail") and opt.overlap == (mode == "overlap")
    opt.begin_step()
    for li in reversed(range(cfg.num_hidden_layers)):        # the model's hook, one call per decoder layer
        lo, hi = lora.layer_range(li)
        opt.on_grad_range_final(lo, hi - lo)
    opt.on_micro_batch_end(is_boundary=True)
    skipped, gnorm = opt.step(1e-3, torch.tensor([7.0]))
    kinds = [c[0] for c in calls]
    if mode == "tail":
        assert kinds == ["rs", "sumsq", "adamw_push"] and opt.loss_sum_global is None
    else:
        ranges = [c for c in calls if c[0] == "range"]
        if mode == "fused":
            assert kinds == ["begin", "range", "step"] and ranges[0][1:4] == (0, 0, lora.padded_numel) and ranges[0][6] is False
        else:
            assert kinds == ["begin"] + ["range"] * cfg.num_hidden_layers + ["step"]
            assert [r[1] for r in ranges] == list(range(cfg.num_hidden_layers)) and all(r[6] for r in ranges)
            assert sum(r[3] for r in ranges) == lora.numel                  # the layers tile the adapter space
        assert all(r[4] == rank * opt.part and r[5] == opt.part for r in ranges)
        assert opt.loss_sum_global == 7.0 and not skipped and abs(gnorm - 2.0) < 1e-6
