"""Multi-GPU serving check (launched with one process per GPU): TP=N greedy decoding must reproduce
TP=1 (same random-init model), on both the fused (K10 push + reduce/norm kernel) and nccl backends."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from dlti_b200.models.config import PRESETS  # noqa: E402
from dlti_b200.parallel import comm  # noqa: E402
from dlti_b200.parallel.fused import FusedComm  # noqa: E402
from dlti_b200.serve.engine import LLMEngine  # noqa: E402
from dlti_b200.serve.model_runner import TPContext, TPWeights  # noqa: E402
from dlti_b200.serve.params import SamplingParams  # noqa: E402
from dlti_b200.serve.tokenizer import ByteTokenizer  # noqa: E402


def run(info, world_tp, backend, fused, graphs):
    cfg = PRESETS["tiny-gqa" if info.world_size <= 2 else "tiny-mha8"]      # 8 KV heads: up to 8 ranks
    if world_tp == 1:
        w = TPWeights(cfg, 0, 1, info.device, torch.bfloat16).init_random(5)
        tp = TPContext()
    else:
        w = TPWeights(cfg, info.rank, info.world_size, info.device, torch.bfloat16).init_random(5)
        tp = TPContext(info, backend, fused, cfg, max_tokens=1024)
    for l in w.layers:                          # widen the logit spread so greedy is stable
        for k in ("w_qkv", "w_o", "w_gu", "w_down"):
            l[k].mul_(4.0)
    w.embed.mul_(4.0); w.lm_head.mul_(4.0)
    eng = LLMEngine(cfg, w, tp, ByteTokenizer(cfg.vocab_size), info=info if world_tp > 1 else None,
                    num_blocks=256, max_model_len=512, use_cuda_graphs=graphs)
    # 469 prompt tokens in the first step: above 256 rows the row-parallel all-reduce takes the reduce-scatter +
    # all-gather path, the decode steps after it the push path
    prompts = [[1] + list(range(40, 40 + n)) for n in (5, 33, 130, 297)]
    sp = SamplingParams(max_tokens=16, temperature=0.0, ignore_eos=True)
    if world_tp > 1 and info.rank != 0:
        eng.worker_loop()
        return None
    out = eng.generate(prompts, sp)
    eng.shutdown_workers()
    if world_tp > 1 and backend == "fused":     # the 469-row prefill step must have taken the GEMM-fused K10/K11 path
        assert getattr(tp, "fused_large_calls", 0) >= 2 * cfg.num_hidden_layers, "prefill did not use the fused path"
    return [o["token_ids"] for o in out]


def main():
    info = comm.init_distributed("cuda")
    fc = FusedComm(info)
    ref = run(info, 1, None, None, False) if info.rank == 0 else None
    ok = True
    for backend, fused, graphs in (("nccl", None, False), ("fused", fc, False), ("fused", fc, True)):
        got = run(info, info.world_size, backend, fused, graphs)
        torch.cuda.synchronize(); dist.barrier()
        if info.rank == 0:
            same = sum(a == b for x, y in zip(got, ref) for a, b in zip(x, y))
            tot = sum(len(x) for x in ref)
            print(f"[tp={info.world_size} {backend} graphs={graphs}] token agreement with TP=1: {same}/{tot}", flush=True)
            ok = ok and same == tot          # token-identical greedy decoding
    fc.check_error()
    flag = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.broadcast(flag, 0)
    if info.rank == 0 and flag.item():
        print("ALL TP SERVING CHECKS PASSED", flush=True)
    comm.shutdown()
    if not flag.item():
        raise SystemExit(1)


if __name__ == "__main__":
    main()
