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
    cfg = PRESETS[os.environ.get("MP_SERVE_PRESET") or ("tiny-gqa" if info.world_size <= 2 else "tiny-mha8")]   # 8 KV heads: up to 8 ranks
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
    # one prefill step (465 rows: the GEMM-fused K10 / K11 path) and one decode step (4 rows: push + barrier-reduce
    # path) run directly, so that the LOGITS can be compared across backends, not only the greedy tokens (which flip
    # on near-ties: every TP path rounds per-rank partial sums to bf16, TP=1 does not)
    bs = eng.blocks.block_size
    lens = [len(p) for p in prompts]
    nb = [-(-(n + 1) // bs) for n in lens]
    first = [sum(nb[:i]) + 200 for i in range(len(prompts))]          # blocks the scheduler will not hand out here
    ids = [t for p in prompts for t in p]
    pos = [j for n in lens for j in range(n)]
    slots = [(first[i] + j // bs) * bs + j % bs for i, n in enumerate(lens) for j in range(n)]
    cu = [0]
    for n in lens:
        cu.append(cu[-1] + n)
    samp = ([0.0] * len(lens), [0] * len(lens), [1.0] * len(lens), [0] * len(lens))

    def run_packed(make):             # (the workers execute the same step from their worker_loop)
        return eng._forward_step(eng._broadcast(make()))[0].float().clone()

    lg_p = run_packed(lambda: eng.pack_step(0, ids, pos, slots, [], [], cu, max(lens), samp))
    nxt = [7 + i for i in range(len(lens))]
    dslots = [(first[i] + n // bs) * bs + n % bs for i, n in enumerate(lens)]
    tables = [list(range(first[i], first[i] + nb[i])) for i in range(len(lens))]
    lg_d = run_packed(lambda: eng.pack_step(len(lens), nxt, lens, dslots, [n + 1 for n in lens], tables, [], 0, samp))
    out = eng.generate(prompts, sp)
    eng.shutdown_workers()
    if world_tp > 1 and backend == "fused" and os.environ.get("DLTI_TP_LARGE_MODE", "fused") == "fused":     # the 469-row prefill step must have taken the GEMM-fused K10/K11 path
        assert getattr(tp, "fused_large_calls", 0) >= 2 * cfg.num_hidden_layers, "prefill did not use the fused path"
    return [o["token_ids"] for o in out], lg_p, lg_d


STRICT = True


def main():
    global STRICT
    info = comm.init_distributed("cuda")
    STRICT = info.world_size <= 2 and not os.environ.get("MP_SERVE_PRESET")
    fc = FusedComm(info)
    ref, ref_p, ref_d = run(info, 1, None, None, False) if info.rank == 0 else (None, None, None)
    ok = True
    for backend, fused, graphs in (("nccl", None, False), ("fused", fc, False), ("fused", fc, True)):
        res = run(info, info.world_size, backend, fused, graphs)
        torch.cuda.synchronize(); dist.barrier()
        if info.rank == 0:
            got, lg_p, lg_d = res
            same = sum(a == b for x, y in zip(got, ref) for a, b in zip(x, y))
            tot = sum(len(x) for x in ref)
            ep = ((lg_p - ref_p).abs().max() / ref_p.abs().max()).item()
            ed = ((lg_d - ref_d).abs().max() / ref_d.abs().max()).item()
            print(f"[tp={info.world_size} {backend} graphs={graphs}] token agreement with TP=1: {same}/{tot}; logits vs TP=1: "
                  f"prefill max err {ep:.2e}, decode max err {ed:.2e} (relative to the largest logit)", flush=True)
            # logits of both kinds of step agree with TP=1 to bf16 accuracy on every backend; greedy decoding is
            # token-identical wherever the model's margins allow it (2 ranks of tiny-gqa: always; more ranks round
            # more partial sums to bf16 — on the NCCL path as much as on the peer-memory one — and may flip near-ties)
            ok = ok and ep < 3e-2 and ed < 3e-2 and (same == tot if STRICT else same >= 0.7 * tot)
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
