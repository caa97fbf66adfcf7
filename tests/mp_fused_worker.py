"""Worker for the multi-GPU tests: one process per GPU (launched by dlti_b200.launch / torchrun).
Checks the peer-memory kernels against NCCL collectives, then fused vs nccl training backends."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from dlti_b200.parallel import comm  # noqa: E402
from dlti_b200.parallel.fused import FusedComm  # noqa: E402


def check(name, ok):
    flag = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if dist.get_rank() == 0:
        print(f"[{'ok' if flag.item() else 'FAIL'}] {name}", flush=True)
    if not flag.item():
        raise SystemExit(f"{name} failed on some rank")


def main():
    info = comm.init_distributed("cuda")
    rank, world = info.rank, info.world_size
    fc = FusedComm(info)
    if rank == 0:
        print(f"arena mode: {fc.arena.mode}, multicast ptr: {fc.arena.chunks[0]['multicast']:#x}", flush=True)
    dev = info.device
    torch.manual_seed(100 + rank)

    # barrier x3
    for _ in range(3):
        fc.barrier()
    torch.cuda.synchronize()
    fc.check_error()
    check("peer barrier", True)

    # all-gather of row shards
    shard = fc.alloc_symmetric((192, 512), torch.bfloat16)
    shard.copy_(torch.randn(192, 512, device=dev))
    torch.cuda.synchronize(); dist.barrier()
    out = torch.empty(world * 192, 512, dtype=torch.bfloat16, device=dev)
    fc.all_gather_rows(shard, out)
    ref = torch.empty_like(out)
    dist.all_gather_into_tensor(ref, shard.contiguous())
    torch.cuda.synchronize()
    check("allgather_rows == nccl all_gather", torch.equal(out, ref))

    # reduce-scatter accumulate
    n = world * 4096
    g = fc.alloc_symmetric((n,), torch.float32)
    g.copy_(torch.randn(n, device=dev))
    part = torch.ones(n // world, device=dev)
    fc.reduce_scatter_accumulate(g, part)
    ref = torch.empty(n // world, device=dev)
    dist.reduce_scatter_tensor(ref, g.clone())
    torch.cuda.synchronize()
    check("rs_accumulate == nccl reduce_scatter (+1)", torch.allclose(part, ref + 1, rtol=1e-5, atol=1e-5))

    # global sum of squares
    x = torch.randn(10000, device=dev) * (rank + 1)
    o = torch.zeros(1, device=dev)
    fc.global_sumsq(x, o)
    r = (x.double() ** 2).sum().float().reshape(1)
    dist.all_reduce(r)
    torch.cuda.synchronize()
    check("global_sumsq == all_reduce", torch.allclose(o, r, rtol=1e-4))
    gathered = [torch.zeros_like(o) for _ in range(world)]
    dist.all_gather(gathered, o)
    check("global_sumsq identical on all ranks", all(torch.equal(gathered[0], t) for t in gathered))

    # adamw push: every rank ends with the same full master
    from dlti_b200.ops import optim as O
    master = fc.alloc_symmetric((n,), torch.float32)
    base = torch.arange(n, device=dev, dtype=torch.float32) / n
    master.copy_(base)
    torch.cuda.synchronize(); dist.barrier()
    pn = n // world
    m, v = torch.zeros(pn, device=dev), torch.zeros(pn, device=dev)
    gp = torch.full((pn,), 0.5 + rank, device=dev)
    nsq = torch.tensor([4.0], device=dev)
    sk = torch.zeros(1, dtype=torch.int32, device=dev)
    fc.adamw_push(master, m, v, gp, rank * pn, pn, lr=1e-2, betas=(0.9, 0.999), eps=1e-8, wd=0.0, step=1,
                  gnorm_sq=nsq, inv_scale=1.0, max_norm=0.0, skipped_out=sk)
    torch.cuda.synchronize()
    ref_full = base.clone()
    for r_ in range(world):
        pr = ref_full[r_ * pn:(r_ + 1) * pn]
        O.adamw_step_ref(pr, torch.zeros(pn, device=dev), torch.zeros(pn, device=dev),
                         torch.full((pn,), 0.5 + r_, device=dev), lr=1e-2, betas=(0.9, 0.999), eps=1e-8,
                         weight_decay=0.0, step=1)
    check("adamw_push: all ranks hold the updated full master", torch.allclose(master, ref_full, rtol=1e-5, atol=1e-6))
    fc.check_error()

    # K1: all-gather fused into the GEMM (NT and NN forms) vs a plain GEMM on the NCCL-gathered weight
    from dlti_b200.ops.gemm import gemm, gemm_ref
    rp, K, M = 256, 512, 384
    wshard = fc.alloc_symmetric((rp, K), torch.bfloat16)
    wshard.copy_(torch.randn(rp, K, device=dev))
    torch.cuda.synchronize(); dist.barrier()
    wfull = torch.empty(world * rp, K, dtype=torch.bfloat16, device=dev)
    dist.all_gather_into_tensor(wfull, wshard.contiguous())
    staging = torch.zeros(world * rp, K, dtype=torch.bfloat16, device=dev)
    counters = torch.zeros(world * rp // 64, dtype=torch.int32, device=dev)
    spec = fc.gather_spec(wshard, staging, counters)
    x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    spec.epoch = 1
    y = gemm(x, staging, gather=spec)                       # NT: pulls while computing
    yr = gemm(x, wfull)
    torch.cuda.synchronize()
    remote = torch.ones(world * rp, dtype=torch.bool, device=dev)
    remote[rank * rp:(rank + 1) * rp] = False          # own rows are read in place, never staged
    check("gather-GEMM NT == GEMM on gathered weight", torch.equal(y, yr) and torch.equal(staging[remote], wfull[remote]))
    dy = torch.randn(M, world * rp, device=dev, dtype=torch.bfloat16)
    dx = gemm(dy, staging, b_mn=True, gather=spec)          # same epoch: staging reused, no pull
    dxr = gemm_ref(dy, wfull, b_mn=True, out_dtype=torch.float32)
    torch.cuda.synchronize()
    check("gather-GEMM NN (reuse epoch)", ((dx.float() - dxr).abs().max() / dxr.abs().max()).item() < 2e-2)
    staging.zero_()
    spec.epoch = 2
    dx2 = gemm(dy, staging, b_mn=True, gather=spec)         # NN form pulling K blocks
    torch.cuda.synchronize()
    check("gather-GEMM NN (fresh epoch)", torch.equal(staging[remote], wfull[remote]) and
          ((dx2.float() - dxr).abs().max() / dxr.abs().max()).item() < 2e-2)
    fc.check_error()

    # training: fused vs nccl backends give the same losses (ZeRO-3 and ZeRO-2)
    from tests.test_trainer_gpu_common import run_short
    for stage in (3, 2):
        a = run_short(info, stage, "fused", fc)
        b = run_short(info, stage, "nccl", None)
        ok = all(abs(x - y) < 3e-2 for x, y in zip(a, b))
        if rank == 0:
            print(f"stage {stage} fused {a} nccl {b}", flush=True)
        check(f"ZeRO-{stage}: fused backend == nccl backend losses", ok)
    if rank == 0:
        print("ALL MULTI-GPU CHECKS PASSED", flush=True)
    comm.shutdown()


if __name__ == "__main__":
    main()
