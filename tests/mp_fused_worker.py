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

    # K1 on the CTA-pair kernel: gather jobs riding on GEMM launches + counter waits (csrc/gemm2.cu)
    from dlti_b200.ops.elementwise import embedding
    from dlti_b200.parallel.fused import GatherQueue, GatherUse, PeerRows, StagedMatrix
    q = GatherQueue(fc)
    for rp2, K2, M2 in ((256, 512, 384), (2752, 1024, 4096)):      # 2752 = Llama-2-7B w_gu rows per rank at 8 GPUs
        sh = fc.alloc_symmetric((rp2, K2), torch.bfloat16)
        sh.copy_(torch.randn(rp2, K2, device=dev) * 0.05)
        torch.cuda.synchronize(); dist.barrier()
        full = torch.empty(world * rp2, K2, dtype=torch.bfloat16, device=dev)
        dist.all_gather_into_tensor(full, sh.contiguous())
        x2 = torch.randn(M2, K2, device=dev, dtype=torch.bfloat16)
        dy2 = torch.randn(M2, world * rp2, device=dev, dtype=torch.bfloat16)
        sm = StagedMatrix(fc, rp2, K2, torch.bfloat16)
        sm.staging.zero_()
        # (a) cold start: the consuming launch pulls its own operand and waits on the counters tile by tile
        sm.bind(sh, "a"); q.push(sm)
        y2 = gemm(x2, sm.staging, gather=GatherUse(sm, q))
        torch.cuda.synchronize()
        check(f"gemm2 self-gather NT rp={rp2}", torch.equal(y2, gemm(x2, full, two_cta=True)) and torch.equal(sm.staging, full)
              and not q.pending)
        sm.staging.zero_()
        sm.bind(sh, "b"); q.push(sm)
        dx2 = gemm(dy2, sm.staging, b_mn=True, gather=GatherUse(sm, q))
        torch.cuda.synchronize()
        check(f"gemm2 self-gather NN rp={rp2}", torch.equal(dx2, gemm(dy2, full, b_mn=True, two_cta=True))
              and torch.equal(sm.staging, full))
        # (b) look-ahead: other launches carry the pulls slice by slice, the consumer then runs without waiting
        sm.staging.zero_()
        sm.bind(sh, "c"); q.push(sm)
        carried = 0
        while q.pending:
            gemm(x2, full, gather=GatherUse(None, q))
            carried += 1
        y3 = gemm(x2, sm.staging, gather=GatherUse(sm, q))
        torch.cuda.synchronize()
        check(f"gemm2 carried gather ({carried} launches) rp={rp2}", torch.equal(y3, y2) and torch.equal(sm.staging, full))
        # (c) leftovers flushed by the stand-alone gather kernel
        sm.staging.zero_()
        sm.bind(sh, "d"); q.push(sm)
        gemm(x2, full, gather=GatherUse(None, q))
        q.flush()
        torch.cuda.synchronize()
        check(f"gather job flush rp={rp2}", torch.equal(sm.staging, full) and not q.pending)
    fc.check_error()
    # K15: embedding rows fetched from their owner ranks
    vsh = fc.alloc_symmetric((192, 512), torch.bfloat16)
    vsh.copy_(torch.randn(192, 512, device=dev))
    torch.cuda.synchronize(); dist.barrier()
    vfull = torch.empty(world * 192, 512, dtype=torch.bfloat16, device=dev)
    dist.all_gather_into_tensor(vfull, vsh.contiguous())
    vocab = world * 192 - 5
    ids = torch.randint(0, vocab, (777,), device=dev)
    check("peer embedding == lookup in the gathered table", torch.equal(embedding(ids, PeerRows(fc, vsh, vocab)), vfull[ids]))

    # training: fused vs nccl backends give the same losses (ZeRO-3 and ZeRO-2)
    from tests.test_trainer_gpu_common import run_short
    for stage in (3, 2, 0):
        a, na = run_short(info, stage, "fused", fc, with_norms=True)
        b, nb = run_short(info, stage, "nccl", None, with_norms=True)
        ok = all(abs(x - y) < 3e-2 for x, y in zip(a, b)) and all(abs(x - y) <= 2e-2 * max(abs(y), 1e-3) for x, y in zip(na, nb))
        if rank == 0:
            print(f"stage {stage} fused {a} nccl {b}\n        grad norms fused {na} nccl {nb}", flush=True)
        check(f"ZeRO-{stage}: fused backend == nccl backend losses and gradient norms", ok)
    # K9, new kernels vs the round-1 sequence on IDENTICAL gradients: bit-identical partition sums, norm,
    # moments and parameters (same reduction order, same arithmetic), and the loss sum rides along
    nn_ = world * 4096 * 4
    pn = nn_ // world
    gsrc = torch.randn(nn_, device=dev) * (0.1 + rank)
    res = []
    for path in ("tail", "fused"):
        gbuf = fc.alloc_symmetric((nn_,), torch.float32)
        mst = fc.alloc_symmetric((nn_,), torch.float32)
        gbuf.copy_(gsrc)
        mst.copy_(torch.linspace(-1, 1, nn_, device=dev))
        torch.cuda.synchronize(); dist.barrier()
        mm, vv = torch.full((pn,), 0.01, device=dev), torch.full((pn,), 0.02, device=dev)
        part = torch.zeros(pn, device=dev)
        sk = torch.zeros(1, dtype=torch.int32, device=dev)
        kw = dict(lr=1e-2, betas=(0.9, 0.999), eps=1e-8, wd=0.01, step=3, inv_scale=0.5, max_norm=1.0, skipped_out=sk)
        if path == "tail":
            nsq = torch.zeros(1, device=dev)
            fc.reduce_scatter_accumulate(gbuf, part)
            red = part.clone()
            fc.global_sumsq(part, nsq)
            fc.adamw_push(mst, mm, vv, part, rank * pn, pn, gnorm_sq=nsq, **kw)
            extra = None
        else:
            out2 = torch.zeros(2, device=dev)
            fc.k9_begin_step()
            fc.grad_range_final(0, gbuf, part, 0, nn_, rank * pn, pn, side=False, ctas=0)
            red = part.clone()
            fc.k9_step(mst, mm, vv, part, rank * pn, pn, loss_sum=torch.tensor([float(rank + 1)], device=dev), out2=out2, **kw)
            nsq, extra = out2[0:1].clone(), out2[1].item()
        torch.cuda.synchronize(); dist.barrier()
        res.append((red, nsq.clone(), mm, vv, mst.clone(), part.clone(), extra))
    (r0, n0, m0, v0, p0, g0, _), (r1, n1, m1, v1, p1, g1, ls) = res
    print(f"[rank {rank}] K9 unit: reduced {torch.equal(r0, r1)} norm {torch.equal(n0, n1)} ({n0.item()!r} vs {n1.item()!r}) "
          f"m {torch.equal(m0, m1)} v {torch.equal(v0, v1)} params {torch.equal(p0, p1)} "
          f"(max diff {(p0 - p1).abs().max().item():.3e}) zeroed {g0.abs().max().item()} {g1.abs().max().item()} loss {ls}", flush=True)
    check("K9 fused step == round-1 sequence, bit for bit (reduced gradients, norm, moments, parameters)",
          torch.equal(r0, r1) and torch.equal(n0, n1) and torch.equal(m0, m1) and torch.equal(v0, v1) and torch.equal(p0, p1)
          and g0.abs().max().item() == 0 and g1.abs().max().item() == 0 and ls == world * (world + 1) / 2)
    fc.check_error()
    # whole training runs: the three K9 modes agree to rounding (they cannot agree bit for bit over many steps: the
    # split-K weight-gradient GEMMs add their partial sums with fp32 atomics, in a different order on every run)
    ref_run = None
    for mode in ("tail", "fused", "overlap"):
        os.environ["DLTI_K9"] = mode
        la, na, ma = run_short(info, 3, "fused", fc, steps=6, with_master=True)
        if ref_run is None:
            ref_run = (la, na, ma)
        ok = all(abs(x - y) < 5e-3 for x, y in zip(la, ref_run[0])) and (ma - ref_run[2]).abs().max().item() < 2e-2
        if rank == 0:
            print(f"K9 {mode}: losses {la} max |dparam| vs tail {(ma - ref_run[2]).abs().max().item():.2e}", flush=True)
        check(f"K9 mode {mode}: 6 ZeRO-3 steps agree with the round-1 sequence", ok)
    os.environ.pop("DLTI_K9", None)
    if rank == 0:
        print("ALL MULTI-GPU CHECKS PASSED", flush=True)
    comm.shutdown()


if __name__ == "__main__":
    main()
