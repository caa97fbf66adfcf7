#!/usr/bin/env python
"""Multi-GPU micro-benchmark of K1 on the CTA-pair GEMM (csrc/gemm2.cu), per Llama-2-7B weight shape:
  plain        the GEMM on an already complete weight
  carried      the same GEMM carrying a gather job for ANOTHER matrix (the steady state of a ZeRO-3 step: the next
               layer's shards ride on this layer's GEMMs); overhead = carried - plain
  self         the GEMM pulling its OWN operand and waiting on the unit counters tile by tile (cold start)
  pull_only    the stand-alone gather kernel: peer-pull bandwidth = remote bytes / time
  nccl_then    NCCL all_gather_into_tensor followed by the plain GEMM (what the NCCL backend does)
Launch: python -m dlti_b200.launch --num_gpus=N --no_local_rank scripts/bench_gather2.py"""
import json, os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops.gemm import gemm
from dlti_b200.parallel import comm
from dlti_b200.parallel.fused import FusedComm, GatherQueue, GatherUse, StagedMatrix

info = comm.init_distributed("cuda"); fc = FusedComm(info); dev = info.device
rank, world = info.rank, info.world_size
M = int(os.environ.get("M", 4096))


def timeit(fn, iters=20, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e3


rows_out = []
q = GatherQueue(fc, pull_gbs=1e9)      # unlimited slice size: one job carries the whole matrix
for name, R, Cc in [("w_qkv", 12288, 4096), ("w_o", 4096, 4096), ("w_gu", 22016, 4096), ("w_down", 4096, 11008)]:
    rp = -(-(-(-R // world)) // 64) * 64
    shard = fc.alloc_symmetric((rp, Cc), torch.bfloat16); shard.normal_(0, 0.02)
    torch.cuda.synchronize(); dist.barrier()
    full = torch.empty(world * rp, Cc, dtype=torch.bfloat16, device=dev)
    dist.all_gather_into_tensor(full, shard.contiguous())
    sm = StagedMatrix(fc, rp, Cc, torch.bfloat16)
    other = StagedMatrix(fc, rp, Cc, torch.bfloat16)
    x = torch.randn(M, Cc, device=dev, dtype=torch.bfloat16); y = torch.empty(M, R, device=dev, dtype=torch.bfloat16)
    res = {"w": name, "M": M, "rows": R, "cols": Cc, "world": world, "remote_MB": rp * Cc * 2 * (world - 1) / 1e6}
    res["plain_us"] = timeit(lambda: gemm(x, full[:R], out=y, two_cta=True))

    def carried():
        other.bind(shard, "x"); q.push(other)
        gemm(x, full[:R], out=y, gather=GatherUse(None, q))
    res["carried_us"] = timeit(carried)

    def cold():
        sm.bind(shard, "y"); q.push(sm)
        gemm(x, sm.staging[:R], out=y, gather=GatherUse(sm, q))
    res["self_us"] = timeit(cold)

    def pull():
        sm.bind(shard, "z"); q.push(sm); q.flush()
    res["pull_only_us"] = timeit(pull)
    res["pull_GBps"] = res["remote_MB"] * 1e6 / (res["pull_only_us"] * 1e-6) / 1e9

    def nccl():
        dist.all_gather_into_tensor(full, shard); gemm(x, full[:R], out=y, two_cta=True)
    res["nccl_then_gemm_us"] = timeit(nccl)
    res["link_floor_us_at_770GBps"] = res["remote_MB"] * 1e6 / 770e9 * 1e6
    rows_out.append(res)
    if rank == 0: print(json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in res.items()}), flush=True)
fc.check_error()
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True); json.dump(rows_out, open(f"gpurun_out/gather2_w{world}.json", "w"), indent=1)
comm.shutdown()
