#!/usr/bin/env python
"""Multi-GPU micro-benchmark of K1 (all-gather fused into the GEMM): for each Llama-2-7B weight
shape, time (a) the plain GEMM on an already-gathered weight, (b) the gather-GEMM when the staging
is already valid (same epoch: pure overhead of the gather machinery), (c) the gather-GEMM pulling
the remote shards while computing (fresh epoch), (d) NCCL all-gather followed by the plain GEMM.
Launch: python -m dlti_b200.launch --num_gpus=N --no_local_rank scripts/bench_gather_gemm.py"""
import json, os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops.gemm import gemm
from dlti_b200.parallel import comm
from dlti_b200.parallel.fused import FusedComm

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
for name, R, Cc in [("w_qkv", 12288, 4096), ("w_o", 4096, 4096), ("w_gu", 22016, 4096), ("w_down", 4096, 11008)]:
    rp = -(-(-(-R // world)) // 64) * 64
    shard = fc.alloc_symmetric((rp, Cc), torch.bfloat16); shard.normal_(0, 0.02)
    torch.cuda.synchronize(); dist.barrier()
    full = torch.empty(world * rp, Cc, dtype=torch.bfloat16, device=dev)
    dist.all_gather_into_tensor(full, shard.contiguous())
    staging = torch.empty_like(full); counters = torch.zeros(world * rp // 64, dtype=torch.int32, device=dev)
    spec = fc.gather_spec(shard, staging, counters)
    x = torch.randn(M, Cc, device=dev, dtype=torch.bfloat16); y = torch.empty(M, R, device=dev, dtype=torch.bfloat16)
    dy = torch.randn(M, R, device=dev, dtype=torch.bfloat16); dx = torch.empty(M, Cc, device=dev, dtype=torch.bfloat16)
    res = {"w": name, "M": M, "rows": R, "cols": Cc, "world": world, "remote_MB": rp * Cc * 2 * (world - 1) / 1e6}
    res["plain_nt_us"] = timeit(lambda: gemm(x, full[:R], out=y))
    res["plain_nn_us"] = timeit(lambda: gemm(dy, full[:R], b_mn=True, out=dx))
    spec.epoch = 1; gemm(x, staging[:R], out=y, gather=spec)
    res["reuse_nt_us"] = timeit(lambda: gemm(x, staging[:R], out=y, gather=spec))
    res["reuse_nn_us"] = timeit(lambda: gemm(dy, staging[:R], b_mn=True, out=dx, gather=spec))
    def fresh():
        spec.epoch += 1; gemm(x, staging[:R], out=y, gather=spec)
    res["fresh_nt_us"] = timeit(fresh)
    def nccl():
        dist.all_gather_into_tensor(full, shard); gemm(x, full[:R], out=y)
    res["nccl_ag_then_gemm_us"] = timeit(nccl)
    res["pull_only_us"] = timeit(lambda: fc.all_gather_rows(shard, staging, blocks_per_peer=max(8, 96 // world)))
    res["link_floor_us"] = res["remote_MB"] * 1e6 / 770e9 * 1e6
    rows_out.append(res)
    if rank == 0: print(json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in res.items()}), flush=True)
fc.check_error()
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True); json.dump(rows_out, open(f"gpurun_out/gather_gemm_w{world}.json", "w"), indent=1)
comm.shutdown()
