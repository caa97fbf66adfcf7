#!/usr/bin/env python
"""Decode-step micro-benchmark (no HTTP): device time of one CUDA-graph decode step of the serving
model at several batch sizes and a fixed context, vs the weight-streaming floor
(local weight bytes / measured HBM copy bandwidth). Single GPU or TP (launch one process per GPU)."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scripts.serve import add_args, build_engine

ap = argparse.ArgumentParser(); add_args(ap)
ap.add_argument("--batches", default="1,8,32,64,128"); ap.add_argument("--ctx", type=int, default=512)
ap.add_argument("--iters", type=int, default=30)
a = ap.parse_args()
a.no_warmup = True          # this script drives the ranks step by step itself
info, cfg, eng = build_engine(a)
bs = eng.blocks.block_size
peaks = {}
try: peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
except Exception: pass
hbm = peaks.get("hbm_gbs", 6650.0)
wbytes = sum(t.numel() * t.element_size() for l in eng.w.layers for t in l.values()) + eng.w.lm_head.numel() * 2
rows = []
for B in [int(x) for x in a.batches.split(",")]:
    nblk = -(-(a.ctx + 1) // bs)
    mk = lambda: eng.pack_step(B, [5] * B, [a.ctx] * B, [(i * nblk + a.ctx // bs) * bs + a.ctx % bs for i in range(B)],
                               [a.ctx + 1] * B, [list(range(i * nblk, (i + 1) * nblk)) for i in range(B)], [], 0,
                               ([0.0] * B, [0] * B, [1.0] * B, [0] * B))
    if info.rank == 0: meta = eng._broadcast(mk())
    else: meta = eng._broadcast(None)
    for _ in range(3): eng._run_step(meta)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(a.iters):
        if info.rank == 0: meta = eng._broadcast(mk())
        else: meta = eng._broadcast(None)
        eng._run_step(meta)
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / a.iters
    kvb = 2 * cfg.num_hidden_layers * (a.ctx + 1) * eng.w.KV * cfg.head_dim * 2 * B
    floor = (wbytes + kvb) / (hbm * 1e9) * 1e3
    rows.append({"batch": B, "ctx": a.ctx, "tp": info.world_size, "ms_per_step": ms, "tok_per_s": B / ms * 1e3,
                 "floor_ms": floor, "frac_of_hbm_floor": floor / ms})
    if info.rank == 0: print(json.dumps(rows[-1]), flush=True)
if info.rank == 0:
    os.makedirs("gpurun_out", exist_ok=True); json.dump(rows, open(f"gpurun_out/decode_bench_tp{info.world_size}.json", "w"), indent=1)

