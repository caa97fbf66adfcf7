"""ZeRO stage 0/1/2/3 optimizer over the flat LoRA parameter space.

DeepSpeed semantics restated for one flat fp32 buffer (models.lora.LoraParams):
  stage 0  every rank holds full optimizer state; gradients are all-reduced           (baseline)
  stage 1  optimizer state + fp32 master are partitioned 1/N; gradients are reduced to their
           owners at the accumulation boundary; updated slices are all-gathered         (D12)
  stage 2  as 1, and the reduced gradient lives only in the owner's partition buffer (grad_part).
           DeepSpeed reduces after every micro-batch; here the accumulation window runs layer-major
           (LoraLlama.forward_backward_multi), so a layer's gradient is final once per optimizer step
           and is reduced once per step. The full flat gradient buffer (67 MB for r=16 on 7B) stays
           allocated on every rank: it is the accumulation target of the weight-gradient GEMMs  (D13)
  stage 3  gradient/optimizer handling as stage 2 (the parameter partitioning of the frozen base
           weights lives in parallel.zero3.Zero3Weights); every LoRA tensor is below
           stage3_param_persistence_threshold, so adapters stay gathered                (D14)
Config keys honoured: reduce_bucket_size (staging-buffer bound for the reduce), allgather_bucket_size,
overlap_comm (reduce on a side stream), reduce_scatter, contiguous_gradients (always true here: the
flat buffer IS the contiguous gradient buffer), offload_optimizer (pinned host state + C++ Adam),
gradient_clipping, fp16 dynamic loss scaling.

Two communication backends: "nccl" = torch.distributed collectives (baseline; also gloo on CPU);
"fused" = one sm_100a kernel doing peer-memory reduce-scatter + norm, and one doing
unscale/clip/AdamW + all-gather push (parallel.fused; K9).
"""
from __future__ import annotations

import ctypes as C
import math
import os

import torch

from ..config.ds_config import DSConfig
from ..models.lora import LoraParams
from ..ops import _lib
from ..ops import optim as O
from . import comm
from .comm import DistInfo


class ZeroOptimizer:
    def __init__(self, lora: LoraParams, ds: DSConfig, info: DistInfo, *, lr: float,
                 betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                 max_grad_norm: float = 1.0, scaler=None, backend: str = "nccl", fused=None):
        self.lora, self.ds, self.info = lora, ds, info
        self.stage = ds.zero_stage
        self.lr, self.betas, self.eps, self.wd = lr, tuple(betas), eps, weight_decay
        self.max_grad_norm = max_grad_norm
        self.scaler = scaler
        self.backend = backend
        self.fused = fused                      # parallel.fused.FusedZeroComm or None
        self.world, self.rank = info.world_size, info.rank
        n = lora.padded_numel
        assert n % self.world == 0, "LoraParams must be padded to a multiple of world_size"
        self.part = n // self.world if self.stage >= 1 else n
        self.lo = self.rank * self.part if self.stage >= 1 else 0
        dev = lora.device
        self.offload = ds.offload_optimizer == "cpu" and self.stage >= 1
        sdev = "cpu" if self.offload else dev
        pin = self.offload and ds.pin_memory and torch.cuda.is_available()
        mk = lambda: (torch.zeros(self.part, dtype=torch.float32, device=sdev).pin_memory()
                      if pin else torch.zeros(self.part, dtype=torch.float32, device=sdev))
        self.m, self.v = mk(), mk()
        if self.offload:
            self.master_host = mk()
            self.master_host.copy_(self.master_part())
            self.grad_host = mk()
        # owner-side reduced gradient accumulator (stages >= 1)
        self.grad_part = torch.zeros(self.part, dtype=torch.float32, device=dev) if self.stage >= 1 else None
        self.gnorm_sq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.skipped_flag = torch.zeros(1, dtype=torch.int32, device=dev)
        self.step_count = 0                     # successful optimizer steps
        self.comm_stream = (torch.cuda.Stream(device=dev) if (dev.type == "cuda" and ds.overlap_comm)
                            else None)
        self._pending_reduce = None
        self.last_grad_norm = 0.0
        # K9 on the fused backend with partitioned gradients (DLTI_K9):
        #   "fused"   (default) at the end of backward: ready flags -> reduce-scatter kernel that waits on them ->
        #             sum of squares -> ONE kernel for norm/loss exchange + unscale/clip/AdamW + push + completion
        #             flags -> wait; no stand-alone barrier launches, 0.4 ms at 2 GPUs on Llama-2-7B r=16;
        #   "overlap" the same kernels, but each layer's range is reduced on a side stream as soon as that layer's
        #             backward is done. Measured SLOWER (2 GPUs: 684 vs 633 ms/step): the waiting reduce kernels
        #             share SMs with the persistent GEMMs and delay their CTA pairs; kept for experiments;
        #   "tail"    the round-1 sequence (barrier, reduce, barrier, norm exchange, AdamW push, barrier).
        mode = os.environ.get("DLTI_K9", "fused")
        k9_ok = fused is not None and self.world > 1 and self.stage >= 1 and not self.offload
        self.k9_fused = k9_ok and mode in ("fused", "overlap")
        self.overlap = k9_ok and mode == "overlap"
        self._ranges_done = 0
        self.out2 = torch.zeros(2, dtype=torch.float32, device=dev)
        self.loss_sum_global = None

    # ---- views -----------------------------------------------------------------------------
    def master_part(self) -> torch.Tensor:
        return self.lora.master[self.lo:self.lo + self.part]

    # ---- gradient reduction ------------------------------------------------------------------
    def _reduce_to_owner(self) -> None:
        """grad_part += sum_ranks grad[lo:hi]; staged in buckets of <= reduce_bucket_size elements."""
        g, info = self.lora.grad, self.info
        if self.world == 1:
            self.grad_part.add_(g)
            return
        if self.fused is not None:
            self.fused.reduce_scatter_accumulate(g, self.grad_part)
            return
        cs = max(1, min(self.part, self.ds.reduce_bucket_size // self.world))
        g2 = g.view(self.world, self.part)
        if cs >= self.part:
            tmp = torch.empty(self.part, dtype=torch.float32, device=g.device)
            comm.reduce_scatter_sum(g, tmp, info)
            self.grad_part.add_(tmp)
            return
        stage_buf = torch.empty(self.world, cs, dtype=torch.float32, device=g.device)
        out = torch.empty(cs, dtype=torch.float32, device=g.device)
        for c0 in range(0, self.part, cs):
            c1 = min(self.part, c0 + cs)
            w = c1 - c0
            sb = stage_buf[:, :w].contiguous() if w != cs else stage_buf
            sb.copy_(g2[:, c0:c1])
            comm.reduce_scatter_sum(sb.view(-1), out[:w], info)
            self.grad_part[c0:c1].add_(out[:w])

    def begin_step(self) -> None:
        if self.k9_fused:
            self.fused.k9_begin_step()
            self._ranges_done = 0

    def on_grad_range_final(self, glo: int, n: int) -> None:
        """Model hook (LoraLlama.grad_hook): the flat gradient range [glo, glo + n) will not be written again in
        this accumulation window (one decoder layer's adapters, all micro-batches done)."""
        if not self.overlap:
            return
        self.fused.grad_range_final(self._ranges_done, self.lora.grad, self.grad_part, glo, n, self.lo, self.part)
        self._ranges_done += 1

    def on_micro_batch_end(self, is_boundary: bool) -> None:
        """Called after each backward. Stage >= 2 reduces now (and frees local grads); stage 1
        reduces at the accumulation boundary; stage 0 all-reduces at the boundary."""
        if self.overlap and self._ranges_done > 0:
            return        # every range was reduced behind its layer's backward; step() joins the side stream
        if self.k9_fused:
            if is_boundary:   # the whole flat buffer as one range, on the main stream
                self.fused.grad_range_final(0, self.lora.grad, self.grad_part, 0, self.lora.padded_numel, self.lo,
                                            self.part, side=False, ctas=0)
                self._ranges_done = 1
            return
        if self.stage >= 2 or (self.stage == 1 and is_boundary):
            if self.comm_stream is not None and self.fused is None:
                self.comm_stream.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(self.comm_stream):
                    self._reduce_to_owner()
                    self.lora.grad.zero_()
                self._pending_reduce = torch.cuda.Event()
                self._pending_reduce.record(self.comm_stream)
            else:
                self._reduce_to_owner()
                self.lora.grad.zero_()
        elif self.stage == 0 and is_boundary and self.world > 1:
            comm.all_reduce_sum_(self.lora.grad, self.info)

    def wait_reduce(self) -> None:
        """Backward of the next micro-batch writes lora.grad: it must follow the side-stream reduce."""
        if self._pending_reduce is not None:
            torch.cuda.current_stream().wait_event(self._pending_reduce)
            self._pending_reduce = None

    # ---- the step ----------------------------------------------------------------------------
    def step(self, lr: float, loss_sum: torch.Tensor | None = None) -> tuple[bool, float]:
        """Returns (skipped, grad_norm). Overflow => skipped, loss scale updated by the caller. loss_sum: this
        rank's loss sum of the step; when the fused step kernel runs, self.loss_sum_global holds the sum over
        ranks afterwards (it travels with the norm exchange), else it stays None."""
        self.wait_reduce()
        info = self.info
        g = self.grad_part if self.stage >= 1 else self.lora.grad
        inv_scale = 1.0 / (self.scaler.cur_scale if self.scaler is not None else 1.0)
        self.loss_sum_global = None
        if self.k9_fused and self._ranges_done > 0:
            step = self.step_count + 1
            ls = loss_sum if loss_sum is not None else torch.zeros(1, dtype=torch.float32, device=g.device)
            self.fused.k9_step(self.lora.master, self.m, self.v, g, self.lo, self.part, lr=lr, betas=self.betas,
                               eps=self.eps, wd=self.wd, step=step, loss_sum=ls.float(), out2=self.out2,
                               inv_scale=inv_scale, max_norm=self.max_grad_norm, skipped_out=self.skipped_flag)
            self.lora.grad.zero_()          # every owner has read this rank's gradients (all pushes are complete)
            vals = torch.cat([self.out2, self.skipped_flag.float()]).tolist()
            nsq, skipped = vals[0], bool(vals[2])
            if loss_sum is not None:
                self.loss_sum_global = vals[1]
            if not skipped:
                self.step_count = step
                self.lora.refresh_compute()
            self.last_grad_norm = math.sqrt(nsq) * inv_scale if math.isfinite(nsq) else float("nan")
            self._ranges_done = 0
            return skipped, self.last_grad_norm
        # global sum of squares (C6 + C7 in one reduction: non-finite => overflow)
        if self.fused is not None and self.world > 1 and self.stage >= 1:
            # partitioned gradients: every rank holds a different slice, the norm is the sum over ranks.
            # (stage 0 has already all-reduced the FULL gradient: each rank's local sum is the global one)
            self.fused.global_sumsq(g, self.gnorm_sq)
        else:
            O.sumsq(g, out=self.gnorm_sq)
            if self.stage >= 1:
                comm.all_reduce_sum_(self.gnorm_sq, info)
        step = self.step_count + 1
        pushed = False
        if self.offload:
            # ZeRO-Offload optimizer step (D15): the partition's gradients go to pinned host memory, the AVX-512
            # Adam steps the host copy of master / moments, the new master comes back — pipelined in chunks on a
            # copy stream so that the D2H of chunk i+1 and the H2D of chunk i-1 overlap the CPU update of chunk i.
            # The 16-bit compute copies are not written by the CPU kernel (its optional p16 output): they live in
            # packed, pre-scaled per-layer buffers that one device kernel produces from the fp32 master after the
            # H2D (LoraParams.refresh_compute), which moves fewer bytes over PCIe than shipping both.
            nsq = float(self.gnorm_sq.item())
            f = _lib.fn("dlti_cpu_adamw", [C.c_void_p] * 5 + [C.c_int, C.c_longlong] + [C.c_float] * 5 +
                        [C.c_int, C.c_float, C.c_int, C.c_float, C.c_float, C.c_int], C.c_int, cpu=True)
            skipped = not math.isfinite(nsq)
            on_gpu = g.is_cuda
            nch = 4 if (on_gpu and self.part >= (1 << 20)) else 1
            bounds = [self.part * i // nch // 64 * 64 for i in range(nch)] + [self.part]
            if on_gpu:
                if getattr(self, "_copy_stream", None) is None:
                    self._copy_stream = torch.cuda.Stream(device=g.device)
                cs = self._copy_stream
                cs.wait_stream(torch.cuda.current_stream())
                d2h = []
                with torch.cuda.stream(cs):
                    for i in range(nch):
                        a, b = bounds[i], bounds[i + 1]
                        self.grad_host[a:b].copy_(g[a:b], non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(cs)
                        d2h.append(ev)
            else:
                self.grad_host.copy_(g)
            for i in range(nch):
                a, b = bounds[i], bounds[i + 1]
                if on_gpu:
                    d2h[i].synchronize()
                if not skipped and b > a:
                    off = 4 * a
                    f(self.master_host.data_ptr() + off, self.m.data_ptr() + off, self.v.data_ptr() + off,
                      self.grad_host.data_ptr() + off, None, 0, b - a, lr, self.betas[0], self.betas[1], self.eps,
                      self.wd, step, nsq, 1, inv_scale, float(self.max_grad_norm), 0)
                    if on_gpu:
                        with torch.cuda.stream(cs):
                            self.master_part()[a:b].copy_(self.master_host[a:b], non_blocking=True)
                    else:
                        self.master_part()[a:b].copy_(self.master_host[a:b])
            if on_gpu:
                torch.cuda.current_stream().wait_stream(cs)
        elif self.fused is not None and self.world > 1 and self.stage >= 1:
            # K9: AdamW on the owned slice, new values pushed into every peer's master buffer
            self.fused.adamw_push(self.lora.master, self.m, self.v, g, self.lo, self.part, lr=lr,
                                  betas=self.betas, eps=self.eps, wd=self.wd, step=step,
                                  gnorm_sq=self.gnorm_sq, inv_scale=inv_scale,
                                  max_norm=self.max_grad_norm, skipped_out=self.skipped_flag)
            nsq, sk = torch.cat([self.gnorm_sq, self.skipped_flag.float()]).tolist()
            skipped = bool(sk)
            pushed = True
        else:
            r = O.adamw_step(self.master_part() if self.stage >= 1 else self.lora.master, self.m, self.v, g,
                             lr=lr, betas=self.betas, eps=self.eps, weight_decay=self.wd, step=step,
                             gnorm_sq=self.gnorm_sq, inv_scale=inv_scale, max_norm=self.max_grad_norm,
                             skipped_out=self.skipped_flag)
            if r is None:   # CUDA: decision is on the device; one small read-back per optimizer step
                nsq, sk = torch.cat([self.gnorm_sq, self.skipped_flag.float()]).tolist()
                skipped = bool(sk)
            else:
                nsq, skipped = float(self.gnorm_sq.item()), bool(r)
        g.zero_()
        if not skipped:
            self.step_count = step
            if self.stage >= 1 and self.world > 1 and not pushed:
                self._all_gather_master()
            self.lora.refresh_compute()
        self.last_grad_norm = math.sqrt(nsq) * inv_scale if math.isfinite(nsq) else float("nan")
        return skipped, self.last_grad_norm

    def _all_gather_master(self) -> None:
        cs = max(1, min(self.part, self.ds.allgather_bucket_size // self.world))
        full, info = self.lora.master, self.info
        if cs >= self.part:
            comm.all_gather_into(full, self.master_part(), info)
            return
        f2 = full.view(self.world, self.part)
        buf = torch.empty(self.world, cs, dtype=torch.float32, device=full.device)
        for c0 in range(0, self.part, cs):
            c1 = min(self.part, c0 + cs)
            w = c1 - c0
            b = buf if w == cs else torch.empty(self.world, w, dtype=torch.float32, device=full.device)
            comm.all_gather_into(b.view(-1), self.master_part()[c0:c1].contiguous(), info)
            f2[:, c0:c1].copy_(b)

    # ---- checkpoint --------------------------------------------------------------------------
    def state_dict(self) -> dict:
        """Per-rank optimizer partition (DeepSpeed writes one *_optim_states.pt per rank)."""
        src = self.master_host if self.offload else (self.master_part() if self.stage >= 1 else self.lora.master)
        return {"zero_stage": self.stage, "partition_count": self.world if self.stage >= 1 else 1,
                "rank": self.rank, "part": self.part, "step_count": self.step_count,
                "fp32_master": src.detach().cpu().clone(), "exp_avg": self.m.detach().cpu().clone(),
                "exp_avg_sq": self.v.detach().cpu().clone(),
                "loss_scaler": self.scaler.state_dict() if self.scaler is not None else None}

    def load_state_dict(self, sd: dict) -> None:
        if sd["part"] != self.part or sd["zero_stage"] != self.stage:
            raise ValueError("optimizer checkpoint was written with a different world size / ZeRO stage "
                             f"(part {sd['part']} vs {self.part}, stage {sd['zero_stage']} vs {self.stage})")
        self.step_count = int(sd["step_count"])
        self.m.copy_(sd["exp_avg"])
        self.v.copy_(sd["exp_avg_sq"])
        if self.offload:
            self.master_host.copy_(sd["fp32_master"])
        (self.master_part() if self.stage >= 1 else self.lora.master).copy_(sd["fp32_master"])
        if self.scaler is not None and sd.get("loss_scaler"):
            self.scaler.load_state_dict(sd["loss_scaler"])
        if self.stage >= 1 and self.world > 1:
            self._all_gather_master()
        self.lora.refresh_compute()
