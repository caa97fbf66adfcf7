// Peer-memory communication kernels over NVLink 5 / NVSwitch (symmetric arena mapped into every
// rank; see dlti_b200/parallel/fused.py). These replace the NCCL collectives the reference's stack
// issues through torch.distributed on the "fused" backend:
//
//   peer_barrier        cross-rank barrier on signal words (st.release.sys / ld.acquire.sys)
//   allgather_rows      pull-based all-gather of row-sharded weights (ZeRO-3 C4, stand-alone form;
//                       the tile-by-tile fused form lives in gemm.cu, GATHER=true)
//   rs_accumulate       reduce-scatter by peer reads: part += sum_r grad_r[lo:hi], fixed rank order
//                       (deterministic) — C2/C5
//   sumsq_exchange      global gradient norm / overflow flag: every rank publishes its partial sum
//                       of squares to all peers and sums the board in rank order — C6 + C7
//   adamw_push          unscale + clip + AdamW on the owned partition, and the updated fp32 values
//                       are stored straight into every peer's parameter buffer (all-gather by
//                       push) — O11/O12 + C3/C7b
//
// Every spin-wait has a clock-based timeout (~10 s) that raises an error word instead of hanging
// the GPU when a peer never arrives.
#include "common.cuh"

namespace {

constexpr int MAX_RANKS = 16;
constexpr long long SPIN_TIMEOUT_CYCLES = 20000000000LL;   // ~10 s at ~2 GHz

struct Peers {
  void* p[MAX_RANKS];
};

__device__ __forceinline__ bool wait_ge_sys(const uint32_t* flag, uint32_t target, int* err) {
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys_u32(flag) - target) < 0) {
    if (clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
      if (err) atomicExch(err, 1);
      return false;
    }
    __nanosleep(64);
  }
#ifdef DLTI_DEBUG_CHECKS
  // Signal-pad sequence check (debug build): every rank runs the same sequence of barriers on a
  // channel and cannot leave barrier k+1 before this rank has entered it, so a peer's flag may be at `target` or
  // one ahead — anything else means two code paths share a channel or a rank skipped / repeated a barrier.
  if (ld_acquire_sys_u32(flag) - target > 1u && err) atomicExch(err, 2);
#endif
  return true;
}

// One CTA, one thread per rank. pads.p[r] = rank r's signal array (uint32), slot layout:
// [channel * MAX_RANKS + source_rank]. Monotonic epochs: a flag only ever grows.
__global__ void peer_barrier_kernel(Peers pads, int rank, int world, int channel, uint32_t epoch,
                                    int* err) {
  const int t = threadIdx.x;
  if (t < world) {
    __threadfence_system();
    uint32_t* remote = reinterpret_cast<uint32_t*>(pads.p[t]) + channel * MAX_RANKS + rank;
    st_release_sys_u32(remote, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(pads.p[rank]) + channel * MAX_RANKS + t;
    wait_ge_sys(mine, epoch, err);
  }
}

// Same barrier with the epoch kept in device memory (incremented by the kernel itself), so the launch
// carries no per-call host value and can be replayed from a CUDA graph. All ranks execute the same
// sequence of barriers on a channel, so their counters advance in lockstep.
__global__ void peer_barrier_auto_kernel(Peers pads, int rank, int world, int channel, uint32_t* counter,
                                         int* err) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = atomicAdd(counter, 1u) + 1u;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  const int t = threadIdx.x;
  if (t < world) {
    __threadfence_system();
    st_release_sys_u32(reinterpret_cast<uint32_t*>(pads.p[t]) + channel * MAX_RANKS + rank, epoch);
    wait_ge_sys(reinterpret_cast<const uint32_t*>(pads.p[rank]) + channel * MAX_RANKS + t, epoch, err);
  }
}

// out[r * shard_bytes ...] <- shard of rank r, 16-byte vectors, 4 loads in flight per thread.
// grid = (blocks_per_peer, world). Weights are immutable during a step: non-coherent loads.
__global__ void __launch_bounds__(256) allgather_rows_kernel(Peers shards, uint8_t* __restrict__ out,
                                                            long long shard_bytes, int rank) {
  const int r = blockIdx.y;
  const int4* src = reinterpret_cast<const int4*>(shards.p[r]);
  int4* dst = reinterpret_cast<int4*>(out + (long long)r * shard_bytes);
  const long long n = shard_bytes >> 4;
  const long long stride = (long long)gridDim.x * 256;
  long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    int4 a = ld_nc_v4(src + i), b = ld_nc_v4(src + i + stride), c = ld_nc_v4(src + i + 2 * stride),
         d = ld_nc_v4(src + i + 3 * stride);
    dst[i] = a; dst[i + stride] = b; dst[i + 2 * stride] = c; dst[i + 3 * stride] = d;
  }
  for (; i < n; i += stride) dst[i] = ld_nc_v4(src + i);
}

// part[i] += sum_{r=0..world-1} grads.p[r][lo + i]   (fp32; fixed order => bitwise reproducible)
__global__ void __launch_bounds__(256) rs_accumulate_kernel(Peers grads, float* __restrict__ part,
                                                           long long lo, long long n, int world) {
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    float4 acc = reinterpret_cast<float4*>(part)[i];
    for (int r = 0; r < world; ++r) {
      const int4 v = ld_relaxed_sys_v4(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(grads.p[r]) + lo) + i);
      acc.x += __int_as_float(v.x); acc.y += __int_as_float(v.y);
      acc.z += __int_as_float(v.z); acc.w += __int_as_float(v.w);
    }
    reinterpret_cast<float4*>(part)[i] = acc;
  }
}

// boards.p[r] = rank r's float board [MAX_RANKS]. Publish my partial to every peer, signal, wait for
// all, sum in rank order. One CTA.
__global__ void sumsq_exchange_kernel(const float* __restrict__ local, Peers boards, Peers pads, int rank,
                                      int world, int channel, uint32_t epoch, float* __restrict__ out,
                                      int* err) {
  const int t = threadIdx.x;
  __shared__ int ok;
  if (t == 0) ok = 1;
  __syncthreads();
  if (t < world) {
    float* rb = reinterpret_cast<float*>(boards.p[t]) + rank;
    asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(rb), "f"(*local) : "memory");
    __threadfence_system();
    st_release_sys_u32(reinterpret_cast<uint32_t*>(pads.p[t]) + channel * MAX_RANKS + rank, epoch);
    if (!wait_ge_sys(reinterpret_cast<const uint32_t*>(pads.p[rank]) + channel * MAX_RANKS + t, epoch, err))
      ok = 0;
  }
  __syncthreads();
  if (t == 0) {
    float s = 0.f;
    const float* mb = reinterpret_cast<const float*>(boards.p[rank]);
    for (int r = 0; r < world; ++r) {
      float v;
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(mb + r) : "memory");
      s += v;
    }
    *out = ok ? s : __int_as_float(0x7fc00000);
  }
}

// AdamW on the owned partition [lo, lo+n) of the flat parameter space; the new fp32 value is written
// to masters.p[r][lo + i] for EVERY rank r (own copy included).
__global__ void __launch_bounds__(256) adamw_push_kernel(Peers masters, float* __restrict__ m,
                                                        float* __restrict__ v, float* __restrict__ g,
                                                        long long lo, long long n, int rank, int world,
                                                        float lr, float b1, float b2, float eps, float wd,
                                                        float bc1, float bc2,
                                                        const float* __restrict__ gnorm_sq, float inv_scale,
                                                        float max_norm, int* __restrict__ skipped_out) {
  float coef = inv_scale;
  const float nsq = *gnorm_sq;
  if (!isfinite(nsq)) {
    if (blockIdx.x == 0 && threadIdx.x == 0 && skipped_out) *skipped_out = 1;
    return;
  }
  if (max_norm > 0.f) coef *= fminf(1.f, max_norm / (sqrtf(nsq) * inv_scale + 1e-6f));
  if (blockIdx.x == 0 && threadIdx.x == 0 && skipped_out) *skipped_out = 0;
  const float ibc1 = 1.f / bc1, ibc2 = 1.f / bc2;
  float* mine = reinterpret_cast<float*>(masters.p[rank]) + lo;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const float ge = __fmul_rn(g[i], coef);
    float mi = m[i], vi = v[i];
    const float pi = adamw_update(mine[i], mi, vi, ge, lr, b1, b2, eps, wd, ibc1, ibc2);
    m[i] = mi; v[i] = vi; g[i] = 0.f;
    for (int r = 0; r < world; ++r) (reinterpret_cast<float*>(masters.p[r]) + lo)[i] = pi;
  }
}

// K10/K11 consumer: residual[row] += sum_{r < world} recv[r][row]  (fp32, fixed rank order => every
// rank computes bit-identical sums), then y[row] = w * T(residual * rstd). One CTA per token row.
//
// BARRIER = true folds the rank barrier that must separate the pushes from the reduction into this kernel
// (one launch less per row-parallel GEMM of a latency-bound decode step): CTA 0 tells every peer "my pushes of
// this round are complete" (it runs after this rank's push GEMM: stream order, or griddep_wait() under PDL),
// and every CTA waits until all peers have said so. The epoch is a device counter read by all CTAs and bumped
// by the last CTA to finish, so the launch carries no per-call host value (CUDA-graph replayable); it shares
// channel and counter with dlti_peer_barrier_auto, whose sequence it continues.
struct TpBarrier {
  Peers pads;
  uint32_t* counter;     // this channel's epoch counter (device)
  uint32_t* done;        // zero-initialised scratch word: CTAs finished in this launch
  int* err;
  int rank, channel;
};

template <typename T, int VPT, bool BARRIER>
__global__ void __launch_bounds__(128) tp_reduce_add_norm_kernel(const T* __restrict__ recv, long long slot_stride,
                                                                int world, T* __restrict__ residual,
                                                                const T* __restrict__ w, T* __restrict__ y,
                                                                int d, float eps, const TpBarrier tb) {
  __shared__ float red[32];
  if constexpr (BARRIER) {
    griddep_wait();
    griddep_launch();
    const int t = threadIdx.x;
    if (t < world) {
      const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(tb.counter) + 1u;
      if (blockIdx.x == 0) {
        __threadfence_system();
        st_release_sys_u32(reinterpret_cast<uint32_t*>(tb.pads.p[t]) + tb.channel * MAX_RANKS + tb.rank, epoch);
      }
      wait_ge_sys(reinterpret_cast<const uint32_t*>(tb.pads.p[tb.rank]) + tb.channel * MAX_RANKS + t, epoch, tb.err);
    }
    __syncthreads();
  }
  const long long row = blockIdx.x;
  const int nvec = d >> 3;
  float f[VPT][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      {
        const int4 rv = reinterpret_cast<const int4*>(residual + row * d)[v];
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&rv);
#pragma unroll
        for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(u[j]); f[i][2 * j] = t.x; f[i][2 * j + 1] = t.y; }
      }
      for (int r = 0; r < world; ++r) {
        const int4 pv = reinterpret_cast<const int4*>(recv + r * slot_stride + row * d)[v];
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&pv);
#pragma unroll
        for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(u[j]); f[i][2 * j] += t.x; f[i][2 * j + 1] += t.y; }
      }
      int4 sv;
      uint32_t* su = reinterpret_cast<uint32_t*>(&sv);
#pragma unroll
      for (int j = 0; j < 4; ++j) su[j] = Cvt<T>::pack2(f[i][2 * j], f[i][2 * j + 1]);
      reinterpret_cast<int4*>(residual + row * d)[v] = sv;
#pragma unroll
      for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(su[j]); f[i][2 * j] = t.x; f[i][2 * j + 1] = t.y; }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += f[i][j] * f[i][j];
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = (threadIdx.x & 31) < 4 ? red[threadIdx.x & 31] : 0.f;
  tot = warp_sum(tot);
  const float rstd = rsqrtf(tot / (float)d + eps);
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      const int4 wv = reinterpret_cast<const int4*>(w)[v];
      const uint32_t* wu = reinterpret_cast<const uint32_t*>(&wv);
      int4 ov;
      uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 wf = Cvt<T>::unpack2(wu[j]);
        const float a = wf.x * Cvt<T>::to_f(Cvt<T>::from_f(f[i][2 * j] * rstd));
        const float b = wf.y * Cvt<T>::to_f(Cvt<T>::from_f(f[i][2 * j + 1] * rstd));
        ou[j] = Cvt<T>::pack2(a, b);
      }
      reinterpret_cast<int4*>(y + row * d)[v] = ov;
    }
  }
  if constexpr (BARRIER) {
    __syncthreads();             // every thread of this CTA has read the counter
    if (threadIdx.x == 0) {
      if (atomicAdd(tb.done, 1u) == gridDim.x - 1) {       // last CTA of the launch: next round
        *tb.done = 0u;
        __threadfence();
        atomicAdd(tb.counter, 1u);
      }
    }
  }
}

// Prefill-sized row-parallel all-reduce as reduce-scatter + all-gather over peer memory. Every rank has
// written its partial [T, d] into its own symmetric buffer (plain fast GEMM, no scattered peer stores from an
// epilogue). This kernel owns rows [row_lo, row_lo + gridDim.x): per row it PULLS the partials of all ranks in
// rank order (coalesced 16-byte peer loads), adds the residual (kept only by the owner rank: the residual
// stream is row-sharded on this path), normalises and PUSHES the normalised row into every rank's activation
// buffer (coalesced 16-byte peer stores). Link traffic per rank: 2 * (world-1)/world * T*d elements instead of
// (world-1) * T*d for the push-everything-to-everyone scheme of the decode path.
template <typename T, int VPT>
__global__ void __launch_bounds__(128) tp_pull_reduce_norm_push_kernel(Peers partials, Peers xn_out, int world,
                                                                      long long row_lo, T* __restrict__ residual,
                                                                      const T* __restrict__ w, int d, float eps) {
  __shared__ float red[32];
  const long long row = row_lo + blockIdx.x;
  const int nvec = d >> 3;
  float f[VPT][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      {
        const int4 rv = reinterpret_cast<const int4*>(residual + row * d)[v];
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&rv);
#pragma unroll
        for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(u[j]); f[i][2 * j] = t.x; f[i][2 * j + 1] = t.y; }
      }
      for (int r = 0; r < world; ++r) {
        const int4 pv = ld_relaxed_sys_v4(reinterpret_cast<const T*>(partials.p[r]) + row * d + v * 8);
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&pv);
#pragma unroll
        for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(u[j]); f[i][2 * j] += t.x; f[i][2 * j + 1] += t.y; }
      }
      int4 sv;
      uint32_t* su = reinterpret_cast<uint32_t*>(&sv);
#pragma unroll
      for (int j = 0; j < 4; ++j) su[j] = Cvt<T>::pack2(f[i][2 * j], f[i][2 * j + 1]);
      reinterpret_cast<int4*>(residual + row * d)[v] = sv;
#pragma unroll
      for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(su[j]); f[i][2 * j] = t.x; f[i][2 * j + 1] = t.y; }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += f[i][j] * f[i][j];
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = (threadIdx.x & 31) < 4 ? red[threadIdx.x & 31] : 0.f;
  tot = warp_sum(tot);
  const float rstd = rsqrtf(tot / (float)d + eps);
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      const int4 wv = reinterpret_cast<const int4*>(w)[v];
      const uint32_t* wu = reinterpret_cast<const uint32_t*>(&wv);
      int4 ov;
      uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 wf = Cvt<T>::unpack2(wu[j]);
        const float a = wf.x * Cvt<T>::to_f(Cvt<T>::from_f(f[i][2 * j] * rstd));
        const float b = wf.y * Cvt<T>::to_f(Cvt<T>::from_f(f[i][2 * j + 1] * rstd));
        ou[j] = Cvt<T>::pack2(a, b);
      }
      for (int r = 0; r < world; ++r) st_v4(reinterpret_cast<T*>(xn_out.p[r]) + row * d + v * 8, ov);
    }
  }
  __threadfence_system();
}

// ---------------------------------------------------------------------------------------------------------
// K9, overlapped form: adapter-gradient reduce-scatter as soon as a layer's gradients are final, and one
// kernel for norm exchange + unscale/clip + AdamW + all-gather push.
//
//   grad_ready_signal     (main stream, after a layer's backward): tell every rank "my gradients of range
//                         `slot` are final for step `epoch`" (one release store per peer).
//   rs_wait_accumulate    (side stream): the owner of (part of) that range waits for all ranks' signals — they
//                         arrive while the remaining layers' backward is still running — and adds the peers'
//                         slices to its partition accumulator with peer loads in fixed rank order (bitwise
//                         reproducible; same order as rs_accumulate_kernel).
//   k9_step               exchange of (sum of squares, loss sum) over peer stores + flags, then unscale / clip /
//                         AdamW on the owned partition with the new fp32 values stored into every rank's master
//                         buffer, then "my pushes are complete" to every peer (last CTA).
//   peer_wait             waits for those flags (before the 16-bit recast of the adapters).
// flags.p[r] = rank r's u32 flag array, entry [slot * MAX_RANKS + source rank], monotonic epochs.
__global__ void grad_ready_signal_kernel(Peers flags, int rank, int world, int slot, uint32_t epoch) {
  const int t = threadIdx.x;
  if (t < world) {
    __threadfence_system();
    st_release_sys_u32(reinterpret_cast<uint32_t*>(flags.p[t]) + slot * MAX_RANKS + rank, epoch);
  }
}

__global__ void __launch_bounds__(256) rs_wait_accumulate_kernel(Peers grads, const uint32_t* __restrict__ my_flags,
                                                                float* __restrict__ part, long long glo,
                                                                long long plo, long long n, int world, int slot,
                                                                uint32_t epoch, int* err) {
  if (threadIdx.x < world) wait_ge_sys(my_flags + slot * MAX_RANKS + threadIdx.x, epoch, err);
  __syncthreads();
  const long long n4 = n >> 2;
  float4* dst = reinterpret_cast<float4*>(part + plo);
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    float4 acc = dst[i];
    for (int r = 0; r < world; ++r) {
      const int4 v = ld_relaxed_sys_v4(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(grads.p[r]) + glo) + i);
      acc.x += __int_as_float(v.x); acc.y += __int_as_float(v.y);
      acc.z += __int_as_float(v.z); acc.w += __int_as_float(v.w);
    }
    dst[i] = acc;
  }
}

struct K9Step {
  Peers masters, boards, pads;
  float *m, *v, *g;
  const float* local2;      // [2]: this rank's sum of squares of its reduced partition, its loss sum
  float* out2;              // [2]: global sums (rank order)
  int* skipped_out;
  unsigned int* done;       // zero-initialised scratch word
  int* err;
  long long lo, n;
  int rank, world, ch_norm, ch_push;
  uint32_t epoch;
  float lr, b1, b2, eps, wd, ibc1, ibc2, inv_scale, max_norm;
};

__global__ void __launch_bounds__(256) k9_step_kernel(const K9Step k) {
  __shared__ float s_nsq;
  const int t = threadIdx.x;
  // ---- exchange: CTA 0 publishes, every CTA waits and sums the board in rank order ----
  if (t < k.world) {
    if (blockIdx.x == 0) {
      float* rb = reinterpret_cast<float*>(k.boards.p[t]) + 2 * k.rank;
      asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(rb), "f"(k.local2[0]) : "memory");
      asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(rb + 1), "f"(k.local2[1]) : "memory");
      __threadfence_system();
      st_release_sys_u32(reinterpret_cast<uint32_t*>(k.pads.p[t]) + k.ch_norm * MAX_RANKS + k.rank, k.epoch);
    }
    wait_ge_sys(reinterpret_cast<const uint32_t*>(k.pads.p[k.rank]) + k.ch_norm * MAX_RANKS + t, k.epoch, k.err);
  }
  __syncthreads();
  if (t == 0) {
    float s = 0.f, l = 0.f;
    const float* mb = reinterpret_cast<const float*>(k.boards.p[k.rank]);
    for (int r = 0; r < k.world; ++r) {
      float a, b;
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(a) : "l"(mb + 2 * r) : "memory");
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(b) : "l"(mb + 2 * r + 1) : "memory");
      s += a; l += b;
    }
    s_nsq = s;
    if (blockIdx.x == 0) { k.out2[0] = s; k.out2[1] = l; }
  }
  __syncthreads();
  const float nsq = s_nsq;
  const bool skip = !isfinite(nsq);
  if (blockIdx.x == 0 && t == 0 && k.skipped_out) *k.skipped_out = skip ? 1 : 0;
  // ---- unscale / clip / AdamW on [lo, lo + n), new values pushed to every rank ----
  float coef = k.inv_scale;
  if (k.max_norm > 0.f) coef *= fminf(1.f, k.max_norm / (sqrtf(nsq) * k.inv_scale + 1e-6f));
  float* mine = reinterpret_cast<float*>(k.masters.p[k.rank]) + k.lo;
  for (long long i = (long long)blockIdx.x * 256 + t; i < k.n; i += (long long)gridDim.x * 256) {
    if (!skip) {
      const float ge = __fmul_rn(k.g[i], coef);
      float mi = k.m[i], vi = k.v[i];
      const float pi = adamw_update(mine[i], mi, vi, ge, k.lr, k.b1, k.b2, k.eps, k.wd, k.ibc1, k.ibc2);
      k.m[i] = mi; k.v[i] = vi;
      for (int r = 0; r < k.world; ++r) (reinterpret_cast<float*>(k.masters.p[r]) + k.lo)[i] = pi;
    }
    k.g[i] = 0.f;
  }
  // ---- "my pushes are complete": the last CTA to finish tells every peer ----
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  if (t == 0) s_last = (atomicAdd(k.done, 1u) == gridDim.x - 1);
  __syncthreads();
  if (s_last) {
    if (t == 0) *k.done = 0u;
    if (t < k.world) {
      __threadfence_system();
      st_release_sys_u32(reinterpret_cast<uint32_t*>(k.pads.p[t]) + k.ch_push * MAX_RANKS + k.rank, k.epoch);
    }
  }
}

__global__ void peer_wait_kernel(const uint32_t* __restrict__ my_pads, int world, int channel, uint32_t epoch, int* err) {
  if (threadIdx.x < world) wait_ge_sys(my_pads + channel * MAX_RANKS + threadIdx.x, epoch, err);
}

// K15: embedding over a vocabulary-sharded table (rank r owns rows [r * rows_per_rank, (r + 1) * rows_per_rank)):
// every token's row is fetched from its owner rank with 16-byte peer loads (the table is immutable during a
// step: non-coherent loads). Replaces "all-gather the table, then look up": per step only the rows that are
// actually used cross NVLink, and no full-size staging copy of the table exists.
template <typename T>
__global__ void __launch_bounds__(256) embedding_peer_kernel(const long long* __restrict__ ids, Peers shards,
                                                            T* __restrict__ out, long long tokens, int d,
                                                            long long rows_per_rank, long long vocab) {
  __shared__ const void* s_sh[MAX_RANKS];
#pragma unroll
  for (int i = 0; i < MAX_RANKS; ++i)
    if (threadIdx.x == i) s_sh[i] = shards.p[i];
  __syncthreads();
  const int vpr = d >> 3;
  const long long total = tokens * vpr;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / vpr;
    const int v = (int)(idx - t * vpr);
    const long long id = ids[t];
    int4 val = make_int4(0, 0, 0, 0);
    if (id >= 0 && id < vocab) {
      const long long owner = id / rows_per_rank;
      val = ld_nc_v4(reinterpret_cast<const T*>(s_sh[owner]) + (id - owner * rows_per_rank) * d + v * 8);
    }
    *reinterpret_cast<int4*>(out + t * d + v * 8) = val;
  }
}

// K10 for prefill-sized steps, owner side. Every rank's row-parallel GEMM has stored the rows this rank owns into
// this rank's receive slots (slots[src * per + local_row], written by the GEMM epilogues over NVLink). One CTA per
// owned row: sum the slots in rank order (bit-identical on every run), add the residual (row-sharded on this path),
// normalise, and leave the normalised row in this rank's symmetric buffer xn_own[local_row], from where the next
// column-parallel GEMM of every rank pulls it (K11). Everything this kernel touches is local memory.
template <typename T, int VPT>
__global__ void __launch_bounds__(128) tp_own_reduce_norm_kernel(const T* __restrict__ slots, int world, long long per,
                                                                long long row_lo, T* __restrict__ residual,
                                                                const T* __restrict__ w, T* __restrict__ xn_own,
                                                                int d, float eps) {
  __shared__ float red[32];
  const long long lrow = blockIdx.x;
  const long long row = row_lo + lrow;
  const int nvec = d >> 3;
  float f[VPT][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      {
        const int4 rv = reinterpret_cast<const int4*>(residual + row * d)[v];
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&rv);
#pragma unroll
        for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(u[j]); f[i][2 * j] = t.x; f[i][2 * j + 1] = t.y; }
      }
      for (int r = 0; r < world; ++r) {
        const int4 pv = ld_relaxed_sys_v4(slots + (r * per + lrow) * d + v * 8);
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&pv);
#pragma unroll
        for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(u[j]); f[i][2 * j] += t.x; f[i][2 * j + 1] += t.y; }
      }
      int4 sv;
      uint32_t* su = reinterpret_cast<uint32_t*>(&sv);
#pragma unroll
      for (int j = 0; j < 4; ++j) su[j] = Cvt<T>::pack2(f[i][2 * j], f[i][2 * j + 1]);
      reinterpret_cast<int4*>(residual + row * d)[v] = sv;
#pragma unroll
      for (int j = 0; j < 4; ++j) { float2 t = Cvt<T>::unpack2(su[j]); f[i][2 * j] = t.x; f[i][2 * j + 1] = t.y; }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += f[i][j] * f[i][j];
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = (threadIdx.x & 31) < 4 ? red[threadIdx.x & 31] : 0.f;
  tot = warp_sum(tot);
  const float rstd = rsqrtf(tot / (float)d + eps);
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      const int4 wv = reinterpret_cast<const int4*>(w)[v];
      const uint32_t* wu = reinterpret_cast<const uint32_t*>(&wv);
      int4 ov;
      uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 wf = Cvt<T>::unpack2(wu[j]);
        const float a = wf.x * Cvt<T>::to_f(Cvt<T>::from_f(f[i][2 * j] * rstd));
        const float b = wf.y * Cvt<T>::to_f(Cvt<T>::from_f(f[i][2 * j + 1] * rstd));
        ou[j] = Cvt<T>::pack2(a, b);
      }
      reinterpret_cast<int4*>(xn_own + lrow * d)[v] = ov;
    }
  }
  __threadfence_system();
}

int fill(Peers& P, const void* const* ptrs, int world) {
  if (world < 1 || world > MAX_RANKS) return DLTI_ERR_BAD_ARG;
  for (int i = 0; i < MAX_RANKS; ++i) P.p[i] = i < world ? const_cast<void*>(ptrs[i]) : nullptr;
  return 0;
}

}  // namespace

DLTI_API int dlti_comm_max_ranks() { return MAX_RANKS; }

DLTI_API int dlti_peer_barrier(const void* const* signal_pads, int rank, int world, int channel,
                               unsigned int epoch, int* err_flag, cudaStream_t st) {
  Peers P;
  if (int rc = fill(P, signal_pads, world)) return rc;
  peer_barrier_kernel<<<1, 32, 0, st>>>(P, rank, world, channel, epoch, err_flag);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_peer_barrier_auto(const void* const* signal_pads, int rank, int world, int channel,
                                    unsigned int* device_counter, int* err_flag, cudaStream_t st) {
  Peers P;
  if (int rc = fill(P, signal_pads, world)) return rc;
  peer_barrier_auto_kernel<<<1, 32, 0, st>>>(P, rank, world, channel, device_counter, err_flag);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_allgather_rows(const void* const* shards, void* out, long long shard_bytes, int rank,
                                 int world, int blocks_per_peer, cudaStream_t st) {
  Peers P;
  if (int rc = fill(P, shards, world)) return rc;
  if (shard_bytes % 16) { dlti_set_error("allgather_rows: shard bytes must be a multiple of 16"); return DLTI_ERR_BAD_ARG; }
  if (blocks_per_peer <= 0) blocks_per_peer = 8;
  dim3 grid(blocks_per_peer, world);
  allgather_rows_kernel<<<grid, 256, 0, st>>>(P, (uint8_t*)out, shard_bytes, rank);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_rs_accumulate(const void* const* grads, float* part, long long lo, long long n,
                                int world, cudaStream_t st) {
  Peers P;
  if (int rc = fill(P, grads, world)) return rc;
  if (n % 4 || lo % 4) { dlti_set_error("rs_accumulate: partition must be 16-byte aligned"); return DLTI_ERR_BAD_ARG; }
  long long want = (n / 4 + 255) / 256;
  const long long cap = (long long)dlti_num_sms() * 4;
  rs_accumulate_kernel<<<(unsigned)(want > cap ? cap : (want < 1 ? 1 : want)), 256, 0, st>>>(P, part, lo, n, world);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_sumsq_exchange(const float* local, const void* const* boards,
                                 const void* const* signal_pads, int rank, int world, int channel,
                                 unsigned int epoch, float* out, int* err_flag, cudaStream_t st) {
  Peers B, S;
  if (int rc = fill(B, boards, world)) return rc;
  if (int rc = fill(S, signal_pads, world)) return rc;
  sumsq_exchange_kernel<<<1, 32, 0, st>>>(local, B, S, rank, world, channel, epoch, out, err_flag);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_adamw_push(const void* const* masters, float* m, float* v, float* g, long long lo,
                             long long n, int rank, int world, float lr, float b1, float b2, float eps,
                             float wd, int step, const float* gnorm_sq, float inv_scale, float max_norm,
                             int* skipped_out, cudaStream_t st) {
  Peers P;
  if (int rc = fill(P, masters, world)) return rc;
  if (n <= 0) return 0;
  if (step < 1 || !gnorm_sq) return DLTI_ERR_BAD_ARG;
  const float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
  long long want = (n + 255) / 256;
  const long long cap = (long long)dlti_num_sms() * 4;
  adamw_push_kernel<<<(unsigned)(want > cap ? cap : want), 256, 0, st>>>(
      P, m, v, g, lo, n, rank, world, lr, b1, b2, eps, wd, bc1, bc2, gnorm_sq, inv_scale, max_norm,
      skipped_out);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_embedding_peer(const long long* ids, const void* const* shards, void* out, long long tokens,
                                 int d, long long rows_per_rank, long long vocab, int world, int dtype,
                                 cudaStream_t st) {
  Peers P;
  if (int rc = fill(P, shards, world)) return rc;
  if (tokens <= 0) return 0;
  if (d % 8 || rows_per_rank <= 0) return DLTI_ERR_BAD_ARG;
  long long want = (tokens * (d / 8) + 255) / 256;
  const long long cap = (long long)dlti_num_sms() * 8;
  const unsigned grid = (unsigned)(want > cap ? cap : want);
  if (dtype == DLTI_BF16)
    embedding_peer_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(ids, P, (__nv_bfloat16*)out, tokens, d, rows_per_rank, vocab);
  else if (dtype == DLTI_FP16)
    embedding_peer_kernel<__half><<<grid, 256, 0, st>>>(ids, P, (__half*)out, tokens, d, rows_per_rank, vocab);
  else
    return DLTI_ERR_UNSUPPORTED;
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_grad_ready_signal(const void* const* flags, int rank, int world, int slot, unsigned int epoch,
                                    cudaStream_t st) {
  Peers F;
  if (int rc = fill(F, flags, world)) return rc;
  grad_ready_signal_kernel<<<1, 32, 0, st>>>(F, rank, world, slot, epoch);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

// part[plo .. plo + n) += sum_r grads[r][glo .. glo + n) once every rank has signalled (slot, epoch).
DLTI_API int dlti_rs_wait_accumulate(const void* const* grads, const unsigned int* my_flags, float* part,
                                     long long glo, long long plo, long long n, int world, int slot,
                                     unsigned int epoch, int ctas, int* err_flag, cudaStream_t st) {
  Peers P;
  if (int rc = fill(P, grads, world)) return rc;
  if (n <= 0) return 0;
  if (n % 4 || glo % 4 || plo % 4) { dlti_set_error("rs_wait_accumulate: ranges must be 16-byte aligned"); return DLTI_ERR_BAD_ARG; }
  long long want = (n / 4 + 255) / 256;
  const long long cap = ctas > 0 ? ctas : (long long)dlti_num_sms() * 4;
  rs_wait_accumulate_kernel<<<(unsigned)(want > cap ? cap : want), 256, 0, st>>>(P, my_flags, part, glo, plo, n, world,
                                                                                 slot, epoch, err_flag);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

// Norm / loss exchange + unscale / clip / AdamW on [lo, lo + n) + push to every rank's master + completion flags.
// local2 = {sum of squares of this rank's reduced partition, this rank's loss sum}; out2 = the global sums.
// boards[r]: rank r's float board [2 * MAX_RANKS]; pads: signal words, channels ch_norm / ch_push.
DLTI_API int dlti_k9_step(const void* const* masters, const void* const* boards, const void* const* signal_pads,
                          float* m, float* v, float* g, long long lo, long long n, int rank, int world, float lr,
                          float b1, float b2, float eps, float wd, int step, const float* local2, float* out2,
                          float inv_scale, float max_norm, int* skipped_out, unsigned int* done_scratch, int ch_norm,
                          int ch_push, unsigned int epoch, int* err_flag, cudaStream_t st) {
  K9Step k{};
  int rc;
  if ((rc = fill(k.masters, masters, world)) || (rc = fill(k.boards, boards, world)) || (rc = fill(k.pads, signal_pads, world)))
    return rc;
  if (step < 1 || !local2 || !out2 || !done_scratch || n < 0) return DLTI_ERR_BAD_ARG;
  k.m = m; k.v = v; k.g = g; k.local2 = local2; k.out2 = out2; k.skipped_out = skipped_out; k.done = done_scratch;
  k.err = err_flag; k.lo = lo; k.n = n; k.rank = rank; k.world = world; k.ch_norm = ch_norm; k.ch_push = ch_push;
  k.epoch = epoch; k.lr = lr; k.b1 = b1; k.b2 = b2; k.eps = eps; k.wd = wd;
  k.ibc1 = 1.f / (1.f - powf(b1, (float)step)); k.ibc2 = 1.f / (1.f - powf(b2, (float)step));
  k.inv_scale = inv_scale; k.max_norm = max_norm;
  long long want = (n + 255) / 256;
  const long long cap = (long long)dlti_num_sms() * 4;
  if (want < 1) want = 1;
  k9_step_kernel<<<(unsigned)(want > cap ? cap : want), 256, 0, st>>>(k);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_peer_wait(const unsigned int* my_pads, int world, int channel, unsigned int epoch, int* err_flag,
                            cudaStream_t st) {
  if (world < 1 || world > MAX_RANKS) return DLTI_ERR_BAD_ARG;
  peer_wait_kernel<<<1, 32, 0, st>>>(my_pads, world, channel, epoch, err_flag);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

// Owner side of the prefill-sized row-parallel all-reduce (see tp_own_reduce_norm_kernel): rows [row_lo, row_lo +
// rows) of the step, slots = this rank's receive buffer [world * per, d], xn_own = this rank's [per, d] output.
DLTI_API int dlti_tp_own_reduce_norm(const void* slots, int world, long long per, long long row_lo, long long rows,
                                     void* residual, const void* w, void* xn_own, int d, float eps, int dtype,
                                     cudaStream_t st) {
  if (rows <= 0) return 0;
  if (d % 8 || d > 8192 || world < 1 || rows > per) return DLTI_ERR_BAD_ARG;
  const int vpt = (d / 8 + 127) / 128;
#define TOR(T, V) tp_own_reduce_norm_kernel<T, V><<<(unsigned)rows, 128, 0, st>>>((const T*)slots, world, per, row_lo, (T*)residual, (const T*)w, (T*)xn_own, d, eps)
  if (dtype == DLTI_BF16) {
    if (vpt <= 1) TOR(__nv_bfloat16, 1); else if (vpt <= 2) TOR(__nv_bfloat16, 2); else if (vpt <= 4) TOR(__nv_bfloat16, 4); else TOR(__nv_bfloat16, 8);
  } else if (dtype == DLTI_FP16) {
    if (vpt <= 1) TOR(__half, 1); else if (vpt <= 2) TOR(__half, 2); else if (vpt <= 4) TOR(__half, 4); else TOR(__half, 8);
  } else {
    return DLTI_ERR_UNSUPPORTED;
  }
#undef TOR
  dlti_count_launch();
  return (int)cudaGetLastError();
}

static int tp_reduce_impl(const void* recv, long long slot_stride, int world, void* residual, const void* w, void* y,
                          long long rows, int d, float eps, int dtype, const TpBarrier* tb, cudaStream_t st) {
  if (rows <= 0) return 0;
  if (d % 8 || d > 8192 || world < 1) return DLTI_ERR_BAD_ARG;
  const int vpt = (d / 8 + 127) / 128;
  const TpBarrier none{};
#define TRN(T, V)                                                                                                    \
  do {                                                                                                               \
    if (tb)                                                                                                          \
      DLTI_CUDA_CHECK(dlti_launch_pdl(rows, tp_reduce_add_norm_kernel<T, V, true>, dim3((unsigned)rows), dim3(128), 0, \
                                      st, (const T*)recv, slot_stride, world, (T*)residual, (const T*)w, (T*)y, d,   \
                                      eps, *tb));                                                                    \
    else                                                                                                             \
      tp_reduce_add_norm_kernel<T, V, false><<<(unsigned)rows, 128, 0, st>>>((const T*)recv, slot_stride, world,     \
                                                                            (T*)residual, (const T*)w, (T*)y, d, eps, \
                                                                            none);                                   \
  } while (0)
  if (dtype == DLTI_BF16) {
    if (vpt <= 1) TRN(__nv_bfloat16, 1); else if (vpt <= 2) TRN(__nv_bfloat16, 2); else if (vpt <= 4) TRN(__nv_bfloat16, 4); else TRN(__nv_bfloat16, 8);
  } else if (dtype == DLTI_FP16) {
    if (vpt <= 1) TRN(__half, 1); else if (vpt <= 2) TRN(__half, 2); else if (vpt <= 4) TRN(__half, 4); else TRN(__half, 8);
  } else {
    return DLTI_ERR_UNSUPPORTED;
  }
#undef TRN
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_tp_reduce_add_norm(const void* recv, long long slot_stride, int world, void* residual,
                                     const void* w, void* y, long long rows, int d, float eps, int dtype,
                                     cudaStream_t st) {
  return tp_reduce_impl(recv, slot_stride, world, residual, w, y, rows, d, eps, dtype, nullptr, st);
}

// Same reduction with the rank barrier inside the kernel (see TpBarrier): replaces dlti_peer_barrier_auto +
// dlti_tp_reduce_add_norm. device_counter: the channel's epoch counter; done_scratch: a zero-initialised u32
// owned by this call site.
DLTI_API int dlti_tp_barrier_reduce_add_norm(const void* recv, long long slot_stride, int world, void* residual,
                                             const void* w, void* y, long long rows, int d, float eps, int dtype,
                                             const void* const* signal_pads, int rank, int channel,
                                             unsigned int* device_counter, unsigned int* done_scratch,
                                             int* err_flag, cudaStream_t st) {
  TpBarrier tb{};
  int rc = fill(tb.pads, signal_pads, world);
  if (rc) return rc;
  if (rank < 0 || rank >= world || !device_counter || !done_scratch) return DLTI_ERR_BAD_ARG;
  tb.counter = device_counter;
  tb.done = done_scratch;
  tb.err = err_flag;
  tb.rank = rank;
  tb.channel = channel;
  return tp_reduce_impl(recv, slot_stride, world, residual, w, y, rows, d, eps, dtype, &tb, st);
}

// Rows [row_lo, row_lo + rows) of the all-reduce + residual add + RMSNorm of a row-parallel GEMM (see
// tp_pull_reduce_norm_push_kernel). partial_peers[r] / xn_peers[r]: rank r's symmetric [T, d] partial /
// activation buffers. Call between two rank barriers (partials complete; activations landed).
DLTI_API int dlti_tp_pull_reduce_norm_push(const void* const* partial_peers, const void* const* xn_peers, int world,
                                           long long row_lo, long long rows, void* residual, const void* w, int d,
                                           float eps, int dtype, cudaStream_t st) {
  if (rows <= 0) return 0;
  if (d % 8 || d > 8192) return DLTI_ERR_BAD_ARG;
  Peers P, X;
  int rc = fill(P, partial_peers, world);
  if (rc) return rc;
  if ((rc = fill(X, xn_peers, world))) return rc;
  const int vpt = (d / 8 + 127) / 128;
#define TPR(T, V) tp_pull_reduce_norm_push_kernel<T, V><<<(unsigned)rows, 128, 0, st>>>(P, X, world, row_lo, (T*)residual, (const T*)w, d, eps)
  if (dtype == DLTI_BF16) {
    if (vpt <= 1) TPR(__nv_bfloat16, 1); else if (vpt <= 2) TPR(__nv_bfloat16, 2); else if (vpt <= 4) TPR(__nv_bfloat16, 4); else TPR(__nv_bfloat16, 8);
  } else if (dtype == DLTI_FP16) {
    if (vpt <= 1) TPR(__half, 1); else if (vpt <= 2) TPR(__half, 2); else if (vpt <= 4) TPR(__half, 4); else TPR(__half, 8);
  } else {
    return DLTI_ERR_UNSUPPORTED;
  }
#undef TPR
  dlti_count_launch();
  return (int)cudaGetLastError();
}
