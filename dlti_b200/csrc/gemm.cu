// K2/K3: persistent warp-specialised tcgen05 GEMM for sm_100a.
//
//   C[M,N] = alpha * ( op(A)[M,K] · op(B)[K,N]  +  A2[M,K2] · op(B2)[K2,N] ) (+ R) (+ C)
//
// * Operands are 16-bit (bf16 or fp16), staged global->shared by TMA with the 128-byte
//   swizzle, multiplied by tcgen05.mma (kind::f16) issued from ONE thread, accumulated in
//   fp32 in tensor memory (TMEM, two accumulator stages so the epilogue of tile i overlaps
//   the MMAs of tile i+1), read back with tcgen05.ld by four epilogue warps.
// * "K-extension": a second operand pair (A2,B2) is streamed after the main K loop into the
//   SAME TMEM tile. This is how LoRA is fused: with A2 = (dropout(x)·A^T)
//   padded to 64 columns and B2 = (alpha/r)·B padded to 64 columns the kernel produces
//   W·x + (alpha/r)·B·(A·x) with one extra 64-deep K block instead of 4 extra kernels
//   (the reference path is PEFT lora.Linear: base GEMM + dropout + 2 GEMMs + scale + add,
//   reference call site training/train_baseline.py:131-140).
// * Operand majorness is a template switch, so forward (x·W^T: B is K-major), input-gradient
//   (dy·W: B is MN-major, read from the SAME weight buffer, no transposed copy) and
//   weight-gradient (dy^T·x: both MN-major, split-K with fp32 atomics) all run here.
// * Roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2..5 = epilogue.
//
// Replaces cuBLAS/cuBLASLt GEMM calls made by torch.nn.functional.linear in the reference
// stack.
#include "common.cuh"

#include <string>
#include <cstring>
#include <cstdio>

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 192;

template <int BN>
struct Cfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int BAR_BYTES = 256;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 1024;
};

// K1: all-gather fused into the GEMM. The B operand (a ZeRO-3 row-sharded weight) is consumed from
// a local staging buffer that the SAME kernel fills: two extra "gather" warps per CTA pull 64-row
// units (in 4-row chunks) from their owner ranks over NVLink (peer loads from the symmetric arena)
// and bump a per-unit counter with a release; the TMA producer acquires the counters of the units
// its next B tile needs before issuing the load. Tiles (NT) / K blocks (NN) are visited starting at
// this rank's own rows, so the math on early tiles overlaps the transfer of later ones.
constexpr int GATHER_CH = 16;              // chunks per 64-row unit (4 rows each)
constexpr int GATHER_THREADS = 128;
struct GatherParams {
  const void* peer[16];     // each rank's shard base (peer-mapped); peer[rank] is the local shard
  uint8_t* staging;         // local full matrix [world * rows_per_rank, row_bytes]
  uint32_t* counters;       // [n_units] completed-chunk counters (monotonic across epochs)
  int* err;                 // set to 1 on spin-wait timeout
  long long row_bytes;
  uint32_t target;          // unit complete  <=>  counters[u] >= target  (= epoch * GATHER_CH)
  int rank, world;
  int units_per_rank;       // rows_per_rank / 64
  int n_units;
  int local_direct;         // this rank's own rows are read straight from its shard (no staging copy)
};

struct GemmParams {
  int M, N;
  int nmb, nnb;           // tile counts
  int nkb_main, nkb_ext;  // 64-deep K blocks in the main and extension operands
  int split_k, kb_per_split;
  void* C;
  const void* R;
  long long ldc, ldr;
  int out_dtype;   // DltiDtype
  int in_dtype;    // DltiDtype (16-bit)
  int accumulate;  // fp32 output only: C += result
  int atomic_out;  // fp32 output only: red.add (split-K)
  int transpose_out;  // atomic fp32 output only: element (row, col) goes to C[col * ldc + row]
  int vec_ok;      // 16-byte vector stores allowed
  float alpha;
  uint32_t idesc;
  int n_rot, k_rot;   // GATHER: first N block / K block owned by this rank
  // K10: row-parallel GEMM -> collective in the epilogue. When push_world > 0 the 16-bit output tile
  // is stored into EVERY rank's receive buffer (peer stores over NVLink) at slot `push_rank`:
  //   dst_r = push_peer[r] + push_rank * push_slot_stride + row * ldc + col
  void* push_peer[16];
  long long push_slot_stride;
  int push_world, push_rank;
  int early_a;     // SK: request the first A (weight) tiles before griddep_wait() (weights must be static)
};

__device__ __forceinline__ uint32_t ld_relaxed_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

__device__ __forceinline__ void gather_wait_unit(const GatherParams& g, int u) {
  if (u >= g.n_units) return;
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_gpu_u32(g.counters + u) - g.target) < 0) {
    if (clock64() - t0 > 20000000000LL) {
      if (g.err) atomicExch(g.err, 1);
      return;
    }
    __nanosleep(32);
  }
}

// Work decomposition shared by the producer, MMA and epilogue roles: a sequence of segments
// (rem = output tile index, K blocks [kb0, kb1)).
//   SK = false: tile-major; one segment per (split, tile), CTAs stride through them.
//   SK = true ("stream-K", atomic fp32 output only): the (tile, K block) pairs are linearised and cut
//   into gridDim.x equal contiguous ranges, so every CTA streams the same number of K blocks whatever
//   the tile count; a range that crosses a tile boundary yields one segment per tile and each segment's
//   partial sum is flushed with red.add.
template <bool SK>
struct SegIter {
  int rem, kb0, kb1;
  int tile;
  int pos, end;
  __device__ __forceinline__ SegIter(const GemmParams& p, int tiles_mn, int nkb_total) {
    tile = blockIdx.x;
    if (SK) {
      const long long total = (long long)tiles_mn * nkb_total;
      pos = (int)(total * blockIdx.x / gridDim.x);
      end = (int)(total * (blockIdx.x + 1) / gridDim.x);
    }
  }
  __device__ __forceinline__ bool next(const GemmParams& p, int tiles_mn, int nkb_total) {
    if (SK) {
      if (pos >= end) return false;
      rem = pos / nkb_total;
      kb0 = pos - rem * nkb_total;
      kb1 = min(nkb_total, kb0 + (end - pos));
      pos += kb1 - kb0;
      return true;
    }
    if (tile >= tiles_mn * p.split_k) return false;
    const int split = tile / tiles_mn;
    rem = tile - split * tiles_mn;
    kb0 = split * p.kb_per_split;
    kb1 = min(nkb_total, kb0 + p.kb_per_split);
    tile += gridDim.x;
    return true;
  }
};

template <int BN, bool A_MN, bool B_MN, bool GATHER, bool PUSH, bool SK = false>
__global__ void __launch_bounds__(NUM_THREADS + (GATHER ? GATHER_THREADS : 0), 1)
gemm_tcgen05_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmA,
                    const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmA2,
                    const __grid_constant__ CUtensorMap tmB2,
                    const __grid_constant__ CUtensorMap tmBL, const GatherParams gp) {
  using C_ = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + C_::STAGES * C_::STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C_::STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C_::STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C_::STAGES + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * C_::STAGES + 4);
  // generic pointer to the TMEM base slot
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  __shared__ void* s_peers[16];   // GATHER: shard bases; PUSH: receive-buffer bases
  if (GATHER || PUSH) {
#pragma unroll
    for (int i = 0; i < 16; ++i)   // static indices: keeps the parameters in the constant bank
      if (threadIdx.x == i) s_peers[i] = GATHER ? const_cast<void*>(gp.peer[i]) : p.push_peer[i];
  }
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // uniform for the compiler
  const int lane = threadIdx.x & 31;
  const int tiles_mn = p.nmb * p.nnb;
  const int nkb_total = p.nkb_main + p.nkb_ext;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.nkb_ext > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
    for (int s = 0; s < C_::STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 4);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C_::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot_ptr, 0);
  if constexpr (SK || PUSH) griddep_launch();   // the consumer of C may be scheduled; it waits for this grid itself

  if (warp == 0) {
    // ===================== TMA producer =====================
    bool all_ready = true;      // GATHER: every unit this kernel could need has already landed
    if (GATHER) {
      for (int u = lane; u < gp.n_units; u += 32) {
        const bool mine = gp.local_direct && (u / gp.units_per_rank) == gp.rank;
        if (!mine && (int32_t)(ld_relaxed_gpu_u32(gp.counters + u) - gp.target) < 0) all_ready = false;
      }
      all_ready = __all_sync(0xffffffffu, all_ready);
      fence_acq_rel_gpu();
      asm volatile("fence.proxy.async.global;" ::: "memory");
    }
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int pre = 0;              // SK: pipeline stages whose A (weight) tile was requested before griddep_wait()
      if constexpr (SK) {
        // Decode path, launched as a programmatic dependent of the kernel that produces B (the activations):
        // the weights do not depend on it, so the first STAGES weight tiles are already in flight while the
        // predecessor drains; only the activation tiles wait for it.
        for (SegIter<SK> it(p, tiles_mn, nkb_total); p.early_a && pre < C_::STAGES && it.next(p, tiles_mn, nkb_total);) {
          const int m0 = (it.rem % p.nmb) * BM;
          for (int kb = it.kb0; kb < it.kb1 && pre < C_::STAGES; ++kb, ++pre) {
            mbar_expect_tx(full_bar(pre), C_::STAGE_BYTES);
            tma_load_2d(smem_base + pre * C_::STAGE_BYTES, &tmA, full_bar(pre), kb * BK, m0);
          }
        }
        griddep_wait();
      }
      if constexpr (PUSH) griddep_wait();   // decode path: A (activations) comes from the kernel before
      int issued = 0;
      for (SegIter<SK> it(p, tiles_mn, nkb_total); it.next(p, tiles_mn, nkb_total);) {
        const int rem = it.rem;
        int n_blk = rem / p.nmb;
        const int m_blk = rem - n_blk * p.nmb;
        if (GATHER && !B_MN) n_blk = (n_blk + p.n_rot) % p.nnb;      // start at this rank's own rows
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        const int kb0 = it.kb0;
        const int kb1 = it.kb1;
        bool b_local = false;       // B tile comes straight from this rank's shard
        int b_row_off = 0;
        const int lo_unit = GATHER ? gp.rank * gp.units_per_rank : 0;
        const int hi_unit = GATHER ? lo_unit + gp.units_per_rank : 0;
        if (GATHER) {
          const int lo_row = gp.rank * gp.units_per_rank * 64;
          b_row_off = lo_row;
          if (!B_MN) {
            b_local = gp.local_direct && n0 >= lo_row && n0 + BN <= lo_row + gp.units_per_rank * 64;
            if (!b_local && !all_ready) {
              uint32_t cv[BN / 64];
#pragma unroll
              for (int j = 0; j < BN / 64; ++j)      // independent relaxed loads, in flight together
                cv[j] = (n0 / 64 + j < gp.n_units) ? ld_relaxed_gpu_u32(gp.counters + n0 / 64 + j) : gp.target;
#pragma unroll
              for (int j = 0; j < BN / 64; ++j)
                if ((int32_t)(cv[j] - gp.target) < 0) gather_wait_unit(gp, n0 / 64 + j);
              fence_acq_rel_gpu();
              asm volatile("fence.proxy.async.global;" ::: "memory");
            }
          }
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * C_::STAGE_BYTES;
          const uint32_t sb = sa + C_::A_BYTES;
          const uint32_t fb = full_bar(stage);
          if constexpr (SK) {
            if (issued < pre) {               // weight tile and expect_tx already issued: add the activations
              ++issued;
              tma_load_2d(sb, &tmB, fb, kb * BK, n0);
              if (++stage == C_::STAGES) { stage = 0; phase ^= 1u; }
              continue;
            }
          }
          mbar_expect_tx(fb, C_::STAGE_BYTES);
          const bool ext = kb >= p.nkb_main;
          const CUtensorMap* ma = ext ? &tmA2 : &tmA;
          const CUtensorMap* mb = ext ? &tmB2 : &tmB;
          int kb_eff = ext ? kb - p.nkb_main : kb;
          if (GATHER && B_MN && !ext) {
            // contraction starts at this rank's own rows; no div/mod here: the producer thread's
            // per-stage budget is only ~500 cycles
            kb_eff = kb + p.k_rot;
            if (kb_eff >= p.nkb_main) kb_eff -= p.nkb_main;
            b_local = gp.local_direct && kb_eff >= lo_unit && kb_eff < hi_unit;
            if (!b_local && !all_ready) {
              gather_wait_unit(gp, kb_eff);
              asm volatile("fence.proxy.async.global;" ::: "memory");
            }
          }
          const int k0 = kb_eff * BK;
          if constexpr (!A_MN) {
            tma_load_2d(sa, ma, fb, k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), ma, fb, m0 + 64 * j, k0);
          }
          if (GATHER && b_local && !ext) {
            if constexpr (!B_MN) {
              tma_load_2d(sb, &tmBL, fb, k0, n0 - b_row_off);
            } else {
#pragma unroll
              for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tmBL, fb, n0 + 64 * j, k0 - b_row_off);
            }
          } else if constexpr (!B_MN) {
            tma_load_2d(sb, mb, fb, k0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), mb, fb, n0 + 64 * j, k0);
          }
          if (++stage == C_::STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp runs the loop (uniform control flow and operands); one elected lane issues.
    {
      const bool el = elect_one_sync();
      const DescLoHi dA = A_MN ? desc_split(smem_base, BK * 128, 1024) : desc_split(smem_base, 0, 1024);
      const DescLoHi dB = B_MN ? desc_split(smem_base + C_::A_BYTES, BK * 128, 1024)
                               : desc_split(smem_base + C_::A_BYTES, 0, 1024);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (SegIter<SK> it(p, tiles_mn, nkb_total); it.next(p, tiles_mn, nkb_total);) {
        const int kb0 = it.kb0;
        const int kb1 = it.kb1;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          DescLoHi a = dA, b = dB;
          a.lo += stage * (C_::STAGE_BYTES >> 4);
          b.lo += stage * (C_::STAGE_BYTES >> 4);
          if (el) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_f16_ss(d_tmem, a.at(A_MN ? k * (UMMA_K * 128) : k * (UMMA_K * 2)),
                          b.at(B_MN ? k * (UMMA_K * 128) : k * (UMMA_K * 2)), p.idesc,
                          (kb > kb0 || k > 0) ? 1u : 0u);
            umma_commit(empty_bar(stage));  // frees the smem slot once these MMAs retire
            if (kb == kb1 - 1) umma_commit(tfull_bar(acc));
          }
          if (++stage == C_::STAGES) { stage = 0; phase ^= 1u; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else if (GATHER && warp >= 6) {
    // ===================== gather warps: pull weight units from their owners ==================
    const int gt = threadIdx.x - NUM_THREADS;                  // 0..127
    const long long chunk_bytes = 4 * gp.row_bytes;            // 4 rows
    const long long unit_bytes = 64 * gp.row_bytes;
    const int skip = gp.local_direct ? gp.units_per_rank : 0;  // own rows are read in place
    const int total_chunks = (gp.n_units - skip) * GATHER_CH;
    for (int c = blockIdx.x; c < total_chunks; c += gridDim.x) {
      const int j = c / GATHER_CH + skip, part = c % GATHER_CH;
      const int u = (gp.rank * gp.units_per_rank + j) % gp.n_units;   // needed-first order
      if ((int32_t)(ld_acquire_gpu_u32(gp.counters + u) - gp.target) >= 0) continue;  // landed earlier this epoch
      const int owner = u / gp.units_per_rank;
      const uint8_t* src = reinterpret_cast<const uint8_t*>(s_peers[owner]) +
                           (long long)(u - owner * gp.units_per_rank) * unit_bytes + part * chunk_bytes;
      uint8_t* dst = gp.staging + (long long)u * unit_bytes + part * chunk_bytes;
      const long long nvec = chunk_bytes >> 4;
      long long i = gt;
      for (; i + 7 * GATHER_THREADS < nvec; i += 8 * GATHER_THREADS) {
        int4 v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = ld_nc_v4(src + ((i + q * GATHER_THREADS) << 4));
#pragma unroll
        for (int q = 0; q < 8; ++q) st_v4(dst + ((i + q * GATHER_THREADS) << 4), v[q]);
      }
      for (; i < nvec; i += GATHER_THREADS) st_v4(dst + (i << 4), ld_nc_v4(src + (i << 4)));
      __threadfence();
      asm volatile("bar.sync 2, 128;" ::: "memory");
      if (gt == 0) red_release_gpu_add_u32(gp.counters + u, 1u);
    }
  } else {
    // ===================== epilogue (4 warps; warp w owns TMEM lanes 32*(w%4)..) =============
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    if constexpr (SK || PUSH) griddep_wait();   // SK: C is re-zeroed by the predecessor; PUSH: peers' slots
    for (SegIter<SK> it(p, tiles_mn, nkb_total); it.next(p, tiles_mn, nkb_total);) {
      const int rem = it.rem;
      int n_blk = rem / p.nmb;
      const int m_blk = rem - n_blk * p.nmb;
      if (GATHER && !B_MN) n_blk = (n_blk + p.n_rot) % p.nnb;
      const int row = m_blk * BM + quarter * 32 + lane;
      const int n0 = n_blk * BN;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t v[32];
        tmem_ld_32x32b_x32(t_row + c * 32, v);
        tmem_wait_ld();
        if (row < p.M) {
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.alpha;
          const bool full = (col0 + 32 <= p.N) && p.vec_ok;
          if (p.out_dtype == DLTI_FP32) {
            float* crow = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + col0;
            if (p.atomic_out) {
              if (p.transpose_out) {      // swap-AB decode GEMM: lanes (= rows) hit consecutive addresses
                float* cb = reinterpret_cast<float*>(p.C) + row;
                for (int i = 0; i < 32; ++i)
                  if (col0 + i < p.N) atomicAdd(cb + (long long)(col0 + i) * p.ldc, f[i]);
              } else {
                for (int i = 0; i < 32; ++i)
                  if (col0 + i < p.N) atomicAdd(crow + i, f[i]);
              }
            } else if (full) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                float4 o = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                float4* dst = reinterpret_cast<float4*>(crow) + i;
                if (p.accumulate) {
                  float4 old = *dst;
                  o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                }
                *dst = o;
              }
            } else {
              for (int i = 0; i < 32; ++i)
                if (col0 + i < p.N) crow[i] = p.accumulate ? crow[i] + f[i] : f[i];
            }
          } else {
            uint16_t* crow = reinterpret_cast<uint16_t*>(p.C) + (long long)row * p.ldc + col0;
            const uint16_t* rrow =
                p.R ? reinterpret_cast<const uint16_t*>(p.R) + (long long)row * p.ldr + col0 : nullptr;
            const bool bf = p.out_dtype == DLTI_BF16;
            if (full) {
              if (rrow) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  int4 rv = *(reinterpret_cast<const int4*>(rrow) + i);
                  const uint32_t* ru = reinterpret_cast<const uint32_t*>(&rv);
#pragma unroll
                  for (int j = 0; j < 4; ++j) {
                    float2 r2 = bf ? Cvt<__nv_bfloat16>::unpack2(ru[j]) : Cvt<__half>::unpack2(ru[j]);
                    f[8 * i + 2 * j] += r2.x;
                    f[8 * i + 2 * j + 1] += r2.y;
                  }
                }
              }
              int4 ov[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                uint32_t* ou = reinterpret_cast<uint32_t*>(&ov[i]);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  ou[j] = bf ? Cvt<__nv_bfloat16>::pack2(f[8 * i + 2 * j], f[8 * i + 2 * j + 1])
                             : Cvt<__half>::pack2(f[8 * i + 2 * j], f[8 * i + 2 * j + 1]);
              }
              if (PUSH) {
                const long long off = (long long)p.push_rank * p.push_slot_stride + (long long)row * p.ldc + col0;
                for (int r = 0; r < p.push_world; ++r) {
                  int4* dst = reinterpret_cast<int4*>(reinterpret_cast<uint16_t*>(s_peers[r]) + off);
#pragma unroll
                  for (int i = 0; i < 4; ++i) dst[i] = ov[i];
                }
              } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) *(reinterpret_cast<int4*>(crow) + i) = ov[i];
              }
            } else {
              for (int i = 0; i < 32; ++i) {
                if (col0 + i < p.N) {
                  float x = f[i];
                  if (rrow) {
                    x += bf ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(rrow + i))
                            : __half2float(*reinterpret_cast<const __half*>(rrow + i));
                  }
                  uint16_t bits;
                  if (bf) { __nv_bfloat16 t = __float2bfloat16_rn(x); bits = *reinterpret_cast<uint16_t*>(&t); }
                  else    { __half t = __float2half_rn(x); bits = *reinterpret_cast<uint16_t*>(&t); }
                  if (PUSH) {
                    const long long off = (long long)p.push_rank * p.push_slot_stride + (long long)row * p.ldc + col0 + i;
                    for (int r = 0; r < p.push_world; ++r) reinterpret_cast<uint16_t*>(s_peers[r])[off] = bits;
                  } else {
                    crow[i] = bits;
                  }
                }
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(acc));
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (PUSH) __threadfence_system();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C_::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------
inline int get_tmap(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t bi,
                    uint32_t bo) {
  return dlti_get_tmap(out, base, inner, outer, ld, bi, bo);
}

template <int BN, bool A_MN, bool B_MN, bool GATHER, bool PUSH = false, bool SK = false>
int launch(const GemmParams& p, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& a2,
           const CUtensorMap& b2, const CUtensorMap& bl, const GatherParams& gp, int grid,
           cudaStream_t stream) {
  using C_ = Cfg<BN>;
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    DLTI_CUDA_CHECK(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, A_MN, B_MN, GATHER, PUSH, SK>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         C_::SMEM_BYTES));
    configured[dev & 63] = true;
  }
  if constexpr (SK || PUSH) {
    DLTI_CUDA_CHECK(dlti_launch_pdl(SK ? p.N : p.M, gemm_tcgen05_kernel<BN, A_MN, B_MN, GATHER, PUSH, SK>, dim3(grid),
                                    dim3(NUM_THREADS), C_::SMEM_BYTES, stream, p, a, b, a2, b2, bl, gp));
    dlti_count_launch();
    return 0;
  }
  gemm_tcgen05_kernel<BN, A_MN, B_MN, GATHER, PUSH, SK>
      <<<grid, NUM_THREADS + (GATHER ? GATHER_THREADS : 0), C_::SMEM_BYTES, stream>>>(p, a, b, a2, b2, bl, gp);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

}  // namespace

static int gemm_impl(const void* A, const void* B, void* C, const void* R, int M, int N, int K,
                     long long lda, long long ldb, long long ldc, long long ldr, int a_mn,
                     int b_mn, const void* A2, const void* B2, int K2, long long lda2,
                     long long ldb2, int in_dtype, int out_dtype, float alpha, int accumulate,
                     int split_k, int block_n, int max_ctas, cudaStream_t stream,
                     const GatherParams* gather, const void* const* push_peers = nullptr,
                     int push_rank = 0, int push_world = 0, long long push_slot_stride = 0,
                     int transpose_out = 0, int stream_k = 0) {
  if (M <= 0 || N <= 0 || K < 0 || (K == 0 && K2 == 0)) return DLTI_ERR_BAD_ARG;
  if (in_dtype != DLTI_BF16 && in_dtype != DLTI_FP16) return DLTI_ERR_UNSUPPORTED;
  if (K2 % BK) { dlti_set_error("K2 must be a multiple of 64"); return DLTI_ERR_BAD_ARG; }
  if (out_dtype != DLTI_FP32 && accumulate) { dlti_set_error("accumulate needs fp32 output"); return DLTI_ERR_BAD_ARG; }
  if (a_mn && !b_mn) return DLTI_ERR_UNSUPPORTED;
  int bn = block_n;
  if (bn == 0) {
    bn = (N <= 64) ? 64 : (N <= 128 ? 128 : 256);
    // decode-sized M (one M tile): the GEMM is a weight stream, so spread N over as many SMs as possible
    if (M <= BM && !gather) {
      const int sms = dlti_num_sms();
      if ((N + 255) / 256 < sms) bn = 128;
      if ((N + 127) / 128 < sms) bn = 64;
    } else if (!gather && bn == 256) {
      // prefill-sized problems (fewer than two waves of 128x256 tiles): narrower tiles fill the SMs better.
      // cost = waves x tile width / relative efficiency of that width (narrow tiles are smem-bandwidth bound)
      const int sms = dlti_num_sms();
      const long long nmb = (M + BM - 1) / BM;
      if (nmb * ((N + 255) / 256) < 2 * sms) {
        const int widths[3] = {256, 128, 64};
        const double eff[3] = {1.0, 0.92, 0.66};
        double best = 1e30;
        for (int i = 0; i < 3; ++i) {
          if (a_mn && widths[i] == 256) continue;
          const long long tiles = nmb * ((N + widths[i] - 1) / widths[i]);
          const double cost = (double)((tiles + sms - 1) / sms) * widths[i] / eff[i];
          if (cost < best - 1e-9) { best = cost; bn = widths[i]; }
        }
      }
    }
  }
  if (a_mn && bn == 256) bn = 128;
  if (bn != 64 && bn != 128 && bn != 256) return DLTI_ERR_BAD_ARG;

  GemmParams p{};
  p.M = M; p.N = N;
  p.nmb = (M + BM - 1) / BM;
  p.nnb = (N + bn - 1) / bn;
  p.nkb_main = (K + BK - 1) / BK;
  p.nkb_ext = (A2 && B2) ? K2 / BK : 0;
  const int nkb_total = p.nkb_main + p.nkb_ext;
  if (split_k < 1) split_k = 1;
  if (split_k > nkb_total) split_k = nkb_total;
  p.kb_per_split = (nkb_total + split_k - 1) / split_k;
  p.split_k = (nkb_total + p.kb_per_split - 1) / p.kb_per_split;
  if (p.split_k > 1 && !(out_dtype == DLTI_FP32 && accumulate)) {
    dlti_set_error("split_k > 1 requires fp32 output with accumulate=1 (atomic adds into C)");
    return DLTI_ERR_BAD_ARG;
  }
  p.C = C; p.R = R; p.ldc = ldc; p.ldr = ldr;
  p.out_dtype = out_dtype; p.in_dtype = in_dtype;
  p.accumulate = accumulate;
  p.atomic_out = p.split_k > 1 || transpose_out;
  p.transpose_out = transpose_out;
  const int ob = out_dtype == DLTI_FP32 ? 4 : 2;
  p.vec_ok = ((reinterpret_cast<uintptr_t>(C) & 15) == 0) && ((ldc * ob) % 16 == 0) &&
             (!R || (((reinterpret_cast<uintptr_t>(R) & 15) == 0) && ((ldr * 2) % 16 == 0)));
  p.alpha = alpha;
  p.idesc = umma_idesc_f16(in_dtype, BM, bn, a_mn != 0, b_mn != 0);
  if (push_world > 0) {
    if (out_dtype == DLTI_FP32 || R || push_world > 16 || !p.vec_ok && false) { dlti_set_error("push GEMM: 16-bit output without residual"); return DLTI_ERR_UNSUPPORTED; }
    for (int i = 0; i < push_world; ++i) p.push_peer[i] = const_cast<void*>(push_peers[i]);
    p.push_world = push_world; p.push_rank = push_rank; p.push_slot_stride = push_slot_stride;
    p.vec_ok = p.vec_ok && ((push_slot_stride * 2) % 16 == 0);
  }
  GatherParams gp{};
  if (gather) {
    if (a_mn || p.split_k != 1) { dlti_set_error("gather GEMM: A must be K-major and split_k = 1"); return DLTI_ERR_UNSUPPORTED; }
    gp = *gather;
    // first N tile lying entirely inside this rank's own rows (ceil): a tile that straddles into the
    // previous rank's rows would need the units that are gathered LAST and stall every CTA at once
    p.n_rot = (int)((((long long)gp.rank * gp.units_per_rank * 64) + bn - 1) / bn) % p.nnb;
    p.k_rot = (gp.rank * gp.units_per_rank) % (p.nkb_main > 0 ? p.nkb_main : 1);
  }

  CUtensorMap ta, tb, ta2, tb2;
  int rc;
  if (p.nkb_main > 0) {
    rc = a_mn ? get_tmap(&ta, A, M, K, lda, 64, 64) : get_tmap(&ta, A, K, M, lda, 64, BM);
    if (rc) return rc;
    rc = b_mn ? get_tmap(&tb, B, N, K, ldb, 64, 64) : get_tmap(&tb, B, K, N, ldb, 64, bn);
    if (rc) return rc;
  }
  if (p.nkb_ext > 0) {
    rc = a_mn ? get_tmap(&ta2, A2, M, K2, lda2, 64, 64) : get_tmap(&ta2, A2, K2, M, lda2, 64, BM);
    if (rc) return rc;
    rc = b_mn ? get_tmap(&tb2, B2, N, K2, ldb2, 64, 64) : get_tmap(&tb2, B2, K2, N, ldb2, 64, bn);
    if (rc) return rc;
    if (p.nkb_main == 0) { ta = ta2; tb = tb2; }
  } else {
    ta2 = ta; tb2 = tb;
  }

  CUtensorMap tbl = tb;
  if (gather) {
    const long long rpr = (long long)gp.units_per_rank * 64;
    gp.local_direct = (b_mn || (rpr % bn) == 0) ? 1 : 0;
    if (gp.local_direct) {
      const void* lsh = gp.peer[gp.rank];
      rc = b_mn ? get_tmap(&tbl, lsh, N, rpr, ldb, 64, 64) : get_tmap(&tbl, lsh, K, rpr, ldb, 64, bn);
      if (rc) return rc;
    }
  }
  const int total = p.nmb * p.nnb * p.split_k;
  int grid = dlti_num_sms();
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (total < grid) grid = total;

  if (stream_k) {
    // equal K-block ranges per CTA over the linearised (tile, K block) space; needs the atomic epilogue
    if (a_mn || b_mn || gather || push_world > 0 || !p.atomic_out || p.split_k != 1) return DLTI_ERR_UNSUPPORTED;
    const long long items = (long long)p.nmb * p.nnb * nkb_total;
    p.early_a = dlti_pdl_enabled() >= 2 && p.N <= DLTI_PDL_MAX_ROWS;   // p.N = token rows (swap-AB)
    grid = dlti_num_sms();
    if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
    if (items < grid) grid = (int)items;
    if (bn == 256) return launch<256, false, false, false, false, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
    if (bn == 128) return launch<128, false, false, false, false, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
    return launch<64, false, false, false, false, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
  }
  if (push_world > 0) {
    if (a_mn || b_mn) return DLTI_ERR_UNSUPPORTED;
    if (bn == 256) return launch<256, false, false, false, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
    if (bn == 128) return launch<128, false, false, false, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
    return launch<64, false, false, false, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
  }
  if (gather) {
    if (bn == 256 && !b_mn) return launch<256, false, false, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
    if (bn == 128 && !b_mn) return launch<128, false, false, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
    if (bn == 256 && b_mn) return launch<256, false, true, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
    if (bn == 128 && b_mn) return launch<128, false, true, true>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
    return DLTI_ERR_UNSUPPORTED;
  }
#define DLTI_GEMM_CASE(BN_, AM_, BM_) \
  if (bn == BN_ && (a_mn != 0) == AM_ && (b_mn != 0) == BM_) \
    return launch<BN_, AM_, BM_, false>(p, ta, tb, ta2, tb2, tbl, gp, grid, stream);
  DLTI_GEMM_CASE(256, false, false)
  DLTI_GEMM_CASE(128, false, false)
  DLTI_GEMM_CASE(64, false, false)
  DLTI_GEMM_CASE(256, false, true)
  DLTI_GEMM_CASE(128, false, true)
  DLTI_GEMM_CASE(64, false, true)
  DLTI_GEMM_CASE(128, true, true)
  DLTI_GEMM_CASE(64, true, true)
#undef DLTI_GEMM_CASE
  return DLTI_ERR_UNSUPPORTED;
}

// See include/dlti_b200.h for the contract.
DLTI_API int dlti_gemm(const void* A, const void* B, void* C, const void* R, int M, int N, int K,
                       long long lda, long long ldb, long long ldc, long long ldr, int a_mn,
                       int b_mn, const void* A2, const void* B2, int K2, long long lda2,
                       long long ldb2, int in_dtype, int out_dtype, float alpha, int accumulate,
                       int split_k, int block_n, int max_ctas, cudaStream_t stream) {
  return gemm_impl(A, B, C, R, M, N, K, lda, ldb, ldc, ldr, a_mn, b_mn, A2, B2, K2, lda2, ldb2, in_dtype,
                   out_dtype, alpha, accumulate, split_k, block_n, max_ctas, stream, nullptr);
}

// K1: same GEMM, with the B operand (a row-sharded weight [world*rows_per_rank, ldb]) gathered from
// its owner ranks by the kernel itself while it computes. B must be the local staging buffer; peer_
// shards[r] is rank r's shard base in the symmetric arena; counters = u32[world*rows_per_rank/64]
// (zero-initialised once, then monotonic); epoch must increase each time the staging is to be
// refilled (same epoch => units already landed are reused: forward recompute -> backward).
DLTI_API int dlti_gemm_gather(const void* A, void* B_staging, void* C, const void* R, int M, int N, int K,
                              long long lda, long long ldb, long long ldc, long long ldr, int b_mn,
                              const void* A2, const void* B2, int K2, long long lda2, long long ldb2,
                              int in_dtype, int out_dtype, float alpha, int block_n, int max_ctas,
                              const void* const* peer_shards, unsigned int* counters, unsigned int epoch,
                              int rank, int world, int rows_per_rank, int* err_flag, cudaStream_t stream) {
  if (world < 1 || world > 16 || rows_per_rank % 64 || !counters) return DLTI_ERR_BAD_ARG;
  GatherParams g{};
  for (int i = 0; i < world; ++i) g.peer[i] = peer_shards[i];
  g.staging = reinterpret_cast<uint8_t*>(B_staging);
  g.counters = counters;
  g.err = err_flag;
  g.row_bytes = ldb * 2;
  g.target = epoch * (unsigned)GATHER_CH;
  g.rank = rank; g.world = world;
  g.units_per_rank = rows_per_rank / 64;
  g.n_units = g.units_per_rank * world;
  if ((g.row_bytes * 4) % 16) return DLTI_ERR_BAD_ARG;
  return gemm_impl(A, B_staging, C, R, M, N, K, lda, ldb, ldc, ldr, 0, b_mn, A2, B2, K2, lda2, ldb2,
                   in_dtype, out_dtype, alpha, 0, 1, block_n, max_ctas, stream, &g);
}

// K10: row-parallel GEMM whose epilogue stores the (partial) 16-bit output tile into every rank's
// receive buffer at slot `rank` (peer stores over NVLink): C_r = recv_peers[r] + rank*slot_stride.
// The consumer (dlti_tp_reduce_add_norm) sums the slots in rank order after a peer barrier.
DLTI_API int dlti_gemm_push(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb,
                            long long ldc, int in_dtype, int block_n, const void* const* recv_peers,
                            int rank, int world, long long slot_stride, cudaStream_t stream) {
  if (world < 1 || world > 16) return DLTI_ERR_BAD_ARG;
  void* local = const_cast<void*>(recv_peers[rank]);
  return gemm_impl(A, B, local, nullptr, M, N, K, lda, ldb, ldc, 0, 0, 0, nullptr, nullptr, 0, 0, 0, in_dtype,
                   in_dtype, 1.f, 0, 1, block_n, 0, stream, nullptr, recv_peers, rank, world, slot_stride);
}

// Decode-sized GEMM (M tokens <= 256): out32[M, N] += X[M, K] · W[N, K]^T computed as (W · X^T)^T so that
// the WEIGHT is the 128-row tcgen05 A operand (every byte streamed from HBM is useful) and the handful
// of token rows is the narrow B operand; K is split across CTAs to put every SM on the weight stream
// and the fp32 partials are reduced with red.global.add into `acc32` (must be zero on entry;
// dlti_skinny_finish or one of the dlti_acc_* consumers converts it to 16 bits and re-zeroes it).
// split_k: 0 = stream-K (equal K-block ranges per CTA, the default), > 0 = that many tile-major splits,
// < 0 = tile-major splits sized automatically.
DLTI_API int dlti_gemm_skinny(const void* X, const void* W, float* acc32, int M, int N, int K, long long ldx,
                              long long ldw, int in_dtype, int split_k, cudaStream_t stream) {
  if (M < 1 || M > 256) return DLTI_ERR_BAD_ARG;
  const int bn = M <= 64 ? 64 : (M <= 128 ? 128 : 256);
  int stream_k = 0;
  if (split_k == 0) {
    stream_k = 1;                 // default: every CTA streams the same number of K blocks
    split_k = 1;
  } else if (split_k < 0) {       // tile-major split-K sized to fill the SMs (ragged last wave)
    const int tiles = (N + BM - 1) / BM;
    const int nkb = (K + BK - 1) / BK;
    split_k = (dlti_num_sms() + tiles - 1) / tiles;
    while (split_k > 1 && nkb / split_k < 8) --split_k;
  }
  return gemm_impl(W, X, acc32, nullptr, N, M, K, ldw, ldx, /*ldc=*/N, 0, 0, 0, nullptr, nullptr, 0, 0, 0, in_dtype,
                   DLTI_FP32, 1.f, 1, split_k, bn, 0, stream, nullptr, nullptr, 0, 0, 0, /*transpose_out=*/1,
                   stream_k);
}
