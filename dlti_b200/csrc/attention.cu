// K6: causal flash attention for sm_100a — forward, and a two-kernel backward (dK/dV and dQ) —
// on tcgen05 tensor cores with TMEM accumulators and TMA-fed, 128B-swizzled shared tiles.
//
// Layout: activations are token-major. `qkv` is [T, (H + 2*KV) * 128] (q heads | k heads | v heads
// per token row, RoPE already applied), sequences are concatenated and described by cu_seqlens
// (varlen: the same kernels serve training batches and serving prefill). head_dim is 128.
// GQA: query head h reads kv head h / (H / KV).
//
// Forward (one CTA = 128 query rows of one head):
//   warp 0: TMA producer (Q once; K/V blocks of 128 keys, 2 stages)
//   warp 1: MMA issuer:  S_j = Q·K_j^T  -> TMEM (double-buffered);  PV_j = P_j·V_j -> TMEM
//   warps 2-5: softmax, one thread per query row (no cross-thread reductions): S from TMEM via
//     tcgen05.ld, online max/sum in the exp2 domain, P written as bf16 into a swizzled smem tile
//     (the A operand of the PV MMA), O kept in registers (o = o*alpha + PV_j).
//   The tensor core computes S_{j+1} while the softmax warps work on S_j.
//
// Backward uses the stored log-sum-exp and D = rowsum(dO*O):
//   dK/dV kernel (one CTA = 128 keys of one kv head; loops over the q heads of the group and over
//   64-row query blocks): computes S^T = K·Q^T and dP^T = V·dO^T so that one thread owns one KEY
//   row, writes P^T and dS^T as bf16 smem tiles, and accumulates dV += P^T·dO, dK += dS^T·Q in TMEM
//   across the whole loop (no rescaling needed in the backward).
//   dQ kernel (one CTA = 128 query rows of one head; loops over 64-key blocks): S = Q·K^T,
//   dP = dO·V^T, dS -> smem, dQ += dS·K in TMEM. Deterministic: no atomics anywhere.
//
// Replaces PyTorch SDPA (mem-efficient / math kernels on the reference's V100; FlashAttention-2
// mma.sync code elsewhere) — SURVEY.md N5, O6, S5.
#include "common.cuh"

#include <cstdlib>
#include <mutex>
#include <unordered_map>

namespace {

constexpr int HD = 128;           // head dim
// depth of the backward kernels' streamed-tile rings (32 KB per stage). dQ runs ~1000-clk iterations and
// needs ~1500 clk of TMA latency hidden -> 5 stages; dK/dV spends the space on its lse/D staging buffers.
constexpr int BWD_NYS_Q = 5, BWD_NYS_KV = 4;
constexpr int BWD_STAT_CHUNK = 8;        // dK/dV: query blocks whose lse/D are staged per bar.sync
constexpr int BWD_STAT_BYTES = 2 * BWD_STAT_CHUNK * 64 * 2 * 4;   // double-buffered {lse, D} x 512 queries
constexpr int AT_THREADS = 320;   // one-tile forward: producer, mma, 8 softmax warps (two threads per row)
constexpr int BWD_THREADS = 576;  // backward: producer, mma, 16 softmax warps (four threads per row)
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

struct AttnParams {
  const int* cu_seqlens;
  int H, KV;
  int W;                 // qkv row width in elements = (H + 2 KV) * 128
  void* O;               // fwd out [T, H*128]
  long long ldo;
  float* lse;            // [H, T_total] natural-log LSE
  const float* D;        // [H, T_total] rowsum(dO * O) (backward)
  void* dqkv;            // backward out [T, W]
  long long T_total;
  float scale;           // 1/sqrt(hd)
  int dtype;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ uint32_t swz_off(int row, int chunk16) {
  // byte offset inside a [rows x 128 B] 128B-swizzled tile
  return (uint32_t)(row * 128 + ((chunk16 ^ (row & 7)) << 4));
}

template <typename T>
__device__ __forceinline__ void store_row32(uint32_t tile_base, int row, int chunk16_base,
                                            const float (&f)[32]) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    uint32_t u0 = Cvt<T>::pack2(f[8 * k + 0], f[8 * k + 1]);
    uint32_t u1 = Cvt<T>::pack2(f[8 * k + 2], f[8 * k + 3]);
    uint32_t u2 = Cvt<T>::pack2(f[8 * k + 4], f[8 * k + 5]);
    uint32_t u3 = Cvt<T>::pack2(f[8 * k + 6], f[8 * k + 7]);
    const uint32_t addr = tile_base + swz_off(row, chunk16_base + k);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(u0), "r"(u1), "r"(u2),
                 "r"(u3)
                 : "memory");
  }
}

// K-major operand [rows x 128 elements] stored as two [rows x 64] swizzled chunks
// (chunk pitch = rows*128 B). Descriptor for k-step kk (16 elements each) of a K=128 operand.
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t base, int rows, int kk) {
  return umma_smem_desc(base + (kk >> 2) * (rows * 128) + (kk & 3) * 32, 0, 1024);
}
// K-major operand with K = 64 (single chunk).
__device__ __forceinline__ uint64_t desc_kmajor64(uint32_t base, int kk) {
  return umma_smem_desc(base + kk * 32, 0, 1024);
}
// MN-major operand: tile [krows x 128 elements] as two [krows x 64] chunks; N (=128) contiguous,
// K = krows. k-step kk covers rows 16kk..16kk+15.
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t base, int krows, int kk) {
  return umma_smem_desc(base + kk * 2048, krows * 128, 1024);
}

// byte offset of k-step kk inside a K-major [rows x 128] operand (two [rows x 64] chunks)
__device__ __forceinline__ constexpr uint32_t koff_kmajor(int rows, int kk) {
  return (uint32_t)((kk >> 2) * (rows * 128) + (kk & 3) * 32);
}

// =============================================================================================
// Forward
// =============================================================================================
template <typename T>
__global__ void __launch_bounds__(AT_THREADS, 1)
attn_fwd_kernel(const AttnParams p, const __grid_constant__ CUtensorMap tm128) {
  const int b = blockIdx.z, h = blockIdx.y;
  const int row0 = p.cu_seqlens[b];
  const int seqlen = p.cu_seqlens[b + 1] - row0;
  const int nqb = (seqlen + 127) >> 7;
  const int qb = nqb - 1 - (int)blockIdx.x;   // heavy (late) query blocks first
  if (qb < 0) return;
  const int kvh = h / (p.H / p.KV);
  const int nkv = qb + 1;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base, sK = base + 32768, sV = base + 3 * 32768, sP = base + 5 * 32768;
  const uint32_t bars = base + 6 * 32768;
  const uint32_t q_full = bars, k_full = bars + 8, v_full = bars + 24, kv_empty = bars + 40,
                 s_full = bars + 56, s_free = bars + 72, p_full = bars + 88, pv_full = bars + 96,
                 tmem_slot = bars + 104;
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm128);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(k_full + 8 * s, 1);
      mbar_init(v_full + 8 * s, 1);
      mbar_init(kv_empty + 8 * s, 1);
      mbar_init(s_full + 8 * s, 1);
      mbar_init(s_free + 8 * s, 8);
    }
    mbar_init(p_full, 8);
    mbar_init(pv_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_ptr;
  const uint32_t TS0 = 0, TPV = 256;  // S buffers at columns 0 and 128, PV at 256

  if (warp == 0) {
    if (lane == 0) {
      const int qcol = h * HD, kcol = p.H * HD + kvh * HD, vcol = (p.H + p.KV) * HD + kvh * HD;
      mbar_expect_tx(q_full, 32768);
      tma_load_2d(sQ, &tm128, q_full, qcol, row0 + qb * 128);
      tma_load_2d(sQ + 16384, &tm128, q_full, qcol + 64, row0 + qb * 128);
      for (int j = 0; j < nkv; ++j) {
        const int s = j & 1;
        if (j >= 2) mbar_wait(kv_empty + 8 * s, ((j >> 1) - 1) & 1);
        const int r = row0 + j * 128;
        mbar_expect_tx(k_full + 8 * s, 32768);
        tma_load_2d(sK + s * 32768, &tm128, k_full + 8 * s, kcol, r);
        tma_load_2d(sK + s * 32768 + 16384, &tm128, k_full + 8 * s, kcol + 64, r);
        mbar_expect_tx(v_full + 8 * s, 32768);
        tma_load_2d(sV + s * 32768, &tm128, v_full + 8 * s, vcol, r);
        tma_load_2d(sV + s * 32768 + 16384, &tm128, v_full + 8 * s, vcol + 64, r);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_f16(p.dtype, 128, 128, false, false);
      const uint32_t idesc_pv = umma_idesc_f16(p.dtype, 128, 128, false, true);
      auto issue_s = [&](int j) {
        const int s = j & 1;
        mbar_wait(k_full + 8 * s, (j >> 1) & 1);
        if (j >= 2) mbar_wait(s_free + 8 * s, ((j >> 1) - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
          umma_f16_ss(tmem + TS0 + s * 128, desc_kmajor(sQ, 128, kk),
                      desc_kmajor(sK + s * 32768, 128, kk), idesc_s, kk > 0);
        umma_commit(s_full + 8 * s);
      };
      mbar_wait(q_full, 0);
      issue_s(0);
      for (int j = 0; j < nkv; ++j) {
        if (j + 1 < nkv) issue_s(j + 1);
        mbar_wait(p_full, j & 1);
        mbar_wait(v_full + 8 * (j & 1), (j >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
          umma_f16_ss(tmem + TPV, desc_kmajor(sP, 128, kk),
                      desc_mnmajor(sV + (j & 1) * 32768, 128, kk), idesc_pv, kk > 0);
        umma_commit(pv_full);
        umma_commit(kv_empty + 8 * (j & 1));
      }
    }
  } else {
    // Two threads per query row: warps 2-5 take S/P/O columns [0,64), warps 6-9 take [64,128). Twice
    // the warps per scheduler hides the ALU/MUFU/TMEM latencies of the per-element softmax chain; the
    // only cross-thread traffic is one row-max exchange per KV block and one row-sum merge at the end.
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;           // column half owned by this thread
    const int r = quarter * 32 + lane;          // query row inside the tile
    const uint32_t trow = tmem + ((uint32_t)(quarter * 32) << 16);
    const float sl2 = p.scale * LOG2E;
    __shared__ float s_xmax[2][2][128];         // [iteration parity][half][row]
    __shared__ float s_xsum[2][128];
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < nkv; ++j) {
      const int s = j & 1;
      const bool diag = (j == qb);
      mbar_wait(s_full + 8 * s, (j >> 1) & 1);
      tc_fence_after();
      float mx = -INFINITY;
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const int c = half * 2 + cc;
        uint32_t v[32];
        tmem_ld_32x32b_x32(trow + TS0 + s * 128 + c * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float sv = __uint_as_float(v[i]) * sl2;
          if (diag && (c * 32 + i > r)) sv = -INFINITY;
          mx = fmaxf(mx, sv);
        }
      }
      s_xmax[s][half][r] = mx;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      mx = fmaxf(m, fmaxf(mx, s_xmax[s][half ^ 1][r]));
      const float alpha = exp2f(m - mx);
      if (j > 0) {
        mbar_wait(pv_full, (j - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(trow + TPV + (half * 2 + cc) * 32, v);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[cc * 32 + i] = (o[cc * 32 + i] + __uint_as_float(v[i])) * alpha;
        }
      }
      l *= alpha;
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const int c = half * 2 + cc;
        uint32_t v[32];
        tmem_ld_32x32b_x32(trow + TS0 + s * 128 + c * 32, v);
        tmem_wait_ld();
        float pr[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float pv = exp2f(__uint_as_float(v[i]) * sl2 - mx);
          if (diag && (c * 32 + i > r)) pv = 0.f;
          // sum the bf16-rounded probabilities the PV MMA will actually use
          pv = Cvt<T>::to_f(Cvt<T>::from_f(pv));
          l += pv;
          pr[i] = pv;
        }
        store_row32<T>(sP + half * 16384, r, cc * 4, pr);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(s_free + 8 * s);
        mbar_arrive(p_full);
      }
      m = mx;
    }
    mbar_wait(pv_full, (nkv - 1) & 1);
    tc_fence_after();
    s_xsum[half][r] = l;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    l += s_xsum[half ^ 1][r];
    const float inv = 1.f / l;
    const int qi = qb * 128 + r;
    T* orow = reinterpret_cast<T*>(p.O) + (long long)(row0 + qi) * p.ldo + h * HD + half * 64;
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(trow + TPV + (half * 2 + cc) * 32, v);
      tmem_wait_ld();
      if (qi < seqlen) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          int4 ov;
          uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int i = 8 * k + 2 * q;
            ou[q] = Cvt<T>::pack2((o[cc * 32 + i] + __uint_as_float(v[i])) * inv,
                                  (o[cc * 32 + i + 1] + __uint_as_float(v[i + 1])) * inv);
          }
          *reinterpret_cast<int4*>(orow + cc * 32 + 8 * k) = ov;
        }
      }
    }
    if (half == 0 && qi < seqlen && p.lse) p.lse[(long long)h * p.T_total + row0 + qi] = (m + log2f(l)) * LN2;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// =============================================================================================
// Backward preprocess: D[h, t] = sum_d dO[t, h, d] * O[t, h, d].  One warp per (t, h).
// =============================================================================================
template <typename T>
__global__ void attn_bwd_prep_kernel(const T* __restrict__ dO, const T* __restrict__ O, float* __restrict__ D,
                                     long long T_total, int H, long long ld) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= T_total * H) return;
  const long long t = w / H;
  const int h = (int)(w - t * H);
  const uint2 a = *reinterpret_cast<const uint2*>(dO + t * ld + h * HD + lane * 4);
  const uint2 c = *reinterpret_cast<const uint2*>(O + t * ld + h * HD + lane * 4);
  float2 a0 = Cvt<T>::unpack2(a.x), a1 = Cvt<T>::unpack2(a.y);
  float2 c0 = Cvt<T>::unpack2(c.x), c1 = Cvt<T>::unpack2(c.y);
  float acc = a0.x * c0.x + a0.y * c0.y + a1.x * c1.x + a1.y * c1.y;
  acc = warp_sum(acc);
  if (lane == 0) D[(long long)h * T_total + t] = acc;
}

// =============================================================================================
// Backward main kernels. One template:
//   DKDV = true : stationary X1 = K_j, X2 = V_j (128 keys); streamed Y1 = Q_i, Y2 = dO_i (64 rows),
//                 looping over the q heads of the GQA group and the causal range of q blocks.
//                 thread = key row. acc1 (dV) += P^T·dO, acc2 (dK) += dS^T·Q.
//   DKDV = false: stationary X1 = Q_i, X2 = dO_i (128 query rows); streamed Y1 = K_j, Y2 = V_j
//                 (64 keys). thread = query row. acc2 (dQ) += dS·K.
// In both, tile A = X1·Y1^T (scores) and tile B = X2·Y2^T (dP) are [128 x 64] fp32 in TMEM,
// double-buffered; P / dS are written back over those tiles as packed 16-bit pairs (tcgen05.st) and
// feed the accumulate MMAs as the TMEM A operand — no shared-memory round trip, and softmax of
// iteration it+1 never waits for the accumulate MMAs of iteration it.
// =============================================================================================
#ifdef DLTI_ATTN_TRACE
// Developer build only: clock64() timeline of one CTA (role x iteration x event) for latency hunting.
__device__ long long g_attn_trace[2 * 3 * 64 * 4];   // [dQ | dK/dV][role][iteration][event]
#define TRACE(role, it, ev) do { if (trace_cta && (it) < 64) g_attn_trace[(((DKDV ? 3 : 0) + (role)) * 64 + (it)) * 4 + (ev)] = clock64(); } while (0)
#else
#define TRACE(role, it, ev) do { } while (0)
#endif

template <typename T, bool DKDV>
__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_kernel(const AttnParams p, const __grid_constant__ CUtensorMap tmQKV128,
                const __grid_constant__ CUtensorMap tmQKV64,
                const __grid_constant__ CUtensorMap tmDO128,
                const __grid_constant__ CUtensorMap tmDO64) {
  const int b = blockIdx.z;
  const int row0 = p.cu_seqlens[b];
  const int seqlen = p.cu_seqlens[b + 1] - row0;
  const int G = p.H / p.KV;
  const int n64 = (seqlen + 63) >> 6;
  // stationary 128-row block index. dK/dV: block 0 has the most query blocks to visit; dQ: the LAST
  // query block has the most key blocks -> walk it in reverse so heavy CTAs are scheduled first.
  const int nblk128 = (seqlen + 127) >> 7;
  const int blk = DKDV ? (int)blockIdx.x : nblk128 - 1 - (int)blockIdx.x;
  if (blk < 0 || blk * 128 >= seqlen) return;
  const int head = blockIdx.y;              // DKDV: kv head; else: q head
#ifdef DLTI_ATTN_TRACE
  const bool trace_cta = blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (threadIdx.x & 31) == 0;
#endif
  // streamed block range and iteration count
  int y0, ny, iters;
  if (DKDV) {
    y0 = 2 * blk; ny = n64 - y0; iters = ny * G;
  } else {
    y0 = 0; ny = min(2 * blk + 2, n64); iters = ny;
  }

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sX1 = base, sX2 = base + 32768;
  // streamed tiles: NYS-deep ring. The load for iteration `it` is issued when the accumulate MMAs of
  // iteration it-NYS retire and is needed when those of it-2 retire, so NYS-2 iterations of TMA latency
  // (~1500 clk measured) are hidden.
  constexpr int NYS = DKDV ? BWD_NYS_KV : BWD_NYS_Q;
  const uint32_t sY1 = base + 65536;                    // NYS stages x 16 KB
  const uint32_t sY2 = base + 65536 + NYS * 16384;      // NYS stages x 16 KB
  const uint32_t sStat = base + 65536 + 2 * NYS * 16384;   // dK/dV: 2 x {512 lse, 512 D} floats
  const uint32_t bars = sStat + (DKDV ? BWD_STAT_BYTES : 1024);
  const uint32_t x_full = bars, y_full = bars + 8, y_empty = y_full + 8 * NYS, ab_full = y_empty + 8 * NYS,
                 pds_full = ab_full + 16, acc_done = pds_full + 16, x_tmem = acc_done + 8, tmem_slot = x_tmem + 8;
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
  float* stat = reinterpret_cast<float*>(smem_raw + (sStat - smem_u32(smem_raw)));
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler

  if (warp == 0 && lane == 0) {
    mbar_init(x_full, 1);
    for (int s = 0; s < NYS; ++s) {
      mbar_init(y_full + 8 * s, 1);
      mbar_init(y_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(ab_full + 8 * s, 1);
      mbar_init(pds_full + 8 * s, 16);
    }
    mbar_init(acc_done, 1);
    mbar_init(x_tmem, 16);
    mbar_fence_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot_ptr, 0);
  // TMEM columns: A tiles at 0/64, B tiles at 128/192, acc1 at 256, acc2 at 384. The dQ kernel has no
  // acc1: it parks its stationary Q and dO tiles there (packed 16-bit, 64 columns each) so the score
  // MMAs take their 128-row operand from TMEM instead of re-reading 4 KB of shared memory per MMA —
  // with N = 64 that re-read made the tensor pipe shared-memory-bound (traced: 950 clk to issue 16 MMAs).
  const uint32_t TA = 0, TB = 128, TACC1 = 256, TACC2 = 384, TX1 = 256, TX2 = 320;

  const int kcol_base = p.H * HD, vcol_base = (p.H + p.KV) * HD;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(x_full, 65536);
      if (DKDV) {
        const int kc = kcol_base + head * HD, vc = vcol_base + head * HD, r = row0 + blk * 128;
        tma_load_2d(sX1, &tmQKV128, x_full, kc, r);
        tma_load_2d(sX1 + 16384, &tmQKV128, x_full, kc + 64, r);
        tma_load_2d(sX2, &tmQKV128, x_full, vc, r);
        tma_load_2d(sX2 + 16384, &tmQKV128, x_full, vc + 64, r);
      } else {
        const int qc = head * HD, r = row0 + blk * 128;
        tma_load_2d(sX1, &tmQKV128, x_full, qc, r);
        tma_load_2d(sX1 + 16384, &tmQKV128, x_full, qc + 64, r);
        tma_load_2d(sX2, &tmDO128, x_full, qc, r);
        tma_load_2d(sX2 + 16384, &tmDO128, x_full, qc + 64, r);
      }
      for (int it = 0; it < iters; ++it) {
        const int s = it % NYS;
        if (it >= NYS) mbar_wait(y_empty + 8 * s, ((it / NYS) - 1) & 1);
        TRACE(2, it, 0);
        mbar_expect_tx(y_full + 8 * s, 32768);
        if (DKDV) {
          const int g = it / ny, i = y0 + (it - g * ny);
          const int qc = (head * G + g) * HD, r = row0 + i * 64;
          tma_load_2d(sY1 + s * 16384, &tmQKV64, y_full + 8 * s, qc, r);
          tma_load_2d(sY1 + s * 16384 + 8192, &tmQKV64, y_full + 8 * s, qc + 64, r);
          tma_load_2d(sY2 + s * 16384, &tmDO64, y_full + 8 * s, qc, r);
          tma_load_2d(sY2 + s * 16384 + 8192, &tmDO64, y_full + 8 * s, qc + 64, r);
        } else {
          const int kvh = head / G;
          const int kc = kcol_base + kvh * HD, vc = vcol_base + kvh * HD, r = row0 + (y0 + it) * 64;
          tma_load_2d(sY1 + s * 16384, &tmQKV64, y_full + 8 * s, kc, r);
          tma_load_2d(sY1 + s * 16384 + 8192, &tmQKV64, y_full + 8 * s, kc + 64, r);
          tma_load_2d(sY2 + s * 16384, &tmQKV64, y_full + 8 * s, vc, r);
          tma_load_2d(sY2 + s * 16384 + 8192, &tmQKV64, y_full + 8 * s, vc + 64, r);
        }
      }
    }
  } else if (warp == 1) {
    {   // whole warp runs the loop (uniform operands stay in uniform registers); one elected lane issues
      const bool leader = elect_one_sync();
      const uint32_t idesc_ab = umma_idesc_f16(p.dtype, 128, 64, false, false);
      const uint32_t idesc_acc = umma_idesc_f16(p.dtype, 128, 128, false, true);
      // descriptor bases (stage 0); a Y stage is 16 KB = 1024 descriptor units further
      const DescLoHi dX1 = desc_split(sX1, 0, 1024), dX2 = desc_split(sX2, 0, 1024);
      const DescLoHi dY1k = desc_split(sY1, 0, 1024), dY2k = desc_split(sY2, 0, 1024);          // K-major view
      const DescLoHi dY1m = desc_split(sY1, 64 * 128, 1024), dY2m = desc_split(sY2, 64 * 128, 1024);  // MN-major view
      auto issue_ab = [&](int it) {
        const int s = it & 1, ys = it % NYS;
        mbar_wait(y_full + 8 * ys, (it / NYS) & 1);
        TRACE(1, it, 0);          // Y tile present
        tc_fence_after();
        const uint32_t tA = tmem + TA + s * 64, tB = tmem + TB + s * 64;
        DescLoHi y1 = dY1k, y2 = dY2k;
        y1.lo += ys * 1024; y2.lo += ys * 1024;
        if (!leader) return;
        if (DKDV) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_f16_ss(tA, dX1.at(koff_kmajor(128, kk)), y1.at(koff_kmajor(64, kk)), idesc_ab, kk > 0);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_f16_ss(tB, dX2.at(koff_kmajor(128, kk)), y2.at(koff_kmajor(64, kk)), idesc_ab, kk > 0);
        } else {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_f16_ts(tA, tmem + TX1 + kk * 8, y1.at(koff_kmajor(64, kk)), idesc_ab, kk > 0);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_f16_ts(tB, tmem + TX2 + kk * 8, y2.at(koff_kmajor(64, kk)), idesc_ab, kk > 0);
        }
        umma_commit(ab_full + 8 * s);
        TRACE(1, it, 1);          // score MMAs issued
      };
      // P / dS come back as packed 16-bit pairs written over the score tiles themselves (each softmax
      // thread overwrites only columns it has already read): k-step kk (16 elements) sits at columns
      // 16*kk .. 16*kk+7 of the 64-column tile. The accumulate MMAs read them as the TMEM A operand.
      mbar_wait(x_full, 0);
      if (!DKDV) {
        mbar_wait(x_tmem, 0);
        tc_fence_after();
      }
      issue_ab(0);
      if (iters > 1) issue_ab(1);
      for (int it = 0; it < iters; ++it) {
        const int s = it & 1, ys = it % NYS;
        mbar_wait(pds_full + 8 * s, (it >> 1) & 1);
        TRACE(1, it, 2);          // P / dS present
        tc_fence_after();
        const uint32_t tA = tmem + TA + s * 64, tB = tmem + TB + s * 64;
        DescLoHi y1 = dY1m, y2 = dY2m;
        y1.lo += ys * 1024; y2.lo += ys * 1024;
        const uint32_t acc_on = it > 0 ? 1u : 0u;
        if (leader) {
        if (DKDV) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)   // dV += P^T · dO
            umma_f16_ts(tmem + TACC1, tA + kk * 16, y2.at(kk * 2048), idesc_acc,
                        kk > 0 ? 1u : acc_on);
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)     // dK += dS^T · Q   /   dQ += dS · K
          umma_f16_ts(tmem + TACC2, tB + kk * 16, y1.at(kk * 2048), idesc_acc,
                      kk > 0 ? 1u : acc_on);
        umma_commit(y_empty + 8 * ys);
        }
        TRACE(1, it, 3);          // accumulate MMAs issued
        // the tensor pipe runs MMAs in issue order, so the next scores may overwrite this buffer now
        if (it + 2 < iters) issue_ab(it + 2);
      }
      if (leader) umma_commit(acc_done);
    }
  } else {
    // FOUR threads per row (16 softmax warps): warp w owns rows 32*(w&3).. and streamed columns
    // [16*part, 16*part+16), part = (w-2)>>2. Four warps per scheduler keep the MUFU/FMA pipes fed where two
    // left the per-iteration softmax (1160 clk traced) longer than the MMA work it overlaps with.
    const int quarter = warp & 3;
    const int part = (warp - 2) >> 2;          // 0..3
    const int r = quarter * 32 + lane;
    const int sidx = blk * 128 + r;            // key index (DKDV) or query index (!DKDV)
    const uint32_t trow = tmem + ((uint32_t)(quarter * 32) << 16);
    const float sl2 = p.scale * LOG2E;
    const int tid = threadIdx.x - 64;          // 0..511
    float my_lse = 0.f, my_D = 0.f;
    if (!DKDV && sidx < seqlen) {
      my_lse = p.lse[(long long)head * p.T_total + row0 + sidx] * LOG2E;
      my_D = p.D[(long long)head * p.T_total + row0 + sidx];
    }
    // DKDV: per-query lse / D are staged through smem in chunks of 8 query blocks (one bar.sync per
    // chunk instead of per iteration); the global loads of the NEXT chunk are issued right after a chunk
    // is published, so they have eight iterations to land.
    constexpr int CH = BWD_STAT_CHUNK;
    float pre_l = 0.f, pre_d = 0.f;
    auto load_chunk = [&](int it0) {           // it0: iteration at which the chunk starts (may be >= iters)
      if (it0 >= iters) return;
      const int g = it0 / ny;
      const int q = (y0 + (it0 - g * ny)) * 64 + tid;
      const long long hb = (long long)(head * G + g) * p.T_total + row0;
      const bool ok = q < seqlen;
      pre_l = ok ? p.lse[hb + q] : 0.f;
      pre_d = ok ? p.D[hb + q] : 0.f;
    };
    int ch_it0 = 0, ch_buf = 0;                // start iteration and smem buffer of the current chunk
    int ch_next = 0;                           // iteration at which the next chunk starts
    if (DKDV) load_chunk(0);
    if (!DKDV) {
      // this thread's row of Q and dO (its 32-element quarter of head_dim) -> TMEM, packed pairs in K order
      mbar_wait(x_full, 0);
#pragma unroll
      for (int x = 0; x < 2; ++x) {
        const uint32_t src = (x ? sX2 : sX1) + (part >> 1) * 16384;
        uint32_t w[16];
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4) {
          const uint32_t addr = src + swz_off(r, (part & 1) * 4 + c4);
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(w[4 * c4]), "=r"(w[4 * c4 + 1]), "=r"(w[4 * c4 + 2]), "=r"(w[4 * c4 + 3])
                       : "r"(addr));
        }
        tmem_st_32x32b_x16(trow + (x ? TX2 : TX1) + part * 16, w);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(x_tmem);
    }
    for (int it = 0; it < iters; ++it) {
      const int s = it & 1;
      int yb;                                   // streamed 64-block index
      if (DKDV) {
        const int g = it / ny;
        const int li = it - g * ny;
        yb = y0 + li;
        if (it == ch_next) {
          ch_it0 = it;
          ch_buf ^= 1;
          const int left = ny - li;             // a chunk never crosses into the next q head
          ch_next = it + (left < CH ? left : CH);
          float* dst = stat + ch_buf * (CH * 128);
          dst[tid] = pre_l * LOG2E;
          dst[CH * 64 + tid] = pre_d;
          load_chunk(ch_next);
          asm volatile("bar.sync 1, 512;" ::: "memory");
        }
      } else {
        yb = y0 + it;
      }
      // Only the blocks that touch the causal diagonal (or the end of the sequence) need a mask.
      const bool need_mask = DKDV ? (yb < 2 * blk + 2 || yb * 64 + 64 > seqlen) : (yb >= 2 * blk);
      if (warp == 2) TRACE(0, it, 0);      // iteration entered
      mbar_wait(ab_full + 8 * s, (it >> 1) & 1);
      if (warp == 2) TRACE(0, it, 1);      // scores present
      tc_fence_after();
      {
        uint32_t sv[16], dv[16];
        tmem_ld_32x32b_x16(trow + TA + s * 64 + part * 16, sv);
        tmem_ld_32x32b_x16(trow + TB + s * 64 + part * 16, dv);
        tmem_wait_ld();
        if (warp == 2) TRACE(0, it, 2);    // scores in registers
        float lse_v[16], D_v[16];
        if (DKDV) {
          const float* cb = stat + ch_buf * (CH * 128) + (it - ch_it0) * 64 + part * 16;
          const float4* l4 = reinterpret_cast<const float4*>(cb);
          const float4* d4 = reinterpret_cast<const float4*>(cb + CH * 64);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 a = l4[i], b4 = d4[i];
            lse_v[4 * i] = a.x; lse_v[4 * i + 1] = a.y; lse_v[4 * i + 2] = a.z; lse_v[4 * i + 3] = a.w;
            D_v[4 * i] = b4.x; D_v[4 * i + 1] = b4.y; D_v[4 * i + 2] = b4.z; D_v[4 * i + 3] = b4.w;
          }
        }
        // p = 2^(s*scale*log2e - lse);  ds = p * (dp - D)   (the 1/sqrt(d) factor of dS is applied once,
        // to the dK / dQ accumulators, in the epilogue)
        float pr[16], ds[16];
        if (!need_mask) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float x = fmaf(__uint_as_float(sv[i]), sl2, -(DKDV ? lse_v[i] : my_lse));
            const float pv = ex2_approx(x);
            pr[i] = pv;
            ds[i] = pv * (__uint_as_float(dv[i]) - (DKDV ? D_v[i] : my_D));
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int yi = yb * 64 + part * 16 + i;  // query index (DKDV) or key index (!DKDV)
            const bool valid = DKDV ? ((yi < seqlen) && (yi >= sidx)) : ((sidx < seqlen) && (yi <= sidx));
            float x = fmaf(__uint_as_float(sv[i]), sl2, -(DKDV ? lse_v[i] : my_lse));
            x = valid ? x : -INFINITY;            // exp2(-inf) = 0 exactly; no NaNs from garbage rows
            const float pv = ex2_approx(x);
            pr[i] = pv;
            ds[i] = valid ? pv * (__uint_as_float(dv[i]) - (DKDV ? D_v[i] : my_D)) : 0.f;
          }
        }
        // packed pairs go back over this thread's own (already read) score columns: k-step `part`
        // (16 elements) occupies columns 16*part .. 16*part+7 of the tile
        uint32_t pk[8];
        if (DKDV) {
#pragma unroll
          for (int i = 0; i < 8; ++i) pk[i] = Cvt<T>::pack2(pr[2 * i], pr[2 * i + 1]);
          tmem_st_32x32b_x8(trow + TA + s * 64 + part * 16, pk);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) pk[i] = Cvt<T>::pack2(ds[2 * i], ds[2 * i + 1]);
        tmem_st_32x32b_x8(trow + TB + s * 64 + part * 16, pk);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_full + 8 * s);
      if (warp == 2) TRACE(0, it, 3);      // P / dS published
    }
    mbar_wait(acc_done, 0);
    tc_fence_after();
    // write the accumulators: this thread's row, columns [32*part, 32*part+32)
    T* outbase = reinterpret_cast<T*>(p.dqkv) + (long long)(row0 + sidx) * p.W;
    const int nacc = DKDV ? 2 : 1;
    for (int a = 0; a < nacc; ++a) {
      const uint32_t tacc = (DKDV && a == 0) ? TACC1 : TACC2;
      const float osc = (tacc == TACC2) ? p.scale : 1.f;     // dK / dQ carry the softmax scale
      int col;
      if (DKDV) col = (a == 0) ? (vcol_base + head * HD) : (kcol_base + head * HD);
      else col = head * HD;
      uint32_t v[32];
      tmem_ld_32x32b_x32(trow + tacc + part * 32, v);
      tmem_wait_ld();
      if (sidx < seqlen) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          int4 ov;
          uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
          for (int q = 0; q < 4; ++q)
            ou[q] = Cvt<T>::pack2(__uint_as_float(v[8 * k + 2 * q]) * osc, __uint_as_float(v[8 * k + 2 * q + 1]) * osc);
          *reinterpret_cast<int4*>(outbase + col + part * 32 + 8 * k) = ov;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// =============================================================================================
// Forward, ping-pong form (the default): one CTA = TWO adjacent 128-row query tiles of one head.
//
// The single-tile kernel above is latency-bound (ncu: tensor pipe 17.6 % active): every
// S -> softmax -> P -> PV hand-off sits on the critical path. Here two query tiles A and B share each
// K/V block (half the K/V traffic) and alternate on the tensor core: while the softmax warps of one
// tile work, the MMAs of the other run.
//   warps 0-3: softmax tile A, warps 4-7: softmax tile B — one thread per query row, the whole
//     128-wide score row in registers (row max/sum need no cross-thread traffic);
//   warp 8: TMA producer (Q_A, Q_B once; K and V blocks of 128 keys, 2 stages each, released separately);
//   warp 9: MMA issuer.
// TMEM (512 columns): S_A | S_B | O_A | O_B. P is written back as packed 16-bit pairs over the first 64
// columns of its own S buffer (tcgen05.st) and read by the PV MMA as the TMEM A operand, so P never
// touches shared memory. O accumulates in TMEM over all KV blocks; the reference maximum is only
// moved (and O rescaled in TMEM) when a row maximum grows by more than 2^8 — exact after the final
// division by the row sum, and rare in practice.
// =============================================================================================
constexpr int F2_THREADS = 320;
constexpr int F2_SMEM = 6 * 32768 + 256 + 1024;
constexpr long long F2_WAIT_LIMIT = 4000000000LL;

__device__ __forceinline__ void mbar_wait_b(uint32_t bar, uint32_t parity) {
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > F2_WAIT_LIMIT) asm volatile("trap;");
  }
}

template <typename T>
__global__ void __launch_bounds__(F2_THREADS, 1)
attn_fwd2_kernel(const AttnParams p, const __grid_constant__ CUtensorMap tm128) {
  const int b = blockIdx.z, h = blockIdx.y;
  const int row0 = p.cu_seqlens[b];
  const int seqlen = p.cu_seqlens[b + 1] - row0;
  const int nqb = (seqlen + 127) >> 7;
  const int npair = (nqb + 1) >> 1;
  const int pr = npair - 1 - (int)blockIdx.x;   // heavy (late) pairs first
  if (pr < 0) return;
  const int qbA = 2 * pr, qbB = 2 * pr + 1;
  const int nA = qbA + 1, nB = (qbB < nqb) ? qbB + 1 : 0;   // KV blocks each tile attends to
  const int nblk = nB > nA ? nB : nA;
  const bool hasB = nB > 0;
  const int kvh = h / (p.H / p.KV);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base, sK = base + 2 * 32768, sV = base + 4 * 32768;
  const uint32_t bars = base + 6 * 32768;
  const uint32_t q_full = bars, k_full = bars + 8, v_full = bars + 24, k_empty = bars + 40, v_empty = bars + 56,
                 s_full = bars + 72, p_full = bars + 88, pv_done = bars + 104, tmem_slot = bars + 120;
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tm128);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(k_full + 8 * s, 1);
      mbar_init(v_full + 8 * s, 1);
      mbar_init(k_empty + 8 * s, 1);
      mbar_init(v_empty + 8 * s, 1);
      mbar_init(s_full + 8 * s, 1);
      mbar_init(p_full + 8 * s, 4);
      mbar_init(pv_done + 8 * s, 1);
    }
    mbar_fence_init();
  }
  if (warp == 9) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot_ptr, 0);

  if (warp == 8) {
    if (lane == 0) {
      const int qcol = h * HD, kcol = p.H * HD + kvh * HD, vcol = (p.H + p.KV) * HD + kvh * HD;
      mbar_expect_tx(q_full, hasB ? 65536 : 32768);
      tma_load_2d(sQ, &tm128, q_full, qcol, row0 + qbA * 128);
      tma_load_2d(sQ + 16384, &tm128, q_full, qcol + 64, row0 + qbA * 128);
      if (hasB) {
        tma_load_2d(sQ + 32768, &tm128, q_full, qcol, row0 + qbB * 128);
        tma_load_2d(sQ + 32768 + 16384, &tm128, q_full, qcol + 64, row0 + qbB * 128);
      }
      for (int j = 0; j < nblk; ++j) {
        const int s = j & 1;
        const int r = row0 + j * 128;
        if (j >= 2) mbar_wait_b(k_empty + 8 * s, ((j >> 1) - 1) & 1);
        mbar_expect_tx(k_full + 8 * s, 32768);
        tma_load_2d(sK + s * 32768, &tm128, k_full + 8 * s, kcol, r);
        tma_load_2d(sK + s * 32768 + 16384, &tm128, k_full + 8 * s, kcol + 64, r);
        if (j >= 2) mbar_wait_b(v_empty + 8 * s, ((j >> 1) - 1) & 1);
        mbar_expect_tx(v_full + 8 * s, 32768);
        tma_load_2d(sV + s * 32768, &tm128, v_full + 8 * s, vcol, r);
        tma_load_2d(sV + s * 32768 + 16384, &tm128, v_full + 8 * s, vcol + 64, r);
      }
    }
  } else if (warp == 9) {
    {   // whole warp runs the loop (uniform operands stay in uniform registers); one elected lane issues
      const bool leader = elect_one_sync();
      const uint32_t idesc_s = umma_idesc_f16(p.dtype, 128, 128, false, false);
      const uint32_t idesc_pv = umma_idesc_f16(p.dtype, 128, 128, false, true);
      const DescLoHi dQ0 = desc_split(sQ, 0, 1024), dK0 = desc_split(sK, 0, 1024);
      const DescLoHi dV0 = desc_split(sV, 128 * 128, 1024);     // MN-major view of a [128 keys x 128] tile
      auto issue_s = [&](int X, int j) {
        DescLoHi q = dQ0, k = dK0;
        q.lo += X * 2048; k.lo += (j & 1) * 2048;                // 32 KB = 2048 descriptor units
        const uint32_t tS = tmem + X * 128;
        if (!leader) return;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
          umma_f16_ss(tS, q.at(koff_kmajor(128, kk)), k.at(koff_kmajor(128, kk)), idesc_s, kk > 0);
        umma_commit(s_full + 8 * X);
      };
      auto issue_pv = [&](int X, int j) {
        DescLoHi v = dV0;
        v.lo += (j & 1) * 2048;
        const uint32_t tO = tmem + 256 + X * 128, tP = tmem + X * 128;
        const uint32_t acc_on = j > 0 ? 1u : 0u;
        if (!leader) return;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
          umma_f16_ts(tO, tP + kk * 8, v.at(kk * 2048), idesc_pv, kk > 0 ? 1u : acc_on);
        umma_commit(pv_done + 8 * X);
      };
      mbar_wait_b(q_full, 0);
      mbar_wait_b(k_full, 0);
      tc_fence_after();
      issue_s(0, 0);
      if (hasB) issue_s(1, 0);
      if (leader) umma_commit(k_empty);
      for (int j = 0; j < nblk; ++j) {
        const int s = j & 1;
        const bool actA = j < nA, actB = j < nB, nxtA = j + 1 < nA, nxtB = j + 1 < nB;
        mbar_wait_b(v_full + 8 * s, (j >> 1) & 1);
        if (nxtA || nxtB) mbar_wait_b(k_full + 8 * (s ^ 1), ((j + 1) >> 1) & 1);
        if (actA) {
          mbar_wait_b(p_full, j & 1);
          tc_fence_after();
          issue_pv(0, j);
          if (!actB && leader) umma_commit(v_empty + 8 * s);
        }
        if (nxtA) {
          issue_s(0, j + 1);
          if (!nxtB && leader) umma_commit(k_empty + 8 * (s ^ 1));
        }
        if (actB) {
          mbar_wait_b(p_full + 8, j & 1);
          tc_fence_after();
          issue_pv(1, j);
          if (leader) umma_commit(v_empty + 8 * s);
        }
        if (nxtB) {
          issue_s(1, j + 1);
          if (leader) umma_commit(k_empty + 8 * (s ^ 1));
        }
      }
    }
  } else {
    const int X = warp >> 2;                    // 0: tile A, 1: tile B
    const int nX = X ? nB : nA;
    const int qbX = X ? qbB : qbA;
    if (nX > 0) {
      const int quarter = warp & 3;
      const int r = quarter * 32 + lane;          // query row inside the tile
      const uint32_t tS = tmem + ((uint32_t)(quarter * 32) << 16) + X * 128;
      const uint32_t tO = tmem + ((uint32_t)(quarter * 32) << 16) + 256 + X * 128;
      const float sl2 = p.scale * LOG2E;
      float m_ref = -INFINITY, l = 0.f;
      for (int j = 0; j < nX; ++j) {
        const bool diag = (j == qbX);
        mbar_wait_b(s_full + 8 * X, j & 1);
        tc_fence_after();
        uint32_t v[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(tS + c * 32, *reinterpret_cast<uint32_t(*)[32]>(v + 32 * c));
        tmem_wait_ld();
        float mx = -INFINITY;
        if (diag) {
#pragma unroll
          for (int i = 0; i < 128; ++i) {
            if (i > r) v[i] = __float_as_uint(-INFINITY);
            mx = fmaxf(mx, __uint_as_float(v[i]));
          }
        } else {
#pragma unroll
          for (int i = 0; i < 128; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
        }
        mx *= sl2;
        if (j == 0) {
          m_ref = mx;
        } else {
          const float mnew = fmaxf(m_ref, mx);
          if (__any_sync(0xffffffffu, (mnew - m_ref) > 8.f)) {
            // move the reference maximum: O (in TMEM) and l shrink by 2^(old - new)
            const float alpha = ex2_approx(m_ref - mnew);
            m_ref = mnew;
            l *= alpha;
            mbar_wait_b(pv_done + 8 * X, (j - 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 8; ++c) {
              uint32_t o[16];
              tmem_ld_32x32b_x16(tO + c * 16, o);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              tmem_st_32x32b_x16(tO + c * 16, o);
            }
            tmem_wait_st();
          }
        }
        const float neg_m = -m_ref;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float p0 = ex2_approx(fmaf(__uint_as_float(v[32 * c + 2 * i]), sl2, neg_m));
            const float p1 = ex2_approx(fmaf(__uint_as_float(v[32 * c + 2 * i + 1]), sl2, neg_m));
            l += p0 + p1;
            pk[i] = Cvt<T>::pack2(p0, p1);
          }
          tmem_st_32x32b_x16(tS + c * 16, pk);
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full + 8 * X);
      }
      mbar_wait_b(pv_done + 8 * X, (nX - 1) & 1);
      tc_fence_after();
      const float inv = 1.f / l;
      const int qi = qbX * 128 + r;
      T* orow = reinterpret_cast<T*>(p.O) + (long long)(row0 + qi) * p.ldo + h * HD;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(tO + c * 32, o);
        tmem_wait_ld();
        if (qi < seqlen) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            int4 ov;
            uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
            for (int q = 0; q < 4; ++q)
              ou[q] = Cvt<T>::pack2(__uint_as_float(o[8 * k + 2 * q]) * inv, __uint_as_float(o[8 * k + 2 * q + 1]) * inv);
            *reinterpret_cast<int4*>(orow + c * 32 + 8 * k) = ov;
          }
        }
      }
      if (qi < seqlen && p.lse) p.lse[(long long)h * p.T_total + row0 + qi] = (m_ref + log2f(l)) * LN2;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

constexpr int FWD_SMEM = 6 * 32768 + 256 + 1024;
constexpr int BWD_SMEM_Q = 65536 + 2 * BWD_NYS_Q * 16384 + 1024 + 256 + 1024;
constexpr int BWD_SMEM_KV = 65536 + 2 * BWD_NYS_KV * 16384 + BWD_STAT_BYTES + 256 + 1024;

int make_map(CUtensorMap* m, const void* base, long long T_total, long long width, int box_rows) {
  return dlti_make_tmap_2d(m, base, (uint64_t)width, (uint64_t)T_total, (uint64_t)width, 64,
                           (uint32_t)box_rows, 2, 128);
}

}  // namespace

#ifdef DLTI_ATTN_TRACE
DLTI_API int dlti_attn_trace_read(long long* host_out /* [2*3*64*4] */) {
  return (int)cudaMemcpyFromSymbol(host_out, g_attn_trace, sizeof(long long) * 2 * 3 * 64 * 4);
}
#endif

// qkv [T, (H+2KV)*128]; out [T, H*128]; lse [H, T] fp32 (may be NULL for inference).
DLTI_API int dlti_attn_fwd(const void* qkv, void* out, float* lse, const int* cu_seqlens, int batch,
                           int max_seqlen, long long T_total, int H, int KV, int hd, float scale,
                           int dtype, cudaStream_t st) {
  if (hd != HD) { dlti_set_error("attention: head_dim must be 128"); return DLTI_ERR_UNSUPPORTED; }
  if (H % KV) return DLTI_ERR_BAD_ARG;
  if (batch <= 0 || T_total <= 0) return 0;
  AttnParams p{};
  p.cu_seqlens = cu_seqlens; p.H = H; p.KV = KV; p.W = (H + 2 * KV) * HD;
  p.O = out; p.ldo = (long long)H * HD; p.lse = lse; p.T_total = T_total; p.scale = scale; p.dtype = dtype;
  CUtensorMap tm;
  int rc = make_map(&tm, qkv, T_total, p.W, 128);
  if (rc) return rc;
  // DLTI_ATTN_FWD=1 selects the older one-tile-per-CTA kernel (kept for A/B measurements)
  static const bool pingpong = []() { const char* e = getenv("DLTI_ATTN_FWD"); return !e || e[0] != '1'; }();
  if (pingpong) {
    dim3 grid2(((max_seqlen + 127) / 128 + 1) / 2, H, batch);
    static bool cfg2[2] = {false, false};
    if (dtype == DLTI_BF16) {
      if (!cfg2[0]) { DLTI_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd2_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, F2_SMEM)); cfg2[0] = true; }
      attn_fwd2_kernel<__nv_bfloat16><<<grid2, F2_THREADS, F2_SMEM, st>>>(p, tm);
    } else if (dtype == DLTI_FP16) {
      if (!cfg2[1]) { DLTI_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd2_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, F2_SMEM)); cfg2[1] = true; }
      attn_fwd2_kernel<__half><<<grid2, F2_THREADS, F2_SMEM, st>>>(p, tm);
    } else {
      return DLTI_ERR_UNSUPPORTED;
    }
    dlti_count_launch();
    return (int)cudaGetLastError();
  }
  dim3 grid((max_seqlen + 127) / 128, H, batch);
  static bool cfg[2] = {false, false};
  if (dtype == DLTI_BF16) {
    if (!cfg[0]) { DLTI_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM)); cfg[0] = true; }
    attn_fwd_kernel<__nv_bfloat16><<<grid, AT_THREADS, FWD_SMEM, st>>>(p, tm);
  } else if (dtype == DLTI_FP16) {
    if (!cfg[1]) { DLTI_CUDA_CHECK(cudaFuncSetAttribute(attn_fwd_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM)); cfg[1] = true; }
    attn_fwd_kernel<__half><<<grid, AT_THREADS, FWD_SMEM, st>>>(p, tm);
  } else {
    return DLTI_ERR_UNSUPPORTED;
  }
  dlti_count_launch();
  return (int)cudaGetLastError();
}

// dqkv [T, (H+2KV)*128] receives dq | dk | dv. D_scratch: [H, T] fp32 workspace.
DLTI_API int dlti_attn_bwd(const void* dout, const void* qkv, const void* out, const float* lse,
                           float* D_scratch, void* dqkv, const int* cu_seqlens, int batch,
                           int max_seqlen, long long T_total, int H, int KV, int hd, float scale,
                           int dtype, cudaStream_t st) {
  if (hd != HD) { dlti_set_error("attention: head_dim must be 128"); return DLTI_ERR_UNSUPPORTED; }
  if (H % KV) return DLTI_ERR_BAD_ARG;
  if (batch <= 0 || T_total <= 0) return 0;
  AttnParams p{};
  p.cu_seqlens = cu_seqlens; p.H = H; p.KV = KV; p.W = (H + 2 * KV) * HD;
  p.lse = const_cast<float*>(lse); p.D = D_scratch; p.dqkv = dqkv; p.T_total = T_total;
  p.scale = scale; p.dtype = dtype; p.ldo = (long long)H * HD;
  CUtensorMap q128, q64, d128, d64;
  int rc;
  if ((rc = make_map(&q128, qkv, T_total, p.W, 128))) return rc;
  if ((rc = make_map(&q64, qkv, T_total, p.W, 64))) return rc;
  if ((rc = make_map(&d128, dout, T_total, (long long)H * HD, 128))) return rc;
  if ((rc = make_map(&d64, dout, T_total, (long long)H * HD, 64))) return rc;
  const long long warps = T_total * H;
  const int prep_blocks = (int)((warps * 32 + 255) / 256);
  dim3 g_kv((max_seqlen + 127) / 128, KV, batch), g_q((max_seqlen + 127) / 128, H, batch);
  static bool cfg[2] = {false, false};
  if (dtype == DLTI_BF16) {
    using T = __nv_bfloat16;
    if (!cfg[0]) {
      DLTI_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM_KV));
      DLTI_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM_Q));
      cfg[0] = true;
    }
    attn_bwd_prep_kernel<T><<<prep_blocks, 256, 0, st>>>((const T*)dout, (const T*)out, D_scratch, T_total, H, p.ldo);
    attn_bwd_kernel<T, true><<<g_kv, BWD_THREADS, BWD_SMEM_KV, st>>>(p, q128, q64, d128, d64);
    attn_bwd_kernel<T, false><<<g_q, BWD_THREADS, BWD_SMEM_Q, st>>>(p, q128, q64, d128, d64);
  } else if (dtype == DLTI_FP16) {
    using T = __half;
    if (!cfg[1]) {
      DLTI_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM_KV));
      DLTI_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM_Q));
      cfg[1] = true;
    }
    attn_bwd_prep_kernel<T><<<prep_blocks, 256, 0, st>>>((const T*)dout, (const T*)out, D_scratch, T_total, H, p.ldo);
    attn_bwd_kernel<T, true><<<g_kv, BWD_THREADS, BWD_SMEM_KV, st>>>(p, q128, q64, d128, d64);
    attn_bwd_kernel<T, false><<<g_q, BWD_THREADS, BWD_SMEM_Q, st>>>(p, q128, q64, d128, d64);
  } else {
    return DLTI_ERR_UNSUPPORTED;
  }
  dlti_count_launch(3);
  return (int)cudaGetLastError();
}
