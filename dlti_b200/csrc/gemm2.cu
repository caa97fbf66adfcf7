// K2 (large-M form): the same GEMM as gemm.cu computed by CTA PAIRS (tcgen05 cta_group::2).
//
// Why: with one CTA per 128x256 tile every k-block moves 48 KB into shared memory (TMA) and the four
// MMAs read 48 KB back, ~96 + 96 B/clk against a 128 B/clk shared-memory port — ncu shows the tensor
// pipe capped near 75 % (profiles/ncu_summary_r1.md). A CTA pair computes a 256x256 tile: each CTA
// stages its own 128 rows of A and only HALF of the B tile (32 KB per k-block instead of 48), the
// tensor cores of both SMs read B from both shared memories, and each CTA ends up with its 128x256
// slice of the accumulator in its own TMEM.
//
//   C[M,N] = alpha * ( A[M,K] · op(B) + A2[M,K2] · op(B2) ) (+ R)      16-bit or fp32 output
//   B K-major ([N,K], x·W^T) or MN-major ([K,N], dy·W); K-extension (LoRA) as in gemm.cu.
//
// Roles per CTA: warp 0 / warp 6 TMA producers of A / B (each CTA loads its own halves; completion bytes of BOTH
// CTAs land on the leader's "full" barrier), warp 1 = TMEM allocator (both CTAs) + MMA issuer (leader only; its
// tcgen05.commit multicasts the "slot free" / "accumulator ready" arrivals to both CTAs), warps 2-5 =
// epilogue on the CTA's own 128 accumulator rows. Every wait is bounded: on timeout the kernel traps
// instead of hanging the GPU.
#include <cstdlib>

#include "common.cuh"

namespace {

constexpr int BM = 128;      // rows per CTA (256 per pair)
constexpr int BN = 256;      // columns per pair (128 B-rows staged per CTA)
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 224;      // A producer, MMA, 4 epilogue warps, B producer
constexpr int A_BYTES = BM * BK * 2;          // 16 KB
constexpr int B_BYTES = (BN / 2) * BK * 2;    // 16 KB (this CTA's half of the B tile)
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int STAGES = 6;
constexpr int TMEM_COLS = 512;                // two accumulator stages of 256 columns
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 256 + 1024;
constexpr int SCATTER_STAGE_BYTES = 4 * 4096;      // one 32 x 64 16-bit block per epilogue warp (K10 scatter epilogue)
constexpr long long WAIT_LIMIT = 4000000000LL;   // ~2 s

struct Gemm2Params {
  int M, N;
  int nmb2, nnb;          // 256-row / 256-col tile counts
  int nkb_main, nkb_ext;
  void* C;
  const void* R;
  long long ldc, ldr;
  int out_dtype, in_dtype;
  int accumulate;
  int vec_ok;
  float alpha;
  uint32_t idesc;
  // Ragged last wave: when the tiles left for the last wave fill at most half of the clusters, each of them
  // is cut into two 256x128 halves (units >= split_from; UMMA N = 128, idesc_half), so the last wave lasts
  // half as long. split_from == nmb2 * nnb: no tile is split.
  uint32_t idesc_half;
  int split_from;
  // K11 (all-gather -> column-parallel GEMM): the A operand is a staging buffer whose 64-row units are pulled from
  // their owner ranks by this launch's gather warp; the A producer waits on the unit counters (WaitSpec::on_a) and
  // the M blocks are visited starting at m_rot (the block that holds this rank's own rows, which land first).
  int m_rot;
  // K10 (row-parallel GEMM -> reduce-scatter), prefill-sized steps: output row r belongs to rank r / scatter_per;
  // the epilogue stores the 16-bit tile straight into that rank's receive slot of this rank:
  //   scatter_peer[owner] + ((long long)scatter_rank * scatter_per + (r - owner * scatter_per)) * ldc + col
  // (peer stores over NVLink, 128 contiguous bytes per thread and row block). The owner sums the slots in rank order.
  void* scatter_peer[16];
  int scatter_world, scatter_rank, scatter_per;
  int b3d;                // MN-major B: tmB is the 3D view (one TMA per stage for a full tile); tmBh stays 2D
  // stream-K: tiles [0, sk_tiles) are cut into `gridDim/2` equal ranges of (tile, k-block) iterations,
  // tiles [sk_tiles, total) are handed out round-robin. sk_tiles == 0 => plain data-parallel schedule.
  // SwiGLU mode (swiglu_ff > 0): B is the [2*ff, K] gate|up weight; CTA rank 0 stages 128 gate rows and
  // rank 1 the matching 128 up rows, so accumulator columns 0-127 are gate and 128-255 are up of the
  // same 128 output columns, and the epilogue writes h = silu(gate) * up  ([M, ff]).
  int swiglu_ff;
  // SwiGLU-backward mode (swiglu_bwd_ff > 0): the accumulator is dh = dy · W_down ([M, ff]); the epilogue
  // reads gate|up from `gu` ([M, 2*ff]) and writes d(gate|up) into C ([M, 2*ff]); dh is never stored.
  int swiglu_bwd_ff;
  const void* gu;
  long long ldgu;
  // RoPE mode (rope_pos != nullptr): output columns [0, rope_cols) are attention heads of 128 that get the
  // rotate-half rotary embedding of their row's position applied in the epilogue (q and k heads of the
  // fused QKV projection); columns >= rope_cols (v heads) are stored unchanged.
  const int* rope_pos;
  const float* rope_cos;    // [max_pos, 64] fp32
  const float* rope_sin;
  int rope_cols;
  int sk_tiles;
  long long sk_iters_per_range;
  float4* sk_ws;            // [ranges][2 CTAs][8 chunks][8 quads][128 rows] float4 partial accumulators
  unsigned int* sk_flags;   // [ranges] ready counters + [ranges] consumed counters
};

// K1 (ZeRO-3 all-gather fused into the GEMM), CTA-pair form. Two independent attachments of a launch:
//  * a gather JOB: one extra warp per CTA pulls 64-row units of a row-sharded weight matrix from their owner
//    ranks (peer loads over NVLink from the symmetric arena; the own rows with local loads) into a local
//    staging buffer and bump a per-unit counter with a release once a 4-row chunk has landed. The job may
//    belong to the B operand of this very launch (cold start) or — the normal case — to a matrix that a LATER
//    launch consumes (the next layer's weights ride on this layer's GEMMs, a slice per launch);
//  * a WAIT spec: the B operand of this launch is such a staging buffer; the TMA producer acquires the
//    counters of the units a tile needs before it loads them (one relaxed sweep at kernel start takes the fast
//    path when everything has already landed).
// TMA only ever reads this GPU's staging buffer; peer memory is read with ld.global by the gather warp.
constexpr int GATHER_CH = 16;              // chunks per 64-row unit (4 rows each)
constexpr int GATHER_THREADS = 32;          // one warp: the CTA stays at 256 threads (ptxas budgets registers per
                                           // 128 threads: 288 threads would cap the epilogue at 168 and spill)
constexpr int GATHER_MLP = 16;              // 16-byte loads in flight per gather thread
struct GatherJob {
  const void* peer[16];     // every rank's shard base (peer-mapped; peer[rank] is local)
  uint8_t* staging;         // local full matrix [world * rows_per_rank, row_bytes]
  uint32_t* counters;       // [n_units] completed-chunk counters, monotonic across epochs
  int* err;
  long long row_bytes;
  uint32_t target;          // unit complete  <=>  counters[u] >= target (= epoch * GATHER_CH)
  int units_per_rank, n_units;
  int rot;                  // job position j is unit (rot + j) % n_units
  int j_begin, j_count;     // positions [j_begin, j_begin + j_count) belong to this launch; 0 = no job
};
struct WaitSpec {
  const uint32_t* counters; // nullptr: the operand is plain memory
  int* err;
  uint32_t target;
  int n_units;
  int on_a;                 // 0: the units are rows of the B operand (K1); 1: rows of the A operand (K11)
};

__device__ __forceinline__ uint32_t ld_relaxed_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void wait_unit(const WaitSpec& w, int u) {
  if (u >= w.n_units) return;
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_gpu_u32(w.counters + u) - w.target) < 0) {
    if (clock64() - t0 > 20000000000LL) {
      if (w.err) atomicExch(w.err, 1);
      return;
    }
    __nanosleep(32);
  }
}

// One unit of work: k-blocks [kb0, kb1) of one 256x256 tile. kb0 > 0: the tile's head belongs to another
// cluster -> this piece is written to the workspace; kb1 < nkb: this cluster owns the tile and adds the
// tail piece from the workspace before the normal epilogue.
struct Seg {
  int tile, kb0, kb1;
  int half;      // -1: the whole 256x256 tile; 0 / 1: its left / right 256x128 half
};
struct SegIter {
  long long it, end;      // stream-K iteration range of this cluster
  int next_dp, total, step, nkb, split_from;
  __device__ SegIter(const Gemm2Params& p, int cid, int ncl) {
    nkb = p.nkb_main + p.nkb_ext;
    split_from = p.split_from;
    total = p.nmb2 * p.nnb;
    total += total - split_from;      // work units: whole tiles, then two halves for every split tile
    step = ncl;
    next_dp = p.sk_tiles + cid;
    if (p.sk_tiles > 0) {
      const long long all = (long long)p.sk_tiles * nkb;
      const int r = ncl - 1 - cid;      // range index: a tile's owner has the HIGHER cluster id than its helper
      it = r * p.sk_iters_per_range;
      end = it + p.sk_iters_per_range;
      if (it > all) it = all;
      if (end > all) end = all;
    } else {
      it = end = 0;
    }
  }
  __device__ bool next(Seg& s) {
    if (it < end) {
      s.tile = (int)(it / nkb);
      s.kb0 = (int)(it - (long long)s.tile * nkb);
      const long long left = end - it;
      s.kb1 = (left < nkb - s.kb0) ? s.kb0 + (int)left : nkb;
      s.half = -1;
      it += s.kb1 - s.kb0;
      return true;
    }
    if (next_dp < total) {
      if (next_dp < split_from) {
        s.tile = next_dp; s.half = -1;
      } else {
        const int h = next_dp - split_from;
        s.tile = split_from + (h >> 1); s.half = h & 1;
      }
      s.kb0 = 0; s.kb1 = nkb;
      next_dp += step;
      return true;
    }
    return false;
  }
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity) {
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > WAIT_LIMIT) asm volatile("trap;");
  }
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion bytes are signalled on a barrier that may live in the peer CTA.
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst_smem, const CUtensorMap* m, uint32_t bar_cluster,
                                                int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d_2sm(uint32_t dst_smem, const CUtensorMap* m, uint32_t bar_cluster,
                                                int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t dst_smem_addr, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem_addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_f16_ss_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this smem offset in BOTH CTAs of the pair once prior MMAs retire
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar) {
  const uint16_t mask = 3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"(mask)
      : "memory");
}

template <bool B_MN>
// __maxnreg__(200) instead of __launch_bounds__: at 256 threads ptxas would otherwise take 248 registers, i.e. the
// whole register file of the SM, and the small peer-memory kernels that run next to the GEMMs on a side stream
// (gradient reduce-scatter overlapped with backward, K9) could only start at GEMM boundaries. 200 spills nothing.
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(200)
gemm2_kernel(const Gemm2Params p, const GatherJob gj, const WaitSpec ws_, const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
             const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
             const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmB2h) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * STAGES + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  __shared__ void* s_scatter[16];
  if (p.scatter_world > 0) {
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (threadIdx.x == 32 + i) s_scatter[i] = p.scatter_peer[i];
  }
  __shared__ const void* s_peers[16];
  if (gj.j_count > 0) {
#pragma unroll
    for (int i = 0; i < 16; ++i)   // static indices keep the by-value parameter in the constant bank
      if (threadIdx.x == i) s_peers[i] = gj.peer[i];
  }
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // uniform for the compiler
  const uint32_t rank = __shfl_sync(0xffffffffu, cluster_ctarank(), 0);
  const bool leader = rank == 0;
  const int cid = blockIdx.x >> 1, ncl = gridDim.x >> 1;
  const int total_tiles = p.nmb2 * p.nnb;
  const int nkb_total = p.nkb_main + p.nkb_ext;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);     // leader's copy is the one in use: 1 arrive (expect_tx) + bytes of both CTAs
      mbar_init(empty_bar(s), 1);    // multicast tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);    // multicast tcgen05.commit
      mbar_init(tempty_bar(s), 8);   // leader's copy: 4 epilogue warps x 2 CTAs
    }
    mbar_fence_init();
  }
  cluster_sync_all();
  if (warp == 1) {
    tmem_alloc2(tmem_slot, TMEM_COLS);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot_ptr, 0);

  if (warp == 0) {
    // ===================== A producer (both CTAs): this CTA's 128 rows of A =====================
    // One elected thread; the loop body is kept minimal — a k-block lasts ~512 clk on the tensor cores and every
    // UTMALDG costs the issuing thread on the order of 100 clk, which is why A and B have separate producer
    // threads (with one thread the input-gradient form, two B boxes per stage, was issue-bound: 1,180 vs cuBLAS
    // 1,296 TFLOP/s sustained).
    bool a_ready = true;          // K11: every unit of A has already landed
    if (ws_.counters && ws_.on_a) {
      for (int u = lane; u < ws_.n_units; u += 32)
        if ((int32_t)(ld_relaxed_gpu_u32(ws_.counters + u) - ws_.target) < 0) a_ready = false;
      a_ready = __all_sync(0xffffffffu, a_ready);
      asm volatile("fence.acq_rel.gpu;" ::: "memory");
      asm volatile("fence.proxy.async.global;" ::: "memory");
    }
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      const uint32_t fb0 = mapa_rank(full_bar(0), 0);          // the leader's barriers (mapping is linear)
      SegIter segs(p, cid, ncl);
      Seg sg;
      while (segs.next(sg)) {
        const int n_blk = sg.tile / p.nmb2;
        int m_blk = sg.tile - n_blk * p.nmb2 + p.m_rot;
        if (m_blk >= p.nmb2) m_blk -= p.nmb2;
        const int m0 = m_blk * 2 * BM + (int)rank * BM;
        if (!a_ready) {               // this CTA's rows [m0, m0 + 128) = 2 units
          wait_unit(ws_, m0 >> 6);
          wait_unit(ws_, (m0 >> 6) + 1);
          asm volatile("fence.proxy.async.global;" ::: "memory");
        }
        const uint32_t tx = sg.half >= 0 ? 2 * (A_BYTES + B_BYTES / 2) : 2 * STAGE_BYTES;
        const int kb_main_end = sg.kb1 < p.nkb_main ? sg.kb1 : p.nkb_main;
        int kb = sg.kb0;
        for (; kb < kb_main_end; ++kb) {
          mbar_wait_bounded(empty_bar(stage), phase ^ 1u);
          if (leader) mbar_expect_tx(full_bar(stage), tx);
          tma_load_2d_2sm(smem_base + stage * STAGE_BYTES, &tmA, fb0 + 8u * stage, kb * BK, m0);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        for (; kb < sg.kb1; ++kb) {                               // K-extension (LoRA) blocks
          mbar_wait_bounded(empty_bar(stage), phase ^ 1u);
          if (leader) mbar_expect_tx(full_bar(stage), tx);
          tma_load_2d_2sm(smem_base + stage * STAGE_BYTES, &tmA2, fb0 + 8u * stage, (kb - p.nkb_main) * BK, m0);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 6) {
    // ===================== B producer (both CTAs): this CTA's half of the B tile =====================
    bool all_ready = true;        // WAIT: every unit of B has already landed (gathered by earlier launches)
    if (ws_.counters && !ws_.on_a) {
      for (int u = lane; u < ws_.n_units; u += 32)
        if ((int32_t)(ld_relaxed_gpu_u32(ws_.counters + u) - ws_.target) < 0) all_ready = false;
      all_ready = __all_sync(0xffffffffu, all_ready);
      asm volatile("fence.acq_rel.gpu;" ::: "memory");
      asm volatile("fence.proxy.async.global;" ::: "memory");
    }
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      const uint32_t fb0 = mapa_rank(full_bar(0), 0);
      const uint32_t sb0 = smem_base + A_BYTES;
      SegIter segs(p, cid, ncl);
      Seg sg;
      while (segs.next(sg)) {
        const int n_blk = sg.tile / p.nmb2;
        // this CTA's half of the B tile (SwiGLU mode: gate rows for rank 0, the matching up rows for rank 1);
        // a half tile needs only 64 B rows per CTA (tmBh / tmB2h: the same matrices with 64-row boxes)
        const bool hf = sg.half >= 0;
        int n0;
        if (!hf) n0 = p.swiglu_ff > 0 ? n_blk * (BN / 2) + (int)rank * p.swiglu_ff : n_blk * BN + (int)rank * (BN / 2);
        else n0 = p.swiglu_ff > 0 ? n_blk * (BN / 2) + sg.half * (BN / 4) + (int)rank * p.swiglu_ff
                                  : n_blk * BN + sg.half * (BN / 2) + (int)rank * (BN / 4);
        if (!B_MN && !all_ready) {      // K-major B: this CTA's rows [n0, n0 + 128 or 64) = 2 or 1 units
          wait_unit(ws_, n0 >> 6);
          if (!hf) wait_unit(ws_, (n0 >> 6) + 1);
          asm volatile("fence.proxy.async.global;" ::: "memory");
        }
        const CUtensorMap* mb = hf ? &tmBh : &tmB;
        const bool two = B_MN && !hf && !p.b3d;                  // MN-major full tile as two 64x64 boxes
        const bool one3d = B_MN && !hf && p.b3d;
        const int kb_main_end = sg.kb1 < p.nkb_main ? sg.kb1 : p.nkb_main;
        int kb = sg.kb0;
        if (all_ready || !B_MN) {
          for (; kb < kb_main_end; ++kb) {
            mbar_wait_bounded(empty_bar(stage), phase ^ 1u);
            const uint32_t sb = sb0 + stage * STAGE_BYTES, fb = fb0 + 8u * stage;
            if constexpr (!B_MN) {
              tma_load_2d_2sm(sb, mb, fb, kb * BK, n0);
            } else if (one3d) {
              tma_load_3d_2sm(sb, mb, fb, 0, kb * BK, n0 >> 6);
            } else {
              tma_load_2d_2sm(sb, mb, fb, n0, kb * BK);
              if (two) tma_load_2d_2sm(sb + BK * 128, mb, fb, n0 + 64, kb * BK);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
        } else {
          for (; kb < kb_main_end; ++kb) {                        // MN-major B still landing: unit kb per k-block
            wait_unit(ws_, kb);
            asm volatile("fence.proxy.async.global;" ::: "memory");
            mbar_wait_bounded(empty_bar(stage), phase ^ 1u);
            const uint32_t sb = sb0 + stage * STAGE_BYTES, fb = fb0 + 8u * stage;
            if (one3d) {
              tma_load_3d_2sm(sb, mb, fb, 0, kb * BK, n0 >> 6);
            } else {
              tma_load_2d_2sm(sb, mb, fb, n0, kb * BK);
              if (two) tma_load_2d_2sm(sb + BK * 128, mb, fb, n0 + 64, kb * BK);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
          }
        }
        const CUtensorMap* mb2 = hf ? &tmB2h : &tmB2;
        for (; kb < sg.kb1; ++kb) {                               // K-extension (LoRA) blocks: plain memory
          mbar_wait_bounded(empty_bar(stage), phase ^ 1u);
          const uint32_t sb = sb0 + stage * STAGE_BYTES, fb = fb0 + 8u * stage;
          const int k0 = (kb - p.nkb_main) * BK;
          if constexpr (!B_MN) {
            tma_load_2d_2sm(sb, mb2, fb, k0, n0);
          } else {
            tma_load_2d_2sm(sb, mb2, fb, n0, k0);
            if (!hf) tma_load_2d_2sm(sb + BK * 128, mb2, fb, n0 + 64, k0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    // The whole warp runs the loop (uniform control flow and operands); one elected lane issues.
    if (leader) {
      const bool el = elect_one_sync();
      const DescLoHi dA = desc_split(smem_base, 0, 1024);
      const DescLoHi dB = B_MN ? desc_split(smem_base + A_BYTES, BK * 128, 1024) : desc_split(smem_base + A_BYTES, 0, 1024);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      SegIter segs(p, cid, ncl);
      Seg sg;
      while (segs.next(sg)) {
        mbar_wait_bounded(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        const uint32_t idesc = sg.half < 0 ? p.idesc : p.idesc_half;
        for (int kb = sg.kb0; kb < sg.kb1; ++kb) {
          mbar_wait_bounded(full_bar(stage), phase);
          tc_fence_after();
          DescLoHi a = dA, b = dB;
          a.lo += stage * (STAGE_BYTES >> 4);
          b.lo += stage * (STAGE_BYTES >> 4);
          if (el) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_f16_ss_2sm(d_tmem, a.at(k * (UMMA_K * 2)), b.at(B_MN ? k * (UMMA_K * 128) : k * (UMMA_K * 2)),
                              idesc, (kb > sg.kb0 || k > 0) ? 1u : 0u);
            umma_commit_2sm(empty_bar(stage));
            if (kb == sg.kb1 - 1) umma_commit_2sm(tfull_bar(acc));
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else if (warp >= 7) {
    // ===================== gather warp (present when a job is attached) =====================
    const int gt = threadIdx.x - NUM_THREADS;                  // 0..31
    const long long chunk_bytes = 4 * gj.row_bytes;            // 4 rows
    const long long unit_bytes = 64 * gj.row_bytes;
    const int total_chunks = gj.j_count * GATHER_CH;
    for (int c = blockIdx.x; c < total_chunks; c += gridDim.x) {
      const int j = gj.j_begin + c / GATHER_CH, part = c % GATHER_CH;
      int u = gj.rot + j;
      u -= (u / gj.n_units) * gj.n_units;
      const int owner = u / gj.units_per_rank;
      const uint8_t* src = reinterpret_cast<const uint8_t*>(s_peers[owner]) +
                           (long long)(u - owner * gj.units_per_rank) * unit_bytes + part * chunk_bytes;
      uint8_t* dst = gj.staging + (long long)u * unit_bytes + part * chunk_bytes;
      const long long nvec = chunk_bytes >> 4;
      long long i = gt;
      for (; i + (GATHER_MLP - 1) * GATHER_THREADS < nvec; i += GATHER_MLP * GATHER_THREADS) {
        int4 v[GATHER_MLP];
#pragma unroll
        for (int q = 0; q < GATHER_MLP; ++q) v[q] = ld_nc_v4(src + ((i + q * GATHER_THREADS) << 4));
#pragma unroll
        for (int q = 0; q < GATHER_MLP; ++q) st_v4(dst + ((i + q * GATHER_THREADS) << 4), v[q]);
      }
      for (; i < nvec; i += GATHER_THREADS) st_v4(dst + (i << 4), ld_nc_v4(src + (i << 4)));
      __threadfence();
      __syncwarp();
      if (gt == 0) red_release_gpu_add_u32(gj.counters + u, 1u);
    }
  } else {
    // ===================== epilogue (both CTAs; own 128 rows x 256 columns) =====================
    const int quarter = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    const uint32_t leader_tempty0 = mapa_rank(tempty_bar(0), 0), leader_tempty1 = mapa_rank(tempty_bar(1), 0);
    SegIter segs(p, cid, ncl);
    Seg sg;
    const int my_range = ncl - 1 - cid;
    while (segs.next(sg)) {
      const int n_blk = sg.tile / p.nmb2;
      int m_blk = sg.tile - n_blk * p.nmb2 + p.m_rot;
      if (m_blk >= p.nmb2) m_blk -= p.nmb2;
      const int row = m_blk * 2 * BM + (int)rank * BM + quarter * 32 + lane;
      const int ncols = sg.half < 0 ? BN : BN / 2;                 // accumulator columns of this unit
      const int n0 = n_blk * BN + (sg.half > 0 ? BN / 2 : 0);
      const bool helper = sg.kb0 > 0;              // write raw partial accumulators to the workspace
      const bool fixup = sg.kb1 < nkb_total;       // add the helper's partial (range my_range + 1)
      // float4 index of (chunk 0, quad 0, this thread's row) in the workspace slot used by this segment
      float4* ws = p.sk_ws + ((size_t)(helper ? my_range : my_range + 1) * 2 + rank) * (size_t)(64 * BM) +
                   (quarter * 32 + lane);
      if (fixup) {
        const unsigned int* flag = p.sk_flags + my_range + 1;
        const long long t0 = clock64();
        while (ld_acquire_gpu_u32(flag) < 8u) {
          if (clock64() - t0 > 5 * WAIT_LIMIT) asm volatile("trap;");
        }
      }
      mbar_wait_bounded(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
      if (p.swiglu_ff > 0) {
        // 16-bit module-boundary semantics of the unfused path: gate, up and silu(gate) are rounded to the
        // storage type before the product (bit-identical to dlti_swiglu_fwd on the stored gate|up tensor).
        const bool bf = p.out_dtype == DLTI_BF16;
        uint16_t* crow = reinterpret_cast<uint16_t*>(p.C) + (long long)row * p.ldc + n_blk * (BN / 2) +
                         (sg.half > 0 ? BN / 4 : 0);
#pragma unroll 1
        for (int c = 0; c < ncols / 64; ++c) {
          uint32_t vg[32], vu[32];
          tmem_ld_32x32b_x32(t_row + c * 32, vg);
          tmem_ld_32x32b_x32(t_row + ncols / 2 + c * 32, vu);
          tmem_wait_ld();
          if (row < p.M) {
            float o[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              float g = __uint_as_float(vg[i]), u = __uint_as_float(vu[i]);
              g = bf ? __bfloat162float(__float2bfloat16_rn(g)) : __half2float(__float2half_rn(g));
              u = bf ? __bfloat162float(__float2bfloat16_rn(u)) : __half2float(__float2half_rn(u));
              float sg = g / (1.f + __expf(-g));
              sg = bf ? __bfloat162float(__float2bfloat16_rn(sg)) : __half2float(__float2half_rn(sg));
              o[i] = sg * u;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              int4 ov;
              uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                ou[j] = bf ? Cvt<__nv_bfloat16>::pack2(o[8 * i + 2 * j], o[8 * i + 2 * j + 1])
                           : Cvt<__half>::pack2(o[8 * i + 2 * j], o[8 * i + 2 * j + 1]);
              *(reinterpret_cast<int4*>(crow + c * 32) + i) = ov;
            }
          }
        }
      } else if (p.rope_pos != nullptr) {
        // Same rounding points as the stand-alone kernel: the projection is rounded to the storage type,
        // cos/sin are rounded to it too (HF casts them), the rotation runs in fp32 and is rounded once.
        const bool bf = p.out_dtype == DLTI_BF16;
        const bool rok = row < p.M;
        const int pos = rok ? p.rope_pos[row] : 0;
        const float4* cosr = reinterpret_cast<const float4*>(p.rope_cos + (long long)pos * 64);
        const float4* sinr = reinterpret_cast<const float4*>(p.rope_sin + (long long)pos * 64);
        uint16_t* crow = reinterpret_cast<uint16_t*>(p.C) + (long long)row * p.ldc;
        auto rnd = [&](float x) { return bf ? __bfloat162float(__float2bfloat16_rn(x)) : __half2float(__float2half_rn(x)); };
        auto store32 = [&](uint16_t* dst, const float (&f)[32]) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            int4 o;
            uint32_t* ou = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              ou[j] = bf ? Cvt<__nv_bfloat16>::pack2(f[8 * i + 2 * j], f[8 * i + 2 * j + 1])
                         : Cvt<__half>::pack2(f[8 * i + 2 * j], f[8 * i + 2 * j + 1]);
            *(reinterpret_cast<int4*>(dst) + i) = o;
          }
        };
#pragma unroll 1
        for (int hb = 0; hb < ncols / 128; ++hb) {
          const int hcol = n0 + hb * 128;
          if (hcol >= p.N) break;
          if (hcol < p.rope_cols) {
#pragma unroll 1
            for (int pr = 0; pr < 2; ++pr) {            // element i pairs with i + 64: chunks pr and pr + 2
              uint32_t v1[32], v2[32];
              tmem_ld_32x32b_x32(t_row + hb * 128 + pr * 32, v1);
              tmem_ld_32x32b_x32(t_row + hb * 128 + 64 + pr * 32, v2);
              float cs[32], sn[32];
              if (rok) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  const float4 c4 = cosr[pr * 8 + i], s4 = sinr[pr * 8 + i];
                  cs[4 * i] = c4.x; cs[4 * i + 1] = c4.y; cs[4 * i + 2] = c4.z; cs[4 * i + 3] = c4.w;
                  sn[4 * i] = s4.x; sn[4 * i + 1] = s4.y; sn[4 * i + 2] = s4.z; sn[4 * i + 3] = s4.w;
                }
              }
              tmem_wait_ld();
              if (rok) {
                float o1[32], o2[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                  const float a = rnd(__uint_as_float(v1[i]) * p.alpha), b = rnd(__uint_as_float(v2[i]) * p.alpha);
                  const float cj = rnd(cs[i]), sj = rnd(sn[i]);
                  o1[i] = a * cj - b * sj;
                  o2[i] = b * cj + a * sj;
                }
                store32(crow + hcol + pr * 32, o1);
                store32(crow + hcol + 64 + pr * 32, o2);
              }
            }
          } else {
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
              uint32_t v[32];
              tmem_ld_32x32b_x32(t_row + hb * 128 + c * 32, v);
              tmem_wait_ld();
              if (rok) {
                float f[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.alpha;
                store32(crow + hcol + c * 32, f);
              }
            }
          }
        }
      } else if (p.swiglu_bwd_ff > 0) {
        // dgate = dh * up * sig * (1 + gate * (1 - sig)),  dup = dh * gate * sig, with dh rounded to the
        // storage type first (bit-identical to dlti_swiglu_bwd on a stored dh). gate|up of chunk c+1 are
        // requested before chunk c is processed so the global-load latency stays off the epilogue's path.
        const bool bf = p.out_dtype == DLTI_BF16;
        const int ff = p.swiglu_bwd_ff;
        const bool rok = row < p.M;
        const uint16_t* grow = reinterpret_cast<const uint16_t*>(p.gu) + (long long)(rok ? row : 0) * p.ldgu;
        uint16_t* orow = reinterpret_cast<uint16_t*>(p.C) + (long long)row * p.ldc;
        int4 gq[4], uq[4];
        auto fetch = [&](int c) {
          const int col0 = n0 + c * 32;
          if (rok && col0 < ff) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              gq[i] = *(reinterpret_cast<const int4*>(grow + col0) + i);
              uq[i] = *(reinterpret_cast<const int4*>(grow + ff + col0) + i);
            }
          }
        };
        fetch(0);
#pragma unroll 1
        for (int c = 0; c < ncols / 32; ++c) {
          const int col0 = n0 + c * 32;
          if (col0 >= ff) break;
          uint32_t v[32];
          tmem_ld_32x32b_x32(t_row + c * 32, v);
          int4 gc[4], uc[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) { gc[i] = gq[i]; uc[i] = uq[i]; }
          if (c + 1 < ncols / 32) fetch(c + 1);
          tmem_wait_ld();
          if (rok) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t* gw = reinterpret_cast<const uint32_t*>(&gc[i]);
              const uint32_t* uw = reinterpret_cast<const uint32_t*>(&uc[i]);
              int4 og, ou4;
              uint32_t* ogw = reinterpret_cast<uint32_t*>(&og);
              uint32_t* ouw = reinterpret_cast<uint32_t*>(&ou4);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float2 g2 = bf ? Cvt<__nv_bfloat16>::unpack2(gw[j]) : Cvt<__half>::unpack2(gw[j]);
                const float2 u2 = bf ? Cvt<__nv_bfloat16>::unpack2(uw[j]) : Cvt<__half>::unpack2(uw[j]);
                float d0 = __uint_as_float(v[8 * i + 2 * j]), d1 = __uint_as_float(v[8 * i + 2 * j + 1]);
                d0 = bf ? __bfloat162float(__float2bfloat16_rn(d0)) : __half2float(__float2half_rn(d0));
                d1 = bf ? __bfloat162float(__float2bfloat16_rn(d1)) : __half2float(__float2half_rn(d1));
                const float s0 = 1.f / (1.f + __expf(-g2.x)), s1 = 1.f / (1.f + __expf(-g2.y));
                const float dg0 = d0 * u2.x * s0 * (1.f + g2.x * (1.f - s0));
                const float dg1 = d1 * u2.y * s1 * (1.f + g2.y * (1.f - s1));
                const float du0 = d0 * (g2.x * s0), du1 = d1 * (g2.y * s1);
                ogw[j] = bf ? Cvt<__nv_bfloat16>::pack2(dg0, dg1) : Cvt<__half>::pack2(dg0, dg1);
                ouw[j] = bf ? Cvt<__nv_bfloat16>::pack2(du0, du1) : Cvt<__half>::pack2(du0, du1);
              }
              *(reinterpret_cast<int4*>(orow + col0) + i) = og;
              *(reinterpret_cast<int4*>(orow + ff + col0) + i) = ou4;
            }
          }
        }
      } else if (p.scatter_world > 0) {
        // K10, prefill-sized steps: the tile's rows go to the receive slots of the ranks that own them. A thread
        // holds one ROW of the accumulator, so storing from registers would put 32 different rows into every
        // store instruction (16-byte pieces on the link: measured 3x slower than pulling). Each warp therefore
        // turns its 32 x 64 block through 4 KB of shared memory (16-byte chunks XOR-swizzled by row: conflict-
        // free both ways) and writes 4 rows x 128 contiguous bytes per instruction.
        const bool bf = p.out_dtype == DLTI_BF16;
        const bool rok = row < p.M;
        const int owner = rok ? row / p.scatter_per : 0;
        const unsigned long long drow = reinterpret_cast<unsigned long long>(
            reinterpret_cast<uint16_t*>(s_scatter[owner]) +
            ((long long)p.scatter_rank * p.scatter_per + (row - owner * p.scatter_per)) * p.ldc);
        const uint32_t st_base = smem_base + STAGES * STAGE_BYTES + 256 + (uint32_t)quarter * 4096u;   // behind the barriers
#pragma unroll 1
        for (int c = 0; c < ncols / 64; ++c) {
          const int col0 = n0 + c * 64;
          if (col0 >= p.N) break;
          uint32_t v0[32], v1[32];
          tmem_ld_32x32b_x32(t_row + c * 64, v0);
          tmem_ld_32x32b_x32(t_row + c * 64 + 32, v1);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint32_t u0[4], u1[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float a0 = __uint_as_float(v0[8 * i + 2 * j]) * p.alpha, a1 = __uint_as_float(v0[8 * i + 2 * j + 1]) * p.alpha;
              const float b0 = __uint_as_float(v1[8 * i + 2 * j]) * p.alpha, b1 = __uint_as_float(v1[8 * i + 2 * j + 1]) * p.alpha;
              u0[j] = bf ? Cvt<__nv_bfloat16>::pack2(a0, a1) : Cvt<__half>::pack2(a0, a1);
              u1[j] = bf ? Cvt<__nv_bfloat16>::pack2(b0, b1) : Cvt<__half>::pack2(b0, b1);
            }
            const uint32_t a0 = st_base + (uint32_t)lane * 128u + (uint32_t)((i ^ (lane & 7)) << 4);
            const uint32_t a1 = st_base + (uint32_t)lane * 128u + (uint32_t)(((i + 4) ^ (lane & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a0), "r"(u0[0]), "r"(u0[1]), "r"(u0[2]), "r"(u0[3]) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a1), "r"(u1[0]), "r"(u1[1]), "r"(u1[2]), "r"(u1[3]) : "memory");
          }
          __syncwarp();
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + (lane >> 3), ch = lane & 7;
            const unsigned long long dr = __shfl_sync(0xffffffffu, drow, r);
            const int ok = __shfl_sync(0xffffffffu, (int)rok, r);
            int4 val;
            const uint32_t sa = st_base + (uint32_t)r * 128u + (uint32_t)((ch ^ (r & 7)) << 4);
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(val.x), "=r"(val.y), "=r"(val.z), "=r"(val.w) : "r"(sa) : "memory");
            if (ok) st_v4(reinterpret_cast<uint16_t*>(dr) + col0 + ch * 8, val);
          }
          __syncwarp();
        }
        __threadfence_system();
      } else
#pragma unroll 1
      for (int c = 0; c < ncols / 32; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= p.N) break;
        uint32_t v[32];
        tmem_ld_32x32b_x32(t_row + c * 32, v);
        tmem_wait_ld();
        if (helper) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            ws[(c * 8 + j) * BM] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                               __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
          continue;
        }
        if (fixup) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 q = __ldcg(ws + (c * 8 + j) * BM);
            v[4 * j] = __float_as_uint(__uint_as_float(v[4 * j]) + q.x);
            v[4 * j + 1] = __float_as_uint(__uint_as_float(v[4 * j + 1]) + q.y);
            v[4 * j + 2] = __float_as_uint(__uint_as_float(v[4 * j + 2]) + q.z);
            v[4 * j + 3] = __float_as_uint(__uint_as_float(v[4 * j + 3]) + q.w);
          }
        }
        if (row < p.M) {
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.alpha;
          const bool full = (col0 + 32 <= p.N) && p.vec_ok;
          if (p.out_dtype == DLTI_FP32) {
            float* crow = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + col0;
            if (full) {
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
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                int4 o;
                uint32_t* ou = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  ou[j] = bf ? Cvt<__nv_bfloat16>::pack2(f[8 * i + 2 * j], f[8 * i + 2 * j + 1])
                             : Cvt<__half>::pack2(f[8 * i + 2 * j], f[8 * i + 2 * j + 1]);
                *(reinterpret_cast<int4*>(crow) + i) = o;
              }
            } else {
              for (int i = 0; i < 32; ++i) {
                if (col0 + i < p.N) {
                  float x = f[i];
                  if (rrow)
                    x += bf ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(rrow + i))
                            : __half2float(*reinterpret_cast<const __half*>(rrow + i));
                  if (bf) *reinterpret_cast<__nv_bfloat16*>(crow + i) = __float2bfloat16_rn(x);
                  else    *reinterpret_cast<__half*>(crow + i) = __float2half_rn(x);
                }
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(acc == 0 ? leader_tempty0 : leader_tempty1);
      if (helper) {
        // publish this warp's part of the partial tile (8 warps per tile: 4 per CTA)
        __threadfence();
        __syncwarp();
        if (lane == 0) red_release_gpu_add_u32(p.sk_flags + my_range, 1u);
      } else if (fixup) {
        // last of the 8 consumer warps re-arms the slot for the next launch on this stream
        __syncwarp();
        if (lane == 0) {
          const unsigned int done = atomicAdd(p.sk_flags + ncl + my_range + 1, 1u);
          if (done == 7u) {
            p.sk_flags[ncl + my_range + 1] = 0u;
            st_release_gpu_u32(p.sk_flags + my_range + 1, 0u);
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  cluster_sync_all();      // the peer's tensor core may still be reading this CTA's shared memory until here
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, TMEM_COLS);
  }
}

// The gather role on its own (no GEMM to ride on: flushing what is left of a matrix, bandwidth probes).
__global__ void __launch_bounds__(GATHER_THREADS) gather_job_kernel(const GatherJob gj) {   // one warp per CTA
  __shared__ const void* s_peers[16];
#pragma unroll
  for (int i = 0; i < 16; ++i)
    if (threadIdx.x == i) s_peers[i] = gj.peer[i];
  __syncthreads();
  const int gt = threadIdx.x;
  const long long chunk_bytes = 4 * gj.row_bytes, unit_bytes = 64 * gj.row_bytes;
  const int total_chunks = gj.j_count * GATHER_CH;
  for (int c = blockIdx.x; c < total_chunks; c += gridDim.x) {
    const int j = gj.j_begin + c / GATHER_CH, part = c % GATHER_CH;
    int u = gj.rot + j;
    u -= (u / gj.n_units) * gj.n_units;
    const int owner = u / gj.units_per_rank;
    const uint8_t* src = reinterpret_cast<const uint8_t*>(s_peers[owner]) +
                         (long long)(u - owner * gj.units_per_rank) * unit_bytes + part * chunk_bytes;
    uint8_t* dst = gj.staging + (long long)u * unit_bytes + part * chunk_bytes;
    const long long nvec = chunk_bytes >> 4;
    long long i = gt;
    for (; i + (GATHER_MLP - 1) * GATHER_THREADS < nvec; i += GATHER_MLP * GATHER_THREADS) {
      int4 v[GATHER_MLP];
#pragma unroll
      for (int q = 0; q < GATHER_MLP; ++q) v[q] = ld_nc_v4(src + ((i + q * GATHER_THREADS) << 4));
#pragma unroll
      for (int q = 0; q < GATHER_MLP; ++q) st_v4(dst + ((i + q * GATHER_THREADS) << 4), v[q]);
    }
    for (; i < nvec; i += GATHER_THREADS) st_v4(dst + (i << 4), ld_nc_v4(src + (i << 4)));
    __threadfence();
    __syncwarp();
    if (gt == 0) red_release_gpu_add_u32(gj.counters + u, 1u);
  }
}

// Attachments for the NEXT CTA-pair launch made from this host thread (dlti_gemm2_attach); cleared by it.
thread_local GatherJob t_job = {};
thread_local WaitSpec t_wait = {};
struct ScatterSpec {
  void* peer[16];
  int world, rank, per;
};
thread_local ScatterSpec t_scatter = {};
thread_local int t_m_rot = 0;

template <bool B_MN>
int launch2(const Gemm2Params& p_in, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& a2,
            const CUtensorMap& b2, const CUtensorMap& bh, const CUtensorMap& b2h, int grid, cudaStream_t stream) {
  const GatherJob job = t_job;
  const WaitSpec wait = t_wait;
  const ScatterSpec sc = t_scatter;
  const int m_rot = t_m_rot;
  t_job = GatherJob{};
  t_wait = WaitSpec{};
  t_scatter = ScatterSpec{};
  t_m_rot = 0;
  Gemm2Params p = p_in;
  if (p.sk_tiles > 0 && (job.j_count > 0 || wait.counters || sc.world > 0)) return DLTI_ERR_UNSUPPORTED;
  if (m_rot > 0) p.m_rot = m_rot % p.nmb2;
  if (sc.world > 0) {
    // plain 16-bit epilogue only (no residual / SwiGLU / RoPE), whole 64-column groups, 16-byte aligned rows
    if (p.out_dtype == DLTI_FP32 || p.R || p.swiglu_ff > 0 || p.swiglu_bwd_ff > 0 || p.rope_pos || p.N % 64 ||
        (p.ldc * 2) % 16)
      return DLTI_ERR_UNSUPPORTED;
    for (int i = 0; i < 16; ++i) p.scatter_peer[i] = sc.peer[i];
    p.scatter_world = sc.world; p.scatter_rank = sc.rank; p.scatter_per = sc.per;
  }
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    DLTI_CUDA_CHECK(cudaFuncSetAttribute(gemm2_kernel<B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         SMEM_BYTES + SCATTER_STAGE_BYTES));
    configured[dev & 63] = true;
  }
  gemm2_kernel<B_MN><<<grid, NUM_THREADS + (job.j_count > 0 ? GATHER_THREADS : 0),
                       SMEM_BYTES + (sc.world > 0 ? SCATTER_STAGE_BYTES : 0), stream>>>(
      p, job, wait, a, b, a2, b2, bh, b2h);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

// Stream-K workspace: one per (device, stream) so GEMMs on different streams never share partial slots.
struct SkWorkspace {
  int dev;
  cudaStream_t stream;
  float4* ws;
  unsigned int* flags;
};
constexpr int MAX_SK_WS = 16;
constexpr int MAX_RANGES = 128;
SkWorkspace g_sk_ws[MAX_SK_WS];
int g_sk_ws_count = 0;

const SkWorkspace* sk_workspace(cudaStream_t stream) {
  int dev = 0;
  cudaGetDevice(&dev);
  for (int i = 0; i < g_sk_ws_count; ++i)
    if (g_sk_ws[i].dev == dev && g_sk_ws[i].stream == stream) return &g_sk_ws[i];
  if (g_sk_ws_count == MAX_SK_WS) return nullptr;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) return nullptr;
  SkWorkspace w{dev, stream, nullptr, nullptr};
  const size_t ws_bytes = (size_t)MAX_RANGES * 2 * 64 * BM * sizeof(float4);   // 256 KB per range
  if (cudaMalloc(&w.ws, ws_bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  if (cudaMalloc(&w.flags, 2 * MAX_RANGES * sizeof(unsigned int)) != cudaSuccess) {
    cudaGetLastError();
    cudaFree(w.ws);
    return nullptr;
  }
  cudaMemset(w.flags, 0, 2 * MAX_RANGES * sizeof(unsigned int));
  g_sk_ws[g_sk_ws_count] = w;
  return &g_sk_ws[g_sk_ws_count++];
}

// Ragged last wave (see Gemm2Params::split_from): with T tiles on C clusters, the T % C tiles of the last wave
// are cut in two when that lets the wave finish in half the time (2 * (T % C) <= C). Llama-2-7B at M = 4096:
// the [4096 x 4096] outputs (o_proj, down_proj, every input gradient of qkv / gate_up) are 256 tiles on 74
// clusters = 3.46 waves -> 3.5 instead of 4. DLTI_GEMM_TAILSPLIT=0 switches it off (A/B).
// MN-major B as one 3D TMA box per stage instead of two 2D boxes: measured SLOWER on B200 (qkv input gradient,
// sustained: 1,068 vs 1,122 TFLOP/s), so it is opt-in (DLTI_GEMM_MN3D=1).
bool mn3d_enabled() {
  static const bool on = []() { const char* e = getenv("DLTI_GEMM_MN3D"); return e && e[0] == '1'; }();
  return on;
}

void set_tail_split(Gemm2Params& p, int grid, int dtype, bool b_mn) {
  static const bool on = []() { const char* e = getenv("DLTI_GEMM_TAILSPLIT"); return !(e && e[0] == '0'); }();
  const int total = p.nmb2 * p.nnb, ncl = grid / 2;
  p.split_from = total;
  p.idesc_half = umma_idesc_f16(dtype, 2 * BM, BN / 2, false, b_mn);
  if (!on || p.sk_tiles > 0 || ncl < 1 || total <= ncl) return;
  const int rem = total % ncl;
  if (rem > 0 && 2 * rem <= ncl) p.split_from = total - rem;
}

}  // namespace

// K1, CTA-pair form: attach a gather job and / or a wait spec to the NEXT dlti_gemm_2cta / dlti_gemm_rope_2cta /
// dlti_gemm_swiglu_2cta / dlti_gemm_swiglu_bwd_2cta launched from this host thread (one launch consumes it).
//   job  (j_count > 0): pull units  (rot_unit + j) % n_units,  j in [j_begin, j_begin + j_count),  of the row-
//         sharded matrix [world * rows_per_rank, row_bytes] from peer_shards[owner] into `staging` and publish
//         counters[u] += 16 chunks; a unit is complete once counters[u] >= epoch * 16. Every unit must be
//         covered exactly once per epoch, by this or other launches.
//   wait (wait_counters != null): the launch's B operand is a staging buffer whose 64-row unit u is valid once
//         wait_counters[u] >= wait_epoch * 16; the TMA producer waits tile by tile.
DLTI_API int dlti_gemm2_attach(const void* const* peer_shards, void* staging, unsigned int* counters,
                               long long row_bytes, int rows_per_rank, int world, int rot_unit, int j_begin,
                               int j_count, unsigned int epoch, const unsigned int* wait_counters, int wait_units,
                               unsigned int wait_epoch, int* err_flag) {
  GatherJob j{};
  WaitSpec w{};
  if (j_count > 0) {
    if (world < 1 || world > 16 || rows_per_rank % 64 || !peer_shards || !staging || !counters || (row_bytes * 4) % 16)
      return DLTI_ERR_BAD_ARG;
    for (int i = 0; i < world; ++i) j.peer[i] = peer_shards[i];
    j.staging = reinterpret_cast<uint8_t*>(staging);
    j.counters = counters;
    j.err = err_flag;
    j.row_bytes = row_bytes;
    j.target = epoch * (unsigned)GATHER_CH;
    j.units_per_rank = rows_per_rank / 64;
    j.n_units = j.units_per_rank * world;
    if (rot_unit < 0 || rot_unit >= j.n_units || j_begin < 0 || j_begin + j_count > j.n_units) return DLTI_ERR_BAD_ARG;
    j.rot = rot_unit;
    j.j_begin = j_begin;
    j.j_count = j_count;
  }
  if (wait_counters) {
    w.counters = wait_counters;
    w.err = err_flag;
    w.target = wait_epoch * (unsigned)GATHER_CH;
    w.n_units = wait_units;
  }
  t_job = j;
  t_wait = w;
  return 0;
}

// K11: the wait spec attached with dlti_gemm2_attach refers to the A operand (activation rows pulled from their
// owner ranks) instead of B, and the M blocks are visited starting at block m_rot (own rows first).
DLTI_API int dlti_gemm2_attach_a_side(int m_rot) {
  if (m_rot < 0) return DLTI_ERR_BAD_ARG;
  t_wait.on_a = 1;
  t_m_rot = m_rot;
  return 0;
}

// K10 for prefill-sized steps: the NEXT CTA-pair launch stores output row r into the receive slot that rank
// r / rows_per_owner keeps for this rank: slot_peers[owner] + (rank * rows_per_owner + r % rows_per_owner) * ldc + col
// (16-bit output, no residual / fused epilogue, N % 64 == 0). ldc is the launch's own ldc argument.
DLTI_API int dlti_gemm2_attach_scatter(const void* const* slot_peers, int world, int rank, int rows_per_owner) {
  if (!slot_peers || world < 1 || world > 16 || rank < 0 || rank >= world || rows_per_owner < 1) return DLTI_ERR_BAD_ARG;
  ScatterSpec sc{};
  for (int i = 0; i < world; ++i) sc.peer[i] = const_cast<void*>(slot_peers[i]);
  sc.world = world; sc.rank = rank; sc.per = rows_per_owner;
  t_scatter = sc;
  return 0;
}

// Run the job attached with dlti_gemm2_attach as a kernel of its own (ctas: 0 = two per SM).
DLTI_API int dlti_gather_job_run(int ctas, cudaStream_t stream) {
  const GatherJob job = t_job;
  t_job = GatherJob{};
  t_wait = WaitSpec{};
  if (job.j_count <= 0) return 0;
  int grid = ctas > 0 ? ctas : 16 * dlti_num_sms();
  if (grid > job.j_count * GATHER_CH) grid = job.j_count * GATHER_CH;
  gather_job_kernel<<<grid, GATHER_THREADS, 0, stream>>>(job);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

// Same contract as dlti_gemm restricted to: A K-major, no split-K. Returns DLTI_ERR_UNSUPPORTED for
// configurations it does not cover (the caller then uses the 1-CTA kernel). stream_k: -1 = automatic
// (ragged last wave only, and only when DLTI_GEMM_STREAMK=1), 0 = never, 1 = whenever the tile count is
// not a multiple of the cluster count.
DLTI_API int dlti_gemm_2cta(const void* A, const void* B, void* C, const void* R, int M, int N, int K,
                            long long lda, long long ldb, long long ldc, long long ldr, int b_mn,
                            const void* A2, const void* B2, int K2, long long lda2, long long ldb2,
                            int in_dtype, int out_dtype, float alpha, int accumulate, int stream_k,
                            cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return DLTI_ERR_BAD_ARG;
  if (in_dtype != DLTI_BF16 && in_dtype != DLTI_FP16) return DLTI_ERR_UNSUPPORTED;
  if (K2 % BK) return DLTI_ERR_BAD_ARG;
  if (out_dtype != DLTI_FP32 && accumulate) return DLTI_ERR_BAD_ARG;
  Gemm2Params p{};
  p.M = M; p.N = N;
  p.nmb2 = (M + 2 * BM - 1) / (2 * BM);
  p.nnb = (N + BN - 1) / BN;
  p.nkb_main = (K + BK - 1) / BK;
  p.nkb_ext = (A2 && B2) ? K2 / BK : 0;
  p.C = C; p.R = R; p.ldc = ldc; p.ldr = ldr;
  p.out_dtype = out_dtype; p.in_dtype = in_dtype; p.accumulate = accumulate;
  const int ob = out_dtype == DLTI_FP32 ? 4 : 2;
  p.vec_ok = ((reinterpret_cast<uintptr_t>(C) & 15) == 0) && ((ldc * ob) % 16 == 0) &&
             (!R || (((reinterpret_cast<uintptr_t>(R) & 15) == 0) && ((ldr * 2) % 16 == 0)));
  p.alpha = alpha;
  p.idesc = umma_idesc_f16(in_dtype, 2 * BM, BN, false, b_mn != 0);
  CUtensorMap ta, tb, ta2, tb2, tbh, tb2h;
  int rc;
  if ((rc = dlti_get_tmap(&ta, A, K, M, lda, 64, BM))) return rc;
  rc = b_mn ? dlti_get_tmap(&tb, B, N, K, ldb, 64, 64) : dlti_get_tmap(&tb, B, K, N, ldb, 64, BN / 2);
  if (rc) return rc;
  tbh = tb;                  // MN-major: a half tile is one of the two 64x64 boxes
  if (!b_mn && (rc = dlti_get_tmap(&tbh, B, K, N, ldb, 64, BN / 4))) return rc;
  if (b_mn && N % 64 == 0 && mn3d_enabled() && dlti_get_tmap_mn3d(&tb, B, N, K, ldb, BK, (BN / 2) / 64) == 0) p.b3d = 1;
  ta2 = ta; tb2 = tb; tb2h = tbh;
  if (p.nkb_ext > 0) {
    if ((rc = dlti_get_tmap(&ta2, A2, K2, M, lda2, 64, BM))) return rc;
    rc = b_mn ? dlti_get_tmap(&tb2, B2, N, K2, ldb2, 64, 64) : dlti_get_tmap(&tb2, B2, K2, N, ldb2, 64, BN / 2);
    if (rc) return rc;
    tb2h = tb2;
    if (!b_mn && (rc = dlti_get_tmap(&tb2h, B2, K2, N, ldb2, 64, BN / 4))) return rc;
  }
  int grid = dlti_num_sms() & ~1;
  const int total = p.nmb2 * p.nnb;
  if (2 * total < grid) grid = 2 * total;
  // Stream-K for the ragged last wave: when total % clusters leaves the last wave under-filled, the
  // last (1 + remainder) waves' tiles are cut into equal (tile, k-block) ranges, one per cluster.
  const int ncl = grid / 2;
  const int nkb = p.nkb_main + p.nkb_ext;
  const int rem = total % ncl;
  // Measured (profiles/streamk_ab_r1.md): on the power-capped B200 the automatic mode LOSES end to end
  // (774 vs 755 ms/step): the idle SMs of a ragged wave let the clocks rise, and filling them costs
  // more in frequency than the schedule gains. So automatic stream-K is opt-in (DLTI_GEMM_STREAMK=1).
  static const bool sk_auto = []() { const char* e = getenv("DLTI_GEMM_STREAMK"); return e && e[0] == '1'; }();
  if ((stream_k > 0 || (stream_k < 0 && sk_auto)) && rem != 0 && total > ncl && ncl <= MAX_RANGES) {
    const int waves = (total + ncl - 1) / ncl;
    const double eff = (double)total / ((double)ncl * waves);
    if (eff < 0.97 || stream_k > 0) {
      if (const SkWorkspace* w = sk_workspace(stream)) {
        p.sk_tiles = rem + ncl;
        p.sk_iters_per_range = ((long long)p.sk_tiles * nkb + ncl - 1) / ncl;
        p.sk_ws = w->ws;
        p.sk_flags = w->flags;
      }
    }
  }
  set_tail_split(p, grid, in_dtype, b_mn != 0);
  return b_mn ? launch2<true>(p, ta, tb, ta2, tb2, tbh, tb2h, grid, stream)
              : launch2<false>(p, ta, tb, ta2, tb2, tbh, tb2h, grid, stream);
}

// h[M, ff] = silu(A · Wg^T) * (A · Wu^T) with W = [gate rows (ff) ; up rows (ff)] stored [2*ff, K]
// (the Llama gate_up projection + SwiGLU in one kernel; the [M, 2*ff] intermediate is never written).
// Requires ff % 128 == 0 and 16-byte aligned output rows; otherwise DLTI_ERR_UNSUPPORTED.
DLTI_API int dlti_gemm_swiglu_2cta(const void* A, const void* W, void* Hout, int M, int ff, int K, long long lda,
                                   long long ldw, long long ldh, int dtype, cudaStream_t stream) {
  if (M <= 0 || ff <= 0 || K <= 0) return DLTI_ERR_BAD_ARG;
  if (dtype != DLTI_BF16 && dtype != DLTI_FP16) return DLTI_ERR_UNSUPPORTED;
  if (ff % (BN / 2) || (reinterpret_cast<uintptr_t>(Hout) & 15) || (ldh * 2) % 16) return DLTI_ERR_UNSUPPORTED;
  Gemm2Params p{};
  p.M = M; p.N = 2 * ff;
  p.nmb2 = (M + 2 * BM - 1) / (2 * BM);
  p.nnb = ff / (BN / 2);
  p.nkb_main = (K + BK - 1) / BK;
  p.nkb_ext = 0;
  p.C = Hout; p.R = nullptr; p.ldc = ldh; p.ldr = 0;
  p.out_dtype = dtype; p.in_dtype = dtype; p.accumulate = 0; p.vec_ok = 1;
  p.alpha = 1.f;
  p.idesc = umma_idesc_f16(dtype, 2 * BM, BN, false, false);
  p.swiglu_ff = ff;
  CUtensorMap ta, tb, tbh;
  int rc;
  if ((rc = dlti_get_tmap(&ta, A, K, M, lda, 64, BM))) return rc;
  if ((rc = dlti_get_tmap(&tb, W, K, 2 * (long long)ff, ldw, 64, BN / 2))) return rc;
  if ((rc = dlti_get_tmap(&tbh, W, K, 2 * (long long)ff, ldw, 64, BN / 4))) return rc;
  int grid = dlti_num_sms() & ~1;
  const int total = p.nmb2 * p.nnb;
  if (2 * total < grid) grid = 2 * total;
  set_tail_split(p, grid, dtype, false);
  return launch2<false>(p, ta, tb, ta, tb, tbh, tbh, grid, stream);
}

// d(gate|up)[M, 2*ff] = swiglu_backward(dh, gate|up) with dh = dY[M, K] · W_down ([K, ff], i.e. the
// down projection weight [d_model, ff] used as the MN-major B operand): the input-gradient GEMM of the
// down projection with the SwiGLU backward applied in its epilogue; dh is never written.
// Requires ff % 32 == 0 and 16-byte aligned rows; otherwise DLTI_ERR_UNSUPPORTED.
DLTI_API int dlti_gemm_swiglu_bwd_2cta(const void* dY, const void* Wdown, const void* GU, void* dGU, int M, int ff,
                                       int K, long long ldy, long long ldw, long long ldgu, long long lddgu,
                                       int dtype, cudaStream_t stream) {
  if (M <= 0 || ff <= 0 || K <= 0) return DLTI_ERR_BAD_ARG;
  if (dtype != DLTI_BF16 && dtype != DLTI_FP16) return DLTI_ERR_UNSUPPORTED;
  if (ff % 32 || (reinterpret_cast<uintptr_t>(dGU) & 15) || (reinterpret_cast<uintptr_t>(GU) & 15) ||
      (ldgu * 2) % 16 || (lddgu * 2) % 16 || (ff * 2) % 16)
    return DLTI_ERR_UNSUPPORTED;
  Gemm2Params p{};
  p.M = M; p.N = ff;
  p.nmb2 = (M + 2 * BM - 1) / (2 * BM);
  p.nnb = (ff + BN - 1) / BN;
  p.nkb_main = (K + BK - 1) / BK;
  p.C = dGU; p.ldc = lddgu;
  p.out_dtype = dtype; p.in_dtype = dtype; p.vec_ok = 1;
  p.alpha = 1.f;
  p.idesc = umma_idesc_f16(dtype, 2 * BM, BN, false, true);
  p.swiglu_bwd_ff = ff; p.gu = GU; p.ldgu = ldgu;
  CUtensorMap ta, tb;
  int rc;
  if ((rc = dlti_get_tmap(&ta, dY, K, M, ldy, 64, BM))) return rc;
  if ((rc = dlti_get_tmap(&tb, Wdown, ff, K, ldw, 64, 64))) return rc;
  CUtensorMap tb3 = tb;
  if (ff % 64 == 0 && mn3d_enabled() && dlti_get_tmap_mn3d(&tb3, Wdown, ff, K, ldw, BK, (BN / 2) / 64) == 0) p.b3d = 1;
  int grid = dlti_num_sms() & ~1;
  const int total = p.nmb2 * p.nnb;
  if (2 * total < grid) grid = 2 * total;
  set_tail_split(p, grid, dtype, true);
  return launch2<true>(p, ta, tb3, ta, tb, tb, tb, grid, stream);
}

// C[M, N] = rope(alpha * (A · B^T + A2 · B2^T)): the fused QKV projection (B = [N, K], optional LoRA
// K-extension) with the rotary embedding of each row's position applied in the epilogue to the first
// rope_cols output columns (heads of 128, rotate-half convention); cos/sin are fp32 tables [max_pos, 64].
// Requires N and rope_cols to be multiples of 128 and 16-byte aligned output rows.
DLTI_API int dlti_gemm_rope_2cta(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                                 long long ldb, long long ldc, const void* A2, const void* B2, int K2,
                                 long long lda2, long long ldb2, int dtype, float alpha, const int* pos,
                                 const float* cos_t, const float* sin_t, int rope_cols, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || !pos || !cos_t || !sin_t) return DLTI_ERR_BAD_ARG;
  if (dtype != DLTI_BF16 && dtype != DLTI_FP16) return DLTI_ERR_UNSUPPORTED;
  if (N % 128 || rope_cols % 128 || K2 % BK || (reinterpret_cast<uintptr_t>(C) & 15) || (ldc * 2) % 16)
    return DLTI_ERR_UNSUPPORTED;
  Gemm2Params p{};
  p.M = M; p.N = N;
  p.nmb2 = (M + 2 * BM - 1) / (2 * BM);
  p.nnb = (N + BN - 1) / BN;
  p.nkb_main = (K + BK - 1) / BK;
  p.nkb_ext = (A2 && B2) ? K2 / BK : 0;
  p.C = C; p.ldc = ldc;
  p.out_dtype = dtype; p.in_dtype = dtype; p.vec_ok = 1;
  p.alpha = alpha;
  p.idesc = umma_idesc_f16(dtype, 2 * BM, BN, false, false);
  p.rope_pos = pos; p.rope_cos = cos_t; p.rope_sin = sin_t; p.rope_cols = rope_cols;
  CUtensorMap ta, tb, ta2, tb2, tbh, tb2h;
  int rc;
  if ((rc = dlti_get_tmap(&ta, A, K, M, lda, 64, BM))) return rc;
  if ((rc = dlti_get_tmap(&tb, B, K, N, ldb, 64, BN / 2))) return rc;
  if ((rc = dlti_get_tmap(&tbh, B, K, N, ldb, 64, BN / 4))) return rc;
  ta2 = ta; tb2 = tb; tb2h = tbh;
  if (p.nkb_ext > 0) {
    if ((rc = dlti_get_tmap(&ta2, A2, K2, M, lda2, 64, BM))) return rc;
    if ((rc = dlti_get_tmap(&tb2, B2, K2, N, ldb2, 64, BN / 2))) return rc;
    if ((rc = dlti_get_tmap(&tb2h, B2, K2, N, ldb2, 64, BN / 4))) return rc;
  }
  int grid = dlti_num_sms() & ~1;
  const int total = p.nmb2 * p.nnb;
  if (2 * total < grid) grid = 2 * total;
  set_tail_split(p, grid, dtype, false);
  return launch2<false>(p, ta, tb, ta2, tb2, tbh, tb2h, grid, stream);
}
