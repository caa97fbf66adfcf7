// Serving kernels: K13 fused RoPE + paged-KV-cache write, K12 paged-KV decode attention (split-KV +
// combine), K14 sampling (greedy / temperature / top-k / top-p, seeded, sort-free).
//
// KV cache layout: k_cache, v_cache = [num_blocks, block_size, KV_heads, 128] (token-major inside a
// block: one token's head vector is 256 contiguous bytes, so both the prefill write and the decode
// read are fully coalesced). slot = block_id * block_size + offset.
//
// These replace vLLM's _C ops rotary_embedding + reshape_and_cache[_flash], paged_attention_v1/v2
// and its sampler — the reference itself only pins `vllm==0.6.0`
// (reference requirements.txt:18) and ships no serving code.
#include "common.cuh"

namespace {

constexpr int HD = 128;

template <typename T>
__device__ __forceinline__ void unpack8(const int4& v, float (&f)[8]) {
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = Cvt<T>::unpack2(u[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
template <typename T>
__device__ __forceinline__ int4 pack8(const float (&f)[8]) {
  int4 v;
  uint32_t* u = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) u[i] = Cvt<T>::pack2(f[2 * i], f[2 * i + 1]);
  return v;
}

// ---------------------------------------------------------------------------------------------
// K13: for every token row of qkv [T, (H+2KV)*128]: rotate q and k heads in place (rotate-half, HF
// convention) and store the rotated k and the v into the paged cache at slot_mapping[t]
// (slot < 0 => token is not cached, e.g. padding). One thread = 8 "i" of one head.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void rope_cache_kernel(T* __restrict__ qkv, const int* __restrict__ pos,
                                  const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                                  const long long* __restrict__ slots, T* __restrict__ k_cache,
                                  T* __restrict__ v_cache, long long tokens, int H, int KV, long long W) {
  const int heads = H + 2 * KV;              // q heads, k heads (rotated), v heads (copied)
  const int vph = HD / 16;                   // 8 threads per head (each: 8 low + 8 high elements)
  const long long total = tokens * heads * vph;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int vi = (int)(idx % vph);
    const long long th = idx / vph;
    const int h = (int)(th % heads);
    const long long t = th / heads;
    T* base = qkv + t * W + (long long)h * HD + vi * 8;
    int4 lo = *reinterpret_cast<int4*>(base);
    int4 hi = *reinterpret_cast<int4*>(base + HD / 2);
    const long long slot = slots ? slots[t] : -1;
    if (h < H + KV) {
      const int p = pos[t];
      float a[8], b[8], o1[8], o2[8];
      unpack8<T>(lo, a);
      unpack8<T>(hi, b);
      const float* cp = cos_t + (long long)p * (HD / 2) + vi * 8;
      const float* sp = sin_t + (long long)p * (HD / 2) + vi * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float c = Cvt<T>::to_f(Cvt<T>::from_f(cp[j]));
        const float s = Cvt<T>::to_f(Cvt<T>::from_f(sp[j]));
        o1[j] = a[j] * c - b[j] * s;
        o2[j] = b[j] * c + a[j] * s;
      }
      lo = pack8<T>(o1);
      hi = pack8<T>(o2);
      *reinterpret_cast<int4*>(base) = lo;
      *reinterpret_cast<int4*>(base + HD / 2) = hi;
      if (h >= H && slot >= 0) {
        T* dst = k_cache + (slot * KV + (h - H)) * HD + vi * 8;
        *reinterpret_cast<int4*>(dst) = lo;
        *reinterpret_cast<int4*>(dst + HD / 2) = hi;
      }
    } else if (slot >= 0) {
      T* dst = v_cache + (slot * KV + (h - H - KV)) * HD + vi * 8;
      *reinterpret_cast<int4*>(dst) = lo;
      *reinterpret_cast<int4*>(dst + HD / 2) = hi;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// K12: decode attention over the paged cache. grid = (splits, H, B), 128 threads (4 warps).
// A warp processes 8 tokens per iteration: lane = (token_in_group = lane/4, dim chunk = lane%4),
// each lane owns 32 dims of its token's K and V (64 B loads, 4 x int4). Scores are reduced over the
// 4 chunk-lanes with 2 shuffles; the running softmax is per token-group lane set; the four warps
// and eight token groups are merged at the end in shared memory. Partial (m, l, o) per split go to
// a workspace and a combine kernel merges the splits.
// ---------------------------------------------------------------------------------------------
#ifndef DLTI_PAGED_MIN_CTAS
#define DLTI_PAGED_MIN_CTAS 4      // resident CTAs per SM the register allocation must allow: 128 regs (56 B of spills) instead of 163; batch 128 decode step 10.90 -> 10.42 ms, batch 64 unchanged
#endif
template <typename T>
__global__ void __launch_bounds__(128, DLTI_PAGED_MIN_CTAS) paged_decode_kernel(
    const T* __restrict__ q, long long q_stride, const T* __restrict__ k_cache,
    const T* __restrict__ v_cache, const int* __restrict__ block_tables, int max_blocks,
    const int* __restrict__ ctx_lens, int block_size, int H, int KV, float scale_log2,
    float* __restrict__ part_o, float* __restrict__ part_ml, int splits, T* __restrict__ out,
    long long out_stride) {
  griddep_wait();            // q and this step's K/V rows come from the kernel before (see dlti_launch_pdl)
  griddep_launch();
  const int split = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int ctx = ctx_lens[b];
  const int kvh = h / (H / KV);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tg = lane >> 2, ch = lane & 3;            // token in group, dim chunk
  // token range of this split (multiples of 32 tokens = one CTA iteration)
  const int iters_total = (ctx + 31) / 32;
  const int per = (iters_total + splits - 1) / splits;
  const int it0 = split * per, it1 = min(iters_total, it0 + per);

  float qf[32];
  {
    // lane `ch` owns dims {i*32 + ch*8 .. +8 : i = 0..3}: for each i the 4 lanes of a token read 64
    // CONTIGUOUS bytes, so every load instruction uses whole 32-byte sectors
    const T* qp = q + (long long)b * q_stride + (long long)h * HD + ch * 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float f[8];
      unpack8<T>(*reinterpret_cast<const int4*>(qp + i * 32), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) qf[i * 8 + j] = f[j] * scale_log2;
    }
  }
  float m = -INFINITY, l = 0.f;
  float o[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) o[i] = 0.f;
  // This split's slice of the block table goes to shared memory first: the loop's K/V addresses then depend
  // on a shared-memory read instead of a second global round trip per iteration.
  constexpr int BT_SMEM = 512;
  __shared__ int s_bt[BT_SMEM];
  const int* btg = block_tables + (long long)b * max_blocks;
  const int blk0 = (it0 * 32) / block_size;
  const int blk1 = min((min(it1 * 32, ctx) + block_size - 1) / block_size, max_blocks);
  const bool bt_staged = blk1 - blk0 <= BT_SMEM;
  if (bt_staged) {
    for (int i = threadIdx.x; i < blk1 - blk0; i += 128) s_bt[i] = btg[blk0 + i];
    __syncthreads();
  }
  const int* bt = bt_staged ? s_bt - blk0 : btg;

  // Software pipeline: the K/V rows of iteration it+1 are requested before iteration it is reduced, so
  // every thread always has 128 bytes in flight (the loop was latency-bound at 38 % of HBM bandwidth).
  int4 kraw[4], vraw[4];
  bool valid = false;
  auto fetch = [&](int it) {
    const int tok = it * 32 + warp * 8 + tg;
    valid = it < it1 && tok < ctx;
    if (valid) {
      const long long slot = (long long)bt[tok / block_size] * block_size + (tok % block_size);
      const T* kp = k_cache + (slot * KV + kvh) * HD + ch * 8;
      const T* vp = v_cache + (slot * KV + kvh) * HD + ch * 8;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        kraw[i] = ld_nc_v4(kp + i * 32);
        vraw[i] = ld_nc_v4(vp + i * 32);
      }
    }
  };
  fetch(it0);
  for (int it = it0; it < it1; ++it) {
    const bool cur_valid = valid;
    int4 kc[4], vc[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { kc[i] = kraw[i]; vc[i] = vraw[i]; }
    fetch(it + 1);
    float s = -INFINITY;
    if (cur_valid) {
      float acc = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float f[8];
        unpack8<T>(kc[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc += f[j] * qf[i * 8 + j];
      }
      s = acc;
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);          // -inf + -inf stays -inf for invalid tokens
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (cur_valid) {
      const float nm = fmaxf(m, s);
      const float alpha = exp2f(m - nm);
      const float p = exp2f(s - nm);
      l = l * alpha + p;
      m = nm;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float f[8];
        unpack8<T>(vc[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[i * 8 + j] = o[i * 8 + j] * alpha + p * f[j];
      }
    }
  }
  // merge the 8 token groups of the warp (lanes with equal ch), then the 4 warps
#pragma unroll
  for (int off = 4; off < 32; off <<= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, off);
    const float l2 = __shfl_xor_sync(0xffffffffu, l, off);
    const float nm = fmaxf(m, m2);
    const float a1 = (m == -INFINITY) ? 0.f : exp2f(m - nm);
    const float a2 = (m2 == -INFINITY) ? 0.f : exp2f(m2 - nm);
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float o2 = __shfl_xor_sync(0xffffffffu, o[i], off);
      o[i] = o[i] * a1 + o2 * a2;
    }
    l = l * a1 + l2 * a2;
    m = nm;
  }
  __shared__ float sm_o[4][HD];
  __shared__ float sm_m[4], sm_l[4];
  if (tg == 0) {
#pragma unroll
    for (int i = 0; i < 32; ++i) sm_o[warp][(i >> 3) * 32 + ch * 8 + (i & 7)] = o[i];
    if (ch == 0) { sm_m[warp] = m; sm_l[warp] = l; }
  }
  __syncthreads();
  {
    const int d = threadIdx.x;                        // 128 threads = 128 dims
    float gm = fmaxf(fmaxf(sm_m[0], sm_m[1]), fmaxf(sm_m[2], sm_m[3]));
    float acc = 0.f, gl = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float a = (sm_m[w] == -INFINITY) ? 0.f : exp2f(sm_m[w] - gm);
      acc += sm_o[w][d] * a;
      gl += sm_l[w] * a;
    }
    if (splits == 1) {                                // nothing to combine: write the normalised row
      out[(long long)b * out_stride + (long long)h * HD + d] = Cvt<T>::from_f(gl > 0.f ? acc / gl : 0.f);
    } else {
      const long long pidx = ((long long)b * H + h) * splits + split;
      part_o[pidx * HD + d] = acc;
      if (d == 0) { part_ml[pidx * 2] = gm; part_ml[pidx * 2 + 1] = gl; }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(128) paged_combine_kernel(const float* __restrict__ part_o,
                                                           const float* __restrict__ part_ml, int splits,
                                                           T* __restrict__ out, long long out_stride, int H) {
  griddep_wait();
  griddep_launch();
  const int h = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
  const long long base = ((long long)b * H + h) * splits;
  float gm = -INFINITY;
  for (int s = 0; s < splits; ++s) gm = fmaxf(gm, part_ml[(base + s) * 2]);
  float acc = 0.f, gl = 0.f;
  for (int s = 0; s < splits; ++s) {
    const float ms = part_ml[(base + s) * 2];
    const float a = (ms == -INFINITY) ? 0.f : exp2f(ms - gm);
    acc += part_o[(base + s) * HD + d] * a;
    gl += part_ml[(base + s) * 2 + 1] * a;
  }
  out[(long long)b * out_stride + (long long)h * HD + d] = Cvt<T>::from_f(gl > 0.f ? acc / gl : 0.f);
}

// ---------------------------------------------------------------------------------------------
// K14: sampling. One CTA (256 threads) per row of logits [B, V] (16-bit or fp32, row pitch ld).
//   temperature <= 0  => greedy argmax
//   else p = softmax(logits / temperature); keep the top_k most likely (0 = all) and the smallest
//   nucleus with mass >= top_p (1 = all), renormalise, draw with a Philox uniform(seed, step, row).
// Sort-free: the top-k / top-p cut is a probability threshold found by bisection (24 rounds over the
// row held in shared memory as exp(x - max)); the draw is an inverse-CDF walk in index order.
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float load_logit(const void* p, long long i) {
  return Cvt<T>::to_f(reinterpret_cast<const T*>(p)[i]);
}

__device__ __forceinline__ float block_reduce_sum256(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = threadIdx.x < 8 ? red[threadIdx.x] : 0.f;
  t = warp_sum(t);
  t = __shfl_sync(0xffffffffu, t, 0);
  __syncthreads();
  if (threadIdx.x == 0) red[8] = t;
  __syncthreads();
  return red[8];
}

template <typename T>
__global__ void __launch_bounds__(256) sample_kernel(const void* __restrict__ logits, long long ld, int V,
                                                    const float* __restrict__ temperature,
                                                    const int* __restrict__ top_k,
                                                    const float* __restrict__ top_p,
                                                    const unsigned long long* __restrict__ seeds,
                                                    unsigned long long step, long long* __restrict__ out,
                                                    float* __restrict__ out_logprob) {
  extern __shared__ float probs[];            // V floats
  __shared__ float red[16];
  __shared__ int s_arg;
  __shared__ float s_argv;
  const int row = blockIdx.x;
  const void* lr = reinterpret_cast<const uint8_t*>(logits) + (long long)row * ld * sizeof(T);
  const float temp = temperature ? temperature[row] : 0.f;
  // pass 1: max + argmax
  float mx = -INFINITY;
  int am = 0;
  for (int i = threadIdx.x; i < V; i += 256) {
    const float x = load_logit<T>(lr, i);
    probs[i] = x;
    if (x > mx) { mx = x; am = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, mx, o);
    const int a2 = __shfl_xor_sync(0xffffffffu, am, o);
    if (m2 > mx || (m2 == mx && a2 < am)) { mx = m2; am = a2; }
  }
  __shared__ float wm[8];
  __shared__ int wa[8];
  if ((threadIdx.x & 31) == 0) { wm[threadIdx.x >> 5] = mx; wa[threadIdx.x >> 5] = am; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float bm = wm[0]; int ba = wa[0];
    for (int w = 1; w < 8; ++w) if (wm[w] > bm || (wm[w] == bm && wa[w] < ba)) { bm = wm[w]; ba = wa[w]; }
    s_arg = ba; s_argv = bm;
  }
  __syncthreads();
  const float gmax = s_argv;
  // softmax denominator at temperature 1 (for the reported logprob) and at the sampling temperature
  const float invt = temp > 0.f ? 1.f / temp : 1.f;
  float z1 = 0.f, zt = 0.f;
  for (int i = threadIdx.x; i < V; i += 256) {
    const float x = probs[i] - gmax;
    z1 += __expf(x);
    const float e = __expf(x * invt);
    zt += e;
    probs[i] = e;                              // unnormalised p at the sampling temperature
  }
  z1 = block_reduce_sum256(z1, red);
  zt = block_reduce_sum256(zt, red);
  if (temp <= 0.f) {
    if (threadIdx.x == 0) {
      out[row] = s_arg;
      if (out_logprob) out_logprob[row] = -__logf(z1);
    }
    return;
  }
  // threshold from top-k and top-p by bisection on e in (0, 1]
  const int k = top_k ? top_k[row] : 0;
  const float tp = top_p ? top_p[row] : 1.f;
  // Kept set as in vLLM / HF: top-k first, then the nucleus of the RENORMALISED top-k distribution. Both cuts are
  // thresholds on e found by bisection in (0, 1]: thr_k = the k-th largest value; thr_p = the largest threshold
  // (>= thr_k) whose kept mass is still >= top_p x (mass of the top-k set).
  float thr = 0.f;
  const bool k_on = (k > 0 && k < V), p_on = (tp < 1.f);
  float base_mass = zt;
  if (k_on) {
    float lo = 0.f, hi = 1.f;
    for (int r = 0; r < 24; ++r) {
      const float mid = 0.5f * (lo + hi);
      float cnt = 0.f;
      for (int i = threadIdx.x; i < V; i += 256) cnt += (probs[i] >= mid) ? 1.f : 0.f;
      cnt = block_reduce_sum256(cnt, red);
      if (cnt >= (float)k) lo = mid; else hi = mid;
    }
    thr = lo;
    if (p_on) {
      float m = 0.f;
      for (int i = threadIdx.x; i < V; i += 256) m += (probs[i] >= thr) ? probs[i] : 0.f;
      base_mass = block_reduce_sum256(m, red);
    }
  }
  if (p_on) {
    float lo = thr, hi = 1.f;
    for (int r = 0; r < 24; ++r) {
      const float mid = 0.5f * (lo + hi);
      float mass = 0.f;
      for (int i = threadIdx.x; i < V; i += 256) mass += (probs[i] >= mid) ? probs[i] : 0.f;
      mass = block_reduce_sum256(mass, red);
      if (mass >= tp * base_mass) lo = mid; else hi = mid;
    }
    thr = lo;
  }
  // kept mass
  float kept = 0.f;
  for (int i = threadIdx.x; i < V; i += 256) kept += (probs[i] >= thr) ? probs[i] : 0.f;
  kept = block_reduce_sum256(kept, red);
  // uniform draw
  const unsigned long long seed = seeds ? seeds[row] : 0ull;
  const uint4 rnd = philox4x32((uint32_t)step, (uint32_t)(step >> 32), (uint32_t)row, 0x5a17u,
                               (uint32_t)seed, (uint32_t)(seed >> 32));
  const float u = ((rnd.x >> 8) + 0.5f) * (1.0f / 16777216.0f) * kept;
  // inverse CDF in index order: each thread owns a contiguous segment
  const int per = (V + 255) / 256;
  const int i0 = threadIdx.x * per, i1 = min(V, i0 + per);
  float seg = 0.f;
  for (int i = i0; i < i1; ++i) seg += (probs[i] >= thr) ? probs[i] : 0.f;
  __shared__ float segs[256];
  segs[threadIdx.x] = seg;
  __syncthreads();
  if (threadIdx.x == 0) {
    float c = 0.f;
    int pick = s_arg;
    bool done = false;
    for (int t = 0; t < 256 && !done; ++t) {
      if (c + segs[t] >= u && segs[t] > 0.f) {
        const int a = t * per, bnd = min(V, a + per);
        for (int i = a; i < bnd; ++i) {
          const float e = (probs[i] >= thr) ? probs[i] : 0.f;
          c += e;
          if (e > 0.f && c >= u) { pick = i; done = true; break; }
        }
        if (!done) c = c;  // numerical slack: fall through to later segments
      } else {
        c += segs[t];
      }
    }
    out[row] = pick;
    if (out_logprob) out_logprob[row] = __logf(fmaxf(probs[pick], 1e-30f)) * temp * 1.0f - __logf(z1) ;
  }
}

inline int grid_for(long long total, int block) {
  long long g = (total + block - 1) / block;
  const long long cap = (long long)dlti_num_sms() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

DLTI_API int dlti_rope_cache(void* qkv, const int* pos, const float* cos_t, const float* sin_t,
                             const long long* slot_mapping, void* k_cache, void* v_cache,
                             long long tokens, int H, int KV, int hd, long long row_stride, int dtype,
                             cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (hd != HD) { dlti_set_error("rope_cache: head_dim must be 128"); return DLTI_ERR_UNSUPPORTED; }
  const long long total = tokens * (H + 2 * KV) * (HD / 16);
  const int g = grid_for(total, 256);
  if (dtype == DLTI_BF16)
    rope_cache_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((__nv_bfloat16*)qkv, pos, cos_t, sin_t, slot_mapping, (__nv_bfloat16*)k_cache, (__nv_bfloat16*)v_cache, tokens, H, KV, row_stride);
  else if (dtype == DLTI_FP16)
    rope_cache_kernel<__half><<<g, 256, 0, st>>>((__half*)qkv, pos, cos_t, sin_t, slot_mapping, (__half*)k_cache, (__half*)v_cache, tokens, H, KV, row_stride);
  else
    return DLTI_ERR_UNSUPPORTED;
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_paged_decode_workspace_floats(int batch, int H, int splits) {
  return batch * H * splits * (HD + 2);
}

// q [B, q_stride] (head h at column h*128); out [B, out_stride]. workspace: fp32 scratch of
// dlti_paged_decode_workspace_floats(batch, H, splits) floats.
DLTI_API int dlti_paged_decode(const void* q, long long q_stride, const void* k_cache, const void* v_cache,
                               const int* block_tables, int max_blocks, const int* ctx_lens, int batch,
                               int block_size, int H, int KV, int hd, float scale, int splits,
                               float* workspace, void* out, long long out_stride, int dtype,
                               cudaStream_t st) {
  if (batch <= 0) return 0;
  if (hd != HD || H % KV || splits < 1) return DLTI_ERR_BAD_ARG;
  float* part_o = workspace;
  float* part_ml = workspace + (long long)batch * H * splits * HD;
  dim3 grid(splits, H, batch), g2(H, batch);
  const float sl2 = scale * 1.4426950408889634f;
  if (dtype == DLTI_BF16) {
    DLTI_CUDA_CHECK(dlti_launch_pdl(batch, paged_decode_kernel<__nv_bfloat16>, grid, dim3(128), 0, st, (const __nv_bfloat16*)q,
                                    q_stride, (const __nv_bfloat16*)k_cache, (const __nv_bfloat16*)v_cache, block_tables,
                                    max_blocks, ctx_lens, block_size, H, KV, sl2, part_o, part_ml, splits,
                                    (__nv_bfloat16*)out, out_stride));
    if (splits > 1)
      DLTI_CUDA_CHECK(dlti_launch_pdl(batch, paged_combine_kernel<__nv_bfloat16>, g2, dim3(128), 0, st, (const float*)part_o,
                                      (const float*)part_ml, splits, (__nv_bfloat16*)out, out_stride, H));
  } else if (dtype == DLTI_FP16) {
    DLTI_CUDA_CHECK(dlti_launch_pdl(batch, paged_decode_kernel<__half>, grid, dim3(128), 0, st, (const __half*)q, q_stride,
                                    (const __half*)k_cache, (const __half*)v_cache, block_tables, max_blocks, ctx_lens,
                                    block_size, H, KV, sl2, part_o, part_ml, splits, (__half*)out, out_stride));
    if (splits > 1)
      DLTI_CUDA_CHECK(dlti_launch_pdl(batch, paged_combine_kernel<__half>, g2, dim3(128), 0, st, (const float*)part_o,
                                      (const float*)part_ml, splits, (__half*)out, out_stride, H));
  } else {
    return DLTI_ERR_UNSUPPORTED;
  }
  dlti_count_launch(splits > 1 ? 2 : 1);
  return 0;
}

// logits [B, V] (dtype 0/1/2), per-row temperature / top_k / top_p / seed device arrays (NULL =>
// greedy / off). out: int64 [B]; out_logprob (optional): log p(token) at temperature 1.
DLTI_API int dlti_sample(const void* logits, long long ld, int batch, int V, const float* temperature,
                         const int* top_k, const float* top_p, const unsigned long long* seeds,
                         unsigned long long step, long long* out, float* out_logprob, int dtype,
                         cudaStream_t st) {
  if (batch <= 0) return 0;
  const size_t smem = (size_t)V * sizeof(float);
  if (smem > 200 * 1024) { dlti_set_error("sample: vocabulary too large for the shared-memory path (max 51200)"); return DLTI_ERR_TOO_LARGE; }
  static bool cfg[64][3] = {};               // the attribute is per device
  int dev = 0;
  DLTI_CUDA_CHECK(cudaGetDevice(&dev));
  dev &= 63;
#define SAMPLE_CASE(IDX, TYPE)                                                                             \
  {                                                                                                         \
    if (!cfg[dev][IDX]) { DLTI_CUDA_CHECK(cudaFuncSetAttribute(sample_kernel<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); cfg[dev][IDX] = true; } \
    sample_kernel<TYPE><<<batch, 256, smem, st>>>(logits, ld, V, temperature, top_k, top_p, seeds, step, out, out_logprob); \
  }
  if (dtype == DLTI_BF16) SAMPLE_CASE(0, __nv_bfloat16)
  else if (dtype == DLTI_FP16) SAMPLE_CASE(1, __half)
  else if (dtype == DLTI_FP32) SAMPLE_CASE(2, float)
  else return DLTI_ERR_UNSUPPORTED;
#undef SAMPLE_CASE
  dlti_count_launch();
  return (int)cudaGetLastError();
}
