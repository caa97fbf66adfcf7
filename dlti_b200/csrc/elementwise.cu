// K4 RMSNorm (fwd/bwd, fused residual add, fused LoRA-input dropout), K5 RoPE (fwd/bwd, in
// place on the merged qkv buffer), K7 SwiGLU (fwd/bwd), dropout, K15 embedding gather.
//
// All are HBM-bound streaming kernels: 16-byte vector loads/stores, fp32 math, one pass over
// the data (the row is held in registers between the reduction and the write).
//
// They replace ~5 (RMSNorm), ~8 (RoPE) and 2 (SiLU·mul) unfused ATen kernels per call site in
// the reference stack (HF LlamaRMSNorm / apply_rotary_pos_emb / LlamaMLP; SURVEY.md N6, O2, O5,
// O7) and vLLM's rms_norm / fused_add_rms_norm / rotary_embedding / silu_and_mul (N9, S2, S4, S8).
#include "common.cuh"

namespace {

// Dropout keep decision for element `idx` of a tensor: pure function of (seed, stream, idx) so
// that forward, recompute and backward agree. Four consecutive elements share one Philox call.
__device__ __forceinline__ uint4 dropout_bits(uint64_t seed, uint64_t stream, uint64_t idx4) {
  return philox4x32((uint32_t)idx4, (uint32_t)(idx4 >> 32), (uint32_t)stream,
                    (uint32_t)(stream >> 32), (uint32_t)seed, (uint32_t)(seed >> 32));
}

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

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (l < nw) ? red[l] : 0.f;
  t = warp_sum(t);
  return t;
}

// ---------------------------------------------------------------------------------------------
// RMSNorm forward. One CTA per row; VPT vectors (8 elements each) per thread held in registers.
//   optional res: residual stream, updated in place: res <- res + x, then normalised (serving).
//   optional ydrop: second output dropout(y) for the LoRA branch (training).
// HF rounding order: y = w * T(x * rstd).
// ---------------------------------------------------------------------------------------------
template <typename T, int VPT>
__global__ void __launch_bounds__(128) rmsnorm_fwd_kernel(const T* __restrict__ x, T* res,
                                                         const T* __restrict__ w, T* __restrict__ y,
                                                         T* __restrict__ ydrop, float* __restrict__ rstd_out,
                                                         int d, float eps, float drop_p, uint64_t seed,
                                                         uint64_t stream) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const int nvec = d >> 3;
  const int4* xr = reinterpret_cast<const int4*>(x + row * d);
  int4* rr = res ? reinterpret_cast<int4*>(res + row * d) : nullptr;
  float f[VPT][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      unpack8<T>(xr[v], f[i]);
      if (rr) {
        float g[8];
        unpack8<T>(rr[v], g);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[i][j] += g[j];
        int4 s = pack8<T>(f[i]);
        rr[v] = s;
        unpack8<T>(s, f[i]);  // normalise the rounded residual, as the unfused path would
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += f[i][j] * f[i][j];
    }
  }
  ss = block_sum(ss, red);
  const float rstd = rsqrtf(ss / (float)d + eps);
  if (threadIdx.x == 0 && rstd_out) rstd_out[row] = rstd;
  const int4* wr = reinterpret_cast<const int4*>(w);
  int4* yr = reinterpret_cast<int4*>(y + row * d);
  int4* dr = ydrop ? reinterpret_cast<int4*>(ydrop + row * d) : nullptr;
  const float keep_scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  const uint32_t thresh = (uint32_t)fminf(drop_p * 4294967296.0f, 4294967295.0f);
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      float wf[8], o[8];
      unpack8<T>(wr[v], wf);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = wf[j] * Cvt<T>::to_f(Cvt<T>::from_f(f[i][j] * rstd));
      int4 ov = pack8<T>(o);
      yr[v] = ov;
      if (dr) {
        unpack8<T>(ov, o);
        const uint64_t e0 = (uint64_t)row * d + (uint64_t)v * 8;
        uint4 r0 = dropout_bits(seed, stream, e0 >> 2);
        uint4 r1 = dropout_bits(seed, stream, (e0 >> 2) + 1);
        const uint32_t rb[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (rb[j] >= thresh) ? o[j] * keep_scale : 0.f;
        dr[v] = pack8<T>(o);
      }
    }
  }
}

// RMSNorm backward (weights frozen: no dw). dx = rstd * (g - xhat * mean(g * xhat)), g = dy * w.
// optional dres: gradient flowing on the residual stream, added to the result.
template <typename T, int VPT>
__global__ void __launch_bounds__(128) rmsnorm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                         const T* __restrict__ w,
                                                         const float* __restrict__ rstd_in,
                                                         const T* __restrict__ dres, T* __restrict__ dx,
                                                         int d) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const int nvec = d >> 3;
  const int4* dyr = reinterpret_cast<const int4*>(dy + row * d);
  const int4* xr = reinterpret_cast<const int4*>(x + row * d);
  const int4* wr = reinterpret_cast<const int4*>(w);
  const float rstd = rstd_in[row];
  float g[VPT][8], xh[VPT][8];
  float dot = 0.f;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      float wf[8];
      unpack8<T>(dyr[v], g[i]);
      unpack8<T>(xr[v], xh[i]);
      unpack8<T>(wr[v], wf);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        g[i][j] *= wf[j];
        xh[i][j] *= rstd;
        dot += g[i][j] * xh[i][j];
      }
    }
  }
  dot = block_sum(dot, red) / (float)d;
  const int4* rr = dres ? reinterpret_cast<const int4*>(dres + row * d) : nullptr;
  int4* dxr = reinterpret_cast<int4*>(dx + row * d);
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = rstd * (g[i][j] - xh[i][j] * dot);
      if (rr) {
        float r[8];
        unpack8<T>(rr[v], r);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += r[j];
      }
      dxr[v] = pack8<T>(o);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// RoPE (rotate-half, HF Llama convention) applied in place to the first `n_rot_heads` heads of
// each token row of a [T, row_stride] buffer (the merged qkv activations: q heads then k heads).
// cos/sin tables are fp32 [max_pos, hd/2]; sign = +1 forward, -1 backward (inverse rotation).
// One thread handles 8 consecutive "i" of one head: elements i..i+7 and i+hd/2..i+hd/2+7.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void rope_kernel(T* __restrict__ buf, const int* __restrict__ pos,
                            const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                            long long T_tokens, int n_rot_heads, int hd, long long row_stride,
                            float sign) {
  const int half = hd >> 1;
  const int vec_per_head = half >> 3;
  const long long total = T_tokens * n_rot_heads * vec_per_head;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int vi = (int)(idx % vec_per_head);
    const long long th = idx / vec_per_head;
    const int h = (int)(th % n_rot_heads);
    const long long t = th / n_rot_heads;
    const int p = pos[t];
    T* base = buf + t * row_stride + (long long)h * hd + vi * 8;
    int4 lo = *reinterpret_cast<int4*>(base);
    int4 hi = *reinterpret_cast<int4*>(base + half);
    float a[8], b[8], c[8], s[8];
    unpack8<T>(lo, a);
    unpack8<T>(hi, b);
    const float4* cp = reinterpret_cast<const float4*>(cos_t + (long long)p * half + vi * 8);
    const float4* sp = reinterpret_cast<const float4*>(sin_t + (long long)p * half + vi * 8);
    float4 c0 = cp[0], c1 = cp[1], s0 = sp[0], s1 = sp[1];
    c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
    s[0] = s0.x; s[1] = s0.y; s[2] = s0.z; s[3] = s0.w; s[4] = s1.x; s[5] = s1.y; s[6] = s1.z; s[7] = s1.w;
    float o1[8], o2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      // HF casts cos/sin to the activation dtype before the multiply.
      const float cj = Cvt<T>::to_f(Cvt<T>::from_f(c[j]));
      const float sj = sign * Cvt<T>::to_f(Cvt<T>::from_f(s[j]));
      o1[j] = a[j] * cj - b[j] * sj;
      o2[j] = b[j] * cj + a[j] * sj;
    }
    *reinterpret_cast<int4*>(base) = pack8<T>(o1);
    *reinterpret_cast<int4*>(base + half) = pack8<T>(o2);
  }
}

// ---------------------------------------------------------------------------------------------
// SwiGLU: gu = [T, 2*ff] (gate columns first, then up). h = silu(g) * u.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void swiglu_fwd_kernel(const T* __restrict__ gu, T* __restrict__ h, long long T_tokens,
                                  int ff, long long gu_stride, long long h_stride) {
  const int vpr = ff >> 3;
  const long long total = T_tokens * vpr;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / vpr;
    const int v = (int)(idx - t * vpr);
    const T* row = gu + t * gu_stride;
    float g[8], u[8], o[8];
    unpack8<T>(*reinterpret_cast<const int4*>(row + v * 8), g);
    unpack8<T>(*reinterpret_cast<const int4*>(row + ff + v * 8), u);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sg = g[j] / (1.f + __expf(-g[j]));
      o[j] = Cvt<T>::to_f(Cvt<T>::from_f(sg)) * u[j];  // HF: act_fn(gate) rounded, then * up
    }
    *reinterpret_cast<int4*>(h + t * h_stride + v * 8) = pack8<T>(o);
  }
}

template <typename T>
__global__ void swiglu_bwd_kernel(const T* __restrict__ dh, const T* __restrict__ gu,
                                  T* __restrict__ dgu, long long T_tokens, int ff,
                                  long long dh_stride, long long gu_stride) {
  const int vpr = ff >> 3;
  const long long total = T_tokens * vpr;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / vpr;
    const int v = (int)(idx - t * vpr);
    const T* row = gu + t * gu_stride;
    float g[8], u[8], d[8], dg[8], du[8];
    unpack8<T>(*reinterpret_cast<const int4*>(row + v * 8), g);
    unpack8<T>(*reinterpret_cast<const int4*>(row + ff + v * 8), u);
    unpack8<T>(*reinterpret_cast<const int4*>(dh + t * dh_stride + v * 8), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sig = 1.f / (1.f + __expf(-g[j]));
      const float sg = g[j] * sig;
      du[j] = d[j] * sg;
      dg[j] = d[j] * u[j] * sig * (1.f + g[j] * (1.f - sig));
    }
    T* orow = dgu + t * gu_stride;
    *reinterpret_cast<int4*>(orow + v * 8) = pack8<T>(dg);
    *reinterpret_cast<int4*>(orow + ff + v * 8) = pack8<T>(du);
  }
}

// Dropout on a [rows, cols] tensor (row pitch given); mask index = row*cols + col.
template <typename T>
__global__ void dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long rows, int cols,
                               long long x_stride, long long y_stride, float p, uint64_t seed,
                               uint64_t stream) {
  const int vpr = cols >> 3;
  const long long total = rows * vpr;
  const float ks = 1.f / (1.f - p);
  const uint32_t thresh = (uint32_t)fminf(p * 4294967296.0f, 4294967295.0f);
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long r = idx / vpr;
    const int v = (int)(idx - r * vpr);
    float f[8];
    unpack8<T>(*reinterpret_cast<const int4*>(x + r * x_stride + v * 8), f);
    const uint64_t e0 = (uint64_t)r * cols + (uint64_t)v * 8;
    uint4 r0 = dropout_bits(seed, stream, e0 >> 2);
    uint4 r1 = dropout_bits(seed, stream, (e0 >> 2) + 1);
    const uint32_t rb[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = (rb[j] >= thresh) ? f[j] * ks : 0.f;
    *reinterpret_cast<int4*>(y + r * y_stride + v * 8) = pack8<T>(f);
  }
}

// y = a + dropout(b)   (LoRA-branch input gradient merged into the base gradient)
template <typename T>
__global__ void add_dropout_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y,
                                   long long rows, int cols, float p, uint64_t seed, uint64_t stream) {
  const int vpr = cols >> 3;
  const long long total = rows * vpr;
  const float ks = p > 0.f ? 1.f / (1.f - p) : 1.f;
  const uint32_t thresh = (uint32_t)fminf(p * 4294967296.0f, 4294967295.0f);
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    float fa[8], fb[8];
    unpack8<T>(reinterpret_cast<const int4*>(a)[idx], fa);
    unpack8<T>(reinterpret_cast<const int4*>(b)[idx], fb);
    const uint64_t e0 = (uint64_t)idx * 8;
    uint4 r0 = dropout_bits(seed, stream, e0 >> 2);
    uint4 r1 = dropout_bits(seed, stream, (e0 >> 2) + 1);
    const uint32_t rb[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) fa[j] += (rb[j] >= thresh) ? fb[j] * ks : 0.f;
    reinterpret_cast<int4*>(y)[idx] = pack8<T>(fa);
  }
}

// Embedding gather: out[t, :] = table[ids[t] - vocab_start, :] (zero row if outside the shard,
// which is what a vocab-parallel embedding contributes before the cross-rank sum).
template <typename T>
__global__ void embedding_kernel(const long long* __restrict__ ids, const T* __restrict__ table,
                                 T* __restrict__ out, long long T_tokens, int d, long long vocab_start,
                                 long long vocab_rows) {
  const int vpr = d >> 3;
  const long long total = T_tokens * vpr;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / vpr;
    const int v = (int)(idx - t * vpr);
    const long long id = ids[t] - vocab_start;
    int4 val = make_int4(0, 0, 0, 0);
    if (id >= 0 && id < vocab_rows) val = *reinterpret_cast<const int4*>(table + id * d + v * 8);
    *reinterpret_cast<int4*>(out + t * d + v * 8) = val;
  }
}

// out[r, c] = T(acc[r, c]); acc[r, c] = 0   (finish step of the split-K decode GEMM)
template <typename T>
__global__ void skinny_finish_kernel(float* __restrict__ acc, T* __restrict__ out, long long rows, int cols,
                                     long long out_stride) {
  const int vpr = cols >> 3;
  const long long total = rows * vpr;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long r = idx / vpr;
    const int v = (int)(idx - r * vpr);
    float4* a = reinterpret_cast<float4*>(acc + r * cols + v * 8);
    const float4 a0 = a[0], a1 = a[1];
    const float f[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    *reinterpret_cast<int4*>(out + r * out_stride + v * 8) = pack8<T>(f);
    a[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    a[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

inline int grid_for(long long total, int block) {
  long long g = (total + block - 1) / block;
  const long long cap = (long long)dlti_num_sms() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

template <typename T>
int rmsnorm_fwd_t(const void* x, void* res, const void* w, void* y, void* ydrop, float* rstd,
                  long long rows, int d, float eps, float p, uint64_t seed, uint64_t stream_id,
                  cudaStream_t st) {
  const int nvec = d / 8;
  const int vpt = (nvec + 127) / 128;
#define RN_CASE(V)                                                                              \
  if (vpt <= V) {                                                                               \
    rmsnorm_fwd_kernel<T, V><<<(unsigned)rows, 128, 0, st>>>((const T*)x, (T*)res, (const T*)w, \
                                                            (T*)y, (T*)ydrop, rstd, d, eps, p,  \
                                                            seed, stream_id);                   \
    dlti_count_launch();                                                                        \
    return (int)cudaGetLastError();                                                             \
  }
  RN_CASE(1) RN_CASE(2) RN_CASE(4) RN_CASE(8)
#undef RN_CASE
  return DLTI_ERR_UNSUPPORTED;
}
template <typename T>
int rmsnorm_bwd_t(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                  void* dx, long long rows, int d, cudaStream_t st) {
  const int nvec = d / 8;
  const int vpt = (nvec + 127) / 128;
#define RN_CASE(V)                                                                               \
  if (vpt <= V) {                                                                                \
    rmsnorm_bwd_kernel<T, V><<<(unsigned)rows, 128, 0, st>>>((const T*)dy, (const T*)x,          \
                                                            (const T*)w, rstd, (const T*)dres,   \
                                                            (T*)dx, d);                          \
    dlti_count_launch();                                                                         \
    return (int)cudaGetLastError();                                                              \
  }
  RN_CASE(1) RN_CASE(2) RN_CASE(4) RN_CASE(8)
#undef RN_CASE
  return DLTI_ERR_UNSUPPORTED;
}

}  // namespace

#define DISPATCH_16(dtype, CALL_BF, CALL_FP)             \
  if ((dtype) == DLTI_BF16) { CALL_BF; }                 \
  else if ((dtype) == DLTI_FP16) { CALL_FP; }            \
  else return DLTI_ERR_UNSUPPORTED;

DLTI_API int dlti_rmsnorm_fwd(const void* x, void* residual_inout, const void* w, void* y,
                              void* ydrop, float* rstd, long long rows, int d, float eps,
                              float drop_p, unsigned long long seed, unsigned long long stream_id,
                              int dtype, cudaStream_t st) {
  if (rows <= 0) return 0;
  if (d % 8 || d > 8192) { dlti_set_error("rmsnorm: d must be a multiple of 8 and <= 8192"); return DLTI_ERR_BAD_ARG; }
  DISPATCH_16(dtype,
              return rmsnorm_fwd_t<__nv_bfloat16>(x, residual_inout, w, y, ydrop, rstd, rows, d, eps, drop_p, seed, stream_id, st),
              return rmsnorm_fwd_t<__half>(x, residual_inout, w, y, ydrop, rstd, rows, d, eps, drop_p, seed, stream_id, st))
}

DLTI_API int dlti_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd,
                              const void* dres, void* dx, long long rows, int d, int dtype,
                              cudaStream_t st) {
  if (rows <= 0) return 0;
  if (d % 8 || d > 8192) return DLTI_ERR_BAD_ARG;
  DISPATCH_16(dtype, return rmsnorm_bwd_t<__nv_bfloat16>(dy, x, w, rstd, dres, dx, rows, d, st),
              return rmsnorm_bwd_t<__half>(dy, x, w, rstd, dres, dx, rows, d, st))
}

DLTI_API int dlti_rope(void* buf, const int* pos, const float* cos_t, const float* sin_t,
                       long long tokens, int n_rot_heads, int hd, long long row_stride, int backward,
                       int dtype, cudaStream_t st) {
  if (tokens <= 0 || n_rot_heads <= 0) return 0;
  if (hd % 16) return DLTI_ERR_BAD_ARG;
  const long long total = tokens * n_rot_heads * (hd / 16);
  const int g = grid_for(total, 256);
  const float sign = backward ? -1.f : 1.f;
  DISPATCH_16(dtype,
              (rope_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((__nv_bfloat16*)buf, pos, cos_t, sin_t, tokens, n_rot_heads, hd, row_stride, sign)),
              (rope_kernel<__half><<<g, 256, 0, st>>>((__half*)buf, pos, cos_t, sin_t, tokens, n_rot_heads, hd, row_stride, sign)))
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_swiglu_fwd(const void* gu, void* h, long long tokens, int ff, long long gu_stride,
                             long long h_stride, int dtype, cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (ff % 8) return DLTI_ERR_BAD_ARG;
  const int g = grid_for(tokens * (ff / 8), 256);
  DISPATCH_16(dtype,
              (swiglu_fwd_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)gu, (__nv_bfloat16*)h, tokens, ff, gu_stride, h_stride)),
              (swiglu_fwd_kernel<__half><<<g, 256, 0, st>>>((const __half*)gu, (__half*)h, tokens, ff, gu_stride, h_stride)))
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_swiglu_bwd(const void* dh, const void* gu, void* dgu, long long tokens, int ff,
                             long long dh_stride, long long gu_stride, int dtype, cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (ff % 8) return DLTI_ERR_BAD_ARG;
  const int g = grid_for(tokens * (ff / 8), 256);
  DISPATCH_16(dtype,
              (swiglu_bwd_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)dh, (const __nv_bfloat16*)gu, (__nv_bfloat16*)dgu, tokens, ff, dh_stride, gu_stride)),
              (swiglu_bwd_kernel<__half><<<g, 256, 0, st>>>((const __half*)dh, (const __half*)gu, (__half*)dgu, tokens, ff, dh_stride, gu_stride)))
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_dropout(const void* x, void* y, long long rows, int cols, long long x_stride,
                          long long y_stride, float p, unsigned long long seed,
                          unsigned long long stream_id, int dtype, cudaStream_t st) {
  if (rows <= 0) return 0;
  if (cols % 8) return DLTI_ERR_BAD_ARG;
  const int g = grid_for(rows * (cols / 8), 256);
  DISPATCH_16(dtype,
              (dropout_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, rows, cols, x_stride, y_stride, p, seed, stream_id)),
              (dropout_kernel<__half><<<g, 256, 0, st>>>((const __half*)x, (__half*)y, rows, cols, x_stride, y_stride, p, seed, stream_id)))
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_add_dropout(const void* a, const void* b, void* y, long long rows, int cols, float p,
                              unsigned long long seed, unsigned long long stream_id, int dtype,
                              cudaStream_t st) {
  if (rows <= 0) return 0;
  if (cols % 8) return DLTI_ERR_BAD_ARG;
  const int g = grid_for(rows * (cols / 8), 256);
  DISPATCH_16(dtype,
              (add_dropout_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)y, rows, cols, p, seed, stream_id)),
              (add_dropout_kernel<__half><<<g, 256, 0, st>>>((const __half*)a, (const __half*)b, (__half*)y, rows, cols, p, seed, stream_id)))
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_embedding(const long long* ids, const void* table, void* out, long long tokens,
                            int d, long long vocab_start, long long vocab_rows, int dtype,
                            cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (d % 8) return DLTI_ERR_BAD_ARG;
  const int g = grid_for(tokens * (d / 8), 256);
  DISPATCH_16(dtype,
              (embedding_kernel<__nv_bfloat16><<<g, 256, 0, st>>>(ids, (const __nv_bfloat16*)table, (__nv_bfloat16*)out, tokens, d, vocab_start, vocab_rows)),
              (embedding_kernel<__half><<<g, 256, 0, st>>>(ids, (const __half*)table, (__half*)out, tokens, d, vocab_start, vocab_rows)))
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_skinny_finish(float* acc32, void* out, long long rows, int cols, long long out_stride,
                                int dtype, cudaStream_t st) {
  if (rows <= 0) return 0;
  if (cols % 8) return DLTI_ERR_BAD_ARG;
  const int g = grid_for(rows * (cols / 8), 256);
  DISPATCH_16(dtype,
              (skinny_finish_kernel<__nv_bfloat16><<<g, 256, 0, st>>>(acc32, (__nv_bfloat16*)out, rows, cols, out_stride)),
              (skinny_finish_kernel<__half><<<g, 256, 0, st>>>(acc32, (__half*)out, rows, cols, out_stride)))
  dlti_count_launch();
  return (int)cudaGetLastError();
}
