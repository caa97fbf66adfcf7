// Decode-step consumers of the swap-AB split-K GEMM (dlti_gemm_skinny): instead of a stand-alone
// "finish" pass (fp32 accumulator -> 16 bits, re-zero) followed by the next elementwise op, the next op
// reads the fp32 accumulator itself, re-zeroes it for the following GEMM and does its own work:
//
//   dlti_acc_add_rmsnorm   o_proj / down_proj accumulator -> residual += x ; y = RMSNorm(residual) * w
//   dlti_acc_swiglu        gate_up accumulator [rows, 2*ff] -> h = silu(gate) * up
//   dlti_acc_rope_cache    qkv accumulator -> rotated q,k + v in the 16-bit qkv buffer, k/v into the paged cache
//
// Each value is first rounded to the model dtype exactly as dlti_skinny_finish would, so the results are
// bit-identical to finish -> op; what goes away is 4 launches and 4 round trips of the activations per
// layer of a latency-bound decode step. vLLM gets the same effect from fused_add_rms_norm /
// silu_and_mul / rotary_embedding + reshape_and_cache running on cuBLAS outputs.
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

// Take 8 consecutive accumulator values, leave zeros behind, return them rounded to T (as fp32).
template <typename T>
__device__ __forceinline__ void take8(float* acc, float (&f)[8]) {
  float4* a = reinterpret_cast<float4*>(acc);
  const float4 a0 = a[0], a1 = a[1];
  a[0] = make_float4(0.f, 0.f, 0.f, 0.f);
  a[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  const float r[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
  unpack8<T>(pack8<T>(r), f);
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

// One CTA per row (same arithmetic as rmsnorm_fwd_kernel with a residual, csrc/elementwise.cu).
template <typename T, int VPT>
__global__ void __launch_bounds__(128) acc_add_rmsnorm_kernel(float* __restrict__ acc, T* __restrict__ res,
                                                              const T* __restrict__ w, T* __restrict__ y,
                                                              int d, float eps) {
  __shared__ float red[32];
  griddep_wait();            // the GEMM that fills acc (decode steps launch these kernels as programmatic dependents)
  griddep_launch();
  const long long row = blockIdx.x;
  const int nvec = d >> 3;
  int4* rr = reinterpret_cast<int4*>(res + row * d);
  const int4* wr = reinterpret_cast<const int4*>(w);
  float f[VPT][8];
  int4 rv[VPT], wv[VPT];
  float ss = 0.f;
  // all global reads of the row (accumulator, residual, norm weights) are issued before anything is consumed:
  // the kernel is one memory round trip + the reduction, and a decode step runs 64 of these back to back
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      rv[i] = rr[v];
      wv[i] = wr[v];
    }
  }
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      take8<T>(acc + row * d + v * 8, f[i]);
      float g[8];
      unpack8<T>(rv[i], g);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[i][j] += g[j];
      const int4 s = pack8<T>(f[i]);
      rr[v] = s;
      unpack8<T>(s, f[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += f[i][j] * f[i][j];
    }
  }
  ss = block_sum(ss, red);
  const float rstd = rsqrtf(ss / (float)d + eps);
  int4* yr = reinterpret_cast<int4*>(y + row * d);
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * 128;
    if (v < nvec) {
      float wf[8], o[8];
      unpack8<T>(wv[i], wf);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = wf[j] * Cvt<T>::to_f(Cvt<T>::from_f(f[i][j] * rstd));
      yr[v] = pack8<T>(o);
    }
  }
}

template <typename T>
__global__ void acc_swiglu_kernel(float* __restrict__ acc, T* __restrict__ h, long long rows, int ff,
                                  long long h_stride) {
  griddep_wait();
  griddep_launch();
  const int vpr = ff >> 3;
  const long long total = rows * vpr;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / vpr;
    const int v = (int)(idx - t * vpr);
    float* row = acc + t * 2 * ff;
    float g[8], u[8], o[8];
    take8<T>(row + v * 8, g);
    take8<T>(row + ff + v * 8, u);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sg = g[j] / (1.f + __expf(-g[j]));
      o[j] = Cvt<T>::to_f(Cvt<T>::from_f(sg)) * u[j];
    }
    *reinterpret_cast<int4*>(h + t * h_stride + v * 8) = pack8<T>(o);
  }
}

// Same thread mapping as rope_cache_kernel (csrc/serving.cu): one thread = 8 low + 8 high elements of a head.
template <typename T>
__global__ void acc_rope_cache_kernel(float* __restrict__ acc, T* __restrict__ qkv, const int* __restrict__ pos,
                                      const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                                      const long long* __restrict__ slots, T* __restrict__ k_cache,
                                      T* __restrict__ v_cache, long long tokens, int H, int KV, long long W) {
  griddep_wait();
  griddep_launch();
  const int heads = H + 2 * KV;
  const int vph = HD / 16;
  const long long total = tokens * heads * vph;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int vi = (int)(idx % vph);
    const long long th = idx / vph;
    const int h = (int)(th % heads);
    const long long t = th / heads;
    float* src = acc + t * (long long)heads * HD + (long long)h * HD + vi * 8;
    T* base = qkv + t * W + (long long)h * HD + vi * 8;
    float a[8], b[8];
    take8<T>(src, a);
    take8<T>(src + HD / 2, b);
    const long long slot = slots ? slots[t] : -1;
    int4 lo, hi;
    if (h < H + KV) {
      const int p = pos[t];
      float o1[8], o2[8];
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
    } else {
      lo = pack8<T>(a);
      hi = pack8<T>(b);
    }
    *reinterpret_cast<int4*>(base) = lo;
    *reinterpret_cast<int4*>(base + HD / 2) = hi;
    if (slot >= 0 && h >= H) {
      T* dst = (h < H + KV) ? k_cache + (slot * KV + (h - H)) * HD + vi * 8
                            : v_cache + (slot * KV + (h - H - KV)) * HD + vi * 8;
      *reinterpret_cast<int4*>(dst) = lo;
      *reinterpret_cast<int4*>(dst + HD / 2) = hi;
    }
  }
}

inline int grid_for(long long total, int block) {
  long long g = (total + block - 1) / block;
  const long long cap = (long long)dlti_num_sms() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

template <typename T>
int acc_add_rmsnorm_t(float* acc, void* res, const void* w, void* y, long long rows, int d, float eps,
                      cudaStream_t st) {
  const int vpt = (d / 8 + 127) / 128;
#define RN_CASE(V)                                                                                         \
  if (vpt <= V) {                                                                                          \
    DLTI_CUDA_CHECK(dlti_launch_pdl(rows, acc_add_rmsnorm_kernel<T, V>, dim3((unsigned)rows), dim3(128), 0, st,  \
                                    acc, (T*)res, (const T*)w, (T*)y, d, eps));                            \
    dlti_count_launch();                                                                                   \
    return 0;                                                                                              \
  }
  RN_CASE(1) RN_CASE(2) RN_CASE(4) RN_CASE(8)
#undef RN_CASE
  return DLTI_ERR_UNSUPPORTED;
}

}  // namespace

// acc32 [rows, d] fp32 (contiguous; zero on return); residual_inout, y [rows, d]; w [d].
DLTI_API int dlti_acc_add_rmsnorm(float* acc32, void* residual_inout, const void* w, void* y, long long rows,
                                  int d, float eps, int dtype, cudaStream_t st) {
  if (rows <= 0) return 0;
  if (d % 8 || d > 8192 || !residual_inout) { dlti_set_error("acc_add_rmsnorm: d must be a multiple of 8 and <= 8192"); return DLTI_ERR_BAD_ARG; }
  if (dtype == DLTI_BF16) return acc_add_rmsnorm_t<__nv_bfloat16>(acc32, residual_inout, w, y, rows, d, eps, st);
  if (dtype == DLTI_FP16) return acc_add_rmsnorm_t<__half>(acc32, residual_inout, w, y, rows, d, eps, st);
  return DLTI_ERR_UNSUPPORTED;
}

// acc32 [rows, 2*ff] fp32 (gate columns then up columns; zero on return); h [rows, h_stride].
DLTI_API int dlti_acc_swiglu(float* acc32, void* h, long long rows, int ff, long long h_stride, int dtype,
                             cudaStream_t st) {
  if (rows <= 0) return 0;
  if (ff % 8) return DLTI_ERR_BAD_ARG;
  const int g = grid_for(rows * (ff / 8), 256);
  if (dtype == DLTI_BF16)
    DLTI_CUDA_CHECK(dlti_launch_pdl(rows, acc_swiglu_kernel<__nv_bfloat16>, dim3(g), dim3(256), 0, st, acc32,
                                    (__nv_bfloat16*)h, rows, ff, h_stride));
  else if (dtype == DLTI_FP16)
    DLTI_CUDA_CHECK(dlti_launch_pdl(rows, acc_swiglu_kernel<__half>, dim3(g), dim3(256), 0, st, acc32, (__half*)h, rows, ff,
                                    h_stride));
  else
    return DLTI_ERR_UNSUPPORTED;
  dlti_count_launch();
  return 0;
}

// acc32 [tokens, (H+2KV)*128] fp32 (zero on return); qkv [tokens, row_stride] receives rope(q), rope(k), v;
// the rotated k and v rows also go to the paged caches at slot_mapping[t] (-1 = skip).
DLTI_API int dlti_acc_rope_cache(float* acc32, void* qkv, const int* pos, const float* cos_t, const float* sin_t,
                                 const long long* slot_mapping, void* k_cache, void* v_cache, long long tokens,
                                 int H, int KV, int hd, long long row_stride, int dtype, cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (hd != HD) { dlti_set_error("acc_rope_cache: head_dim must be 128"); return DLTI_ERR_UNSUPPORTED; }
  const long long total = tokens * (H + 2 * KV) * (HD / 16);
  const int g = grid_for(total, 256);
  if (dtype == DLTI_BF16)
    DLTI_CUDA_CHECK(dlti_launch_pdl(tokens, acc_rope_cache_kernel<__nv_bfloat16>, dim3(g), dim3(256), 0, st, acc32,
                                    (__nv_bfloat16*)qkv, pos, cos_t, sin_t, slot_mapping, (__nv_bfloat16*)k_cache,
                                    (__nv_bfloat16*)v_cache, tokens, H, KV, row_stride));
  else if (dtype == DLTI_FP16)
    DLTI_CUDA_CHECK(dlti_launch_pdl(tokens, acc_rope_cache_kernel<__half>, dim3(g), dim3(256), 0, st, acc32, (__half*)qkv, pos,
                                    cos_t, sin_t, slot_mapping, (__half*)k_cache, (__half*)v_cache, tokens, H, KV,
                                    row_stride));
  else
    return DLTI_ERR_UNSUPPORTED;
  dlti_count_launch();
  return 0;
}
