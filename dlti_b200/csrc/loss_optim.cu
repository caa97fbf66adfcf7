// K8 (loss part) fused softmax cross-entropy fwd+bwd over 16-bit logits, K9 (single-GPU part)
// gradient sum-of-squares / overflow detection and fused unscale + clip + AdamW + 16-bit
// write-back, plus the batched fp32->16-bit "pack" cast that refreshes the padded LoRA compute
// buffers after an optimizer step.
//
// Replaces: ATen log_softmax + nll_loss (+ backward) [SURVEY O8]; torch.optim.AdamW / DeepSpeed
// FusedAdam multi_tensor_adam.cu [N1, O11]; GradScaler's _amp_foreach_non_finite_check_and_unscale_
// and clip_grad_norm_ [N7, O12] (reference configs: configs/ds_config_zero1.json:6-14,44).
#include "common.cuh"

#include <cstdio>

namespace {

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
// Cross entropy. One CTA (256 threads) per token row.
//   loss[row] = logsumexp(logits[row]) - logits[row, label]      (0 if label == ignore_index)
//   dlogits[row, j] = (softmax_j - [j == label]) * (*gscale)      (0 row if ignored)
// dlogits may alias logits (in-place gradient). fp32 math, online max/sum in one read pass.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) ce_kernel(const T* __restrict__ logits, T* dlogits,
                                                const long long* __restrict__ labels,
                                                float* __restrict__ loss, const float* __restrict__ gscale,
                                                int V, long long ld, long long ldd,
                                                long long ignore_index) {
  __shared__ float sm[8], ss[8];
  __shared__ float s_max, s_sum;
  const long long row = blockIdx.x;
  const long long label = labels[row];
  const T* lr = logits + row * ld;
  T* dr = dlogits ? dlogits + row * ldd : nullptr;
  const int nvec = V >> 3;
  if (label == ignore_index) {
    if (threadIdx.x == 0) loss[row] = 0.f;
    if (dr) {
      const int4 z = make_int4(0, 0, 0, 0);
      for (int v = threadIdx.x; v < nvec; v += 256) reinterpret_cast<int4*>(dr)[v] = z;
    }
    return;
  }
  float m = -INFINITY, s = 0.f;
  for (int v = threadIdx.x; v < nvec; v += 256) {
    float f[8];
    unpack8<T>(reinterpret_cast<const int4*>(lr)[v], f);
    float lm = f[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) lm = fmaxf(lm, f[j]);
    const float nm = fmaxf(m, lm);
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += __expf(f[j] - nm);
    s = s * __expf(m - nm) + acc;
    m = nm;
  }
  // combine (m, s) pairs across the block
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  float wm = warp_max(m);
  float wsum = warp_sum(m == -INFINITY ? 0.f : s * __expf(m - wm));
  if (l == 0) { sm[w] = wm; ss[w] = wsum; }
  __syncthreads();
  if (w == 0) {
    float bm = l < 8 ? sm[l] : -INFINITY;
    float bs = l < 8 ? ss[l] : 0.f;
    const float gm = warp_max(bm);
    const float gs = warp_sum(bm == -INFINITY ? 0.f : bs * __expf(bm - gm));
    if (l == 0) { s_max = gm; s_sum = gs; }
  }
  __syncthreads();
  const float gm = s_max, gs = s_sum;
  const float lse = gm + __logf(gs);
  if (threadIdx.x == 0) loss[row] = lse - Cvt<T>::to_f(lr[label]);
  if (dr) {
    const float g = *gscale;
    const float inv = g / gs;
    for (int v = threadIdx.x; v < nvec; v += 256) {
      float f[8];
      unpack8<T>(reinterpret_cast<const int4*>(lr)[v], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float pj = __expf(f[j] - gm) * inv;
        if ((long long)v * 8 + j == label) pj -= g;
        f[j] = pj;
      }
      reinterpret_cast<int4*>(dr)[v] = pack8<T>(f);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Sum of squares of an fp32 vector (gradient norm + overflow detection: a non-finite element
// makes the sum non-finite). Two launches, deterministic: per-CTA partials -> final reduce.
// `partials` is caller-provided scratch of >= SUMSQ_MAX_BLOCKS floats.
// ---------------------------------------------------------------------------------------------
constexpr int SUMSQ_MAX_BLOCKS = 1024;

__global__ void __launch_bounds__(256) sumsq_partial_kernel(const float* __restrict__ g, long long n,
                                                           float* __restrict__ partials) {
  __shared__ float red[8];
  float acc = 0.f;
  const long long n4 = n >> 2;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = g4[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0)
    for (long long i = (n4 << 2) + threadIdx.x; i < n; i += 256) acc += g[i] * g[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < 8 ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) partials[blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(256) sumsq_final_kernel(const float* __restrict__ partials, int nblocks,
                                                         float* __restrict__ out, int accumulate) {
  __shared__ float red[8];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nblocks; i += 256) acc += partials[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < 8 ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) out[0] = accumulate ? out[0] + t : t;
  }
}

// ---------------------------------------------------------------------------------------------
// Fused unscale + global-norm clip + AdamW on a flat fp32 partition, with optional 16-bit copy.
//   gnorm_sq: device scalar = sum over ALL ranks of squares of the (still loss-scaled) grads.
//   If it is not finite the whole step is skipped (fp16 overflow) and *skipped_out = 1.
//   g_eff = g * inv_scale * min(1, max_norm / (sqrt(gnorm_sq) * inv_scale + 1e-6))
//   m = b1 m + (1-b1) g_eff ; v = b2 v + (1-b2) g_eff^2
//   p = p - lr * ( (m / bc1) / (sqrt(v / bc2) + eps) + wd * p )        (decoupled weight decay)
// ---------------------------------------------------------------------------------------------
template <typename T16>
__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p, float* __restrict__ m,
                                                   float* __restrict__ v, const float* __restrict__ g,
                                                   T16* __restrict__ p16, long long n, float lr, float b1,
                                                   float b2, float eps, float wd, float bc1, float bc2,
                                                   const float* __restrict__ gnorm_sq, float inv_scale,
                                                   float max_norm, int* __restrict__ skipped_out,
                                                   int zero_grad, float* __restrict__ g_mut) {
  float coef = inv_scale;
  if (gnorm_sq) {
    const float nsq = *gnorm_sq;
    if (!isfinite(nsq)) {
      if (blockIdx.x == 0 && threadIdx.x == 0 && skipped_out) *skipped_out = 1;
      return;
    }
    if (max_norm > 0.f) {
      const float gn = sqrtf(nsq) * inv_scale;
      coef *= fminf(1.f, max_norm / (gn + 1e-6f));
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && skipped_out) *skipped_out = 0;
  const float ibc1 = 1.f / bc1, ibc2 = 1.f / bc2;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const float ge = __fmul_rn(g[i], coef);
    float mi = m[i], vi = v[i];
    const float pi = adamw_update(p[i], mi, vi, ge, lr, b1, b2, eps, wd, ibc1, ibc2);
    m[i] = mi; v[i] = vi; p[i] = pi;
    if (p16) p16[i] = Cvt<T16>::from_f(pi);
    if (zero_grad) g_mut[i] = 0.f;
  }
}

// ---------------------------------------------------------------------------------------------
// Batched 2-D cast fp32 -> 16-bit with scale (one launch for every LoRA compute buffer).
// ---------------------------------------------------------------------------------------------
struct CastJob {
  const float* src;
  void* dst;
  int rows, cols;
  long long src_pitch, dst_pitch;
  float scale;
  int pad;
};
template <typename T16>
__global__ void __launch_bounds__(256) multi_cast_kernel(const CastJob* __restrict__ jobs, int njobs) {
  for (int j = blockIdx.y; j < njobs; j += gridDim.y) {
    const CastJob job = jobs[j];
    const long long total = (long long)job.rows * job.cols;
    T16* dst = reinterpret_cast<T16*>(job.dst);
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
      const int r = (int)(i / job.cols), c = (int)(i - (long long)r * job.cols);
      dst[r * job.dst_pitch + c] = Cvt<T16>::from_f(job.src[r * job.src_pitch + c] * job.scale);
    }
  }
}

}  // namespace

DLTI_API int dlti_cross_entropy(const void* logits, void* dlogits, const long long* labels,
                                float* loss_rows, const float* gscale, long long rows, int V,
                                long long ld, long long ldd, long long ignore_index, int dtype,
                                cudaStream_t st) {
  if (rows <= 0) return 0;
  if (V % 8 || ld % 8 || (dlogits && ldd % 8)) { dlti_set_error("cross_entropy: V and pitches must be multiples of 8"); return DLTI_ERR_BAD_ARG; }
  if (dlogits && !gscale) return DLTI_ERR_BAD_ARG;
  if (dtype == DLTI_BF16)
    ce_kernel<__nv_bfloat16><<<(unsigned)rows, 256, 0, st>>>((const __nv_bfloat16*)logits, (__nv_bfloat16*)dlogits, labels, loss_rows, gscale, V, ld, ldd, ignore_index);
  else if (dtype == DLTI_FP16)
    ce_kernel<__half><<<(unsigned)rows, 256, 0, st>>>((const __half*)logits, (__half*)dlogits, labels, loss_rows, gscale, V, ld, ldd, ignore_index);
  else
    return DLTI_ERR_UNSUPPORTED;
  dlti_count_launch();
  return (int)cudaGetLastError();
}

DLTI_API int dlti_sumsq_scratch_floats() { return SUMSQ_MAX_BLOCKS; }

// out[0] (+)= sum(g[i]^2). `partials` = scratch of dlti_sumsq_scratch_floats() floats.
DLTI_API int dlti_sumsq(const float* g, long long n, float* partials, float* out, int accumulate,
                        cudaStream_t st) {
  if (n < 0 || !partials || !out) return DLTI_ERR_BAD_ARG;
  long long want = (n / 4 + 255) / 256;
  int nb = (int)(want < 1 ? 1 : (want > SUMSQ_MAX_BLOCKS ? SUMSQ_MAX_BLOCKS : want));
  const int cap = dlti_num_sms() * 4;
  if (nb > cap) nb = cap;
  sumsq_partial_kernel<<<nb, 256, 0, st>>>(g, n, partials);
  sumsq_final_kernel<<<1, 256, 0, st>>>(partials, nb, out, accumulate);
  dlti_count_launch(2);
  return (int)cudaGetLastError();
}

DLTI_API int dlti_adamw(float* p, float* m, float* v, float* g, void* p16, int p16_dtype, long long n,
                        float lr, float b1, float b2, float eps, float wd, int step,
                        const float* gnorm_sq, float inv_scale, float max_norm, int* skipped_out,
                        int zero_grad, cudaStream_t st) {
  if (n <= 0) return 0;
  if (step < 1) return DLTI_ERR_BAD_ARG;
  const float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
  long long want = (n + 255) / 256;
  const long long cap = (long long)dlti_num_sms() * 8;
  const int nb = (int)(want > cap ? cap : want);
  if (p16 && p16_dtype == DLTI_FP16)
    adamw_kernel<__half><<<nb, 256, 0, st>>>(p, m, v, g, (__half*)p16, n, lr, b1, b2, eps, wd, bc1, bc2, gnorm_sq, inv_scale, max_norm, skipped_out, zero_grad, g);
  else
    adamw_kernel<__nv_bfloat16><<<nb, 256, 0, st>>>(p, m, v, g, (__nv_bfloat16*)p16, n, lr, b1, b2, eps, wd, bc1, bc2, gnorm_sq, inv_scale, max_norm, skipped_out, zero_grad, g);
  dlti_count_launch();
  return (int)cudaGetLastError();
}

// jobs: device array of {src fp32*, dst*, rows, cols, src_pitch, dst_pitch, scale, pad} (40 bytes
// padded to 48: see CastJob). One launch casts them all.
DLTI_API int dlti_multi_cast(const void* jobs_dev, int njobs, int dst_dtype, cudaStream_t st) {
  if (njobs <= 0) return 0;
  dim3 grid(32, njobs < 1024 ? njobs : 1024);
  if (dst_dtype == DLTI_BF16)
    multi_cast_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const CastJob*)jobs_dev, njobs);
  else if (dst_dtype == DLTI_FP16)
    multi_cast_kernel<__half><<<grid, 256, 0, st>>>((const CastJob*)jobs_dev, njobs);
  else
    return DLTI_ERR_UNSUPPORTED;
  dlti_count_launch();
  return (int)cudaGetLastError();
}
DLTI_API int dlti_cast_job_bytes() { return (int)sizeof(CastJob); }

// =============================================================================================
// dlti_sumsq_noalloc — the same reduction as dlti_sumsq for callers that cannot supply (or create)
// a device workspace.
//
// Who needs it: a process that has already handed every free byte of HBM to its own pool (the
// serving engine's KV-cache arena does exactly that at start-up), so a later cudaMalloc of even a
// few KB can fail; a latency-critical loop that must not call cudaMalloc/cudaFree because both
// synchronise the whole device; and a plain C host that links this library without any tensor
// allocator from which to carve the `partials` buffer dlti_sumsq asks for.
//
// How: the per-CTA partial sums go to a fixed array that is part of the library image
// (32 KB per GPU that loads the library, held until the process exits). Each CTA reduces one fixed
// chunk of NOALLOC_CHUNK elements, so the array bounds the input:
//     n <= dlti_sumsq_noalloc_max_elems() = 16384 * 8192 = 134,217,728 floats (512 MB of gradients)
// Larger inputs are rejected with DLTI_ERR_TOO_LARGE (use dlti_sumsq with your own scratch).
//
// Set-up: the FIRST call on each device resolves the array's device address with
// cudaGetSymbolAddress (the array's name is only a host-side handle, not a pointer a kernel can
// use) and caches it per device. Make that first call before starting any stream capture.
//
// Sharing: every call on a device uses the same array and leaves its partials behind. Nothing has
// to be cleared between calls — a launch overwrites every slot it later reads — but two calls whose
// kernels overlap on the GPU (different streams, no ordering between them) race on the array and
// give wrong sums. Serialise them (same stream, or an event).
//
// The function itself calls no cudaMalloc / cudaFree / allocating library routine.
// =============================================================================================
namespace {

constexpr int NOALLOC_CHUNK = 16384;       // elements reduced by one CTA
constexpr int NOALLOC_MAX_BLOCKS = 8192;   // slots in the built-in scratch array

static __device__ __align__(256) float g_sumsq_noalloc_scratch[NOALLOC_MAX_BLOCKS];

__global__ void __launch_bounds__(256) sumsq_chunk_kernel(const float* __restrict__ g, long long n,
                                                         float* __restrict__ partials, int vec_ok) {
  __shared__ float red[8];
  const long long lo = (long long)blockIdx.x * NOALLOC_CHUNK;
  const long long hi = (lo + NOALLOC_CHUNK < n) ? lo + NOALLOC_CHUNK : n;
  float acc = 0.f;
  if (vec_ok && hi - lo == NOALLOC_CHUNK) {
    const float4* g4 = reinterpret_cast<const float4*>(g + lo);
    for (int i = threadIdx.x; i < NOALLOC_CHUNK / 4; i += 256) {
      const float4 v = g4[i];
      acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
  } else {
    for (long long i = lo + threadIdx.x; i < hi; i += 256) acc += g[i] * g[i];
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < 8 ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) partials[blockIdx.x] = t;
  }
}

constexpr int NOALLOC_MAX_DEVICES = 64;
float* g_noalloc_scratch_addr[NOALLOC_MAX_DEVICES] = {nullptr};

}  // namespace

DLTI_API long long dlti_sumsq_noalloc_max_elems() {
  return (long long)NOALLOC_CHUNK * NOALLOC_MAX_BLOCKS;
}

DLTI_API int dlti_sumsq_noalloc(const float* g, long long n, float* out, int accumulate,
                                cudaStream_t st) {
  if (n < 0 || !out || (n > 0 && !g)) return DLTI_ERR_BAD_ARG;
  if (n > dlti_sumsq_noalloc_max_elems()) {
    char buf[256];
    snprintf(buf, sizeof(buf),
             "dlti_sumsq_noalloc: n = %lld exceeds the built-in scratch capacity of %lld elements; "
             "use dlti_sumsq with caller-provided scratch",
             n, dlti_sumsq_noalloc_max_elems());
    dlti_set_error(buf);
    return DLTI_ERR_TOO_LARGE;
  }
  int dev = 0;
  DLTI_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return DLTI_ERR_UNSUPPORTED;
  float* scratch = __atomic_load_n(&g_noalloc_scratch_addr[dev], __ATOMIC_ACQUIRE);
  if (!scratch) {
    // once per device, with that device current: name -> device address (two host threads racing here both
    // obtain the same address; the atomic store only makes that well-defined)
    DLTI_CUDA_CHECK(cudaGetSymbolAddress((void**)&scratch, g_sumsq_noalloc_scratch));
    __atomic_store_n(&g_noalloc_scratch_addr[dev], scratch, __ATOMIC_RELEASE);
  }
  long long nb = (n + NOALLOC_CHUNK - 1) / NOALLOC_CHUNK;
  if (nb < 1) nb = 1;
  const int vec_ok = (reinterpret_cast<uintptr_t>(g) & 15) == 0;
  sumsq_chunk_kernel<<<(unsigned)nb, 256, 0, st>>>(g, n, scratch, vec_ok);
  sumsq_final_kernel<<<1, 256, 0, st>>>(scratch, (int)nb, out, accumulate);
  dlti_count_launch(2);
  return (int)cudaGetLastError();
}
