// K9c: host AdamW for ZeRO-Offload (optimizer state and fp32 master partition in pinned host
// memory). OpenMP across cores, AVX-512 (runtime-dispatched) or compiler-vectorised AVX2 inner
// loop, fused unscale + clip coefficient and fused fp32 -> bf16/fp16 copy-out of the updated
// partition (the buffer that is then copied H2D and all-gathered).
//
// Replaces DeepSpeedCPUAdam (csrc/adam/cpu_adam.cpp + simd.h in DeepSpeed 0.15), which the
// reference selects with zero_optimization.offload_optimizer.device = "cpu"
// (reference configs/ds_config_zero3.json:19-22,39-47).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <immintrin.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define DLTI_CPU_API extern "C" __attribute__((visibility("default")))

namespace {

inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);                                            // round-nearest-even
  return (uint16_t)(u >> 16);
}

inline uint16_t f32_to_f16(float f) {
  return _cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT);
}

struct Hyper {
  float lr, b1, b2, eps, wd, ibc1, ibc2, coef;
};

void adam_scalar(float* p, float* m, float* v, const float* g, long long i0, long long i1,
                 const Hyper& h, uint16_t* p16, int p16_kind) {
#pragma omp simd
  for (long long i = i0; i < i1; ++i) {
    const float ge = g[i] * h.coef;
    const float mi = h.b1 * m[i] + (1.f - h.b1) * ge;
    const float vi = h.b2 * v[i] + (1.f - h.b2) * ge * ge;
    float pi = p[i];
    pi -= h.lr * ((mi * h.ibc1) / (std::sqrt(vi * h.ibc2) + h.eps) + h.wd * pi);
    m[i] = mi; v[i] = vi; p[i] = pi;
  }
  if (p16) {
    if (p16_kind == 0) for (long long i = i0; i < i1; ++i) p16[i] = f32_to_bf16(p[i]);
    else               for (long long i = i0; i < i1; ++i) p16[i] = f32_to_f16(p[i]);
  }
}

__attribute__((target("avx512f,avx512bw,avx512vl")))
void adam_avx512(float* p, float* m, float* v, const float* g, long long i0, long long i1,
                 const Hyper& h, uint16_t* p16, int p16_kind) {
  const __m512 b1 = _mm512_set1_ps(h.b1), ob1 = _mm512_set1_ps(1.f - h.b1);
  const __m512 b2 = _mm512_set1_ps(h.b2), ob2 = _mm512_set1_ps(1.f - h.b2);
  const __m512 lr = _mm512_set1_ps(h.lr), eps = _mm512_set1_ps(h.eps), wd = _mm512_set1_ps(h.wd);
  const __m512 ibc1 = _mm512_set1_ps(h.ibc1), ibc2 = _mm512_set1_ps(h.ibc2), coef = _mm512_set1_ps(h.coef);
  long long i = i0;
  for (; i + 16 <= i1; i += 16) {
    const __m512 ge = _mm512_mul_ps(_mm512_loadu_ps(g + i), coef);
    const __m512 mi = _mm512_fmadd_ps(b1, _mm512_loadu_ps(m + i), _mm512_mul_ps(ob1, ge));
    const __m512 vi = _mm512_fmadd_ps(b2, _mm512_loadu_ps(v + i), _mm512_mul_ps(ob2, _mm512_mul_ps(ge, ge)));
    __m512 pi = _mm512_loadu_ps(p + i);
    const __m512 den = _mm512_add_ps(_mm512_sqrt_ps(_mm512_mul_ps(vi, ibc2)), eps);
    const __m512 upd = _mm512_fmadd_ps(wd, pi, _mm512_div_ps(_mm512_mul_ps(mi, ibc1), den));
    pi = _mm512_fnmadd_ps(lr, upd, pi);
    _mm512_storeu_ps(m + i, mi);
    _mm512_storeu_ps(v + i, vi);
    _mm512_storeu_ps(p + i, pi);
    if (p16) {
      if (p16_kind == 0) {
        // bf16 round-to-nearest-even on 16 lanes
        __m512i u = _mm512_castps_si512(pi);
        __m512i lsb = _mm512_and_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(1));
        u = _mm512_add_epi32(u, _mm512_add_epi32(lsb, _mm512_set1_epi32(0x7fff)));
        _mm256_storeu_si256((__m256i*)(p16 + i), _mm512_cvtepi32_epi16(_mm512_srli_epi32(u, 16)));
      } else {
        _mm256_storeu_si256((__m256i*)(p16 + i), _mm512_cvtps_ph(pi, _MM_FROUND_TO_NEAREST_INT));
      }
    }
  }
  if (i < i1) adam_scalar(p, m, v, g, i, i1, h, p16, p16_kind);
}

bool have_avx512() {
  static int cached = -1;
  if (cached < 0) cached = (__builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") &&
                            __builtin_cpu_supports("avx512vl")) ? 1 : 0;
  return cached == 1;
}

}  // namespace

DLTI_CPU_API int dlti_cpu_simd_width() { return have_avx512() ? 16 : 8; }

// Returns 1 if the step was skipped (non-finite gnorm_sq: fp16 overflow), else 0.
//   g_eff = g * inv_scale * min(1, max_norm / (sqrt(gnorm_sq) * inv_scale + 1e-6))
// p16: optional 16-bit copy of the updated parameters (p16_kind 0 = bf16, 1 = fp16).
DLTI_CPU_API int dlti_cpu_adamw(float* p, float* m, float* v, const float* g, void* p16, int p16_kind,
                                long long n, float lr, float b1, float b2, float eps, float wd, int step,
                                float gnorm_sq, int use_gnorm, float inv_scale, float max_norm,
                                int num_threads) {
  Hyper h;
  h.lr = lr; h.b1 = b1; h.b2 = b2; h.eps = eps; h.wd = wd;
  h.ibc1 = 1.f / (1.f - std::pow(b1, (float)step));
  h.ibc2 = 1.f / (1.f - std::pow(b2, (float)step));
  h.coef = inv_scale;
  if (use_gnorm) {
    if (!std::isfinite(gnorm_sq)) return 1;
    if (max_norm > 0.f) {
      const float gn = std::sqrt(gnorm_sq) * inv_scale;
      h.coef *= std::fmin(1.f, max_norm / (gn + 1e-6f));
    }
  }
  const bool wide = have_avx512();
  const long long chunk = 1 << 16;
  const long long nchunks = (n + chunk - 1) / chunk;
#ifdef _OPENMP
  if (num_threads > 0) omp_set_num_threads(num_threads);
#endif
#pragma omp parallel for schedule(static)
  for (long long c = 0; c < nchunks; ++c) {
    const long long i0 = c * chunk, i1 = (i0 + chunk < n) ? i0 + chunk : n;
    if (wide) adam_avx512(p, m, v, g, i0, i1, h, (uint16_t*)p16, p16_kind);
    else      adam_scalar(p, m, v, g, i0, i1, h, (uint16_t*)p16, p16_kind);
  }
  return 0;
}
