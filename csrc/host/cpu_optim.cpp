// cpu_optim.cpp — host-side optimizers for ZeRO-Offload: Adam/AdamW, Adagrad, Lion on pinned fp32
// state, AVX-512 (falls back to the compiler's auto-vectoriser elsewhere) + OpenMP, with an optional
// fused bf16/fp16 copy of the updated parameters so the H2D transfer ships 2 bytes per element.
//
// Replaces the absent csrc behind CPUAdamBuilder / CPUAdagradBuilder / CPULionBuilder
// (deepspeed/ops/adam/cpu_adam.py:94-102,163-165; ops/adagrad/cpu_adagrad.py:22-29,90-96;
// ops/lion/cpu_lion.py:57-64,125).  The id-keyed create/destroy registry is kept for API parity.
#include <immintrin.h>
#include <omp.h>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>

#define SXBH_API extern "C" __attribute__((visibility("default")))

namespace {

struct AdamCfg { float lr, b1, b2, eps, wd; int adamw; };
struct AdagradCfg { float lr, eps, wd; };
struct LionCfg { float lr, b1, b2, wd; };
std::map<int, AdamCfg> g_adam;
std::map<int, AdagradCfg> g_adagrad;
std::map<int, LionCfg> g_lion;
std::mutex g_mu;

inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
inline uint16_t f32_to_f16(float f) { return _cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT); }

// lp_kind: 0 none, 1 bf16, 2 fp16
inline void store_lp(void* lp, int lp_kind, int64_t i, float v) {
  if (lp_kind == 1) reinterpret_cast<uint16_t*>(lp)[i] = f32_to_bf16(v);
  else if (lp_kind == 2) reinterpret_cast<uint16_t*>(lp)[i] = f32_to_f16(v);
}

void adam_range(float* p, const float* g, float* m, float* v, void* lp, int lp_kind, int64_t lo, int64_t hi,
                float lr, float b1, float b2, float eps, float wd, int adamw, float bc1, float bc2,
                float grad_scale) {
  int64_t i = lo;
#if defined(__AVX512F__)
  const __m512 vb1 = _mm512_set1_ps(b1), vb1m = _mm512_set1_ps(1.f - b1);
  const __m512 vb2 = _mm512_set1_ps(b2), vb2m = _mm512_set1_ps(1.f - b2);
  const __m512 veps = _mm512_set1_ps(eps), vwd = _mm512_set1_ps(wd), vlr = _mm512_set1_ps(lr);
  const __m512 vibc1 = _mm512_set1_ps(1.f / bc1), vibc2 = _mm512_set1_ps(1.f / bc2);
  const __m512 vgs = _mm512_set1_ps(grad_scale);
  for (; i + 16 <= hi; i += 16) {
    __m512 pp = _mm512_loadu_ps(p + i);
    __m512 gg = _mm512_mul_ps(_mm512_loadu_ps(g + i), vgs);
    if (!adamw) gg = _mm512_fmadd_ps(vwd, pp, gg);
    __m512 mm = _mm512_fmadd_ps(vb1, _mm512_loadu_ps(m + i), _mm512_mul_ps(vb1m, gg));
    __m512 vv = _mm512_fmadd_ps(vb2, _mm512_loadu_ps(v + i), _mm512_mul_ps(vb2m, _mm512_mul_ps(gg, gg)));
    __m512 den = _mm512_add_ps(_mm512_sqrt_ps(_mm512_mul_ps(vv, vibc2)), veps);
    __m512 upd = _mm512_div_ps(_mm512_mul_ps(mm, vibc1), den);
    if (adamw) upd = _mm512_fmadd_ps(vwd, pp, upd);
    pp = _mm512_fnmadd_ps(vlr, upd, pp);
    _mm512_storeu_ps(p + i, pp);
    _mm512_storeu_ps(m + i, mm);
    _mm512_storeu_ps(v + i, vv);
    if (lp_kind == 1) {
      // round-to-nearest-even bf16 of 16 lanes
      __m512i u = _mm512_castps_si512(pp);
      __m512i lsb = _mm512_and_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(1));
      u = _mm512_add_epi32(u, _mm512_add_epi32(lsb, _mm512_set1_epi32(0x7fff)));
      __m256i h = _mm512_cvtepi32_epi16(_mm512_srli_epi32(u, 16));
      _mm256_storeu_si256(reinterpret_cast<__m256i*>(reinterpret_cast<uint16_t*>(lp) + i), h);
    } else if (lp_kind == 2) {
      __m256i h = _mm512_cvtps_ph(pp, _MM_FROUND_TO_NEAREST_INT);
      _mm256_storeu_si256(reinterpret_cast<__m256i*>(reinterpret_cast<uint16_t*>(lp) + i), h);
    }
  }
#endif
  for (; i < hi; ++i) {
    float pp = p[i], gg = g[i] * grad_scale;
    if (!adamw) gg += wd * pp;
    const float mm = b1 * m[i] + (1.f - b1) * gg;
    const float vv = b2 * v[i] + (1.f - b2) * gg * gg;
    float upd = (mm / bc1) / (std::sqrt(vv / bc2) + eps);
    if (adamw) upd += wd * pp;
    pp -= lr * upd;
    p[i] = pp; m[i] = mm; v[i] = vv;
    store_lp(lp, lp_kind, i, pp);
  }
}

template <typename F>
void parallel_chunks(int64_t n, F&& fn) {
  const int64_t chunk = 1 << 16;
  const int64_t nchunks = (n + chunk - 1) / chunk;
#pragma omp parallel for schedule(static)
  for (int64_t c = 0; c < nchunks; ++c) {
    const int64_t lo = c * chunk;
    const int64_t hi = lo + chunk < n ? lo + chunk : n;
    fn(lo, hi);
  }
}

}  // namespace

SXBH_API int sxbh_create_adam(int id, float lr, float b1, float b2, float eps, float wd, int adamw, int log) {
  std::lock_guard<std::mutex> l(g_mu);
  g_adam[id] = AdamCfg{lr, b1, b2, eps, wd, adamw};
  (void)log;
  return 0;
}
SXBH_API int sxbh_destroy_adam(int id) {
  std::lock_guard<std::mutex> l(g_mu);
  g_adam.erase(id);
  return 0;
}
// One Adam step on n fp32 elements.  lp (may be NULL) receives the updated params as bf16 (lp_kind=1)
// or fp16 (lp_kind=2).  grad_scale folds 1/loss_scale and the clip coefficient.
SXBH_API int sxbh_adam_update(int id, long long step, float lr, float b1, float b2, float eps, float wd,
                              int bias_correction, float* p, const float* g, float* m, float* v, long long n,
                              void* lp, int lp_kind, float grad_scale) {
  int adamw = 1;
  {
    std::lock_guard<std::mutex> l(g_mu);
    auto it = g_adam.find(id);
    if (it != g_adam.end()) adamw = it->second.adamw;
  }
  const float bc1 = bias_correction ? 1.f - std::pow(b1, (float)step) : 1.f;
  const float bc2 = bias_correction ? 1.f - std::pow(b2, (float)step) : 1.f;
  parallel_chunks(n, [&](int64_t lo, int64_t hi) {
    adam_range(p, g, m, v, lp, lp_kind, lo, hi, lr, b1, b2, eps, wd, adamw, bc1, bc2, grad_scale);
  });
  return 0;
}

SXBH_API int sxbh_create_adagrad(int id, float lr, float eps, float wd, int log) {
  std::lock_guard<std::mutex> l(g_mu);
  g_adagrad[id] = AdagradCfg{lr, eps, wd};
  (void)log;
  return 0;
}
SXBH_API int sxbh_destroy_adagrad(int id) {
  std::lock_guard<std::mutex> l(g_mu);
  g_adagrad.erase(id);
  return 0;
}
SXBH_API int sxbh_adagrad_update(int id, long long step, float lr, float eps, float wd, float* p, const float* g,
                                 float* h, long long n, void* lp, int lp_kind) {
  (void)id; (void)step;
  parallel_chunks(n, [&](int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) {
      float pp = p[i];
      const float gg = g[i] + wd * pp;
      const float hh = h[i] + gg * gg;
      h[i] = hh;
      pp -= lr * gg / (std::sqrt(hh) + eps);
      p[i] = pp;
      store_lp(lp, lp_kind, i, pp);
    }
  });
  return 0;
}

SXBH_API int sxbh_create_lion(int id, float lr, float b1, float b2, float wd, int log) {
  std::lock_guard<std::mutex> l(g_mu);
  g_lion[id] = LionCfg{lr, b1, b2, wd};
  (void)log;
  return 0;
}
SXBH_API int sxbh_destroy_lion(int id) {
  std::lock_guard<std::mutex> l(g_mu);
  g_lion.erase(id);
  return 0;
}
SXBH_API int sxbh_lion_update(int id, long long step, float lr, float b1, float b2, float wd, float* p,
                              const float* g, float* m, long long n, void* lp, int lp_kind) {
  (void)id; (void)step;
  parallel_chunks(n, [&](int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) {
      float pp = p[i];
      const float gg = g[i], mm = m[i];
      const float c = b1 * mm + (1.f - b1) * gg;
      const float s = (c > 0.f) ? 1.f : ((c < 0.f) ? -1.f : 0.f);
      pp = pp * (1.f - lr * wd) - lr * s;
      m[i] = b2 * mm + (1.f - b2) * gg;
      p[i] = pp;
      store_lp(lp, lp_kind, i, pp);
    }
  });
  return 0;
}

// bf16 <-> fp32 bulk converters used by the offload paths (gradients arrive from the GPU in bf16).
SXBH_API int sxbh_bf16_to_f32(const uint16_t* src, float* dst, long long n, float scale, int accumulate) {
  parallel_chunks(n, [&](int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) {
      const float v = bf16_to_f32(src[i]) * scale;
      dst[i] = accumulate ? dst[i] + v : v;
    }
  });
  return 0;
}
SXBH_API int sxbh_num_threads() { return omp_get_max_threads(); }
SXBH_API int sxbh_simd_width() {
#if defined(__AVX512F__)
  return 16;
#elif defined(__AVX2__)
  return 8;
#else
  return 1;
#endif
}
