// Host-side AdamW over bf16 parameter / moment shards in (pinned) host memory, used when FSDP
// parameters and optimizer state are offloaded to the CPU (and by the chapter-01 CPU configuration).
// OpenMP over cores; inside a thread 16 elements per AVX-512 iteration when the CPU has it (bf16 <-> fp32
// by shifts, fp32 math, round-to-nearest-even back to bf16), a scalar loop otherwise -- the same update
// as the device kernel. The update is memory bound on the host: 4 reads + 3 writes of 2 bytes per element.
//
// Reference: CPUOffload(offload_params=True) makes torch.optim.AdamW(fused=True) run its CPU kernel
// on the host shards (04-fully-sharded-data-parallel/train_llm.py:94, 05-training-llama-405b/README.md:147-153;
// OMP_NUM_THREADS=26 in 05-training-llama-405b/launch.sh:18).
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

struct AdamConsts {
  float decay, step_size, inv_bc2_sqrt, beta1, beta2, eps, grad_scale;
};

static void adamw_scalar(uint16_t* p, uint16_t* m, uint16_t* v, const uint16_t* g, long long lo, long long hi,
                         const AdamConsts& c) {
  for (long long i = lo; i < hi; ++i) {
    const float gr = bf16_to_f32(g[i]) * c.grad_scale;
    const float mf = c.beta1 * bf16_to_f32(m[i]) + (1.f - c.beta1) * gr;
    const float vf = c.beta2 * bf16_to_f32(v[i]) + (1.f - c.beta2) * gr * gr;
    const float pf = bf16_to_f32(p[i]) * c.decay - c.step_size * (mf / (sqrtf(vf) * c.inv_bc2_sqrt + c.eps));
    p[i] = f32_to_bf16(pf);
    m[i] = f32_to_bf16(mf);
    v[i] = f32_to_bf16(vf);
  }
}

#if defined(__x86_64__)
__attribute__((target("avx512f,avx512bw"))) static inline __m512 load_bf16x16(const uint16_t* ptr) {
  const __m256i h = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(ptr));
  return _mm512_castsi512_ps(_mm512_slli_epi32(_mm512_cvtepu16_epi32(h), 16));
}
__attribute__((target("avx512f,avx512bw"))) static inline void store_bf16x16(uint16_t* ptr, __m512 x) {
  const __m512i u = _mm512_castps_si512(x);
  const __m512i lsb = _mm512_and_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(1));
  __m512i r = _mm512_srli_epi32(_mm512_add_epi32(u, _mm512_add_epi32(lsb, _mm512_set1_epi32(0x7fff))), 16);
  const __mmask16 nan = _mm512_cmp_ps_mask(x, x, _CMP_UNORD_Q);
  r = _mm512_mask_or_epi32(r, nan, _mm512_srli_epi32(u, 16), _mm512_set1_epi32(0x40));
  _mm256_storeu_si256(reinterpret_cast<__m256i*>(ptr), _mm512_cvtepi32_epi16(r));
}
__attribute__((target("avx512f,avx512bw"))) static void adamw_avx512(uint16_t* p, uint16_t* m, uint16_t* v,
                                                                    const uint16_t* g, long long lo,
                                                                    long long hi, const AdamConsts& c) {
  const __m512 b1 = _mm512_set1_ps(c.beta1), omb1 = _mm512_set1_ps(1.f - c.beta1);
  const __m512 b2 = _mm512_set1_ps(c.beta2), omb2 = _mm512_set1_ps(1.f - c.beta2);
  const __m512 decay = _mm512_set1_ps(c.decay), step = _mm512_set1_ps(c.step_size);
  const __m512 ibc2 = _mm512_set1_ps(c.inv_bc2_sqrt), eps = _mm512_set1_ps(c.eps), gs = _mm512_set1_ps(c.grad_scale);
  long long i = lo;
  for (; i + 16 <= hi; i += 16) {
    const __m512 gr = _mm512_mul_ps(load_bf16x16(g + i), gs);
    const __m512 mf = _mm512_fmadd_ps(b1, load_bf16x16(m + i), _mm512_mul_ps(omb1, gr));
    const __m512 vf = _mm512_fmadd_ps(b2, load_bf16x16(v + i), _mm512_mul_ps(omb2, _mm512_mul_ps(gr, gr)));
    const __m512 denom = _mm512_fmadd_ps(_mm512_sqrt_ps(vf), ibc2, eps);
    const __m512 pf = _mm512_fnmadd_ps(step, _mm512_div_ps(mf, denom), _mm512_mul_ps(load_bf16x16(p + i), decay));
    store_bf16x16(p + i, pf);
    store_bf16x16(m + i, mf);
    store_bf16x16(v + i, vf);
  }
  adamw_scalar(p, m, v, g, i, hi, c);
}
static bool has_avx512() {
  static const bool ok = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw");
  return ok;
}
#endif

#if defined(_OPENMP)
#include <omp.h>
#endif
// Threads of the host optimizer. torchrun exports OMP_NUM_THREADS=1 to every rank, which would leave the
// offloaded AdamW on one core; the FSDP strategy calls this with cores / local ranks when it offloads
// (the reference sets OMP_NUM_THREADS=26 by hand: 05-training-llama-405b/launch.sh:18).
extern "C" int b2_cpu_set_threads(int n) {
#if defined(_OPENMP)
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  return 1;
#endif
}

extern "C" int b2_cpu_adamw_bf16(void* p_, void* m_, void* v_, const void* g_, size_t n, float lr,
                                 float beta1, float beta2, float eps, float wd, int step,
                                 float grad_scale) {
  uint16_t* p = static_cast<uint16_t*>(p_);
  uint16_t* m = static_cast<uint16_t*>(m_);
  uint16_t* v = static_cast<uint16_t*>(v_);
  const uint16_t* g = static_cast<const uint16_t*>(g_);
  AdamConsts c;
  c.decay = 1.f - lr * wd;
  c.step_size = lr / (1.f - powf(beta1, (float)step));
  c.inv_bc2_sqrt = 1.f / sqrtf(1.f - powf(beta2, (float)step));
  c.beta1 = beta1;
  c.beta2 = beta2;
  c.eps = eps;
  c.grad_scale = grad_scale;
  const long long N = (long long)n;
  const long long kBlock = 1 << 16;  // elements per work item: long enough to stream, short enough to balance
  const long long nblocks = (N + kBlock - 1) / kBlock;
#pragma omp parallel for schedule(static)
  for (long long b = 0; b < nblocks; ++b) {
    const long long lo = b * kBlock, hi = lo + kBlock < N ? lo + kBlock : N;
#if defined(__x86_64__)
    if (has_avx512()) {
      adamw_avx512(p, m, v, g, lo, hi, c);
      continue;
    }
#endif
    adamw_scalar(p, m, v, g, lo, hi, c);
  }
  return 0;
}
