// Host-side dropout with the SAME counter-based mask as the device kernel (csrc/elementwise.cu:dropout_kernel):
// vector i of eight consecutive elements takes two Philox4x32-10 blocks (counter (i, 0 | 1), key from
// (seed, offset)); element j is kept when its 32-bit draw is >= p * 2^32 and scaled by 1 / (1 - p). The
// mask is a pure function of (seed, offset, element index), so backward calls the same function on the
// gradient instead of loading a stored mask, and a CPU run and a GPU run with the same (seed, offset) drop
// the same elements.
//
// Why not ATen's: torch.nn.functional.dropout on the CPU draws its Bernoulli samples from ONE sequential
// mt19937 stream: 134 ms for the [12, 1024, 1024] attention probabilities of a GPT-2 layer on 8 cores, half of
// the chapter-01 CPU step (BASELINE.json configuration 1) in this framework and in the reference alike.
// Philox needs no shared state, so every core generates its own part, 16 counters per AVX-512 instruction
// when the CPU has it (scalar loop otherwise, and for the tail): 2.4 ms for the same tensor.
//
// Reference: GPT-2's three dropout sites reached from 01-single-gpu/train_llm.py:140 (HF GPT2Config
// embd_pdrop / attn_pdrop / resid_pdrop = 0.1, module left in training mode by from_config).
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace {

struct U4 {
  uint32_t x, y, z, w;
};

inline U4 philox4x32_10(U4 c, uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c.x, p1 = (uint64_t)0xCD9E8D57u * c.z;
    const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    const uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    c = U4{hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0};
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c;
}

inline float bf16_to_f32(uint16_t h) {
  const uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

inline float load(const uint16_t* p) { return bf16_to_f32(*p); }
inline float load(const float* p) { return *p; }
inline void store(uint16_t* p, float v) { *p = f32_to_bf16(v); }
inline void store(float* p, float v) { *p = v; }

#if defined(__x86_64__)
// ---- AVX-512: 16 vectors of eight elements (128 elements) per iteration -------------------------------------
// Lane l works on vector i0 + l. A Philox round needs the high and low halves of two 32 x 32 bit products per
// lane: low halves from vpmulld, high halves from two vpmuludq (even / odd lanes) blended back together.
#define B2_AVX512 __attribute__((target("avx512f,avx512bw,avx512dq,avx512vl,bmi2")))

B2_AVX512 static inline __m512i mulhi_u32(__m512i a, __m512i m) {
  const __m512i pe = _mm512_mul_epu32(a, m);                          // even lanes: 64-bit products
  const __m512i po = _mm512_mul_epu32(_mm512_srli_epi64(a, 32), m);   // odd lanes
  return _mm512_mask_blend_epi32(0xAAAA, _mm512_srli_epi64(pe, 32), po);
}

struct V4 {
  __m512i x, y, z, w;
};

B2_AVX512 static inline V4 philox_v(V4 c, uint32_t k0s, uint32_t k1s) {
  const __m512i M0 = _mm512_set1_epi32((int)0xD2511F53u), M1 = _mm512_set1_epi32((int)0xCD9E8D57u);
  for (int r = 0; r < 10; ++r) {
    const __m512i k0 = _mm512_set1_epi32((int)k0s), k1 = _mm512_set1_epi32((int)k1s);
    const __m512i hi0 = mulhi_u32(c.x, M0), lo0 = _mm512_mullo_epi32(c.x, M0);
    const __m512i hi1 = mulhi_u32(c.z, M1), lo1 = _mm512_mullo_epi32(c.z, M1);
    c = V4{_mm512_xor_si512(_mm512_xor_si512(hi1, c.y), k0), lo1,
           _mm512_xor_si512(_mm512_xor_si512(hi0, c.w), k1), lo0};
    k0s += 0x9E3779B9u;
    k1s += 0xBB67AE85u;
  }
  return c;
}

// keep bits of 128 consecutive elements starting at vector i0 (element e = 8 l + j: bit e of the pair lo / hi)
B2_AVX512 static inline void keep_bits_128(long long i0, uint32_t thr, uint32_t k0, uint32_t k1, uint64_t* lo,
                                           uint64_t* hi) {
  const __m512i lane = _mm512_setr_epi32(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15);
  // 64-bit vector index i0 + l split into its halves per lane
  const __m512i ilo = _mm512_add_epi32(_mm512_set1_epi32((int)(uint32_t)i0), lane);
  const __mmask16 carry = _mm512_cmplt_epu32_mask(ilo, lane);   // wrapped around 2^32
  const __m512i ihi = _mm512_mask_add_epi32(_mm512_set1_epi32((int)(uint32_t)((uint64_t)i0 >> 32)), carry,
                                            _mm512_set1_epi32((int)(uint32_t)((uint64_t)i0 >> 32)),
                                            _mm512_set1_epi32(1));
  const __m512i zero = _mm512_setzero_si512(), one = _mm512_set1_epi32(1);
  const V4 r0 = philox_v(V4{ilo, ihi, zero, zero}, k0, k1);
  const V4 r1 = philox_v(V4{ilo, ihi, one, zero}, k0, k1);
  const __m512i t = _mm512_set1_epi32((int)thr);
  const __m512i r[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
  uint64_t l = 0, h = 0;
  for (int j = 0; j < 8; ++j) {
    const uint32_t m = _mm512_cmpge_epu32_mask(r[j], t);   // bit lane: element 8 * lane + j is kept
    l |= _pdep_u64(m & 0xffu, 0x0101010101010101ULL << j);
    h |= _pdep_u64(m >> 8, 0x0101010101010101ULL << j);
  }
  *lo = l;
  *hi = h;
}

B2_AVX512 static void dropout_avx512_bf16(uint16_t* out, const uint16_t* x, long long groups, uint32_t thr,
                                          float scale, uint32_t k0, uint32_t k1) {
  const __m512 sc = _mm512_set1_ps(scale);
#pragma omp parallel for schedule(static)
  for (long long g = 0; g < groups; ++g) {
    uint64_t keep[2];
    keep_bits_128(g * 16, thr, k0, k1, &keep[0], &keep[1]);
    for (int q = 0; q < 8; ++q) {   // 16 elements at a time: bf16 -> fp32, scale, round to nearest even, mask
      const size_t e = (size_t)g * 128 + (size_t)q * 16;
      const __m256i hraw = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(x + e));
      const __m512 f = _mm512_mul_ps(_mm512_castsi512_ps(_mm512_slli_epi32(_mm512_cvtepu16_epi32(hraw), 16)), sc);
      const __m512i u = _mm512_castps_si512(f);
      const __m512i lsb = _mm512_and_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(1));
      __m512i r = _mm512_srli_epi32(_mm512_add_epi32(u, _mm512_add_epi32(lsb, _mm512_set1_epi32(0x7fff))), 16);
      const __mmask16 nan = _mm512_cmp_ps_mask(f, f, _CMP_UNORD_Q);
      r = _mm512_mask_or_epi32(r, nan, _mm512_srli_epi32(u, 16), _mm512_set1_epi32(0x40));
      const __mmask16 k = (__mmask16)(keep[q >> 2] >> ((q & 3) * 16));
      _mm256_storeu_si256(reinterpret_cast<__m256i*>(out + e), _mm512_cvtepi32_epi16(_mm512_maskz_mov_epi32(k, r)));
    }
  }
}

B2_AVX512 static void dropout_avx512_f32(float* out, const float* x, long long groups, uint32_t thr, float scale,
                                         uint32_t k0, uint32_t k1) {
  const __m512 sc = _mm512_set1_ps(scale);
#pragma omp parallel for schedule(static)
  for (long long g = 0; g < groups; ++g) {
    uint64_t keep[2];
    keep_bits_128(g * 16, thr, k0, k1, &keep[0], &keep[1]);
    for (int q = 0; q < 8; ++q) {
      const size_t e = (size_t)g * 128 + (size_t)q * 16;
      const __mmask16 k = (__mmask16)(keep[q >> 2] >> ((q & 3) * 16));
      _mm512_storeu_ps(out + e, _mm512_maskz_mul_ps(k, _mm512_loadu_ps(x + e), sc));
    }
  }
}

static bool have_avx512() {
  static const bool ok = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") &&
                         __builtin_cpu_supports("avx512dq") && __builtin_cpu_supports("avx512vl") &&
                         __builtin_cpu_supports("bmi2");
  return ok;
}
#endif

template <class T>
void dropout_impl(T* out, const T* x, size_t n, uint32_t thr, float scale, uint64_t seed, uint64_t offset) {
  const uint32_t k0 = (uint32_t)seed ^ (uint32_t)offset;
  const uint32_t k1 = (uint32_t)(seed >> 32) ^ (uint32_t)(offset >> 32) ^ 0x27D4EB2Fu;
  const long long nvec = (long long)((n + 7) / 8);
  long long first = 0;   // vectors [0, first) are done by the AVX-512 path, the rest (tail) by the scalar loop
#if defined(__x86_64__)
  if (have_avx512()) {
    const long long groups = (long long)(n / 128);
    if (sizeof(T) == 2)
      dropout_avx512_bf16(reinterpret_cast<uint16_t*>(out), reinterpret_cast<const uint16_t*>(x), groups, thr, scale,
                          k0, k1);
    else
      dropout_avx512_f32(reinterpret_cast<float*>(out), reinterpret_cast<const float*>(x), groups, thr, scale, k0, k1);
    first = groups * 16;
  }
#endif
#pragma omp parallel for schedule(static)
  for (long long i = first; i < nvec; ++i) {
    const U4 r0 = philox4x32_10(U4{(uint32_t)i, (uint32_t)((uint64_t)i >> 32), 0u, 0u}, k0, k1);
    const U4 r1 = philox4x32_10(U4{(uint32_t)i, (uint32_t)((uint64_t)i >> 32), 1u, 0u}, k0, k1);
    const uint32_t r[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    const size_t base = (size_t)i * 8;
    const int m = base + 8 <= n ? 8 : (int)(n - base);
    for (int j = 0; j < m; ++j) store(out + base + j, r[j] >= thr ? load(x + base + j) * scale : 0.f);
  }
}

}  // namespace

// out = dropout(x) for n elements of elem_size 2 (bf16) or 4 (fp32); out may alias x. Returns 0, or 1 for an
// unsupported element size / probability.
extern "C" int b2_cpu_dropout(void* out, const void* x, size_t n, int elem_size, float p, unsigned long long seed,
                              unsigned long long offset) {
  if (!(p >= 0.f && p < 1.f) || (elem_size != 2 && elem_size != 4)) return 1;
  const double t = (double)p * 4294967296.0;
  const uint32_t thr = t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
  const float scale = 1.f / (1.f - p);
  if (elem_size == 2)
    dropout_impl(static_cast<uint16_t*>(out), static_cast<const uint16_t*>(x), n, thr, scale, seed, offset);
  else
    dropout_impl(static_cast<float*>(out), static_cast<const float*>(x), n, thr, scale, seed, offset);
  return 0;
}
