// Host-side AdamW over bf16 parameter / moment shards in (pinned) host memory, used when FSDP
// parameters and optimizer state are offloaded to the CPU. OpenMP over cores, fp32 math per element,
// round-to-nearest-even back to bf16 -- the same update as the device kernel.
//
// Reference: CPUOffload(offload_params=True) makes torch.optim.AdamW(fused=True) run its CPU kernel
// on the host shards (04-fully-sharded-data-parallel/train_llm.py:94, 05-training-llama-405b/README.md:147-153;
// OMP_NUM_THREADS=26 in 05-training-llama-405b/launch.sh:18).
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

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

extern "C" int b2_cpu_adamw_bf16(void* p_, void* m_, void* v_, const void* g_, size_t n, float lr,
                                 float beta1, float beta2, float eps, float wd, int step,
                                 float grad_scale) {
  uint16_t* p = static_cast<uint16_t*>(p_);
  uint16_t* m = static_cast<uint16_t*>(m_);
  uint16_t* v = static_cast<uint16_t*>(v_);
  const uint16_t* g = static_cast<const uint16_t*>(g_);
  const float decay = 1.f - lr * wd;
  const float step_size = lr / (1.f - powf(beta1, (float)step));
  const float inv_bc2_sqrt = 1.f / sqrtf(1.f - powf(beta2, (float)step));
  const long long N = (long long)n;
#pragma omp parallel for schedule(static)
  for (long long i = 0; i < N; ++i) {
    const float gr = bf16_to_f32(g[i]) * grad_scale;
    const float mf = beta1 * bf16_to_f32(m[i]) + (1.f - beta1) * gr;
    const float vf = beta2 * bf16_to_f32(v[i]) + (1.f - beta2) * gr * gr;
    const float pf = bf16_to_f32(p[i]) * decay - step_size * (mf / (sqrtf(vf) * inv_bc2_sqrt + eps));
    p[i] = f32_to_bf16(pf);
    m[i] = f32_to_bf16(mf);
    v[i] = f32_to_bf16(vf);
  }
  return 0;
}
