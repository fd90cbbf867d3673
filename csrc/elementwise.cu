// Memory-bound training ops for sm_100a, each one pass over its operands with 16-byte accesses:
// RMSNorm / LayerNorm (+fused residual add) fwd+bwd, SwiGLU, gelu_new, RoPE, embedding gather /
// scatter-add, cross-entropy (online softmax, in-place dlogits) and AdamW over flat bf16 buffers.
//
// These replace the chains of ATen eager kernels that the reference reaches through HF modeling
// code (reference: 01-single-gpu/train_llm.py:65,140-149).
#include "common.cuh"
#include <math.h>
#include <mutex>

namespace b2 {

constexpr int kNormThreads = 256;
constexpr int kNormMaxVec = 8;  // per-thread 16B vectors -> H <= 256*8*8 = 16384

struct alignas(16) BF8 { uint32_t u[4]; };

__device__ __forceinline__ void bf8_to_f(const BF8& v, float (&f)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = unpack_bf16(v.u[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ BF8 f_to_bf8(const float (&f)[8]) {
  BF8 v;
#pragma unroll
  for (int i = 0; i < 4; ++i) v.u[i] = pack_bf16(f[2 * i], f[2 * i + 1]);
  return v;
}
// streaming (evict-first) 16-byte accesses for data touched exactly once per kernel
__device__ __forceinline__ BF8 ld_stream(const BF8* p) {
  const uint4 v = __ldcs(reinterpret_cast<const uint4*>(p));
  BF8 r;
  r.u[0] = v.x; r.u[1] = v.y; r.u[2] = v.z; r.u[3] = v.w;
  return r;
}
__device__ __forceinline__ void st_stream(BF8* p, const BF8& v) {
  __stcs(reinterpret_cast<uint4*>(p), make_uint4(v.u[0], v.u[1], v.u[2], v.u[3]));
}
__device__ __forceinline__ float bf16_round(float x) {
  return __bfloat162float(__float2bfloat16_rn(x));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// block-wide sum; `red` is >= 33 floats of shared memory. All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    float t = lane < nw ? red[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    float t = lane < nw ? red[lane] : -INFINITY;
    t = warp_max(t);
    if (lane == 0) red[32] = t;
  }
  __syncthreads();
  return red[32];
}

// ------------------------------------------------------------------------------------------------
// RMSNorm.  y = w * bf16(h * rsqrt(mean(h^2) + eps)),  h = x (+ res)   [HF LlamaRMSNorm rounding]
// ------------------------------------------------------------------------------------------------
template <int MV>  // 16-byte vectors per thread: sized to H so a row lives in few registers
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_fwd_kernel(__nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out,
                   __nv_bfloat16* __restrict__ h_out, const __nv_bfloat16* __restrict__ x,
                   const __nv_bfloat16* __restrict__ res, const __nv_bfloat16* __restrict__ w,
                   int T, int H, float eps) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  // persistent blocks (4 per SM) walking rows blockIdx.x, +gridDim.x, ...: T rows split evenly over the
  // SMs (one block per row leaves a 13 % tail at T = 4096), the next row's loads are in flight while
  // this row is reduced and stored
  BF8 nx[MV], nr[MV];
  int row = blockIdx.x;
  if (row < T) {
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = threadIdx.x + i * kNormThreads;
      if (idx < nvec) {
        nx[i] = ld_stream(reinterpret_cast<const BF8*>(x + (size_t)row * H) + idx);
        if (res) nr[i] = ld_stream(reinterpret_cast<const BF8*>(res + (size_t)row * H) + idx);
      }
    }
  }
  for (; row < T; row += gridDim.x) {
    float v[MV][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = threadIdx.x + i * kNormThreads;
      if (idx < nvec) {
        bf8_to_f(nx[i], v[i]);
        if (res) {
          float r[8];
          bf8_to_f(nr[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[i][j] = bf16_round(v[i][j] + r[j]);
          if (h_out) reinterpret_cast<BF8*>(h_out + (size_t)row * H)[idx] = f_to_bf8(v[i]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += v[i][j] * v[i][j];
      }
    }
    const int next = row + gridDim.x;
    if (next < T) {
#pragma unroll
      for (int i = 0; i < MV; ++i) {
        const int idx = threadIdx.x + i * kNormThreads;
        if (idx < nvec) {
          nx[i] = ld_stream(reinterpret_cast<const BF8*>(x + (size_t)next * H) + idx);
          if (res) nr[i] = ld_stream(reinterpret_cast<const BF8*>(res + (size_t)next * H) + idx);
        }
      }
    }
    ss = block_sum(ss, red);
    const float rstd = rsqrtf(ss / (float)H + eps);
    if (threadIdx.x == 0 && rstd_out) rstd_out[row] = rstd;
    const BF8* wr = reinterpret_cast<const BF8*>(w);
    BF8* yr = reinterpret_cast<BF8*>(y + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = threadIdx.x + i * kNormThreads;
      if (idx < nvec) {
        float wf[8], o[8];
        bf8_to_f(wr[idx], wf);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = wf[j] * bf16_round(v[i][j] * rstd);
        yr[idx] = f_to_bf8(o);
      }
    }
  }
}

// Backward. Each block walks rows blockIdx.x, +gridDim.x, ... keeping its dw partial in registers,
// then writes one fp32 partial row into `partial` [gridDim.x, H]; a second kernel folds those rows.
//   dx = rstd * (dy*w - xhat * mean(dy*w*xhat)) (+ dres)
template <int MV>
__global__ void __launch_bounds__(kNormThreads, MV <= 2 ? 3 : 1)
rmsnorm_bwd_kernel(__nv_bfloat16* __restrict__ dx, float* __restrict__ partial,
                   const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h,
                   const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd,
                   const __nv_bfloat16* __restrict__ dres, int T, int H) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  float dw[MV][8];
#pragma unroll
  for (int i = 0; i < MV; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dw[i][j] = 0.f;
  // The loads of the next row are issued before this row's block reduction: a block keeps two rows
  // (2 x 2 x H bf16) in flight, which is what the HBM latency-bandwidth product asks of the two or
  // three blocks an SM holds at this register count.
  BF8 nh[MV], ng[MV], nr[MV];
  int row = blockIdx.x;
  // rstd was written by the forward pass long ago: its 4 bytes per row come from DRAM, so they are fetched one
  // row ahead like the row itself instead of at the point of use (a dependent miss per row on the critical path)
  float rs_next = 0.f;
  if (row < T) {
    rs_next = rstd[row];
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = threadIdx.x + i * kNormThreads;
      if (idx < nvec) {
        nh[i] = ld_stream(reinterpret_cast<const BF8*>(h + (size_t)row * H) + idx);
        ng[i] = ld_stream(reinterpret_cast<const BF8*>(dy + (size_t)row * H) + idx);
        if (dres) nr[i] = ld_stream(reinterpret_cast<const BF8*>(dres + (size_t)row * H) + idx);
      }
    }
  }
  for (; row < T; row += gridDim.x) {
    const float rs = rs_next;
    float xh[MV][8], g[MV][8];
    BF8 rcur[MV];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = threadIdx.x + i * kNormThreads;
      if (idx < nvec) {
        float wf[8];
        bf8_to_f(reinterpret_cast<const BF8*>(w)[idx], wf);  // 2 * H bytes: stays in L1
        bf8_to_f(nh[i], xh[i]);
        bf8_to_f(ng[i], g[i]);
        rcur[i] = nr[i];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] *= rs;
          dw[i][j] += g[i][j] * xh[i][j];
          g[i][j] *= wf[j];
          dot += g[i][j] * xh[i][j];
        }
      }
    }
    const int next = row + gridDim.x;
    if (next < T) {
      rs_next = rstd[next];
#pragma unroll
      for (int i = 0; i < MV; ++i) {
        const int idx = threadIdx.x + i * kNormThreads;
        if (idx < nvec) {
          nh[i] = ld_stream(reinterpret_cast<const BF8*>(h + (size_t)next * H) + idx);
          ng[i] = ld_stream(reinterpret_cast<const BF8*>(dy + (size_t)next * H) + idx);
          if (dres) nr[i] = ld_stream(reinterpret_cast<const BF8*>(dres + (size_t)next * H) + idx);
        }
      }
    }
    dot = block_sum(dot, red) / (float)H;
    BF8* dxr = reinterpret_cast<BF8*>(dx + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = threadIdx.x + i * kNormThreads;
      if (idx < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (g[i][j] - xh[i][j] * dot);
        if (dres) {
          float r[8];
          bf8_to_f(rcur[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        dxr[idx] = f_to_bf8(o);
      }
    }
  }
  float* prow = partial + (size_t)blockIdx.x * H;
#pragma unroll
  for (int i = 0; i < MV; ++i) {
    const int idx = threadIdx.x + i * kNormThreads;
    if (idx < nvec) {
      float4* p4 = reinterpret_cast<float4*>(prow + (size_t)idx * 8);
      p4[0] = make_float4(dw[i][0], dw[i][1], dw[i][2], dw[i][3]);
      p4[1] = make_float4(dw[i][4], dw[i][5], dw[i][6], dw[i][7]);
    }
  }
}

// out[h] (+)= sum_b partial[b, h].  Block = 32 columns x 8 row slices (coalesced 128 B row reads).
constexpr int kFoldSlices = 32;
__global__ void __launch_bounds__(32 * kFoldSlices)
fold_partials_kernel(__nv_bfloat16* __restrict__ out, const float* __restrict__ partial,
                     int nblocks, int H, int accumulate) {
  // 32 row slices per block: a thread sums ~nblocks / 32 rows (14 of 444) with four independent chains, so the
  // 7 MB of partials are read with all loads in flight instead of 56 dependent steps per thread (the fold was
  // a fifth of the backward's time). The order of the additions is fixed: results repeat run to run.
  __shared__ float red[kFoldSlices][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int hcol = blockIdx.x * 32 + tx;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (hcol < H) {
    int b = ty;
    for (; b + 3 * kFoldSlices < nblocks; b += 4 * kFoldSlices) {
      s0 += partial[(size_t)b * H + hcol];
      s1 += partial[(size_t)(b + kFoldSlices) * H + hcol];
      s2 += partial[(size_t)(b + 2 * kFoldSlices) * H + hcol];
      s3 += partial[(size_t)(b + 3 * kFoldSlices) * H + hcol];
    }
    for (; b < nblocks; b += kFoldSlices) s0 += partial[(size_t)b * H + hcol];
  }
  red[ty][tx] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (ty == 0 && hcol < H) {
    float s = red[0][tx];
#pragma unroll
    for (int k = 1; k < kFoldSlices; ++k) s += red[k][tx];
    if (accumulate) s += __bfloat162float(out[hcol]);
    out[hcol] = __float2bfloat16_rn(s);
  }
}

// ------------------------------------------------------------------------------------------------
// LayerNorm (GPT-2): y = (x - mean) * rstd * w + b, optional residual add in front
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kNormThreads)
layernorm_fwd_kernel(__nv_bfloat16* __restrict__ y, float* __restrict__ mean_out,
                     float* __restrict__ rstd_out, __nv_bfloat16* __restrict__ h_out,
                     const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                     const __nv_bfloat16* __restrict__ w, const __nv_bfloat16* __restrict__ b,
                     int H, float eps) {
  __shared__ float red[33];
  const int row = blockIdx.x;
  const int nvec = H >> 3;
  const BF8* xr = reinterpret_cast<const BF8*>(x + (size_t)row * H);
  const BF8* rr = res ? reinterpret_cast<const BF8*>(res + (size_t)row * H) : nullptr;
  float v[kNormMaxVec][8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    const int idx = threadIdx.x + i * kNormThreads;
    if (idx < nvec) {
      bf8_to_f(xr[idx], v[i]);
      if (rr) {
        float r[8];
        bf8_to_f(rr[idx], r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] = bf16_round(v[i][j] + r[j]);
        if (h_out) reinterpret_cast<BF8*>(h_out + (size_t)row * H)[idx] = f_to_bf8(v[i]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) s += v[i][j];
    }
  }
  const float mean = block_sum(s, red) / (float)H;
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    const int idx = threadIdx.x + i * kNormThreads;
    if (idx < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = v[i][j] - mean;
        ss += d * d;
      }
    }
  }
  const float rstd = rsqrtf(block_sum(ss, red) / (float)H + eps);
  if (threadIdx.x == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
  BF8* yr = reinterpret_cast<BF8*>(y + (size_t)row * H);
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    const int idx = threadIdx.x + i * kNormThreads;
    if (idx < nvec) {
      float wf[8], bf[8], o[8];
      bf8_to_f(reinterpret_cast<const BF8*>(w)[idx], wf);
      bf8_to_f(reinterpret_cast<const BF8*>(b)[idx], bf);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * wf[j] + bf[j];
      yr[idx] = f_to_bf8(o);
    }
  }
}

// partial layout: [gridDim.x, 2, H]  (dw row then db row)
__global__ void __launch_bounds__(kNormThreads)
layernorm_bwd_kernel(__nv_bfloat16* __restrict__ dx, float* __restrict__ partial,
                     const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h,
                     const __nv_bfloat16* __restrict__ w, const float* __restrict__ mean,
                     const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres, int T,
                     int H) {
  __shared__ float red[33];
  const int nvec = H >> 3;
  // GPT-2 scale hidden sizes only (H <= 4096): keep register use bounded
  constexpr int MV = 2;
  float dw[MV][8], db[MV][8], wf[MV][8];
#pragma unroll
  for (int i = 0; i < MV; ++i) {
    const int idx = threadIdx.x + i * kNormThreads;
#pragma unroll
    for (int j = 0; j < 8; ++j) { dw[i][j] = 0.f; db[i][j] = 0.f; }
    if (idx < nvec) bf8_to_f(reinterpret_cast<const BF8*>(w)[idx], wf[i]);
  }
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const BF8* hr = reinterpret_cast<const BF8*>(h + (size_t)row * H);
    const BF8* gr = reinterpret_cast<const BF8*>(dy + (size_t)row * H);
    const float mu = mean[row], rs = rstd[row];
    float xh[MV][8], g[MV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = threadIdx.x + i * kNormThreads;
      if (idx < nvec) {
        bf8_to_f(hr[idx], xh[i]);
        bf8_to_f(gr[idx], g[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xh[i][j] - mu) * rs;
          dw[i][j] += g[i][j] * xh[i][j];
          db[i][j] += g[i][j];
          g[i][j] *= wf[i][j];
          s1 += g[i][j];
          s2 += g[i][j] * xh[i][j];
        }
      }
    }
    s1 = block_sum(s1, red) / (float)H;
    s2 = block_sum(s2, red) / (float)H;
    BF8* dxr = reinterpret_cast<BF8*>(dx + (size_t)row * H);
    const BF8* drr = dres ? reinterpret_cast<const BF8*>(dres + (size_t)row * H) : nullptr;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = threadIdx.x + i * kNormThreads;
      if (idx < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (g[i][j] - s1 - xh[i][j] * s2);
        if (drr) {
          float r[8];
          bf8_to_f(drr[idx], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        dxr[idx] = f_to_bf8(o);
      }
    }
  }
  float* pw = partial + (size_t)blockIdx.x * 2 * H;
  float* pb = pw + H;
#pragma unroll
  for (int i = 0; i < MV; ++i) {
    const int idx = threadIdx.x + i * kNormThreads;
    if (idx < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        pw[(size_t)idx * 8 + j] = dw[i][j];
        pb[(size_t)idx * 8 + j] = db[i][j];
      }
    }
  }
}

// partial [nblocks, 2H] (dw row | db row): block = 32 columns x 8 row slices, like fold_partials_kernel
__global__ void __launch_bounds__(256)
fold_partials2_kernel(__nv_bfloat16* __restrict__ dw, __nv_bfloat16* __restrict__ db,
                      const float* __restrict__ partial, int nblocks, int H, int accumulate) {
  __shared__ float red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + tx;  // column in [0, 2H)
  float s = 0.f;
  if (col < 2 * H)
    for (int b = ty; b < nblocks; b += 8) s += partial[(size_t)b * 2 * H + col];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && col < 2 * H) {
#pragma unroll
    for (int k = 1; k < 8; ++k) s += red[k][tx];
    __nv_bfloat16* out = col < H ? dw + col : db + (col - H);
    if (accumulate) s += __bfloat162float(*out);
    *out = __float2bfloat16_rn(s);
  }
}

// ---- H <= 1024 (GPT-2 scale): one warp per row, the row in registers, shuffles only ----------------
// The block-per-row kernels above leave most of a 256-thread block idle at H = 768 (96 vectors) and pay
// two block reductions per row; here a warp owns a row (MV <= 4 vectors per lane) and a block of eight
// warps walks eight rows at a time.
template <int MV>
__global__ void __launch_bounds__(256)
layernorm_fwd_warp_kernel(__nv_bfloat16* __restrict__ y, float* __restrict__ mean_out,
                          float* __restrict__ rstd_out, __nv_bfloat16* __restrict__ h_out,
                          const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                          const __nv_bfloat16* __restrict__ w, const __nv_bfloat16* __restrict__ b,
                          int T, int H, float eps) {
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int nvec = H >> 3;
  const float inv_h = 1.f / (float)H;
  for (int row = blockIdx.x * 8 + wib; row < T; row += gridDim.x * 8) {
    const BF8* xr = reinterpret_cast<const BF8*>(x + (size_t)row * H);
    const BF8* rr = res ? reinterpret_cast<const BF8*>(res + (size_t)row * H) : nullptr;
    float v[MV][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = lane + i * 32;
      if (idx < nvec) {
        bf8_to_f(ld_stream(xr + idx), v[i]);
        if (rr) {
          float r[8];
          bf8_to_f(ld_stream(rr + idx), r);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[i][j] = bf16_round(v[i][j] + r[j]);
          if (h_out) reinterpret_cast<BF8*>(h_out + (size_t)row * H)[idx] = f_to_bf8(v[i]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[i][j];
      }
    }
    const float mean = warp_sum(s) * inv_h;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MV; ++i)
      if (lane + i * 32 < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[i][j] - mean;
          ss += d * d;
        }
      }
    const float rstd = rsqrtf(warp_sum(ss) * inv_h + eps);
    if (lane == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    BF8* yr = reinterpret_cast<BF8*>(y + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = lane + i * 32;
      if (idx < nvec) {
        float wf[8], bf[8], o[8];
        bf8_to_f(reinterpret_cast<const BF8*>(w)[idx], wf);
        bf8_to_f(reinterpret_cast<const BF8*>(b)[idx], bf);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * wf[j] + bf[j];
        yr[idx] = f_to_bf8(o);
      }
    }
  }
}

// Backward, same layout. Each warp keeps its dw / db partial sums in registers over its rows; at the end
// the eight warps of a block add them into a shared [2H] fp32 row, which becomes one partial row.
template <int MV>
__global__ void __launch_bounds__(256, 2)
layernorm_bwd_warp_kernel(__nv_bfloat16* __restrict__ dx, float* __restrict__ partial,
                          const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h,
                          const __nv_bfloat16* __restrict__ w, const float* __restrict__ mean,
                          const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres,
                          int T, int H) {
  extern __shared__ float acc[];  // [2H]
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int nvec = H >> 3;
  const float inv_h = 1.f / (float)H;
  for (int i = threadIdx.x; i < 2 * H; i += 256) acc[i] = 0.f;
  __syncthreads();
  float dw[MV][8], db[MV][8];
#pragma unroll
  for (int i = 0; i < MV; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) { dw[i][j] = 0.f; db[i][j] = 0.f; }
  for (int row = blockIdx.x * 8 + wib; row < T; row += gridDim.x * 8) {
    const BF8* hr = reinterpret_cast<const BF8*>(h + (size_t)row * H);
    const BF8* gr = reinterpret_cast<const BF8*>(dy + (size_t)row * H);
    const BF8* drr = dres ? reinterpret_cast<const BF8*>(dres + (size_t)row * H) : nullptr;
    const float mu = mean[row], rs = rstd[row];
    float xh[MV][8], g[MV][8];
    BF8 rraw[MV];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = lane + i * 32;
      if (idx < nvec) {
        bf8_to_f(ld_stream(hr + idx), xh[i]);
        bf8_to_f(ld_stream(gr + idx), g[i]);
        if (drr) rraw[i] = ld_stream(drr + idx);
        float wf[8];
        bf8_to_f(reinterpret_cast<const BF8*>(w)[idx], wf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xh[i][j] - mu) * rs;
          dw[i][j] += g[i][j] * xh[i][j];
          db[i][j] += g[i][j];
          g[i][j] *= wf[j];
          s1 += g[i][j];
          s2 += g[i][j] * xh[i][j];
        }
      }
    }
    s1 = warp_sum(s1) * inv_h;
    s2 = warp_sum(s2) * inv_h;
    BF8* dxr = reinterpret_cast<BF8*>(dx + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < MV; ++i) {
      const int idx = lane + i * 32;
      if (idx < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (g[i][j] - s1 - xh[i][j] * s2);
        if (drr) {
          float r[8];
          bf8_to_f(rraw[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        dxr[idx] = f_to_bf8(o);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < MV; ++i) {
    const int idx = lane + i * 32;
    if (idx < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(&acc[idx * 8 + j], dw[i][j]);
        atomicAdd(&acc[H + idx * 8 + j], db[i][j]);
      }
    }
  }
  __syncthreads();
  float* prow = partial + (size_t)blockIdx.x * 2 * H;
  for (int i = threadIdx.x; i < 2 * H; i += 256) prow[i] = acc[i];
}

// ------------------------------------------------------------------------------------------------
// Column sums of a [T, N] bf16 matrix (bias gradients of GPT-2's Conv1D layers: db = sum_t dy[t, :]).
// grid (ceil(N / 256), R): a warp reads 256 consecutive columns of one row (16 bytes per lane), the eight
// warps of a block take every eighth row of the block's slice, the block writes one fp32 partial row;
// fold_partials_kernel adds the R rows up (and accumulates into an existing gradient).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
colsum_partial_kernel(float* __restrict__ partial, const __nv_bfloat16* __restrict__ x, int T, int N,
                      int ld) {
  __shared__ float red[8][257];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + lane * 8;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col < N) {
    for (int row = blockIdx.y * 8 + wib; row < T; row += gridDim.y * 8) {
      float v[8];
      bf8_to_f(ld_stream(reinterpret_cast<const BF8*>(x + (size_t)row * ld + col)), v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[wib][lane * 8 + j] = acc[j];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < N) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[k][threadIdx.x];
    partial[(size_t)blockIdx.y * N + c] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// SwiGLU on a fused [T, 2I] gate|up buffer, and gelu_new
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ BF8 swiglu_vec(const BF8& gv, const BF8& uv) {
  float g[8], u[8], o[8];
  bf8_to_f(gv, g);
  bf8_to_f(uv, u);
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j] = bf16_round(__fdividef(g[j], 1.f + __expf(-g[j]))) * u[j];
  return f_to_bf8(o);
}

// one block walks rows (grid-stride), its threads walk the 16-byte columns: no index division
__global__ void __launch_bounds__(256)
swiglu_fwd_kernel(__nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ gu, int T, int I) {
  const int nvec = I >> 3;
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const BF8* g = reinterpret_cast<const BF8*>(gu + (size_t)row * 2 * I);
    const BF8* u = g + nvec;
    BF8* o = reinterpret_cast<BF8*>(out + (size_t)row * I);
    int c = threadIdx.x;
    for (; c + 256 < nvec; c += 512) {  // two independent vectors (4 loads) in flight
      const BF8 g0 = ld_stream(g + c), u0 = ld_stream(u + c);
      const BF8 g1 = ld_stream(g + c + 256), u1 = ld_stream(u + c + 256);
      o[c] = swiglu_vec(g0, u0);
      o[c + 256] = swiglu_vec(g1, u1);
    }
    if (c < nvec) o[c] = swiglu_vec(ld_stream(g + c), ld_stream(u + c));
  }
}

__global__ void __launch_bounds__(256)
swiglu_bwd_kernel(__nv_bfloat16* __restrict__ dgu, const __nv_bfloat16* __restrict__ dout,
                  const __nv_bfloat16* __restrict__ gu, int T, int I) {
  const int nvec = I >> 3;
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const BF8* gp = reinterpret_cast<const BF8*>(gu + (size_t)row * 2 * I);
    const BF8* dp = reinterpret_cast<const BF8*>(dout + (size_t)row * I);
    BF8* ob = reinterpret_cast<BF8*>(dgu + (size_t)row * 2 * I);
    for (int c = threadIdx.x; c < nvec; c += 256) {
      const BF8 gv = ld_stream(gp + c), uv = ld_stream(gp + nvec + c), dv = ld_stream(dp + c);
      float g[8], u[8], d[8], dg[8], du[8];
      bf8_to_f(gv, g);
      bf8_to_f(uv, u);
      bf8_to_f(dv, d);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float sig = __fdividef(1.f, 1.f + __expf(-g[j]));
        const float silu = g[j] * sig;
        du[j] = d[j] * silu;
        dg[j] = d[j] * u[j] * (sig * (1.f + g[j] * (1.f - sig)));
      }
      ob[c] = f_to_bf8(dg);
      ob[nvec + c] = f_to_bf8(du);
    }
  }
}

// tanh as the single MUFU instruction (|error| ~ 2^-11, below bf16 resolution); tanhf() costs ~20
// instructions and made the GELU kernels compute bound
__device__ __forceinline__ float tanh_approx(float x) {
  float r;
  asm("tanh.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float gelu_new_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  return 0.5f * x * (1.f + tanh_approx(k0 * (x + k1 * x * x * x)));
}
__device__ __forceinline__ float gelu_new_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  const float t = tanh_approx(k0 * (x + k1 * x * x * x));
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * k0 * (1.f + 3.f * k1 * x * x);
}
// mode 0: out = gelu(x);  mode 1: out = dout * gelu'(x)
__global__ void gelu_kernel(__nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ x,
                            const __nv_bfloat16* __restrict__ dout, size_t nvec, int mode) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (size_t)gridDim.x * blockDim.x) {
    float a[8], o[8];
    bf8_to_f(reinterpret_cast<const BF8*>(x)[i], a);
    if (mode == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = gelu_new_f(a[j]);
    } else {
      float d[8];
      bf8_to_f(reinterpret_cast<const BF8*>(dout)[i], d);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = d[j] * gelu_new_grad(a[j]);
    }
    reinterpret_cast<BF8*>(out)[i] = f_to_bf8(o);
  }
}

// ------------------------------------------------------------------------------------------------
// RoPE, in place on the first `nheads` heads of each row of a fused QKV buffer.
// x: [T, row_stride] ; cos/sin: [P, D] bf16 tables ; pos: [T] int32 (nullptr -> t % seq_len)
// out = x*cos + rotate_half(x)*sin ; sign = -1 applies the transpose (backward).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rope_kernel(__nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ cosT,
            const __nv_bfloat16* __restrict__ sinT, const int* __restrict__ pos, int T, int nheads,
            int D, int row_stride, int seq_len, float sign) {
  const int half_vec = D >> 4;  // 16B vectors per half head
  const int per_row = nheads * half_vec;
  for (int t = blockIdx.x; t < T; t += gridDim.x) {  // one token row per block trip: no index division
    const int p = pos ? pos[t] : t % seq_len;
    const __nv_bfloat16* crow = cosT + (size_t)p * D;
    const __nv_bfloat16* srow = sinT + (size_t)p * D;
    for (int i = threadIdx.x; i < per_row; i += blockDim.x) {
      const int head = i / half_vec, c = i - head * half_vec;
      __nv_bfloat16* base = x + (size_t)t * row_stride + (size_t)head * D;
      BF8* lo = reinterpret_cast<BF8*>(base) + c;
      BF8* hi = reinterpret_cast<BF8*>(base + (D >> 1)) + c;
      float a[8], b[8], c0[8], c1[8], s0[8], s1[8], oa[8], ob[8];
      bf8_to_f(*lo, a);
      bf8_to_f(*hi, b);
      bf8_to_f(reinterpret_cast<const BF8*>(crow)[c], c0);
      bf8_to_f(reinterpret_cast<const BF8*>(crow + (D >> 1))[c], c1);
      bf8_to_f(reinterpret_cast<const BF8*>(srow)[c], s0);
      bf8_to_f(reinterpret_cast<const BF8*>(srow + (D >> 1))[c], s1);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (sign > 0.f) {
          // rotate_half(x) = [-x_hi, x_lo]
          oa[j] = a[j] * c0[j] - b[j] * s0[j];
          ob[j] = b[j] * c1[j] + a[j] * s1[j];
        } else {
          // transpose: d_lo = g_lo*cos_lo + g_hi*sin_hi ; d_hi = g_hi*cos_hi - g_lo*sin_lo
          oa[j] = a[j] * c0[j] + b[j] * s1[j];
          ob[j] = b[j] * c1[j] - a[j] * s0[j];
        }
      }
      *lo = f_to_bf8(oa);
      *hi = f_to_bf8(ob);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Embedding
// ------------------------------------------------------------------------------------------------
__global__ void embedding_fwd_kernel(__nv_bfloat16* __restrict__ out,
                                     const __nv_bfloat16* __restrict__ table,
                                     const long long* __restrict__ ids, int T, int H) {
  const int nvec = H >> 3;
  const size_t total = (size_t)T * nvec;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t t = i / nvec;
    const int c = (int)(i - t * nvec);
    reinterpret_cast<BF8*>(out + t * (size_t)H)[c] =
        reinterpret_cast<const BF8*>(table + (size_t)ids[t] * H)[c];
  }
}
// dtable[id, :] (+)= sum over the positions holding `id` of dout[pos, :], accumulated in fp32 and
// without atomics: the token ids are sorted on the host side (sorted_ids / perm), block i does work only
// when position i starts a run of equal ids and then sums the whole run (deterministic, and exact like
// the reference's sort-based embedding_dense_backward, however often a token repeats).
__global__ void __launch_bounds__(128)
embedding_bwd_kernel(__nv_bfloat16* __restrict__ dtable, const __nv_bfloat16* __restrict__ dout,
                     const long long* __restrict__ sorted_ids, const long long* __restrict__ perm,
                     int T, int H, int accumulate) {
  const int i = blockIdx.x;
  const long long id = sorted_ids[i];
  if (i > 0 && sorted_ids[i - 1] == id) return;  // not the first position of its run
  const int nvec = H >> 3;
  for (int c = threadIdx.x; c < nvec; c += blockDim.x) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k = i; k < T && sorted_ids[k] == id; ++k) {
      float v[8];
      bf8_to_f(reinterpret_cast<const BF8*>(dout + (size_t)perm[k] * H)[c], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
    BF8* dst = reinterpret_cast<BF8*>(dtable + (size_t)id * H) + c;
    if (accumulate) {
      float o[8];
      bf8_to_f(*dst, o);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += o[j];
    }
    *dst = f_to_bf8(acc);
  }
}

// ------------------------------------------------------------------------------------------------
// Cross entropy over bf16 logits [T, ld] (first V columns valid), fp32 math (HF upcasts to fp32).
// fwd: lse[t], loss[t] (0 where label == ignore). bwd: dlogits = (softmax - onehot) * scale, in place
// allowed; columns >= V and ignored rows are zeroed so the buffer can feed the wgrad GEMM directly.
// `vocab_start` supports vocabulary-parallel shards: labels are global, columns are local.
// ------------------------------------------------------------------------------------------------
constexpr int kCEThreads = 512;

__global__ void __launch_bounds__(kCEThreads)
ce_fwd_kernel(float* __restrict__ lse_out, float* __restrict__ loss_out,
              float* __restrict__ max_out, float* __restrict__ sum_out,
              float* __restrict__ tgt_out, const __nv_bfloat16* __restrict__ logits,
              const long long* __restrict__ labels, int V, int ld, int ignore_index,
              int vocab_start) {
  __shared__ float red[33];
  const int row = blockIdx.x;
  const __nv_bfloat16* lr = logits + (size_t)row * ld;
  const int nvec = V >> 3;
  float m = -INFINITY, s = 0.f;
  for (int i = threadIdx.x; i < nvec; i += kCEThreads) {
    float f[8];
    bf8_to_f(reinterpret_cast<const BF8*>(lr)[i], f);
    float lm = f[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) lm = fmaxf(lm, f[j]);
    if (lm > m) {
      s *= __expf(m - lm);
      m = lm;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) s += __expf(f[j] - m);
  }
  for (int i = (nvec << 3) + threadIdx.x; i < V; i += kCEThreads) {
    const float f = __bfloat162float(lr[i]);
    if (f > m) {
      s *= __expf(m - f);
      m = f;
    }
    s += __expf(f - m);
  }
  const float gm = block_max(m, red);
  s = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
  const float gs = block_sum(s, red);
  if (threadIdx.x == 0) {
    const long long lab = labels[row];
    const long long local = lab - vocab_start;
    const bool mine = lab != ignore_index && local >= 0 && local < V;
    const float tgt = mine ? __bfloat162float(lr[local]) : 0.f;
    if (max_out) {  // vocabulary-parallel: export the shard statistics, finish after the exchange
      max_out[row] = gm;
      sum_out[row] = gs;
      tgt_out[row] = tgt;
    } else {
      const float lse = gm + logf(gs);
      lse_out[row] = lse;
      loss_out[row] = (lab == ignore_index) ? 0.f : (lse - tgt);
    }
  }
}

__global__ void __launch_bounds__(kCEThreads)
ce_bwd_kernel(__nv_bfloat16* __restrict__ dlogits, const __nv_bfloat16* __restrict__ logits,
              const float* __restrict__ lse, const long long* __restrict__ labels, int V, int ld,
              int ignore_index, int vocab_start, float scale_in,
              const float* __restrict__ scale_ptr) {
  const float scale = scale_ptr ? scale_in * (*scale_ptr) : scale_in;
  const int row = blockIdx.x;
  const __nv_bfloat16* lr = logits + (size_t)row * ld;
  __nv_bfloat16* dr = dlogits + (size_t)row * ld;
  const long long lab = labels[row];
  const bool ignored = lab == ignore_index;
  const long long local = lab - vocab_start;
  const float l = lse[row];
  const int nvec_all = ld >> 3;
  for (int i = threadIdx.x; i < nvec_all; i += kCEThreads) {
    float f[8], o[8];
    bf8_to_f(reinterpret_cast<const BF8*>(lr)[i], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int col = i * 8 + j;
      float g = 0.f;
      if (!ignored && col < V) {
        g = __expf(f[j] - l);
        if (col == local) g -= 1.f;
        g *= scale;
      }
      o[j] = g;
    }
    reinterpret_cast<BF8*>(dr)[i] = f_to_bf8(o);
  }
}

// ------------------------------------------------------------------------------------------------
// AdamW over a flat bf16 parameter range with bf16 moments (matches torch fused AdamW on bf16
// params: fp32 math per element, decoupled weight decay, bias correction). `grad_scale` folds the
// 1/world factor of a data-parallel mean into the update.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__device__ __forceinline__ void adamw_vec(float (&pf)[8], float (&mf)[8], float (&vf)[8],
                                          const float (&gf)[8], float decay, float beta1,
                                          float beta2, float eps, float step_size,
                                          float inv_bc2_sqrt, float grad_scale) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float gr = gf[j] * grad_scale;
    mf[j] = beta1 * mf[j] + (1.f - beta1) * gr;
    vf[j] = beta2 * vf[j] + (1.f - beta2) * gr * gr;
    const float denom = sqrt_approx(vf[j]) * inv_bc2_sqrt + eps;
    pf[j] = pf[j] * decay - step_size * __fdividef(mf[j], denom);
  }
}

// store of an updated parameter vector: plain, or (ZeRO-1) one multimem.st on the multicast mapping of the
// replicated flat parameter buffer, which lands the owner's update in every rank's copy at once -- the
// "broadcast of updated params" of ZeroRedundancyOptimizer (N3) without a separate all-gather
__device__ __forceinline__ void st_param(BF8* p, BF8* p_mc, size_t i, const BF8& v) {
  if (p_mc == nullptr) {
    st_stream(p + i, v);
  } else {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p_mc + i), "r"(v.u[0]),
                 "r"(v.u[1]), "r"(v.u[2]), "r"(v.u[3])
                 : "memory");
  }
}

// kSide only makes two distinct functions: the one launched beside the GEMMs of backward asks for the GEMM's
// shared-memory carveout (see B2_SIDE_KERNEL); the stand-alone one keeps the default, because with 28 KB of
// L1 left a streaming kernel cannot keep enough loads in flight (0.61 vs 0.50 ms per 218 M parameters)
template <bool kSide>
__global__ void __launch_bounds__(256)
adamw_kernel(__nv_bfloat16* __restrict__ p, __nv_bfloat16* __restrict__ m,
             __nv_bfloat16* __restrict__ v, const __nv_bfloat16* __restrict__ g, size_t n, float lr,
             float beta1, float beta2, float eps, float wd, float bc1, float bc2_sqrt,
             float grad_scale, __nv_bfloat16* p_mc) {
  const size_t nvec = n >> 3;
  const float decay = 1.f - lr * wd, step_size = lr / bc1, inv_bc2_sqrt = 1.f / bc2_sqrt;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  // two independent 16-byte vectors per thread per trip: 8 loads in flight before any math
  for (; i + stride < nvec; i += 2 * stride) {
    const size_t i1 = i + stride;
    const BF8 p0 = ld_stream(reinterpret_cast<const BF8*>(p) + i), p1 = ld_stream(reinterpret_cast<const BF8*>(p) + i1);
    const BF8 m0 = ld_stream(reinterpret_cast<const BF8*>(m) + i), m1 = ld_stream(reinterpret_cast<const BF8*>(m) + i1);
    const BF8 v0 = ld_stream(reinterpret_cast<const BF8*>(v) + i), v1 = ld_stream(reinterpret_cast<const BF8*>(v) + i1);
    const BF8 g0 = ld_stream(reinterpret_cast<const BF8*>(g) + i), g1 = ld_stream(reinterpret_cast<const BF8*>(g) + i1);
    float pf[8], mf[8], vf[8], gf[8];
    bf8_to_f(p0, pf); bf8_to_f(m0, mf); bf8_to_f(v0, vf); bf8_to_f(g0, gf);
    adamw_vec(pf, mf, vf, gf, decay, beta1, beta2, eps, step_size, inv_bc2_sqrt, grad_scale);
    st_param(reinterpret_cast<BF8*>(p), reinterpret_cast<BF8*>(p_mc), i, f_to_bf8(pf));
    st_stream(reinterpret_cast<BF8*>(m) + i, f_to_bf8(mf));
    st_stream(reinterpret_cast<BF8*>(v) + i, f_to_bf8(vf));
    bf8_to_f(p1, pf); bf8_to_f(m1, mf); bf8_to_f(v1, vf); bf8_to_f(g1, gf);
    adamw_vec(pf, mf, vf, gf, decay, beta1, beta2, eps, step_size, inv_bc2_sqrt, grad_scale);
    st_param(reinterpret_cast<BF8*>(p), reinterpret_cast<BF8*>(p_mc), i1, f_to_bf8(pf));
    st_stream(reinterpret_cast<BF8*>(m) + i1, f_to_bf8(mf));
    st_stream(reinterpret_cast<BF8*>(v) + i1, f_to_bf8(vf));
  }
  if (i < nvec) {
    float pf[8], mf[8], vf[8], gf[8];
    bf8_to_f(reinterpret_cast<const BF8*>(p)[i], pf);
    bf8_to_f(reinterpret_cast<const BF8*>(m)[i], mf);
    bf8_to_f(reinterpret_cast<const BF8*>(v)[i], vf);
    bf8_to_f(reinterpret_cast<const BF8*>(g)[i], gf);
    adamw_vec(pf, mf, vf, gf, decay, beta1, beta2, eps, step_size, inv_bc2_sqrt, grad_scale);
    st_param(reinterpret_cast<BF8*>(p), reinterpret_cast<BF8*>(p_mc), i, f_to_bf8(pf));
    reinterpret_cast<BF8*>(m)[i] = f_to_bf8(mf);
    reinterpret_cast<BF8*>(v)[i] = f_to_bf8(vf);
  }
  // tail (n % 8) handled by the first block
  if (blockIdx.x == 0) {
    for (size_t i = (nvec << 3) + threadIdx.x; i < n; i += blockDim.x) {
      const float gr = __bfloat162float(g[i]) * grad_scale;
      float pf = __bfloat162float(p[i]) * (1.f - lr * wd);
      const float mf = beta1 * __bfloat162float(m[i]) + (1.f - beta1) * gr;
      const float vf = beta2 * __bfloat162float(v[i]) + (1.f - beta2) * gr * gr;
      pf -= (lr / bc1) * (mf / (sqrtf(vf) / bc2_sqrt + eps));
      p[i] = __float2bfloat16_rn(pf);
      m[i] = __float2bfloat16_rn(mf);
      v[i] = __float2bfloat16_rn(vf);
    }
  }
}

// Elementwise dropout: out = keep ? x / (1 - p) : 0 with the keep mask a pure function of (seed, offset,
// element index) -- backward calls the same kernel on the gradient. Eight elements per thread from two
// Philox4x32-10 blocks; an element is kept when its 32-bit draw is >= p * 2^32 (exact drop rate).
__global__ void __launch_bounds__(256)
dropout_kernel(__nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ x, size_t nvec, uint32_t thr,
               float scale, unsigned long long seed, unsigned long long offset) {
  const uint2 key = make_uint2((uint32_t)seed ^ (uint32_t)offset,
                               (uint32_t)(seed >> 32) ^ (uint32_t)(offset >> 32) ^ 0x27D4EB2Fu);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 r0 = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)(i >> 32), 0u, 0u), key);
    const uint4 r1 = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)(i >> 32), 1u, 0u), key);
    const uint32_t r[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    float v[8];
    bf8_to_f(ld_stream(reinterpret_cast<const BF8*>(x) + i), v);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = r[j] >= thr ? v[j] * scale : 0.f;
    st_stream(reinterpret_cast<BF8*>(out) + i, f_to_bf8(v));
  }
}

// out = a + b (bf16), used for gradient accumulation into flat buffers and residual joins
__global__ void add_kernel(__nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ a,
                           const __nv_bfloat16* __restrict__ b, size_t nvec) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (size_t)gridDim.x * blockDim.x) {
    float x[8], y[8];
    bf8_to_f(reinterpret_cast<const BF8*>(a)[i], x);
    bf8_to_f(reinterpret_cast<const BF8*>(b)[i], y);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] += y[j];
    reinterpret_cast<BF8*>(out)[i] = f_to_bf8(x);
  }
}

static inline int ew_grid(size_t work_items, int threads) {
  size_t b = (work_items + threads - 1) / threads;
  const size_t cap = (size_t)sm_count() * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace b2

using namespace b2;
typedef __nv_bfloat16 bf16;

extern "C" {

int b2_rmsnorm_bwd_partial_rows(int T) {
  int nb = sm_count() * 4;
  return nb < T ? nb : T;
}

int b2_rmsnorm_fwd(void* y, float* rstd, void* h_out, const void* x, const void* res,
                   const void* w, int T, int H, float eps, cudaStream_t stream) {
  B2_REQUIRE(H % 8 == 0 && H <= kNormThreads * kNormMaxVec * 8, "rmsnorm: unsupported H=%d", H);
  if (T == 0) return 0;
  const int grid = T < sm_count() * 4 ? T : sm_count() * 4;
#define B2_RMS_FWD(MV)                                                                       \
  rmsnorm_fwd_kernel<MV><<<grid, kNormThreads, 0, stream>>>((bf16*)y, rstd, (bf16*)h_out,    \
                                                            (const bf16*)x, (const bf16*)res, \
                                                            (const bf16*)w, T, H, eps)
  if (H <= kNormThreads * 8) B2_RMS_FWD(1);
  else if (H <= kNormThreads * 8 * 2) B2_RMS_FWD(2);
  else if (H <= kNormThreads * 8 * 4) B2_RMS_FWD(4);
  else B2_RMS_FWD(8);
#undef B2_RMS_FWD
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// one wave of resident blocks (3 per SM up to H = 4096, 1 above: register footprint); the scratch is
// sized for b2_rmsnorm_bwd_partial_rows(T) >= this many rows
static int rms_bwd_blocks(int T, int H) {
  int nb = sm_count() * (H <= kNormThreads * 8 * 2 ? 3 : 1);
  return nb < T ? nb : T;
}

// partial: caller-provided fp32 scratch of b2_rmsnorm_bwd_partial_rows(T) * H floats
int b2_rmsnorm_bwd(void* dx, void* dw, float* partial, const void* dy, const void* h,
                   const void* w, const float* rstd, const void* dres, int T, int H,
                   int accumulate_dw, cudaStream_t stream) {
  B2_REQUIRE(H % 8 == 0 && H <= kNormThreads * kNormMaxVec * 8, "rmsnorm: unsupported H=%d", H);
  if (T == 0) return 0;
  const int nb = rms_bwd_blocks(T, H);
#define B2_RMS_BWD(MV)                                                                     \
  rmsnorm_bwd_kernel<MV><<<nb, kNormThreads, 0, stream>>>(                                 \
      (bf16*)dx, partial, (const bf16*)dy, (const bf16*)h, (const bf16*)w, rstd,           \
      (const bf16*)dres, T, H)
  if (H <= kNormThreads * 8 * 2) B2_RMS_BWD(2);
  else if (H <= kNormThreads * 8 * 4) B2_RMS_BWD(4);
  else B2_RMS_BWD(8);
#undef B2_RMS_BWD
  B2_CUDA_CHECK(cudaGetLastError());
  fold_partials_kernel<<<(H + 31) / 32, 32 * kFoldSlices, 0, stream>>>((bf16*)dw, partial, nb, H,
                                                          accumulate_dw);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_layernorm_fwd(void* y, float* mean, float* rstd, void* h_out, const void* x,
                     const void* res, const void* w, const void* b, int T, int H, float eps,
                     cudaStream_t stream) {
  B2_REQUIRE(H % 8 == 0 && H <= kNormThreads * kNormMaxVec * 8, "layernorm: unsupported H=%d", H);
  if (T == 0) return 0;
  if (H <= 1024) {
    const int want = (T + 7) / 8, cap = sm_count() * 8;
    const int grid = want < cap ? want : cap;
#define B2_LN_FWD(MV)                                                                          \
  layernorm_fwd_warp_kernel<MV><<<grid, 256, 0, stream>>>(                                     \
      (bf16*)y, mean, rstd, (bf16*)h_out, (const bf16*)x, (const bf16*)res, (const bf16*)w,    \
      (const bf16*)b, T, H, eps)
    if (H <= 256) B2_LN_FWD(1);
    else if (H <= 512) B2_LN_FWD(2);
    else if (H <= 768) B2_LN_FWD(3);
    else B2_LN_FWD(4);
#undef B2_LN_FWD
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
  }
  layernorm_fwd_kernel<<<T, kNormThreads, 0, stream>>>((bf16*)y, mean, rstd, (bf16*)h_out,
                                                       (const bf16*)x, (const bf16*)res,
                                                       (const bf16*)w, (const bf16*)b, H, eps);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// partial: fp32 scratch of b2_rmsnorm_bwd_partial_rows(T) * 2 * H floats
int b2_layernorm_bwd(void* dx, void* dw, void* db, float* partial, const void* dy, const void* h,
                     const void* w, const float* mean, const float* rstd, const void* dres, int T,
                     int H, int accumulate, cudaStream_t stream) {
  B2_REQUIRE(H % 8 == 0 && H <= kNormThreads * 2 * 8, "layernorm bwd: unsupported H=%d", H);
  if (T == 0) return 0;
  int nb = b2_rmsnorm_bwd_partial_rows(T);
  if (H <= 1024) {
    // warp-per-row: at most 2 blocks per SM (each keeps its rows' dw / db in registers), nb <= scratch rows
    const int want = (T + 7) / 8, cap = sm_count() * 2;
    nb = want < cap ? want : cap;
#define B2_LN_BWD(MV)                                                                            \
  layernorm_bwd_warp_kernel<MV><<<nb, 256, 2 * H * sizeof(float), stream>>>(                     \
      (bf16*)dx, partial, (const bf16*)dy, (const bf16*)h, (const bf16*)w, mean, rstd,           \
      (const bf16*)dres, T, H)
    if (H <= 256) B2_LN_BWD(1);
    else if (H <= 512) B2_LN_BWD(2);
    else if (H <= 768) B2_LN_BWD(3);
    else B2_LN_BWD(4);
#undef B2_LN_BWD
  } else {
    layernorm_bwd_kernel<<<nb, kNormThreads, 0, stream>>>((bf16*)dx, partial, (const bf16*)dy,
                                                          (const bf16*)h, (const bf16*)w, mean, rstd,
                                                          (const bf16*)dres, T, H);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  fold_partials2_kernel<<<(2 * H + 31) / 32, 256, 0, stream>>>((bf16*)dw, (bf16*)db, partial, nb, H,
                                                               accumulate);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// out[N] (bf16) (+)= sum over rows of x[T, N] (row stride ld). partial: fp32 scratch of
// b2_colsum_partial_rows(T, N) * N floats.
int b2_colsum_partial_rows(int T, int N) {
  const int tiles = (N + 255) / 256;
  int r = (sm_count() * 4 + tiles - 1) / tiles;
  const int max_r = (T + 7) / 8;
  if (r > max_r) r = max_r;
  return r < 1 ? 1 : r;
}
int b2_colsum(void* out, float* partial, const void* x, int T, int N, int ld, int accumulate,
              cudaStream_t stream) {
  B2_REQUIRE(N % 8 == 0 && ld % 8 == 0, "colsum: N=%d ld=%d must be multiples of 8", N, ld);
  if (T == 0 || N == 0) return 0;
  const int R = b2_colsum_partial_rows(T, N);
  colsum_partial_kernel<<<dim3((N + 255) / 256, R), 256, 0, stream>>>(partial, (const bf16*)x, T, N, ld);
  B2_CUDA_CHECK(cudaGetLastError());
  fold_partials_kernel<<<(N + 31) / 32, 32 * kFoldSlices, 0, stream>>>((bf16*)out, partial, R, N, accumulate);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_swiglu_fwd(void* out, const void* gu, int T, int I, cudaStream_t stream) {
  B2_REQUIRE(I % 8 == 0, "swiglu: I=%d not a multiple of 8", I);
  if (T == 0) return 0;
  swiglu_fwd_kernel<<<T < sm_count() * 8 ? T : sm_count() * 8, 256, 0, stream>>>((bf16*)out,
                                                                                 (const bf16*)gu, T, I);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}
int b2_swiglu_bwd(void* dgu, const void* dout, const void* gu, int T, int I, cudaStream_t stream) {
  B2_REQUIRE(I % 8 == 0, "swiglu: I=%d not a multiple of 8", I);
  if (T == 0) return 0;
  swiglu_bwd_kernel<<<T < sm_count() * 8 ? T : sm_count() * 8, 256, 0, stream>>>(
      (bf16*)dgu, (const bf16*)dout, (const bf16*)gu, T, I);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}
int b2_gelu_fwd(void* out, const void* x, size_t n, cudaStream_t stream) {
  B2_REQUIRE(n % 8 == 0, "gelu: n not a multiple of 8");
  if (n == 0) return 0;
  gelu_kernel<<<ew_grid(n / 8, 256), 256, 0, stream>>>((bf16*)out, (const bf16*)x, nullptr, n / 8,
                                                       0);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}
int b2_gelu_bwd(void* dx, const void* x, const void* dout, size_t n, cudaStream_t stream) {
  B2_REQUIRE(n % 8 == 0, "gelu: n not a multiple of 8");
  if (n == 0) return 0;
  gelu_kernel<<<ew_grid(n / 8, 256), 256, 0, stream>>>((bf16*)dx, (const bf16*)x,
                                                       (const bf16*)dout, n / 8, 1);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_rope(void* x, const void* cosT, const void* sinT, const int* pos, int T, int nheads, int D,
            int row_stride, int seq_len, int backward, cudaStream_t stream) {
  B2_REQUIRE(D % 16 == 0 && row_stride % 8 == 0, "rope: D=%d row_stride=%d unsupported", D,
             row_stride);
  if (T == 0 || nheads == 0) return 0;
  rope_kernel<<<T < sm_count() * 8 ? T : sm_count() * 8, 256, 0, stream>>>(
      (bf16*)x, (const bf16*)cosT, (const bf16*)sinT, pos, T, nheads, D, row_stride, seq_len,
      backward ? -1.f : 1.f);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_embedding_fwd(void* out, const void* table, const long long* ids, int T, int H,
                     cudaStream_t stream) {
  B2_REQUIRE(H % 8 == 0, "embedding: H=%d not a multiple of 8", H);
  if (T == 0) return 0;
  embedding_fwd_kernel<<<ew_grid((size_t)T * (H / 8), 256), 256, 0, stream>>>(
      (bf16*)out, (const bf16*)table, ids, T, H);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}
// sorted_ids: the T token ids in ascending order; perm[k]: original position of sorted_ids[k].
// Rows of tokens that do not occur are left untouched (zero dtable first unless accumulating).
int b2_embedding_bwd(void* dtable, const void* dout, const long long* sorted_ids,
                     const long long* perm, int T, int H, int accumulate, cudaStream_t stream) {
  B2_REQUIRE(H % 8 == 0, "embedding: H=%d not a multiple of 8", H);
  if (T == 0) return 0;
  embedding_bwd_kernel<<<T, 128, 0, stream>>>((bf16*)dtable, (const bf16*)dout, sorted_ids, perm, T, H,
                                              accumulate);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_ce_fwd(float* lse, float* loss, const void* logits, const long long* labels, int T, int V,
              int ld, int ignore_index, cudaStream_t stream) {
  B2_REQUIRE(ld % 8 == 0 && ld >= V, "ce: ld=%d V=%d unsupported", ld, V);
  if (T == 0) return 0;
  ce_fwd_kernel<<<T, kCEThreads, 0, stream>>>(lse, loss, nullptr, nullptr, nullptr,
                                              (const bf16*)logits, labels, V, ld, ignore_index, 0);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}
// vocabulary-parallel shard statistics: row max, sum exp(l - max), target logit (0 if not local)
int b2_ce_shard_stats(float* row_max, float* row_sum, float* row_tgt, const void* logits,
                      const long long* labels, int T, int V, int ld, int ignore_index,
                      int vocab_start, cudaStream_t stream) {
  B2_REQUIRE(ld % 8 == 0 && ld >= V, "ce: ld=%d V=%d unsupported", ld, V);
  if (T == 0) return 0;
  ce_fwd_kernel<<<T, kCEThreads, 0, stream>>>(nullptr, nullptr, row_max, row_sum, row_tgt,
                                              (const bf16*)logits, labels, V, ld, ignore_index,
                                              vocab_start);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}
// effective scale = scale * (*scale_dev) when scale_dev (a device float) is given
int b2_ce_bwd(void* dlogits, const void* logits, const float* lse, const long long* labels, int T,
              int V, int ld, int ignore_index, int vocab_start, float scale, const float* scale_dev,
              cudaStream_t stream) {
  B2_REQUIRE(ld % 8 == 0 && ld >= V, "ce: ld=%d V=%d unsupported", ld, V);
  if (T == 0) return 0;
  ce_bwd_kernel<<<T, kCEThreads, 0, stream>>>((bf16*)dlogits, (const bf16*)logits, lse, labels, V,
                                              ld, ignore_index, vocab_start, scale, scale_dev);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

static int adamw_impl(void* p, void* m, void* v, const void* g, size_t n, float lr, float beta1,
                      float beta2, float eps, float wd, int step, float grad_scale, int max_blocks, void* p_mc,
                      cudaStream_t stream) {
  if (n == 0) return 0;
  B2_REQUIRE(p_mc == nullptr || (n % 8 == 0 && (uintptr_t)p_mc % 16 == 0),
             "adamw: a multicast parameter store needs n %% 8 == 0 (n=%zu) and a 16B aligned address", n);
  B2_REQUIRE(((uintptr_t)p | (uintptr_t)m | (uintptr_t)v | (uintptr_t)g) % 16 == 0,
             "adamw: buffers must be 16B aligned");
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float bc2 = 1.f - powf(beta2, (float)step);
  int grid = ew_grid(n / 8 + 1, 256);
  if (max_blocks > 0 && grid > max_blocks) grid = max_blocks;
  if (max_blocks > 0) {  // capped grid = launched beside the GEMMs of backward on the optimizer stream
    B2_SIDE_KERNEL((adamw_kernel<true>));
    adamw_kernel<true><<<grid, 256, 0, stream>>>((bf16*)p, (bf16*)m, (bf16*)v, (const bf16*)g, n, lr, beta1,
                                                 beta2, eps, wd, bc1, sqrtf(bc2), grad_scale, (bf16*)p_mc);
  } else {
    adamw_kernel<false><<<grid, 256, 0, stream>>>((bf16*)p, (bf16*)m, (bf16*)v, (const bf16*)g, n, lr, beta1,
                                                  beta2, eps, wd, bc1, sqrtf(bc2), grad_scale, (bf16*)p_mc);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_adamw(void* p, void* m, void* v, const void* g, size_t n, float lr, float beta1,
             float beta2, float eps, float wd, int step, float grad_scale, int max_blocks,
             cudaStream_t stream) {
  return adamw_impl(p, m, v, g, n, lr, beta1, beta2, eps, wd, step, grad_scale, max_blocks, nullptr, stream);
}

// ZeRO-1: p is this rank's slice of a flat parameter buffer that every rank replicates in symmetric memory and
// p_mc the multicast address of the same slice: the updated parameters are written with multimem.st, i.e.
// into every rank's copy (including this one's). Follow with b2_peer_barrier before anyone reads them.
int b2_adamw_multicast(void* p, void* p_mc, void* m, void* v, const void* g, size_t n, float lr, float beta1,
                       float beta2, float eps, float wd, int step, float grad_scale, int max_blocks,
                       cudaStream_t stream) {
  B2_REQUIRE(p_mc != nullptr, "adamw_multicast: no multicast mapping%s", "");
  return adamw_impl(p, m, v, g, n, lr, beta1, beta2, eps, wd, step, grad_scale, max_blocks, p_mc, stream);
}

// out[n] = dropout(x) with drop probability p (n % 8 == 0); out may alias x. Same (seed, offset) -> same mask.
int b2_dropout(void* out, const void* x, size_t n, float p, unsigned long long seed, unsigned long long offset,
               cudaStream_t stream) {
  B2_REQUIRE(n % 8 == 0, "dropout: n not a multiple of 8");
  B2_REQUIRE(p >= 0.f && p < 1.f, "dropout: p=%f out of range", p);
  if (n == 0) return 0;
  const double t = (double)p * 4294967296.0;
  const uint32_t thr = t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
  dropout_kernel<<<ew_grid(n / 8, 256), 256, 0, stream>>>((bf16*)out, (const bf16*)x, n / 8, thr, 1.f / (1.f - p),
                                                          seed, offset);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_add(void* out, const void* a, const void* b, size_t n, cudaStream_t stream) {
  B2_REQUIRE(n % 8 == 0, "add: n not a multiple of 8");
  if (n == 0) return 0;
  add_kernel<<<ew_grid(n / 8, 256), 256, 0, stream>>>((bf16*)out, (const bf16*)a, (const bf16*)b,
                                                      n / 8);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// RMSNorm backward for callers that cannot allocate device memory.
//
// b2_rmsnorm_bwd needs an fp32 scratch buffer that its caller allocates. Some callers have no way to
// do that: a plain-C plugin loaded into a host application that has already handed all remaining
// device memory to its own pool (so cudaMalloc fails and the pool's allocator is not reachable from
// here), or a latency-critical section where the implicit device-wide synchronisation of cudaMalloc
// is not acceptable. For them the scratch lives in the library image itself, in the fixed-size array
// below. Every GPU that loads libb200train.so holds these 32 MiB from load until process exit.
// ------------------------------------------------------------------------------------------------
namespace b2 {
constexpr size_t kStaticScratchFloats = size_t(8) << 20;  // 8 Mi floats = 32 MiB
static __device__ __align__(256) float g_rmsnorm_static_scratch[kStaticScratchFloats];

// Host code must not pass `g_rmsnorm_static_scratch` to a kernel: in host code that name is only a
// host-side handle for the symbol, not a device address. The device address is looked up once per
// device (with that device current) and cached here.
constexpr int kMaxDevices = 64;
static float* g_static_scratch_ptr[kMaxDevices] = {};
static std::mutex g_static_scratch_mutex;

static int static_scratch_for_current_device(float** out) {
  int dev = -1;
  B2_CUDA_CHECK(cudaGetDevice(&dev));
  B2_REQUIRE(dev >= 0 && dev < kMaxDevices, "static scratch: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (g_static_scratch_ptr[dev] == nullptr) {
    float* p = nullptr;
    B2_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_rmsnorm_static_scratch));
    B2_REQUIRE(p != nullptr, "static scratch: cudaGetSymbolAddress returned a null pointer");
    g_static_scratch_ptr[dev] = p;
  }
  *out = g_static_scratch_ptr[dev];
  return 0;
}
}  // namespace b2

extern "C" {

// Largest scratch (in floats) the no-allocation entry point can serve: partial_rows(T) * H must fit.
size_t b2_rmsnorm_bwd_static_capacity(void) { return b2::kStaticScratchFloats; }

// Same computation, same launch shape and same results as b2_rmsnorm_bwd, but without a scratch
// argument and without any allocation (no cudaMalloc / cudaFree / allocating library call): the fp32
// partial sums go to the library's built-in 32 MiB device array.
//   * Limit: b2_rmsnorm_bwd_partial_rows(T) * H <= b2_rmsnorm_bwd_static_capacity() floats; larger
//     problems are refused with an error (use b2_rmsnorm_bwd with your own scratch instead).
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress); make
//     that first call during start-up, before any stream capture begins.
//   * All calls on one device share the one array. Nothing in it needs to be zero on entry: each call
//     overwrites every element it later reads, so leftovers of the previous call are harmless. Two
//     calls that overlap on the GPU (different streams, no ordering between them) would race on the
//     array and corrupt each other's dw; keep calls on one stream or order them with events.
int b2_rmsnorm_bwd_static(void* dx, void* dw, const void* dy, const void* h, const void* w,
                          const float* rstd, const void* dres, int T, int H, int accumulate_dw,
                          cudaStream_t stream) {
  B2_REQUIRE(H % 8 == 0 && H <= kNormThreads * kNormMaxVec * 8, "rmsnorm: unsupported H=%d", H);
  if (T == 0) return 0;
  const int nb = b2_rmsnorm_bwd_partial_rows(T);
  B2_REQUIRE((size_t)nb * (size_t)H <= kStaticScratchFloats,
             "b2_rmsnorm_bwd_static: needs %zu scratch floats (partial rows %d x H %d) but the built-in "
             "array holds %zu; call b2_rmsnorm_bwd with caller-provided scratch instead",
             (size_t)nb * (size_t)H, nb, H, (size_t)kStaticScratchFloats);
  float* partial = nullptr;
  if (int rc = static_scratch_for_current_device(&partial)) return rc;
  const int grid = rms_bwd_blocks(T, H);  // same launch as b2_rmsnorm_bwd (<= nb partial rows)
#define B2_RMS_BWD(MV)                                                                     \
  rmsnorm_bwd_kernel<MV><<<grid, kNormThreads, 0, stream>>>(                                 \
      (bf16*)dx, partial, (const bf16*)dy, (const bf16*)h, (const bf16*)w, rstd,           \
      (const bf16*)dres, T, H)
  if (H <= kNormThreads * 8 * 2) B2_RMS_BWD(2);
  else if (H <= kNormThreads * 8 * 4) B2_RMS_BWD(4);
  else B2_RMS_BWD(8);
#undef B2_RMS_BWD
  B2_CUDA_CHECK(cudaGetLastError());
  fold_partials_kernel<<<(H + 31) / 32, 32 * kFoldSlices, 0, stream>>>((bf16*)dw, partial, grid, H,
                                                          accumulate_dw);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
