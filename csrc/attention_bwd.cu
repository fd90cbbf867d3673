// Causal flash-attention backward for sm_100a on tcgen05 tensor cores with TMEM accumulators.
//
// One CTA per (document, kv head, 64-row kv block j). It keeps K_j and V_j in shared memory and walks
// every (query head of the GQA group, 128-row query block i that can see block j). Per step:
//     S    = Q_i  K_j^T            [128 x 64]   TMEM, double buffered
//     dP   = dO_i V_j^T            [128 x 64]   TMEM, double buffered
//     P    = exp2(S*c - lse), dS = P * (dP - delta) * scale        (softmax warpgroup, registers)
//     dV^T += dO_i^T P             [D x 64]     TMEM accumulator over all steps
//     dK^T += Q_i^T  dS            [D x 64]     TMEM accumulator over all steps
//     dQ_i  = dS K_j               [128 x D]    TMEM -> smem -> cp.reduce.async.bulk.tensor (fp32 add) into dq_accum
// dV and dK are accumulated transposed so that every MMA has M = 128 (full tensor-core rate); the same
// 128B-swizzled smem tile serves as a K-major operand ([rows, 64] atoms) and as the MN-major operand of
// the transposed product, so Q, dO, K, P and dS are each stored once.
//   warp 0     TMA producer (K_j, V_j once; 2-stage ring of Q_i / dO_i)
//   warp 1     MMA issuer
//   warps 2-5  softmax warpgroup A (one thread per query row, kv columns 0..31)
//   warps 6-9  dQ warpgroup (TMEM -> swizzled smem -> TMA reduce-add into dq_accum), drains dK^T at the end
//   warps 10-13 softmax warpgroup B (kv columns 32..63): the exp2 / dS arithmetic is the longest stage
//              of the per-step chain, two warps per scheduler halve it
//
// Replaces flash_attn_2_cuda's backward (reference: 05-training-llama-405b/train_llm.py:87 +
// loss.backward() at :237).
#include "common.cuh"
#include <math.h>

namespace b2 {

constexpr int QB = 128;  // query rows per step
constexpr int KB = 64;   // kv rows per CTA
constexpr int kBwdThreads = 448;

template <int D>
struct AttnBwdSmem {
  static constexpr int kK = KB * D * 2;
  static constexpr int kV = KB * D * 2;
  static constexpr int kQ = QB * D * 2;
  static constexpr int kDO = QB * D * 2;
  static constexpr int kP = QB * KB * 2;   // 16 KB
  static constexpr int kDS = QB * KB * 2;  // 16 KB
  static constexpr int kDQStage = QB * 32 * 4;  // 128 rows x 32 fp32 columns = 16 KB
  static constexpr int kTotal = kK + kV + 2 * (kQ + kDO) + (kP + kDS) + 2 * kDQStage + 256 + 1024;
};

struct AttnBwdParams {
  float* dq_accum;  // [T, Hq*D] fp32, zeroed
  __nv_bfloat16* dk;
  __nv_bfloat16* dv;
  const float* lse;    // [Hq, T]
  const float* delta;  // [Hq, T]
  const int* cu_seqlens;
  int dk_stride, dv_stride;
  int T, Hq, Hkv, ndocs;
  float scale, scale_log2;
  int causal;
  AttnDropout drop;   // attention-probability dropout of the forward (thr == 0: none)
  int n_items = 0;      // attn_bwd2_kernel<2>: work items (kv blocks x kv heads x documents) ...
  int split_items = 0;  // ... of which the first (heaviest) split_items are shared by a CTA pair
};

__global__ void attn_dropout_mask_kernel(unsigned char* keep, int S, int doc, AttnDropout d) {
  const int g = blockIdx.x, q = blockIdx.y, h = blockIdx.z, i = threadIdx.x;
  const uint4 bits = attn_dropout_bits(d, q, g, h, doc);
  const int k = g * 16 + i;
  if (k < S) keep[((size_t)h * S + q) * S + k] = dropout_keep(bits, i, d.thr) ? 1 : 0;
}

// delta[h, t] = sum_d dO[t, h, d] * O[t, h, d]
// D / 8 lanes per (t, h) row, one 16-byte load of each tensor per lane, four rows in flight per lane
// group (the first version used a whole warp per row with 4-byte loads and ran at a third of the
// memory bandwidth).
template <int D>
__global__ void __launch_bounds__(256)
attn_bwd_delta_kernel(float* __restrict__ delta, const __nv_bfloat16* __restrict__ dO,
                      const __nv_bfloat16* __restrict__ O, int T, int Hq, int do_stride,
                      int o_stride) {
  constexpr int kLanes = D / 8;                 // lanes per row: 16 (D = 128) or 8 (D = 64)
  constexpr int kRowsPerBlock = 256 / kLanes;   // rows handled concurrently by a block
  constexpr int kUnroll = 4;
  const int sub = threadIdx.x % kLanes;
  const int grp = threadIdx.x / kLanes;
  const int total = T * Hq;  // rows (checked on the host to fit 31 bits)
  for (int base = blockIdx.x * kRowsPerBlock * kUnroll; base < total;
       base += gridDim.x * kRowsPerBlock * kUnroll) {
    uint4 a[kUnroll], b[kUnroll];
    int rows[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      rows[u] = base + u * kRowsPerBlock + grp;
      if (rows[u] < total) {
        const int t = rows[u] / Hq, h = rows[u] - t * Hq;
        a[u] = *reinterpret_cast<const uint4*>(dO + (size_t)t * do_stride + (size_t)h * D + sub * 8);
        b[u] = *reinterpret_cast<const uint4*>(O + (size_t)t * o_stride + (size_t)h * D + sub * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      float s = 0.f;
      if (rows[u] < total) {
        const uint32_t av[4] = {a[u].x, a[u].y, a[u].z, a[u].w}, bv[4] = {b[u].x, b[u].y, b[u].z, b[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 x = unpack_bf16(av[j]), y = unpack_bf16(bv[j]);
          s += x.x * y.x + x.y * y.y;
        }
      }
#pragma unroll
      for (int o = kLanes / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (sub == 0 && rows[u] < total) {
        const int t = rows[u] / Hq, h = rows[u] - t * Hq;
        delta[(size_t)h * T + t] = s;
      }
    }
  }
}

// dq[t, :] = bf16(dq_accum[t, :])
__global__ void attn_bwd_convert_dq_kernel(__nv_bfloat16* __restrict__ dq,
                                           const float* __restrict__ acc, int T, int cols,
                                           int dq_stride, float scale) {
  const int nvec = cols >> 2;
  const size_t total = (size_t)T * nvec;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t t = i / nvec;
    const int c = (int)(i - t * nvec);
    const float4 v = reinterpret_cast<const float4*>(acc + t * cols)[c];
    uint2 o;
    o.x = pack_bf16(v.x * scale, v.y * scale);
    o.y = pack_bf16(v.z * scale, v.w * scale);
    *reinterpret_cast<uint2*>(dq + t * (size_t)dq_stride + c * 4) = o;
  }
}

// smem tile -> global tensor, element-wise fp32 add performed by the TMA unit / L2
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* src, int c0,
                                                  int c1) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::
          "l"(reinterpret_cast<uint64_t>(m)),
      "r"(smem_u32(src)), "r"(c0), "r"(c1)
      : "memory");
}

template <int D, bool kDrop = false>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                const __grid_constant__ CUtensorMap tmDQ, AttnBwdParams p) {
  using SM = AttnBwdSmem<D>;
  constexpr int kAtoms = D / 64;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + SM::kK;
  uint8_t* sQ = sV + SM::kV;            // 2 stages
  uint8_t* sDO = sQ + 2 * SM::kQ;       // 2 stages
  uint8_t* sP = sDO + 2 * SM::kDO;
  uint8_t* sDS = sP + SM::kP;
  uint8_t* sDQ = sDS + SM::kDS;         // 2 staging buffers for the TMA reduce-add of dQ
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDQ + 2 * SM::kDQStage);
  uint64_t* kv_full = bars;            // 1
  uint64_t* qdo_full = bars + 1;       // 2
  uint64_t* qdo_empty = bars + 3;      // 2
  uint64_t* sdp_full = bars + 5;       // 2
  uint64_t* sdp_free = bars + 7;       // 2
  uint64_t* pds_full = bars + 9;       // 1 (+1 unused)
  uint64_t* pds_empty = bars + 11;     // 1 (+1 unused)
  uint64_t* dq_full = bars + 13;       // 1
  uint64_t* dq_free = bars + 14;       // 1
  uint64_t* done_bar = bars + 15;      // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5;
  // linear block id -> (kv block, kv head, document) with the kv block slowest: causal blocks with a
  // small index do the most work and are launched first
  const int per_kb = p.Hkv * p.ndocs;
  const int jb = (int)blockIdx.x / per_kb;
  const int kvh = ((int)blockIdx.x % per_kb) % p.Hkv;
  const int doc = ((int)blockIdx.x % per_kb) / p.Hkv;
  const int doc_start = p.cu_seqlens[doc];
  const int doc_len = p.cu_seqlens[doc + 1] - doc_start;
  const int num_kb = (doc_len + KB - 1) / KB;
  if (jb >= num_kb) return;
  const int kv0 = jb * KB;
  const int num_qb = (doc_len + QB - 1) / QB;
  const int i0 = p.causal ? kv0 / QB : 0;   // first query block that sees this kv block
  const int nqi = num_qb - i0;
  const int group = p.Hq / p.Hkv;
  const int n_iter = group * nqi;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDQ);
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
      mbar_init(&sdp_full[i], 1);
      mbar_init(&sdp_free[i], 8);   // 4 warps of each softmax warpgroup
      mbar_init(&pds_full[i], 8);
      mbar_init(&pds_empty[i], 1);
    }
    mbar_init(dq_full, 1);
    mbar_init(dq_free, 4);
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<1>(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tm_S[2] = {tmem_base, tmem_base + 64};
  const uint32_t tm_dP[2] = {tmem_base + 128, tmem_base + 192};
  const uint32_t tm_dQ = tmem_base + 256;
  const uint32_t tm_dVt = tmem_base + 384;
  const uint32_t tm_dKt = tmem_base + 448;

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {
      mbar_expect_tx(kv_full, SM::kK + SM::kV);
#pragma unroll
      for (int a = 0; a < kAtoms; ++a) {
        tma_load_2d<1>(&tmK, kv_full, sK + a * (KB * 128), kvh * D + a * 64, doc_start + kv0,
                       kEvictFirst);
        tma_load_2d<1>(&tmV, kv_full, sV + a * (KB * 128), kvh * D + a * 64, doc_start + kv0,
                       kEvictFirst);
      }
      for (int it = 0; it < n_iter; ++it) {
        const int s = it & 1;
        const int head = kvh * group + it / nqi;
        const int row = doc_start + (i0 + it % nqi) * QB;
        mbar_wait(&qdo_empty[s], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&qdo_full[s], SM::kQ + SM::kDO);
#pragma unroll
        for (int a = 0; a < kAtoms; ++a) {
          tma_load_2d<1>(&tmQ, &qdo_full[s], sQ + s * SM::kQ + a * (QB * 128), head * D + a * 64, row,
                         kEvictLast);
          tma_load_2d<1>(&tmDO, &qdo_full[s], sDO + s * SM::kDO + a * (QB * 128), head * D + a * 64,
                         row, kEvictLast);
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc_s = make_idesc_bf16(QB, KB, false, false);   // S, dP: K-major x K-major
    constexpr uint32_t idesc_t = make_idesc_bf16(128, KB, true, true);    // dV^T, dK^T: MN x MN
    constexpr uint32_t idesc_q = make_idesc_bf16(QB, D, false, true);     // dQ: K-major x MN-major
    // D = 64: the M = 128 transposed products read the single 64-row atom twice (LBO = 0); the
    // duplicate accumulator lanes 64..127 are never read back.
    constexpr uint32_t lbo_t = (D == 128) ? QB * 128 : 0;
    auto issue_sdp = [&](int it) {
      const int s = it & 1;
      mbar_wait(&qdo_full[s], (it >> 1) & 1);
      mbar_wait(&sdp_free[s], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t qa = smem_u32(sQ + s * SM::kQ), da = smem_u32(sDO + s * SM::kDO);
        const uint32_t ka = smem_u32(sK), va = smem_u32(sV);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t offa = (kk >> 2) * (QB * 128) + (kk & 3) * 32;
          const uint32_t offb = (kk >> 2) * (KB * 128) + (kk & 3) * 32;
          umma_f16<1>(tm_S[s], make_smem_desc_sw128(qa + offa, 16, 1024),
                      make_smem_desc_sw128(ka + offb, 16, 1024), idesc_s, kk != 0);
        }
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t offa = (kk >> 2) * (QB * 128) + (kk & 3) * 32;
          const uint32_t offb = (kk >> 2) * (KB * 128) + (kk & 3) * 32;
          umma_f16<1>(tm_dP[s], make_smem_desc_sw128(da + offa, 16, 1024),
                      make_smem_desc_sw128(va + offb, 16, 1024), idesc_s, kk != 0);
        }
        umma_commit<1>(&sdp_full[s]);
      }
      __syncwarp();
    };
    mbar_wait(kv_full, 0);
    if (n_iter > 0) issue_sdp(0);
    for (int it = 0; it < n_iter; ++it) {
      if (it + 1 < n_iter) issue_sdp(it + 1);
      const int s = it & 1;
      mbar_wait(&pds_full[0], it & 1);
      mbar_wait(dq_free, (it & 1) ^ 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t qa = smem_u32(sQ + s * SM::kQ), da = smem_u32(sDO + s * SM::kDO);
        const uint32_t pa = smem_u32(sP), dsa = smem_u32(sDS);
        const uint32_t ka = smem_u32(sK);
        // dV^T[D, kv] += dO^T[D, q] P[q, kv]   (reduction over the 128 query rows)
#pragma unroll
        for (int kk = 0; kk < QB / 16; ++kk)
          umma_f16<1>(tm_dVt, make_smem_desc_sw128(da + kk * 2048, lbo_t, 1024),
                      make_smem_desc_sw128(pa + kk * 2048, 16, 1024), idesc_t, (it | kk) != 0);
        // dK^T[D, kv] += Q^T[D, q] dS[q, kv]
#pragma unroll
        for (int kk = 0; kk < QB / 16; ++kk)
          umma_f16<1>(tm_dKt, make_smem_desc_sw128(qa + kk * 2048, lbo_t, 1024),
                      make_smem_desc_sw128(dsa + kk * 2048, 16, 1024), idesc_t, (it | kk) != 0);
        // dQ[q, D] = dS[q, kv] K[kv, D]        (reduction over the 64 kv rows)
#pragma unroll
        for (int kk = 0; kk < KB / 16; ++kk)
          umma_f16<1>(tm_dQ, make_smem_desc_sw128(dsa + kk * 32, 16, 1024),
                      make_smem_desc_sw128(ka + kk * 2048, KB * 128, 1024), idesc_q, kk != 0);
        umma_commit<1>(&qdo_empty[s]);
        umma_commit<1>(&pds_empty[0]);
        umma_commit<1>(dq_full);
        if (it == n_iter - 1) umma_commit<1>(done_bar);
      }
      __syncwarp();
    }
  } else if (warp < 6 || warp >= 10) {
    // ================= softmax warpgroups (A: columns 0..31, B: columns 32..63) =================
    const int hf = warp >= 10 ? 1 : 0;
    const int quarter = warp & 3;
    const int row = quarter * 32 + (int)lane_id();
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    for (int it = 0; it < n_iter; ++it) {
      const int s = it & 1;
      const int head = kvh * group + it / nqi;
      const int q_idx = (i0 + it % nqi) * QB + row;
      const bool valid_row = q_idx < doc_len;
      float lse2 = 0.f, delta = 0.f;
      if (valid_row) {
        lse2 = p.lse[(size_t)head * p.T + doc_start + q_idx] * 1.4426950408889634f;
        delta = p.delta[(size_t)head * p.T + doc_start + q_idx];
      }
      mbar_wait(&sdp_full[s], (it >> 1) & 1);
      tc_fence_after();
      // last kv index this query row may attend
      const int lim = valid_row ? (p.causal ? min(q_idx, doc_len - 1) : doc_len - 1) : -1;
      const uint32_t prow = smem_u32(sP) + row * 128;
      const uint32_t drow = smem_u32(sDS) + row * 128;
      {
        uint32_t sr[32], dr[32];
        tmem_ld_32x32b_x32(tm_S[s] + lane_off + hf * 32, sr);
        tmem_ld_32x32b_x32(tm_dP[s] + lane_off + hf * 32, dr);
        tmem_ld_wait();
        {  // this warp's share is in registers: the S / dP buffers may be refilled once all 8 arrive
          tc_fence_before();
          __syncwarp();
          if (lane_id() == 0) mbar_arrive(&sdp_free[s]);
        }
        uint32_t pp[16], dd[16];
        uint4 bits0, bits1;
        if constexpr (kDrop) {  // the forward's keep decisions for key positions kv0 + 32 hf + [0, 16), + [16, 32)
          const int g = (kv0 + hf * 32) >> 4;
          bits0 = attn_dropout_bits(p.drop, q_idx, g, head, doc);
          bits1 = attn_dropout_bits(p.drop, q_idx, g + 1, head, doc);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float pv[2], dv[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int c = 2 * i + e;
            const bool on = (kv0 + hf * 32 + c) <= lim;
            const float pe = on ? fast_exp2(__uint_as_float(sr[c]) * p.scale_log2 - lse2) : 0.f;
            if constexpr (kDrop) {
              // O = (P o M / (1 - p)) V: dV sees the dropped, rescaled P; dS = P o (dP o M / (1 - p) - delta)
              const bool keep = dropout_keep(c < 16 ? bits0 : bits1, c & 15, p.drop.thr);
              pv[e] = keep ? pe * p.drop.scale : 0.f;
              dv[e] = pe * ((keep ? __uint_as_float(dr[c]) * p.drop.scale : 0.f) - delta) * p.scale;
            } else {
              pv[e] = pe;
              dv[e] = pe * (__uint_as_float(dr[c]) - delta) * p.scale;
            }
          }
          pp[i] = pack_bf16(pv[0], pv[1]);
          dd[i] = pack_bf16(dv[0], dv[1]);
        }
        mbar_wait(&pds_empty[0], (it & 1) ^ 1);  // previous step's MMAs finished reading P / dS
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t off = ((hf * 4 + c) ^ (row & 7)) << 4;
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(prow + off), "r"(pp[4 * c]),
                       "r"(pp[4 * c + 1]), "r"(pp[4 * c + 2]), "r"(pp[4 * c + 3])
                       : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(drow + off), "r"(dd[4 * c]),
                       "r"(dd[4 * c + 1]), "r"(dd[4 * c + 2]), "r"(dd[4 * c + 3])
                       : "memory");
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&pds_full[0]);
    }
    // ---- drain dV^T: lane = head-dim index d, columns = kv rows ----
    if (n_iter > 0) {
      mbar_wait(done_bar, 0);
      tc_fence_after();
      const int d = row;
      {
        const int c = hf;  // each warpgroup drains its own 32 kv columns
        uint32_t r[32];
        tmem_ld_32x32b_x32(tm_dVt + lane_off + c * 32, r);
        tmem_ld_wait();
        if (d < D) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int kv = kv0 + c * 32 + i;
            if (kv < doc_len)
              p.dv[(size_t)(doc_start + kv) * p.dv_stride + (size_t)kvh * D + d] =
                  __float2bfloat16_rn(__uint_as_float(r[i]));
          }
        }
      }
      tc_fence_before();
    }
  } else {
    // ================= dQ warpgroup =================
    const int quarter = warp & 3;
    const int row = quarter * 32 + (int)lane_id();
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const bool dq_leader = threadIdx.x == 6 * 32;
    uint32_t stage_ctr = 0;
    for (int it = 0; it < n_iter; ++it) {
      const int head = kvh * group + it / nqi;
      mbar_wait(dq_full, it & 1);
      tc_fence_after();
      const int q_row0 = doc_start + (i0 + it % nqi) * QB;
#pragma unroll 1
      for (int c = 0; c < D / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tm_dQ + lane_off + c * 32, r);
        tmem_ld_wait();
        if (c == D / 32 - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane_id() == 0) mbar_arrive(dq_free);
        }
        // stage 128 rows x 32 fp32 (128 B per row, 128B swizzle) and add it into dq_accum with one
        // TMA reduce; rows past the document end hold exact zeros (P = dS = 0 there)
        uint8_t* sb = sDQ + (stage_ctr & 1) * SM::kDQStage;
        if (dq_leader) tma_store_wait_read<1>();
        named_bar_sync(2, 128);
        const uint32_t row_addr = smem_u32(sb) + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t addr = row_addr + ((j ^ (row & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(r[4 * j]),
                       "r"(r[4 * j + 1]), "r"(r[4 * j + 2]), "r"(r[4 * j + 3])
                       : "memory");
        }
        fence_proxy_async_smem();
        named_bar_sync(2, 128);
        if (dq_leader) {
          tma_reduce_add_2d(&tmDQ, sb, head * D + c * 32, q_row0);
          tma_store_commit();
        }
        ++stage_ctr;
      }
    }
    if (dq_leader) tma_store_wait<0>();
    // ---- drain dK^T ----
    if (n_iter > 0) {
      mbar_wait(done_bar, 0);
      tc_fence_after();
      const int d = row;
#pragma unroll
      for (int c = 0; c < KB / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tm_dKt + lane_off + c * 32, r);
        tmem_ld_wait();
        if (d < D) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int kv = kv0 + c * 32 + i;
            if (kv < doc_len)
              p.dk[(size_t)(doc_start + kv) * p.dk_stride + (size_t)kvh * D + d] =
                  __float2bfloat16_rn(__uint_as_float(r[i]));
          }
        }
      }
      tc_fence_before();
    }
  }

  __syncwarp();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}


// ------------------------------------------------------------------------------------------------
// Second-generation backward (head_dim 128): transposed scores with P^T / dS^T kept in tensor memory.
//
// One CTA per (document, kv head, 128-row kv block j); a step is one (query head, 64-row query block i).
//     S^T  = K_j  Q_i^T            [128 kv x 64 q]   TMEM, double buffered
//     dP^T = V_j  dO_i^T           [128 kv x 64 q]   TMEM, double buffered
//     P^T  = exp2(S^T c - lse_q),  dS^T = P^T (dP^T - delta_q) scale     (lane = kv row)
//     dV  += P^T  dO_i             A operand read from TMEM (P^T written back over S^T as bf16)
//     dK  += dS^T Q_i              A operand read from TMEM (dS^T next to P^T)
//     dQ^T = K_j^T dS              [128 d x 64 q] into the columns dP^T occupied; dS^T also goes to smem
// Against the first kernel (S = Q K^T, 64-row kv block, P and dS through shared memory) a unit of work
// moves ~1.5x fewer shared-memory bytes: no P tile, half the dQ staging (one reduce per 128 kv rows
// instead of per 64), half the Q / dO re-loads, and two of the five MMAs read A from TMEM. That
// kernel was bound by shared-memory bandwidth (ncu: tensor pipe 30 % active).
//   warp 0       TMA producer        warp 1        MMA issuer
//   warps 2-5    softmax warpgroup A (q columns 0..31)     warps 10-13  softmax warpgroup B (32..63)
//   warps 6-9    dQ^T drain: warp w owns head-dim lanes 32w..32w+31 = one [64 q x 32 d] fp32 box per
//                step, staged in its private buffer and added into dq_accum by its own TMA reduce
constexpr int QB2 = 64;    // query rows per step
constexpr int KB2 = 128;   // kv rows per CTA
struct AttnBwd2Smem {
  static constexpr int D = 128;
  static constexpr int kK = KB2 * D * 2;        // 32 KB
  static constexpr int kV = KB2 * D * 2;        // 32 KB
  static constexpr int kQ = QB2 * D * 2;        // 16 KB per stage
  static constexpr int kDO = QB2 * D * 2;       // 16 KB per stage
  static constexpr int kDS = KB2 * QB2 * 2;     // 16 KB per stage
  static constexpr int kDQBox = QB2 * 32 * 4;   // 8 KB: one warp's [64 q x 32 d] fp32 box
  static constexpr int kStats = 2 * 2 * 64 * 4; // [warpgroup][stage][lse2 x32 | delta x32] fp32 = 1 KB
  static constexpr int kTotal = kK + kV + 2 * (kQ + kDO + kDS) + 4 * 2 * kDQBox + kStats + 256 + 1024;
};

// kSplit == 2: the grid is made of CTA pairs (clusters of 2). The first p.split_items work items are SHARED by
// a pair, each CTA taking half of the query heads of the kv head's group; any remaining items go one per CTA
// as in the plain kernel (the launcher shares all items or none; the mixed form exists for experiments and
// measured slower than both). Sharing pays when whole items cannot fill or balance 148 SMs: few kv heads per
// rank under tensor parallelism (Llama-70B with tp=4 at S=4096 has 64 items), short sequences, or large
// groups whose heaviest item (causal: kv block 0) is longer than the per-SM average. dK / dV are sums over
// all query heads of the group: in a shared item CTA 1 hands its partial sums to CTA 0 through distributed
// shared memory at the end (128 KB, once per item), and CTA 0 adds them to its own while draining to global
// memory. Every additional CTA costs about 16 steps of turnover (launch, TMEM allocation, pipeline fill,
// drain), which the launcher's cost model weighs against the better balance.
template <int kSplit>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                 const __grid_constant__ CUtensorMap tmDQ, AttnBwdParams p) {
  using SM = AttnBwd2Smem;
  constexpr int D = 128;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + SM::kK;
  uint8_t* sQ = sV + SM::kV;             // 2 stages
  uint8_t* sDO = sQ + 2 * SM::kQ;        // 2 stages
  uint8_t* sDS = sDO + 2 * SM::kDO;      // 2 stages
  uint8_t* sDQ = sDS + 2 * SM::kDS;      // [4 warps][2] boxes
  float* sStats = reinterpret_cast<float*>(sDQ + 8 * SM::kDQBox);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDQ + 8 * SM::kDQBox + SM::kStats);
  uint64_t* kv_full = bars;            // 1
  uint64_t* qdo_full = bars + 1;       // 2
  uint64_t* qdo_empty = bars + 3;      // 2
  uint64_t* s_full = bars + 5;         // 2  S^T of a step is in TMEM
  uint64_t* p_full = bars + 7;         // 2  P^T written back (8 warps)
  uint64_t* ds_full = bars + 9;        // 2  dS^T in TMEM and smem (8 warps)
  uint64_t* dq_full = bars + 11;       // 2  dQ^T of a step is in TMEM
  uint64_t* dq_free = bars + 13;       // 2  ... and has been read out (4 warps)
  uint64_t* done_bar = bars + 15;      // 1
  uint64_t* dp_full = bars + 16;       // 2  dP^T of a step is in TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);

  const int warp = threadIdx.x >> 5;
  const int per_kb = p.Hkv * p.ndocs;
  const int crank = kSplit == 2 ? (int)cluster_ctarank() : 0;
  int item = (int)blockIdx.x;
  bool shared_item = false;
  if constexpr (kSplit == 2) {
    const int cl = (int)blockIdx.x >> 1;
    shared_item = cl < p.split_items;
    item = shared_item ? cl : p.split_items + 2 * (cl - p.split_items) + crank;
    if (item >= p.n_items) return;   // odd number of whole items: the last pair has one idle CTA
  }
  const int jb = item / per_kb;
  const int kvh = (item % per_kb) % p.Hkv;
  const int doc = (item % per_kb) / p.Hkv;
  const int doc_start = p.cu_seqlens[doc];
  const int doc_len = p.cu_seqlens[doc + 1] - doc_start;
  const int num_kb = (doc_len + KB2 - 1) / KB2;
  if (jb >= num_kb) return;   // the CTAs of a shared item leave together; the others never meet
  const int kv0 = jb * KB2;
  const int num_qb = (doc_len + QB2 - 1) / QB2;
  const int i0 = p.causal ? kv0 / QB2 : 0;  // first query block with a row >= kv0
  const int nqi = num_qb - i0;
  const int group = p.Hq / p.Hkv;
  const int my_heads = shared_item ? group / 2 : group;   // query heads this CTA walks
  const int head0 = kvh * group + (shared_item ? crank * my_heads : 0);
  const int n_iter = my_heads * nqi;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDQ);
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&dp_full[i], 1);
      mbar_init(&p_full[i], 8);
      mbar_init(&ds_full[i], 8);
      mbar_init(&dq_full[i], 1);
      mbar_init(&dq_free[i], 4);
    }
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<1>(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // columns: [0,128) S^T x2 (64 each; later P^T at +0/+32 and dS^T at +16/+48, bf16 pairs),
  //          [128,256) dP^T x2 (64 each; later dQ^T), [256,384) dV, [384,512) dK
  const uint32_t tm_S[2] = {tmem_base, tmem_base + 64};
  const uint32_t tm_dP[2] = {tmem_base + 128, tmem_base + 192};
  const uint32_t tm_dV = tmem_base + 256;
  const uint32_t tm_dK = tmem_base + 384;

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {
      mbar_expect_tx(kv_full, SM::kK + SM::kV);
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        tma_load_2d<1>(&tmK, kv_full, sK + a * (KB2 * 128), kvh * D + a * 64, doc_start + kv0, kEvictFirst);
        tma_load_2d<1>(&tmV, kv_full, sV + a * (KB2 * 128), kvh * D + a * 64, doc_start + kv0, kEvictFirst);
      }
      for (int it = 0; it < n_iter; ++it) {
        const int s = it & 1;
        const int head = head0 + it / nqi;
        const int row = doc_start + (i0 + it % nqi) * QB2;
        mbar_wait(&qdo_empty[s], ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(&qdo_full[s], SM::kQ + SM::kDO);
#pragma unroll
        for (int a = 0; a < 2; ++a) {
          tma_load_2d<1>(&tmQ, &qdo_full[s], sQ + s * SM::kQ + a * (QB2 * 128), head * D + a * 64, row, kEvictLast);
          tma_load_2d<1>(&tmDO, &qdo_full[s], sDO + s * SM::kDO + a * (QB2 * 128), head * D + a * 64, row, kEvictLast);
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc_st = make_idesc_bf16(KB2, QB2, false, false);  // S^T, dP^T: K-major x K-major
    constexpr uint32_t idesc_dv = make_idesc_bf16(KB2, D, false, true);     // dV, dK: TMEM A x MN-major B
    constexpr uint32_t idesc_dq = make_idesc_bf16(D, QB2, true, true);      // dQ^T: MN-major x MN-major
    // S^T of the next step goes out as early as possible (its columns are free once the previous
    // user's dV / dK were issued); dP^T waits for the dQ^T that lived in its columns to be drained
    auto issue_s = [&](int it) {
      const int s = it & 1;
      mbar_wait(&qdo_full[s], (it >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t qa = smem_u32(sQ + s * SM::kQ), ka = smem_u32(sK);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t offa = (kk >> 2) * (KB2 * 128) + (kk & 3) * 32;
          const uint32_t offb = (kk >> 2) * (QB2 * 128) + (kk & 3) * 32;
          umma_f16<1>(tm_S[s], make_smem_desc_sw128(ka + offa, 16, 1024),
                      make_smem_desc_sw128(qa + offb, 16, 1024), idesc_st, kk != 0);
        }
        umma_commit<1>(&s_full[s]);
      }
      __syncwarp();
    };
    auto issue_dp = [&](int it) {
      const int s = it & 1;
      mbar_wait(&dq_free[s], ((it >> 1) & 1) ^ 1);  // dQ^T of step it-2 has left the dP^T columns
      tc_fence_after();
      if (elect_one()) {
        const uint32_t da = smem_u32(sDO + s * SM::kDO), va = smem_u32(sV);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t offa = (kk >> 2) * (KB2 * 128) + (kk & 3) * 32;
          const uint32_t offb = (kk >> 2) * (QB2 * 128) + (kk & 3) * 32;
          umma_f16<1>(tm_dP[s], make_smem_desc_sw128(va + offa, 16, 1024),
                      make_smem_desc_sw128(da + offb, 16, 1024), idesc_st, kk != 0);
        }
        umma_commit<1>(&dp_full[s]);
      }
      __syncwarp();
    };
    mbar_wait(kv_full, 0);
    if (n_iter > 0) {
      issue_s(0);
      issue_dp(0);
    }
    for (int it = 0; it < n_iter; ++it) {
      if (it + 1 < n_iter) issue_s(it + 1);
      const int s = it & 1;
      const uint32_t par = (it >> 1) & 1;
      const uint32_t qa = smem_u32(sQ + s * SM::kQ), da = smem_u32(sDO + s * SM::kDO);
      mbar_wait(&p_full[s], par);
      tc_fence_after();
      if (elect_one()) {
        // dV[kv, d] += P^T[kv, q] dO[q, d]: K = 64 query rows, 16 per instruction = 8 TMEM columns
#pragma unroll
        for (int kk = 0; kk < QB2 / 16; ++kk)
          umma_f16_ts(tm_dV, tm_S[s] + (kk >> 1) * 32 + (kk & 1) * 8,
                      make_smem_desc_sw128(da + kk * 2048, QB2 * 128, 1024), idesc_dv, (it | kk) != 0);
      }
      __syncwarp();
      if (it + 1 < n_iter) issue_dp(it + 1);
      mbar_wait(&ds_full[s], par);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t ka = smem_u32(sK), dsa = smem_u32(sDS + s * SM::kDS);
#pragma unroll
        for (int kk = 0; kk < QB2 / 16; ++kk)
          umma_f16_ts(tm_dK, tm_S[s] + 16 + (kk >> 1) * 32 + (kk & 1) * 8,
                      make_smem_desc_sw128(qa + kk * 2048, QB2 * 128, 1024), idesc_dv, (it | kk) != 0);
        // dQ^T[d, q] = K^T[d, kv] dS[kv, q]: reduction over the 128 kv rows
#pragma unroll
        for (int kk = 0; kk < KB2 / 16; ++kk)
          umma_f16<1>(tm_dP[s], make_smem_desc_sw128(ka + kk * 2048, KB2 * 128, 1024),
                      make_smem_desc_sw128(dsa + kk * 2048, KB2 * 128, 1024), idesc_dq, kk != 0);
        umma_commit<1>(&qdo_empty[s]);
        umma_commit<1>(&dq_full[s]);
        if (it == n_iter - 1) umma_commit<1>(done_bar);
      }
      __syncwarp();
    }
  } else if (warp < 6 || warp >= 10) {
    // ================= softmax warpgroups: lane = kv row, h selects 32 of the 64 q columns =========
    const int h = warp >= 10 ? 1 : 0;
    const int quarter = warp & 3;
    const int row = quarter * 32 + (int)lane_id();   // kv row inside the block
    const int kv_idx = kv0 + row;
    const bool kv_ok = kv_idx < doc_len;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    for (int it = 0; it < n_iter; ++it) {
      const int s = it & 1;
      const uint32_t par = (it >> 1) & 1;
      const int head = head0 + it / nqi;
      const int q0 = (i0 + it % nqi) * QB2 + 32 * h;   // first query row of this warpgroup's columns
      // lane j fetches the row statistics of query q0 + j; after the wait below they go to a small
      // shared table (one per warpgroup and stage; the four warps store identical values, and a warp
      // only reads after its own stores) that every lane reads back with broadcast 16-byte loads
      float lse2_l = 0.f, delta_l = 0.f;
      if (q0 + (int)lane_id() < doc_len) {
        const size_t idx = (size_t)head * p.T + doc_start + q0 + lane_id();
        lse2_l = p.lse[idx] * 1.4426950408889634f;
        delta_l = p.delta[idx];
      }
      const int n_q = min(32, doc_len - q0);           // valid columns
      // column c is visible iff q0 + c >= kv_idx (causal) and c < n_q
      const int c_lo = p.causal ? max(kv_idx - q0, 0) : 0;
      const bool all_on = __all_sync(0xffffffffu, kv_ok && c_lo == 0 && n_q >= 32);
      mbar_wait(&s_full[s], par);
      tc_fence_after();
      // past this wait every warp has finished step it-2, the previous user of this table
      float* st = sStats + (h * 2 + s) * 64;
      st[lane_id()] = lse2_l;
      st[32 + lane_id()] = delta_l;
      __syncwarp();
      uint32_t sr[32];
      float pf[32];
      tmem_ld_32x32b_x32(tm_S[s] + lane_off + 32 * h, sr);
      tmem_ld_wait();
      uint32_t pk[16];
      if (all_on) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 l4 = reinterpret_cast<const float4*>(st)[i];
          pf[4 * i] = fast_exp2(fmaf(__uint_as_float(sr[4 * i]), p.scale_log2, -l4.x));
          pf[4 * i + 1] = fast_exp2(fmaf(__uint_as_float(sr[4 * i + 1]), p.scale_log2, -l4.y));
          pf[4 * i + 2] = fast_exp2(fmaf(__uint_as_float(sr[4 * i + 2]), p.scale_log2, -l4.z));
          pf[4 * i + 3] = fast_exp2(fmaf(__uint_as_float(sr[4 * i + 3]), p.scale_log2, -l4.w));
        }
      } else {
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const bool on = kv_ok && c >= c_lo && c < n_q;
          pf[c] = on ? fast_exp2(fmaf(__uint_as_float(sr[c]), p.scale_log2, -st[c])) : 0.f;
        }
      }
#pragma unroll
      for (int i = 0; i < 16; ++i) pk[i] = pack_bf16(pf[2 * i], pf[2 * i + 1]);
      tmem_st_32x32b_x16(tm_S[s] + lane_off + 32 * h, pk);       // P^T over this warpgroup's S^T columns
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&p_full[s]);
      // ---- dS^T = P^T (dP^T - delta); the softmax scale is applied by the dK drain and the dQ convert ----
      mbar_wait(&dp_full[s], par);
      tc_fence_after();
      tmem_ld_32x32b_x32(tm_dP[s] + lane_off + 32 * h, sr);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 d4 = reinterpret_cast<const float4*>(st + 32)[i];
        const float a0 = pf[4 * i] * (__uint_as_float(sr[4 * i]) - d4.x);
        const float a1 = pf[4 * i + 1] * (__uint_as_float(sr[4 * i + 1]) - d4.y);
        const float a2 = pf[4 * i + 2] * (__uint_as_float(sr[4 * i + 2]) - d4.z);
        const float a3 = pf[4 * i + 3] * (__uint_as_float(sr[4 * i + 3]) - d4.w);
        pk[2 * i] = pack_bf16(a0, a1);
        pk[2 * i + 1] = pack_bf16(a2, a3);
      }
      tmem_st_32x32b_x16(tm_S[s] + lane_off + 32 * h + 16, pk);  // dS^T next to P^T
      // the same values as a [kv rows][64 q] swizzled tile: B operand of dQ^T
      const uint32_t drow = smem_u32(sDS + s * SM::kDS) + row * 128;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint32_t off = ((h * 4 + c) ^ (row & 7)) << 4;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(drow + off), "r"(pk[4 * c]),
                     "r"(pk[4 * c + 1]), "r"(pk[4 * c + 2]), "r"(pk[4 * c + 3])
                     : "memory");
      }
      tmem_st_wait();
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&ds_full[s]);
    }
  } else {
    // ================= dQ^T drain: warp w owns head-dim lanes 32w .. 32w+31 =================
    const int w = warp & 3;   // warps 6..9 -> TMEM lane quarters 2,3,0,1
    const uint32_t lane_off = (uint32_t)(w * 32) << 16;
    uint8_t* my_box = sDQ + w * 2 * SM::kDQBox;
    const uint32_t chunk = lane_id() >> 2, sub = (lane_id() & 3) * 4;
    for (int it = 0; it < n_iter; ++it) {
      const int s = it & 1;
      const int head = head0 + it / nqi;
      const int q_row0 = doc_start + (i0 + it % nqi) * QB2;
      mbar_wait(&dq_full[s], (it >> 1) & 1);
      tc_fence_after();
      uint32_t r0[32], r1[32];
      tmem_ld_32x32b_x32(tm_dP[s] + lane_off, r0);
      tmem_ld_32x32b_x32(tm_dP[s] + lane_off + 32, r1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) {
        mbar_arrive(&dq_free[s]);
        tma_store_wait_read<1>();   // the reduce issued two steps ago has finished reading this box
      }
      __syncwarp();
      // box [64 q rows][32 d] fp32, 128 B rows, 128B swizzle; lane = d: one row per store instruction
      const uint32_t box = smem_u32(my_box + (it & 1) * SM::kDQBox);
#pragma unroll
      for (int q = 0; q < 32; ++q) {
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(box + q * 128 + ((chunk ^ (q & 7)) << 4) + sub), "r"(r0[q]) : "memory");
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(box + (q + 32) * 128 + ((chunk ^ (q & 7)) << 4) + sub), "r"(r1[q]) : "memory");
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane_id() == 0) {
        tma_reduce_add_2d(&tmDQ, my_box + (it & 1) * SM::kDQBox, head * D + w * 32, q_row0);
        tma_store_commit();
      }
    }
    if (lane_id() == 0) tma_store_wait<0>();
  }
  __syncwarp();

  // ================= drain dV (softmax warps) and dK (dQ warps): lane = kv row =================
  // The accumulators cover 32-column chunks `ch` of the two [128 kv x 128 d] tiles (dV: ch 0..3, dK: ch
  // 4..7). Split items: CTA 1 writes its chunks as fp32 into CTA 0's shared memory (the K / V / Q / dO
  // buffers, free once every MMA has completed), laid out [16-byte column group][kv row] so that the 32
  // lanes of a warp touch consecutive 16-byte words on both sides; CTA 0 adds them while it drains.
  const bool is_sm = (warp >= 2 && warp < 6) || warp >= 10;
  const bool is_dq = warp >= 6 && warp < 10;
  if (n_iter > 0 && (is_sm || is_dq)) {
    mbar_wait(done_bar, 0);
    tc_fence_after();
  }
  auto drain = [&](int mode) {   // 0: to global, 1: to CTA 0's shared memory, 2: add CTA 1's sums, to global
    if (n_iter == 0 || !(is_sm || is_dq)) return;
    const int quarter = warp & 3;   // TMEM lane quarter of this warp
    const int row = quarter * 32 + (int)lane_id();
    const int kv_idx = kv0 + row;
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const int c_lo = is_sm ? (warp >= 10 ? 2 : 0) : 0, c_n = is_sm ? 2 : 4;
    const uint32_t tm = is_sm ? tm_dV : tm_dK;
    const float sc = is_sm ? 1.f : p.scale;
    __nv_bfloat16* out = is_sm ? p.dv : p.dk;
    const int ostride = is_sm ? p.dv_stride : p.dk_stride;
    uint32_t xbase = smem_u32(sK);
    if (mode == 1) asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(xbase) : "r"(xbase));
#pragma unroll 1
    for (int c = c_lo; c < c_lo + c_n; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(tm + lane_off + c * 32, r);
      tmem_ld_wait();
      const uint32_t xoff = xbase + (uint32_t)((((is_sm ? 0 : 32) + c * 8) * KB2 + row) * 16);
      if (mode == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(xoff + j * (KB2 * 16)),
                       "r"(r[4 * j]), "r"(r[4 * j + 1]), "r"(r[4 * j + 2]), "r"(r[4 * j + 3])
                       : "memory");
        continue;
      }
      if (mode == 2) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint32_t a0, a1, a2, a3;
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3)
                       : "r"(xoff + j * (KB2 * 16))
                       : "memory");
          r[4 * j] = __float_as_uint(__uint_as_float(r[4 * j]) + __uint_as_float(a0));
          r[4 * j + 1] = __float_as_uint(__uint_as_float(r[4 * j + 1]) + __uint_as_float(a1));
          r[4 * j + 2] = __float_as_uint(__uint_as_float(r[4 * j + 2]) + __uint_as_float(a2));
          r[4 * j + 3] = __float_as_uint(__uint_as_float(r[4 * j + 3]) + __uint_as_float(a3));
        }
      }
      if (kv_idx < doc_len) {
        uint4* dst = reinterpret_cast<uint4*>(out + (size_t)(doc_start + kv_idx) * ostride + (size_t)kvh * D +
                                              c * 32);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 o;
          o.x = pack_bf16(__uint_as_float(r[8 * j]) * sc, __uint_as_float(r[8 * j + 1]) * sc);
          o.y = pack_bf16(__uint_as_float(r[8 * j + 2]) * sc, __uint_as_float(r[8 * j + 3]) * sc);
          o.z = pack_bf16(__uint_as_float(r[8 * j + 4]) * sc, __uint_as_float(r[8 * j + 5]) * sc);
          o.w = pack_bf16(__uint_as_float(r[8 * j + 6]) * sc, __uint_as_float(r[8 * j + 7]) * sc);
          dst[j] = o;
        }
      }
    }
    tc_fence_before();
  };
  if (kSplit == 2 && shared_item) {
    cluster_sync_all();   // every MMA of both CTAs has completed: CTA 0's operand buffers are free
    if (crank == 1) drain(1);
    cluster_sync_all();   // CTA 1's sums are in CTA 0's shared memory
    if (crank == 0) drain(2);
  } else {
    drain(0);
  }

  __syncwarp();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

static int launch_attn_bwd2(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                            const CUtensorMap& tmDO, const CUtensorMap& tmDQ, const AttnBwdParams& prm,
                            int max_kblocks, int ndocs, cudaStream_t stream) {
  using SM = AttnBwd2Smem;
  static bool attr_set = false;
  if (!attr_set) {
    B2_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd2_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       SM::kTotal));
    B2_CUDA_CHECK(cudaFuncSetAttribute(attn_bwd2_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       SM::kTotal));
    attr_set = true;
  }
  const int per_kb = prm.Hkv * ndocs;
  const int items = max_kblocks * per_kb;
  // Share every item between two CTAs, or none: the choice minimises the estimated makespan in steps,
  // max(average per SM, longest piece) + CTAs per SM x 16 (a CTA's turnover: launch, TMEM allocation, pipeline
  // fill, drain). Measured whole / shared, ms (profiles/attn_bwd_split_ab.jsonl): 8B S=4096 0.534 / 0.538,
  // S=2048 0.273 / 0.177, B=2 1.017 / 1.101; 70B S=4096 1.098 / 1.030, with tp=4 (64 items) 0.925 / 0.493; 8B
  // tp=8 B=16 S=1024 0.161 / 0.124 -- the model picks the faster one in each case. Sharing only the heaviest
  // kv blocks (B200TRAIN_ATTN_BWD_SPLIT_KB=n) measured slower than either: pairs of independent whole items
  // wait for each other's SM. B200TRAIN_ATTN_BWD_SPLIT=0 never shares, =1 always does.
  static const int force = [] {
    const char* e = getenv("B200TRAIN_ATTN_BWD_SPLIT");
    return e ? (e[0] == '1' ? 1 : 0) : -1;
  }();
  const int group = prm.Hq / prm.Hkv;
  int split_kb = 0;
  if (group % 2 == 0 && force != 0) {
    if (force == 1) {
      split_kb = max_kblocks;
    } else {
      const int J = max_kblocks, sms = sm_count();
      const int nq = J * (KB2 / QB2);                                  // query blocks of the longest document
      double total = 0;
      for (int j = 0; j < J; ++j) total += (double)group * (prm.causal ? nq - j * (KB2 / QB2) : nq) * per_kb;
      const double avg = total / sms, longest = (double)group * nq;
      const double whole = (avg > longest ? avg : longest) + 16.0 * items / sms;
      const double shared = (avg > longest / 2 ? avg : longest / 2) + 32.0 * items / sms;
      if (shared < 0.98 * whole) split_kb = J;
    }
    static const int fixed_kb = [] {   // B200TRAIN_ATTN_BWD_SPLIT_KB=n: share exactly the first n kv blocks
      const char* e = getenv("B200TRAIN_ATTN_BWD_SPLIT_KB");
      return e ? atoi(e) : -1;
    }();
    if (fixed_kb >= 0) split_kb = fixed_kb < max_kblocks ? fixed_kb : max_kblocks;
  }
  if (split_kb == 0) {
    attn_bwd2_kernel<1><<<dim3(items), kBwdThreads, SM::kTotal, stream>>>(tmQ, tmK, tmV, tmDO, tmDQ, prm);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
  }
  AttnBwdParams prm2 = prm;
  prm2.n_items = items;
  prm2.split_items = split_kb * per_kb;
  const int clusters = prm2.split_items + (items - prm2.split_items + 1) / 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(kBwdThreads);
  cfg.dynamicSmemBytes = SM::kTotal;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = 2;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B2_CUDA_CHECK(cudaLaunchKernelEx(&cfg, attn_bwd2_kernel<2>, tmQ, tmK, tmV, tmDO, tmDQ, prm2));
  return 0;
}

template <int D, bool kDrop = false>
static int launch_attn_bwd(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                           const CUtensorMap& tmDO, const CUtensorMap& tmDQ,
                           const AttnBwdParams& prm, int max_kblocks, int ndocs,
                           cudaStream_t stream) {
  using SM = AttnBwdSmem<D>;
  auto kern = attn_bwd_kernel<D, kDrop>;
  static bool attr_set = false;
  if (!attr_set) {
    B2_CUDA_CHECK(
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::kTotal));
    attr_set = true;
  }
  dim3 grid(max_kblocks * prm.Hkv * ndocs);
  kern<<<grid, kBwdThreads, SM::kTotal, stream>>>(tmQ, tmK, tmV, tmDO, tmDQ, prm);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b2

extern "C" {

// Gradients of b2_attn_fwd. dq/dk/dv are strided [T, heads*D] bf16 views (e.g. slices of one fused
// dQKV buffer). Scratch supplied by the caller: dq_accum fp32 [T, Hq*D] (zeroed here, contiguous),
// delta fp32 [Hq, T].
static int attn_bwd_impl(void* dq, void* dk, void* dv, float* dq_accum, float* delta, const void* dout,
                const void* q, const void* k, const void* v, const void* o, const float* lse,
                const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                int q_stride, int k_stride, int v_stride, int o_stride, int do_stride,
                int dq_stride, int dk_stride, int dv_stride, float softmax_scale, int causal,
                const b2::AttnDropout& drop, cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(D == 64 || D == 128, "attn bwd: head_dim %d unsupported (64 or 128)", D);
  B2_REQUIRE(Hq % Hkv == 0, "attn bwd: Hq=%d not a multiple of Hkv=%d", Hq, Hkv);
  B2_REQUIRE(dq_stride % 4 == 0, "attn bwd: dq stride must be a multiple of 4");
  if (T == 0 || ndocs == 0) return 0;
  B2_CUDA_CHECK(cudaMemsetAsync(dq_accum, 0, (size_t)T * Hq * D * sizeof(float), stream));
  {
    B2_REQUIRE(do_stride % 8 == 0 && o_stride % 8 == 0 && ((uintptr_t)dout & 15) == 0 && ((uintptr_t)o & 15) == 0,
               "attn bwd: dout / o rows must be 16-byte aligned (strides %d, %d)", do_stride, o_stride);
    B2_REQUIRE((long long)T * Hq < (1ll << 31), "attn bwd: T * Hq = %lld rows exceed 31 bits", (long long)T * Hq);
    const long long rows = (long long)T * Hq;
    const int rows_per_block = (256 / (D / 8)) * 4;
    long long blocks = (rows + rows_per_block - 1) / rows_per_block;
    const int cap = sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (D == 128)
      attn_bwd_delta_kernel<128><<<(int)blocks, 256, 0, stream>>>(
          delta, static_cast<const __nv_bfloat16*>(dout), static_cast<const __nv_bfloat16*>(o), T, Hq,
          do_stride, o_stride);
    else
      attn_bwd_delta_kernel<64><<<(int)blocks, 256, 0, stream>>>(
          delta, static_cast<const __nv_bfloat16*>(dout), static_cast<const __nv_bfloat16*>(o), T, Hq,
          do_stride, o_stride);
    B2_CUDA_CHECK(cudaGetLastError());
  }
  // head_dim 128 runs the second-generation kernel (B200TRAIN_ATTN_BWD_GEN=1 selects the first one; the
  // dk / dv rows are written with 16-byte stores there, so their strides must allow it)
  static const bool want_v2 = [] {
    const char* e = getenv("B200TRAIN_ATTN_BWD_GEN");
    return !(e && e[0] == '1');
  }();
  const bool v2 = want_v2 && drop.thr == 0 && D == 128 && dk_stride % 8 == 0 && dv_stride % 8 == 0 &&
                  (reinterpret_cast<uintptr_t>(dk) & 15) == 0 && (reinterpret_cast<uintptr_t>(dv) & 15) == 0;
  const int qb = v2 ? QB2 : QB, kb = v2 ? KB2 : KB;
  CUtensorMap tmQ, tmK, tmV, tmDO, tmDQ;
  int rc;
  if ((rc = make_tmap_2d_f32(&tmDQ, dq_accum, T, (uint64_t)Hq * D, (uint64_t)Hq * D, qb, 32))) return rc;
  if ((rc = make_tmap_2d_bf16(&tmQ, q, T, (uint64_t)Hq * D, q_stride, qb, 64))) return rc;
  if ((rc = make_tmap_2d_bf16(&tmDO, dout, T, (uint64_t)Hq * D, do_stride, qb, 64))) return rc;
  if ((rc = make_tmap_2d_bf16(&tmK, k, T, (uint64_t)Hkv * D, k_stride, kb, 64))) return rc;
  if ((rc = make_tmap_2d_bf16(&tmV, v, T, (uint64_t)Hkv * D, v_stride, kb, 64))) return rc;
  AttnBwdParams prm;
  prm.dq_accum = dq_accum;
  prm.dk = static_cast<__nv_bfloat16*>(dk);
  prm.dv = static_cast<__nv_bfloat16*>(dv);
  prm.lse = lse;
  prm.delta = delta;
  prm.cu_seqlens = cu_seqlens;
  prm.dk_stride = dk_stride;
  prm.dv_stride = dv_stride;
  prm.T = T;
  prm.Hq = Hq;
  prm.Hkv = Hkv;
  prm.ndocs = ndocs;
  prm.scale = softmax_scale;
  prm.scale_log2 = softmax_scale * 1.4426950408889634f;
  prm.causal = causal;
  prm.drop = drop;
  const int max_kb = (max_seqlen + kb - 1) / kb;
  if (v2) rc = launch_attn_bwd2(tmQ, tmK, tmV, tmDO, tmDQ, prm, max_kb, ndocs, stream);
  else if (drop.thr != 0 && D == 64) rc = launch_attn_bwd<64, true>(tmQ, tmK, tmV, tmDO, tmDQ, prm, max_kb, ndocs, stream);
  else if (drop.thr != 0) rc = launch_attn_bwd<128, true>(tmQ, tmK, tmV, tmDO, tmDQ, prm, max_kb, ndocs, stream);
  else if (D == 64) rc = launch_attn_bwd<64>(tmQ, tmK, tmV, tmDO, tmDQ, prm, max_kb, ndocs, stream);
  else rc = launch_attn_bwd<128>(tmQ, tmK, tmV, tmDO, tmDQ, prm, max_kb, ndocs, stream);
  if (rc) return rc;
  {
    const int cols = Hq * D;
    const size_t work = (size_t)T * (cols / 4);
    int blocks = (int)((work + 255) / 256);
    const int cap = sm_count() * 16;
    if (blocks > cap) blocks = cap;
    // the second-generation kernel leaves the softmax scale out of dS: it is applied here and in its
    // dK drain
    attn_bwd_convert_dq_kernel<<<blocks, 256, 0, stream>>>(static_cast<__nv_bfloat16*>(dq), dq_accum,
                                                           T, cols, dq_stride, v2 ? softmax_scale : 1.f);
    B2_CUDA_CHECK(cudaGetLastError());
  }
  return 0;
}

int b2_attn_bwd(void* dq, void* dk, void* dv, float* dq_accum, float* delta, const void* dout,
                const void* q, const void* k, const void* v, const void* o, const float* lse,
                const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                int q_stride, int k_stride, int v_stride, int o_stride, int do_stride,
                int dq_stride, int dk_stride, int dv_stride, float softmax_scale, int causal,
                cudaStream_t stream) {
  b2::AttnDropout none = {};
  return attn_bwd_impl(dq, dk, dv, dq_accum, delta, dout, q, k, v, o, lse, cu_seqlens, ndocs, max_seqlen, T, Hq,
                       Hkv, D, q_stride, k_stride, v_stride, o_stride, do_stride, dq_stride, dk_stride, dv_stride,
                       softmax_scale, causal, none, stream);
}

// Gradients of b2_attn_fwd_dropout: the same (drop_p, seed, offset) regenerate the forward's mask.
int b2_attn_bwd_dropout(void* dq, void* dk, void* dv, float* dq_accum, float* delta, const void* dout,
                        const void* q, const void* k, const void* v, const void* o, const float* lse,
                        const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                        int q_stride, int k_stride, int v_stride, int o_stride, int do_stride,
                        int dq_stride, int dk_stride, int dv_stride, float softmax_scale, int causal,
                        float drop_p, unsigned long long seed, unsigned long long offset,
                        cudaStream_t stream) {
  return attn_bwd_impl(dq, dk, dv, dq_accum, delta, dout, q, k, v, o, lse, cu_seqlens, ndocs, max_seqlen, T, Hq,
                       Hkv, D, q_stride, k_stride, v_stride, o_stride, do_stride, dq_stride, dk_stride, dv_stride,
                       softmax_scale, causal, b2::make_attn_dropout(drop_p, seed, offset), stream);
}

// Test helper: keep[h][q][k] (1 byte each) of one document of length S -- the mask b2_attn_fwd_dropout applies.
int b2_attn_dropout_mask(unsigned char* keep, int Hq, int S, int doc, float drop_p, unsigned long long seed,
                         unsigned long long offset, cudaStream_t stream) {
  b2::AttnDropout d = b2::make_attn_dropout(drop_p, seed, offset);
  dim3 grid((S + 15) / 16, S, Hq);
  b2::attn_dropout_mask_kernel<<<grid, 16, 0, stream>>>(keep, S, doc, d);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
