// Causal flash-attention forward for sm_100a on tcgen05 tensor cores with TMEM accumulators.
//
// One CTA per (document, query head, 128-row query block); 320 threads:
//   warp 0     TMA producer  : Q once, then a 2-stage ring of K (K-major) and V (MN-major) tiles
//   warp 1     MMA issuer    : S_j = Q K_j^T into one of two TMEM S buffers, O += P_j V_j into TMEM O
//   warps 2-9  softmax       : two warpgroups, each thread owns one query row x 64 of the block's 128 kv
//                              columns; tcgen05.ld S, online softmax in the exp2 domain (row max
//                              exchanged through smem) with lazy rescale of the TMEM O accumulator,
//                              P_j -> bf16 -> 128B-swizzled smem (A operand of the PV MMA); O / l, LSE.
// S_{j+1} is issued before P_j is consumed so the tensor core overlaps the softmax of block j.
// GQA (Hq = g * Hkv) and packed variable-length batches (cu_seqlens) are handled by indexing; the
// operands are strided views (e.g. slices of one fused QKV activation), nothing is copied.
//
// Replaces flash_attn_2_cuda / SDPA behind HF attention
// (reference: 05-training-llama-405b/train_llm.py:87, 01-single-gpu/train_llm.py:140;
//  packed variant 00-rime/packed_dataset.py:35-48).
#include "common.cuh"
#include <math.h>

namespace b2 {

constexpr int ABM = 128;  // query rows per CTA
constexpr int ABN = 128;  // kv rows per block
constexpr int kAttnThreads = 320;  // TMA warp, MMA warp, 2 softmax warpgroups (64 columns each)

template <int D>
struct AttnFwdSmem {
  static constexpr int kQ = ABM * D * 2;
  static constexpr int kK = ABN * D * 2;
  static constexpr int kV = ABN * D * 2;
  static constexpr int kP = ABM * ABN * 2;
  static constexpr int kStats = 2 * 2 * ABM * 2;  // bf16 row-max exchange between the two softmax halves
  static constexpr int kTotal = kQ + 2 * kK + 2 * kV + 2 * kP + kStats + 256 + 1024;
};

struct AttnFwdParams {
  __nv_bfloat16* o;      // [T, o_stride]
  float* lse;            // [Hq, T]
  const int* cu_seqlens; // [ndocs + 1]
  int o_stride;
  int T;
  int Hq, Hkv;
  int q_col0, k_col0, v_col0;  // element offsets of head 0 inside the q/k/v row
  float scale_log2;            // softmax_scale * log2(e)
  int causal;
};

template <int D>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, AttnFwdParams p) {
  using SM = AttnFwdSmem<D>;
  constexpr int kAtoms = D / 64;  // 64-element (128B) swizzle atoms per head row
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + SM::kQ;
  uint8_t* sV = sK + 2 * SM::kK;
  uint8_t* sP = sV + 2 * SM::kV;
  __nv_bfloat16* sStat = reinterpret_cast<__nv_bfloat16*>(sP + 2 * SM::kP);  // [2 (parity)][2 (half)][ABM]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * SM::kP + SM::kStats);
  uint64_t* q_full = bars;          // 1
  uint64_t* k_full = bars + 1;      // 2
  uint64_t* k_empty = bars + 3;     // 2
  uint64_t* v_full = bars + 5;      // 2
  uint64_t* v_empty = bars + 7;     // 2
  uint64_t* s_full = bars + 9;      // 2
  uint64_t* s_free = bars + 11;     // 2
  uint64_t* p_full = bars + 13;     // 2
  uint64_t* p_empty = bars + 15;    // 2
  uint64_t* o_done = bars + 17;     // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);

  const int warp = threadIdx.x >> 5;
  const int doc = blockIdx.z;
  const int head = blockIdx.y;
  const int doc_start = p.cu_seqlens[doc];
  const int doc_len = p.cu_seqlens[doc + 1] - doc_start;
  const int num_qb = (doc_len + ABM - 1) / ABM;
  const int qb = num_qb - 1 - (int)blockIdx.x;  // heaviest (latest) query blocks first
  if (qb < 0) return;
  const int q0 = qb * ABM;  // row offset inside the document
  const int kv_len = p.causal ? min(doc_len, q0 + ABM) : doc_len;
  const int nkv = (kv_len + ABN - 1) / ABN;
  const int kvh = head / (p.Hq / p.Hkv);

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_free[i], 8);
      mbar_init(&p_full[i], 8);
      mbar_init(&p_empty[i], 1);
    }
    mbar_init(o_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<1>(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S[2] = {tmem_base, tmem_base + 128};
  const uint32_t tmem_O = tmem_base + 256;

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {
      mbar_expect_tx(q_full, SM::kQ);
#pragma unroll
      for (int a = 0; a < kAtoms; ++a)
        tma_load_2d<1>(&tmQ, q_full, sQ + a * (ABM * 128), p.q_col0 + head * D + a * 64,
                       doc_start + q0, kEvictFirst);
      for (int j = 0; j < nkv; ++j) {
        const int s = j & 1;
        const uint32_t ph = ((j >> 1) & 1) ^ 1;
        const int row = doc_start + j * ABN;
        mbar_wait(&k_empty[s], ph);
        mbar_expect_tx(&k_full[s], SM::kK);
#pragma unroll
        for (int a = 0; a < kAtoms; ++a)
          tma_load_2d<1>(&tmK, &k_full[s], sK + s * SM::kK + a * (ABN * 128),
                         p.k_col0 + kvh * D + a * 64, row, kEvictLast);
        mbar_wait(&v_empty[s], ph);
        mbar_expect_tx(&v_full[s], SM::kV);
        tma_load_3d<1>(&tmV, &v_full[s], sV + s * SM::kV, 0, row, (p.v_col0 + kvh * D) / 64,
                       kEvictLast);
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc_s = make_idesc_bf16(ABM, ABN, false, false);
    constexpr uint32_t idesc_o = make_idesc_bf16(ABM, D, false, true);
    auto issue_s = [&](int j) {
      const int s = j & 1;
      mbar_wait(&k_full[s], (j >> 1) & 1);
      mbar_wait(&s_free[s], ((j >> 1) & 1) ^ 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK + s * SM::kK);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (ABM * 128) + (kk & 3) * 32;
          umma_f16<1>(tmem_S[s], make_smem_desc_sw128(qa + off, 16, 1024),
                      make_smem_desc_sw128(ka + off, 16, 1024), idesc_s, kk != 0);
        }
        umma_commit<1>(&k_empty[s]);
        umma_commit<1>(&s_full[s]);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    issue_s(0);
    for (int j = 0; j < nkv; ++j) {
      if (j + 1 < nkv) issue_s(j + 1);
      const int s = j & 1;
      mbar_wait(&v_full[s], (j >> 1) & 1);
      mbar_wait(&p_full[s], (j >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t pa = smem_u32(sP + s * SM::kP), va = smem_u32(sV + s * SM::kV);
#pragma unroll
        for (int kk = 0; kk < ABN / 16; ++kk) {
          const uint64_t da =
              make_smem_desc_sw128(pa + (kk >> 2) * (ABM * 128) + (kk & 3) * 32, 16, 1024);
          const uint64_t db = make_smem_desc_sw128(va + kk * 2048, ABN * 128, 1024);
          umma_f16<1>(tmem_O, da, db, idesc_o, (j | kk) != 0);
        }
        umma_commit<1>(&v_empty[s]);
        umma_commit<1>(&p_empty[s]);
        umma_commit<1>(o_done);
      }
      __syncwarp();
    }
  } else {
    // ================= softmax / correction / epilogue =================
    // Two warpgroups share every query row: half 0 owns kv columns [0, 64) of the block and the first
    // D/2 columns of O, half 1 the rest. Only the row maximum is exchanged per block (through smem);
    // the running sums are combined once at the end.
    constexpr int HC = ABN / 2;            // score columns per thread
    constexpr int OC = D / 2;              // O columns per thread
    const int half = (warp - 2) >> 2;
    const int quarter = warp & 3;  // TMEM lane quarter this warp may access
    const int row = quarter * 32 + (int)lane_id();
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const int q_idx = q0 + row;  // query position inside the document
    float m_used = -INFINITY;    // running max (scaled, log2 domain) used as the exponent base
    float l = 0.f;               // this half's share of the running sum
    for (int j = 0; j < nkv; ++j) {
      const int s = j & 1;
      mbar_wait(&s_full[s], (j >> 1) & 1);
      tc_fence_after();
      float sv[HC];
#pragma unroll
      for (int c = 0; c < HC / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_S[s] + lane_off + half * HC + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) sv[c * 32 + i] = __uint_as_float(r[i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&s_free[s]);
      // masking: causal diagonal and document tail
      const int kv0 = j * ABN + half * HC;
      const bool need_mask = (j * ABN + ABN > kv_len) || (p.causal && j * ABN + ABN > q0);
      if (need_mask) {
        const int lim = p.causal ? min(q_idx, doc_len - 1) : doc_len - 1;  // last visible kv index
#pragma unroll
        for (int i = 0; i < HC; ++i)
          if (kv0 + i > lim) sv[i] = -INFINITY;
      }
      float mx = sv[0];
#pragma unroll
      for (int i = 1; i < HC; ++i) mx = fmaxf(mx, sv[i]);
      // exchange the row maximum with the other half (double-buffered by block parity: one barrier)
      // (both halves use the bf16-rounded maxima, so they agree bit for bit on the exponent base; a base
      // that is a rounding below the true maximum only makes the largest p slightly exceed 1)
      __nv_bfloat16* stat = sStat + (j & 1) * (2 * ABM);
      const __nv_bfloat16 mine = __float2bfloat16_rn(mx);
      stat[half * ABM + row] = mine;
      named_bar_sync(3, 256);
      mx = fmaxf(__bfloat162float(mine), __bfloat162float(stat[(half ^ 1) * ABM + row])) * p.scale_log2;
      // lazy rescale: keep a stale exponent base until the row max grows by more than 2^8
      float alpha = 1.f;
      const bool need = mx > m_used + 8.f;
      if (need) {
        alpha = (m_used == -INFINITY) ? 0.f : exp2f(m_used - mx);
        l *= alpha;
        m_used = mx;
      }
      if (j > 0 && __any_sync(0xffffffffu, need)) {
        mbar_wait(o_done, (j - 1) & 1);  // PV_{j-1} retired: O is stable
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < OC / 16; ++c) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(tmem_O + lane_off + half * OC + c * 16, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st_32x32b_x16(tmem_O + lane_off + half * OC + c * 16, r);
        }
        tmem_st_wait();
        tc_fence_before();
      }
      const float base = (m_used == -INFINITY) ? 0.f : m_used;
      float psum = 0.f;
      uint32_t pk[HC / 2];
#pragma unroll
      for (int i = 0; i < HC / 2; ++i) {
        const float p0 = exp2f(fmaf(sv[2 * i], p.scale_log2, -base));
        const float p1 = exp2f(fmaf(sv[2 * i + 1], p.scale_log2, -base));
        psum += p0 + p1;
        pk[i] = pack_bf16(p0, p1);
      }
      l += psum;
      mbar_wait(&p_empty[s], ((j >> 1) & 1) ^ 1);  // PV_{j-2} finished reading this P buffer
      // this half's 64 columns are one 128-byte swizzle atom of the K-major P tile
      const uint32_t pbase = smem_u32(sP + s * SM::kP) + half * (ABM * 128) + row * 128;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint32_t addr = pbase + ((c ^ (row & 7)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(pk[4 * c]),
                     "r"(pk[4 * c + 1]), "r"(pk[4 * c + 2]), "r"(pk[4 * c + 3])
                     : "memory");
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&p_full[s]);
    }
    // ---- epilogue: combine the two partial sums, O / l -> bf16, LSE ----
    // (the two parity buffers together hold one fp32 per (half, row); every thread is past its last
    // row-max exchange once it arrives at the barrier below, and reads only after it)
    named_bar_sync(3, 256);
    float* lsum = reinterpret_cast<float*>(sStat);
    lsum[half * ABM + row] = l;
    named_bar_sync(3, 256);
    l += lsum[(half ^ 1) * ABM + row];
    mbar_wait(o_done, (nkv - 1) & 1);
    tc_fence_after();
    const float inv_l = l > 0.f ? 1.f / l : 0.f;
    const bool valid = q_idx < doc_len;
    __nv_bfloat16* orow =
        p.o + (size_t)(doc_start + q_idx) * p.o_stride + (size_t)head * D + half * OC;
#pragma unroll
    for (int c = 0; c < OC / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(tmem_O + lane_off + half * OC + c * 32, r);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 v;
          v.x = pack_bf16(__uint_as_float(r[8 * i]) * inv_l, __uint_as_float(r[8 * i + 1]) * inv_l);
          v.y = pack_bf16(__uint_as_float(r[8 * i + 2]) * inv_l,
                          __uint_as_float(r[8 * i + 3]) * inv_l);
          v.z = pack_bf16(__uint_as_float(r[8 * i + 4]) * inv_l,
                          __uint_as_float(r[8 * i + 5]) * inv_l);
          v.w = pack_bf16(__uint_as_float(r[8 * i + 6]) * inv_l,
                          __uint_as_float(r[8 * i + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = v;
        }
      }
    }
    if (valid && half == 0)
      p.lse[(size_t)head * p.T + doc_start + q_idx] =
          (l > 0.f) ? (m_used + log2f(l)) * 0.6931471805599453f : -INFINITY;
    tc_fence_before();
  }

  __syncwarp();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

template <int D>
static int launch_attn_fwd(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                           const AttnFwdParams& prm, int max_qblocks, int ndocs,
                           cudaStream_t stream) {
  using SM = AttnFwdSmem<D>;
  auto kern = attn_fwd_kernel<D>;
  static bool attr_set = false;
  if (!attr_set) {
    B2_CUDA_CHECK(
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::kTotal));
    attr_set = true;
  }
  dim3 grid(max_qblocks, prm.Hq, ndocs);
  kern<<<grid, kAttnThreads, SM::kTotal, stream>>>(tmQ, tmK, tmV, prm);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace b2

extern "C" {

// q/k/v are row-major [T, *_stride] bf16 views whose head h starts at column h * D (q: Hq heads,
// k/v: Hkv heads); they may alias one fused QKV buffer. o: [T, o_stride], lse: [Hq, T] fp32.
// cu_seqlens: device int32 [ndocs+1] document boundaries; max_seqlen bounds the longest document.
int b2_attn_fwd(void* o, float* lse, const void* q, const void* k, const void* v,
                const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                int q_stride, int k_stride, int v_stride, int o_stride, float softmax_scale,
                int causal, cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(D == 64 || D == 128, "attn: head_dim %d unsupported (64 or 128)", D);
  B2_REQUIRE(Hq % Hkv == 0, "attn: Hq=%d not a multiple of Hkv=%d", Hq, Hkv);
  B2_REQUIRE(v_stride % 64 == 0, "attn: v row stride %d must be a multiple of 64", v_stride);
  B2_REQUIRE(((uintptr_t)o % 16) == 0 && o_stride % 8 == 0, "attn: o must be 16B aligned");
  if (T == 0 || ndocs == 0) return 0;
  CUtensorMap tmQ, tmK, tmV;
  int rc;
  if ((rc = make_tmap_2d_bf16(&tmQ, q, T, (uint64_t)Hq * D, q_stride, ABM, 64))) return rc;
  if ((rc = make_tmap_2d_bf16(&tmK, k, T, (uint64_t)Hkv * D, k_stride, ABN, 64))) return rc;
  if ((rc = make_tmap_mn_bf16(&tmV, v, T, (uint64_t)Hkv * D, v_stride, ABN, D / 64))) return rc;
  AttnFwdParams prm;
  prm.o = static_cast<__nv_bfloat16*>(o);
  prm.lse = lse;
  prm.cu_seqlens = cu_seqlens;
  prm.o_stride = o_stride;
  prm.T = T;
  prm.Hq = Hq;
  prm.Hkv = Hkv;
  prm.q_col0 = prm.k_col0 = prm.v_col0 = 0;
  prm.scale_log2 = softmax_scale * 1.4426950408889634f;
  prm.causal = causal;
  const int max_qb = (max_seqlen + ABM - 1) / ABM;
  if (D == 64) return launch_attn_fwd<64>(tmQ, tmK, tmV, prm, max_qb, ndocs, stream);
  return launch_attn_fwd<128>(tmQ, tmK, tmV, prm, max_qb, ndocs, stream);
}

}  // extern "C"
