// Causal flash-attention forward for sm_100a on tcgen05 tensor cores with TMEM accumulators.
//
// One CTA per (document, query head, 128-row query block); 320 threads:
//   warp 0     TMA producer  : Q once, then a 2-stage ring of K (K-major) and V (MN-major) tiles
//   warp 1     MMA issuer    : S_j = Q K_j^T into one of two TMEM S buffers, O += P_j V_j into TMEM O
//   warps 2-9  softmax       : two warpgroups in ping-pong: warpgroup w handles the kv blocks j with
//                              j % 2 == w, one thread per query row, with its own S buffer, its own O
//                              accumulator in TMEM and its own running (max, sum). While one warpgroup
//                              drains S from TMEM the other is in its exp2 phase, so the TMEM read
//                              port, the MUFU pipe and the tensor pipe work concurrently. tcgen05.ld S,
//                              online softmax in the exp2 domain with lazy rescale of O, P_j -> bf16 ->
//                              tcgen05.st back into tensor memory over S (the PV MMA reads its A operand
//                              from TMEM, so P never touches shared memory). The two partial results
//                              are merged at the end (through a staging buffer) into O / l and the LSE.
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
  static constexpr int kStats = 2 * ABM * 4;      // (max, sum) of warpgroup 1 for the final merge
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
  AttnDropout drop;            // attention-probability dropout (thr == 0: none)
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
  float* sStat = reinterpret_cast<float*>(sP + 2 * SM::kP);  // [2 (max, sum)][ABM]
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
  uint64_t* o_done = bars + 17;     // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 19);

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
      mbar_init(&s_free[i], 4);
      mbar_init(&p_full[i], 4);
      mbar_init(&o_done[i], 1);
      mbar_init(&p_empty[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<1>(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S[2] = {tmem_base, tmem_base + 128};
  const uint32_t tmem_O[2] = {tmem_base + 256, tmem_base + 256 + D};  // one accumulator per warpgroup

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
        // P_j (bf16) sits in the first 64 columns of this warpgroup's S buffer: the A operand is read
        // from tensor memory, 8 columns (16 K-elements) per MMA
        const uint32_t va = smem_u32(sV + s * SM::kV);
#pragma unroll
        for (int kk = 0; kk < ABN / 16; ++kk) {
          const uint64_t db = make_smem_desc_sw128(va + kk * 2048, ABN * 128, 1024);
          umma_f16_ts(tmem_O[s], tmem_S[s] + kk * 8, db, idesc_o, (j >= 2 || kk != 0) ? 1u : 0u);
        }
        umma_commit<1>(&v_empty[s]);
        umma_commit<1>(&o_done[s]);
      }
      __syncwarp();
    }
  } else {
    // ================= softmax / correction / epilogue =================
    const int wg = (warp - 2) >> 2;  // ping-pong warpgroup: handles kv blocks j with (j & 1) == wg
    const int quarter = warp & 3;    // TMEM lane quarter this warp may access
    const int row = quarter * 32 + (int)lane_id();
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const int q_idx = q0 + row;  // query position inside the document
    float m_used = -INFINITY;    // running max (scaled, log2 domain) used as the exponent base
    float l = 0.f;
    int k = 0;                   // number of blocks this warpgroup has finished
    for (int j = wg; j < nkv; j += 2, ++k) {
      mbar_wait(&s_full[wg], k & 1);
      tc_fence_after();
      // all four TMEM loads are in flight together (one wait), the row lives in registers from here on
      uint32_t su[ABN];
#pragma unroll
      for (int c = 0; c < ABN / 32; ++c)
        tmem_ld_32x32b_x32(tmem_S[wg] + lane_off + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&su[c * 32]));
      tmem_ld_wait();
      float sv[ABN];
#pragma unroll
      for (int i = 0; i < ABN; ++i) sv[i] = __uint_as_float(su[i]);
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&s_free[wg]);
      // masking: causal diagonal and document tail
      const int kv0 = j * ABN;
      const bool need_mask = (kv0 + ABN > kv_len) || (p.causal && kv0 + ABN > q0);
      if (need_mask) {
        const int lim = p.causal ? min(q_idx, doc_len - 1) : doc_len - 1;  // last visible kv index
#pragma unroll
        for (int i = 0; i < ABN; ++i)
          if (kv0 + i > lim) sv[i] = -INFINITY;
      }
      // eight independent running maxima, then a tree: a single chain of 127 dependent FMNMX would
      // cost ~500 cycles of latency that two warps per scheduler cannot hide
      float m8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) m8[i] = sv[i];
#pragma unroll
      for (int i = 8; i < ABN; ++i) m8[i & 7] = fmaxf(m8[i & 7], sv[i]);
      float mx = fmaxf(fmaxf(fmaxf(m8[0], m8[1]), fmaxf(m8[2], m8[3])),
                       fmaxf(fmaxf(m8[4], m8[5]), fmaxf(m8[6], m8[7])));
      mx *= p.scale_log2;
      // lazy rescale: keep a stale exponent base until the row max grows by more than 2^8
      float alpha = 1.f;
      const bool need = mx > m_used + 8.f;
      if (need) {
        alpha = (m_used == -INFINITY) ? 0.f : fast_exp2(m_used - mx);
        l *= alpha;
        m_used = mx;
      }
      if (k > 0 && __any_sync(0xffffffffu, need)) {
        mbar_wait(&o_done[wg], (k - 1) & 1);  // this warpgroup's previous PV retired: its O is stable
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < D / 16; ++c) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(tmem_O[wg] + lane_off + c * 16, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st_32x32b_x16(tmem_O[wg] + lane_off + c * 16, r);
        }
        tmem_st_wait();
        tc_fence_before();
      }
      const float base = (m_used == -INFINITY) ? 0.f : m_used;
      float ps4[4] = {0.f, 0.f, 0.f, 0.f};  // independent partial row sums (same latency argument)
      // P (bf16, two K-elements per 32-bit column) overwrites the first 64 columns of this warpgroup's S
      // buffer; the next S MMA into this buffer is issued after PV_j, so tensor-pipe order protects it
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t pk[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const int i = c * 32 + e * 2;
          const float p0 = fast_exp2(fmaf(sv[i], p.scale_log2, -base));
          const float p1 = fast_exp2(fmaf(sv[i + 1], p.scale_log2, -base));
          ps4[e & 3] += p0 + p1;
          pk[e] = pack_bf16(p0, p1);
        }
        tmem_st_32x32b_x16(tmem_S[wg] + lane_off + c * 16, pk);
      }
      l += (ps4[0] + ps4[1]) + (ps4[2] + ps4[3]);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&p_full[wg]);
    }
    // ---- merge the two warpgroups' partial results: O = (f0 O0 + f1 O1) / (f0 l0 + f1 l1) ----
    if (k > 0) {
      mbar_wait(&o_done[wg], (k - 1) & 1);  // last PV of this warpgroup (which also read P last)
      tc_fence_after();
    }
    if (wg == 1) {
      sStat[row] = m_used;
      sStat[ABM + row] = l;
    }
    // all PV MMAs that read the P buffers have retired for both warpgroups once both pass this barrier
    named_bar_sync(3, 256);
    const float m1 = sStat[row], l1 = sStat[ABM + row];
    float m_all, f_mine, l_all;
    if (wg == 0) {
      m_all = fmaxf(m_used, m1);
      const float f0 = (m_used == -INFINITY) ? 0.f : fast_exp2(m_used - m_all);
      const float f1 = (m1 == -INFINITY) ? 0.f : fast_exp2(m1 - m_all);
      l_all = f0 * l + f1 * l1;
      f_mine = f0;
    } else {
      m_all = 0.f;
      l_all = 0.f;
      f_mine = 1.f;  // warpgroup 1 hands over its raw accumulator; warpgroup 0 applies both factors
    }
    float* xbuf = reinterpret_cast<float*>(sP);  // [D][ABM] fp32, column-major: conflict-free per warp
    const bool valid = q_idx < doc_len;
    if (wg == 1) {
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t r[32];
        if (k > 0) {
          tmem_ld_32x32b_x32(tmem_O[1] + lane_off + c * 32, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = 0u;
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) xbuf[(c * 32 + i) * ABM + row] = __uint_as_float(r[i]);
      }
    }
    named_bar_sync(3, 256);
    if (wg == 0) {
      const float f1 = (m1 == -INFINITY) ? 0.f : fast_exp2(m1 - m_all);
      const float inv_l = l_all > 0.f ? 1.f / l_all : 0.f;
      const float a0 = f_mine * inv_l, a1 = f1 * inv_l;
      __nv_bfloat16* orow = p.o + (size_t)(doc_start + q_idx) * p.o_stride + (size_t)head * D;
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_O[0] + lane_off + c * 32, r);
        tmem_ld_wait();
        float o[32];
#pragma unroll
        for (int i = 0; i < 32; ++i)
          o[i] = __uint_as_float(r[i]) * a0 + xbuf[(c * 32 + i) * ABM + row] * a1;
        if (valid) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 v;
            v.x = pack_bf16(o[8 * i], o[8 * i + 1]);
            v.y = pack_bf16(o[8 * i + 2], o[8 * i + 3]);
            v.z = pack_bf16(o[8 * i + 4], o[8 * i + 5]);
            v.w = pack_bf16(o[8 * i + 6], o[8 * i + 7]);
            *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = v;
          }
        }
      }
      if (valid)
        p.lse[(size_t)head * p.T + doc_start + q_idx] =
            (l_all > 0.f) ? (m_all + log2f(l_all)) * 0.6931471805599453f : -INFINITY;
    }
    tc_fence_before();
  }

  __syncwarp();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}


// ------------------------------------------------------------------------------------------------
// One warpgroup per CTA, two CTAs per SM (default).
//
// Timing the ping-pong kernel above at S = 4096 and 8192 splits its run time into ~0.9 us per kv block
// and ~13.6 us per CTA: Q load and pipeline fill, the final merge of the two warpgroups' accumulators
// through shared memory, TMEM release and the launch of the next CTA, none of which overlap anything
// because that kernel fills the SM (231 KB of shared memory, all 512 TMEM columns). At S = 4096 that fixed
// part is half of the kernel. This variant keeps the same per-block chain (S_j -> softmax_j -> P_j in
// TMEM -> PV_j) but gives a CTA one softmax warpgroup, 256 TMEM columns (S | O) and single-stage K / V
// tiles (97 KB), so two CTAs share an SM: while one is in its softmax the other owns the tensor pipe
// (what the two warpgroups did before), and the prologue / epilogue of one CTA runs under the steady
// state of its neighbour. One accumulator per query tile: no merge.
constexpr int kAttn1Threads = 192;  // TMA warp, MMA warp, one softmax warpgroup

template <int D>
struct AttnFwd1Smem {
  static constexpr int kQ = ABM * D * 2;
  static constexpr int kK = ABN * D * 2;
  static constexpr int kV = ABN * D * 2;
  static constexpr int kTotal = kQ + kK + kV + 256 + 1024;
};

template <int D, bool kDrop = false>
__global__ void __launch_bounds__(kAttn1Threads, 2)
attn_fwd1_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, AttnFwdParams p) {
  using SM = AttnFwd1Smem<D>;
  constexpr int kAtoms = D / 64;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + SM::kQ;
  uint8_t* sV = sK + SM::kK;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + SM::kV);
  uint64_t* q_full = bars;       // Q tile landed
  uint64_t* k_full = bars + 1;   // K_j landed / S_j has read it
  uint64_t* k_empty = bars + 2;
  uint64_t* v_full = bars + 3;   // V_j landed / PV_j has read it
  uint64_t* v_empty = bars + 4;
  uint64_t* s_full = bars + 5;   // S_j is in TMEM
  uint64_t* p_full = bars + 6;   // P_j written back over S_j (4 warps)
  uint64_t* o_done = bars + 7;   // PV_j retired
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5;
  const int doc = blockIdx.z;
  const int head = blockIdx.y;
  const int doc_start = p.cu_seqlens[doc];
  const int doc_len = p.cu_seqlens[doc + 1] - doc_start;
  const int num_qb = (doc_len + ABM - 1) / ABM;
  const int qb = num_qb - 1 - (int)blockIdx.x;  // heaviest (latest) query blocks first
  if (qb < 0) return;
  const int q0 = qb * ABM;
  const int kv_len = p.causal ? min(doc_len, q0 + ABM) : doc_len;
  const int nkv = (kv_len + ABN - 1) / ABN;
  const int kvh = head / (p.Hq / p.Hkv);

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    mbar_init(q_full, 1);
    mbar_init(k_full, 1);
    mbar_init(k_empty, 1);
    mbar_init(v_full, 1);
    mbar_init(v_empty, 1);
    mbar_init(s_full, 1);
    mbar_init(p_full, 4);
    mbar_init(o_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<1>(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base, tmem_O = tmem_base + 128;

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {
      mbar_expect_tx(q_full, SM::kQ);
#pragma unroll
      for (int a = 0; a < kAtoms; ++a)
        tma_load_2d<1>(&tmQ, q_full, sQ + a * (ABM * 128), p.q_col0 + head * D + a * 64,
                       doc_start + q0, kEvictFirst);
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = (j & 1) ^ 1;
        const int row = doc_start + j * ABN;
        mbar_wait(k_empty, ph);
        mbar_expect_tx(k_full, SM::kK);
#pragma unroll
        for (int a = 0; a < kAtoms; ++a)
          tma_load_2d<1>(&tmK, k_full, sK + a * (ABN * 128), p.k_col0 + kvh * D + a * 64, row,
                         kEvictLast);
        mbar_wait(v_empty, ph);
        mbar_expect_tx(v_full, SM::kV);
        tma_load_3d<1>(&tmV, v_full, sV, 0, row, (p.v_col0 + kvh * D) / 64, kEvictLast);
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc_s = make_idesc_bf16(ABM, ABN, false, false);
    constexpr uint32_t idesc_o = make_idesc_bf16(ABM, D, false, true);
    mbar_wait(q_full, 0);
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(k_full, j & 1);
      tc_fence_after();
      if (elect_one()) {
        // S_j overwrites P_{j-1}: PV_{j-1} was issued before it, the tensor pipe keeps the order
        const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (ABM * 128) + (kk & 3) * 32;
          umma_f16<1>(tmem_S, make_smem_desc_sw128(qa + off, 16, 1024),
                      make_smem_desc_sw128(ka + off, 16, 1024), idesc_s, kk != 0);
        }
        umma_commit<1>(k_empty);
        umma_commit<1>(s_full);
      }
      __syncwarp();
      mbar_wait(v_full, j & 1);
      mbar_wait(p_full, j & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t va = smem_u32(sV);
#pragma unroll
        for (int kk = 0; kk < ABN / 16; ++kk) {
          const uint64_t db = make_smem_desc_sw128(va + kk * 2048, ABN * 128, 1024);
          umma_f16_ts(tmem_O, tmem_S + kk * 8, db, idesc_o, (j > 0 || kk != 0) ? 1u : 0u);
        }
        umma_commit<1>(v_empty);
        umma_commit<1>(o_done);
      }
      __syncwarp();
    }
  } else {
    // ================= softmax / correction / epilogue: one thread per query row =================
    const int quarter = warp & 3;
    const int row = quarter * 32 + (int)lane_id();
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    const int q_idx = q0 + row;
    float m_used = -INFINITY;  // running max (scaled, log2 domain) used as the exponent base
    float l = 0.f;
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      uint32_t su[ABN];
#pragma unroll
      for (int c = 0; c < ABN / 32; ++c)
        tmem_ld_32x32b_x32(tmem_S + lane_off + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&su[c * 32]));
      tmem_ld_wait();
      float sv[ABN];
#pragma unroll
      for (int i = 0; i < ABN; ++i) sv[i] = __uint_as_float(su[i]);
      const int kv0 = j * ABN;
      const bool need_mask = (kv0 + ABN > kv_len) || (p.causal && kv0 + ABN > q0);
      if (need_mask) {
        const int lim = p.causal ? min(q_idx, doc_len - 1) : doc_len - 1;  // last visible kv index
#pragma unroll
        for (int i = 0; i < ABN; ++i)
          if (kv0 + i > lim) sv[i] = -INFINITY;
      }
      float m8[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) m8[i] = sv[i];
#pragma unroll
      for (int i = 8; i < ABN; ++i) m8[i & 7] = fmaxf(m8[i & 7], sv[i]);
      float mx = fmaxf(fmaxf(fmaxf(m8[0], m8[1]), fmaxf(m8[2], m8[3])),
                       fmaxf(fmaxf(m8[4], m8[5]), fmaxf(m8[6], m8[7])));
      mx *= p.scale_log2;
      // lazy rescale: keep a stale exponent base until the row max grows by more than 2^8
      float alpha = 1.f;
      const bool need = mx > m_used + 8.f;
      if (need) {
        alpha = (m_used == -INFINITY) ? 0.f : fast_exp2(m_used - mx);
        l *= alpha;
        m_used = mx;
      }
      if (j > 0 && __any_sync(0xffffffffu, need)) {
        mbar_wait(o_done, (j - 1) & 1);  // PV_{j-1} retired (it was issued before S_j): O is stable
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < D / 16; ++c) {
          uint32_t r[16];
          tmem_ld_32x32b_x16(tmem_O + lane_off + c * 16, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st_32x32b_x16(tmem_O + lane_off + c * 16, r);
        }
        tmem_st_wait();
        tc_fence_before();
      }
      const float base = (m_used == -INFINITY) ? 0.f : m_used;
      float ps4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t pk[16];
        uint4 bits0, bits1;
        if constexpr (kDrop) {  // keep decisions of key positions kv0 + 32 c + [0, 16) and + [16, 32)
          const int g = (kv0 + c * 32) >> 4;
          bits0 = attn_dropout_bits(p.drop, q_idx, g, head, doc);
          bits1 = attn_dropout_bits(p.drop, q_idx, g + 1, head, doc);
        }
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const int i = c * 32 + e * 2;
          float p0 = fast_exp2(fmaf(sv[i], p.scale_log2, -base));
          float p1 = fast_exp2(fmaf(sv[i + 1], p.scale_log2, -base));
          ps4[e & 3] += p0 + p1;   // the normaliser is the sum of the un-dropped probabilities
          if constexpr (kDrop) {
            const uint4& b = e < 8 ? bits0 : bits1;
            if (!dropout_keep(b, (2 * e) & 15, p.drop.thr)) p0 = 0.f;
            if (!dropout_keep(b, (2 * e + 1) & 15, p.drop.thr)) p1 = 0.f;
          }
          pk[e] = pack_bf16(p0, p1);
        }
        tmem_st_32x32b_x16(tmem_S + lane_off + c * 16, pk);  // P_j over the first 64 columns of S_j
      }
      l += (ps4[0] + ps4[1]) + (ps4[2] + ps4[3]);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(p_full);
    }
    // ---- epilogue: O / l -> bf16, LSE ----
    if (nkv > 0) {
      mbar_wait(o_done, (nkv - 1) & 1);
      tc_fence_after();
    }
    const bool valid = q_idx < doc_len;
    float inv_l = l > 0.f ? 1.f / l : 0.f;
    if constexpr (kDrop) inv_l *= p.drop.scale;   // kept probabilities are scaled by 1 / (1 - p)
    __nv_bfloat16* orow = p.o + (size_t)(doc_start + q_idx) * p.o_stride + (size_t)head * D;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t r[32];
      if (nkv > 0) {
        tmem_ld_32x32b_x32(tmem_O + lane_off + c * 32, r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) r[i] = 0u;
      }
      if (valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 v;
          v.x = pack_bf16(__uint_as_float(r[8 * i]) * inv_l, __uint_as_float(r[8 * i + 1]) * inv_l);
          v.y = pack_bf16(__uint_as_float(r[8 * i + 2]) * inv_l, __uint_as_float(r[8 * i + 3]) * inv_l);
          v.z = pack_bf16(__uint_as_float(r[8 * i + 4]) * inv_l, __uint_as_float(r[8 * i + 5]) * inv_l);
          v.w = pack_bf16(__uint_as_float(r[8 * i + 6]) * inv_l, __uint_as_float(r[8 * i + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = v;
        }
      }
    }
    if (valid)
      p.lse[(size_t)head * p.T + doc_start + q_idx] =
          (l > 0.f) ? (m_used + log2f(l)) * 0.6931471805599453f : -INFINITY;
    tc_fence_before();
  }

  __syncwarp();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 256);
  }
}

template <int D, bool kDrop = false>
static int launch_attn_fwd1(const CUtensorMap& tmQ, const CUtensorMap& tmK, const CUtensorMap& tmV,
                            const AttnFwdParams& prm, int max_qblocks, int ndocs,
                            cudaStream_t stream) {
  using SM = AttnFwd1Smem<D>;
  auto kern = attn_fwd1_kernel<D, kDrop>;
  static bool attr_set = false;
  if (!attr_set) {
    B2_CUDA_CHECK(
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::kTotal));
    attr_set = true;
  }
  dim3 grid(max_qblocks, prm.Hq, ndocs);
  kern<<<grid, kAttn1Threads, SM::kTotal, stream>>>(tmQ, tmK, tmV, prm);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
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
static int attn_fwd_impl(void* o, float* lse, const void* q, const void* k, const void* v,
                         const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                         int q_stride, int k_stride, int v_stride, int o_stride, float softmax_scale,
                         int causal, const b2::AttnDropout& drop, cudaStream_t stream) {
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
  prm.drop = drop;
  const int max_qb = (max_seqlen + ABM - 1) / ABM;
  if (drop.thr != 0) {
    if (D == 64) return launch_attn_fwd1<64, true>(tmQ, tmK, tmV, prm, max_qb, ndocs, stream);
    return launch_attn_fwd1<128, true>(tmQ, tmK, tmV, prm, max_qb, ndocs, stream);
  }
  // default: one warpgroup per CTA, two CTAs per SM; B200TRAIN_ATTN_FWD_PINGPONG=1 selects the ping-pong kernel
  static const bool pingpong = [] {
    const char* e = getenv("B200TRAIN_ATTN_FWD_PINGPONG");
    return e && e[0] == '1';
  }();
  if (pingpong) {
    if (D == 64) return launch_attn_fwd<64>(tmQ, tmK, tmV, prm, max_qb, ndocs, stream);
    return launch_attn_fwd<128>(tmQ, tmK, tmV, prm, max_qb, ndocs, stream);
  }
  if (D == 64) return launch_attn_fwd1<64>(tmQ, tmK, tmV, prm, max_qb, ndocs, stream);
  return launch_attn_fwd1<128>(tmQ, tmK, tmV, prm, max_qb, ndocs, stream);
}

int b2_attn_fwd(void* o, float* lse, const void* q, const void* k, const void* v,
                const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                int q_stride, int k_stride, int v_stride, int o_stride, float softmax_scale,
                int causal, cudaStream_t stream) {
  b2::AttnDropout none = {};
  return attn_fwd_impl(o, lse, q, k, v, cu_seqlens, ndocs, max_seqlen, T, Hq, Hkv, D, q_stride, k_stride,
                       v_stride, o_stride, softmax_scale, causal, none, stream);
}

// The same forward with dropout on the attention probabilities (GPT-2's attn_pdrop): a probability is kept
// when its Philox byte is >= round(256 p), so the effective rate is round(256 p) / 256 and the kept ones are
// scaled by its reciprocal complement. b2_attn_bwd_dropout regenerates the mask from (seed, offset).
int b2_attn_fwd_dropout(void* o, float* lse, const void* q, const void* k, const void* v,
                        const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                        int q_stride, int k_stride, int v_stride, int o_stride, float softmax_scale,
                        int causal, float drop_p, unsigned long long seed, unsigned long long offset,
                        cudaStream_t stream) {
  b2::AttnDropout d = b2::make_attn_dropout(drop_p, seed, offset);
  return attn_fwd_impl(o, lse, q, k, v, cu_seqlens, ndocs, max_seqlen, T, Hq, Hkv, D, q_stride, k_stride,
                       v_stride, o_stride, softmax_scale, causal, d, stream);
}

}  // extern "C"
