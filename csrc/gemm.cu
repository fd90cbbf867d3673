// bf16 GEMM for sm_100a: C[M,N] = epilogue(A[M,K] * B[N,K]^T), fp32 accumulation in tensor memory.
//
// Structure (one persistent CTA, or CTA pair, per SM):
//   warp 0      TMA producer   : cp.async.bulk.tensor loads of A/B k-blocks into a multi-stage ring
//   warp 1      MMA issuer     : one elected lane issues tcgen05.mma (cta_group::1 or ::2), fp32 D in TMEM
//   warp 2      TMEM allocator
//   warps 4..7  epilogue       : tcgen05.ld TMEM -> registers -> (bias / GELU / accumulate) -> bf16 ->
//                                128B-swizzled smem staging -> TMA store
// The accumulator is double buffered in TMEM (2 x BN columns) so the epilogue of tile i overlaps the
// main loop of tile i+1. Operands may be K-major ([rows, K], K contiguous) or MN-major ([K, rows],
// rows contiguous): forward uses K/K, dgrad K/MN, wgrad MN/MN, so no transposes are materialised.
//
// Replaces the cuBLAS calls behind every nn.Linear / Conv1D of the reference
// (reference: 01-single-gpu/train_llm.py:140,145 -> HF modeling -> F.linear).
#include "common.cuh"
#include <mutex>
#include <unordered_map>
#include <string>
#include <cstring>
#include <cstdarg>
#include <cstdio>

namespace b2 {

constexpr int BM = 128;  // rows of C per CTA
constexpr int BK = 64;   // k-block: one 128B swizzle span of bf16
constexpr int kEpiChunk = 64;
constexpr int kEpiStageBytes = BM * kEpiChunk * 2;  // 16 KB
constexpr int kNumEpiStages = 2;
constexpr int kSmemBudget = 232448 - 1024;  // max dynamic smem minus alignment slack

// EPI_DSWIGLU: C is d(gate|up) [M, 2I] of a SwiGLU whose output gradient dh [M, I] is THIS GEMM's
//   result (down-projection dgrad): each dh tile is combined in registers with the saved gate / up
//   pre-activations and stored as dgate (columns j) and dup (columns I + j); dh is never written.
// EPI_ROPE: rotary embedding applied to the first `rope_heads` 128-wide heads of each output row
//   (q and k heads of a fused QKV projection) before the store; the remaining heads (v) pass through.
enum Epilogue { EPI_NONE = 0, EPI_BIAS = 1, EPI_BIAS_GELU = 2, EPI_ACCUM = 3, EPI_DSWIGLU = 4, EPI_ROPE = 5 };

struct EpiArgs {
  const __nv_bfloat16* bias;  // [N] (EPI_BIAS*)
  const __nv_bfloat16* c_old; // [M, ldc] (EPI_ACCUM)
  int ldc;
  const __nv_bfloat16* gu;    // [M, ld_gu] saved gate | up (EPI_DSWIGLU)
  int ld_gu, inter;           // row stride of gu, I
  const __nv_bfloat16 *cosT, *sinT;  // [max_pos, 128] (EPI_ROPE)
  const int* pos;             // [M] positions or nullptr (row % seq_len)
  int rope_heads, seq_len;
};

// ---- fused compute + collective variants (tensor parallel, see b2_gemm_ag / b2_gemm_rs) ----------
// kFuse = 0: plain GEMM.
// kFuse = 1 (all-gather -> GEMM): the rows of A arrive from the peers in `W` chunks of `rows_per_rank`
//   rows while the kernel runs; chunk r is usable once flags[r] >= epoch. Tiles are visited starting
//   with the local chunk and the TMA producer waits for a chunk's flag before its first load.
// kFuse = 2 (GEMM -> reduce-scatter): row block r of C belongs to rank r. Each tile is stored with TMA
//   straight into slot `rank` of the owner's (peer-mapped) staging buffer, so the scatter traffic
//   overlaps the remaining main loop; the owner sums its W slots afterwards.
constexpr int kMaxTP = 8;
struct NoFuse {};
struct FuseAWait {
  const uint32_t* flags;
  uint32_t epoch;
  int rows_per_rank, rank, W;
};
struct FuseScatterC {
  CUtensorMap cm[kMaxTP];  // cm[r]: rank r's slot `rank`, shape [rows_per_rank, N]
  int rows_per_rank, rank, W;
};
// FSDP parameter all-gather -> GEMM (every variant of the kernel, switched on at run time by
//   gw.flags != nullptr): the weight (B operand) is a slice of a flat unit buffer that the copy engines are
//   still filling from the peers' shards, chunk by chunk, while the kernel runs; chunk c is present once
//   flags[c] >= epoch (the gathering stream writes the flag after the chunk's copy). The TMA producer waits
//   for the chunks that hold a tile's weight rows before the tile's first load, and the n-blocks are
//   visited starting with the rows that arrive first (the rank's own shard, then the peers in the order
//   they are pulled). All loads are from local memory.
struct GatherWaitArgs {
  const uint32_t* flags;  // nullptr: B is complete, no waiting
  uint32_t epoch;
  long long b_off;        // element offset of B[0][0] in the flat buffer the flags describe
  long long chunk_elems;  // elements per flag
  int n_chunks;
  int nb_start;           // first n-block to visit
  int ldb;
  unsigned long long timeout_ns;
};
// kFuse = 4 (gate/up projection + SwiGLU): B is the stacked [2I, K] weight. The two CTAs of a pair load
//   DIFFERENT row blocks of it -- CTA 0 the gate rows [j, j+128), CTA 1 the up rows [I+j, I+j+128) -- so
//   the 256-column accumulator of a tile holds gate in columns 0..127 and up in 128..255 for the same 128
//   outputs; the epilogue stores both pre-activations (backward needs them) into C [M, 2I] and
//   h = silu(gate) * up into H [M, I] through the tensor map `th`.
struct FuseSwiglu {
  CUtensorMap th;
  int inter;
};
template <class F> struct FuseKind { static constexpr int v = 0; };
template <> struct FuseKind<FuseSwiglu> { static constexpr int v = 4; };
template <> struct FuseKind<FuseAWait> { static constexpr int v = 1; };
template <> struct FuseKind<FuseScatterC> { static constexpr int v = 2; };

template <int kCta, int BN>
struct GemmCfg {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = (BN / kCta) * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiBytes = kNumEpiStages * kEpiStageBytes;
  static constexpr int kStagesRaw = (kSmemBudget - kEpiBytes - 256) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 256 + 1024;
  static constexpr int kTmemCols = 2 * BN;  // double-buffered accumulator (power of two)
};

__device__ __forceinline__ float gelu_new(float x) {
  // 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))  (HF "gelu_new")
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t = 1.f - 2.f / (1.f + __expf(2.f * u));
  return 0.5f * x * (1.f + t);
}

template <int kCta, bool kAK, bool kBK, int kEpi, int BN, class Fuse = NoFuse>
__global__ void __maxnreg__(168)   // 256 threads x 168 registers leave room for one 256 x 80 side-stream block
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, int M, int N, int K, EpiArgs epi,
                 const __grid_constant__ Fuse fuse, const __grid_constant__ CUtensorMap tmBh,
                 int split_from, const __grid_constant__ GatherWaitArgs gw) {
  using Cfg = GemmCfg<kCta, BN>;
  constexpr int kFuse = FuseKind<Fuse>::v;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // swizzle-128B operands need 1024B alignment; align by hand (slack is included in kSmemBytes)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * Cfg::kABytes;
  uint8_t* smem_c = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_c + Cfg::kEpiBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tfull_bar = bars + 2 * kStages;
  uint64_t* tempty_bar = bars + 2 * kStages + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const uint32_t cta_rank = (kCta == 2) ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    tma_prefetch_desc(&tmBh);
  }
  if (warp == 1 && elect_one()) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4 * kCta);  // one arrival per epilogue warp of each CTA
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<kCta>(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // ---- persistent tile schedule (identical in every role) ----
  const int tile_m = BM * kCta;
  const int num_m = (M + tile_m - 1) / tile_m;
  int num_n_ = (N + BN - 1) / BN;
  if constexpr (kFuse == 4) num_n_ = fuse.inter / (BN / 2);   // one tile = 128 gate + 128 up columns
  const int num_n = num_n_;
  const int num_tiles = num_m * num_n;
  const int num_k = (K + BK - 1) / BK;
  const int num_workers = gridDim.x / kCta;
  const int worker = blockIdx.x / kCta;
  // Tail split: tiles [split_from, num_tiles) -- the partial last wave -- are each cut into two
  // half-width (BN/2) units so the remainder spreads over twice as many workers (wave quantisation:
  // 256 tiles on 74 workers cost 4 waves whole, 3.5 split). split_from == num_tiles: no split.
  const int num_units = split_from + 2 * (num_tiles - split_from);
  auto unit_tile = [&](int u, int& t, int& n_off, int& bn_cur) {
    if (u < split_from) {
      t = u; n_off = 0; bn_cur = BN;
    } else {
      const int h = u - split_from;
      t = split_from + (h >> 1); n_off = (h & 1) * (BN / 2); bn_cur = BN / 2;
    }
  };
  constexpr int kGroupM = 8;
  auto tile_coord = [&](int t, int& mb, int& nb) {
    const int per_group = kGroupM * num_n;
    const int g = t / per_group;
    const int first_m = g * kGroupM;
    const int gsz = min(num_m - first_m, kGroupM);
    const int r = t - g * per_group;
    mb = first_m + r % gsz;
    nb = r / gsz;
    if constexpr (kFuse == 1) {
      // start with the m-blocks of the local chunk (already present), then the peers in ring order
      const int blocks_per_rank = fuse.rows_per_rank / (BM * kCta);
      mb = (mb + fuse.rank * blocks_per_rank) % num_m;
    }
    nb += gw.nb_start;   // 0 unless the weight is still being gathered
    if (nb >= num_n) nb -= num_n;
  };

  if (warp == 0) {
    // ================= TMA producer =================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      unsigned long long known = 0;  // kFuse == 3: chunks already seen present
      for (int u = worker; u < num_units; u += num_workers) {
        int t, n_off, bn_cur, mb, nb;
        unit_tile(u, t, n_off, bn_cur);
        tile_coord(t, mb, nb);
        const int m0 = mb * tile_m + (int)cta_rank * BM;
        int n0 = nb * BN + n_off + (int)cta_rank * (bn_cur / kCta);
        if constexpr (kFuse == 4) n0 = nb * (BN / 2) + (int)cta_rank * fuse.inter;
        const CUtensorMap* tb = bn_cur == BN ? &tmB : &tmBh;
        const uint32_t stage_tx = (uint32_t)(Cfg::kABytes + (bn_cur / kCta) * BK * 2) * kCta;
        if constexpr (kFuse == 1) {
          // rows of this m-block come from rank `owner`: wait until its chunk has landed
          const int owner = min(m0 / fuse.rows_per_rank, fuse.W - 1);
          if (owner != fuse.rank) {
            wait_flag_geq(fuse.flags + owner, fuse.epoch, 0);
            // the chunk was written by the copy engine: order it before the async-proxy (TMA) reads
            asm volatile("fence.proxy.async.global;" ::: "memory");
          }
        }
        if (gw.flags != nullptr) {
          // flat range of the weight rows this CTA loads for the tile (K-major: rows n0.. of length
          // ldb; MN-major: the tile's columns run through every row, i.e. the whole matrix)
          long long lo, hi;
          if constexpr (kBK) {
            lo = gw.b_off + (long long)n0 * gw.ldb;
            hi = gw.b_off + (long long)(min(n0 + bn_cur / kCta, kFuse == 4 ? 2 * N : N) - 1) * gw.ldb + K;
          } else {
            lo = gw.b_off;
            hi = gw.b_off + (long long)(K - 1) * gw.ldb + N;
          }
          int c_lo = (int)(lo / gw.chunk_elems);
          int c_hi = (int)((hi - 1) / gw.chunk_elems);
          if (c_hi > gw.n_chunks - 1) c_hi = gw.n_chunks - 1;
          bool waited = false;
          for (int c = c_lo; c <= c_hi; ++c) {
            if (!((known >> c) & 1ull)) {
              wait_flag_geq(gw.flags + c, gw.epoch, gw.timeout_ns);
              known |= 1ull << c;
              waited = true;
            }
          }
          if (waited) asm volatile("fence.proxy.async.global;" ::: "memory");
        }
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader) mbar_expect_tx(&full_bar[stage], stage_tx);
          const int k0 = kb * BK;
          uint8_t* sa = smem_a + stage * Cfg::kABytes;
          uint8_t* sb = smem_b + stage * Cfg::kBBytes;
          if constexpr (kAK) tma_load_2d<kCta>(&tmA, &full_bar[stage], sa, k0, m0, kEvictNormal);
          else tma_load_3d<kCta>(&tmA, &full_bar[stage], sa, 0, k0, m0 / 64, kEvictNormal);
          if constexpr (kBK) tma_load_2d<kCta>(tb, &full_bar[stage], sb, k0, n0, kEvictNormal);
          else tma_load_3d<kCta>(tb, &full_bar[stage], sb, 0, k0, n0 / 64, kEvictNormal);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (pair leader only) =================
    if (leader) {
      constexpr uint32_t idesc_full = make_idesc_bf16(BM * kCta, BN, !kAK, !kBK);
      constexpr uint32_t idesc_half = make_idesc_bf16(BM * kCta, BN / 2, !kAK, !kBK);
      // K-major: rows are 128B apart, 8-row groups 1024B apart (SBO); LBO unused.
      // MN-major: 64-element MN atoms are BK*128 B apart (LBO), 8-row K groups 1024B apart (SBO).
      constexpr uint32_t lbo_a = kAK ? 16 : BK * 128, lbo_b = kBK ? 16 : BK * 128;
      constexpr uint32_t kstep_a = kAK ? 32 : 2048, kstep_b = kBK ? 32 : 2048;  // bytes per UMMA_K=16
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int u = worker; u < num_units; u += num_workers) {
        const uint32_t idesc = u < split_from ? idesc_full : idesc_half;
        if constexpr (kCta == 2) mbar_wait_cluster(&tempty_bar[acc], acc_phase ^ 1);
        else mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          if constexpr (kCta == 2) mbar_wait_cluster(&full_bar[stage], phase);
          else mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_addr = smem_u32(smem_a + stage * Cfg::kABytes);
            const uint32_t b_addr = smem_u32(smem_b + stage * Cfg::kBBytes);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = make_smem_desc_sw128(a_addr + k * kstep_a, lbo_a, 1024);
              const uint64_t db = make_smem_desc_sw128(b_addr + k * kstep_b, lbo_b, 1024);
              umma_f16<kCta>(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
            }
            umma_commit<kCta>(&empty_bar[stage]);  // frees the smem slot once these MMAs retire
            if (kb == num_k - 1) umma_commit<kCta>(&tfull_bar[acc]);
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue =================
    const int ew = warp - 4;                // TMEM lane quarter owned by this warp
    const int row = ew * 32 + (int)lane_id();
    const bool epi_leader = threadIdx.x == 128;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t chunk_ctr = 0;
    for (int u = worker; u < num_units; u += num_workers) {
      int t, n_off, bn_cur, mb, nb;
      unit_tile(u, t, n_off, bn_cur);
      tile_coord(t, mb, nb);
      const int m0 = mb * tile_m + (int)cta_rank * BM;
      const int n0 = nb * BN + n_off;
      const int n_chunks = bn_cur / kEpiChunk;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + acc * BN;
      // hand the drained TMEM stage back to the MMA warp
      auto release_tmem = [&]() {
        tc_fence_before();
        __syncwarp();
        if (lane_id() == 0) {
          if constexpr (kCta == 2) mbar_arrive_cluster(&tempty_bar[acc], 0);
          else mbar_arrive(&tempty_bar[acc]);
        }
      };
      // one 64-column chunk (32 packed bf16x2 per row) -> 128B-swizzled staging -> TMA store
      auto emit = [&](const uint32_t (&packed)[32], const CUtensorMap* map, int col, int row0, bool valid) {
        // staging buffer reuse: the TMA store issued two chunks ago must have finished reading
        uint8_t* sc = smem_c + (chunk_ctr & 1) * kEpiStageBytes;
        if (epi_leader) tma_store_wait_read<kNumEpiStages - 1>();
        named_bar_sync(1, 128);
        const uint32_t row_addr = smem_u32(sc) + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t addr = row_addr + ((j ^ (row & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(packed[4 * j]),
                       "r"(packed[4 * j + 1]), "r"(packed[4 * j + 2]), "r"(packed[4 * j + 3])
                       : "memory");
        }
        fence_proxy_async_smem();
        named_bar_sync(1, 128);
        if (epi_leader) {
          if (valid) tma_store_2d(map, sc, col, row0);
          tma_store_commit();
        }
        ++chunk_ctr;
      };

      if constexpr (kFuse == 4) {
        // ---- gate | up accumulator -> gate, up, silu(gate) * up ----
        const int j0 = nb * (BN / 2);
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          uint32_t pg[32], pu[32], ph[32];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t rg[32], ru[32];
            tmem_ld_32x32b_x32(taddr + c * kEpiChunk + half * 32, rg);
            tmem_ld_32x32b_x32(taddr + BN / 2 + c * kEpiChunk + half * 32, ru);
            tmem_ld_wait();
            if (c == 1 && half == 1) release_tmem();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const uint32_t gb = pack_bf16(__uint_as_float(rg[2 * j]), __uint_as_float(rg[2 * j + 1]));
              const uint32_t ub = pack_bf16(__uint_as_float(ru[2 * j]), __uint_as_float(ru[2 * j + 1]));
              const float2 g = unpack_bf16(gb), u = unpack_bf16(ub);
              // the standalone kernel's arithmetic on the bf16-rounded pre-activations: silu rounded to
              // bf16 (as HF's act_fn output is), then the product
              const float s0 = unpack_bf16(pack_bf16(__fdividef(g.x, 1.f + __expf(-g.x)), 0.f)).x;
              const float s1 = unpack_bf16(pack_bf16(__fdividef(g.y, 1.f + __expf(-g.y)), 0.f)).x;
              pg[half * 16 + j] = gb;
              pu[half * 16 + j] = ub;
              ph[half * 16 + j] = pack_bf16(s0 * u.x, s1 * u.y);
            }
          }
          const bool valid = m0 < M;
          emit(pg, &tmC, j0 + c * kEpiChunk, m0, valid);
          emit(pu, &tmC, fuse.inter + j0 + c * kEpiChunk, m0, valid);
          emit(ph, &fuse.th, j0 + c * kEpiChunk, m0, valid);
        }
      } else if constexpr (kEpi == EPI_DSWIGLU) {
        // ---- dh tile (+ saved gate / up) -> dgate, dup ----
        const int grow = m0 + row;
#pragma unroll 1
        for (int c = 0; c < n_chunks; ++c) {
          const int col0 = n0 + c * kEpiChunk;
          uint32_t pdg[32], pdu[32];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t gq[16], uq[16];   // 32 saved gate / up values each, packed bf16x2
            const int colh = col0 + half * 32;
            if (grow < M && colh + 32 <= epi.inter) {
              const uint4* gp = reinterpret_cast<const uint4*>(epi.gu + (size_t)grow * epi.ld_gu + colh);
              const uint4* up = reinterpret_cast<const uint4*>(epi.gu + (size_t)grow * epi.ld_gu + epi.inter + colh);
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint4 a = gp[q], bq = up[q];
                gq[4 * q] = a.x; gq[4 * q + 1] = a.y; gq[4 * q + 2] = a.z; gq[4 * q + 3] = a.w;
                uq[4 * q] = bq.x; uq[4 * q + 1] = bq.y; uq[4 * q + 2] = bq.z; uq[4 * q + 3] = bq.w;
              }
            } else {
#pragma unroll
              for (int q = 0; q < 16; ++q) { gq[q] = 0u; uq[q] = 0u; }
            }
            uint32_t rd[32];
            tmem_ld_32x32b_x32(taddr + c * kEpiChunk + half * 32, rd);
            tmem_ld_wait();
            if (c == n_chunks - 1 && half == 1) release_tmem();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float2 g = unpack_bf16(gq[j]), u = unpack_bf16(uq[j]);
              const float d0 = __uint_as_float(rd[2 * j]), d1 = __uint_as_float(rd[2 * j + 1]);
              const float sg0 = __fdividef(1.f, 1.f + __expf(-g.x)), sg1 = __fdividef(1.f, 1.f + __expf(-g.y));
              pdu[half * 16 + j] = pack_bf16(d0 * (g.x * sg0), d1 * (g.y * sg1));
              pdg[half * 16 + j] = pack_bf16(d0 * u.x * (sg0 * (1.f + g.x * (1.f - sg0))),
                                             d1 * u.y * (sg1 * (1.f + g.y * (1.f - sg1))));
            }
          }
          const bool valid = col0 < N && m0 < M;
          emit(pdg, &tmC, col0, m0, valid);
          emit(pdu, &tmC, epi.inter + col0, m0, valid);
        }
      } else if constexpr (kEpi == EPI_ROPE) {
        // ---- rotary embedding on 128-wide heads (two 64-column chunks: low and high half of a head) ----
        const int grow = m0 + row;
        int p = 0;
        if (grow < M) p = epi.pos ? epi.pos[grow] : grow % epi.seq_len;
        const __nv_bfloat16* crow = epi.cosT + (size_t)p * 128;
        const __nv_bfloat16* srow = epi.sinT + (size_t)p * 128;
        const int n_heads_tile = n_chunks / 2;
#pragma unroll 1
        for (int h = 0; h < n_heads_tile; ++h) {
          const int col0 = n0 + h * 128;
          const bool rot = (col0 >> 7) < epi.rope_heads;
          uint32_t plo[32], phi[32];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t rl[32], rh[32];
            tmem_ld_32x32b_x32(taddr + h * 128 + half * 32, rl);
            tmem_ld_32x32b_x32(taddr + h * 128 + 64 + half * 32, rh);
            tmem_ld_wait();
            if (h == n_heads_tile - 1 && half == 1) release_tmem();
            if (rot) {
              const uint4* c0 = reinterpret_cast<const uint4*>(crow + half * 32);
              const uint4* c1 = reinterpret_cast<const uint4*>(crow + 64 + half * 32);
              const uint4* s0 = reinterpret_cast<const uint4*>(srow + half * 32);
              const uint4* s1 = reinterpret_cast<const uint4*>(srow + 64 + half * 32);
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint4 vc0 = c0[q], vc1 = c1[q], vs0 = s0[q], vs1 = s1[q];
                const uint32_t ac0[4] = {vc0.x, vc0.y, vc0.z, vc0.w}, ac1[4] = {vc1.x, vc1.y, vc1.z, vc1.w};
                const uint32_t as0[4] = {vs0.x, vs0.y, vs0.z, vs0.w}, as1[4] = {vs1.x, vs1.y, vs1.z, vs1.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const int j = 4 * q + e;
                  // the standalone kernel rotates the bf16-rounded projection output
                  const float2 a = unpack_bf16(pack_bf16(__uint_as_float(rl[2 * j]), __uint_as_float(rl[2 * j + 1])));
                  const float2 b = unpack_bf16(pack_bf16(__uint_as_float(rh[2 * j]), __uint_as_float(rh[2 * j + 1])));
                  const float2 cl = unpack_bf16(ac0[e]), ch = unpack_bf16(ac1[e]);
                  const float2 sl = unpack_bf16(as0[e]), sh = unpack_bf16(as1[e]);
                  plo[half * 16 + j] = pack_bf16(a.x * cl.x - b.x * sl.x, a.y * cl.y - b.y * sl.y);
                  phi[half * 16 + j] = pack_bf16(b.x * ch.x + a.x * sh.x, b.y * ch.y + a.y * sh.y);
                }
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                plo[half * 16 + j] = pack_bf16(__uint_as_float(rl[2 * j]), __uint_as_float(rl[2 * j + 1]));
                phi[half * 16 + j] = pack_bf16(__uint_as_float(rh[2 * j]), __uint_as_float(rh[2 * j + 1]));
              }
            }
          }
          emit(plo, &tmC, col0, m0, col0 < N && m0 < M);
          emit(phi, &tmC, col0 + 64, m0, col0 + 64 < N && m0 < M);
        }
      } else {
#pragma unroll 1
      for (int c = 0; c < n_chunks; ++c) {
        uint32_t r0[32], r1[32];
        tmem_ld_32x32b_x32(taddr + c * kEpiChunk, r0);
        tmem_ld_32x32b_x32(taddr + c * kEpiChunk + 32, r1);
        tmem_ld_wait();
        if (c == n_chunks - 1) release_tmem();   // accumulator fully drained to registers
        const int col0 = n0 + c * kEpiChunk;
        uint32_t packed[32];
        if constexpr (kEpi == EPI_NONE) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            packed[j] = pack_bf16(__uint_as_float(r0[2 * j]), __uint_as_float(r0[2 * j + 1]));
            packed[16 + j] = pack_bf16(__uint_as_float(r1[2 * j]), __uint_as_float(r1[2 * j + 1]));
          }
        } else if constexpr (kEpi == EPI_BIAS || kEpi == EPI_BIAS_GELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int col = col0 + 2 * j;
            float b0 = 0.f, b1 = 0.f;
            if (col < N) b0 = __bfloat162float(epi.bias[col]);
            if (col + 1 < N) b1 = __bfloat162float(epi.bias[col + 1]);
            float v0 = __uint_as_float(j < 16 ? r0[2 * j] : r1[2 * (j - 16)]) + b0;
            float v1 = __uint_as_float(j < 16 ? r0[2 * j + 1] : r1[2 * (j - 16) + 1]) + b1;
            if constexpr (kEpi == EPI_BIAS_GELU) { v0 = gelu_new(v0); v1 = gelu_new(v1); }
            packed[j] = pack_bf16(v0, v1);
          }
        } else {  // EPI_ACCUM: C = C_old + acc
          const int grow = m0 + row;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int col = col0 + 2 * j;
            float o0 = 0.f, o1 = 0.f;
            if (grow < M && col + 1 < N) {
              float2 o = unpack_bf16(*reinterpret_cast<const uint32_t*>(
                  epi.c_old + (size_t)grow * epi.ldc + col));
              o0 = o.x; o1 = o.y;
            } else if (grow < M && col < N) {
              o0 = __bfloat162float(epi.c_old[(size_t)grow * epi.ldc + col]);
            }
            float v0 = __uint_as_float(j < 16 ? r0[2 * j] : r1[2 * (j - 16)]) + o0;
            float v1 = __uint_as_float(j < 16 ? r0[2 * j + 1] : r1[2 * (j - 16) + 1]) + o1;
            packed[j] = pack_bf16(v0, v1);
          }
        }
        if constexpr (kFuse == 2) {
          const int owner = min(m0 / fuse.rows_per_rank, fuse.W - 1);
          emit(packed, &fuse.cm[owner], col0, m0 - owner * fuse.rows_per_rank, col0 < N && m0 < M);
        } else {
          emit(packed, &tmC, col0, m0, col0 < N && m0 < M);
        }
      }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (epi_leader) tma_store_wait<0>();
  }

  // ---- teardown ----
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) tmem_dealloc<kCta>(tmem_base, Cfg::kTmemCols);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static char g_err[1024] = "";
void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// cuTensorMapEncodeTiled is a driver call and needs a current context on the calling thread. A thread whose
// first CUDA work is one of our launches (an autograd worker whose first backward node is one of these ops)
// has had no runtime call yet, so bind the primary context once per thread.
static void ensure_context() {
  static thread_local bool ready = false;
  if (!ready) {
    cudaFree(nullptr);
    ready = true;
  }
}

static EncodeTiledFn get_encode() {
  ensure_context();
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// Encoded tensor maps are cached: a training step launches ~1,500 GEMMs on buffers whose addresses and
// shapes recur every step (flat parameter / gradient buffers, the caching allocator's activation blocks),
// and cuTensorMapEncodeTiled is a driver call per map. The key is everything the encoding depends on.
struct TmapKey {
  const void* ptr;
  uint64_t rows, cols, stride;
  uint32_t box_rows, box_cols, kind;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && stride == o.stride &&
           box_rows == o.box_rows && box_cols == o.box_cols && kind == o.kind;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = reinterpret_cast<uint64_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
    h ^= (k.rows + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
    h ^= (k.cols + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
    h ^= (k.stride + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
    h ^= (((uint64_t)k.box_rows << 40) ^ ((uint64_t)k.box_cols << 8) ^ k.kind) * 0xC2B2AE3D27D4EB4Full;
    return (size_t)h;
  }
};
static thread_local std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmap_cache;
constexpr size_t kTmapCacheMax = 8192;

static bool tmap_cache_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200TRAIN_TMAP_CACHE");
    return !(e && e[0] == '0');
  }();
  return on;
}
static bool tmap_lookup(const TmapKey& k, CUtensorMap* out) {
  if (!tmap_cache_enabled()) return false;
  auto it = g_tmap_cache.find(k);
  if (it == g_tmap_cache.end()) return false;
  *out = it->second;
  return true;
}
static void tmap_store(const TmapKey& k, const CUtensorMap& m) {
  if (!tmap_cache_enabled()) return;
  if (g_tmap_cache.size() >= kTmapCacheMax) g_tmap_cache.clear();
  g_tmap_cache.emplace(k, m);
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols,
                      uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_cols) {
  const TmapKey key{ptr, rows, cols, row_stride_elems, box_rows, box_cols, 0u};
  if (tmap_lookup(key, out)) return 0;
  EncodeTiledFn enc = get_encode();
  B2_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  B2_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA base %p not 16B aligned", ptr);
  B2_REQUIRE((row_stride_elems * 2) % 16 == 0, "TMA row stride %llu elems not 16B multiple",
             (unsigned long long)row_stride_elems);
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B2_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(2d) failed: %d (rows=%llu cols=%llu ld=%llu)",
             (int)r, (unsigned long long)rows, (unsigned long long)cols,
             (unsigned long long)row_stride_elems);
  tmap_store(key, *out);
  return 0;
}

int make_tmap_mn_bf16(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols,
                      uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_chunks) {
  const TmapKey key{ptr, rows, cols, row_stride_elems, box_rows, box_chunks, 1u};
  if (tmap_lookup(key, out)) return 0;
  EncodeTiledFn enc = get_encode();
  B2_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  B2_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA base %p not 16B aligned", ptr);
  B2_REQUIRE((row_stride_elems * 2) % 16 == 0, "TMA row stride not 16B multiple");
  B2_REQUIRE(cols % 64 == 0, "MN-major operand needs its contiguous extent (%llu) %% 64 == 0",
             (unsigned long long)cols);
  cuuint64_t dims[3] = {64, rows, cols / 64};
  cuuint64_t strides[2] = {row_stride_elems * 2, 128};
  cuuint32_t box[3] = {64, box_rows, box_chunks};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B2_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(3d) failed: %d", (int)r);
  tmap_store(key, *out);
  return 0;
}

int make_tmap_2d_f32(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols,
                     uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_cols) {
  const TmapKey key{ptr, rows, cols, row_stride_elems, box_rows, box_cols, 2u};
  if (tmap_lookup(key, out)) return 0;
  EncodeTiledFn enc = get_encode();
  B2_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  B2_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA base %p not 16B aligned", ptr);
  B2_REQUIRE((row_stride_elems * 4) % 16 == 0, "TMA row stride not 16B multiple");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * 4};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B2_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(f32) failed: %d", (int)r);
  tmap_store(key, *out);
  return 0;
}

int sm_count() {
  static int n[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (n[dev] == 0) cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
  return n[dev];
}

static bool tail_split_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200TRAIN_GEMM_TAIL_SPLIT");
    return !(e && e[0] == '0');
  }();
  return on;
}

template <int kCta, bool kAK, bool kBK, int kEpi, int BN, class Fuse = NoFuse>
static int launch_gemm(void* C, const void* A, const void* B, int M, int N, int K, int lda, int ldb,
                       int ldc, const EpiArgs& epi, int max_ctas, cudaStream_t stream,
                       const Fuse& fuse = Fuse(), bool allow_split = true,
                       const GatherWaitArgs& gw = GatherWaitArgs()) {
  using Cfg = GemmCfg<kCta, BN>;
  CUtensorMap tmA, tmB, tmC;
  int rc;
  if constexpr (kAK) rc = make_tmap_2d_bf16(&tmA, A, M, K, lda, BM, BK);
  else rc = make_tmap_mn_bf16(&tmA, A, K, M, lda, BK, BM / 64);
  if (rc) return rc;
  // SwiGLU: B is the stacked [2I, K] weight while the GEMM's N is I
  if constexpr (kBK) rc = make_tmap_2d_bf16(&tmB, B, FuseKind<Fuse>::v == 4 ? 2 * N : N, K, ldb, BN / kCta, BK);
  else rc = make_tmap_mn_bf16(&tmB, B, K, N, ldb, BK, BN / kCta / 64);
  if (rc) return rc;
  if constexpr (FuseKind<Fuse>::v == 2) {
    tmC = fuse.cm[0];  // unused by the scatter variant; keep the parameter well formed
  } else if constexpr (FuseKind<Fuse>::v == 4 || kEpi == EPI_DSWIGLU) {
    // C is the [M, 2I] gate | up (or d gate | d up) tensor while the GEMM's N is I
    rc = make_tmap_2d_bf16(&tmC, C, M, 2 * N, ldc, BM, kEpiChunk);
    if (rc) return rc;
  } else {
    rc = make_tmap_2d_bf16(&tmC, C, M, N, ldc, BM, kEpiChunk);
    if (rc) return rc;
  }

  auto kern = gemm_bf16_kernel<kCta, kAK, kBK, kEpi, BN, Fuse>;
  static bool attr_set[64] = {};   // per device: a process may drive several GPUs
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!attr_set[dev]) {
    B2_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::kSmemBytes));
    attr_set[dev] = true;
  }
  const int tile_m = BM * kCta;
  const int num_n_host = FuseKind<Fuse>::v == 4 ? N / (BN / 2) : (N + BN - 1) / BN;
  const int num_tiles = ((M + tile_m - 1) / tile_m) * num_n_host;
  int sms = sm_count();
  if (max_ctas > 0 && max_ctas < sms) sms = max_ctas;
  int workers = sms / kCta;
  if (workers > num_tiles) workers = num_tiles;
  if (workers < 1) workers = 1;

  // tail split (see the kernel): only when the last wave is at most half full, so the halves fit in one
  // more round; the fused variants order their tiles by chunk arrival and keep whole tiles
  int split_from = num_tiles;
  CUtensorMap tmBh = tmB;
  if (allow_split && tail_split_enabled() && BN == 256 && FuseKind<Fuse>::v == 0 && num_tiles > workers) {
    const int rem = num_tiles % workers;
    if (rem > 0 && 2 * rem <= workers) {
      if constexpr (kBK) rc = make_tmap_2d_bf16(&tmBh, B, N, K, ldb, BN / 2 / kCta, BK);
      else rc = make_tmap_mn_bf16(&tmBh, B, K, N, ldb, BK, BN / 2 / kCta / 64);
      if (rc) return rc;
      split_from = num_tiles - rem;
    }
  }

  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(workers * kCta);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = kCta;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B2_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, M, N, K, epi, fuse, tmBh, split_from, gw));
  return 0;
}

template <int kCta, int BN>
static int dispatch_layout(void* C, const void* A, const void* B, int M, int N, int K, int a_mn,
                           int b_mn, int lda, int ldb, int ldc, int epilogue, const EpiArgs& epi,
                           int max_ctas, cudaStream_t stream, bool allow_split = true) {
#define B2_GEMM_CASE(AK, BKM, EPI)                                                                 \
  return launch_gemm<kCta, AK, BKM, EPI, BN>(C, A, B, M, N, K, lda, ldb, ldc, epi, max_ctas, stream, \
                                             NoFuse(), allow_split)
  if (!a_mn && !b_mn) {
    switch (epilogue) {
      case EPI_NONE: B2_GEMM_CASE(true, true, EPI_NONE);
      case EPI_BIAS: B2_GEMM_CASE(true, true, EPI_BIAS);
      case EPI_BIAS_GELU: B2_GEMM_CASE(true, true, EPI_BIAS_GELU);
      case EPI_ACCUM: B2_GEMM_CASE(true, true, EPI_ACCUM);
    }
  } else if (!a_mn && b_mn) {
    switch (epilogue) {
      case EPI_NONE: B2_GEMM_CASE(true, false, EPI_NONE);
      case EPI_BIAS: B2_GEMM_CASE(true, false, EPI_BIAS);
      case EPI_BIAS_GELU: B2_GEMM_CASE(true, false, EPI_BIAS_GELU);
      case EPI_ACCUM: B2_GEMM_CASE(true, false, EPI_ACCUM);
    }
  } else if (a_mn && b_mn) {
    switch (epilogue) {
      case EPI_NONE: B2_GEMM_CASE(false, false, EPI_NONE);
      case EPI_ACCUM: B2_GEMM_CASE(false, false, EPI_ACCUM);
    }
  } else {
    switch (epilogue) {
      case EPI_NONE: B2_GEMM_CASE(false, true, EPI_NONE);
    }
  }
#undef B2_GEMM_CASE
  set_last_error("b2_gemm_bf16: unsupported layout/epilogue combination a_mn=%d b_mn=%d epi=%d",
                 a_mn, b_mn, epilogue);
  return -2;
}

typedef CUresult (*WriteValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static WriteValue32Fn get_write_value() {
  static WriteValue32Fn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<WriteValue32Fn>(p);
  });
  return fn;
}

typedef CUresult (*WaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static WaitValue32Fn get_wait_value() {
  static WaitValue32Fn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<WaitValue32Fn>(p);
  });
  return fn;
}

// bound of the in-kernel waits on arrival flags (B200TRAIN_FLAG_TIMEOUT_MS, default 30 s): a gather that
// never arrives turns into a trap with a message instead of a GEMM that spins for ever
static unsigned long long flag_wait_timeout_ns() {
  static const unsigned long long ns = [] {
    const char* e = getenv("B200TRAIN_FLAG_TIMEOUT_MS");
    const unsigned long long ms = e ? strtoull(e, nullptr, 10) : 30000ull;
    return ms * 1000000ull;
  }();
  return ns;
}

// Describe a weight that is still being gathered (flags == nullptr: none). num_n_blocks / rows_per_block:
// the K-major n-block structure used to pick the first block to visit (0 blocks: MN-major, no rotation).
static int make_gather_wait(GatherWaitArgs* gw, const unsigned int* flags, unsigned int epoch, long long b_off,
                            long long chunk_elems, int n_chunks, int chunks_per_rank, int rank, int W, int ldb,
                            int num_n_blocks, int rows_per_block) {
  *gw = GatherWaitArgs();
  if (flags == nullptr) return 0;
  B2_REQUIRE(n_chunks >= 1 && n_chunks <= 64 && chunk_elems > 0 && chunks_per_rank >= 1 && W >= 1,
             "gather wait: %d chunks (at most 64) of %lld elements", n_chunks, chunk_elems);
  gw->flags = flags;
  gw->epoch = epoch;
  gw->b_off = b_off;
  gw->chunk_elems = chunk_elems;
  gw->n_chunks = n_chunks;
  gw->ldb = ldb;
  gw->timeout_ns = flag_wait_timeout_ns();
  gw->nb_start = 0;
  int best = W + 1;  // visit first the n-blocks whose rows arrive first: own shard, then the peers in pull order
  for (int nb = 0; nb < num_n_blocks; ++nb) {
    long long c = (b_off + (long long)nb * rows_per_block * ldb) / chunk_elems;
    if (c > n_chunks - 1) c = n_chunks - 1;
    const int owner = (int)(c / chunks_per_rank) % W;
    const int dist = (owner - rank + W) % W;
    if (dist < best) { best = dist; gw->nb_start = nb; }
  }
  return 0;
}

// out[i] = sum_w slots[w * n + i]   (bf16, fp32 accumulate): the local half of a scattered reduction
__global__ void sum_slots_kernel(uint4* __restrict__ out, const uint4* __restrict__ slots, size_t n16,
                                 int W) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16;
       i += (size_t)gridDim.x * blockDim.x) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int w = 0; w < W; ++w) {
      const uint4 v = slots[(size_t)w * n16 + i];
      const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16(u[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
    uint4 o;
    o.x = pack_bf16(acc[0], acc[1]);
    o.y = pack_bf16(acc[2], acc[3]);
    o.z = pack_bf16(acc[4], acc[5]);
    o.w = pack_bf16(acc[6], acc[7]);
    out[i] = o;
  }
}

}  // namespace b2

extern "C" {

int b2_local_copy(void* dst, const void* src, size_t bytes, int blocks, cudaStream_t stream);  // comm.cu

const char* b2_last_error() { return b2::last_error(); }

// ---- tensor-parallel fused GEMMs ---------------------------------------------------------------
// All-gather -> GEMM. A [M, K] is a local buffer whose row chunks (rows_per_rank each) are filled by
// b2_peer_all_gather_chunks on another stream while this kernel runs; flags[r] >= epoch marks chunk r
// as present. C[M, N] = A * B^T (b_mn = 1: B stored [K, N]).
int b2_gemm_ag(void* C, const void* A, const void* B, int M, int N, int K, int b_mn, int lda, int ldb,
               int ldc, const unsigned int* flags, unsigned int epoch, int rows_per_rank, int rank,
               int W, cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(W >= 1 && W <= kMaxTP && rows_per_rank * W == M, "gemm_ag: M=%d != W=%d x rows=%d", M, W,
             rows_per_rank);
  B2_REQUIRE(rows_per_rank % (2 * BM) == 0, "gemm_ag: rows per rank (%d) must be a multiple of %d",
             rows_per_rank, 2 * BM);
  FuseAWait f;
  f.flags = flags;
  f.epoch = epoch;
  f.rows_per_rank = rows_per_rank;
  f.rank = rank;
  f.W = W;
  EpiArgs epi = {};
  if (b_mn)
    return launch_gemm<2, true, false, EPI_NONE, 256, FuseAWait>(C, A, B, M, N, K, lda, ldb, ldc, epi,
                                                                 0, stream, f);
  return launch_gemm<2, true, true, EPI_NONE, 256, FuseAWait>(C, A, B, M, N, K, lda, ldb, ldc, epi, 0,
                                                              stream, f);
}

// GEMM -> reduce-scatter (scatter half). slot_bases[r]: address of slot `rank` in rank r's staging
// buffer (peer-mapped), each slot [rows_per_rank, ld_slot >= N]. Row block r of A * B^T is TMA-stored
// into slot_bases[r]. After a cross-rank barrier the owner sums its W slots (b2_sum_slots).
int b2_gemm_rs(void* const* slot_bases, const void* A, const void* B, int M, int N, int K, int b_mn,
               int lda, int ldb, int ld_slot, int rows_per_rank, int rank, int W,
               cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(W >= 1 && W <= kMaxTP && rows_per_rank * W == M, "gemm_rs: M=%d != W=%d x rows=%d", M, W,
             rows_per_rank);
  B2_REQUIRE(rows_per_rank % (2 * BM) == 0, "gemm_rs: rows per rank (%d) must be a multiple of %d",
             rows_per_rank, 2 * BM);
  FuseScatterC f;
  for (int r = 0; r < kMaxTP; ++r) {
    const int src = r < W ? r : 0;
    if (int rc = make_tmap_2d_bf16(&f.cm[r], slot_bases[src], rows_per_rank, N, ld_slot, BM, kEpiChunk))
      return rc;
  }
  f.rows_per_rank = rows_per_rank;
  f.rank = rank;
  f.W = W;
  EpiArgs epi = {};
  if (b_mn)
    return launch_gemm<2, true, false, EPI_NONE, 256, FuseScatterC>(nullptr, A, B, M, N, K, lda, ldb,
                                                                    ld_slot, epi, 0, stream, f);
  return launch_gemm<2, true, true, EPI_NONE, 256, FuseScatterC>(nullptr, A, B, M, N, K, lda, ldb,
                                                                 ld_slot, epi, 0, stream, f);
}

// FSDP all-gather -> GEMM. B is a weight inside a flat unit buffer that is being gathered chunk by chunk
// (b2_peer_all_gather_flagged): flags[c] >= epoch marks chunk c (chunk_elems elements, chunks_per_rank
// consecutive chunks per owning rank) as present in local memory. C[M, N] = epilogue(A * B^T), A [M, K]
// K-major; b_mn = 0: B stored [N, K] (forward), 1: B stored [K, N] (dgrad; waits for the whole matrix).
// epilogue: 0 none, 1 + bias[N]. b_off: element offset of B inside the flat buffer.
int b2_gemm_bf16_bwait(void* C, const void* A, const void* B, int M, int N, int K, int b_mn, int lda,
                       int ldb, int ldc, int epilogue, const void* bias, const unsigned int* flags,
                       unsigned int epoch, long long b_off, long long chunk_elems, int n_chunks,
                       int chunks_per_rank, int rank, int W, cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(M > BM, "gemm_bwait: M=%d must exceed %d (2-CTA tiles)", M, BM);
  B2_REQUIRE(epilogue == EPI_NONE || epilogue == EPI_BIAS, "gemm_bwait: unsupported epilogue %d", epilogue);
  GatherWaitArgs gw;
  if (int rc = make_gather_wait(&gw, flags, epoch, b_off, chunk_elems, n_chunks, chunks_per_rank, rank, W, ldb,
                                b_mn ? 0 : (N + 255) / 256, 256))
    return rc;
  EpiArgs epi = {};
  epi.bias = static_cast<const __nv_bfloat16*>(bias);
  if (b_mn) {
    if (epilogue == EPI_BIAS)
      return launch_gemm<2, true, false, EPI_BIAS, 256>(C, A, B, M, N, K, lda, ldb, ldc, epi, 0, stream, NoFuse(), true, gw);
    return launch_gemm<2, true, false, EPI_NONE, 256>(C, A, B, M, N, K, lda, ldb, ldc, epi, 0, stream, NoFuse(), true, gw);
  }
  if (epilogue == EPI_BIAS)
    return launch_gemm<2, true, true, EPI_BIAS, 256>(C, A, B, M, N, K, lda, ldb, ldc, epi, 0, stream, NoFuse(), true, gw);
  return launch_gemm<2, true, true, EPI_NONE, 256>(C, A, B, M, N, K, lda, ldb, ldc, epi, 0, stream, NoFuse(), true, gw);
}

// ---- GEMMs with fused elementwise neighbours -------------------------------------------------------
// Every function below takes the optional gather-wait description of its weight (flags == nullptr: none).
//
// gate/up projection + SwiGLU: GU[M, 2I] = x W^T (gate | up pre-activations, kept for backward) and
// H[M, I] = silu(gate) * up, W stacked [2I, K]. Requires I % 128 == 0 and M > 128.
int b2_gemm_swiglu(void* GU, void* H, const void* A, const void* B, int M, int I, int K, int lda, int ldb,
                   int ld_gu, int ld_h, const unsigned int* flags, unsigned int epoch, long long b_off,
                   long long chunk_elems, int n_chunks, int chunks_per_rank, int rank, int W,
                   cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(M > BM && I % 128 == 0, "gemm_swiglu: M=%d (> %d) I=%d (%% 128)", M, BM, I);
  GatherWaitArgs gw;
  if (int rc = make_gather_wait(&gw, flags, epoch, b_off, chunk_elems, n_chunks, chunks_per_rank, rank, W, ldb,
                                I / 128, 128))
    return rc;
  FuseSwiglu f;
  if (int rc = make_tmap_2d_bf16(&f.th, H, M, I, ld_h, BM, kEpiChunk)) return rc;
  f.inter = I;
  EpiArgs epi = {};
  return launch_gemm<2, true, true, EPI_NONE, 256, FuseSwiglu>(GU, A, B, M, I, K, lda, ldb, ld_gu, epi, 0, stream,
                                                               f, false, gw);
}

// down-projection dgrad + SwiGLU backward: dGU[M, 2I] = dswiglu(dY W_down, GU) with W_down stored [H, I]
// (the dgrad operand layout); dh = dY W_down never leaves the chip.
int b2_gemm_dswiglu(void* dGU, const void* dY, const void* Wd, const void* GU, int M, int I, int Hdim, int lda,
                    int ldb, int ld_dgu, int ld_gu, const unsigned int* flags, unsigned int epoch,
                    long long b_off, long long chunk_elems, int n_chunks, int chunks_per_rank, int rank,
                    int W, cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(M > BM && I % 64 == 0, "gemm_dswiglu: M=%d (> %d) I=%d (%% 64)", M, BM, I);
  GatherWaitArgs gw;
  if (int rc = make_gather_wait(&gw, flags, epoch, b_off, chunk_elems, n_chunks, chunks_per_rank, rank, W, ldb,
                                0, 256))
    return rc;
  EpiArgs epi = {};
  epi.gu = static_cast<const __nv_bfloat16*>(GU);
  epi.ld_gu = ld_gu;
  epi.inter = I;
  // C's tensor map spans [M, 2I]; the GEMM's N is I
  return launch_gemm<2, true, false, EPI_DSWIGLU, 256>(dGU, dY, Wd, M, I, Hdim, lda, ldb, ld_dgu, epi, 0, stream,
                                                       NoFuse(), true, gw);
}

// QKV projection + rotary embedding: C[M, N] = x W^T with the first rope_heads 128-wide heads of every
// row rotated by the row's position (pos[row], or row % seq_len when pos is null); cos / sin [max_pos, 128].
int b2_gemm_rope(void* C, const void* A, const void* B, int M, int N, int K, int lda, int ldb, int ldc,
                 const void* cosT, const void* sinT, const int* pos, int rope_heads, int seq_len,
                 const unsigned int* flags, unsigned int epoch, long long b_off, long long chunk_elems,
                 int n_chunks, int chunks_per_rank, int rank, int W, cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(M > BM && N % 128 == 0, "gemm_rope: M=%d (> %d) N=%d (%% 128)", M, BM, N);
  GatherWaitArgs gw;
  if (int rc = make_gather_wait(&gw, flags, epoch, b_off, chunk_elems, n_chunks, chunks_per_rank, rank, W, ldb,
                                (N + 255) / 256, 256))
    return rc;
  EpiArgs epi = {};
  epi.cosT = static_cast<const __nv_bfloat16*>(cosT);
  epi.sinT = static_cast<const __nv_bfloat16*>(sinT);
  epi.pos = pos;
  epi.rope_heads = rope_heads;
  epi.seq_len = seq_len;
  return launch_gemm<2, true, true, EPI_ROPE, 256>(C, A, B, M, N, K, lda, ldb, ldc, epi, 0, stream, NoFuse(), true,
                                                   gw);
}

int b2_sum_slots(void* out, const void* slots, size_t n_elems, int W, cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(n_elems % 8 == 0, "sum_slots: n not a multiple of 8");
  const size_t n16 = n_elems / 8;
  int blocks = (int)((n16 + 255) / 256);
  if (blocks > sm_count() * 8) blocks = sm_count() * 8;
  if (blocks < 1) blocks = 1;
  B2_SIDE_KERNEL((sum_slots_kernel));
  sum_slots_kernel<<<blocks, 256, 0, stream>>>((uint4*)out, (const uint4*)slots, n16, W);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Copy-engine all-gather in chunks with a completion flag per chunk: for every peer r != rank the rows
// of chunk r are copied from peer_bases[r] + r * chunk_bytes into dst + r * chunk_bytes and then
// flags[r] = epoch is written by the stream (cuStreamWriteValue32, no SM involved), which is what
// b2_gemm_ag polls.
int b2_peer_all_gather_chunks(void* dst, void* const* peer_bases, size_t chunk_bytes, int rank, int W,
                              unsigned int* flags, unsigned int epoch, cudaStream_t stream) {
  using namespace b2;
  WriteValue32Fn wv = get_write_value();
  B2_REQUIRE(wv != nullptr, "cuStreamWriteValue32 entry point unavailable");
  for (int k = 1; k < W; ++k) {
    const int peer = (rank + k) % W;
    B2_CUDA_CHECK(cudaMemcpyAsync(static_cast<char*>(dst) + (size_t)peer * chunk_bytes,
                                  static_cast<const char*>(peer_bases[peer]) + (size_t)peer * chunk_bytes,
                                  chunk_bytes, cudaMemcpyDeviceToDevice, stream));
    CUresult r = wv((CUstream)stream, (CUdeviceptr)(flags + peer), epoch, 0);
    B2_REQUIRE(r == CUDA_SUCCESS, "cuStreamWriteValue32 failed: %d", (int)r);
  }
  return 0;
}

// FSDP parameter gather with an arrival flag per chunk. dst: local full buffer [W * shard_bytes];
// srcs[r]: rank r's shard (peer-mapped). Every shard is cut into `sub` chunks; chunk (r, j) is copied by
// the copy engines and then flags[r * sub + j] = epoch is written by the stream (no SM involved), in the
// order own shard, rank+1, rank+2, ... The own shard is placed by `self_blocks` copy blocks.
// part: 0 = own shard only, 1 = peers only, 2 = both (callers record an event between the two parts so that
// consumers order themselves behind the local copy KERNEL with a normal event and only ever spin on flags
// that copy engines and stream memory operations produce).
int b2_peer_all_gather_flagged(void* dst, void* const* srcs, size_t shard_bytes, int rank, int W, int sub,
                               int self_blocks, int part, unsigned int* flags, unsigned int epoch,
                               cudaStream_t stream) {
  using namespace b2;
  WriteValue32Fn wv = get_write_value();
  B2_REQUIRE(wv != nullptr, "cuStreamWriteValue32 entry point unavailable");
  B2_REQUIRE(sub >= 1 && W * sub <= 64 && shard_bytes % ((size_t)sub * 16) == 0,
             "all_gather_flagged: shard of %zu bytes in %d chunks", shard_bytes, sub);
  const size_t piece = shard_bytes / sub;
  for (int k = 0; k < W; ++k) {
    const int peer = (rank + k) % W;
    char* d = static_cast<char*>(dst) + (size_t)peer * shard_bytes;
    const char* sp = static_cast<const char*>(srcs[peer]);
    if (peer == rank && part == 1) continue;
    if (peer != rank && part == 0) continue;
    if (peer == rank) {
      if (int rc = b2_local_copy(d, sp, shard_bytes, self_blocks, stream)) return rc;
      for (int j = 0; j < sub; ++j) {
        CUresult r = wv((CUstream)stream, (CUdeviceptr)(flags + peer * sub + j), epoch, 0);
        B2_REQUIRE(r == CUDA_SUCCESS, "cuStreamWriteValue32 failed: %d", (int)r);
      }
      continue;
    }
    for (int j = 0; j < sub; ++j) {
      B2_CUDA_CHECK(cudaMemcpyAsync(d + (size_t)j * piece, sp + (size_t)j * piece, piece,
                                    cudaMemcpyDeviceToDevice, stream));
      CUresult r = wv((CUstream)stream, (CUdeviceptr)(flags + peer * sub + j), epoch, 0);
      B2_REQUIRE(r == CUDA_SUCCESS, "cuStreamWriteValue32 failed: %d", (int)r);
    }
  }
  return 0;
}

// Make `stream` wait (cuStreamWaitValue32, no SM involved) until flags[first .. first + count) >= epoch.
int b2_stream_wait_flags(const unsigned int* flags, int first, int count, unsigned int epoch,
                         cudaStream_t stream) {
  using namespace b2;
  WaitValue32Fn wf = get_wait_value();
  B2_REQUIRE(wf != nullptr, "cuStreamWaitValue32 entry point unavailable");
  for (int c = first; c < first + count; ++c) {
    CUresult r = wf((CUstream)stream, (CUdeviceptr)(flags + c), epoch, CU_STREAM_WAIT_VALUE_GEQ);
    B2_REQUIRE(r == CUDA_SUCCESS, "cuStreamWaitValue32 failed: %d", (int)r);
  }
  return 0;
}

// C[M,N] (row-major, ldc) = epilogue(A * B^T) with A logically [M,K], B logically [N,K].
//   a_mn = 0: A stored [M,K] (lda = row stride); a_mn = 1: A stored [K,M] (rows of length M).
//   b_mn = 0: B stored [N,K];                    b_mn = 1: B stored [K,N].
//   epilogue: 0 none, 1 +bias[N], 2 gelu_new(+bias), 3 C += (reads c_old == C, same ldc).
//   config: 0 auto; otherwise (cta_group) | (BN << 8), e.g. 2 | (256 << 8); bit 24 keeps the last
//   partial wave in whole tiles (no tail split).
//   max_ctas: cap on the persistent grid (0 = all SMs) so a concurrent kernel keeps some SMs.
int b2_gemm_bf16(void* C, const void* A, const void* B, int M, int N, int K, int a_mn, int b_mn,
                 int lda, int ldb, int ldc, int epilogue, const void* bias, int config,
                 int max_ctas, cudaStream_t stream) {
  using namespace b2;
  B2_REQUIRE(M > 0 && N > 0 && K > 0, "b2_gemm_bf16: empty problem %dx%dx%d", M, N, K);
  EpiArgs epi;
  epi.bias = static_cast<const __nv_bfloat16*>(bias);
  epi.c_old = static_cast<const __nv_bfloat16*>(C);
  epi.ldc = ldc;
  int cta = config & 0xff, bn = (config >> 8) & 0xffff;
  const bool allow_split = !(config & (1 << 24));
  if (cta == 0) cta = 2;
  if (bn == 0) bn = 256;
  // MN-major operands are fetched in 64-element atoms: each CTA's slice must be atom aligned
  if (cta == 2 && (M <= BM)) cta = 1;
  if (cta == 1 && bn == 256)
    return dispatch_layout<1, 256>(C, A, B, M, N, K, a_mn, b_mn, lda, ldb, ldc, epilogue, epi,
                                   max_ctas, stream, allow_split);
  if (cta == 1 && bn == 128)
    return dispatch_layout<1, 128>(C, A, B, M, N, K, a_mn, b_mn, lda, ldb, ldc, epilogue, epi,
                                   max_ctas, stream, allow_split);
  if (cta == 2 && bn == 256)
    return dispatch_layout<2, 256>(C, A, B, M, N, K, a_mn, b_mn, lda, ldb, ldc, epilogue, epi,
                                   max_ctas, stream, allow_split);
  if (cta == 2 && bn == 128)
    return dispatch_layout<2, 128>(C, A, B, M, N, K, a_mn, b_mn, lda, ldb, ldc, epilogue, epi,
                                   max_ctas, stream, allow_split);
  set_last_error("b2_gemm_bf16: bad config %d", config);
  return -2;
}

}  // extern "C"
