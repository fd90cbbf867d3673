/* C ABI of libb200train.so: hand-written sm_100a kernels for the b200train framework.
 *
 * Conventions
 *   - every function returns 0 on success; on failure it returns non-zero and b2_last_error()
 *     describes the problem (argument validation happens before any launch).
 *   - all tensors are device pointers to bf16 unless the type says otherwise; `stream` is the CUDA
 *     stream to launch on (a cudaStream_t passed as an opaque pointer).
 *   - no function allocates device memory unless its comment says so; scratch buffers are passed in.
 *   - parity citations point into the reference tree (rimelabs/lambda-labs_distributed-training-guide),
 *     which reaches the equivalent library kernels through PyTorch / HF transformers / flash-attn.
 */
#ifndef B200TRAIN_H_
#define B200TRAIN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b2_stream_t;

const char* b2_last_error(void);

/* ---- tcgen05 GEMM (csrc/gemm.cu) ----------------------------------------------------------
 * C[M,N] = epilogue(A * B^T), A logically [M,K], B logically [N,K], fp32 accumulate in TMEM.
 * a_mn / b_mn = 0: operand stored [rows, K] (K contiguous); = 1: stored [K, rows] (rows contiguous).
 * epilogue: 0 none, 1 + bias[N], 2 gelu_new(. + bias[N]), 3 C += result.
 * config: 0 auto, else cta_group | (BN << 8).  max_ctas: 0 = all SMs.
 * Reference: every nn.Linear / Conv1D matmul, 01-single-gpu/train_llm.py:140,145 (cuBLAS via ATen). */
int b2_gemm_bf16(void* C, const void* A, const void* B, int M, int N, int K, int a_mn, int b_mn,
                 int lda, int ldb, int ldc, int epilogue, const void* bias, int config,
                 int max_ctas, b2_stream_t stream);

/* ---- flash attention forward (csrc/attention_fwd.cu) --------------------------------------
 * Reference: attn_implementation="flash_attention_2", 05-training-llama-405b/train_llm.py:87;
 * varlen via restarting position ids, 00-rime/packed_dataset.py:35-48. */
int b2_attn_fwd(void* o, float* lse, const void* q, const void* k, const void* v,
                const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                int q_stride, int k_stride, int v_stride, int o_stride, float softmax_scale,
                int causal, b2_stream_t stream);

/* ---- flash attention backward (csrc/attention_bwd.cu) -------------------------------------
 * dq_accum: fp32 scratch [T, Hq*D] (zeroed by the call); delta: fp32 scratch [Hq, T]. */
int b2_attn_bwd(void* dq, void* dk, void* dv, float* dq_accum, float* delta, const void* dout,
                const void* q, const void* k, const void* v, const void* o, const float* lse,
                const int* cu_seqlens, int ndocs, int max_seqlen, int T, int Hq, int Hkv, int D,
                int q_stride, int k_stride, int v_stride, int o_stride, int do_stride,
                int dq_stride, int dk_stride, int dv_stride, float softmax_scale, int causal,
                b2_stream_t stream);

/* ---- norms (csrc/elementwise.cu) -----------------------------------------------------------
 * RMSNorm with the HF LlamaRMSNorm rounding; `res` (optional) is added first and the sum is
 * written to h_out. Backward needs an fp32 scratch `partial` of
 * b2_rmsnorm_bwd_partial_rows(T) * H floats (LayerNorm: * 2 * H).
 * Reference: HF LlamaRMSNorm / GPT-2 LayerNorm eager kernels behind 01-single-gpu/train_llm.py:140. */
int b2_rmsnorm_bwd_partial_rows(int T);
int b2_rmsnorm_fwd(void* y, float* rstd, void* h_out, const void* x, const void* res,
                   const void* w, int T, int H, float eps, b2_stream_t stream);
int b2_rmsnorm_bwd(void* dx, void* dw, float* partial, const void* dy, const void* h,
                   const void* w, const float* rstd, const void* dres, int T, int H,
                   int accumulate_dw, b2_stream_t stream);
/* RMSNorm backward with no scratch argument and no allocation: the fp32 partial sums go to a fixed
 * 32 MiB device array inside the library (held by every GPU that loads it, until exit). For callers
 * that cannot allocate: plugins in an application that has pooled all remaining device memory, or
 * latency-critical code that cannot take cudaMalloc's device-wide sync. Same results as
 * b2_rmsnorm_bwd. Refuses problems with b2_rmsnorm_bwd_partial_rows(T) * H above
 * b2_rmsnorm_bwd_static_capacity() floats. The first call on a device resolves the array's address
 * (make it before any stream capture). All calls share the array: overlapping calls on one GPU race.
 * No counterpart in the reference (it never manages scratch memory itself). */
size_t b2_rmsnorm_bwd_static_capacity(void);
int b2_rmsnorm_bwd_static(void* dx, void* dw, const void* dy, const void* h, const void* w,
                          const float* rstd, const void* dres, int T, int H, int accumulate_dw,
                          b2_stream_t stream);
int b2_layernorm_fwd(void* y, float* mean, float* rstd, void* h_out, const void* x,
                     const void* res, const void* w, const void* b, int T, int H, float eps,
                     b2_stream_t stream);
int b2_layernorm_bwd(void* dx, void* dw, void* db, float* partial, const void* dy, const void* h,
                     const void* w, const float* mean, const float* rstd, const void* dres, int T,
                     int H, int accumulate, b2_stream_t stream);

/* ---- activations / rope / embedding ---------------------------------------------------------
 * Reference: HF LlamaMLP act_fn(gate)*up, apply_rotary_pos_emb, nn.Embedding (eager ATen). */
int b2_swiglu_fwd(void* out, const void* gu, int T, int I, b2_stream_t stream);
int b2_swiglu_bwd(void* dgu, const void* dout, const void* gu, int T, int I, b2_stream_t stream);
int b2_gelu_fwd(void* out, const void* x, size_t n, b2_stream_t stream);
int b2_gelu_bwd(void* dx, const void* x, const void* dout, size_t n, b2_stream_t stream);
int b2_rope(void* x, const void* cosT, const void* sinT, const int* pos, int T, int nheads, int D,
            int row_stride, int seq_len, int backward, b2_stream_t stream);
int b2_embedding_fwd(void* out, const void* table, const long long* ids, int T, int H,
                     b2_stream_t stream);
int b2_embedding_bwd(void* dtable, const void* dout, const long long* sorted_ids,
                     const long long* perm, int T, int H, int accumulate, b2_stream_t stream);

/* ---- cross entropy ---------------------------------------------------------------------------
 * Reference: HF ForCausalLMLoss (fp32 log-softmax + nll), loss parallel 06-tensor-parallel/train_llm.py:239. */
int b2_ce_fwd(float* lse, float* loss, const void* logits, const long long* labels, int T, int V,
              int ld, int ignore_index, b2_stream_t stream);
int b2_ce_shard_stats(float* row_max, float* row_sum, float* row_tgt, const void* logits,
                      const long long* labels, int T, int V, int ld, int ignore_index,
                      int vocab_start, b2_stream_t stream);
int b2_ce_bwd(void* dlogits, const void* logits, const float* lse, const long long* labels, int T,
              int V, int ld, int ignore_index, int vocab_start, float scale, const float* scale_dev,
              b2_stream_t stream);

/* ---- optimizer --------------------------------------------------------------------------------
 * Reference: torch.optim.AdamW(fused=True) on bf16 params, 01-single-gpu/train_llm.py:65. */
int b2_adamw(void* p, void* m, void* v, const void* g, size_t n, float lr, float beta1,
             float beta2, float eps, float wd, int step, float grad_scale, int max_blocks,
             b2_stream_t stream);
int b2_add(void* out, const void* a, const void* b, size_t n, b2_stream_t stream);

/* ---- column sums (bias gradients) ---------------------------------------------------------------
 * out[N] (+)= sum_t x[t, :]; partial: fp32 scratch of b2_colsum_partial_rows(T, N) * N floats.
 * Reference: the bias gradients of GPT-2's Conv1D layers, computed by ATen reductions under
 * loss.backward() (02-distributed-data-parallel/train_llm.py:160). */
int b2_colsum_partial_rows(int T, int N);
int b2_colsum(void* out, float* partial, const void* x, int T, int N, int ld, int accumulate,
              b2_stream_t stream);

/* ---- host AdamW for CPU-offloaded shards (OpenMP) -----------------------------------------------
 * Reference: AdamW(fused=True) running on CPU under CPUOffload, 05-training-llama-405b/README.md:147-153. */
int b2_cpu_adamw_bf16(void* p, void* m, void* v, const void* g, size_t n, float lr, float beta1,
                      float beta2, float eps, float wd, int step, float grad_scale);
int b2_cpu_set_threads(int n);   /* OpenMP threads of the host kernels (n <= 0: query); returns the count */

/* ---- host dropout (OpenMP + AVX-512), same Philox mask as b2_dropout ---------------------------------
 * out = dropout(x) over n HOST elements of elem_size 2 (bf16) or 4 (fp32); out may alias x. Element e is kept
 * iff the device kernel keeps it for the same (p, seed, offset): backward calls this on the gradient.
 * Reference: GPT-2's three dropout sites on the CPU path of 01-single-gpu/train_llm.py (ATen bernoulli_). */
int b2_cpu_dropout(void* out, const void* x, size_t n, int elem_size, float p, unsigned long long seed,
                   unsigned long long offset);

/* ---- GEMMs with fused elementwise neighbours ----------------------------------------------------------
 * b2_gemm_swiglu : GU[M, 2I] = x W^T (gate | up, kept for backward) and H[M, I] = silu(gate) * up in one GEMM
 *                  on the stacked weight W [2I, K] (I % 128 == 0, M > 128).
 * b2_gemm_dswiglu: dGU[M, 2I] = SwiGLU backward of dh = dY W_down (W_down stored [H, I]) and the saved GU; dh is
 *                  formed in tensor memory and never written.
 * b2_gemm_rope   : C = x W^T with the first rope_heads 128-wide heads of each row rotated (cos / sin
 *                  [max_pos, 128], position = pos[row] or row % seq_len).
 * The trailing (flags .. W) arguments describe an in-flight all-gather of the weight (see below); flags == NULL
 * means the weight is complete. Reference: HF LlamaMLP.forward / apply_rotary_pos_emb behind
 * 01-single-gpu/train_llm.py:140 (Liger-patched in 00-rime/train_llm_01-single-gpu.py:51-54). */
int b2_gemm_swiglu(void* GU, void* H, const void* A, const void* B, int M, int I, int K, int lda, int ldb,
                   int ld_gu, int ld_h, const unsigned int* flags, unsigned int epoch, long long b_off,
                   long long chunk_elems, int n_chunks, int chunks_per_rank, int rank, int W, b2_stream_t stream);
int b2_gemm_dswiglu(void* dGU, const void* dY, const void* Wd, const void* GU, int M, int I, int Hdim, int lda,
                    int ldb, int ld_dgu, int ld_gu, const unsigned int* flags, unsigned int epoch,
                    long long b_off, long long chunk_elems, int n_chunks, int chunks_per_rank, int rank,
                    int W, b2_stream_t stream);
int b2_gemm_rope(void* C, const void* A, const void* B, int M, int N, int K, int lda, int ldb, int ldc,
                 const void* cosT, const void* sinT, const int* pos, int rope_heads, int seq_len,
                 const unsigned int* flags, unsigned int epoch, long long b_off, long long chunk_elems,
                 int n_chunks, int chunks_per_rank, int rank, int W, b2_stream_t stream);

/* ---- FSDP parameter all-gather fused into the consuming GEMM ---------------------------------------
 * b2_peer_all_gather_flagged pulls the peers' shards of a flat unit into the local full buffer with the
 * copy engines, `sub` chunks per shard, and writes flags[r * sub + j] = epoch after chunk (r, j) with a
 * stream memory operation. b2_gemm_bf16_bwait is the tcgen05 GEMM whose weight operand B lives in that
 * buffer (element offset b_off): its TMA producer waits for the chunks that hold a tile's weight rows
 * and visits first the n-blocks that arrive first. b2_stream_wait_flags makes a stream wait for chunks
 * (for consumers that are not GEMMs, e.g. norm weights). Every load is from local memory.
 * Reference: the implicit per-unit all-gather of FullyShardedDataParallel before forward / backward
 * (04-fully-sharded-data-parallel/train_llm.py:87-95,206,210). */
int b2_peer_all_gather_flagged(void* dst, void* const* srcs, size_t shard_bytes, int rank, int W, int sub,
                               int self_blocks, int part, unsigned int* flags, unsigned int epoch,
                               b2_stream_t stream);
int b2_gemm_bf16_bwait(void* C, const void* A, const void* B, int M, int N, int K, int b_mn, int lda,
                       int ldb, int ldc, int epilogue, const void* bias, const unsigned int* flags,
                       unsigned int epoch, long long b_off, long long chunk_elems, int n_chunks,
                       int chunks_per_rank, int rank, int W, b2_stream_t stream);
int b2_stream_wait_flags(const unsigned int* flags, int first, int count, unsigned int epoch,
                         b2_stream_t stream);
int b2_local_copy(void* dst, const void* src, size_t bytes, int blocks, b2_stream_t stream);

/* ZeRO-1 parameter re-assembly inside the optimizer kernel: b2_adamw on this rank's slice, the updated
 * parameters stored through the multicast address p_mc of the same slice (multimem.st), so every rank's
 * replica is updated by the owner -- the broadcast of ZeroRedundancyOptimizer.step()
 * (02-distributed-data-parallel/train_llm.py:81-83,163). Follow with b2_peer_barrier. */
int b2_adamw_multicast(void* p, void* p_mc, void* m, void* v, const void* g, size_t n, float lr, float beta1,
                       float beta2, float eps, float wd, int step, float grad_scale, int max_blocks,
                       b2_stream_t stream);

/* ---- NVLink / NVSwitch peer-memory collectives (one process per GPU, one node) -------------------
 * Pointer arrays hold W peer-mapped addresses of the same symmetric buffer, index = rank. `pads` is the
 * symmetric signal pad (uint32 [channel][16]); `epoch` increases by one per call and channel pair;
 * `counter` is one zeroed uint32 in local memory per stream.
 * Reference: the NCCL collectives issued by DistributedDataParallel / ZeroRedundancyOptimizer /
 * FullyShardedDataParallel (02-distributed-data-parallel/train_llm.py:65,81-83,
 * 04-fully-sharded-data-parallel/train_llm.py:87-95) and the separate optimizer.step() (04-...:213). */
int b2_peer_barrier(void* const* pads, int rank, int W, int channel, unsigned int epoch,
                    b2_stream_t stream);
int b2_peer_all_gather(void* dst, void* const* srcs, size_t shard_bytes, int rank, int W,
                       int skip_self, int max_blocks, b2_stream_t stream);
/* copy-engine gather; the rank's own shard is placed by `self_blocks` copy blocks (0: cudaMemcpyAsync) */
int b2_peer_all_gather_ce(void* dst, void* const* srcs, size_t shard_bytes, int rank, int W,
                          int skip_self, int self_blocks, b2_stream_t stream);
/* Copy-engine reduce-scatter: the peers' copies of this rank's slice are pulled into `staging` by the copy
 * engines (b2_ce_pull_slices) between a signal / stream-wait handshake on the symmetric pad
 * (b2_signal_peers + b2_stream_wait_pad: no kernel ever spins), then one local kernel sums the W slices
 * and either stores the reduced shard (b2_reduce_slices) or applies AdamW to the owned parameters with the
 * reduced gradient held in registers (b2_reduce_adamw). Reference: FSDP's reduce-scatter followed by a
 * separate optimizer.step() (04-fully-sharded-data-parallel/train_llm.py:210,213). */
int b2_signal_peers(void* const* pads, int rank, int W, int channel, unsigned int epoch, b2_stream_t stream);
int b2_stream_wait_pad(const void* pad_local, int rank, int W, int channel, unsigned int epoch,
                       b2_stream_t stream);
int b2_ce_pull_slices(void* staging, void* const* fulls, size_t shard_bytes, int rank, int W,
                      b2_stream_t stream);
int b2_reduce_slices(void* out, const void* g_local, const void* staging, int nstaged, size_t shard_elems,
                     float scale, int max_blocks, b2_stream_t stream);
int b2_reduce_adamw(void* p, void* m, void* v, const void* g_local, const void* staging, int nstaged,
                    size_t shard_elems, float grad_scale, float lr, float beta1, float beta2, float eps,
                    float wd, int step, int max_blocks, b2_stream_t stream);
/* Tensor-parallel plan, small collectives in device code / on the copy engines.
 * b2_ce_combine_peer: vocabulary-parallel cross entropy -- each rank combines the W shard statistics
 * (fp32 [3][T]: row max, sum exp(x - max), target logit) by reading the peers' symmetric buffers, instead of
 * the all-reduce(MAX) + all-reduce(SUM) of tp.loss_parallel (06-tensor-parallel/train_llm.py:239-243).
 * b2_peer_pull_2d: strided copy-engine pulls from every peer, the hidden-sharded -> sequence-sharded
 * exchange behind ColwiseParallel(output_layouts=Shard(1)) on embed_tokens (06-.../train_llm.py:84-88). */
int b2_ce_combine_peer(float* lse, float* loss_row, void* const* stats, const long long* labels, int T, int W,
                       int ignore_index, b2_stream_t stream);
int b2_peer_pull_2d(void* dst, size_t dst_pitch, size_t dst_step_bytes, void* const* srcs, size_t src_offset_bytes,
                    size_t src_pitch, size_t width_bytes, size_t rows, int rank, int W, b2_stream_t stream);
/* bound (ms) of every cross-rank spin wait on the current device; past it the waiting kernel reports the
 * rank / channel / epoch it is stuck on and traps (a dead peer becomes a launch failure, not a hang) */
int b2_comm_set_timeout_ms(unsigned long long ms);
int b2_peer_reduce_scatter(void* out_shard, void* const* fulls, size_t shard_elems, int rank, int W,
                           float scale, void* const* pads, int chan_base, unsigned int epoch,
                           unsigned int* counter, int max_blocks, b2_stream_t stream);
int b2_peer_reduce_scatter_adamw(void* p, void* m, void* v, void* const* fulls, size_t shard_elems,
                                 int rank, int W, float grad_scale, float lr, float beta1,
                                 float beta2, float eps, float wd, int step, void* const* pads,
                                 int chan_base, unsigned int epoch, unsigned int* counter,
                                 int max_blocks, b2_stream_t stream);
int b2_peer_all_reduce(void* const* bufs, size_t elems, int rank, int W, float scale,
                       void* const* pads, int chan_base, unsigned int epoch, unsigned int* counter,
                       int max_blocks, b2_stream_t stream);
/* The same reductions done inside the NVSwitch (NVLS): `mc_*` is the multicast address of the
 * symmetric buffer; multimem.ld_reduce returns the sum over ranks, multimem.st writes to every rank. */
int b2_nvls_reduce_scatter(void* out_shard, const void* mc_full, size_t shard_elems, int rank, int W,
                           float scale, void* const* pads, int chan_base, unsigned int epoch,
                           unsigned int* counter, int max_blocks, b2_stream_t stream);
int b2_nvls_reduce_scatter_adamw(void* p, void* m, void* v, const void* mc_full, size_t shard_elems,
                                 int rank, int W, float grad_scale, float lr, float beta1,
                                 float beta2, float eps, float wd, int step, void* const* pads,
                                 int chan_base, unsigned int epoch, unsigned int* counter,
                                 int max_blocks, b2_stream_t stream);
int b2_nvls_all_reduce(void* mc_buf, size_t elems, int rank, int W, float scale, void* const* pads,
                       int chan_base, unsigned int epoch, unsigned int* counter, int max_blocks,
                       b2_stream_t stream);

/* ---- tensor-parallel linears fused with their collective ------------------------------------------
 * Reference: PrepareModuleInput(Shard(1) -> Replicate()) + ColwiseParallel, and
 * RowwiseParallel(output_layouts=Shard(1)) (06-tensor-parallel/train_llm.py:89-115), i.e. NCCL
 * all-gather -> cuBLAS and cuBLAS -> NCCL reduce-scatter as separate steps. */
int b2_peer_all_gather_chunks(void* dst, void* const* peer_bases, size_t chunk_bytes, int rank, int W,
                              unsigned int* flags, unsigned int epoch, b2_stream_t stream);
int b2_gemm_ag(void* C, const void* A, const void* B, int M, int N, int K, int b_mn, int lda, int ldb,
               int ldc, const unsigned int* flags, unsigned int epoch, int rows_per_rank, int rank,
               int W, b2_stream_t stream);
int b2_gemm_rs(void* const* slot_bases, const void* A, const void* B, int M, int N, int K, int b_mn,
               int lda, int ldb, int ld_slot, int rows_per_rank, int rank, int W, b2_stream_t stream);
int b2_sum_slots(void* out, const void* slots, size_t n_elems, int W, b2_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200TRAIN_H_ */
