// NVLink / NVSwitch peer-memory collectives for one node (one process per GPU).
//
// Every rank maps every other rank's symmetric buffers into its address space (allocation and handle
// exchange happen in parallel/peer.py); the kernels below move data with plain 16-byte loads from
// peer pointers, reduce in fp32 registers, and synchronise ranks through flags in a symmetric signal
// pad (st.release.sys / ld.acquire.sys), so a collective is a single kernel per GPU with no NCCL call.
//
//   b2_peer_barrier                 all ranks reach a point (replaces dist.barrier, N12)
//   b2_peer_all_gather              pull the peers' shards into a local full buffer (FSDP N4, ZeRO-1 N3)
//   b2_peer_reduce_scatter          read slice r of every peer's full buffer, sum, scale (FSDP N5, DDP N2)
//   b2_peer_reduce_scatter_adamw    the same reduction feeding AdamW on the owned parameter shard in
//                                   registers: the reduced gradient shard never touches memory
//   b2_peer_all_reduce              two-phase all-reduce on a symmetric buffer (DDP buckets, TP norm grads)
//   b2_nvls_reduce_scatter(_adamw)  the reduction done by the NVSwitch: one multimem.ld_reduce on the
//   b2_nvls_all_reduce              multicast mapping returns the sum over all ranks (NVLS), and
//                                   multimem.st writes a result into every rank's copy at once
//
// Reference: the NCCL collectives issued implicitly by DistributedDataParallel / FullyShardedDataParallel
// / ZeroRedundancyOptimizer (02-distributed-data-parallel/train_llm.py:65,81-83,
// 04-fully-sharded-data-parallel/train_llm.py:87-95) and the separate optimizer.step() at :213.
#include "common.cuh"
#include <cstdio>

namespace b2 {

constexpr int kMaxPeers = 16;
struct PeerPtrs {
  void* p[kMaxPeers];
};

struct alignas(16) V4 { uint32_t u[4]; };

__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__device__ __forceinline__ V4 ld_peer(const V4* ptr) {
  V4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.u[0]), "=r"(r.u[1]), "=r"(r.u[2]), "=r"(r.u[3])
               : "l"(ptr));
  return r;
}
// peer data that another GPU may have just written: must not use the non-coherent path
__device__ __forceinline__ V4 ld_peer_coherent(const V4* ptr) {
  V4 r;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.u[0]), "=r"(r.u[1]), "=r"(r.u[2]), "=r"(r.u[3])
               : "l"(ptr)
               : "memory");
  return r;
}
__device__ __forceinline__ void st_flag_sys(uint32_t* ptr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(ptr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_flag_sys(const uint32_t* ptr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(ptr) : "memory");
  return v;
}

// Signal pad layout per rank: uint32 flags[channel][kMaxPeers]. A barrier on `channel` with sequence
// number `epoch` (monotonically increasing per channel): rank r stores epoch into slot [channel][r] of
// every peer's pad, then waits until all slots of its own pad reach epoch.
__device__ __forceinline__ void signal_peers(const PeerPtrs& pads, int rank, int W, int channel,
                                             uint32_t epoch) {
  if (threadIdx.x < W) {
    uint32_t* pad = static_cast<uint32_t*>(pads.p[threadIdx.x]);
    st_flag_sys(pad + channel * kMaxPeers + rank, epoch);
  }
}
// A peer that died (or never launched its half of the collective) would leave this spin running for
// ever, as an unkillable kernel with no diagnostic. The wait is therefore bounded: after
// g_spin_timeout_ns (default 120 s, b2_comm_set_timeout_ms) the waiting thread prints who it is waiting
// for and traps, which surfaces on the host as a launch failure of this rank.
__device__ unsigned long long g_spin_timeout_ns = 120ull * 1000ull * 1000ull * 1000ull;

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __noinline__ void spin_timeout_trap(int rank, int peer, int channel, uint32_t epoch,
                                               uint32_t seen) {
  printf("b200train comm: rank %d timed out waiting for rank %d on channel %d (epoch %u, last seen %u)\n",
         rank, peer, channel, epoch, seen);
  __trap();
}

__device__ __forceinline__ void wait_peers(const PeerPtrs& pads, int rank, int W, int channel,
                                           uint32_t epoch) {
  if (threadIdx.x < W) {
    const uint32_t* pad = static_cast<const uint32_t*>(pads.p[rank]);
    uint32_t seen = ld_flag_sys(pad + channel * kMaxPeers + threadIdx.x);
    if ((int32_t)(seen - epoch) < 0) {
      const unsigned long long t0 = global_timer_ns();
      uint32_t polls = 0;
      do {
        seen = ld_flag_sys(pad + channel * kMaxPeers + threadIdx.x);
        if ((++polls & 0x3ff) == 0 && global_timer_ns() - t0 > g_spin_timeout_ns)
          spin_timeout_trap(rank, (int)threadIdx.x, channel, epoch, seen);
      } while ((int32_t)(seen - epoch) < 0);
    }
  }
  __syncthreads();
}

__global__ void peer_barrier_kernel(PeerPtrs pads, int rank, int W, int channel, uint32_t epoch) {
  __threadfence_system();
  signal_peers(pads, rank, W, channel, epoch);
  wait_peers(pads, rank, W, channel, epoch);
}

// dst[peer * n .. ] = peer's shard, for every peer (own shard copied locally unless it already
// aliases dst). n16 = shard size in 16-byte units.
// 256 threads x <= 85 registers: a comm block fits beside a resident GEMM CTA (256 x <= 168 registers,
// all of the SM's shared memory), so collectives really overlap the persistent GEMMs
constexpr int kCommThreads = 256;

__global__ void __launch_bounds__(kCommThreads, 3)
peer_all_gather_kernel(V4* __restrict__ dst, PeerPtrs src, size_t n16, int rank, int W,
                       int skip_self) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int k = 0; k < W; ++k) {
    const int peer = (rank + k) % W;  // start with the local shard, then walk the ring
    if (peer == rank && skip_self) continue;
    const V4* s = static_cast<const V4*>(src.p[peer]);
    V4* d = dst + (size_t)peer * n16;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    // 8 independent 16-byte loads in flight per thread: NVLink round trips are ~4 us
    for (; i + 7 * stride < n16; i += 8 * stride) {
      V4 t[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) t[u] = ld_peer(s + i + u * stride);
#pragma unroll
      for (int u = 0; u < 8; ++u) d[i + u * stride] = t[u];
    }
    for (; i < n16; i += stride) d[i] = ld_peer(s + i);
  }
}

__device__ __forceinline__ void acc_bf16x8(float (&acc)[8], const V4& v) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = unpack_bf16(v.u[j]);
    acc[2 * j] += f.x;
    acc[2 * j + 1] += f.y;
  }
}

// counter-based "last block out" used to publish the completion flag once per kernel
__device__ __forceinline__ bool last_block_done(unsigned int* counter) {
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int prev = atomicAdd(counter, 1u);
    is_last = (prev == gridDim.x - 1);
    if (is_last) *counter = 0;  // ready for the next launch (launches on one stream are ordered)
  }
  __syncthreads();
  return is_last;
}

enum { kChanIn = 0, kChanOut = 1 };

// out_shard[i] = scale * sum_peers full_peer[rank * n + i]   (bf16 in/out, fp32 accumulate)
// If adam != 0 the reduced value is consumed by AdamW on (p, m, v) instead of being stored.
struct AdamArgs {
  __nv_bfloat16 *p, *m, *v;
  float decay, beta1, beta2, eps, step_size, inv_bc2_sqrt;
};

template <bool kAdam>
__device__ __forceinline__ void rs_finish(float (&acc)[8], size_t i, float scale, V4* out_shard,
                                          const AdamArgs& adam, V4 pv, V4 mv, V4 vv) {
  if constexpr (!kAdam) {
    V4 o;
#pragma unroll
    for (int j = 0; j < 4; ++j) o.u[j] = pack_bf16(acc[2 * j] * scale, acc[2 * j + 1] * scale);
    out_shard[i] = o;
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 pf = unpack_bf16(pv.u[j]), mf = unpack_bf16(mv.u[j]), vf = unpack_bf16(vv.u[j]);
      const float g0 = acc[2 * j] * scale, g1 = acc[2 * j + 1] * scale;
      mf.x = adam.beta1 * mf.x + (1.f - adam.beta1) * g0;
      mf.y = adam.beta1 * mf.y + (1.f - adam.beta1) * g1;
      vf.x = adam.beta2 * vf.x + (1.f - adam.beta2) * g0 * g0;
      vf.y = adam.beta2 * vf.y + (1.f - adam.beta2) * g1 * g1;
      pf.x = pf.x * adam.decay -
             adam.step_size * __fdividef(mf.x, sqrt_approx(vf.x) * adam.inv_bc2_sqrt + adam.eps);
      pf.y = pf.y * adam.decay -
             adam.step_size * __fdividef(mf.y, sqrt_approx(vf.y) * adam.inv_bc2_sqrt + adam.eps);
      pv.u[j] = pack_bf16(pf.x, pf.y);
      mv.u[j] = pack_bf16(mf.x, mf.y);
      vv.u[j] = pack_bf16(vf.x, vf.y);
    }
    reinterpret_cast<V4*>(adam.p)[i] = pv;
    reinterpret_cast<V4*>(adam.m)[i] = mv;
    reinterpret_cast<V4*>(adam.v)[i] = vv;
  }
}

// U vectors per thread per trip, all W peer loads of all U vectors issued before any reduction
template <bool kAdam, int U>
__global__ void __launch_bounds__(kCommThreads, 3)
peer_reduce_scatter_kernel(V4* __restrict__ out_shard, PeerPtrs full, size_t n16, int rank, int W,
                           float scale, PeerPtrs pads, int chan_base, uint32_t epoch,
                           unsigned int* counter, AdamArgs adam) {
  // every rank has finished producing its full buffer
  if (blockIdx.x == 0) {
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanIn, epoch);
  }
  wait_peers(pads, rank, W, chan_base + kChanIn, epoch);

  constexpr int kPeersPerTrip = 8 / U;  // U * kPeersPerTrip = 8 loads in flight
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n16; i0 += U * stride) {
    float acc[U][8];
    V4 pv[U], mv[U], vv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[u][j] = 0.f;
      if constexpr (kAdam) {  // optimizer state loads travel together with the peer loads
        const size_t i = i0 + u * stride;
        if (i < n16) {
          pv[u] = reinterpret_cast<V4*>(adam.p)[i];
          mv[u] = reinterpret_cast<V4*>(adam.m)[i];
          vv[u] = reinterpret_cast<V4*>(adam.v)[i];
        }
      }
    }
    for (int base = 0; base < W; base += kPeersPerTrip) {
      V4 vals[U][kPeersPerTrip];
#pragma unroll
      for (int k = 0; k < kPeersPerTrip; ++k)
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const size_t i = i0 + u * stride;
          if (base + k < W && i < n16) {
            const int peer = (rank + base + k) % W;
            vals[u][k] =
                ld_peer_coherent(static_cast<const V4*>(full.p[peer]) + (size_t)rank * n16 + i);
          }
        }
#pragma unroll
      for (int k = 0; k < kPeersPerTrip; ++k)
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (base + k < W && i0 + u * stride < n16) acc_bf16x8(acc[u], vals[u][k]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + u * stride;
      if (i < n16) rs_finish<kAdam>(acc[u], i, scale, out_shard, adam, pv[u], mv[u], vv[u]);
    }
  }
  // every rank has finished reading the full buffers: they may be overwritten
  if (last_block_done(counter)) {
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanOut, epoch);
    wait_peers(pads, rank, W, chan_base + kChanOut, epoch);
  }
}

template <bool kAdam>
static int launch_rs(int W, int blocks, cudaStream_t stream, V4* out, PeerPtrs full, size_t n16,
                     int rank, float scale, PeerPtrs pads, int chan_base, uint32_t epoch,
                     unsigned int* counter, AdamArgs a) {
  if (W <= 2 && !kAdam) {
    B2_SIDE_KERNEL((peer_reduce_scatter_kernel<kAdam, 4>));
    peer_reduce_scatter_kernel<kAdam, 4><<<blocks, kCommThreads, 0, stream>>>(
        out, full, n16, rank, W, scale, pads, chan_base, epoch, counter, a);
  } else if (W <= 4) {
    B2_SIDE_KERNEL((peer_reduce_scatter_kernel<kAdam, 2>));
    peer_reduce_scatter_kernel<kAdam, 2><<<blocks, kCommThreads, 0, stream>>>(
        out, full, n16, rank, W, scale, pads, chan_base, epoch, counter, a);
  } else {
    B2_SIDE_KERNEL((peer_reduce_scatter_kernel<kAdam, 1>));
    peer_reduce_scatter_kernel<kAdam, 1><<<blocks, kCommThreads, 0, stream>>>(
        out, full, n16, rank, W, scale, pads, chan_base, epoch, counter, a);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// In-place all-reduce of a symmetric buffer: phase 1 reduces slice `rank` from all peers into the
// local copy, phase 2 (after a barrier) pulls the other reduced slices.
// Contract: the kernel itself has no exit handshake. A rank may return while peers are still pulling
// its reduced slice, so `buf` must not be overwritten before a cross-rank ordering point on the same
// stream. PeerComm._all_reduce therefore launches b2_peer_barrier on a channel of its own right behind
// this kernel; a caller of the C entry point has to do the same.
__global__ void __launch_bounds__(256)
peer_all_reduce_kernel(PeerPtrs buf, size_t n16_slice, int rank, int W, float scale, PeerPtrs pads,
                       int chan_base, uint32_t epoch, unsigned int* counter) {
  if (blockIdx.x == 0) {
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanIn, epoch);
  }
  wait_peers(pads, rank, W, chan_base + kChanIn, epoch);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  V4* mine = static_cast<V4*>(buf.p[rank]);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16_slice; i += stride) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < W; ++k) {
      const int peer = (rank + k) % W;
      acc_bf16x8(acc, ld_peer_coherent(static_cast<const V4*>(buf.p[peer]) + (size_t)rank * n16_slice + i));
    }
    V4 o;
#pragma unroll
    for (int j = 0; j < 4; ++j) o.u[j] = pack_bf16(acc[2 * j] * scale, acc[2 * j + 1] * scale);
    mine[(size_t)rank * n16_slice + i] = o;
  }
  // all ranks have written their reduced slice (grid-wide + cross-rank barrier)
  __shared__ bool released;
  if (last_block_done(counter)) {
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanOut, epoch);
  }
  wait_peers(pads, rank, W, chan_base + kChanOut, epoch);
  (void)released;
  for (int k = 1; k < W; ++k) {
    const int peer = (rank + k) % W;
    const V4* s = static_cast<const V4*>(buf.p[peer]) + (size_t)peer * n16_slice;
    V4* d = mine + (size_t)peer * n16_slice;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16_slice; i += stride)
      d[i] = ld_peer_coherent(s + i);
  }
}

// ---- NVLS (in-switch reduction / multicast) variants -----------------------------------------------
// `mc` is the multicast mapping of a symmetric buffer: a load-reduce on it makes the NVSwitch fetch the
// same offset from every rank and return the fp32-accumulated sum (rounded to bf16 once), a store on
// it lands in every rank's copy. Per reduced 16 bytes a rank issues ONE load instead of W, and only
// 1/W of the unit crosses its NVLink receive side.
__device__ __forceinline__ V4 multimem_ld_reduce_bf16x8(const V4* mc) {
  V4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.u[0]), "=r"(r.u[1]), "=r"(r.u[2]), "=r"(r.u[3])
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st(V4* mc, const V4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.u[0]),
               "r"(v.u[1]), "r"(v.u[2]), "r"(v.u[3])
               : "memory");
}

template <bool kAdam>
__global__ void __launch_bounds__(kCommThreads, 3)
nvls_reduce_scatter_kernel(V4* __restrict__ out_shard, const V4* __restrict__ mc_full, size_t n16,
                           int rank, int W, float scale, PeerPtrs pads, int chan_base, uint32_t epoch,
                           unsigned int* counter, AdamArgs adam) {
  if (blockIdx.x == 0) {
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanIn, epoch);
  }
  wait_peers(pads, rank, W, chan_base + kChanIn, epoch);
  constexpr int U = kAdam ? 2 : 4;
  const V4* src = mc_full + (size_t)rank * n16;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n16; i0 += U * stride) {
    V4 g[U], pv[U], mv[U], vv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + u * stride;
      if (i < n16) {
        g[u] = multimem_ld_reduce_bf16x8(src + i);
        if constexpr (kAdam) {
          pv[u] = reinterpret_cast<V4*>(adam.p)[i];
          mv[u] = reinterpret_cast<V4*>(adam.m)[i];
          vv[u] = reinterpret_cast<V4*>(adam.v)[i];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + u * stride;
      if (i < n16) {
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        acc_bf16x8(acc, g[u]);
        rs_finish<kAdam>(acc, i, scale, out_shard, adam, pv[u], mv[u], vv[u]);
      }
    }
  }
  if (last_block_done(counter)) {  // every rank has finished reading: the full buffers may be reused
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanOut, epoch);
    wait_peers(pads, rank, W, chan_base + kChanOut, epoch);
  }
}

// In-place all-reduce: rank r load-reduces slice r through the switch and multicasts the result into
// every rank's buffer; one pass over 1/W of the data per rank.
__global__ void __launch_bounds__(kCommThreads, 3)
nvls_all_reduce_kernel(V4* __restrict__ mc, size_t n16_slice, int rank, int W, float scale,
                       PeerPtrs pads, int chan_base, uint32_t epoch, unsigned int* counter) {
  if (blockIdx.x == 0) {
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanIn, epoch);
  }
  wait_peers(pads, rank, W, chan_base + kChanIn, epoch);
  V4* slice = mc + (size_t)rank * n16_slice;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n16_slice; i0 += 4 * stride) {
    V4 g[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i0 + u * stride < n16_slice) g[u] = multimem_ld_reduce_bf16x8(slice + i0 + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i0 + u * stride < n16_slice) {
        if (scale != 1.f) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_bf16(g[u].u[j]);
            g[u].u[j] = pack_bf16(f.x * scale, f.y * scale);
          }
        }
        multimem_st(slice + i0 + u * stride, g[u]);
      }
  }
  // my slice has reached every rank, and every rank's slice has reached me
  if (last_block_done(counter)) {
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanOut, epoch);
    wait_peers(pads, rank, W, chan_base + kChanOut, epoch);
  }
}


// ---- local shard copy ---------------------------------------------------------------------------------
// The FSDP gather also places the rank's own shard into the full buffer. A cudaMemcpyAsync for that
// local device-to-device copy holds back every kernel launch of the other streams for its whole
// duration (measured: the next GEMM of the compute stream starts when the 218 MB copy ends, 75 us per
// unit at 2 ranks); a few copy blocks beside the GEMM CTAs do not.
__global__ void __launch_bounds__(kCommThreads, 3)
local_copy_kernel(V4* __restrict__ dst, const V4* __restrict__ src, size_t n16) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 7 * stride < n16; i += 8 * stride) {
    V4 t[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) t[u] = ld_peer(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < 8; ++u) dst[i + u * stride] = t[u];
  }
  for (; i < n16; i += stride) dst[i] = ld_peer(src + i);
}

// ---- software-pipelined reduce-scatter (+ AdamW) ------------------------------------------------------
// Per 16-byte output vector there is ONE long-latency fetch -- the peer's copy of the slice at 2 ranks,
// or an in-switch multimem.ld_reduce over all ranks (NVLS) -- and up to four short-latency local ones
// (own gradient slice, p, m, v). The long fetch runs D trips ahead in a register ring, the local loads
// are issued just in time, so an NVLink round trip (2-4 us) no longer gates the HBM traffic (~1 us) and
// the kernel needs neither many blocks nor many registers to keep both busy. 256 threads x <= 85
// registers: one such block fits beside a resident GEMM CTA on every SM.
__device__ __forceinline__ V4 ld_local(const V4* ptr) {
  V4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.u[0]), "=r"(r.u[1]), "=r"(r.u[2]), "=r"(r.u[3])
               : "l"(ptr));
  return r;
}

template <bool kAdam, bool kNvls, int U, int D>
__global__ void __launch_bounds__(kCommThreads, 3)
rs_pipe_kernel(V4* __restrict__ out_shard, const V4* __restrict__ remote, const V4* __restrict__ local,
               size_t n16, int rank, int W, float scale, PeerPtrs pads, int chan_base, uint32_t epoch,
               unsigned int* counter, AdamArgs adam) {
  if (blockIdx.x == 0) {
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanIn, epoch);
  }
  wait_peers(pads, rank, W, chan_base + kChanIn, epoch);

  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t lane0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t trips = (n16 + U * stride - 1) / (U * stride);
  auto index = [&](size_t t, int u) { return (t * U + u) * stride + lane0; };
  auto fetch = [&](size_t i) {
    if constexpr (kNvls) return multimem_ld_reduce_bf16x8(remote + i);
    else return ld_peer_coherent(remote + i);
  };
  V4 ring[D][U];
#pragma unroll
  for (int d = 0; d < D; ++d)
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = index(d, u);
      if (d < trips && i < n16) ring[d][u] = fetch(i);
    }
  for (size_t t0 = 0; t0 < trips; t0 += D) {
#pragma unroll
    for (int d = 0; d < D; ++d) {
      const size_t t = t0 + d;
      if (t >= trips) break;
      V4 gl[U], pv[U], mv[U], vv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t i = index(t, u);
        if (i < n16) {
          if constexpr (!kNvls) gl[u] = ld_local(local + i);
          if constexpr (kAdam) {
            pv[u] = ld_local(reinterpret_cast<const V4*>(adam.p) + i);
            mv[u] = ld_local(reinterpret_cast<const V4*>(adam.m) + i);
            vv[u] = ld_local(reinterpret_cast<const V4*>(adam.v) + i);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t i = index(t, u);
        if (i < n16) {
          float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          acc_bf16x8(acc, ring[d][u]);
          if constexpr (!kNvls) acc_bf16x8(acc, gl[u]);
          rs_finish<kAdam>(acc, i, scale, out_shard, adam, pv[u], mv[u], vv[u]);
        }
        const size_t inext = index(t + D, u);
        if (t + D < trips && inext < n16) ring[d][u] = fetch(inext);
      }
    }
  }
  if (last_block_done(counter)) {  // every rank has finished reading: the full buffers may be reused
    __threadfence_system();
    signal_peers(pads, rank, W, chan_base + kChanOut, epoch);
    wait_peers(pads, rank, W, chan_base + kChanOut, epoch);
  }
}


// ---- copy-engine reduce-scatter (+ AdamW) -------------------------------------------------------------
// Transport and arithmetic split by what each engine is good at. The peers' copies of this rank's slice
// are PULLED by the copy engines into a local staging area (full NVLink rate, no SM, nothing resident
// beside the GEMMs); cross-rank ordering is a one-warp signal kernel plus cuStreamWaitValue32 on the
// local signal pad; the reduction is then a purely local, HBM-bound kernel that sums the W slices and --
// fused -- applies AdamW to the owned shard, so the reduced gradient never exists in memory.
//   b2_signal_peers      store `epoch` into slot [channel][rank] of every peer's pad (release, system scope)
//   b2_stream_wait_pad   make the stream wait until every peer's slot of the LOCAL pad reached `epoch`
//   b2_ce_pull_slices    staging[k-1] = slice `rank` of rank (rank+k)%W's full buffer, k = 1..W-1
//   b2_reduce_adamw      p/m/v <- AdamW(scale * (g_local + sum staged))          (b2_reduce_slices: store)
__global__ void signal_peers_kernel(PeerPtrs pads, int rank, int W, int channel, uint32_t epoch) {
  __threadfence_system();
  if (threadIdx.x < W && (int)threadIdx.x != rank) {
    uint32_t* pad = static_cast<uint32_t*>(pads.p[threadIdx.x]);
    st_flag_sys(pad + channel * kMaxPeers + rank, epoch);
  }
}

constexpr int kMaxStaged = 15;
struct StagedPtrs {
  const V4* p[kMaxStaged];
};

// NST staged slices known at compile time (0 = runtime count): every load of a vector -- optimizer state,
// own slice, all staged slices -- is issued before the first use, one memory round trip per vector
// instead of one per pair of slices.
template <bool kAdam, int NST>
__global__ void __launch_bounds__(kCommThreads, 3)
reduce_local_kernel(V4* __restrict__ out, const V4* __restrict__ g_local, StagedPtrs staged, int nstaged,
                    size_t n16, float scale, AdamArgs adam) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
    V4 pv, mv, vv;
    if constexpr (kAdam) {
      pv = ld_local(reinterpret_cast<const V4*>(adam.p) + i);
      mv = ld_local(reinterpret_cast<const V4*>(adam.m) + i);
      vv = ld_local(reinterpret_cast<const V4*>(adam.v) + i);
    }
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const V4 gl = ld_local(g_local + i);
    if constexpr (NST > 0) {
      V4 st[NST];
#pragma unroll
      for (int k = 0; k < NST; ++k) st[k] = ld_local(staged.p[k] + i);
      acc_bf16x8(acc, gl);
#pragma unroll
      for (int k = 0; k < NST; ++k) acc_bf16x8(acc, st[k]);
    } else {
      acc_bf16x8(acc, gl);
      int k = 0;
      for (; k + 1 < nstaged; k += 2) {
        const V4 a = ld_local(staged.p[k] + i), b = ld_local(staged.p[k + 1] + i);
        acc_bf16x8(acc, a);
        acc_bf16x8(acc, b);
      }
      if (k < nstaged) acc_bf16x8(acc, ld_local(staged.p[k] + i));
    }
    rs_finish<kAdam>(acc, i, scale, out, adam, pv, mv, vv);
  }
}

template <bool kAdam>
static int launch_reduce_local(int blocks, cudaStream_t stream, V4* out, const V4* g_local, StagedPtrs sp,
                               int nstaged, size_t n16, float scale, const AdamArgs& a) {
#define B2_RL_CASE(N)                                                                                  \
  case N:                                                                                              \
    B2_SIDE_KERNEL((reduce_local_kernel<kAdam, N>));                                                   \
    reduce_local_kernel<kAdam, N><<<blocks, kCommThreads, 0, stream>>>(out, g_local, sp, nstaged, n16, \
                                                                        scale, a);                     \
    break;
  switch (nstaged) {
    B2_RL_CASE(1)
    B2_RL_CASE(3)
    B2_RL_CASE(7)
    default:
      B2_SIDE_KERNEL((reduce_local_kernel<kAdam, 0>));
      reduce_local_kernel<kAdam, 0><<<blocks, kCommThreads, 0, stream>>>(out, g_local, sp, nstaged, n16,
                                                                         scale, a);
  }
#undef B2_RL_CASE
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}


// ---- vocabulary-parallel cross entropy: statistics exchange through peer memory -------------------------
// Every rank has written its per-row shard statistics -- row max, sum exp(x - max), target logit -- as fp32
// [3][T] into a symmetric buffer. After a barrier each rank combines the W shards itself, reading the
// peers' buffers directly: global max, rescaled sum, log-sum-exp and the row loss. This replaces the
// all-reduce(MAX) + all-reduce(SUM) pair of the reference's loss_parallel (N10).
__global__ void __launch_bounds__(256)
ce_combine_peer_kernel(float* __restrict__ lse, float* __restrict__ loss_row, PeerPtrs stats,
                       const long long* __restrict__ labels, int T, int W, int ignore_index) {
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < T; t += gridDim.x * blockDim.x) {
    float mx[kMaxPeers], sm[kMaxPeers];
    float gmax = -INFINITY, tgt = 0.f;
    for (int r = 0; r < W; ++r) {
      const float* st = static_cast<const float*>(stats.p[r]);
      float a, b, c;
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(a) : "l"(st + t) : "memory");
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(b) : "l"(st + T + t) : "memory");
      asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(c) : "l"(st + 2 * (size_t)T + t) : "memory");
      mx[r] = a;
      sm[r] = b;
      tgt += c;
      gmax = fmaxf(gmax, a);
    }
    float total = 0.f;
    for (int r = 0; r < W; ++r) total += sm[r] * __expf(mx[r] - gmax);
    const float l = gmax + __logf(total);
    lse[t] = l;
    loss_row[t] = labels[t] == ignore_index ? 0.f : l - tgt;
  }
}

static bool rs_pipe_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200TRAIN_RS_PIPE");
    return !(e && e[0] == '0');
  }();
  return on;
}

static int clamp_blocks(size_t work16, int max_blocks) {
  size_t b = (work16 + kCommThreads - 1) / kCommThreads;
  int cap = max_blocks > 0 ? max_blocks : 64;
  if (b > (size_t)cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

static PeerPtrs to_ptrs(void* const* arr, int W) {
  PeerPtrs p;
  for (int i = 0; i < kMaxPeers; ++i) p.p[i] = i < W ? arr[i] : nullptr;
  return p;
}

}  // namespace b2

using namespace b2;

extern "C" {

int b2_peer_barrier(void* const* pads, int rank, int W, int channel, unsigned int epoch,
                    cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_SIDE_KERNEL((peer_barrier_kernel));
  peer_barrier_kernel<<<1, 32, 0, stream>>>(to_ptrs(pads, W), rank, W, channel, epoch);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// dst: local full buffer [W * shard_bytes]; srcs[r]: rank r's shard (peer-mapped pointers).
int b2_peer_all_gather(void* dst, void* const* srcs, size_t shard_bytes, int rank, int W,
                       int skip_self, int max_blocks, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_REQUIRE(shard_bytes % 16 == 0, "all_gather: shard size %zu not a multiple of 16 bytes", shard_bytes);
  if (shard_bytes == 0) return 0;
  const size_t n16 = shard_bytes / 16;
  B2_SIDE_KERNEL((peer_all_gather_kernel));
  peer_all_gather_kernel<<<clamp_blocks(n16, max_blocks), kCommThreads, 0, stream>>>(
      static_cast<V4*>(dst), to_ptrs(srcs, W), n16, rank, W, skip_self);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Same gather driven by the copy engines: one cudaMemcpyAsync per peer from the mapped peer pointer
// (DMA over NVLink, no SM time). The source shards must already be final (see b2_peer_barrier).
// The rank's own shard is placed by `self_blocks` copy blocks (see local_copy_kernel; 0 = cudaMemcpyAsync).
int b2_peer_all_gather_ce(void* dst, void* const* srcs, size_t shard_bytes, int rank, int W,
                          int skip_self, int self_blocks, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  for (int k = 0; k < W; ++k) {
    const int peer = (rank + k) % W;
    if (peer == rank && skip_self) continue;
    char* d = static_cast<char*>(dst) + (size_t)peer * shard_bytes;
    if (peer == rank && self_blocks > 0 && shard_bytes % 16 == 0 &&
        (reinterpret_cast<uintptr_t>(d) & 15) == 0 && (reinterpret_cast<uintptr_t>(srcs[peer]) & 15) == 0) {
      const size_t n16 = shard_bytes / 16;
      if (n16 > 0) {
        B2_SIDE_KERNEL((local_copy_kernel));
        local_copy_kernel<<<clamp_blocks(n16, self_blocks), kCommThreads, 0, stream>>>(
            reinterpret_cast<V4*>(d), static_cast<const V4*>(srcs[peer]), n16);
        B2_CUDA_CHECK(cudaGetLastError());
      }
      continue;
    }
    B2_CUDA_CHECK(cudaMemcpyAsync(d, srcs[peer], shard_bytes, cudaMemcpyDeviceToDevice, stream));
  }
  return 0;
}

// dst[0 .. bytes) = src on `blocks` copy blocks (0 or unaligned: cudaMemcpyAsync).
int b2_local_copy(void* dst, const void* src, size_t bytes, int blocks, cudaStream_t stream) {
  if (bytes == 0) return 0;
  if (blocks > 0 && bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0 &&
      (reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const size_t n16 = bytes / 16;
    B2_SIDE_KERNEL((local_copy_kernel));
    local_copy_kernel<<<clamp_blocks(n16, blocks), kCommThreads, 0, stream>>>(
        static_cast<V4*>(dst), static_cast<const V4*>(src), n16);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
  }
  B2_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, stream));
  return 0;
}

int b2_signal_peers(void* const* pads, int rank, int W, int channel, unsigned int epoch,
                    cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_SIDE_KERNEL((signal_peers_kernel));
  signal_peers_kernel<<<1, 32, 0, stream>>>(to_ptrs(pads, W), rank, W, channel, epoch);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

typedef CUresult (*CommWaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static CommWaitValue32Fn comm_wait_value() {
  static CommWaitValue32Fn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      return reinterpret_cast<CommWaitValue32Fn>(p);
    return static_cast<CommWaitValue32Fn>(nullptr);
  }();
  return fn;
}

int b2_stream_wait_pad(const void* pad_local, int rank, int W, int channel, unsigned int epoch,
                       cudaStream_t stream) {
  CommWaitValue32Fn wf = comm_wait_value();
  B2_REQUIRE(wf != nullptr, "cuStreamWaitValue32 entry point unavailable%s", "");
  const uint32_t* pad = static_cast<const uint32_t*>(pad_local) + channel * kMaxPeers;
  for (int q = 0; q < W; ++q) {
    if (q == rank) continue;
    CUresult r = wf((CUstream)stream, (CUdeviceptr)(pad + q), epoch, CU_STREAM_WAIT_VALUE_GEQ);
    B2_REQUIRE(r == CUDA_SUCCESS, "cuStreamWaitValue32 failed: %d", (int)r);
  }
  return 0;
}

int b2_ce_pull_slices(void* staging, void* const* fulls, size_t shard_bytes, int rank, int W,
                      cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  for (int k = 1; k < W; ++k) {
    const int peer = (rank + k) % W;
    B2_CUDA_CHECK(cudaMemcpyAsync(static_cast<char*>(staging) + (size_t)(k - 1) * shard_bytes,
                                  static_cast<const char*>(fulls[peer]) + (size_t)rank * shard_bytes,
                                  shard_bytes, cudaMemcpyDeviceToDevice, stream));
  }
  return 0;
}

static StagedPtrs staged_ptrs(const void* staging, size_t n16, int nstaged) {
  StagedPtrs sp;
  for (int k = 0; k < kMaxStaged; ++k)
    sp.p[k] = k < nstaged ? static_cast<const V4*>(staging) + (size_t)k * n16 : nullptr;
  return sp;
}

// out[shard_elems] = scale * (g_local + sum_k staging[k])   (bf16, fp32 accumulation)
int b2_reduce_slices(void* out, const void* g_local, const void* staging, int nstaged, size_t shard_elems,
                     float scale, int max_blocks, cudaStream_t stream) {
  B2_REQUIRE(nstaged >= 0 && nstaged <= kMaxStaged, "reduce_slices: %d staged slices", nstaged);
  B2_REQUIRE(shard_elems % 8 == 0, "reduce_slices: %zu elements is not a multiple of 8", shard_elems);
  const size_t n16 = shard_elems / 8;
  if (n16 == 0) return 0;
  AdamArgs none = {};
  return launch_reduce_local<false>(clamp_blocks(n16, max_blocks > 0 ? max_blocks : 296), stream,
                                    static_cast<V4*>(out), static_cast<const V4*>(g_local),
                                    staged_ptrs(staging, n16, nstaged), nstaged, n16, scale, none);
}

// AdamW on the owned shard with the gradient scale * (g_local + sum_k staging[k]) formed in registers.
int b2_reduce_adamw(void* p, void* m, void* v, const void* g_local, const void* staging, int nstaged,
                    size_t shard_elems, float grad_scale, float lr, float beta1, float beta2, float eps,
                    float wd, int step, int max_blocks, cudaStream_t stream) {
  B2_REQUIRE(nstaged >= 0 && nstaged <= kMaxStaged, "reduce_adamw: %d staged slices", nstaged);
  B2_REQUIRE(shard_elems % 8 == 0, "reduce_adamw: %zu elements is not a multiple of 8", shard_elems);
  const size_t n16 = shard_elems / 8;
  if (n16 == 0) return 0;
  AdamArgs a;
  a.p = static_cast<__nv_bfloat16*>(p);
  a.m = static_cast<__nv_bfloat16*>(m);
  a.v = static_cast<__nv_bfloat16*>(v);
  a.decay = 1.f - lr * wd;
  a.beta1 = beta1;
  a.beta2 = beta2;
  a.eps = eps;
  a.step_size = lr / (1.f - powf(beta1, (float)step));
  a.inv_bc2_sqrt = 1.f / sqrtf(1.f - powf(beta2, (float)step));
  return launch_reduce_local<true>(clamp_blocks(n16, max_blocks > 0 ? max_blocks : 296), stream, nullptr,
                                   static_cast<const V4*>(g_local), staged_ptrs(staging, n16, nstaged), nstaged,
                                   n16, grad_scale, a);
}

// lse[T], loss_row[T] from the W ranks' shard statistics (stats[r]: fp32 [3][T] in rank r's symmetric
// memory, complete on every rank -- issue b2_peer_barrier first).
int b2_ce_combine_peer(float* lse, float* loss_row, void* const* stats, const long long* labels, int T, int W,
                       int ignore_index, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  if (T == 0) return 0;
  int blocks = (T + 255) / 256;
  if (blocks > 296) blocks = 296;
  ce_combine_peer_kernel<<<blocks, 256, 0, stream>>>(lse, loss_row, to_ptrs(stats, W), labels, T, W,
                                                     ignore_index);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Copy-engine strided pulls from the peers: for every rank j, `rows` rows of `width_bytes` from
// srcs[j] + src_offset_bytes (pitch src_pitch) to dst + j * dst_step_bytes (pitch dst_pitch). Used for the
// hidden-sharded -> sequence-sharded exchange of the tensor-parallel embedding output and its transpose.
int b2_peer_pull_2d(void* dst, size_t dst_pitch, size_t dst_step_bytes, void* const* srcs, size_t src_offset_bytes,
                    size_t src_pitch, size_t width_bytes, size_t rows, int rank, int W, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  for (int k = 0; k < W; ++k) {
    const int j = (rank + k) % W;
    B2_CUDA_CHECK(cudaMemcpy2DAsync(static_cast<char*>(dst) + (size_t)j * dst_step_bytes, dst_pitch,
                                    static_cast<const char*>(srcs[j]) + src_offset_bytes, src_pitch, width_bytes,
                                    rows, cudaMemcpyDeviceToDevice, stream));
  }
  return 0;
}

// Spin bound of every cross-rank wait in this file (milliseconds; applies to the current device).
int b2_comm_set_timeout_ms(unsigned long long ms) {
  const unsigned long long ns = ms * 1000000ull;
  B2_CUDA_CHECK(cudaMemcpyToSymbol(g_spin_timeout_ns, &ns, sizeof(ns)));
  return 0;
}

// out_shard[shard_elems] (bf16) = scale * sum over ranks of fulls[r][rank * shard_elems + i].
// counter: one zero-initialised uint32 in local device memory, reused across calls on one stream.
int b2_peer_reduce_scatter(void* out_shard, void* const* fulls, size_t shard_elems, int rank, int W,
                           float scale, void* const* pads, int chan_base, unsigned int epoch,
                           unsigned int* counter, int max_blocks, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_REQUIRE(shard_elems % 8 == 0, "reduce_scatter: shard of %zu elements is not a multiple of 8", shard_elems);
  const size_t n16 = shard_elems / 8;
  AdamArgs none = {};
  if (W == 2 && rs_pipe_enabled()) {
    B2_SIDE_KERNEL((rs_pipe_kernel<false, false, 4, 2>));
    rs_pipe_kernel<false, false, 4, 2><<<clamp_blocks(n16, max_blocks), kCommThreads, 0, stream>>>(
        static_cast<V4*>(out_shard), static_cast<const V4*>(fulls[1 - rank]) + (size_t)rank * n16,
        static_cast<const V4*>(fulls[rank]) + (size_t)rank * n16, n16, rank, W, scale, to_ptrs(pads, W),
        chan_base, epoch, counter, none);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
  }
  return launch_rs<false>(W, clamp_blocks(n16, max_blocks), stream, static_cast<V4*>(out_shard),
                          to_ptrs(fulls, W), n16, rank, scale, to_ptrs(pads, W), chan_base, epoch,
                          counter, none);
}

// Reduce-scatter fused with the sharded AdamW update: p/m/v are this rank's shard (shard_elems).
int b2_peer_reduce_scatter_adamw(void* p, void* m, void* v, void* const* fulls, size_t shard_elems,
                                 int rank, int W, float grad_scale, float lr, float beta1,
                                 float beta2, float eps, float wd, int step, void* const* pads,
                                 int chan_base, unsigned int epoch, unsigned int* counter,
                                 int max_blocks, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_REQUIRE(shard_elems % 8 == 0, "reduce_scatter: shard of %zu elements is not a multiple of 8", shard_elems);
  const size_t n16 = shard_elems / 8;
  AdamArgs a;
  a.p = static_cast<__nv_bfloat16*>(p);
  a.m = static_cast<__nv_bfloat16*>(m);
  a.v = static_cast<__nv_bfloat16*>(v);
  a.decay = 1.f - lr * wd;
  a.beta1 = beta1;
  a.beta2 = beta2;
  a.eps = eps;
  a.step_size = lr / (1.f - powf(beta1, (float)step));
  a.inv_bc2_sqrt = 1.f / sqrtf(1.f - powf(beta2, (float)step));
  if (W == 2 && rs_pipe_enabled()) {
    B2_SIDE_KERNEL((rs_pipe_kernel<true, false, 2, 3>));
    rs_pipe_kernel<true, false, 2, 3><<<clamp_blocks(n16, max_blocks), kCommThreads, 0, stream>>>(
        nullptr, static_cast<const V4*>(fulls[1 - rank]) + (size_t)rank * n16,
        static_cast<const V4*>(fulls[rank]) + (size_t)rank * n16, n16, rank, W, grad_scale,
        to_ptrs(pads, W), chan_base, epoch, counter, a);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
  }
  return launch_rs<true>(W, clamp_blocks(n16, max_blocks), stream, nullptr, to_ptrs(fulls, W), n16,
                         rank, grad_scale, to_ptrs(pads, W), chan_base, epoch, counter, a);
}

// In-place all-reduce (scale * sum) of symmetric buffers bufs[r] of `elems` bf16 (elems % (8*W) == 0).
int b2_peer_all_reduce(void* const* bufs, size_t elems, int rank, int W, float scale,
                       void* const* pads, int chan_base, unsigned int epoch, unsigned int* counter,
                       int max_blocks, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_REQUIRE(elems % (8 * (size_t)W) == 0, "all_reduce: %zu elements not a multiple of 8*W", elems);
  const size_t n16_slice = elems / 8 / W;
  B2_SIDE_KERNEL((peer_all_reduce_kernel));
  peer_all_reduce_kernel<<<clamp_blocks(n16_slice, max_blocks), 256, 0, stream>>>(
      to_ptrs(bufs, W), n16_slice, rank, W, scale, to_ptrs(pads, W), chan_base, epoch, counter);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// NVLS variants: `mc_full` / `mc_buf` is the multicast address of the symmetric buffer (0 when the
// platform has no multicast support: callers then use the peer-load kernels above).
int b2_nvls_reduce_scatter(void* out_shard, const void* mc_full, size_t shard_elems, int rank, int W,
                           float scale, void* const* pads, int chan_base, unsigned int epoch,
                           unsigned int* counter, int max_blocks, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_REQUIRE(mc_full != nullptr, "nvls reduce_scatter: no multicast mapping%s", "");
  B2_REQUIRE(shard_elems % 8 == 0, "reduce_scatter: shard of %zu elements is not a multiple of 8", shard_elems);
  const size_t n16 = shard_elems / 8;
  AdamArgs none = {};
  if (rs_pipe_enabled()) {
    B2_SIDE_KERNEL((rs_pipe_kernel<false, true, 4, 3>));
    rs_pipe_kernel<false, true, 4, 3><<<clamp_blocks(n16, max_blocks), kCommThreads, 0, stream>>>(
        static_cast<V4*>(out_shard), static_cast<const V4*>(mc_full) + (size_t)rank * n16, nullptr, n16,
        rank, W, scale, to_ptrs(pads, W), chan_base, epoch, counter, none);
  } else {
    B2_SIDE_KERNEL((nvls_reduce_scatter_kernel<false>));
    nvls_reduce_scatter_kernel<false><<<clamp_blocks(n16, max_blocks), kCommThreads, 0, stream>>>(
        static_cast<V4*>(out_shard), static_cast<const V4*>(mc_full), n16, rank, W, scale,
        to_ptrs(pads, W), chan_base, epoch, counter, none);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_nvls_reduce_scatter_adamw(void* p, void* m, void* v, const void* mc_full, size_t shard_elems,
                                 int rank, int W, float grad_scale, float lr, float beta1,
                                 float beta2, float eps, float wd, int step, void* const* pads,
                                 int chan_base, unsigned int epoch, unsigned int* counter,
                                 int max_blocks, cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_REQUIRE(mc_full != nullptr, "nvls reduce_scatter: no multicast mapping%s", "");
  B2_REQUIRE(shard_elems % 8 == 0, "reduce_scatter: shard of %zu elements is not a multiple of 8", shard_elems);
  const size_t n16 = shard_elems / 8;
  AdamArgs a;
  a.p = static_cast<__nv_bfloat16*>(p);
  a.m = static_cast<__nv_bfloat16*>(m);
  a.v = static_cast<__nv_bfloat16*>(v);
  a.decay = 1.f - lr * wd;
  a.beta1 = beta1;
  a.beta2 = beta2;
  a.eps = eps;
  a.step_size = lr / (1.f - powf(beta1, (float)step));
  a.inv_bc2_sqrt = 1.f / sqrtf(1.f - powf(beta2, (float)step));
  if (rs_pipe_enabled()) {
    B2_SIDE_KERNEL((rs_pipe_kernel<true, true, 2, 4>));
    rs_pipe_kernel<true, true, 2, 4><<<clamp_blocks(n16, max_blocks), kCommThreads, 0, stream>>>(
        nullptr, static_cast<const V4*>(mc_full) + (size_t)rank * n16, nullptr, n16, rank, W, grad_scale,
        to_ptrs(pads, W), chan_base, epoch, counter, a);
  } else {
    B2_SIDE_KERNEL((nvls_reduce_scatter_kernel<true>));
    nvls_reduce_scatter_kernel<true><<<clamp_blocks(n16, max_blocks), kCommThreads, 0, stream>>>(
        nullptr, static_cast<const V4*>(mc_full), n16, rank, W, grad_scale, to_ptrs(pads, W), chan_base,
        epoch, counter, a);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int b2_nvls_all_reduce(void* mc_buf, size_t elems, int rank, int W, float scale, void* const* pads,
                       int chan_base, unsigned int epoch, unsigned int* counter, int max_blocks,
                       cudaStream_t stream) {
  B2_REQUIRE(W >= 1 && W <= kMaxPeers, "peer collectives support up to %d ranks (got %d)", kMaxPeers, W);
  B2_REQUIRE(mc_buf != nullptr, "nvls all_reduce: no multicast mapping%s", "");
  B2_REQUIRE(elems % (8 * (size_t)W) == 0, "all_reduce: %zu elements not a multiple of 8*W", elems);
  const size_t n16_slice = elems / 8 / W;
  B2_SIDE_KERNEL((nvls_all_reduce_kernel));
  nvls_all_reduce_kernel<<<clamp_blocks(n16_slice, max_blocks), kCommThreads, 0, stream>>>(
      static_cast<V4*>(mc_buf), n16_slice, rank, W, scale, to_ptrs(pads, W), chan_base, epoch, counter);
  B2_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
