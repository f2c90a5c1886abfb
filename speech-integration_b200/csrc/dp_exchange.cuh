// Data-parallel exchange of the tied LM-head / embedding gradient over NVLink peer memory (one process per GPU,
// one NVSwitch node).  Every rank exposes one exchange buffer to its peers (CUDA IPC):
//
//   [ctrl 4 KiB: arrive_a[8] | arrive_b[8] | local block counter]  [dW_acc fp32 V x D]  [stage bf16 Vs x D]
//
// dW_acc is where the rank's own dW contraction (and its embedding scatter-add) accumulate in fp32; Vs =
// ceil(V / world) is the height of the vocabulary slice a rank owns.  The exchange is two small-footprint kernels
// (no shared memory, 256 threads per block) that co-reside with the persistent tensor-core CTAs of the dX
// contraction still running on the main stream, so the transfer hides behind math:
//
//   dp_reduce_kernel   rank r pulls slice r of every peer's fp32 dW_acc through NVLink, sums the `world` terms in
//                      rank order in fp32 (bit-identical on every rank and run to run), applies the single-pass
//                      ratio, adds the existing gradient when accumulating, rounds ONCE to bf16 and writes the
//                      result to its own dW slice and to its stage buffer.
//   dp_gather_kernel   every rank pulls the other ranks' finished bf16 slices from their stage buffers.
//
// Cross-GPU ordering uses epoch flags in the ctrl block: a rank announces "my dW_acc is complete" (arrive_a) /
// "my slice is reduced" (arrive_b) by storing the step's epoch into ITS slot of every peer's ctrl block with
// system-scope release semantics; consumers spin on their own ctrl block.  No block ever waits for another block
// of its own grid, so no co-residency of the whole grid is required.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "aux_kernels.cuh"

namespace ssi {

constexpr int DP_MAX_WORLD = 8;
constexpr size_t DP_CTRL_BYTES = 4096;
constexpr int DP_THREADS = 256;

struct DpCtrl {
  unsigned int arrive_a[DP_MAX_WORLD];  // written by peers: epoch of their last "dW_acc complete"
  unsigned int arrive_b[DP_MAX_WORLD];  // written by peers: epoch of their last "slice reduced"
  unsigned int blocks_done;             // local: blocks of dp_reduce_kernel that finished (reset by the last one)
  unsigned int timed_out;               // local: set when a wait gave up (a peer never arrived): results are invalid
};
constexpr unsigned long long DP_WAIT_TIMEOUT_NS = 20ull * 1000 * 1000 * 1000;  // a dead peer must not hang the GPU

struct DpPeers {
  char* buf[DP_MAX_WORLD];  // exchange buffer of every rank as mapped in THIS process (own buffer included)
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Peer data is read exactly once per launch and only after the acquire on the peer's flag, and L1 is invalidated
// at every launch boundary, so plain (weak) vector loads are safe; L1::no_allocate keeps the stream out of L1.
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ld_peer_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// wait until every peer's slot in `flags` (this rank's own ctrl block) has reached `epoch`; gives up after
// DP_WAIT_TIMEOUT_NS and records it in *timed_out (checked by the host with ssi_b200_dp_status)
__device__ __forceinline__ void dp_wait_all(const unsigned int* flags, int world, int rank, unsigned int epoch,
                                            unsigned int* timed_out) {
  if (threadIdx.x < static_cast<unsigned>(world) && static_cast<int>(threadIdx.x) != rank) {
    // epochs only grow; the signed difference keeps the comparison valid across 32-bit wrap-around
    const unsigned long long t0 = global_timer_ns();
    unsigned int spins = 0;
    while (static_cast<int>(ld_acquire_sys(flags + threadIdx.x) - epoch) < 0) {
      __nanosleep(64);
      if ((++spins & 1023u) == 0 && global_timer_ns() - t0 > DP_WAIT_TIMEOUT_NS) {
        *timed_out = 1u;
        break;
      }
    }
  }
  __syncthreads();
}

template <int WORLD>
__global__ void __launch_bounds__(DP_THREADS) dp_reduce_kernel(DpPeers peers, int rank, long long slice_f4,
                                                               long long total_f4, unsigned int epoch,
                                                               const float* __restrict__ grad_out,
                                                               const float* __restrict__ grad_hint, int accumulate,
                                                               uint2* __restrict__ dW) {
  DpCtrl* ctrl = reinterpret_cast<DpCtrl*>(peers.buf[rank]);
  // ---- announce: everything this rank contributes (its dW_acc) was finished by earlier kernels of this stream
  if (blockIdx.x == 0 && threadIdx.x < WORLD && static_cast<int>(threadIdx.x) != rank) {
    __threadfence_system();
    st_release_sys(&reinterpret_cast<DpCtrl*>(peers.buf[threadIdx.x])->arrive_a[rank], epoch);
  }
  dp_wait_all(ctrl->arrive_a, WORLD, rank, epoch, &ctrl->timed_out);

  const float ratio = grad_ratio(grad_out, grad_hint);
  const long long begin = static_cast<long long>(rank) * slice_f4;
  const long long end = min(begin + slice_f4, total_f4);
  const float4* src[WORLD];
#pragma unroll
  for (int p = 0; p < WORLD; ++p) src[p] = reinterpret_cast<const float4*>(peers.buf[p] + DP_CTRL_BYTES);
  uint2* stage = reinterpret_cast<uint2*>(peers.buf[rank] + DP_CTRL_BYTES + static_cast<size_t>(total_f4) * 16);
  constexpr int UNROLL = (WORLD <= 2) ? 4 : 2;
  const long long stride = static_cast<long long>(gridDim.x) * DP_THREADS;
  for (long long i0 = begin + blockIdx.x * static_cast<long long>(DP_THREADS) + threadIdx.x; i0 < end;
       i0 += stride * UNROLL) {
    float4 v[UNROLL][WORLD];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = i0 + u * stride;
      if (i < end) {
#pragma unroll
        for (int p = 0; p < WORLD; ++p) v[u][p] = (p == rank) ? src[p][i] : ld_peer_f4(src[p] + i);
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = i0 + u * stride;
      if (i >= end) continue;
      float4 a = v[u][0];
#pragma unroll
      for (int p = 1; p < WORLD; ++p) a.x += v[u][p].x, a.y += v[u][p].y, a.z += v[u][p].z, a.w += v[u][p].w;
      a.x *= ratio, a.y *= ratio, a.z *= ratio, a.w *= ratio;
      if (accumulate) {
        const uint2 old = dW[i];
        a.x += bf16lo(old.x), a.y += bf16hi(old.x), a.z += bf16lo(old.y), a.w += bf16hi(old.y);
      }
      uint2 o;
      o.x = pack_bf16x2(a.x, a.y);
      o.y = pack_bf16x2(a.z, a.w);
      dW[i] = o;
      stage[i - begin] = o;
    }
  }
  // ---- the last block to finish announces "slice reduced" to every peer
  __threadfence_system();
  __syncthreads();
  __shared__ unsigned int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(&ctrl->blocks_done, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (s_last) {
    if (threadIdx.x == 0) ctrl->blocks_done = 0;
    if (threadIdx.x < WORLD && static_cast<int>(threadIdx.x) != rank) {
      __threadfence_system();
      st_release_sys(&reinterpret_cast<DpCtrl*>(peers.buf[threadIdx.x])->arrive_b[rank], epoch);
    }
  }
}

template <int WORLD>
__global__ void __launch_bounds__(DP_THREADS) dp_gather_kernel(DpPeers peers, int rank, long long slice_f4,
                                                               long long total_f4, unsigned int epoch,
                                                               uint2* __restrict__ dW) {
  DpCtrl* ctrl = reinterpret_cast<DpCtrl*>(peers.buf[rank]);
  dp_wait_all(ctrl->arrive_b, WORLD, rank, epoch, &ctrl->timed_out);
  // bf16 slices move as 16-byte words: 2 of the fp32-quad units per word (slice_f4 is even, see the host side)
  const long long slice_u4 = slice_f4 / 2, total_u4 = total_f4 / 2;
  const long long stride = static_cast<long long>(gridDim.x) * DP_THREADS;
  uint4* out = reinterpret_cast<uint4*>(dW);
  for (int q = 1; q < WORLD; ++q) {
    const int p = (rank + q) % WORLD;  // start with a different peer on every rank
    const uint4* stage =
        reinterpret_cast<const uint4*>(peers.buf[p] + DP_CTRL_BYTES + static_cast<size_t>(total_f4) * 16);
    const long long begin = static_cast<long long>(p) * slice_u4;
    const long long n = min(slice_u4, total_u4 - begin);
    for (long long i0 = blockIdx.x * static_cast<long long>(DP_THREADS) + threadIdx.x; i0 < n; i0 += stride * 4) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < n) v[u] = ld_peer_u4(stage + i0 + u * stride);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < n) out[begin + i0 + u * stride] = v[u];
    }
  }
}

}  // namespace ssi
