// Data-parallel exchange of the tied LM-head / embedding gradient over NVLink peer memory (one process per GPU,
// one NVSwitch node).  Every rank exposes one exchange buffer to its peers (CUDA IPC):
//
//   [ctrl 4 KiB: flags, token counts]  [dW_acc fp32 V x D]  [stage bf16 Vs x D]
//   [slot u64 V]  [emb_ids int32 cap]  [emb_rows bf16 cap x D]  [own_flag u32 Vs]  [own_rows bf16 Vs x D]
//
// dW_acc is where the rank's own dW contractions accumulate in fp32; Vs = ceil(V / world) is the height of the
// vocabulary slice a rank owns.  The step has THREE exchanges, in the order a training step produces their inputs:
//
//   dp_counts_kernel   (start of the step) every rank stores its two token counts into every peer's ctrl block and
//                      sums the `world` pairs: the global counts without a collective call on the critical path.
//   dense exchange     (LM-head part, as soon as the last dW contraction is done -- behind the last dX contraction
//                      and, in a full model, behind the decoder backward): dp_reduce_kernel + dp_gather_kernel below.
//   sparse exchange    (embedding part; it only exists after the decoder backward has produced dH0): every rank
//                      segment-reduces its dH0 rows by id into a compact list (scatter.cuh, SCAT_COMPACT: emb_rows,
//                      emb_ids, slot[id] = epoch | index) and dp_sparse_add_kernel then adds, for every id present
//                      on any rank, the ranks' rows in rank order (fp32) to the bf16 gradient row -- the other ranks'
//                      distinct rows per rank instead of a second dense V x D pass.  For world > 4 in two phases
//                      (dp_sparse_reduce_kernel + dp_sparse_apply_kernel at the end of this file): the owner of an id
//                      (id % world) sums its rows once, every rank then adds the reduced row.
//
// The dense exchange is two small-footprint kernels (no shared memory, 256 threads per block) that co-reside with the
// persistent tensor-core CTAs of the dX contraction still running on the main stream, so the transfer hides behind
// math:
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
  unsigned int arrive_c[DP_MAX_WORLD];  // written by peers: epoch of their last "compact embedding rows published"
  unsigned int arrive_n[DP_MAX_WORLD];  // written by peers: epoch of their last "token counts posted"
  unsigned int arrive_d[DP_MAX_WORLD];  // written by peers: epoch of their last "owned embedding rows reduced"
  unsigned long long counts[2][DP_MAX_WORLD][2];  // [epoch & 1][rank] = {unshifted, shifted} valid-token counts
  unsigned int blocks_done;             // local: blocks of dp_reduce_kernel that finished (reset by the last one)
  unsigned int timed_out;               // set when a wait gave up, here OR on a peer (which then stores 1 into every
                                        // rank's word): this step's results are invalid, see dp_wait_all
  unsigned long long timeout_ns;        // local: how long a wait may last (set by dp_alloc / dp_reset; 0 = the default)
};
static_assert(sizeof(DpCtrl) <= DP_CTRL_BYTES, "ctrl block too large");
// A dead peer must not hang the GPU, but ordinary rank skew (a checkpoint written by rank 0, an uneven eval shard,
// a dataloader stall, first-step compilation) must not trip the wait either: the default matches the 10 minutes a
// torch.distributed process group waits by default.  Option "dp_timeout_s" / SSI_B200_DP_TIMEOUT_S, read at dp_alloc.
constexpr unsigned long long DP_WAIT_TIMEOUT_DEFAULT_NS = 600ull * 1000 * 1000 * 1000;
constexpr uint32_t BF16X2_NAN = 0x7FC07FC0u;

struct DpPeers {
  char* buf[DP_MAX_WORLD];  // exchange buffer of every rank as mapped in THIS process (own buffer included)
};

// byte offsets inside an exchange buffer (identical on every rank)
struct DpLayout {
  size_t off_acc, off_stage, off_slot, off_ids, off_rows, off_own_flag, off_own_rows, bytes;
  long long cap;  // rows of the compact embedding list
  long long vs;   // ceil(V / world): ids a rank owns in the two-phase sparse exchange (id % world == rank)
};
__host__ __device__ inline DpLayout dp_layout(long long V, long long D, int world, long long cap) {
  DpLayout l;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  const long long vs = (V + world - 1) / world;
  size_t o = DP_CTRL_BYTES;
  l.off_acc = o, o += up(static_cast<size_t>(V) * D * 4);
  l.off_stage = o, o += up(static_cast<size_t>(vs) * D * 2);
  l.off_slot = o, o += up(static_cast<size_t>(V) * 8);
  l.off_ids = o, o += up(static_cast<size_t>(cap) * 4);
  l.off_rows = o, o += up(static_cast<size_t>(cap) * D * 2);
  // two-phase sparse exchange: the reduced row of owned id v = k * world + rank lives at own_rows[k], valid for the
  // exchange whose epoch is in own_flag[k]
  l.off_own_flag = o, o += up(static_cast<size_t>(vs) * 4);
  l.off_own_rows = o, o += up(static_cast<size_t>(cap > 0 ? vs : 0) * D * 2);
  l.bytes = o, l.cap = cap, l.vs = vs;
  return l;
}

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

// NVLS: one load from the MULTICAST mapping of the exchange buffers returns the sum over all ranks, formed inside
// the NVSwitch (fp32 adds in the switch's fixed order: the same value whichever rank asks, run to run on one topology)
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float4* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// timed_out stays set -- every exchange writes NaN -- until all ranks clear it together (ssi_b200_dp_reset between two
// host barriers, TiedGradExchange.reset())
__device__ __forceinline__ bool dp_is_bad(const DpCtrl* ctrl) {
  return *reinterpret_cast<const volatile unsigned int*>(&ctrl->timed_out) != 0u;
}

__device__ __forceinline__ unsigned long long dp_timeout_ns(const DpCtrl* ctrl) {
  const unsigned long long t = *reinterpret_cast<const volatile unsigned long long*>(&ctrl->timeout_ns);
  return t ? t : DP_WAIT_TIMEOUT_DEFAULT_NS;
}

// A wait gave up: mark EVERY rank's ctrl block (own block included), not only this one.  The peer that was late
// finds all flags of the step already posted, so its own waits succeed at once -- without the poison it would read
// buffers this rank may already be rewriting for its next step and carry on with wrong, finite gradients.
__device__ __forceinline__ void dp_poison_all(const DpPeers& peers, int world, unsigned int* host_flag) {
  for (int p = 0; p < world; ++p)
    *reinterpret_cast<volatile unsigned int*>(&reinterpret_cast<DpCtrl*>(peers.buf[p])->timed_out) = 1u;
  if (host_flag) *reinterpret_cast<volatile unsigned int*>(host_flag) = 1u;  // pinned host word: no sync to read
  __threadfence_system();
}

// wait until every peer's slot in `flags` (this rank's own ctrl block) has reached `epoch`; gives up after the ctrl
// block's timeout and then poisons all ranks (dp_poison_all) and raises the pinned host word `host_flag` (the host
// polls that word without synchronising; the kernels write NaN instead of sums so that no optimizer step can use them)
__device__ __forceinline__ void dp_wait_all(const DpPeers& peers, const unsigned int* flags, int world, int rank,
                                            unsigned int epoch, unsigned int* host_flag) {
  if (threadIdx.x < static_cast<unsigned>(world) && static_cast<int>(threadIdx.x) != rank) {
    // epochs only grow; the signed difference keeps the comparison valid across 32-bit wrap-around
    const unsigned long long t0 = global_timer_ns();
    const unsigned long long limit = dp_timeout_ns(reinterpret_cast<const DpCtrl*>(peers.buf[rank]));
    unsigned int spins = 0;
    while (static_cast<int>(ld_acquire_sys(flags + threadIdx.x) - epoch) < 0) {
      __nanosleep(64);
      if ((++spins & 1023u) == 0) {
        if (dp_is_bad(reinterpret_cast<const DpCtrl*>(peers.buf[rank]))) break;  // the step is already lost: do not wait again
        if (global_timer_ns() - t0 > limit) {
          dp_poison_all(peers, world, host_flag);
          break;
        }
      }
    }
  }
  __syncthreads();
  // a PEER's wait gave up (it marked this rank's block): tell this rank's host too
  if (threadIdx.x == 0 && host_flag &&
      *reinterpret_cast<const volatile unsigned int*>(&reinterpret_cast<const DpCtrl*>(peers.buf[rank])->timed_out))
    *reinterpret_cast<volatile unsigned int*>(host_flag) = 1u;
}


// ---------------------------------------------------------------------------------------------------
// Token counts: one warp.  Lane p stores this rank's pair into rank p's ctrl block (own block included) and
// releases arrive_n; then lane p waits for rank p's pair in the own block; out[0..1] = the sums over ranks,
// hint_out (optional) = float(out[0]), the multiplier of ssi/trainer.py:393 on the global batch.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) dp_counts_kernel(DpPeers peers, int rank, int world, unsigned int epoch,
                                                       const int64_t* __restrict__ local, int64_t* __restrict__ out,
                                                       float* __restrict__ hint_out, unsigned int* host_flag) {
  DpCtrl* ctrl = reinterpret_cast<DpCtrl*>(peers.buf[rank]);
  const int p = threadIdx.x;
  const int par = epoch & 1u;
  if (p < world) {
    DpCtrl* pc = reinterpret_cast<DpCtrl*>(peers.buf[p]);
    volatile unsigned long long* dst = pc->counts[par][rank];
    dst[0] = static_cast<unsigned long long>(local[0]);
    dst[1] = static_cast<unsigned long long>(local[1]);
    __threadfence_system();
    st_release_sys(&pc->arrive_n[rank], epoch);
  }
  unsigned long long c0 = 0, c1 = 0;
  if (p < world) {
    const unsigned long long t0 = global_timer_ns();
    const unsigned long long limit = dp_timeout_ns(ctrl);
    unsigned int spins = 0;
    while (static_cast<int>(ld_acquire_sys(&ctrl->arrive_n[p]) - epoch) < 0) {
      __nanosleep(64);
      if ((++spins & 1023u) == 0) {
        if (dp_is_bad(reinterpret_cast<const DpCtrl*>(peers.buf[rank]))) break;  // the step is already lost: do not wait again
        if (global_timer_ns() - t0 > limit) {
          dp_poison_all(peers, world, host_flag);
          break;
        }
      }
    }
    const volatile unsigned long long* src = ctrl->counts[par][p];
    c0 = src[0], c1 = src[1];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    c0 += __shfl_xor_sync(0xffffffffu, c0, o);
    c1 += __shfl_xor_sync(0xffffffffu, c1, o);
  }
  if (p == 0) {
    out[0] = static_cast<int64_t>(c0);
    out[1] = static_cast<int64_t>(c1);
    if (hint_out) hint_out[0] = static_cast<float>(static_cast<long long>(c0));
  }
}

// ---------------------------------------------------------------------------------------------------
// Dense exchange of the fp32 LM-head sums
// ---------------------------------------------------------------------------------------------------
// NVLS = true: `mc_acc` is the multicast address of dW_acc (all ranks' buffers bound to one multicast object); the
// `world` peer loads per element become ONE multimem.ld_reduce, i.e. 1/world of the NVLink bytes into this GPU.
template <int WORLD, bool NVLS = false>
__global__ void __launch_bounds__(DP_THREADS) dp_reduce_kernel(DpPeers peers, DpLayout lay, int rank,
                                                               long long slice_f4, long long total_f4,
                                                               unsigned int epoch, const float* __restrict__ grad_out,
                                                               const float* __restrict__ grad_hint, int accumulate,
                                                               uint2* __restrict__ dW, unsigned int* host_flag,
                                                               GradFold fold, const float4* mc_acc = nullptr) {
  __shared__ float s_fold[DP_THREADS / 32];
  DpCtrl* ctrl = reinterpret_cast<DpCtrl*>(peers.buf[rank]);
  // ---- announce: everything this rank contributes (its dW_acc) was finished by earlier kernels of this stream
  if (blockIdx.x == 0 && threadIdx.x < WORLD && static_cast<int>(threadIdx.x) != rank) {
    __threadfence_system();
    st_release_sys(&reinterpret_cast<DpCtrl*>(peers.buf[threadIdx.x])->arrive_a[rank], epoch);
  }
  dp_wait_all(peers, ctrl->arrive_a, WORLD, rank, epoch, host_flag);
  const bool bad = dp_is_bad(ctrl);  // a peer never arrived: NaN out, so that no optimizer step can use the result

  const float ratio = grad_ratio(grad_out, grad_hint);
  const float ps = fold_scale(fold);
  float sq = 0.f;
  const long long begin = static_cast<long long>(rank) * slice_f4;
  const long long end = min(begin + slice_f4, total_f4);
  const float4* src[WORLD];
#pragma unroll
  for (int p = 0; p < WORLD; ++p) src[p] = reinterpret_cast<const float4*>(peers.buf[p] + lay.off_acc);
  uint2* stage = reinterpret_cast<uint2*>(peers.buf[rank] + lay.off_stage);
  constexpr int UNROLL = NVLS ? 8 : (WORLD <= 2) ? 4 : 2;
  constexpr int TERMS = NVLS ? 1 : WORLD;
  const long long stride = static_cast<long long>(gridDim.x) * DP_THREADS;
  for (long long i0 = begin + blockIdx.x * static_cast<long long>(DP_THREADS) + threadIdx.x; i0 < end;
       i0 += stride * UNROLL) {
    float4 v[UNROLL][TERMS];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = i0 + u * stride;
      if (i < end) {
        if constexpr (NVLS) {
          v[u][0] = multimem_ld_reduce_f4(mc_acc + i);
        } else {
#pragma unroll
          for (int p = 0; p < WORLD; ++p) v[u][p] = (p == rank) ? src[p][i] : ld_peer_f4(src[p] + i);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = i0 + u * stride;
      if (i >= end) continue;
      float4 a = v[u][0];
#pragma unroll
      for (int p = 1; p < TERMS; ++p) a.x += v[u][p].x, a.y += v[u][p].y, a.z += v[u][p].z, a.w += v[u][p].w;
      a.x *= ratio, a.y *= ratio, a.z *= ratio, a.w *= ratio;
      if (accumulate) {
        const uint2 old = dW[i];
        a.x += bf16lo(old.x), a.y += bf16hi(old.x), a.z += bf16lo(old.y), a.w += bf16hi(old.y);
      }
      if (fold.post_scale) a.x *= ps, a.y *= ps, a.z *= ps, a.w *= ps;
      uint2 o;
      o.x = pack_bf16x2(a.x, a.y);
      o.y = pack_bf16x2(a.z, a.w);
      if (bad) o.x = o.y = BF16X2_NAN;
      dW[i] = o;
      stage[i - begin] = o;
      if (fold.sq) sq += sq_of_bf16x2(o.x) + sq_of_bf16x2(o.y);
    }
  }
  if (fold.sq) fold_block_partial(sq, fold.sq, s_fold);  // this rank's own slice
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
__global__ void __launch_bounds__(DP_THREADS) dp_gather_kernel(DpPeers peers, DpLayout lay, int rank,
                                                               long long slice_f4, long long total_f4,
                                                               unsigned int epoch, uint2* __restrict__ dW,
                                                               unsigned int* host_flag, GradFold fold) {
  __shared__ float s_fold[DP_THREADS / 32];
  DpCtrl* ctrl = reinterpret_cast<DpCtrl*>(peers.buf[rank]);
  dp_wait_all(peers, ctrl->arrive_b, WORLD, rank, epoch, host_flag);
  const bool bad = dp_is_bad(ctrl);
  float sq = 0.f;
  // bf16 slices move as 16-byte words: 2 of the fp32-quad units per word (slice_f4 is even, see the host side)
  const long long slice_u4 = slice_f4 / 2, total_u4 = total_f4 / 2;
  const long long stride = static_cast<long long>(gridDim.x) * DP_THREADS;
  uint4* out = reinterpret_cast<uint4*>(dW);
  for (int q = 1; q < WORLD; ++q) {
    const int p = (rank + q) % WORLD;  // start with a different peer on every rank
    const uint4* stage = reinterpret_cast<const uint4*>(peers.buf[p] + lay.off_stage);
    const long long begin = static_cast<long long>(p) * slice_u4;
    const long long n = min(slice_u4, total_u4 - begin);
    for (long long i0 = blockIdx.x * static_cast<long long>(DP_THREADS) + threadIdx.x; i0 < n; i0 += stride * 4) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < n) v[u] = ld_peer_u4(stage + i0 + u * stride);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < n) {
          if (bad) v[u] = make_uint4(BF16X2_NAN, BF16X2_NAN, BF16X2_NAN, BF16X2_NAN);
          out[begin + i0 + u * stride] = v[u];
          if (fold.sq)
            sq += (sq_of_bf16x2(v[u].x) + sq_of_bf16x2(v[u].y)) + (sq_of_bf16x2(v[u].z) + sq_of_bf16x2(v[u].w));
        }
    }
  }
  if (fold.sq) fold_block_partial(sq, fold.sq, s_fold);  // the other ranks' slices
}

// ---------------------------------------------------------------------------------------------------
// Sparse exchange of the embedding part.  Every rank has published (earlier kernels of this stream: scatter.cuh in
// SCAT_COMPACT form) its compact list: emb_rows[u, :] = bf16 sum of its dH0 rows with the u-th distinct id and
// slot[id] = (epoch << 32 | u).  Here every rank adds, for every vocabulary id present on ANY rank, the ranks' rows in
// rank order in fp32 to its bf16 gradient row: one read-modify-write (one rounding) per touched row, the same
// arithmetic in the same order on every rank, so the gradients stay bit-identical across ranks.
// A warp looks at 32 / WP2 ids x WP2 ranks per iteration (one 8-byte peer load per lane).
// ---------------------------------------------------------------------------------------------------
constexpr int SPARSE_G = 4;                      // 16-byte column groups per lane and pass
constexpr int SPARSE_COLS = 32 * SPARSE_G * 8;   // 1024 columns per pass

template <int WORLD>
__global__ void __launch_bounds__(DP_THREADS) dp_sparse_add_kernel(DpPeers peers, DpLayout lay, int rank, int V, int D,
                                                                   unsigned int epoch, __nv_bfloat16* __restrict__ dW,
                                                                   unsigned int* host_flag, GradFold fold) {
  constexpr int WP2 = WORLD <= 1 ? 1 : WORLD <= 2 ? 2 : WORLD <= 4 ? 4 : 8;
  constexpr int PER = 32 / WP2;  // ids per warp iteration
  DpCtrl* ctrl = reinterpret_cast<DpCtrl*>(peers.buf[rank]);
  if (blockIdx.x == 0 && threadIdx.x < WORLD && static_cast<int>(threadIdx.x) != rank) {
    __threadfence_system();
    st_release_sys(&reinterpret_cast<DpCtrl*>(peers.buf[threadIdx.x])->arrive_c[rank], epoch);
  }
  dp_wait_all(peers, ctrl->arrive_c, WORLD, rank, epoch, host_flag);
  if (dp_is_bad(ctrl)) {  // poison the first gradient row: grad-norm and loss checks trip at once
    if (blockIdx.x == 0)
      for (int j = threadIdx.x; j < D / 2; j += DP_THREADS) reinterpret_cast<uint32_t*>(dW)[j] = BF16X2_NAN;
    return;
  }
  const uint32_t lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * DP_THREADS + threadIdx.x) >> 5;
  const int nw = (gridDim.x * DP_THREADS) >> 5;
  const int sub = lane / WP2, pl = lane % WP2;
  const float ps = fold_scale(fold);
  float dsq = 0.f;
  const unsigned long long* slot_p =
      (pl < WORLD) ? reinterpret_cast<const unsigned long long*>(peers.buf[pl] + lay.off_slot) : nullptr;
  for (long long v0 = static_cast<long long>(gw) * PER; v0 < V; v0 += static_cast<long long>(nw) * PER) {
    const long long v = v0 + sub;
    unsigned long long s = 0;
    if (slot_p != nullptr && v < V) {
      asm volatile("ld.global.L1::no_allocate.u64 %0, [%1];" : "=l"(s) : "l"(slot_p + v));
    }
    const bool present = (static_cast<unsigned int>(s >> 32) == epoch) && slot_p != nullptr && v < V &&
                         static_cast<long long>(s & 0xFFFFFFFFull) < lay.cap;
    const unsigned pm = __ballot_sync(0xffffffffu, present);
    if (pm == 0) continue;
    const unsigned int my_idx = static_cast<unsigned int>(s & 0xFFFFFFFFull);
    for (int q = 0; q < PER; ++q) {
      const unsigned m = (pm >> (q * WP2)) & ((1u << WP2) - 1u);
      if (m == 0) continue;
      __nv_bfloat16* drow = dW + static_cast<size_t>(v0 + q) * D;
      for (int d0 = 0; d0 < D; d0 += SPARSE_COLS) {
        const int dcols = min(SPARSE_COLS, D - d0);
        float acc[SPARSE_G][8];
#pragma unroll
        for (int g = 0; g < SPARSE_G; ++g)
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[g][e] = 0.f;
#pragma unroll
        for (int pb = 0; pb < WORLD; pb += 4) {  // 4 peers' loads in flight, summed in rank order
          uint4 u[4][SPARSE_G];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int p = pb + k;
            if (p >= WORLD) break;
            const unsigned int idx = __shfl_sync(0xffffffffu, my_idx, q * WP2 + p);
            const bool has = (m >> p) & 1u;
            const __nv_bfloat16* src =
                reinterpret_cast<const __nv_bfloat16*>(peers.buf[p] + lay.off_rows) + static_cast<size_t>(idx) * D + d0;
#pragma unroll
            for (int g = 0; g < SPARSE_G; ++g) {
              const int col = g * 256 + static_cast<int>(lane) * 8;
              u[k][g] = (has && col < dcols) ? ld_peer_u4(reinterpret_cast<const uint4*>(src + col))
                                             : make_uint4(0, 0, 0, 0);
            }
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (pb + k >= WORLD) break;
#pragma unroll
            for (int g = 0; g < SPARSE_G; ++g) {
              acc[g][0] += bf16lo(u[k][g].x), acc[g][1] += bf16hi(u[k][g].x);
              acc[g][2] += bf16lo(u[k][g].y), acc[g][3] += bf16hi(u[k][g].y);
              acc[g][4] += bf16lo(u[k][g].z), acc[g][5] += bf16hi(u[k][g].z);
              acc[g][6] += bf16lo(u[k][g].w), acc[g][7] += bf16hi(u[k][g].w);
            }
          }
        }
#pragma unroll
        for (int g = 0; g < SPARSE_G; ++g) {
          const int col = g * 256 + static_cast<int>(lane) * 8;
          if (col < dcols) {
            uint4* ptr = reinterpret_cast<uint4*>(drow + d0 + col);
            const uint4 old = *ptr;
            uint4 o;
            // ps = 1 without a fold: the plain sum
            o.x = pack_bf16x2(fmaf(ps, acc[g][0], bf16lo(old.x)), fmaf(ps, acc[g][1], bf16hi(old.x)));
            o.y = pack_bf16x2(fmaf(ps, acc[g][2], bf16lo(old.y)), fmaf(ps, acc[g][3], bf16hi(old.y)));
            o.z = pack_bf16x2(fmaf(ps, acc[g][4], bf16lo(old.z)), fmaf(ps, acc[g][5], bf16hi(old.z)));
            o.w = pack_bf16x2(fmaf(ps, acc[g][6], bf16lo(old.w)), fmaf(ps, acc[g][7], bf16hi(old.w)));
            *ptr = o;
            if (fold.sq)
              dsq += (sq_of_bf16x2(o.x) - sq_of_bf16x2(old.x)) + (sq_of_bf16x2(o.y) - sq_of_bf16x2(old.y)) +
                     (sq_of_bf16x2(o.z) - sq_of_bf16x2(old.z)) + (sq_of_bf16x2(o.w) - sq_of_bf16x2(old.w));
          }
        }
      }
    }
  }
  if (fold.sq) {  // one partial per warp: sum(new^2 - old^2) over the rows it rewrote
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dsq += __shfl_xor_sync(0xffffffffu, dsq, o);
    if (lane == 0) fold.sq[gw] = dsq;
  }
}


// ---------------------------------------------------------------------------------------------------
// Two-phase form of the sparse exchange (default for world > 4).  The one-phase kernel above makes every rank pull
// every other rank's rows: sum_p U_p rows per rank (c5 at 8 ranks: 56 k rows = 229 MB, of which two thirds are the
// ~5000 DSU ids that EVERY rank holds).  Here id v is owned by rank v % world:
//   dp_sparse_reduce_kernel  the owner sums the ranks' rows of its ids in rank order in fp32, rounds once to bf16 and
//                            leaves the row in its own_rows[v / world] (own_flag = epoch);
//   dp_sparse_apply_kernel   every rank reads each present id's reduced row from its owner (one row per id instead
//                            of one per holder) and adds it to its bf16 gradient row.
// c5 at 8 ranks: 7 k + 26 k rows = 135 MB per rank.  The same bf16 row is added by every rank, so the gradients stay
// bit-identical across ranks; one more rounding of the embedding sum than the one-phase form.
// ---------------------------------------------------------------------------------------------------
template <int WORLD>
__global__ void __launch_bounds__(DP_THREADS) dp_sparse_reduce_kernel(DpPeers peers, DpLayout lay, int rank, int V,
                                                                      int D, unsigned int epoch,
                                                                      unsigned int* host_flag) {
  constexpr int WP2 = WORLD <= 1 ? 1 : WORLD <= 2 ? 2 : WORLD <= 4 ? 4 : 8;
  constexpr int PER = 32 / WP2;  // owned ids per warp iteration
  DpCtrl* ctrl = reinterpret_cast<DpCtrl*>(peers.buf[rank]);
  if (blockIdx.x == 0 && threadIdx.x < WORLD && static_cast<int>(threadIdx.x) != rank) {
    __threadfence_system();
    st_release_sys(&reinterpret_cast<DpCtrl*>(peers.buf[threadIdx.x])->arrive_c[rank], epoch);
  }
  dp_wait_all(peers, ctrl->arrive_c, WORLD, rank, epoch, host_flag);
  if (dp_is_bad(ctrl)) return;  // the apply kernel poisons the gradient
  const uint32_t lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * DP_THREADS + threadIdx.x) >> 5;
  const int nw = (gridDim.x * DP_THREADS) >> 5;
  const int sub = lane / WP2, pl = lane % WP2;
  const unsigned long long* slot_p =
      (pl < WORLD) ? reinterpret_cast<const unsigned long long*>(peers.buf[pl] + lay.off_slot) : nullptr;
  unsigned int* own_flag = reinterpret_cast<unsigned int*>(peers.buf[rank] + lay.off_own_flag);
  __nv_bfloat16* own_rows = reinterpret_cast<__nv_bfloat16*>(peers.buf[rank] + lay.off_own_rows);
  for (long long k0 = static_cast<long long>(gw) * PER; k0 < lay.vs; k0 += static_cast<long long>(nw) * PER) {
    const long long v = (k0 + sub) * WORLD + rank;
    unsigned long long s = 0;
    if (slot_p != nullptr && v < V) {
      asm volatile("ld.global.L1::no_allocate.u64 %0, [%1];" : "=l"(s) : "l"(slot_p + v));
    }
    const bool present = (static_cast<unsigned int>(s >> 32) == epoch) && slot_p != nullptr && v < V &&
                         static_cast<long long>(s & 0xFFFFFFFFull) < lay.cap;
    const unsigned pm = __ballot_sync(0xffffffffu, present);
    if (pm == 0) continue;
    const unsigned int my_idx = static_cast<unsigned int>(s & 0xFFFFFFFFull);
    for (int q = 0; q < PER; ++q) {
      const unsigned m = (pm >> (q * WP2)) & ((1u << WP2) - 1u);
      if (m == 0) continue;
      __nv_bfloat16* orow = own_rows + static_cast<size_t>(k0 + q) * D;
      for (int d0 = 0; d0 < D; d0 += SPARSE_COLS) {
        const int dcols = min(SPARSE_COLS, D - d0);
        float acc[SPARSE_G][8];
#pragma unroll
        for (int g = 0; g < SPARSE_G; ++g)
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[g][e] = 0.f;
#pragma unroll
        for (int pb = 0; pb < WORLD; pb += 4) {  // 4 peers' loads in flight, summed in rank order
          uint4 u[4][SPARSE_G];
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const int p = pb + kk;
            if (p >= WORLD) break;
            const unsigned int idx = __shfl_sync(0xffffffffu, my_idx, q * WP2 + p);
            const bool has = (m >> p) & 1u;
            const __nv_bfloat16* src =
                reinterpret_cast<const __nv_bfloat16*>(peers.buf[p] + lay.off_rows) + static_cast<size_t>(idx) * D + d0;
#pragma unroll
            for (int g = 0; g < SPARSE_G; ++g) {
              const int col = g * 256 + static_cast<int>(lane) * 8;
              u[kk][g] = (has && col < dcols) ? ld_peer_u4(reinterpret_cast<const uint4*>(src + col))
                                              : make_uint4(0, 0, 0, 0);
            }
          }
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            if (pb + kk >= WORLD) break;
#pragma unroll
            for (int g = 0; g < SPARSE_G; ++g) {
              acc[g][0] += bf16lo(u[kk][g].x), acc[g][1] += bf16hi(u[kk][g].x);
              acc[g][2] += bf16lo(u[kk][g].y), acc[g][3] += bf16hi(u[kk][g].y);
              acc[g][4] += bf16lo(u[kk][g].z), acc[g][5] += bf16hi(u[kk][g].z);
              acc[g][6] += bf16lo(u[kk][g].w), acc[g][7] += bf16hi(u[kk][g].w);
            }
          }
        }
#pragma unroll
        for (int g = 0; g < SPARSE_G; ++g) {
          const int col = g * 256 + static_cast<int>(lane) * 8;
          if (col < dcols) *reinterpret_cast<uint4*>(orow + d0 + col) = f32x8_to_bf16(acc[g]);
        }
      }
      if (lane == 0) own_flag[k0 + q] = epoch;
    }
  }
}

constexpr int APPLY_IDS = 8;  // ids per warp iteration of the apply kernel (balances the dense DSU id range)

template <int WORLD>
__global__ void __launch_bounds__(DP_THREADS) dp_sparse_apply_kernel(DpPeers peers, DpLayout lay, int rank, int V, int D,
                                                                     unsigned int epoch, __nv_bfloat16* __restrict__ dW,
                                                                     unsigned int* host_flag, GradFold fold) {
  DpCtrl* ctrl = reinterpret_cast<DpCtrl*>(peers.buf[rank]);
  // everything this rank owns was reduced by the previous kernel of this stream
  if (blockIdx.x == 0 && threadIdx.x < WORLD && static_cast<int>(threadIdx.x) != rank) {
    __threadfence_system();
    st_release_sys(&reinterpret_cast<DpCtrl*>(peers.buf[threadIdx.x])->arrive_d[rank], epoch);
  }
  dp_wait_all(peers, ctrl->arrive_d, WORLD, rank, epoch, host_flag);
  if (dp_is_bad(ctrl)) {  // poison the first gradient row: grad-norm and loss checks trip at once
    if (blockIdx.x == 0)
      for (int j = threadIdx.x; j < D / 2; j += DP_THREADS) reinterpret_cast<uint32_t*>(dW)[j] = BF16X2_NAN;
    return;
  }
  const uint32_t lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * DP_THREADS + threadIdx.x) >> 5;
  const int nw = (gridDim.x * DP_THREADS) >> 5;
  const float ps = fold_scale(fold);
  float dsq = 0.f;
  const int row_u4 = D / 8;
  for (long long v0 = static_cast<long long>(gw) * APPLY_IDS; v0 < V; v0 += static_cast<long long>(nw) * APPLY_IDS) {
    const long long v = v0 + lane;
    unsigned int fl = epoch + 1u;
    if (lane < APPLY_IDS && v < V) {
      const unsigned int* fp =
          reinterpret_cast<const unsigned int*>(peers.buf[v % WORLD] + lay.off_own_flag) + v / WORLD;
      asm volatile("ld.global.L1::no_allocate.u32 %0, [%1];" : "=r"(fl) : "l"(fp));
    }
    unsigned pm = __ballot_sync(0xffffffffu, fl == epoch);
    while (pm) {
      const int q = __ffs(pm) - 1;
      pm &= pm - 1;
      const long long vq = v0 + q;
      const uint4* src = reinterpret_cast<const uint4*>(
          reinterpret_cast<const __nv_bfloat16*>(peers.buf[vq % WORLD] + lay.off_own_rows) +
          static_cast<size_t>(vq / WORLD) * D);
      uint4* dst = reinterpret_cast<uint4*>(dW + static_cast<size_t>(vq) * D);
      for (int c0 = 0; c0 < row_u4; c0 += 32 * 8) {  // 8 x 16 bytes in flight per lane (D = 2048: one trip)
        uint4 r[8], o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = c0 + j * 32 + static_cast<int>(lane);
          if (c < row_u4) r[j] = ld_peer_u4(src + c);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = c0 + j * 32 + static_cast<int>(lane);
          if (c < row_u4) o[j] = dst[c];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = c0 + j * 32 + static_cast<int>(lane);
          if (c >= row_u4) continue;
          uint4 w;
          w.x = pack_bf16x2(fmaf(ps, bf16lo(r[j].x), bf16lo(o[j].x)), fmaf(ps, bf16hi(r[j].x), bf16hi(o[j].x)));
          w.y = pack_bf16x2(fmaf(ps, bf16lo(r[j].y), bf16lo(o[j].y)), fmaf(ps, bf16hi(r[j].y), bf16hi(o[j].y)));
          w.z = pack_bf16x2(fmaf(ps, bf16lo(r[j].z), bf16lo(o[j].z)), fmaf(ps, bf16hi(r[j].z), bf16hi(o[j].z)));
          w.w = pack_bf16x2(fmaf(ps, bf16lo(r[j].w), bf16lo(o[j].w)), fmaf(ps, bf16hi(r[j].w), bf16hi(o[j].w)));
          dst[c] = w;
          if (fold.sq)
            dsq += (sq_of_bf16x2(w.x) - sq_of_bf16x2(o[j].x)) + (sq_of_bf16x2(w.y) - sq_of_bf16x2(o[j].y)) +
                   (sq_of_bf16x2(w.z) - sq_of_bf16x2(o[j].z)) + (sq_of_bf16x2(w.w) - sq_of_bf16x2(o[j].w));
        }
      }
    }
  }
  if (fold.sq) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dsq += __shfl_xor_sync(0xffffffffu, dsq, o);
    if (lane == 0) fold.sq[gw] = dsq;
  }
}

}  // namespace ssi
