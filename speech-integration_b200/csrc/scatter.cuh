// Tied-embedding gradient (K3): dW[ids[i], :] += dH[i, :]  -- the backward of the nn.Embedding built at
// tt/torchtune/models/llama3_2/_component_builders.py:114 and called at tt/torchtune/modules/transformer.py:628.
//
// Three launches, no floating-point atomics, bit-identical run to run:
//   sort_grid_kernel       up to 64 co-resident 1024-thread blocks (one per 2048 ids, cooperative launch, grid
//                          barriers inside): stable LSD radix sort of (id, row) with 9-bit digits (two passes cover
//                          V < 2^18), ids read as int64 in the first pass, then the run index of every sorted
//                          position ("urank": how many distinct ids precede it) and the number of distinct ids.
//   scatter_reduce_kernel  one warp per 32 consecutive sorted positions ("range") and 256-column slab: runs of equal ids
//                          that lie inside the warp's range are summed in fp32 in sorted order and written with one
//                          access per touched row; runs that cross the range boundary leave an fp32 partial row.
//   scatter_fixup_kernel   chains the partial rows of a crossing run in order and writes that row.
// Output forms (ScatMode): read-modify-write add into a dense bf16 or fp32 [V, D] buffer, or -- for the sparse
// data-parallel exchange (dp_exchange.cuh) -- a COMPACT list: rows[u, :] = bf16 sum of the u-th distinct id,
// ids[u] = that id, slot[id] = (epoch << 32 | u).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "aux_kernels.cuh"
#include "ptx.cuh"

namespace ssi {

constexpr int SORT_THREADS = 1024;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_BITS = 9;
constexpr int SORT_DIGITS = 1 << SORT_BITS;
constexpr int SORT_BATCH = 8;  // keys in flight per lane before they are ranked (memory-level parallelism)
constexpr int SORT_SMEM_BYTES = SORT_WARPS * SORT_DIGITS * 4;
constexpr uint32_t KEY_NONE = 0xFFFFFFFFu;

// exclusive prefix of `v` over the 1024 threads of the block; `total` = block sum.  s_w: 33 words of shared memory.
__device__ __forceinline__ uint32_t block_excl_scan_1024(uint32_t v, uint32_t* s_w, uint32_t& total) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= static_cast<uint32_t>(o)) inc += t;
  }
  if (lane == 31) s_w[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const uint32_t w = s_w[lane];
    uint32_t winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= static_cast<uint32_t>(o)) winc += t;
    }
    s_w[lane] = winc - w;
    if (lane == 31) s_w[32] = winc;
  }
  __syncthreads();
  const uint32_t r = s_w[warp] + inc - v;
  total = s_w[32];
  __syncthreads();  // s_w may be reused by the caller
  return r;
}

// Grid-wide barrier of the sort kernel (launched cooperatively, so all blocks are resident): arrival counter in
// global memory, k-th barrier waits for k * gridDim.x arrivals.  A wait of 10 s means a lost block: trap (the launch
// fails loudly) instead of hanging the stream.
__device__ __forceinline__ void sort_grid_barrier(unsigned int* ctr, unsigned int target) {
  __syncthreads();
  if (gridDim.x > 1 && threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
    unsigned long long t0 = 0;
    unsigned int spins = 0;
    while (static_cast<int>(ld_acquire_gpu(ctr) - target) < 0) {
      if ((++spins & 1023u) == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t0 == 0) t0 = t;
        else if (t - t0 > 10000000000ull) __trap();
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// Stable sort of the token ids by a grid of G <= SORT_MAX_BLOCKS co-resident blocks.  Pass p reads buffer
// (p - 1) & 1 (pass 0: the int64 ids; out-of-range ids become the sentinel key V, which sorts last) and writes buffer
// p & 1; the result is in buffer (passes - 1) & 1.  Every block owns one contiguous segment of the input and every
// warp one contiguous part of it: positions within a digit are ordered by (block, warp, round, lane), i.e. by input
// position, so each pass is stable.  Per pass: per-warp digit counts in shared memory -> per-block counts to ghist
// [digit][block] -> grid barrier -> every block scans ghist (digit-major, block-minor) for its bases -> scatter ->
// grid barrier.  Buffers and ghist are read with ld.global.cg: they are rewritten inside the kernel.
constexpr int SORT_MAX_BLOCKS = 64;
#ifndef SSI_SORT_KEYS_PER_BLOCK
#define SSI_SORT_KEYS_PER_BLOCK 2048
#endif
constexpr int SORT_KEYS_PER_BLOCK = SSI_SORT_KEYS_PER_BLOCK;

__global__ void __launch_bounds__(SORT_THREADS, 1) sort_grid_kernel(const int64_t* __restrict__ ids, int n,
                                                                    long long V, int passes, uint32_t* kbuf0,
                                                                    uint32_t* vbuf0, uint32_t* kbuf1,
                                                                    uint32_t* vbuf1, int* __restrict__ urank,
                                                                    int* __restrict__ n_unique, uint32_t* ghist,
                                                                    unsigned int* bar) {
  extern __shared__ uint32_t s_cnt[];  // [SORT_WARPS][SORT_DIGITS]
  __shared__ uint32_t s_w[33];
  __shared__ uint32_t s_tot[SORT_DIGITS], s_bef[SORT_DIGITS];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int G = static_cast<int>(gridDim.x), blk = static_cast<int>(blockIdx.x);
  // block segments are multiples of 1024 keys: every warp part is a multiple of 32
  const int bseg = ((n + G - 1) / G + SORT_THREADS - 1) / SORT_THREADS * SORT_THREADS;
  const int wseg = bseg / SORT_WARPS;
  const long long w0pos = static_cast<long long>(blk) * bseg + static_cast<long long>(warp) * wseg;
  const int wbeg = static_cast<int>(w0pos < n ? w0pos : n);
  const int wend = static_cast<int>(w0pos + wseg < n ? w0pos + wseg : n);
  uint32_t* mycnt = s_cnt + warp * SORT_DIGITS;
  unsigned int barrier_no = 0;

  for (int pass = 0; pass < passes; ++pass) {
    const int shift = pass * SORT_BITS;
    const uint32_t* kin = (pass & 1) ? kbuf0 : kbuf1;
    const uint32_t* vin = (pass & 1) ? vbuf0 : vbuf1;
    uint32_t* kout = (pass & 1) ? kbuf1 : kbuf0;
    uint32_t* vout = (pass & 1) ? vbuf1 : vbuf0;
    auto load_key = [&](int i) -> uint32_t {
      if (pass == 0) {
        const long long id = ids[i];
        return (id >= 0 && id < V) ? static_cast<uint32_t>(id) : static_cast<uint32_t>(V);
      }
      return __ldcg(kin + i);
    };
    for (int i = threadIdx.x; i < SORT_WARPS * SORT_DIGITS; i += SORT_THREADS) s_cnt[i] = 0;
    __syncthreads();
    // ---- digit counts of this warp's part
    for (int i0 = wbeg; i0 < wend; i0 += 32 * SORT_BATCH) {
      uint32_t key[SORT_BATCH];
#pragma unroll
      for (int b = 0; b < SORT_BATCH; ++b) {
        const int i = i0 + b * 32 + static_cast<int>(lane);
        key[b] = (i < wend) ? load_key(i) : 0u;
      }
#pragma unroll
      for (int b = 0; b < SORT_BATCH; ++b) {
        if (i0 + b * 32 >= wend) break;  // warp-uniform
        const bool act = i0 + b * 32 + static_cast<int>(lane) < wend;
        const unsigned amask = __ballot_sync(0xffffffffu, act);
        if (act) {
          const uint32_t d = (key[b] >> shift) & (SORT_DIGITS - 1);
          const unsigned peers = __match_any_sync(amask, d);
          if ((peers & lt_mask) == 0) mycnt[d] += __popc(peers);
        }
        __syncwarp();
      }
    }
    __syncthreads();
    // ---- thread t owns digit t / 2 and warps 16 (t & 1) .. + 15 of this block, and -- in ghist -- blocks of half t & 1
    const int d = threadIdx.x >> 1, h = threadIdx.x & 1, w0 = h * 16;
    uint32_t c[16], sumc = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      c[j] = s_cnt[(w0 + j) * SORT_DIGITS + d];
      sumc += c[j];
    }
    const uint32_t partner_sumc = __shfl_xor_sync(0xffffffffu, sumc, 1);
    uint32_t half_total = h ? 0u : sumc + partner_sumc, before = 0;  // G == 1: the block's own counts
    if (G > 1) {
      if (h == 0) ghist[d * G + blk] = sumc + partner_sumc;
      sort_grid_barrier(bar, ++barrier_no * static_cast<unsigned int>(G));
      // warp w sums digits [16 w, 16 w + 16) over the blocks: lane = block index (and + 32), all loads in flight
      uint32_t x0[16], x1[16];
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        const uint32_t* row = ghist + (warp * 16 + q) * G;
        x0[q] = (static_cast<int>(lane) < G) ? __ldcg(row + lane) : 0u;
        x1[q] = (static_cast<int>(lane) + 32 < G) ? __ldcg(row + lane + 32) : 0u;
      }
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        uint32_t tot = x0[q] + x1[q];
        uint32_t bef = (static_cast<int>(lane) < blk ? x0[q] : 0u) + (static_cast<int>(lane) + 32 < blk ? x1[q] : 0u);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          tot += __shfl_xor_sync(0xffffffffu, tot, o);
          bef += __shfl_xor_sync(0xffffffffu, bef, o);
        }
        if (lane == 0) s_tot[warp * 16 + q] = tot, s_bef[warp * 16 + q] = bef;
      }
      __syncthreads();
      half_total = h ? 0u : s_tot[d];
      before = h ? 0u : s_bef[d];
    }
    uint32_t total;
    const uint32_t start = block_excl_scan_1024(half_total, s_w, total);
    const uint32_t digit_start = __shfl_sync(0xffffffffu, start, lane & ~1u);
    before += __shfl_xor_sync(0xffffffffu, before, 1);
    uint32_t run = digit_start + before + (h ? partner_sumc : 0u);
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      s_cnt[(w0 + j) * SORT_DIGITS + d] = run;
      run += c[j];
    }
    __syncthreads();
    // ---- stable scatter
    for (int i0 = wbeg; i0 < wend; i0 += 32 * SORT_BATCH) {
      uint32_t key[SORT_BATCH], val[SORT_BATCH];
#pragma unroll
      for (int b = 0; b < SORT_BATCH; ++b) {
        const int i = i0 + b * 32 + static_cast<int>(lane);
        key[b] = 0u, val[b] = 0u;
        if (i < wend) {
          key[b] = load_key(i);
          val[b] = (pass == 0) ? static_cast<uint32_t>(i) : __ldcg(vin + i);
        }
      }
#pragma unroll
      for (int b = 0; b < SORT_BATCH; ++b) {
        if (i0 + b * 32 >= wend) break;
        const bool act = i0 + b * 32 + static_cast<int>(lane) < wend;
        const unsigned amask = __ballot_sync(0xffffffffu, act);
        uint32_t dg = 0, base = 0;
        unsigned peers = 0;
        if (act) {
          dg = (key[b] >> shift) & (SORT_DIGITS - 1);
          peers = __match_any_sync(amask, dg);
          base = mycnt[dg];
          const uint32_t pos = base + __popc(peers & lt_mask);
          kout[pos] = key[b];
          vout[pos] = val[b];
        }
        __syncwarp();
        if (act && (peers & lt_mask) == 0) mycnt[dg] = base + __popc(peers);
        __syncwarp();
      }
    }
    __threadfence();
    sort_grid_barrier(bar, ++barrier_no * static_cast<unsigned int>(G));
  }

  // ---- run index of every sorted position and the number of distinct in-range ids
  const uint32_t* ks = ((passes - 1) & 1) ? kbuf1 : kbuf0;
  uint32_t heads = 0;
  for (int i0 = wbeg; i0 < wend; i0 += 32) {
    const int i = i0 + static_cast<int>(lane);
    bool hd = false;
    if (i < wend) hd = (i == 0) || (__ldcg(ks + i) != __ldcg(ks + i - 1));
    heads += __popc(__ballot_sync(0xffffffffu, hd));
  }
  uint32_t total;
  uint32_t run = block_excl_scan_1024(lane == 0 ? heads : 0u, s_w, total);
  run = __shfl_sync(0xffffffffu, run, 0);  // heads before this warp's part, within the block
  uint32_t all = total;
  if (G > 1) {
    if (threadIdx.x == 0) ghist[blk] = total;
    sort_grid_barrier(bar, ++barrier_no * static_cast<unsigned int>(G));
    const uint32_t y0 = (static_cast<int>(lane) < G) ? __ldcg(ghist + lane) : 0u;
    const uint32_t y1 = (static_cast<int>(lane) + 32 < G) ? __ldcg(ghist + lane + 32) : 0u;
    all = y0 + y1;
    uint32_t bef = (static_cast<int>(lane) < blk ? y0 : 0u) + (static_cast<int>(lane) + 32 < blk ? y1 : 0u);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      all += __shfl_xor_sync(0xffffffffu, all, o);
      bef += __shfl_xor_sync(0xffffffffu, bef, o);
    }
    run += bef;
  }
  for (int i0 = wbeg; i0 < wend; i0 += 32) {
    const int i = i0 + static_cast<int>(lane);
    bool hd = false;
    if (i < wend) hd = (i == 0) || (__ldcg(ks + i) != __ldcg(ks + i - 1));
    const unsigned m = __ballot_sync(0xffffffffu, hd);
    if (i < wend) urank[i] = static_cast<int>(run + __popc(m & (lt_mask | (1u << lane)))) - 1;
    run += __popc(m);
  }
  if (blk == 0 && threadIdx.x == 0) {
    const bool has_sentinel = n > 0 && __ldcg(ks + n - 1) >= static_cast<uint32_t>(V);
    n_unique[0] = static_cast<int>(all) - (has_sentinel ? 1 : 0);
  }
}

// ---------------------------------------------------------------------------------------------------
constexpr int SCAT_COLS = 32 * 8;  // columns per slab (blockIdx.y): one 16-byte group per lane and row
#ifndef SSI_SCAT_AHEAD
#define SSI_SCAT_AHEAD 3
#endif
constexpr int SCAT_AHEAD = SSI_SCAT_AHEAD;  // rows in flight per warp ahead of the one being accumulated
constexpr int SCAT_THREADS = 128;
enum ScatMode { SCAT_DENSE_BF16 = 0, SCAT_DENSE_F32 = 1, SCAT_COMPACT = 2 };

struct ScatOut {
  void* dense;               // dense modes: dW [V, D] (bf16 or fp32), rows are added to
  __nv_bfloat16* rows;       // compact mode: [cap, D] bf16 sums, written (not added)
  int* ids;                  // compact mode: [cap]
  unsigned long long* slot;  // compact mode: [V], (epoch << 32 | index into rows)
  unsigned int epoch;
  int cap;
  // fp32 dense mode: rows are multiplied by scale_sign * (*scale_num) / (*scale_den) before they are added
  // (null scale_num: plain add).  Used for the onehot term of the LM-head gradient, keyed by the labels.
  const float* scale_num;
  const long long* scale_den;
  float scale_sign;
  // fp32 dense mode: optional per-SOURCE-row weight (indexed like dH's rows; null = 1): row r joins its id's sum as
  // row_w[r] * dH[r, :].  The label-keyed term of the LM-head gradient weights row r with 1 - p_label[r].
  const float* row_w;
  // bf16 dense mode: scale_grads / grad-norm fold (aux_kernels.cuh GradFold): rows are multiplied by *post_scale
  // before they are added; fold.sq[warp] (scatter_reduce) and fold.sq[warps + warp] (scatter_fixup) receive
  // sum(new^2 - old^2) over the elements the warp rewrote
  GradFold fold;
};

struct ScatInfo {
  uint32_t slot0_key;  // KEY_NONE if no left-open run
  uint32_t slot0_open_right;
  uint32_t slot1_key;  // KEY_NONE if no right-open run that starts inside
  int slot1_urank;     // run index of that run
};

// The 8 columns of a dense output row this lane adds to, fetched when the run STARTS so that the read-modify-write
// at its end does not wait for memory.
struct ScatOld {
  uint4 a, b;  // bf16 mode: a; fp32 mode: a, b (two float4)
};

template <int MODE>
__device__ __forceinline__ void fetch_old(const ScatOut& o, uint32_t key, uint32_t V, int D, int col, bool active,
                                          ScatOld& old) {
  if constexpr (MODE == SCAT_COMPACT) return;
  if (!active || key >= V) return;
  if constexpr (MODE == SCAT_DENSE_BF16) {
    old.a = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(o.dense) + static_cast<size_t>(key) * D +
                                            col);
  } else {
    const uint4* ptr = reinterpret_cast<const uint4*>(static_cast<const float*>(o.dense) + static_cast<size_t>(key) * D +
                                                      col);
    old.a = ptr[0], old.b = ptr[1];
  }
}

__device__ __forceinline__ float scat_scale(const ScatOut& o) {
  if (!o.scale_num) return 1.f;
  const long long den = *o.scale_den;
  return (den > 0) ? o.scale_sign * (*o.scale_num / static_cast<float>(den)) : 0.f;
}

// one finished run -> its output row (columns [col, col + 8) of it); `old` = what fetch_old read for this key
template <int MODE>
__device__ __forceinline__ void emit_run(const ScatOut& o, uint32_t key, int ur, uint32_t V, int D, int col,
                                         bool active, bool first_slab_lane0, const float (&acc)[8],
                                         const ScatOld& old, float sc, float& dsq) {
  if (key >= V) return;  // sentinel run of the out-of-range ids
  if constexpr (MODE == SCAT_COMPACT) {
    if (ur < 0 || ur >= o.cap) return;
    if (active) *reinterpret_cast<uint4*>(o.rows + static_cast<size_t>(ur) * D + col) = f32x8_to_bf16(acc);
    if (first_slab_lane0) {
      o.ids[ur] = static_cast<int>(key);
      o.slot[key] = (static_cast<unsigned long long>(o.epoch) << 32) | static_cast<unsigned int>(ur);
    }
  } else if constexpr (MODE == SCAT_DENSE_BF16) {
    if (!active) return;
    float f[8];
    bf16x8_to_f32(old.a, f);
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = fmaf(sc, acc[e], f[e]);  // sc = 1 without a fold: the plain sum
    const uint4 nw = f32x8_to_bf16(f);
    *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(o.dense) + static_cast<size_t>(key) * D + col) = nw;
    if (o.fold.sq)
      dsq += (sq_of_bf16x2(nw.x) - sq_of_bf16x2(old.a.x)) + (sq_of_bf16x2(nw.y) - sq_of_bf16x2(old.a.y)) +
             (sq_of_bf16x2(nw.z) - sq_of_bf16x2(old.a.z)) + (sq_of_bf16x2(nw.w) - sq_of_bf16x2(old.a.w));
  } else {
    if (!active) return;
    float4* ptr = reinterpret_cast<float4*>(static_cast<float*>(o.dense) + static_cast<size_t>(key) * D + col);
    float4 a, b;
    a.x = fmaf(sc, acc[0], __uint_as_float(old.a.x)), a.y = fmaf(sc, acc[1], __uint_as_float(old.a.y));
    a.z = fmaf(sc, acc[2], __uint_as_float(old.a.z)), a.w = fmaf(sc, acc[3], __uint_as_float(old.a.w));
    b.x = fmaf(sc, acc[4], __uint_as_float(old.b.x)), b.y = fmaf(sc, acc[5], __uint_as_float(old.b.y));
    b.z = fmaf(sc, acc[6], __uint_as_float(old.b.z)), b.w = fmaf(sc, acc[7], __uint_as_float(old.b.w));
    ptr[0] = a, ptr[1] = b;
  }
}

// 8 resident blocks of 128 threads per SM (64 registers) and 3 rows in flight per warp measured best on B200
// (profiles/r02_k3_variants.txt): more rows in flight cost occupancy and lose
#ifndef SSI_SCAT_MINBLOCKS
#define SSI_SCAT_MINBLOCKS 8
#endif
template <int MODE>
__global__ void __launch_bounds__(SCAT_THREADS, SSI_SCAT_MINBLOCKS) scatter_reduce_kernel(const __nv_bfloat16* __restrict__ dH,
                                                             const uint32_t* __restrict__ skeys,
                                                             const uint32_t* __restrict__ srows,
                                                             const int* __restrict__ urank, int n, uint32_t V, int D,
                                                             int slabs, ScatOut out, float* __restrict__ partial,
                                                             ScatInfo* __restrict__ info) {
  // warp index = slab * ranges + range (measured: faster than running the slabs of one range side by side)
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nranges = (n + 31) >> 5;
  const int slab = wid / nranges, wg = wid - slab * nranges;
  const uint32_t lane = threadIdx.x & 31;
  const int p0 = wg * 32;
  if (slab >= slabs) return;
  const int col = slab * SCAT_COLS + static_cast<int>(lane) * 8;
  const bool active = col < D;
  const bool lead = slab == 0 && lane == 0;
  const int cnt = min(32, n - p0);
  const bool in = static_cast<int>(lane) < cnt;
  const uint32_t my_key = in ? skeys[p0 + lane] : KEY_NONE;
  const uint32_t my_row = in ? srows[p0 + lane] : 0u;
  const int my_ur = in ? urank[p0 + lane] : -1;
  float my_w = 1.f;
  if constexpr (MODE == SCAT_DENSE_F32) {
    if (in && out.row_w) my_w = out.row_w[my_row];
  }
  const uint32_t prev_key = (p0 > 0) ? skeys[p0 - 1] : KEY_NONE;
  const uint32_t next_key = (p0 + cnt < n) ? skeys[p0 + cnt] : KEY_NONE;
  const float sc = (MODE == SCAT_DENSE_F32) ? scat_scale(out) : (MODE == SCAT_DENSE_BF16) ? fold_scale(out.fold) : 1.f;
  float dsq = 0.f;
  ScatInfo inf;
  inf.slot0_key = KEY_NONE;
  inf.slot0_open_right = 0;
  inf.slot1_key = KEY_NONE;
  inf.slot1_urank = -1;

  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  int run_start = 0;
  // software pipeline: the loads of rows i + 1 .. i + SCAT_AHEAD are in flight while row i is accumulated
  uint4 u[SCAT_AHEAD + 1];
  auto load_row = [&](int i) -> uint4 {
    const uint32_t row = __shfl_sync(0xffffffffu, my_row, i & 31);
    return (i < cnt && active) ? __ldg(reinterpret_cast<const uint4*>(dH + static_cast<size_t>(row) * D + col))
                               : make_uint4(0, 0, 0, 0);
  };
#pragma unroll
  for (int a = 0; a < SCAT_AHEAD; ++a) u[a] = load_row(a);
  ScatOld old;
  old.a = old.b = make_uint4(0, 0, 0, 0);
  for (int i = 0; i < cnt; ++i) {
    const uint32_t key = __shfl_sync(0xffffffffu, my_key, i);
    const uint32_t nkey = (i + 1 < cnt) ? __shfl_sync(0xffffffffu, my_key, (i + 1) & 31) : KEY_NONE;
    const int ur = __shfl_sync(0xffffffffu, my_ur, i);
    u[SCAT_AHEAD] = load_row(i + SCAT_AHEAD);
    const bool open_left = (run_start == 0) && (prev_key == key);
    if (i == run_start && !open_left) fetch_old<MODE>(out, key, V, D, col, active, old);  // unused if open right
    {
      float f[8];
      bf16x8_to_f32(u[0], f);
      if constexpr (MODE == SCAT_DENSE_F32) {
        const float w = __shfl_sync(0xffffffffu, my_w, i);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(w, f[e], acc[e]);  // w == 1: the plain sum, bit for bit
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += f[e];
      }
    }
#pragma unroll
    for (int a = 0; a < SCAT_AHEAD; ++a) u[a] = u[a + 1];
    const bool end_of_run = (i == cnt - 1) || (nkey != key);
    if (end_of_run) {
      const bool open_right = (i == cnt - 1) && (next_key == key);
      if (!open_left && !open_right) {
        emit_run<MODE>(out, key, ur, V, D, col, active, lead, acc, old, sc, dsq);
      } else {
        const int slot = open_left ? 0 : 1;
        if (active) {
          float* pr = partial + (static_cast<size_t>(wg) * 2 + slot) * D + col;
          *reinterpret_cast<float4*>(pr) = make_float4(acc[0], acc[1], acc[2], acc[3]);
          *reinterpret_cast<float4*>(pr + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
        }
        if (open_left) {
          inf.slot0_key = key;
          inf.slot0_open_right = open_right ? 1u : 0u;
        } else {
          inf.slot1_key = key;
          inf.slot1_urank = ur;
        }
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.f;
      run_start = i + 1;
    }
  }
  if (lead) info[wg] = inf;
  if constexpr (MODE == SCAT_DENSE_BF16) {
    if (out.fold.sq) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dsq += __shfl_xor_sync(0xffffffffu, dsq, o);
      if (lane == 0) out.fold.sq[wid] = dsq;
    }
  }
}

// One warp per range (and column slab) whose last run continues into the next range(s): sum the chain of partials
// in range order and write that run's row.  The chain is found 32 ranges at a time (one info record per lane).
template <int MODE>
__global__ void __launch_bounds__(SCAT_THREADS) scatter_fixup_kernel(const float* __restrict__ partial,
                                                            const ScatInfo* __restrict__ info, int nwarps, uint32_t V,
                                                            int D, int slabs, ScatOut out) {
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int slab = wid / nwarps, wg = wid - slab * nwarps;
  const uint32_t lane = threadIdx.x & 31;
  if (slab >= slabs) return;
  const uint32_t key = info[wg].slot1_key;
  if (key == KEY_NONE) return;
  const int col = slab * SCAT_COLS + static_cast<int>(lane) * 8;
  const bool active = col < D;
  ScatOld old;
  old.a = old.b = make_uint4(0, 0, 0, 0);
  fetch_old<MODE>(out, key, V, D, col, active, old);
  const float sc = (MODE == SCAT_DENSE_F32) ? scat_scale(out) : (MODE == SCAT_DENSE_BF16) ? fold_scale(out.fold) : 1.f;
  float dsq = 0.f;
  float acc[8];
  auto add_partial = [&](const float* pr, bool first) {
    float4 a = make_float4(0, 0, 0, 0), b = a;
    if (active) {
      a = *reinterpret_cast<const float4*>(pr + col);
      b = *reinterpret_cast<const float4*>(pr + col + 4);
    }
    if (first) {
      acc[0] = a.x, acc[1] = a.y, acc[2] = a.z, acc[3] = a.w, acc[4] = b.x, acc[5] = b.y, acc[6] = b.z, acc[7] = b.w;
    } else {
      acc[0] += a.x, acc[1] += a.y, acc[2] += a.z, acc[3] += a.w;
      acc[4] += b.x, acc[5] += b.y, acc[6] += b.z, acc[7] += b.w;
    }
  };
  add_partial(partial + (static_cast<size_t>(wg) * 2 + 1) * D, true);
  for (int base = wg + 1; base < nwarps; base += 32) {
    const int w = base + static_cast<int>(lane);
    bool cont = false, open = false;
    if (w < nwarps) {
      const ScatInfo inf = info[w];
      cont = inf.slot0_key == key;  // defensive: the chain must continue with the same id
      open = cont && inf.slot0_open_right != 0;
    }
    const unsigned m_cont = __ballot_sync(0xffffffffu, cont), m_open = __ballot_sync(0xffffffffu, open);
    const int f = (~m_open) ? __ffs(~m_open) - 1 : 32;  // first range that does not continue to its right
    const int n_inc = (f < 32) ? (((m_cont >> f) & 1u) ? f + 1 : f) : 32;
    for (int j = 0; j < n_inc; ++j) add_partial(partial + (static_cast<size_t>(base + j) * 2) * D, false);
    if (f < 32) break;
  }
  emit_run<MODE>(out, key, info[wg].slot1_urank, V, D, col, active, slab == 0 && lane == 0, acc, old, sc, dsq);
  if constexpr (MODE == SCAT_DENSE_BF16) {
    if (out.fold.sq) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dsq += __shfl_xor_sync(0xffffffffu, dsq, o);
      if (lane == 0) out.fold.sq[static_cast<size_t>(nwarps) * slabs + wid] = dsq;
    }
  }
}

}  // namespace ssi
