// Tied-embedding gradient (K3): dW[ids[i], :] += dH[i, :]  -- the backward of the nn.Embedding built at
// tt/torchtune/models/llama3_2/_component_builders.py:114 and called at tt/torchtune/modules/transformer.py:628.
//
// Three launches, no floating-point atomics, bit-identical run to run:
//   sort_block_kernel      ONE 1024-thread block: stable LSD radix sort of (id, row) with 9-bit digits (two passes
//                          cover V < 2^18), ids read as int64 in the first pass, then the run index of every sorted
//                          position ("urank": how many distinct ids precede it) and the number of distinct ids.
//   scatter_reduce_kernel  one warp per 32 consecutive sorted positions and 1024-column slab: runs of equal ids
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

// Stable sort of the token ids.  Pass p reads buffer (p - 1) & 1 (pass 0: the int64 ids; out-of-range ids become the
// sentinel key V, which sorts last) and writes buffer p & 1; the result is in buffer (passes - 1) & 1.  Every warp
// owns one contiguous segment of the input: positions within a digit are ordered by (warp, round, lane), i.e. by
// input position, so each pass is stable.  Buffers are read with ld.global.cg: they are rewritten inside the kernel.
__global__ void __launch_bounds__(SORT_THREADS, 1) sort_block_kernel(const int64_t* __restrict__ ids, int n,
                                                                     long long V, int passes, uint32_t* kbuf0,
                                                                     uint32_t* vbuf0, uint32_t* kbuf1,
                                                                     uint32_t* vbuf1, int* __restrict__ urank,
                                                                     int* __restrict__ n_unique) {
  extern __shared__ uint32_t s_cnt[];  // [SORT_WARPS][SORT_DIGITS]
  __shared__ uint32_t s_w[33];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int seg = (((n + SORT_WARPS - 1) / SORT_WARPS) + 31) / 32 * 32;
  const int wbeg = min(static_cast<int>(warp) * seg, n), wend = min(wbeg + seg, n);
  uint32_t* mycnt = s_cnt + warp * SORT_DIGITS;

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
    // ---- digit counts of this warp's segment
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
    // ---- exclusive scan over (digit-major, warp-minor): thread t owns digit t / 2, warps 16 (t & 1) .. + 15
    {
      const int d = threadIdx.x >> 1, w0 = (threadIdx.x & 1) * 16;
      uint32_t c[16], sum = 0;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        c[j] = s_cnt[(w0 + j) * SORT_DIGITS + d];
        sum += c[j];
      }
      uint32_t total;
      uint32_t run = block_excl_scan_1024(sum, s_w, total);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        s_cnt[(w0 + j) * SORT_DIGITS + d] = run;
        run += c[j];
      }
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
        uint32_t d = 0, base = 0;
        unsigned peers = 0;
        if (act) {
          d = (key[b] >> shift) & (SORT_DIGITS - 1);
          peers = __match_any_sync(amask, d);
          base = mycnt[d];
          const uint32_t pos = base + __popc(peers & lt_mask);
          kout[pos] = key[b];
          vout[pos] = val[b];
        }
        __syncwarp();
        if (act && (peers & lt_mask) == 0) mycnt[d] = base + __popc(peers);
        __syncwarp();
      }
    }
    __threadfence_block();
    __syncthreads();
  }

  // ---- run index of every sorted position and the number of distinct in-range ids
  const uint32_t* ks = ((passes - 1) & 1) ? kbuf1 : kbuf0;
  uint32_t heads = 0;
  for (int i0 = wbeg; i0 < wend; i0 += 32) {
    const int i = i0 + static_cast<int>(lane);
    bool h = false;
    if (i < wend) h = (i == 0) || (__ldcg(ks + i) != __ldcg(ks + i - 1));
    heads += __popc(__ballot_sync(0xffffffffu, h));
  }
  uint32_t total;
  uint32_t run = block_excl_scan_1024(lane == 0 ? heads : 0u, s_w, total);
  run = __shfl_sync(0xffffffffu, run, 0);  // heads before this warp's segment
  for (int i0 = wbeg; i0 < wend; i0 += 32) {
    const int i = i0 + static_cast<int>(lane);
    bool h = false;
    if (i < wend) h = (i == 0) || (__ldcg(ks + i) != __ldcg(ks + i - 1));
    const unsigned m = __ballot_sync(0xffffffffu, h);
    if (i < wend) urank[i] = static_cast<int>(run + __popc(m & (lt_mask | (1u << lane)))) - 1;
    run += __popc(m);
  }
  if (threadIdx.x == 0) {
    const bool has_sentinel = n > 0 && __ldcg(ks + n - 1) >= static_cast<uint32_t>(V);
    n_unique[0] = static_cast<int>(total) - (has_sentinel ? 1 : 0);
  }
}

// ---------------------------------------------------------------------------------------------------
constexpr int SCAT_G = 4;                   // 16-byte column groups per lane and pass
constexpr int SCAT_COLS = 32 * SCAT_G * 8;  // 1024 columns per slab (blockIdx.y)
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
};

struct ScatInfo {
  uint32_t slot0_key;  // KEY_NONE if no left-open run
  uint32_t slot0_open_right;
  uint32_t slot1_key;  // KEY_NONE if no right-open run that starts inside
  int slot1_urank;     // run index of that run
};

// one finished run -> its output row (columns [d0, d0 + dcols) of it)
template <int MODE>
__device__ __forceinline__ void emit_run(const ScatOut& o, uint32_t key, int ur, uint32_t V, int D, int d0, int dcols,
                                         uint32_t lane, const float (&acc)[SCAT_G][8]) {
  if (key >= V) return;  // sentinel run of the out-of-range ids
  if constexpr (MODE == SCAT_COMPACT) {
    if (ur < 0 || ur >= o.cap) return;
    __nv_bfloat16* dst = o.rows + static_cast<size_t>(ur) * D + d0;
#pragma unroll
    for (int g = 0; g < SCAT_G; ++g) {
      const int col = g * 256 + static_cast<int>(lane) * 8;
      if (col < dcols) *reinterpret_cast<uint4*>(dst + col) = f32x8_to_bf16(acc[g]);
    }
    if (d0 == 0 && lane == 0) {
      o.ids[ur] = static_cast<int>(key);
      o.slot[key] = (static_cast<unsigned long long>(o.epoch) << 32) | static_cast<unsigned int>(ur);
    }
  } else if constexpr (MODE == SCAT_DENSE_BF16) {
    __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(o.dense) + static_cast<size_t>(key) * D + d0;
#pragma unroll
    for (int g = 0; g < SCAT_G; ++g) {
      const int col = g * 256 + static_cast<int>(lane) * 8;
      if (col < dcols) {
        uint4* ptr = reinterpret_cast<uint4*>(dst + col);
        float f[8];
        bf16x8_to_f32(*ptr, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] += acc[g][e];
        *ptr = f32x8_to_bf16(f);
      }
    }
  } else {
    float* dst = static_cast<float*>(o.dense) + static_cast<size_t>(key) * D + d0;
    float sc = 1.f;
    if (o.scale_num) {
      const long long den = *o.scale_den;
      sc = (den > 0) ? o.scale_sign * (*o.scale_num / static_cast<float>(den)) : 0.f;
    }
#pragma unroll
    for (int g = 0; g < SCAT_G; ++g) {
      const int col = g * 256 + static_cast<int>(lane) * 8;
      if (col < dcols) {
        float4* ptr = reinterpret_cast<float4*>(dst + col);
        float4 a = ptr[0], b = ptr[1];
        a.x = fmaf(sc, acc[g][0], a.x), a.y = fmaf(sc, acc[g][1], a.y), a.z = fmaf(sc, acc[g][2], a.z);
        a.w = fmaf(sc, acc[g][3], a.w);
        b.x = fmaf(sc, acc[g][4], b.x), b.y = fmaf(sc, acc[g][5], b.y), b.z = fmaf(sc, acc[g][6], b.z);
        b.w = fmaf(sc, acc[g][7], b.w);
        ptr[0] = a, ptr[1] = b;
      }
    }
  }
}

template <int MODE>
__global__ void __launch_bounds__(128) scatter_reduce_kernel(const __nv_bfloat16* __restrict__ dH,
                                                             const uint32_t* __restrict__ skeys,
                                                             const uint32_t* __restrict__ srows,
                                                             const int* __restrict__ urank, int n, uint32_t V, int D,
                                                             ScatOut out, float* __restrict__ partial,
                                                             ScatInfo* __restrict__ info) {
  const int wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  const int p0 = wg * 32;
  if (p0 >= n) return;
  const int d0 = blockIdx.y * SCAT_COLS;
  const int dcols = min(SCAT_COLS, D - d0);
  const int cnt = min(32, n - p0);
  const bool in = static_cast<int>(lane) < cnt;
  const uint32_t my_key = in ? skeys[p0 + lane] : KEY_NONE;
  const uint32_t my_row = in ? srows[p0 + lane] : 0u;
  const int my_ur = in ? urank[p0 + lane] : -1;
  const uint32_t prev_key = (p0 > 0) ? skeys[p0 - 1] : KEY_NONE;
  const uint32_t next_key = (p0 + cnt < n) ? skeys[p0 + cnt] : KEY_NONE;
  ScatInfo inf;
  inf.slot0_key = KEY_NONE;
  inf.slot0_open_right = 0;
  inf.slot1_key = KEY_NONE;
  inf.slot1_urank = -1;

  float acc[SCAT_G][8];
#pragma unroll
  for (int g = 0; g < SCAT_G; ++g)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[g][e] = 0.f;
  int run_start = 0;
  // software pipeline: the loads of rows i + 1 and i + 2 are in flight while row i is accumulated
  uint4 u0[SCAT_G], u1[SCAT_G], u2[SCAT_G];
  auto load_row = [&](int i, uint4 (&dst)[SCAT_G]) {
    const uint32_t row = __shfl_sync(0xffffffffu, my_row, i & 31);
    const __nv_bfloat16* src = dH + static_cast<size_t>(row) * D + d0;
#pragma unroll
    for (int g = 0; g < SCAT_G; ++g) {
      const int col = g * 256 + static_cast<int>(lane) * 8;
      dst[g] = (i < cnt && col < dcols) ? __ldg(reinterpret_cast<const uint4*>(src + col)) : make_uint4(0, 0, 0, 0);
    }
  };
  load_row(0, u0);
  load_row(1, u1);
  for (int i = 0; i < cnt; ++i) {
    const uint32_t key = __shfl_sync(0xffffffffu, my_key, i);
    const uint32_t nkey = (i + 1 < cnt) ? __shfl_sync(0xffffffffu, my_key, (i + 1) & 31) : KEY_NONE;
    const int ur = __shfl_sync(0xffffffffu, my_ur, i);
    load_row(i + 2, u2);
#pragma unroll
    for (int g = 0; g < SCAT_G; ++g) {
      float f[8];
      bf16x8_to_f32(u0[g], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[g][e] += f[e];
    }
#pragma unroll
    for (int g = 0; g < SCAT_G; ++g) u0[g] = u1[g], u1[g] = u2[g];
    const bool end_of_run = (i == cnt - 1) || (nkey != key);
    if (end_of_run) {
      const bool open_left = (run_start == 0) && (prev_key == key);
      const bool open_right = (i == cnt - 1) && (next_key == key);
      if (!open_left && !open_right) {
        emit_run<MODE>(out, key, ur, V, D, d0, dcols, lane, acc);
      } else {
        const int slot = open_left ? 0 : 1;
        float* pr = partial + (static_cast<size_t>(wg) * 2 + slot) * D + d0;
#pragma unroll
        for (int g = 0; g < SCAT_G; ++g) {
          const int col = g * 256 + static_cast<int>(lane) * 8;
          if (col < dcols) {
            *reinterpret_cast<float4*>(pr + col) = make_float4(acc[g][0], acc[g][1], acc[g][2], acc[g][3]);
            *reinterpret_cast<float4*>(pr + col + 4) = make_float4(acc[g][4], acc[g][5], acc[g][6], acc[g][7]);
          }
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
      for (int g = 0; g < SCAT_G; ++g)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[g][e] = 0.f;
      run_start = i + 1;
    }
  }
  if (lane == 0 && blockIdx.y == 0) info[wg] = inf;
}

// One warp per range (and column slab) whose last run continues into the next range(s): sum the chain of partials
// in order and write that run's row.
template <int MODE>
__global__ void __launch_bounds__(128) scatter_fixup_kernel(const float* __restrict__ partial,
                                                            const ScatInfo* __restrict__ info, int nwarps, uint32_t V,
                                                            int D, ScatOut out) {
  const int wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (wg >= nwarps) return;
  const uint32_t key = info[wg].slot1_key;
  if (key == KEY_NONE) return;
  const int d0 = blockIdx.y * SCAT_COLS;
  const int dcols = min(SCAT_COLS, D - d0);
  float acc[SCAT_G][8];
  const float* pr = partial + (static_cast<size_t>(wg) * 2 + 1) * D + d0;
#pragma unroll
  for (int g = 0; g < SCAT_G; ++g) {
    const int col = g * 256 + static_cast<int>(lane) * 8;
    float4 a = make_float4(0, 0, 0, 0), b = a;
    if (col < dcols) {
      a = *reinterpret_cast<const float4*>(pr + col);
      b = *reinterpret_cast<const float4*>(pr + col + 4);
    }
    acc[g][0] = a.x, acc[g][1] = a.y, acc[g][2] = a.z, acc[g][3] = a.w;
    acc[g][4] = b.x, acc[g][5] = b.y, acc[g][6] = b.z, acc[g][7] = b.w;
  }
  for (int w = wg + 1; w < nwarps; ++w) {
    if (info[w].slot0_key != key) break;  // defensive: the chain must continue with the same id
    const float* q = partial + (static_cast<size_t>(w) * 2) * D + d0;
#pragma unroll
    for (int g = 0; g < SCAT_G; ++g) {
      const int col = g * 256 + static_cast<int>(lane) * 8;
      if (col < dcols) {
        const float4 a = *reinterpret_cast<const float4*>(q + col);
        const float4 b = *reinterpret_cast<const float4*>(q + col + 4);
        acc[g][0] += a.x, acc[g][1] += a.y, acc[g][2] += a.z, acc[g][3] += a.w;
        acc[g][4] += b.x, acc[g][5] += b.y, acc[g][6] += b.z, acc[g][7] += b.w;
      }
    }
    if (!info[w].slot0_open_right) break;
  }
  emit_run<MODE>(out, key, info[wg].slot1_urank, V, D, d0, dcols, lane, acc);
}

}  // namespace ssi
