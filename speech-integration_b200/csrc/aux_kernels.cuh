// Small HBM-bound helper kernels around the tensor-core stages: split-V log-sum-exp combine and loss
// reduction (fixed order, deterministic), fp32 -> bf16 cast of the dX accumulator, label shift + counts.
// (The tied-embedding gradient scatter-add lives in scatter.cuh.)
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "ptx.cuh"

namespace ssi {

// ---------------------------------------------------------------------------------------------------
// Row compaction: rows whose label is ignore_index contribute nothing to the loss or to any gradient, so the
// tensor-core stages only see the valid rows, gathered into a dense X_c.  The valid count lives on the device;
// the tile schedule that depends on it (DevPlan) is computed on the device too, so there is no host sync.
// ---------------------------------------------------------------------------------------------------
struct DevPlan {
  int n_valid;           // rows with label != ignore_index
  int tiles_m;           // ceil(n_valid / 128)
  int splits;            // K1: vocabulary splits
  int vtiles_per_split;  // K1: 256-column tiles per split
  int k_blocks_tok;      // K2b: K loop over tokens = tiles_m * 2 (64-row blocks)
  int part_stride;       // K1 partial arrays: floats per split = tiles_m * 128
  int row0;              // first compact row of this plan (0 for the whole-problem plan; chunk plans: chunk start)
  int first;             // chunk plans: 1 = first non-empty round (its dW contraction overwrites dW_acc, later ones add)
};
constexpr int MAX_ROW_CHUNKS = 120;  // plan area = (1 + MAX_ROW_CHUNKS) DevPlans (single-pass mode)
constexpr int PLAN_AREA_BYTES = 4096;
constexpr int MAX_LSE_ITEMS = 8192;  // cap on tiles_m * splits (bounds the partial arrays)

// K1 split count: smallest s whose static-schedule makespan (waves x tiles per split) is within 1% of the best
// `cg` = CTAs per work item (1, or 2 for CTA pairs owning two m-tiles); `sms / cg` items run concurrently.
__device__ inline void choose_splits(int tiles_m, int vtiles_total, int sms, int cg, int& splits, int& per_split) {
  int best_s = 1;
  const int units_m = (tiles_m + cg - 1) / cg;
  sms = max(sms / cg, 1);
  if (tiles_m > 0) {
    const int max_s = min(vtiles_total, 148);
    long long best = -1;
    for (int pass = 0; pass < 2; ++pass) {
      for (int s2 = 1; s2 <= max_s; ++s2) {
        if (s2 > 1 && static_cast<long long>(s2) * tiles_m > MAX_LSE_ITEMS) break;
        const int pr = (vtiles_total + s2 - 1) / s2;
        if ((vtiles_total + pr - 1) / pr != s2) continue;
        const long long items = static_cast<long long>(units_m) * s2;
        const long long c = ((items + sms - 1) / sms) * pr;
        if (pass == 0) {
          if (best < 0 || c < best) best = c;
        } else if (c * 100 <= best * 101) {
          best_s = s2;
          break;
        }
      }
    }
  }
  splits = best_s;
  per_split = (vtiles_total + best_s - 1) / best_s;
}

// One block: stable compaction of the valid rows. pos[r] = compact position or -1, idx[c] = original row,
// labels_c[c] = label (ignore_index beyond n_valid up to n_pad), then the device-side plan.
__global__ void __launch_bounds__(1024) compact_scan_kernel(const int64_t* __restrict__ labels, int n, int n_pad,
                                                            long long ignore_index, int* __restrict__ pos,
                                                            int* __restrict__ idx, int64_t* __restrict__ labels_c,
                                                            DevPlan* __restrict__ plan, int vtiles_total, int sms,
                                                            int chunk_t_max = 0, int chunks_max = 0,
                                                            int d_tiles = 0, int64_t* n_valid_out = nullptr,
                                                            int cg = 1, int half_tail = 0) {
  // Every warp owns one contiguous segment of the rows (a multiple of 32) and walks it 32 rows at a time with
  // coalesced loads, four of them in flight: positions come from ballots, so the compaction stays stable.
  __shared__ int s_wsum[32];
  __shared__ int s_chunk_t;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  const int wseg = ((n + 31) / 32 + 31) / 32 * 32;
  const int b = min(warp * wseg, n);
  const int e = min(b + wseg, n);
  int a = 0;
  for (int i0 = b; i0 < e; i0 += 128) {
    long long l[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 32 + lane;
      l[u] = (i < e) ? labels[i] : ignore_index;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) a += (l[u] != ignore_index);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  if (lane == 0) s_wsum[warp] = a;
  __syncthreads();
  int run = 0, n_valid = 0;
  for (int w = 0; w < 32; ++w) {
    const int c = s_wsum[w];
    if (w < warp) run += c;
    n_valid += c;
  }
  for (int i0 = b; i0 < e; i0 += 128) {
    long long l[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 32 + lane;
      l[u] = (i < e) ? labels[i] : ignore_index;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 32 + lane;
      const bool ok = l[u] != ignore_index;
      const unsigned m = __ballot_sync(0xffffffffu, ok);
      if (i < e) {
        if (ok) {
          const int r = run + __popc(m & lt_mask);
          pos[i] = r;
          idx[r] = i;
          labels_c[r] = l[u];
        } else {
          pos[i] = -1;
        }
      }
      run += __popc(m);
    }
  }
  for (int i = n_valid + threadIdx.x; i < n_pad; i += 1024) labels_c[i] = ignore_index;
  if (threadIdx.x == 0) {
    DevPlan d;
    d.n_valid = n_valid;
    d.tiles_m = (n_valid + 127) / 128;
    d.k_blocks_tok = d.tiles_m * 2;
    d.part_stride = d.tiles_m * 128;
    choose_splits(d.tiles_m, vtiles_total, sms, cg, d.splits, d.vtiles_per_split);
    d.row0 = d.first = 0;
    *plan = d;
    if (n_valid_out) *n_valid_out = n_valid;
    // Single-pass mode: m-tiles per row chunk.  The dX contraction of a chunk has t * d_tiles output tiles on
    // `sms` persistent CTAs, so t is chosen to waste the fewest waves; a per-chunk charge prefers fewer chunks.
    int best_t = max(chunk_t_max, 1);
    if (chunks_max > 0 && d.tiles_m > 0) {
      float best_cost = -1.f;
      for (int t = 1; t <= chunk_t_max; ++t) {
        const int full = d.tiles_m / t, rem = d.tiles_m % t;
        const int nch = full + (rem ? 1 : 0);
        if (nch > chunks_max) continue;
        const int wk = max(sms / cg, 1);  // concurrent work items (CTAs or CTA pairs)
        // waves of the dX contraction; with tail splitting an at most half-full last wave costs half a wave
        auto waves = [&](int items) {
          const int tail = items % wk;
          return static_cast<float>(items / wk) + (tail == 0 ? 0.f : (half_tail && 2 * tail <= wk) ? 0.5f : 1.f);
        };
        float cost = static_cast<float>(full) * waves(((t + cg - 1) / cg) * d_tiles) + 0.3f * nch;
        if (rem) cost += waves(((rem + cg - 1) / cg) * d_tiles);
        if (best_cost < 0.f || cost <= best_cost) best_cost = cost, best_t = t;
      }
    }
    s_chunk_t = best_t;
  }
  if (chunks_max > 0) {
    __syncthreads();
    const int t = max(s_chunk_t, 1);
    const int tiles_all = (n_valid + 127) / 128;
    // The host launches `chunks_max` rounds; the real chunks occupy the LAST rounds (the empty ones come first),
    // so that the final round's dW contraction is always the last work on dW_acc (the caller's "dW ready" event
    // is recorded right after it).  With no valid row at all the final round still zero-fills dW_acc.
    const int nch = max((tiles_all + t - 1) / t, 1);
    const int lead = max(chunks_max - nch, 0);
    for (int c = threadIdx.x; c < chunks_max; c += 1024) {
      DevPlan d;
      const int cc = c - lead;
      const int m0 = (cc < 0) ? tiles_all : min(cc * t, tiles_all);
      d.tiles_m = min(t, tiles_all - m0);
      d.row0 = (cc < 0) ? 0 : m0 * 128;
      d.n_valid = max(0, min(n_valid - m0 * 128, d.tiles_m * 128));
      d.k_blocks_tok = d.tiles_m * 2;
      d.part_stride = d.tiles_m * 128;
      choose_splits(d.tiles_m, vtiles_total, sms, cg, d.splits, d.vtiles_per_split);
      d.first = (cc == 0) ? 1 : 0;
      plan[1 + c] = d;
    }
  }
}

// X_c[c, :] = X[idx[c], :] for c < n_valid, zeros for the padding rows of the last 128-row tile; optionally
// lse_c[c] = lse[idx[c]].  One warp per compact row, 16-byte accesses.
__global__ void __launch_bounds__(256) gather_rows_kernel(const uint4* __restrict__ X, const int* __restrict__ idx,
                                                          const DevPlan* __restrict__ plan, int row_u4,
                                                          uint4* __restrict__ Xc, const float* __restrict__ lse,
                                                          float* __restrict__ lse_c) {
  const int n_valid = plan->n_valid;
  const int rows = plan->tiles_m * 128;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < rows; c += warps) {
    uint4* dst = Xc + static_cast<size_t>(c) * row_u4;
    if (c < n_valid) {
      const int r = idx[c];
      const uint4* src = X + static_cast<size_t>(r) * row_u4;
      for (int j = lane; j < row_u4; j += 32) dst[j] = __ldg(src + j);
      if (lse_c && lane == 0) lse_c[c] = lse[r];
    } else {
      for (int j = lane; j < row_u4; j += 32) dst[j] = make_uint4(0, 0, 0, 0);
      if (lse_c && lane == 0) lse_c[c] = 0.f;
    }
  }
}

// dX[r, :] = bf16(dXacc_c[pos[r], :]) for valid rows, zeros for ignored rows. One warp per original row.
__global__ void __launch_bounds__(256) dx_finalize_kernel(const float4* __restrict__ acc, const int* __restrict__ pos,
                                                          int n, int row_f4, uint2* __restrict__ dX) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n; r += warps) {
    const int c = pos[r];
    uint2* dst = dX + static_cast<size_t>(r) * row_f4;
    if (c >= 0) {
      const float4* src = acc + static_cast<size_t>(c) * row_f4;
      for (int j = lane; j < row_f4; j += 32) {
        const float4 v = src[j];
        uint2 o;
        o.x = pack_bf16x2(v.x, v.y);
        o.y = pack_bf16x2(v.z, v.w);
        dst[j] = o;
      }
    } else {
      for (int j = lane; j < row_f4; j += 32) dst[j] = make_uint2(0, 0);
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Forward tail: combine the per-split (max, sum, label-logit) partials of each row in split order.
// ---------------------------------------------------------------------------------------------------
constexpr int COMBINE_THREADS = 256;

__global__ void __launch_bounds__(COMBINE_THREADS) lse_combine_kernel(
    const float* __restrict__ part_m, const float* __restrict__ part_l, const float* __restrict__ part_z,
    const DevPlan* __restrict__ plan, const int* __restrict__ pos, int n_rows, float* __restrict__ lse,
    float* __restrict__ blk_loss, int* __restrict__ blk_cnt) {
  __shared__ float s_loss[COMBINE_THREADS];
  __shared__ int s_cnt[COMBINE_THREADS];
  const int row = blockIdx.x * COMBINE_THREADS + threadIdx.x;  // ORIGINAL row
  const int splits = plan->splits, part_stride = plan->part_stride;
  float loss = 0.f;
  int cnt = 0;
  if (row < n_rows) {
    const int c = pos[row];  // compact position, -1 for ignored rows
    float l = 0.f;
    if (c >= 0) {
      float M = -INFINITY;
      for (int s = 0; s < splits; ++s) M = fmaxf(M, part_m[static_cast<size_t>(s) * part_stride + c]);
      float L = 0.f, z = -INFINITY;
      for (int s = 0; s < splits; ++s) {
        const size_t o = static_cast<size_t>(s) * part_stride + c;
        L += part_l[o] * exp2f(part_m[o] - M);
        z = fmaxf(z, part_z[o]);
      }
      l = (M + log2f(L)) * 0.6931471805599453f;
      loss = l - z;
      cnt = 1;
    }
    lse[row] = l;
  }
  s_loss[threadIdx.x] = loss;
  s_cnt[threadIdx.x] = cnt;
  __syncthreads();
  for (int off = COMBINE_THREADS / 2; off > 0; off >>= 1) {
    if (threadIdx.x < off) {
      s_loss[threadIdx.x] += s_loss[threadIdx.x + off];
      s_cnt[threadIdx.x] += s_cnt[threadIdx.x + off];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    blk_loss[blockIdx.x] = s_loss[0];
    blk_cnt[blockIdx.x] = s_cnt[0];
  }
}

__global__ void __launch_bounds__(256) loss_finalize_kernel(const float* __restrict__ blk_loss,
                                                            const int* __restrict__ blk_cnt, int nblk,
                                                            float* __restrict__ loss_out,
                                                            int64_t* __restrict__ n_valid) {
  __shared__ double s_loss[256];
  __shared__ long long s_cnt[256];
  double a = 0.0;
  long long c = 0;
  for (int i = threadIdx.x; i < nblk; i += 256) {
    a += static_cast<double>(blk_loss[i]);
    c += blk_cnt[i];
  }
  s_loss[threadIdx.x] = a;
  s_cnt[threadIdx.x] = c;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if (threadIdx.x < off) {
      s_loss[threadIdx.x] += s_loss[threadIdx.x + off];
      s_cnt[threadIdx.x] += s_cnt[threadIdx.x + off];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double sum = s_loss[0];
    const long long n = s_cnt[0];
    loss_out[0] = static_cast<float>(sum);
    // 0/0 -> NaN exactly like total_loss / total_elements in ce_chunked_output_loss.py:83
    loss_out[1] = static_cast<float>(sum) / static_cast<float>(n);
    n_valid[0] = n;
  }
}

// ---------------------------------------------------------------------------------------------------
// Single-pass ("stash") mode helpers.  Per row chunk: K1 stores the bf16 logits tile block S and the split
// partials; chunk_combine_kernel folds the partials into lse_c / loss_c for the chunk's compact rows;
// softmax_grad_kernel rewrites S in place into dlogits G (HBM-bound, 2 B read + 2 B written per logit).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) chunk_combine_kernel(const float* __restrict__ part_m,
                                                            const float* __restrict__ part_l,
                                                            const float* __restrict__ part_z,
                                                            const DevPlan* __restrict__ plan,
                                                            float* __restrict__ lse_c, float* __restrict__ loss_c) {
  const int rows = plan->tiles_m * 128, nv = plan->n_valid, row0 = plan->row0;
  const int splits = plan->splits, part_stride = plan->part_stride;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += gridDim.x * blockDim.x) {
    float l = 0.f, loss = 0.f;
    if (r < nv) {
      float M = -INFINITY;
      for (int s = 0; s < splits; ++s) M = fmaxf(M, part_m[static_cast<size_t>(s) * part_stride + r]);
      float L = 0.f, z = -INFINITY;
      for (int s = 0; s < splits; ++s) {
        const size_t o = static_cast<size_t>(s) * part_stride + r;
        L += part_l[o] * exp2f(part_m[o] - M);
        z = fmaxf(z, part_z[o]);
      }
      l = (M + log2f(L)) * 0.6931471805599453f;
      loss = l - z;
    }
    lse_c[row0 + r] = l;
    loss_c[row0 + r] = loss;
  }
}

__device__ __forceinline__ float bf16lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

// G[r, v] = (exp(S[r, v] - lse[r]) - [v == label[r]]) * grad_hint / n_valid_total, bf16, in place; padding
// columns (v >= V) and padding rows (r >= n_valid of the chunk) become exact zeros.  One block per row at a
// time, 16-byte accesses, 4 loads in flight per thread.
__global__ void __launch_bounds__(256) softmax_grad_kernel(uint4* __restrict__ S, int row_u4,
                                                           const DevPlan* __restrict__ plan,
                                                           const float* __restrict__ lse_c,
                                                           const int64_t* __restrict__ labels_c, int V,
                                                           const float* __restrict__ grad_hint,
                                                           const int64_t* __restrict__ n_valid_total) {
  const int rows = plan->tiles_m * 128, nv = plan->n_valid, row0 = plan->row0;
  const long long nvt = *n_valid_total;
  const float gscale = (nvt > 0) ? (*grad_hint / static_cast<float>(nvt)) : 0.f;
  constexpr float kLog2e = 1.4426950408889634f;
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    uint4* row = S + static_cast<size_t>(r) * row_u4;
    if (r >= nv) {
      for (int j = threadIdx.x; j < row_u4; j += 256) row[j] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const float lse2 = lse_c[row0 + r] * kLog2e;
    const int label = static_cast<int>(labels_c[row0 + r]);
    for (int j0 = threadIdx.x; j0 < row_u4; j0 += 4 * 256) {
      uint4 u[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int j = j0 + k * 256;
        if (j < row_u4) u[k] = row[j];
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int j = j0 + k * 256;
        if (j >= row_u4) continue;
        const int c0 = j * 8;
        uint32_t w[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
        float g[8];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          g[2 * e] = fast_exp2(fmaf(bf16lo(w[e]), kLog2e, -lse2)) * gscale;
          g[2 * e + 1] = fast_exp2(fmaf(bf16hi(w[e]), kLog2e, -lse2)) * gscale;
        }
        if (label >= c0 && label < c0 + 8) {
#pragma unroll
          for (int e = 0; e < 8; ++e)
            if (c0 + e == label) g[e] -= gscale;
        }
        if (c0 + 8 > V) {
#pragma unroll
          for (int e = 0; e < 8; ++e)
            if (c0 + e >= V) g[e] = 0.f;
        }
        uint4 o;
        o.x = pack_bf16x2(g[0], g[1]);
        o.y = pack_bf16x2(g[2], g[3]);
        o.z = pack_bf16x2(g[4], g[5]);
        o.w = pack_bf16x2(g[6], g[7]);
        row[j] = o;
      }
    }
  }
}

__device__ __forceinline__ void bf16x8_to_f32(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}

__device__ __forceinline__ uint4 f32x8_to_bf16(const float (&f)[8]) {
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]);
  o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]);
  o.w = pack_bf16x2(f[6], f[7]);
  return o;
}

// ---------------------------------------------------------------------------------------------------
// Exp-stash form of single-pass mode (default).  K1 left E[r, v] = bf16(2^(s log2e - ref2[r])) in the block and the
// per-split row sums of the fp32 values; softmax[r, v] = E[r, v] / L[r] with L = sum over splits, so with
// g = hint / n_valid:   dlogits = g * softmax - g * onehot = rowscale[r] * E[r, :] - g * onehot,  rowscale = g / L.
// The two terms are kept apart.  The softmax term goes through the tensor cores as E times a per-ROW factor:
//   exp_combine_kernel   lse = (ref2 + log2 L) ln2, loss = lse - z_label, rowscale[r] = g / L[r]
//   scale_rows_kernel    Xs[r, :] = bf16(rowscale[r] * X_c[r, :]): the rows dW = E^T Xs contracts with
//   dX contraction       rowscale[r] * (E W)[r, :] in its epilogue.
// The LABEL column of every row -- softmax and onehot together, g (p_label - 1), the dominant and for a confident
// row nearly cancelling part of the gradient -- never goes through a bf16 matrix element: K1 stashes 0 there (the
// label's exponential still counts in L), exp_combine_kernel forms labw[r] = 1 - p_label = (L - e_label) / L in fp32,
// and the term is exact arithmetic on the operands:
//   dX[r, :] -= g * labw[r] * W[label_r, :]   while the compact rows go back to their positions
//                                          (dx_scatter_bf16_kernel, or dx_label_term_kernel in place before the fused
//                                          norm backward)
//   dW_acc[v, :] -= g * sum_{label_r = v} labw[r] * X_c[r, :]   by the sorted fp32 segmented reduction of scatter.cuh,
//                                          keyed by the labels (no atomics: bit-identical run to run).
// No pass over the [rows, V] block.  A row whose L is not finite (a logit more than ~88 nats above the pilot's
// reference: E overflowed) raises bit 1 of the status word that ssi_b200_single_pass_lockstep_status reads; its loss
// is inf and its gradients are NaN, never silently wrong.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ float grad_scale(const float* grad_hint, const int64_t* n_valid_total) {
  const long long nvt = *n_valid_total;
  return (nvt > 0) ? (*grad_hint / static_cast<float>(nvt)) : 0.f;
}

__global__ void __launch_bounds__(256) exp_combine_kernel(const float* __restrict__ ref2,
                                                          const float* __restrict__ part_l,
                                                          const float* __restrict__ part_z,
                                                          const DevPlan* __restrict__ plan,
                                                          const float* __restrict__ grad_hint,
                                                          const int64_t* __restrict__ n_valid_total,
                                                          float* __restrict__ lse_c, float* __restrict__ loss_c,
                                                          float* __restrict__ rowscale_c,
                                                          float* __restrict__ labw_c,
                                                          unsigned int* __restrict__ overflow_word) {
  const int rows = plan->tiles_m * 128, nv = plan->n_valid, row0 = plan->row0;
  const int splits = plan->splits, part_stride = plan->part_stride;
  const float gscale = grad_scale(grad_hint, n_valid_total);
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += gridDim.x * blockDim.x) {
    float l = 0.f, loss = 0.f, rs = 0.f, lw = 0.f;
    if (r < nv) {
      float L = 0.f, z = -INFINITY;
      for (int s = 0; s < splits; ++s) {  // fixed order: deterministic
        const size_t o = static_cast<size_t>(s) * part_stride + r;
        L += part_l[o];
        z = fmaxf(z, part_z[o]);
      }
      l = (ref2[r] + log2f(L)) * 0.6931471805599453f;
      loss = l - z;
      rs = gscale / L;
      // the label's exponential exactly as K1 formed it (same instruction, same operands), so L - e_lab is the sum
      // of the OTHER columns up to the rounding of the additions: 1 - p_label without cancellation against 1
      const float e_lab = (z > -INFINITY) ? fast_exp2(fmaf(z, 1.4426950408889634f, -ref2[r])) : 0.f;
      lw = fminf(fmaxf((L - e_lab) / L, 0.f), 1.f);
      if (!(L > 0.f && L < INFINITY)) {
        lw = NAN;  // E overflowed (or every exponential underflowed): poison instead of a plausible-looking gradient
        atomicOr(overflow_word, 1u);
      }
    }
    lse_c[row0 + r] = l;
    loss_c[row0 + r] = loss;
    rowscale_c[row0 + r] = rs;
    labw_c[row0 + r] = lw;
  }
}

// Xs[row0 + r, :] = bf16(rowscale[row0 + r] * X_c[row0 + r, :]) for the chunk's rows (padding rows: scale 0 -> zeros).
__global__ void __launch_bounds__(256) scale_rows_kernel(const uint4* __restrict__ Xc,
                                                         const float* __restrict__ rowscale_c,
                                                         const DevPlan* __restrict__ plan, int row_u4,
                                                         uint4* __restrict__ Xs) {
  const int rows = plan->tiles_m * 128, row0 = plan->row0;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows; r += warps) {
    const float rs = rowscale_c[row0 + r];
    const uint4* src = Xc + static_cast<size_t>(row0 + r) * row_u4;
    uint4* dst = Xs + static_cast<size_t>(row0 + r) * row_u4;
    for (int j = lane; j < row_u4; j += 32) {
      const uint4 u = src[j];
      uint4 o;
      o.x = pack_bf16x2(bf16lo(u.x) * rs, bf16hi(u.x) * rs);
      o.y = pack_bf16x2(bf16lo(u.y) * rs, bf16hi(u.y) * rs);
      o.z = pack_bf16x2(bf16lo(u.z) * rs, bf16hi(u.z) * rs);
      o.w = pack_bf16x2(bf16lo(u.w) * rs, bf16hi(u.w) * rs);
      dst[j] = o;
    }
  }
}

// lse[row] / per-block loss partials over ORIGINAL rows from the compact lse_c / loss_c (fixed order).
__global__ void __launch_bounds__(COMBINE_THREADS) stash_finish_kernel(const float* __restrict__ lse_c,
                                                                       const float* __restrict__ loss_c,
                                                                       const int* __restrict__ pos, int n_rows,
                                                                       float* __restrict__ lse,
                                                                       float* __restrict__ blk_loss,
                                                                       int* __restrict__ blk_cnt) {
  __shared__ float s_loss[COMBINE_THREADS];
  __shared__ int s_cnt[COMBINE_THREADS];
  const int row = blockIdx.x * COMBINE_THREADS + threadIdx.x;
  float loss = 0.f;
  int cnt = 0;
  if (row < n_rows) {
    const int c = pos[row];
    float l = 0.f;
    if (c >= 0) {
      l = lse_c[c];
      loss = loss_c[c];
      cnt = 1;
    }
    lse[row] = l;
  }
  s_loss[threadIdx.x] = loss;
  s_cnt[threadIdx.x] = cnt;
  __syncthreads();
  for (int off = COMBINE_THREADS / 2; off > 0; off >>= 1) {
    if (threadIdx.x < off) {
      s_loss[threadIdx.x] += s_loss[threadIdx.x + off];
      s_cnt[threadIdx.x] += s_cnt[threadIdx.x + off];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    blk_loss[blockIdx.x] = s_loss[0];
    blk_cnt[blockIdx.x] = s_cnt[0];
  }
}

// out = bf16(a - g * w), eight bf16 values at a time
__device__ __forceinline__ uint4 bf16x8_sub_scaled(const uint4& a, const uint4& w, float g) {
  float fa[8], fw[8];
  bf16x8_to_f32(a, fa);
  bf16x8_to_f32(w, fw);
#pragma unroll
  for (int e = 0; e < 8; ++e) fa[e] = fmaf(-g, fw[e], fa[e]);
  return f32x8_to_bf16(fa);
}

// dX[r, :] = dXc[pos[r], :] for valid rows, zeros for ignored rows (bf16 rows, 16-byte accesses).  With W (exp-stash
// form): dX[r, :] = dXc[pos[r], :] - g * labw[pos[r]] * W[label, :], the label column's term of the row.
__global__ void __launch_bounds__(256) dx_scatter_bf16_kernel(const uint4* __restrict__ dXc,
                                                              const int* __restrict__ pos, int n, int row_u4,
                                                              uint4* __restrict__ dX,
                                                              const uint4* __restrict__ W,
                                                              const int64_t* __restrict__ labels_c, int V,
                                                              const float* __restrict__ grad_hint,
                                                              const int64_t* __restrict__ n_valid_total,
                                                              const float* __restrict__ labw_c) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float g = W ? grad_scale(grad_hint, n_valid_total) : 0.f;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n; r += warps) {
    const int c = pos[r];
    uint4* dst = dX + static_cast<size_t>(r) * row_u4;
    if (c >= 0) {
      const uint4* src = dXc + static_cast<size_t>(c) * row_u4;
      const long long lab = W ? labels_c[c] : -1;
      if (lab >= 0 && lab < V) {
        const uint4* wrow = W + static_cast<size_t>(lab) * row_u4;
        const float gl = g * labw_c[c];
        for (int j = lane; j < row_u4; j += 32) dst[j] = bf16x8_sub_scaled(src[j], __ldg(wrow + j), gl);
      } else {
        for (int j = lane; j < row_u4; j += 32) dst[j] = src[j];
      }
    } else {
      for (int j = lane; j < row_u4; j += 32) dst[j] = make_uint4(0, 0, 0, 0);
    }
  }
}

// In place on the compact rows: dXc[c, :] -= g * labw[c] * W[labels_c[c], :] (exp-stash form ahead of the fused norm
// backward).
__global__ void __launch_bounds__(256) dx_label_term_kernel(uint4* __restrict__ dXc,
                                                            const int64_t* __restrict__ labels_c, int n_pad,
                                                            int row_u4, const uint4* __restrict__ W, int V,
                                                            const float* __restrict__ grad_hint,
                                                            const int64_t* __restrict__ n_valid_total,
                                                            const float* __restrict__ labw_c) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float g = grad_scale(grad_hint, n_valid_total);
  for (int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < n_pad; c += warps) {
    const long long lab = labels_c[c];
    if (lab < 0 || lab >= V) continue;
    uint4* row = dXc + static_cast<size_t>(c) * row_u4;
    const uint4* wrow = W + static_cast<size_t>(lab) * row_u4;
    const float gl = g * labw_c[c];
    for (int j = lane; j < row_u4; j += 32) row[j] = bf16x8_sub_scaled(row[j], __ldg(wrow + j), gl);
  }
}

// Backward of single-pass mode: the gradients were formed in forward for an upstream gradient `grad_hint`;
// ratio = grad_out / grad_hint rescales them (exactly 1 when the caller announced the right hint).
__device__ __forceinline__ float grad_ratio(const float* grad_out, const float* grad_hint) {
  const float g = *grad_out, h = *grad_hint;
  return (g == h) ? 1.f : g / h;
}

__global__ void __launch_bounds__(256) apply_dx_kernel(uint4* __restrict__ dX, size_t n_u4,
                                                       const float* __restrict__ grad_out,
                                                       const float* __restrict__ grad_hint) {
  const float ratio = grad_ratio(grad_out, grad_hint);
  if (ratio == 1.f) return;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n_u4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const uint4 u = dX[i];
    uint4 o;
    o.x = pack_bf16x2(bf16lo(u.x) * ratio, bf16hi(u.x) * ratio);
    o.y = pack_bf16x2(bf16lo(u.y) * ratio, bf16hi(u.y) * ratio);
    o.z = pack_bf16x2(bf16lo(u.z) * ratio, bf16hi(u.z) * ratio);
    o.w = pack_bf16x2(bf16lo(u.w) * ratio, bf16hi(u.w) * ratio);
    dX[i] = o;
  }
}

// Fold of the trainer's scale_grads + grad-norm passes into the kernels that write the tied gradient anyway
// (tt/torchtune/training/_grad_scaler.py:25-31, ssi/trainer.py:404-408).  post_scale: the scaler of
// scale_grads, applied in fp32 BEFORE the one rounding to bf16 (null = 1).  sq: where a kernel leaves its share of
// sum(g^2) -- dense writers one partial per block (sum of the new values), sparse read-modify-write sites one
// partial per warp (sum of new^2 - old^2); grad_fold_finalize_kernel adds everything up in double, in order.
struct GradFold {
  const float* post_scale;
  float* sq;
};
__device__ __forceinline__ float fold_scale(const GradFold& f) { return f.post_scale ? *f.post_scale : 1.f; }
__device__ __forceinline__ float sq_of_bf16x2(uint32_t w) {
  const float a = __uint_as_float(w << 16), b = __uint_as_float(w & 0xFFFF0000u);
  return fmaf(a, a, b * b);
}
// block sum of `v` -> sq[blockIdx.x] (256-thread blocks; s_w: 8 floats of shared memory)
__device__ __forceinline__ void fold_block_partial(float v, float* sq, float* s_w) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
    for (int w = 0; w < static_cast<int>(blockDim.x >> 5); ++w) a += s_w[w];
    sq[blockIdx.x] = a;
  }
}

// dW = bf16(post_scale * ((accumulate ? dW : 0) + ratio * acc)): the fp32 accumulator over ALL tokens is rounded once.
__global__ void __launch_bounds__(256) apply_dw_kernel(const float4* __restrict__ acc, size_t n_f4,
                                                       const float* __restrict__ grad_out,
                                                       const float* __restrict__ grad_hint, int accumulate,
                                                       uint2* __restrict__ dW, GradFold fold) {
  __shared__ float s_w[8];
  const float ratio = grad_ratio(grad_out, grad_hint);
  const float ps = fold_scale(fold);
  float sq = 0.f;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n_f4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float4 a = acc[i];
    a.x *= ratio, a.y *= ratio, a.z *= ratio, a.w *= ratio;
    if (accumulate) {
      const uint2 old = dW[i];
      a.x += bf16lo(old.x), a.y += bf16hi(old.x), a.z += bf16lo(old.y), a.w += bf16hi(old.y);
    }
    if (fold.post_scale) a.x *= ps, a.y *= ps, a.z *= ps, a.w *= ps;
    uint2 o;
    o.x = pack_bf16x2(a.x, a.y);
    o.y = pack_bf16x2(a.z, a.w);
    dW[i] = o;
    if (fold.sq) sq += sq_of_bf16x2(o.x) + sq_of_bf16x2(o.y);
  }
  if (fold.sq) fold_block_partial(sq, fold.sq, s_w);
}

// sum of the fold partials (fixed order, double) -> out[0]
__global__ void __launch_bounds__(256) grad_fold_finalize_kernel(const float* __restrict__ partial, int n,
                                                                 float* __restrict__ out) {
  __shared__ double s[256];
  double a = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) a += static_cast<double>(partial[i]);
  s[threadIdx.x] = a;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if (threadIdx.x < off) s[threadIdx.x] += s[threadIdx.x + off];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = static_cast<float>(s[0] > 0.0 ? s[0] : 0.0);
}

// ---------------------------------------------------------------------------------------------------
// Label shift (ssi/loss.py:16) + valid-token counts (ssi/trainer.py:391, ce_chunked_output_loss.py:66)
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) label_shift_count_kernel(const int64_t* __restrict__ labels, long long B,
                                                                long long S, long long ignore_index,
                                                                int64_t* __restrict__ out,
                                                                unsigned long long* __restrict__ counts) {
  __shared__ unsigned int s_c[2];
  if (threadIdx.x < 2) s_c[threadIdx.x] = 0;
  __syncthreads();
  const long long total = B * S;
  unsigned int c0 = 0, c1 = 0;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long s = i % S;
    const long long cur = labels[i];
    const long long nxt = (s + 1 < S) ? labels[i + 1] : ignore_index;
    out[i] = nxt;
    c0 += (cur != ignore_index);
    c1 += (nxt != ignore_index);
  }
  atomicAdd(&s_c[0], c0);
  atomicAdd(&s_c[1], c1);
  __syncthreads();
  if (threadIdx.x < 2 && s_c[threadIdx.x]) atomicAdd(&counts[threadIdx.x], static_cast<unsigned long long>(s_c[threadIdx.x]));
}

// counts[0] = counts[1] = #(labels != ignore_index) over a flat label array (labels that are already shifted: the
// data-parallel loss module posts this as its local count when the trainer did not announce one)
__global__ void __launch_bounds__(256) count_valid_kernel(const int64_t* __restrict__ labels, long long n,
                                                          long long ignore_index,
                                                          unsigned long long* __restrict__ counts) {
  unsigned int c = 0;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    c += (labels[i] != ignore_index);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0 && c) {
    atomicAdd(&counts[0], static_cast<unsigned long long>(c));
    atomicAdd(&counts[1], static_cast<unsigned long long>(c));
  }
}

}  // namespace ssi
