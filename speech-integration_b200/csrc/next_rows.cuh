// HBM-bound kernels of the rows SURVEY.md 8(f) marks "next": the callers and data formats either side of the
// fused LM-head + CE path.
//   f1  final RMSNorm forward / backward            tt/torchtune/modules/rms_norm.py:38-43, transformer.py:653
//   f2  label shift + token counts + token-type histogram in one pass
//                                                   ssi/loss.py:16, ssi/trainer.py:388-391, ssi/train_utils.py:150-165
//   f3  scale_grads + squared grad-norm of the tied weight in one pass
//                                                   tt/torchtune/training/_grad_scaler.py:25-31, ssi/trainer.py:404-408
//   f4  forward embedding gather                    tt/torchtune/modules/transformer.py:628
// All reductions use fixed-order partials (no floating-point atomics): results are run-to-run deterministic.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "aux_kernels.cuh"

namespace ssi {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------------------------------------------
// f1: y = bf16((float(x) * rsqrt(mean(float(x)^2) + eps)) * float(scale)), one warp per row, 16-byte accesses.
// rstd [N] fp32 is kept for backward.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rmsnorm_fwd_kernel(const uint4* __restrict__ x, const uint4* __restrict__ scale,
                                                          float eps, int n, int row_u4, uint4* __restrict__ y,
                                                          float* __restrict__ rstd) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float inv_d = 1.f / static_cast<float>(row_u4 * 8);
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n; r += warps) {
    const uint4* src = x + static_cast<size_t>(r) * row_u4;
    float ss = 0.f;
    for (int j = lane; j < row_u4; j += 32) {
      float f[8];
      bf16x8_to_f32(src[j], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(f[e], f[e], ss);
    }
    ss = warp_sum(ss);
    const float rs = rsqrtf(ss * inv_d + eps);
    if (lane == 0 && rstd) rstd[r] = rs;
    uint4* dst = y + static_cast<size_t>(r) * row_u4;
    for (int j = lane; j < row_u4; j += 32) {
      float f[8], g[8];
      bf16x8_to_f32(src[j], f);  // second read of the row hits L1 / L2
      bf16x8_to_f32(__ldg(scale + j), g);
      uint4 o;
      o.x = pack_bf16x2((f[0] * rs) * g[0], (f[1] * rs) * g[1]);
      o.y = pack_bf16x2((f[2] * rs) * g[2], (f[3] * rs) * g[3]);
      o.z = pack_bf16x2((f[4] * rs) * g[4], (f[5] * rs) * g[5]);
      o.w = pack_bf16x2((f[6] * rs) * g[6], (f[7] * rs) * g[7]);
      dst[j] = o;
    }
  }
}

// Backward in fp32: g = dy * scale, c = mean(g * x), dx = rstd * (g - x * rstd^2 * c); dscale[j] = sum_rows dy * x
// * rstd.  Each warp keeps its dscale partial for the columns it owns in registers over all its rows and writes
// ONE partial row; rmsnorm_dscale_reduce_kernel sums the partial rows in order.  MAXJ = ceil(row_u4 / 32).
template <int MAXJ>
__global__ void __launch_bounds__(256) rmsnorm_bwd_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ x,
                                                          const uint4* __restrict__ scale,
                                                          const float* __restrict__ rstd, int n, int row_u4,
                                                          uint4* __restrict__ dx, float* __restrict__ partial) {
  const int lane = threadIdx.x & 31;
  const int warp_g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float inv_d = 1.f / static_cast<float>(row_u4 * 8);
  float ds[MAXJ][8];
#pragma unroll
  for (int k = 0; k < MAXJ; ++k)
#pragma unroll
    for (int e = 0; e < 8; ++e) ds[k][e] = 0.f;
  for (int r = warp_g; r < n; r += warps) {
    const uint4* py = dy + static_cast<size_t>(r) * row_u4;
    const uint4* px = x + static_cast<size_t>(r) * row_u4;
    const float rs = rstd[r];
    float c = 0.f;
#pragma unroll
    for (int k = 0; k < MAXJ; ++k) {
      const int j = lane + 32 * k;
      if (j < row_u4) {
        float fy[8], fx[8], g[8];
        bf16x8_to_f32(py[j], fy);
        bf16x8_to_f32(px[j], fx);
        bf16x8_to_f32(__ldg(scale + j), g);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          c = fmaf(fy[e] * g[e], fx[e], c);
          ds[k][e] = fmaf(fy[e] * fx[e], rs, ds[k][e]);
        }
      }
    }
    c = warp_sum(c) * inv_d;
    const float a = rs * rs * c;
    uint4* pd = dx + static_cast<size_t>(r) * row_u4;
#pragma unroll
    for (int k = 0; k < MAXJ; ++k) {
      const int j = lane + 32 * k;
      if (j < row_u4) {
        float fy[8], fx[8], g[8];
        bf16x8_to_f32(py[j], fy);  // L1 / L2 hits
        bf16x8_to_f32(px[j], fx);
        bf16x8_to_f32(__ldg(scale + j), g);
        uint4 o;
        o.x = pack_bf16x2(rs * (fy[0] * g[0] - fx[0] * a), rs * (fy[1] * g[1] - fx[1] * a));
        o.y = pack_bf16x2(rs * (fy[2] * g[2] - fx[2] * a), rs * (fy[3] * g[3] - fx[3] * a));
        o.z = pack_bf16x2(rs * (fy[4] * g[4] - fx[4] * a), rs * (fy[5] * g[5] - fx[5] * a));
        o.w = pack_bf16x2(rs * (fy[6] * g[6] - fx[6] * a), rs * (fy[7] * g[7] - fx[7] * a));
        pd[j] = o;
      }
    }
  }
  float* pp = partial + static_cast<size_t>(warp_g) * row_u4 * 8;
#pragma unroll
  for (int k = 0; k < MAXJ; ++k) {
    const int j = lane + 32 * k;
    if (j < row_u4) {
      *reinterpret_cast<float4*>(pp + j * 8) = make_float4(ds[k][0], ds[k][1], ds[k][2], ds[k][3]);
      *reinterpret_cast<float4*>(pp + j * 8 + 4) = make_float4(ds[k][4], ds[k][5], ds[k][6], ds[k][7]);
    }
  }
}

// dscale[j] = sum over the partial rows, in a fixed order: a block owns 32 columns, its 8 warps each sum every
// 8th partial row (coalesced 128-byte reads, 8 loads in flight), then the 8 sums are added in warp order.
__global__ void __launch_bounds__(256) rmsnorm_dscale_reduce_kernel(const float* __restrict__ partial, int nparts,
                                                                    int d, float* __restrict__ dscale) {
  __shared__ float s_acc[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  float a = 0.f;
  if (j < d) {
    int p = w;
    for (; p + 56 < nparts; p += 64) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = partial[static_cast<size_t>(p + 8 * u) * d + j];
#pragma unroll
      for (int u = 0; u < 8; ++u) a += v[u];
    }
    for (; p < nparts; p += 8) a += partial[static_cast<size_t>(p) * d + j];
  }
  s_acc[w][lane] = a;
  __syncthreads();
  if (w == 0 && j < d) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += s_acc[k][lane];
    dscale[j] = t;
  }
}

// ---------------------------------------------------------------------------------------------------
// f2: one pass over the micro-batch's tokens and labels: shifted labels (ssi/loss.py:16), valid-token counts
// before / after the shift (ssi/trainer.py:391, ce_chunked_output_loss.py:66) and the token-type histogram
// (ssi/train_utils.py:150-165: inclusive id ranges, plus "total" = tokens != pad_idx as the last entry).
// Integer atomics only (exact, order-independent).
// ---------------------------------------------------------------------------------------------------
constexpr int MAX_TOKEN_RANGES = 15;
struct TokenRanges {
  long long lo[MAX_TOKEN_RANGES], hi[MAX_TOKEN_RANGES];
  int n;
};

__global__ void __launch_bounds__(256) label_shift_count_types_kernel(
    const int64_t* __restrict__ labels, const int64_t* __restrict__ tokens, long long B, long long S,
    long long ignore_index, TokenRanges ranges, long long pad_idx, int64_t* __restrict__ out,
    unsigned long long* __restrict__ counts, unsigned long long* __restrict__ type_counts) {
  __shared__ unsigned int s_c[2 + MAX_TOKEN_RANGES + 1];
  if (threadIdx.x < 2 + MAX_TOKEN_RANGES + 1) s_c[threadIdx.x] = 0;
  __syncthreads();
  const long long total = B * S;
  unsigned int c0 = 0, c1 = 0, ct = 0;
  unsigned int cr[MAX_TOKEN_RANGES];
#pragma unroll
  for (int k = 0; k < MAX_TOKEN_RANGES; ++k) cr[k] = 0;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long s = i % S;
    const long long cur = labels[i];
    const long long nxt = (s + 1 < S) ? labels[i + 1] : ignore_index;
    out[i] = nxt;
    c0 += (cur != ignore_index);
    c1 += (nxt != ignore_index);
    const long long t = tokens[i];
    ct += (t != pad_idx);
#pragma unroll
    for (int k = 0; k < MAX_TOKEN_RANGES; ++k)
      if (k < ranges.n) cr[k] += (t >= ranges.lo[k] && t <= ranges.hi[k]);
  }
  atomicAdd(&s_c[0], c0);
  atomicAdd(&s_c[1], c1);
#pragma unroll
  for (int k = 0; k < MAX_TOKEN_RANGES; ++k)
    if (k < ranges.n && cr[k]) atomicAdd(&s_c[2 + k], cr[k]);
  atomicAdd(&s_c[2 + MAX_TOKEN_RANGES], ct);
  __syncthreads();
  if (threadIdx.x < 2 && s_c[threadIdx.x])
    atomicAdd(&counts[threadIdx.x], static_cast<unsigned long long>(s_c[threadIdx.x]));
  if (threadIdx.x >= 2 && threadIdx.x < 2 + ranges.n && s_c[threadIdx.x])
    atomicAdd(&type_counts[threadIdx.x - 2], static_cast<unsigned long long>(s_c[threadIdx.x]));
  if (threadIdx.x == 2 + MAX_TOKEN_RANGES && s_c[threadIdx.x])
    atomicAdd(&type_counts[ranges.n], static_cast<unsigned long long>(s_c[threadIdx.x]));
}

// ---------------------------------------------------------------------------------------------------
// f3: g = bf16(float(g) * scale) in place (scale_grads: p.grad *= scaler) and block partials of sum(g_new^2)
// (what clip_grad_norm_ then needs of this parameter) in the same pass; sqnorm_finalize_kernel sums the
// partials in order.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) grad_scale_sqnorm_kernel(uint4* __restrict__ g, size_t n_u4,
                                                                const float* __restrict__ scale,
                                                                float* __restrict__ partial) {
  __shared__ float s_w[8];
  const float sc = *scale;
  float acc = 0.f;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n_u4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const uint4 u = g[i];
    uint4 o;
    o.x = pack_bf16x2(bf16lo(u.x) * sc, bf16hi(u.x) * sc);
    o.y = pack_bf16x2(bf16lo(u.y) * sc, bf16hi(u.y) * sc);
    o.z = pack_bf16x2(bf16lo(u.z) * sc, bf16hi(u.z) * sc);
    o.w = pack_bf16x2(bf16lo(u.w) * sc, bf16hi(u.w) * sc);
    g[i] = o;
    float f[8];
    bf16x8_to_f32(o, f);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc = fmaf(f[e], f[e], acc);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += s_w[w];
    partial[blockIdx.x] = a;
  }
}

// dst = bf16(float(dst) + float(src)): one micro-batch's all-reduced contribution joins an existing gradient
// (NCCL exchange with gradient accumulation: only the NEW contribution may be summed over ranks)
__global__ void __launch_bounds__(256) grad_add_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src,
                                                       size_t n_u4) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n_u4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float a[8], b[8];
    bf16x8_to_f32(dst[i], a);
    bf16x8_to_f32(src[i], b);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] += b[e];
    dst[i] = f32x8_to_bf16(a);
  }
}

__global__ void __launch_bounds__(256) sqnorm_finalize_kernel(const float* __restrict__ partial, int n,
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
  if (threadIdx.x == 0) out[0] = static_cast<float>(s[0]);
}

// ---------------------------------------------------------------------------------------------------
// f4: out[i, :] = W[ids[i], :] (zeros for ids outside [0, V)), one warp per row, 16-byte accesses.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) embedding_gather_kernel(const uint4* __restrict__ W,
                                                               const int64_t* __restrict__ ids, int n, long long V,
                                                               int row_u4, uint4* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n; r += warps) {
    const long long id = ids[r];
    uint4* dst = out + static_cast<size_t>(r) * row_u4;
    if (id >= 0 && id < V) {
      const uint4* src = W + static_cast<size_t>(id) * row_u4;
      for (int j = lane; j < row_u4; j += 32) dst[j] = __ldg(src + j);
    } else {
      for (int j = lane; j < row_u4; j += 32) dst[j] = make_uint4(0, 0, 0, 0);
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// f1 fused into the path (single-pass mode with a norm scale): the final RMSNorm is applied WHILE the valid rows
// are gathered into X_c (no separate pass over the hidden states), and its backward runs WHILE the compact dX rows
// are scattered back to their positions, so the path hands back d(pre-norm h) directly.
// ---------------------------------------------------------------------------------------------------
// X_c[c, :] = rmsnorm(X[idx[c], :]) for c < n_valid (zeros for tile padding rows); rstd_c[c] kept for backward.
__global__ void __launch_bounds__(256) gather_rows_norm_kernel(const uint4* __restrict__ X, const int* __restrict__ idx,
                                                               const DevPlan* __restrict__ plan, int row_u4,
                                                               const uint4* __restrict__ scale, float eps,
                                                               uint4* __restrict__ Xc, float* __restrict__ rstd_c) {
  const int n_valid = plan->n_valid;
  const int rows = plan->tiles_m * 128;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float inv_d = 1.f / static_cast<float>(row_u4 * 8);
  for (int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < rows; c += warps) {
    uint4* dst = Xc + static_cast<size_t>(c) * row_u4;
    if (c >= n_valid) {
      for (int j = lane; j < row_u4; j += 32) dst[j] = make_uint4(0, 0, 0, 0);
      if (lane == 0) rstd_c[c] = 0.f;
      continue;
    }
    const uint4* src = X + static_cast<size_t>(idx[c]) * row_u4;
    float ss = 0.f;
    for (int j = lane; j < row_u4; j += 32) {
      float f[8];
      bf16x8_to_f32(__ldg(src + j), f);
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(f[e], f[e], ss);
    }
    ss = warp_sum(ss);
    const float rs = rsqrtf(ss * inv_d + eps);
    if (lane == 0) rstd_c[c] = rs;
    for (int j = lane; j < row_u4; j += 32) {
      float f[8], g[8];
      bf16x8_to_f32(__ldg(src + j), f);
      bf16x8_to_f32(__ldg(scale + j), g);
      uint4 o;
      o.x = pack_bf16x2((f[0] * rs) * g[0], (f[1] * rs) * g[1]);
      o.y = pack_bf16x2((f[2] * rs) * g[2], (f[3] * rs) * g[3]);
      o.z = pack_bf16x2((f[4] * rs) * g[4], (f[5] * rs) * g[5]);
      o.w = pack_bf16x2((f[6] * rs) * g[6], (f[7] * rs) * g[7]);
      dst[j] = o;
    }
  }
}

// dX[r, :] = rmsnorm_backward(dy = dXc[pos[r], :], x = X[r, :]) for valid rows, zeros for ignored rows; dscale
// partials per warp as in rmsnorm_bwd_kernel (summed by rmsnorm_dscale_reduce_kernel).
template <int MAXJ>
__global__ void __launch_bounds__(256) dx_scatter_norm_bwd_kernel(const uint4* __restrict__ dXc,
                                                                  const int* __restrict__ pos,
                                                                  const uint4* __restrict__ X,
                                                                  const uint4* __restrict__ scale,
                                                                  const float* __restrict__ rstd_c, int n, int row_u4,
                                                                  uint4* __restrict__ dX, float* __restrict__ partial) {
  const int lane = threadIdx.x & 31;
  const int warp_g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float inv_d = 1.f / static_cast<float>(row_u4 * 8);
  float ds[MAXJ][8];
#pragma unroll
  for (int k = 0; k < MAXJ; ++k)
#pragma unroll
    for (int e = 0; e < 8; ++e) ds[k][e] = 0.f;
  for (int r = warp_g; r < n; r += warps) {
    const int c = pos[r];
    uint4* pd = dX + static_cast<size_t>(r) * row_u4;
    if (c < 0) {
      for (int j = lane; j < row_u4; j += 32) pd[j] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const uint4* py = dXc + static_cast<size_t>(c) * row_u4;
    const uint4* px = X + static_cast<size_t>(r) * row_u4;
    const float rs = rstd_c[c];
    float cc = 0.f;
#pragma unroll
    for (int k = 0; k < MAXJ; ++k) {
      const int j = lane + 32 * k;
      if (j < row_u4) {
        float fy[8], fx[8], g[8];
        bf16x8_to_f32(py[j], fy);
        bf16x8_to_f32(__ldg(px + j), fx);
        bf16x8_to_f32(__ldg(scale + j), g);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          cc = fmaf(fy[e] * g[e], fx[e], cc);
          ds[k][e] = fmaf(fy[e] * fx[e], rs, ds[k][e]);
        }
      }
    }
    cc = warp_sum(cc) * inv_d;
    const float a = rs * rs * cc;
#pragma unroll
    for (int k = 0; k < MAXJ; ++k) {
      const int j = lane + 32 * k;
      if (j < row_u4) {
        float fy[8], fx[8], g[8];
        bf16x8_to_f32(py[j], fy);
        bf16x8_to_f32(__ldg(px + j), fx);
        bf16x8_to_f32(__ldg(scale + j), g);
        uint4 o;
        o.x = pack_bf16x2(rs * (fy[0] * g[0] - fx[0] * a), rs * (fy[1] * g[1] - fx[1] * a));
        o.y = pack_bf16x2(rs * (fy[2] * g[2] - fx[2] * a), rs * (fy[3] * g[3] - fx[3] * a));
        o.z = pack_bf16x2(rs * (fy[4] * g[4] - fx[4] * a), rs * (fy[5] * g[5] - fx[5] * a));
        o.w = pack_bf16x2(rs * (fy[6] * g[6] - fx[6] * a), rs * (fy[7] * g[7] - fx[7] * a));
        pd[j] = o;
      }
    }
  }
  float* pp = partial + static_cast<size_t>(warp_g) * row_u4 * 8;
#pragma unroll
  for (int k = 0; k < MAXJ; ++k) {
    const int j = lane + 32 * k;
    if (j < row_u4) {
      *reinterpret_cast<float4*>(pp + j * 8) = make_float4(ds[k][0], ds[k][1], ds[k][2], ds[k][3]);
      *reinterpret_cast<float4*>(pp + j * 8 + 4) = make_float4(ds[k][4], ds[k][5], ds[k][6], ds[k][7]);
    }
  }
}

}  // namespace ssi
