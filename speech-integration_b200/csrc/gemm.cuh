// Persistent, warp-specialised tcgen05 contraction used by every tensor-core stage of the path:
//   EPI_LSE   logits tile -> online log-sum-exp partials              (forward,  K1)
//   EPI_GRAD  logits tile -> dlogits block G in bf16                  (backward, K2a)
//   EPI_BF16  accumulator -> bf16 tile store / reduce-add             (dW = G^T X, K2b; dX when one block)
//   EPI_F32   accumulator -> fp32 tile store / reduce-add             (dX accumulation, K2c; test hook)
//   EPI_LSE_STASH  logits tile -> bf16 logits block + online LSE      (single-pass K1, separate softmax-grad pass)
//   EPI_EXP_STASH  logits tile -> bf16 block of E = exp(logit - ref[row]) + its row sums
//                  (single-pass K1, default: the softmax normalisation 1 / sum E is a per-ROW factor, so it moves
//                  into the dX epilogue and into the rows of X that dW contracts with -- no pass over the block)
// One CTA per SM: warp 0 = TMA producer, warp 1 = MMA issuer (+TMEM owner), warps 2..5 = epilogue
// (one thread per accumulator row).  128x256 fp32 accumulators, double-buffered in the 512 TMEM columns, so
// the epilogue of tile j overlaps the MMAs of tile j+1.  Operands are staged by TMA with 128-byte swizzle.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>

#include "aux_kernels.cuh"
#include "ptx.cuh"

namespace ssi {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB
// CG = CTAs per MMA (tcgen05 cta_group). CG 1: one CTA owns a 128x256 tile, stage = A 16 KiB + B 32 KiB, 4 stages.
// CG 2: a CTA pair (one TPC) owns a 256x256 tile; each CTA stages its own 128 rows of A and HALF of B (128 of
// the 256 columns), so a stage is 32 KiB per CTA (6 stages) and each SM pulls a third less operand data
// through L2 / shared memory per flop.
// NW = n-tiles per work item (CG 2 only).  NW 2: the pair owns a 256x512 tile -- BOTH 256-column accumulators of
// the 512 TMEM columns belong to one tile and share every A stage, so a stage is A 16 KiB + 2 x 16 KiB of B per
// CTA (4 stages) and each SM pulls a quarter less operand data through L2 per flop than with 256x256 tiles.  The
// accumulator is then single-buffered (the epilogue no longer overlaps the next tile's MMAs), which only suits
// the long-K contractions (dW, dX).
template <int CG, int NW = 1> struct StageCfg {
  static_assert(NW == 1 || (NW == 2 && CG == 2), "wide tiles need a CTA pair");
  static constexpr int B_HALF_BYTES = (BN / CG) * BK * 2;
  static constexpr int B_BYTES = NW * B_HALF_BYTES;
  static constexpr int BYTES = A_STAGE_BYTES + B_BYTES;
  static constexpr int STAGES = (CG == 1) ? 4 : (NW == 2 ? 4 : 6);
};
constexpr int MAX_STAGES = 6;
constexpr int PIPE_BYTES = 4 * (A_STAGE_BYTES + BN * BK * 2);  // 192 KiB either way
static_assert(StageCfg<1>::STAGES * StageCfg<1>::BYTES == PIPE_BYTES && StageCfg<2>::STAGES * StageCfg<2>::BYTES == PIPE_BYTES &&
              StageCfg<2, 2>::STAGES * StageCfg<2, 2>::BYTES == PIPE_BYTES, "");
constexpr int MN_BOX_BYTES = 64 * BK * 2;  // one {64 MN, 64 K} TMA box of an MN-major operand = 8 KiB
constexpr int EPI_BUF_BYTES = 128 * 128;   // 128 rows x 128 B staging tile for TMA stores
constexpr int NUM_EPI_BUFS = 2;
constexpr int SMEM_BAR_BYTES = 256;
constexpr int GEMM_SMEM_BYTES = 1024 + PIPE_BYTES + NUM_EPI_BUFS * EPI_BUF_BYTES + SMEM_BAR_BYTES;
constexpr int GEMM_THREADS = 192;
constexpr int EPI_THREADS = 128;
constexpr int TMEM_COLS = 512;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

enum Epi { EPI_LSE = 0, EPI_GRAD = 1, EPI_BF16 = 2, EPI_F32 = 3, EPI_LSE_STASH = 4, EPI_EXP_STASH = 5 };
template <int EPI>
__host__ __device__ constexpr bool is_lse_epi() {
  return EPI == EPI_LSE || EPI == EPI_LSE_STASH || EPI == EPI_EXP_STASH;
}
// GemmParams::plan_rows bits: add dplan->row0 (first compact row of a row chunk) to ...
enum PlanRows { PR_A_ROWS = 1, PR_B_K = 2, PR_C_ROWS = 4, PR_LABEL_ROWS = 8 };

struct GemmParams {
  CUtensorMap tmA;
  CUtensorMap tmB;
  CUtensorMap tmC;
  int tiles_m, tiles_n, k_blocks, group_m;
  int a_r0, a_k0, b_r0, b_k0, c_r0, c_c0;
  int c_reduce;  // EPI_BF16 / EPI_F32: 1 = TMA reduce-add into C, 0 = plain store
  int c_reduce_plan;  // 1 = take it from the chunk plan instead: reduce-add unless dplan->first
  int tail_split;     // wide tiles (NW 2): run the items of a last, at most half-full wave as 256x256 halves
  // Shape gate (sizes that depend on the number of valid rows are only known on the device): the host enqueues the
  // 256x512 kernel (gate 1) AND the 256x256 kernel (gate 2) for the same contraction; each evaluates the same
  // predicate from the plan and exactly one of them runs, the other exits at once.  0 = no gate.
  int shape_gate, gate_workers;
  unsigned long long store_policy;  // L2 cache policy for plain C / G stores (0 = default)
  unsigned long long a_policy, b_policy;  // L2 cache policy of the A / B operand loads (0 = default)
  // Lockstep throttle (EPI_BF16 / EPI_F32 only, uniform K loops): every `sync_every` k-blocks the producers of
  // all co-running tiles meet at a split-phase counter barrier, so the tiles of one wave stream the same
  // K window of their shared operands through L2 together instead of drifting apart.  nullptr = off.
  // The throttle is a performance hint only: a wait that is not satisfied within LOCKSTEP_TIMEOUT_NS (part of the
  // grid not resident: MPS, a debugger, a smaller device) sets *sync_err and the CTA runs un-throttled from then on.
  unsigned int* sync_ctr;
  unsigned int* sync_err;
  int sync_every;
  // Device-side schedule (row compaction): when set, token-dependent sizes are read from *dplan at kernel start
  const DevPlan* dplan;
  int m_from_plan;  // tiles_m (and for EPI_LSE the split schedule and partial stride) come from dplan
  int k_from_plan;  // k_blocks = dplan->k_blocks_tok (the K loop runs over compacted token rows)
  int plan_rows;    // PlanRows bits (single-pass mode: the plan describes one row chunk)
  // ---- EPI_LSE / EPI_GRAD
  const int64_t* labels;
  long long ignore_index;
  int n_rows;  // N (rows >= n_rows are padding)
  int vocab;   // V (columns >= vocab are padding)
  int col0;    // global vocabulary index of column 0 of n-tile 0 (EPI_GRAD: start of the vocab block)
  // EPI_LSE: tiles_n = number of vocabulary splits; split s sweeps n-tiles [s*per, min((s+1)*per, total))
  int vtiles_per_split, vtiles_total;
  float* part_m;  // [splits][part_stride]  running max (log2 domain)
  float* part_l;  // [splits][part_stride]  running sum of 2^(s*log2e - m)
  float* part_z;  // [splits][part_stride]  logit of the label if it fell in this split, else -inf
  int part_stride;
  // EPI_GRAD
  const float* lse;          // [N] natural log
  const float* grad_out;     // [1]
  const int64_t* n_valid;    // [1]
  // EPI_LSE: 1 = ONE split of p.vtiles_per_split tiles whatever the plan says (the pilot sweep over the first tile)
  int lse_one_split;
  // LSE epilogues: m-tiles (128 rows) per raster group; 0 = all of them.  The X rows of one group are re-read from L2
  // for every vocabulary tile, so a group should fit ONE L2 partition next to the stash write stream.
  int lse_group_m;
  // EPI_EXP_STASH: per-row reference in the log2 domain, indexed by the plan-relative row (the pilot's running max)
  const float* row_ref2;
  // EPI_BF16: optional per-row factor applied before the rounding, indexed like the C rows (c_r0 + plan offset + row)
  const float* row_scale;
};

struct TileItem {
  int m, n0, ncount;
};

constexpr unsigned long long LOCKSTEP_TIMEOUT_NS = 2ull * 1000 * 1000;  // normal waits are a few microseconds

// bounded spin of the lockstep throttle; returns false (and raises *err) when it gave up
__device__ __forceinline__ bool lockstep_wait(const unsigned int* ctr, unsigned int target, unsigned int* err) {
  unsigned int spins = 0;
  unsigned long long t0 = 0;
  while (static_cast<int>(ld_acquire_gpu(ctr) - target) < 0) {
    if ((++spins & 255u) == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (t0 == 0) t0 = t;
      else if (t - t0 > LOCKSTEP_TIMEOUT_NS) {
        if (err) atomicExch(err, 1u);
        return false;
      }
    }
  }
  return true;
}

struct Sched {
  int tiles_m, tiles_n, group_m, k_blocks, vtiles_per_split, vtiles_total, part_stride;
  int a_row_off, b_k_off, c_row_off, label_off;
  int c_reduce;
};

template <int EPI>
__device__ __forceinline__ Sched make_sched(const GemmParams& p) {
  Sched s;
  s.tiles_m = p.tiles_m, s.tiles_n = p.tiles_n, s.group_m = p.group_m, s.k_blocks = p.k_blocks;
  s.vtiles_per_split = p.vtiles_per_split, s.vtiles_total = p.vtiles_total, s.part_stride = p.part_stride;
  s.a_row_off = s.b_k_off = s.c_row_off = s.label_off = 0;
  s.c_reduce = p.c_reduce;
  if (p.dplan != nullptr) {
    if (p.c_reduce_plan) s.c_reduce = (p.dplan->first == 0) ? 1 : 0;
    const int row0 = p.dplan->row0;
    if (p.plan_rows & PR_A_ROWS) s.a_row_off = row0;
    if (p.plan_rows & PR_B_K) s.b_k_off = row0;
    if (p.plan_rows & PR_C_ROWS) s.c_row_off = row0;
    if (p.plan_rows & PR_LABEL_ROWS) s.label_off = row0;
    if (p.m_from_plan) {
      s.tiles_m = p.dplan->tiles_m;
      if constexpr (is_lse_epi<EPI>()) {
        s.tiles_n = p.dplan->splits;
        s.vtiles_per_split = p.dplan->vtiles_per_split;
        if (p.lse_one_split) s.tiles_n = 1, s.vtiles_per_split = p.vtiles_per_split;
        s.part_stride = p.dplan->part_stride;
        s.group_m = max(p.lse_group_m > 0 ? min(p.lse_group_m, s.tiles_m) : s.tiles_m, 1);
      }
    }
    if (p.k_from_plan) s.k_blocks = p.dplan->k_blocks_tok;
  }
  // accumulating an empty contraction (an empty row chunk) is a no-op
  if (s.k_blocks == 0 && s.c_reduce) s.tiles_m = 0;
  return s;
}

template <int EPI>
__device__ __forceinline__ TileItem decode_item(const Sched& p, int item) {
  const int per_group = p.group_m * p.tiles_n;
  const int g = item / per_group;
  const int r = item - g * per_group;
  const int m0 = g * p.group_m;
  const int gm = min(p.group_m, p.tiles_m - m0);
  TileItem it;
  it.m = m0 + r % gm;
  const int n = r / gm;
  if constexpr (is_lse_epi<EPI>()) {
    it.n0 = n * p.vtiles_per_split;
    it.ncount = min(p.vtiles_per_split, p.vtiles_total - it.n0);
  } else {
    it.n0 = n;
    it.ncount = 1;
  }
  return it;
}

// 128-byte-swizzled staging row write: 8 x 16-byte chunks of row `row`, chunk j lands at (j ^ (row & 7)).
__device__ __forceinline__ void stage_row_128B(uint32_t buf, int row, const uint32_t (&w)[32]) {
  const uint32_t base = buf + row * 128;
  const int sw = row & 7;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    st_shared_v4(base + ((j ^ sw) << 4), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
}

// epilogue warp -> MMA issuer: "this accumulator buffer has been read" (CG 2: the barrier lives in the leader CTA)
template <int CG>
__device__ __forceinline__ void release_acc(uint32_t bar) {
  if constexpr (CG == 2)
    mbar_arrive_cluster(bar, 0);
  else
    mbar_arrive(bar);
}

template <int EPI, bool A_MN, bool B_MN, int CG = 1, int NW = 1>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_kernel(const __grid_constant__ GemmParams p) {
  static_assert(NW == 1 || EPI == EPI_BF16 || EPI == EPI_F32, "wide tiles: plain store epilogues only");
  constexpr int STAGES = StageCfg<CG, NW>::STAGES;
  constexpr int STAGE_BYTES = StageCfg<CG, NW>::BYTES;
  constexpr int BN_CTA = BN / CG;  // B columns of ONE 256-column n-tile staged by this CTA
  constexpr int B_HALF_BYTES = StageCfg<CG, NW>::B_HALF_BYTES;
  constexpr int BNW = BN * NW;     // output columns per work item
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_stage0 = smem_base;
  const uint32_t smem_epi0 = smem_base + PIPE_BYTES;
  const uint32_t smem_bar0 = smem_epi0 + NUM_EPI_BUFS * EPI_BUF_BYTES;
  // barriers: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], then the TMEM base address slot.
  // CG 2: full[] and tmem_empty[] are only used in the leader CTA (rank 0) of the pair; empty[] and
  // tmem_full[] exist in both and are arrived on by the leader's multicast tcgen05.commit.
  auto full_bar = [&](int s) { return smem_bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return smem_bar0 + 8u * (MAX_STAGES + s); };
  auto tfull_bar = [&](int s) { return smem_bar0 + 8u * (2 * MAX_STAGES + s); };
  auto tempty_bar = [&](int s) { return smem_bar0 + 8u * (2 * MAX_STAGES + 2 + s); };
  const uint32_t tmem_slot = smem_bar0 + 8u * (2 * MAX_STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;
  const int worker = (CG == 2) ? (blockIdx.x >> 1) : blockIdx.x;       // CTA (pair) index
  const int workers = (CG == 2) ? (gridDim.x >> 1) : gridDim.x;
  Sched sc = make_sched<EPI>(p);
  if (p.shape_gate != 0) {
    // wide tiles pay off when they fill more than half a wave (otherwise every item is a latency-starved half
    // item) and the K loop is long enough to amortise their single-buffered epilogue
    const int pairs = (sc.tiles_m + 1) >> 1;
    const int wide_items_g = pairs * ((sc.tiles_n + 1) >> 1);
    const bool wide_ok = (2 * wide_items_g > p.gate_workers) && (sc.k_blocks >= 48);
    if ((p.shape_gate == 1) != wide_ok) return;  // uniform over the whole grid
  }
  const int tiles_m_real = sc.tiles_m;  // in 128-row tiles
  if constexpr (CG == 2) {
    sc.tiles_m = (sc.tiles_m + 1) >> 1;  // the schedule runs over 256-row tile pairs
    sc.group_m = max(1, sc.group_m >> 1);
  }
  if constexpr (NW == 2) sc.tiles_n = (sc.tiles_n + 1) >> 1;  // ... and over 512-column tile pairs
  // Wide tiles: when the last wave of the static schedule would be at most half full, its items are split into their
  // two 256-column halves (one MMA per k-step instead of two, on the same stage layout), so that wave takes half
  // as long: e.g. 256 items on 74 CTA pairs = 3 waves + 34 items -> 3 waves + 68 half items.
  const int wide_items = sc.tiles_m * sc.tiles_n;
  int split_from = wide_items, total_items = wide_items;
  if constexpr (NW == 2) {
    if (p.tail_split && workers > 0) {
      const int tail = wide_items % workers;
      if (tail > 0 && 2 * tail <= workers) split_from = wide_items - tail, total_items = split_from + 2 * tail;
    }
  }
  // item -> (tile, first 256-column half, number of halves)
  auto get_item = [&](int item, int& h0, int& nh) {
    if (NW == 2 && item >= split_from) {
      const int j = item - split_from;
      h0 = j & 1, nh = 1;
      return decode_item<EPI>(sc, split_from + (j >> 1));
    }
    h0 = 0, nh = NW;
    return decode_item<EPI>(sc, item);
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    if constexpr (EPI != EPI_LSE) tma_prefetch_desc(&p.tmC);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 4 * CG);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc<CG>(tmem_slot, TMEM_COLS);
    tmem_relinquish<CG>();
  }
  tc_fence_before();
  if constexpr (CG == 2) {
    cluster_arrive();
    cluster_wait();
  } else {
    __syncthreads();
  }
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    // ===================================================== TMA producer (one per CTA)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      // lockstep throttle: only the leader producers take part (the peer follows through the shared pipeline)
      const bool throttle = !is_lse_epi<EPI>() && EPI != EPI_GRAD && p.sync_ctr != nullptr && p.sync_every > 0 &&
                            cta_rank == 0 && workers > 1;
      const int sync_T = throttle ? p.sync_every : 0x7fffffff;
      const int sync_pts = throttle ? (sc.k_blocks - 1) / sync_T : 0;  // points at kb = T, 2T, ... < k_blocks
      unsigned int sync_base = 0;                                       // arrivals of all earlier rounds
      // K1 (LSE epilogues): the items of one round sweep their vocabulary tiles together -- one sync point per tile
      // after the first (arrive at its first k-block, wait at its middle).  Splits are ragged, so an item with
      // fewer tiles than the longest posts the arrivals it will never make when it ends ("catch-up").
      const bool throttle_v = is_lse_epi<EPI>() && p.sync_ctr != nullptr && p.sync_every > 0 && cta_rank == 0 &&
                              workers > 1 && sc.k_blocks >= 2;
      const int v_pts = throttle_v ? sc.vtiles_per_split - 1 : 0;
      bool wait_ok = true;  // false once a wait timed out: keep posting arrivals (others may still wait), stop waiting
      for (int item = worker; item < total_items; item += workers) {
        int h0, nh;
        const TileItem it = get_item(item, h0, nh);
        const int m_tile = (CG == 2) ? (2 * it.m + static_cast<int>(cta_rank)) : it.m;
        const unsigned int round_workers = static_cast<unsigned int>(min(workers, total_items - (item - worker)));
        for (int t = 0; t < it.ncount; ++t) {
          const int nt = it.n0 + t;
          const int b_row = p.b_r0 + nt * BNW + h0 * BN + static_cast<int>(cta_rank) * BN_CTA;
          int next_arrive = sync_T, next_wait = 0x7fffffff, sync_j = 0;
          for (int kb = 0; kb < sc.k_blocks; ++kb) {
            if (throttle_v && t >= 1) {
              if (kb == 0) atomicAdd(p.sync_ctr, 1u);
              if (kb == (sc.k_blocks >> 1) && wait_ok) {
                const unsigned int target = sync_base + static_cast<unsigned int>(t) * round_workers;
                wait_ok = lockstep_wait(p.sync_ctr, target, p.sync_err);
              }
            }
            if (throttle) {
              if (kb == next_arrive) {
                atomicAdd(p.sync_ctr, 1u);
                next_wait = min(kb + sync_T / 2, sc.k_blocks - 1);
                next_arrive = (sync_j + 1 < sync_pts) ? kb + sync_T : 0x7fffffff;
              }
              if (kb == next_wait) {
                const unsigned int target = sync_base + static_cast<unsigned int>(sync_j + 1) * round_workers;
                if (wait_ok) wait_ok = lockstep_wait(p.sync_ctr, target, p.sync_err);
                ++sync_j;
                next_wait = 0x7fffffff;
              }
            }
            mbar_wait(empty_bar(stage), phase ^ 1u);
            // CG 2: both CTAs signal the LEADER's full barrier; only the leader posts the expected bytes
            const uint32_t fb = (CG == 2) ? (full_bar(stage) & 0xFEFFFFFFu) : full_bar(stage);
            if (CG == 1 || cta_rank == 0)
              mbar_expect_tx(full_bar(stage), CG * (A_STAGE_BYTES + nh * B_HALF_BYTES));
            const uint32_t sA = smem_stage0 + stage * STAGE_BYTES;
            const uint32_t sB = sA + A_STAGE_BYTES;
            if constexpr (!A_MN) {
              tma_load_2d_g<CG>(sA, &p.tmA, p.a_k0 + kb * BK, p.a_r0 + sc.a_row_off + m_tile * BM, fb, p.a_policy);
            } else {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j)
                tma_load_2d_g<CG>(sA + j * MN_BOX_BYTES, &p.tmA, p.a_r0 + m_tile * BM + 64 * j, p.a_k0 + kb * BK, fb,
                                  p.a_policy);
            }
#pragma unroll
            for (int h = 0; h < NW; ++h) {
              if (h >= nh) break;
              if constexpr (!B_MN) {
                tma_load_2d_g<CG>(sB + h * B_HALF_BYTES, &p.tmB, p.b_k0 + kb * BK, b_row + h * BN, fb, p.b_policy);
              } else {
#pragma unroll
                for (int j = 0; j < BN_CTA / 64; ++j)
                  tma_load_2d_g<CG>(sB + h * B_HALF_BYTES + j * MN_BOX_BYTES, &p.tmB, b_row + h * BN + 64 * j,
                                    p.b_k0 + sc.b_k_off + kb * BK, fb, p.b_policy);
              }
            }
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1u;
            }
          }
        }
        sync_base += static_cast<unsigned int>(sync_pts) * round_workers;
        if (throttle_v) {
          const int missing = v_pts - (it.ncount - 1);
          if (missing > 0) atomicAdd(p.sync_ctr, static_cast<unsigned int>(missing));
          sync_base += static_cast<unsigned int>(v_pts) * round_workers;
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer (single thread; CG 2: leader CTA only)
    if (lane == 0 && cta_rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM * CG, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      constexpr uint32_t A_LBO = A_MN ? MN_BOX_BYTES : 16;
      constexpr uint32_t B_LBO = B_MN ? MN_BOX_BYTES : 16;
      constexpr uint32_t A_KADV = A_MN ? (UMMA_K * 128) : (UMMA_K * 2);  // bytes per UMMA_K step
      constexpr uint32_t B_KADV = B_MN ? (UMMA_K * 128) : (UMMA_K * 2);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int item = worker; item < total_items; item += workers) {
        int h0, nh;
        const TileItem it = get_item(item, h0, nh);
        for (int t = 0; t < it.ncount; ++t) {
          if constexpr (CG == 2)
            mbar_wait_cluster(tempty_bar(acc), acc_phase ^ 1u);
          else
            mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (NW == 2 ? 0 : acc * BN);
          for (int kb = 0; kb < sc.k_blocks; ++kb) {
            mbar_wait(full_bar(stage), phase);
            tc_fence_after();
            const uint32_t sA = smem_stage0 + stage * STAGE_BYTES;
            const uint32_t sB = sA + A_STAGE_BYTES;
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t adesc = make_smem_desc(sA + k * A_KADV, A_LBO, 1024);
#pragma unroll
              for (int h = 0; h < NW; ++h) {
                if (h >= nh) break;
                const uint64_t bdesc = make_smem_desc(sB + h * B_HALF_BYTES + k * B_KADV, B_LBO, 1024);
                umma_bf16<CG>(d_tmem + h * BN, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
              }
            }
            if constexpr (CG == 2)
              umma_commit_cg2(empty_bar(stage), 3);
            else
              umma_commit(empty_bar(stage));
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1u;
            }
          }
          if constexpr (CG == 2)
            umma_commit_cg2(tfull_bar(acc), 3);
          else
            umma_commit(tfull_bar(acc));
          if (NW == 2 || ++acc == 2) {  // wide tiles: one accumulator set, the phase flips every tile
            acc = 0;
            acc_phase ^= 1u;
          }
        }
      }
    }
  } else {
    // ===================================================== epilogue warps (one thread per accumulator row)
    const int q = warp & 3;               // TMEM lane quadrant this warp may access
    const int row_in_tile = q * 32 + lane;
    const bool issuer = (threadIdx.x == 64);
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t chunk_ctr = 0;  // alternates the staging buffers across the whole kernel
    (void)issuer;
    (void)chunk_ctr;

    const bool zero_acc = (sc.k_blocks == 0);  // empty contraction (no valid rows): the result is exactly zero
    (void)zero_acc;
    float gscale = 0.f;
    if constexpr (EPI == EPI_GRAD) {
      const long long nv = *p.n_valid;
      gscale = (nv > 0) ? (*p.grad_out / static_cast<float>(nv)) : 0.f;
    }

    for (int item = worker; item < total_items; item += workers) {
      int h0, nh;
      const TileItem it = get_item(item, h0, nh);
      const int m_tile = (CG == 2) ? (2 * it.m + static_cast<int>(cta_rank)) : it.m;
      const bool m_ok = (CG == 1) || (m_tile < tiles_m_real);  // odd tile count: the pair's 2nd tile is a phantom
      const int row = m_tile * BM + row_in_tile;  // row of this thread (relative to the plan's row0)

      // per-item row state
      long long label = p.ignore_index;
      bool valid = false;
      float m_run = -INFINITY, l_run = 0.f, z_lab = -INFINITY;  // EPI_LSE
      float lse2 = 0.f;                                         // EPI_GRAD
      float ref2 = 0.f;                                         // EPI_EXP_STASH
      if constexpr (EPI == EPI_EXP_STASH) {
        if (m_ok) ref2 = p.row_ref2[row];
      }
      float rscale = 1.f;                                       // EPI_BF16 with row_scale
      if constexpr (EPI == EPI_BF16) {
        if (p.row_scale != nullptr) rscale = m_ok ? p.row_scale[p.c_r0 + sc.c_row_off + row] : 0.f;
      }
      (void)ref2, (void)rscale;
      if constexpr (is_lse_epi<EPI>() || EPI == EPI_GRAD) {
        if (sc.label_off + row < p.n_rows) {
          label = p.labels[sc.label_off + row];
          valid = (label != p.ignore_index);
        }
        if constexpr (EPI == EPI_GRAD) {
          if (valid) lse2 = p.lse[row] * LOG2E;
        }
      }

      for (int t = 0; t < it.ncount; ++t) {
        const int nt = it.n0 + t;
        mbar_wait(tfull_bar(acc), acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + (NW == 2 ? 0 : acc * BN);

        if constexpr (EPI == EPI_LSE) {
          const int cbase = p.col0 + nt * BN;
          const bool tile_has_pad = (cbase + BN > p.vocab);
#pragma unroll 1
          for (int c = 0; c < BN / 32; ++c) {
            const int c0 = cbase + c * 32;
            if (c0 >= p.vocab) break;
            float v[32];
            tmem_ld32(taddr + c * 32, v);
            // the reference's logits are bf16 (tied_linear.py:25) and upcast before the CE
            // (ce_chunked_output_loss.py:41): round here too, so every mode sees the same values
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const uint32_t pk = pack_bf16x2(v[2 * j], v[2 * j + 1]);
              v[2 * j] = bf16lo(pk);
              v[2 * j + 1] = bf16hi(pk);
            }
            if (tile_has_pad) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (c0 + j >= p.vocab) v[j] = -INFINITY;
            }
            if (valid && label >= c0 && label < c0 + 32) {
              const int jl = static_cast<int>(label - c0);
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j == jl) z_lab = v[j];
            }
            float cmax = v[0];
#pragma unroll
            for (int j = 1; j < 32; ++j) cmax = fmaxf(cmax, v[j]);
            const float m_new = fmaxf(m_run, cmax * LOG2E);
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              s0 += fast_exp2(fmaf(v[j], LOG2E, -m_new));
              s1 += fast_exp2(fmaf(v[j + 1], LOG2E, -m_new));
              s2 += fast_exp2(fmaf(v[j + 2], LOG2E, -m_new));
              s3 += fast_exp2(fmaf(v[j + 3], LOG2E, -m_new));
            }
            l_run = l_run * fast_exp2(m_run - m_new) + ((s0 + s1) + (s2 + s3));
            m_run = m_new;
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) release_acc<CG>(tempty_bar(acc));
        } else if constexpr (EPI == EPI_LSE_STASH) {
          // Single-pass mode: the logits tile is rounded to bf16 (the reference's logits dtype,
          // tied_linear.py:25) and stored into the row chunk's block S; the online log-sum-exp runs on the
          // ROUNDED values, like ce_chunked_output_loss.py:41 upcasting bf16 logits.
          const int cbase = p.col0 + nt * BN;
#pragma unroll 1
          for (int c = 0; c < BN / 64; ++c) {
            uint32_t w[32];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              float v[32];
              tmem_ld32(taddr + c * 64 + h * 32, v);
              const int ch0 = cbase + c * 64 + h * 32;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const uint32_t pk = pack_bf16x2(v[2 * j], v[2 * j + 1]);
                w[h * 16 + j] = pk;
                v[2 * j] = bf16lo(pk);
                v[2 * j + 1] = bf16hi(pk);
              }
              if (ch0 < p.vocab) {
                if (ch0 + 32 > p.vocab) {
#pragma unroll
                  for (int j = 0; j < 32; ++j)
                    if (ch0 + j >= p.vocab) v[j] = -INFINITY;
                }
                if (valid && label >= ch0 && label < ch0 + 32) {
                  const int jl = static_cast<int>(label - ch0);
#pragma unroll
                  for (int j = 0; j < 32; ++j)
                    if (j == jl) z_lab = v[j];
                }
                float cmax = v[0];
#pragma unroll
                for (int j = 1; j < 32; ++j) cmax = fmaxf(cmax, v[j]);
                const float m_new = fmaxf(m_run, cmax * LOG2E);
                float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                  s0 += fast_exp2(fmaf(v[j], LOG2E, -m_new));
                  s1 += fast_exp2(fmaf(v[j + 1], LOG2E, -m_new));
                  s2 += fast_exp2(fmaf(v[j + 2], LOG2E, -m_new));
                  s3 += fast_exp2(fmaf(v[j + 3], LOG2E, -m_new));
                }
                l_run = l_run * fast_exp2(m_run - m_new) + ((s0 + s1) + (s2 + s3));
                m_run = m_new;
              }
            }
            if (c == BN / 64 - 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) release_acc<CG>(tempty_bar(acc));
            }
            const uint32_t buf = smem_epi0 + (chunk_ctr & 1u) * EPI_BUF_BYTES;
            if (issuer) tma_wait_group_read<1>();
            named_bar_sync(1, EPI_THREADS);
            stage_row_128B(buf, row_in_tile, w);
            fence_proxy_async_smem();
            named_bar_sync(1, EPI_THREADS);
            if (issuer && m_ok) {
              const int cc = p.c_c0 + nt * BN + c * 64, cr = p.c_r0 + sc.c_row_off + m_tile * BM;
              if (p.store_policy)
                tma_store_2d_hint(&p.tmC, buf, cc, cr, p.store_policy);
              else
                tma_store_2d(&p.tmC, buf, cc, cr);
              tma_commit_group();
            }
            ++chunk_ctr;
          }
        } else if constexpr (EPI == EPI_EXP_STASH) {
          // Single-pass mode, default: the logits tile is rounded to bf16 (the reference's logits dtype,
          // tied_linear.py:25), E = 2^(s log2e - ref2[row]) is formed in fp32 from the ROUNDED value (what
          // ce_chunked_output_loss.py:41 upcasts), summed per row in fp32 and stored as bf16 into the row chunk's
          // block.  ref2 is a per-row constant (the pilot's max over the first vocabulary tile), so no running max
          // and no rescaling: softmax = E / sum(E) is applied later as a per-row factor.  Padding columns store 0.
          const int cbase = p.col0 + nt * BN;
#pragma unroll 1
          for (int c = 0; c < BN / 64; ++c) {
            uint32_t w[32];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              float v[32];
              tmem_ld32(taddr + c * 64 + h * 32, v);
              const int ch0 = cbase + c * 64 + h * 32;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const uint32_t pk = pack_bf16x2(v[2 * j], v[2 * j + 1]);
                v[2 * j] = bf16lo(pk);
                v[2 * j + 1] = bf16hi(pk);
              }
              if (valid && label >= ch0 && label < ch0 + 32) {
                const int jl = static_cast<int>(label - ch0);
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (j == jl) z_lab = v[j];
              }
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = fast_exp2(fmaf(v[j], LOG2E, -ref2));
              if (ch0 + 32 > p.vocab) {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (ch0 + j >= p.vocab) v[j] = 0.f;
              }
              float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
              for (int j = 0; j < 32; j += 4) s0 += v[j], s1 += v[j + 1], s2 += v[j + 2], s3 += v[j + 3];
              l_run += (s0 + s1) + (s2 + s3);
              // The label's own exponential stays in the row sum but is NOT stashed: its share of the gradient,
              // g (p_label - 1), nearly cancels for a confident row and would lose its leading bits to the bf16
              // rounding of E.  exp_combine_kernel forms 1 - p_label in fp32 and the label-keyed passes apply it.
              if (valid && label >= ch0 && label < ch0 + 32) {
                const int jl = static_cast<int>(label - ch0);
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (j == jl) v[j] = 0.f;
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) w[h * 16 + j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
            }
            if (c == BN / 64 - 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) release_acc<CG>(tempty_bar(acc));
            }
            const uint32_t buf = smem_epi0 + (chunk_ctr & 1u) * EPI_BUF_BYTES;
            if (issuer) tma_wait_group_read<1>();
            named_bar_sync(1, EPI_THREADS);
            stage_row_128B(buf, row_in_tile, w);
            fence_proxy_async_smem();
            named_bar_sync(1, EPI_THREADS);
            if (issuer && m_ok) {
              const int cc = p.c_c0 + nt * BN + c * 64, cr = p.c_r0 + sc.c_row_off + m_tile * BM;
              if (p.store_policy)
                tma_store_2d_hint(&p.tmC, buf, cc, cr, p.store_policy);
              else
                tma_store_2d(&p.tmC, buf, cc, cr);
              tma_commit_group();
            }
            ++chunk_ctr;
          }
        } else if constexpr (EPI == EPI_GRAD) {
          const int cbase = p.col0 + nt * BN;  // global vocab index of tile column 0
#pragma unroll 1
          for (int c = 0; c < BN / 64; ++c) {
            const int c0 = cbase + c * 64;
            uint32_t w[32];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              float v[32];
              tmem_ld32(taddr + c * 64 + h * 32, v);
              const int ch0 = c0 + h * 32;
#pragma unroll
              for (int j = 0; j < 16; ++j) {  // bf16 logits, as in forward
                const uint32_t pk = pack_bf16x2(v[2 * j], v[2 * j + 1]);
                v[2 * j] = bf16lo(pk);
                v[2 * j + 1] = bf16hi(pk);
              }
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                float g = valid ? fast_exp2(fmaf(v[j], LOG2E, -lse2)) * gscale : 0.f;
                if (ch0 + j >= p.vocab) g = 0.f;
                v[j] = g;
              }
              if (valid && label >= ch0 && label < ch0 + 32) {
                const int jl = static_cast<int>(label - ch0);
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (j == jl) v[j] -= gscale;
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) w[h * 16 + j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
            }
            if (c == BN / 64 - 1) {  // accumulator fully read: release it to the MMA warp
              tc_fence_before();
              __syncwarp();
              if (lane == 0) release_acc<CG>(tempty_bar(acc));
            }
            const uint32_t buf = smem_epi0 + (chunk_ctr & 1u) * EPI_BUF_BYTES;
            if (issuer) tma_wait_group_read<1>();
            named_bar_sync(1, EPI_THREADS);
            stage_row_128B(buf, row_in_tile, w);
            fence_proxy_async_smem();
            named_bar_sync(1, EPI_THREADS);
            if (issuer && m_ok) {
              if (p.store_policy)
                tma_store_2d_hint(&p.tmC, buf, p.c_c0 + nt * BN + c * 64, p.c_r0 + m_tile * BM, p.store_policy);
              else
                tma_store_2d(&p.tmC, buf, p.c_c0 + nt * BN + c * 64, p.c_r0 + m_tile * BM);
              tma_commit_group();
            }
            ++chunk_ctr;
          }
        } else if constexpr (EPI == EPI_BF16) {
#pragma unroll 1
          for (int c = 0; c < nh * (BN / 64); ++c) {
            uint32_t w[32];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              float v[32];
              tmem_ld32(taddr + c * 64 + h * 32, v);
              if (zero_acc) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0.f;
              }
              if (p.row_scale != nullptr) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] *= rscale;
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) w[h * 16 + j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
            }
            if (c == nh * (BN / 64) - 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) release_acc<CG>(tempty_bar(acc));
            }
            const uint32_t buf = smem_epi0 + (chunk_ctr & 1u) * EPI_BUF_BYTES;
            if (issuer) tma_wait_group_read<1>();
            named_bar_sync(1, EPI_THREADS);
            stage_row_128B(buf, row_in_tile, w);
            fence_proxy_async_smem();
            named_bar_sync(1, EPI_THREADS);
            if (issuer && m_ok) {
              const int cc = p.c_c0 + nt * BNW + h0 * BN + c * 64, cr = p.c_r0 + sc.c_row_off + m_tile * BM;
              if (sc.c_reduce)
                tma_reduce_add_2d(&p.tmC, buf, cc, cr);
              else if (p.store_policy)
                tma_store_2d_hint(&p.tmC, buf, cc, cr, p.store_policy);
              else
                tma_store_2d(&p.tmC, buf, cc, cr);
              tma_commit_group();
            }
            ++chunk_ctr;
          }
        } else {  // EPI_F32
#pragma unroll 1
          for (int c = 0; c < nh * (BN / 32); ++c) {
            uint32_t w[32];
            {
              float v[32];
              tmem_ld32(taddr + c * 32, v);
#pragma unroll
              for (int j = 0; j < 32; ++j) w[j] = zero_acc ? 0u : __float_as_uint(v[j]);
            }
            if (c == nh * (BN / 32) - 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) release_acc<CG>(tempty_bar(acc));
            }
            const uint32_t buf = smem_epi0 + (chunk_ctr & 1u) * EPI_BUF_BYTES;
            if (issuer) tma_wait_group_read<1>();
            named_bar_sync(1, EPI_THREADS);
            stage_row_128B(buf, row_in_tile, w);
            fence_proxy_async_smem();
            named_bar_sync(1, EPI_THREADS);
            if (issuer && m_ok) {
              const int cc = p.c_c0 + nt * BNW + h0 * BN + c * 32, cr = p.c_r0 + sc.c_row_off + m_tile * BM;
              if (sc.c_reduce)
                tma_reduce_add_2d(&p.tmC, buf, cc, cr);
              else if (p.store_policy)
                tma_store_2d_hint(&p.tmC, buf, cc, cr, p.store_policy);
              else
                tma_store_2d(&p.tmC, buf, cc, cr);
              tma_commit_group();
            }
            ++chunk_ctr;
          }
        }
        if (NW == 2 || ++acc == 2) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }

      if constexpr (is_lse_epi<EPI>()) {
        // item.n0 / vtiles_per_split is the split index
        const int split = it.n0 / sc.vtiles_per_split;
        if (m_ok) {
          const size_t o = static_cast<size_t>(split) * sc.part_stride + row;
          p.part_m[o] = m_run;
          p.part_l[o] = l_run;
          p.part_z[o] = z_lab;
        }
      }
    }
    if constexpr (EPI != EPI_LSE) {
      if (issuer) tma_wait_group<0>();
    }
  }

  // ---- teardown (CG 2: neither CTA may leave while its peer can still touch its smem / barriers / TMEM)
  tc_fence_before();
  if constexpr (CG == 2) {
    cluster_arrive();
    cluster_wait();
  } else {
    __syncthreads();
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<CG>(tmem_base, TMEM_COLS);
  }
}

}  // namespace ssi
