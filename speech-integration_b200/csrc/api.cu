// C ABI of libssi_b200.so (see include/ssi_b200.h).  Host side only builds TMA descriptors, picks the
// tile schedule and enqueues kernels on the caller's stream: no allocation, no synchronisation.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/ssi_b200.h"
#include "aux_kernels.cuh"
#include "dp_exchange.cuh"
#include "gemm.cuh"
#include "next_rows.cuh"
#include "scatter.cuh"

#include <nvtx3/nvToolsExt.h>

namespace {

thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

// ---- optional per-kernel-class timing with CUDA events on the launching stream (bench.py's roofline leg)
enum ProfCat { CAT_LSE = 0, CAT_DLOGITS = 1, CAT_DW = 2, CAT_DX = 3, CAT_SCATTER = 4, CAT_AUX = 5, CAT_SOFTMAX = 6,
               CAT_EXCHANGE = 7, CAT_COUNT = 8 };
struct ProfRec {
  int cat;
  cudaEvent_t a, b;
};
struct ProfState {
  std::mutex mu;
  bool on = false;
  std::vector<ProfRec> recs;
  std::vector<cudaEvent_t> pool;
  size_t used = 0;
  cudaEvent_t get() {
    if (used == pool.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      pool.push_back(e);
    }
    return pool[used++];
  }
};
ProfState g_prof;
const char* const kCatNames[CAT_COUNT] = {"ssi_b200:K1 logits+LSE", "ssi_b200:K2a dlogits", "ssi_b200:K2b dW",
                                          "ssi_b200:K2c dX",        "ssi_b200:K3 embedding scatter",
                                          "ssi_b200:aux",           "ssi_b200:softmax-grad / onehot term",
                                          "ssi_b200:K4 exchange"};
struct ProfScope {
  cudaStream_t st;
  cudaEvent_t b = nullptr;
  bool nvtx = false;
  ProfScope(int cat, cudaStream_t s) : st(s) {
    if (cat >= 0 && cat < CAT_COUNT) {  // NVTX range around the enqueue (free when no tool is attached)
      nvtxRangePushA(kCatNames[cat]);
      nvtx = true;
    }
    if (!g_prof.on || cat < 0) return;
    std::lock_guard<std::mutex> lk(g_prof.mu);
    cudaEvent_t a = g_prof.get();
    b = g_prof.get();
    g_prof.recs.push_back({cat, a, b});
    cudaEventRecord(a, st);
  }
  ~ProfScope() {
    if (b) cudaEventRecord(b, st);
    if (nvtx) nvtxRangePop();
  }
};

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define SSI_CHECK_ARG(cond, ...) \
  do {                           \
    if (!(cond)) return fail(-1, __VA_ARGS__); \
  } while (0)

#define SSI_CUDA(expr)                                                                          \
  do {                                                                                          \
    cudaError_t e_ = (expr);                                                                    \
    if (e_ != cudaSuccess) return fail(static_cast<int>(e_), "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

#define SSI_LAUNCH_OK(name)                                                                      \
  do {                                                                                           \
    cudaError_t e_ = cudaGetLastError();                                                         \
    if (e_ != cudaSuccess) return fail(static_cast<int>(e_), "launch %s: %s", name, cudaGetErrorString(e_)); \
    g_launches.fetch_add(1, std::memory_order_relaxed);                                          \
  } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

struct DeviceInfo {
  int ok = 0;  // 1 = sm_100, -1 = unusable
  int dev = 0;
  int sms = 0;
  int cc_major = 0, cc_minor = 0;
  char why[256] = "";
};

int device_info(DeviceInfo** out) {
  static DeviceInfo infos[64];
  static std::mutex mu;
  int dev = 0;
  SSI_CUDA(cudaGetDevice(&dev));
  SSI_CHECK_ARG(dev >= 0 && dev < 64, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(mu);
  DeviceInfo& d = infos[dev];
  if (d.ok == 0) {
    cudaDeviceProp prop;
    SSI_CUDA(cudaGetDeviceProperties(&prop, dev));
    d.dev = dev;
    d.sms = prop.multiProcessorCount;
    d.cc_major = prop.major;
    d.cc_minor = prop.minor;
    if (prop.major != 10) {
      d.ok = -1;
      snprintf(d.why, sizeof(d.why), "device %d is sm_%d%d; ssi_b200 kernels are built for sm_100a only (no fallback)",
               dev, prop.major, prop.minor);
    } else {
      d.ok = 1;
    }
  }
  if (d.ok < 0) return fail(-2, "%s", d.why);
  *out = &d;
  return 0;
}

PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

// 2D row-major tensor [outer, inner] of `esize`-byte elements, 128-byte swizzle, OOB reads -> zeros.
int make_tmap(CUtensorMap* m, CUtensorMapDataType dt, int esize, const void* ptr, uint64_t inner, uint64_t outer,
              uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_outer) {
  auto enc = get_encode();
  if (!enc) return fail(-3, "cuTensorMapEncodeTiled entry point not available");
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "tensor base %p is not 16-byte aligned", ptr);
  SSI_CHECK_ARG(row_stride_bytes % 16 == 0, "row stride %llu B is not a multiple of 16",
                static_cast<unsigned long long>(row_stride_bytes));
  SSI_CHECK_ARG(box_inner * esize == 128 && box_outer <= 256, "bad TMA box %u x %u", box_inner, box_outer);
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(-4, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return 0;
}

int cta_group();       // tuning knobs, defined below
int tail_split_opt();
int shape_gate_opt();

template <int EPI, bool A_MN, bool B_MN, int CG, int NW = 1>
int launch_gemm_cg(const ssi::GemmParams& p, int sms, cudaStream_t st, const char* name, int cat, int grid_override,
                   bool prof = true) {
  auto kern = ssi::gemm_kernel<EPI, A_MN, B_MN, CG, NW>;
  // cudaFuncSetAttribute and the occupancy answer are per DEVICE: cache them per (instantiation, device ordinal)
  struct PerDev {
    bool attr_set = false;
    int max_workers = 0;  // CTAs (CG 1) or CTA pairs (CG 2) that can be co-resident
  };
  static PerDev per_dev[64];
  static std::mutex mu;
  int dev = 0;
  SSI_CUDA(cudaGetDevice(&dev));
  SSI_CHECK_ARG(dev >= 0 && dev < 64, "device ordinal %d out of range", dev);
  int max_workers = 0;
  {
    std::lock_guard<std::mutex> lk(mu);
    PerDev& pd = per_dev[dev];
    if (!pd.attr_set) {
      SSI_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, ssi::GEMM_SMEM_BYTES));
      pd.max_workers = sms / CG;
      if (CG == 2) {
        cudaLaunchConfig_t qc = {};
        qc.gridDim = dim3(sms / 2 * 2), qc.blockDim = dim3(ssi::GEMM_THREADS);
        qc.dynamicSmemBytes = ssi::GEMM_SMEM_BYTES;
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeClusterDimension;
        qa[0].val.clusterDim.x = 2, qa[0].val.clusterDim.y = 1, qa[0].val.clusterDim.z = 1;
        qc.attrs = qa, qc.numAttrs = 1;
        int n = 0;
        SSI_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &qc));
        if (n < 1) return fail(-5, "%s: no CTA pair can be resident (cudaOccupancyMaxActiveClusters = %d)", name, n);
        if (n < pd.max_workers) pd.max_workers = n;
      }
      pd.attr_set = true;
    }
    max_workers = pd.max_workers;
  }
  const int tiles_m = (CG == 2) ? (p.tiles_m + 1) / 2 : p.tiles_m;
  const int items = tiles_m * ((p.tiles_n + NW - 1) / NW);  // host-side upper bound when the schedule is device-side
  if (items <= 0 && grid_override <= 0) return 0;
  int workers = grid_override > 0 ? max_workers : (items < max_workers ? items : max_workers);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(workers * CG), cfg.blockDim = dim3(ssi::GEMM_THREADS);
  cfg.dynamicSmemBytes = ssi::GEMM_SMEM_BYTES, cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CG, at[0].val.clusterDim.y = 1, at[0].val.clusterDim.z = 1;
  cfg.attrs = at, cfg.numAttrs = 1;
  {
    ProfScope ps(prof ? cat : -1, st);
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
    if (e != cudaSuccess) return fail(static_cast<int>(e), "launch %s: %s", name, cudaGetErrorString(e));
  }
  SSI_LAUNCH_OK(name);
  return 0;
}

template <int EPI, bool A_MN, bool B_MN>
int launch_gemm(const ssi::GemmParams& p, int sms, cudaStream_t st, const char* name, int cat,
                int grid_override = 0) {
  if (cta_group() == 2) return launch_gemm_cg<EPI, A_MN, B_MN, 2>(p, sms, st, name, cat, grid_override);
  return launch_gemm_cg<EPI, A_MN, B_MN, 1>(p, sms, st, name, cat, grid_override);
}

// Long-K contractions (single-pass dW / dX): 256x512 tiles per CTA pair when `wide` == 2 ("wide_dw" / "wide_dx").
template <int EPI, bool A_MN, bool B_MN>
int launch_gemm_wide(const ssi::GemmParams& p, int sms, cudaStream_t st, const char* name, int cat, int wide) {
  if (cta_group() == 2 && wide == 2) {
    ssi::GemmParams q = p;
    q.tail_split = tail_split_opt();
    if (!shape_gate_opt()) return launch_gemm_cg<EPI, A_MN, B_MN, 2, 2>(q, sms, st, name, cat, 0);
    // both tile shapes are enqueued; the device-side plan decides which one runs (GemmParams::shape_gate)
    q.shape_gate = 1, q.gate_workers = sms / 2;
    ProfScope ps(cat, st);  // ONE timing record for the pair: only one of the two kernels does the work
    if (int r = launch_gemm_cg<EPI, A_MN, B_MN, 2, 2>(q, sms, st, name, cat, 0, false)) return r;
    q.shape_gate = 2;
    return launch_gemm_cg<EPI, A_MN, B_MN, 2, 1>(q, sms, st, name, cat, 0, false);
  }
  return launch_gemm<EPI, A_MN, B_MN>(p, sms, st, name, cat);
}

// ---------------------------------------------------------------------------------------------------
// Schedules
// ---------------------------------------------------------------------------------------------------
// Workspace layouts.  Sizes depend on (N, V, D) only; everything that depends on the number of VALID rows is
// decided on the device (ssi::DevPlan).
struct CompactWs {  // shared head of the forward and backward workspaces
  size_t off_plan, off_pos, off_idx, off_labels_c, off_xc, end;
};
CompactWs plan_compact(int64_t N, int64_t n_pad, int64_t D, size_t o) {
  CompactWs c;
  c.off_plan = o, o += ssi::PLAN_AREA_BYTES;
  c.off_pos = o, o += align_up(static_cast<size_t>(N) * sizeof(int), 256);
  c.off_idx = o, o += align_up(static_cast<size_t>(n_pad) * sizeof(int), 256);
  c.off_labels_c = o, o += align_up(static_cast<size_t>(n_pad) * sizeof(int64_t), 256);
  c.off_xc = o, o += align_up(static_cast<size_t>(n_pad) * D * 2, 1024);
  c.end = o;
  return c;
}

struct FwdPlan {
  int64_t n_pad;
  int tiles_m, vtiles_total, combine_blocks;
  CompactWs cw;
  size_t off_m, off_l, off_z, off_bl, off_bc, off_sync, bytes;
};

FwdPlan plan_fwd(int64_t N, int64_t V, int64_t D) {
  FwdPlan f;
  f.n_pad = ceil_div(N, ssi::BM) * ssi::BM;
  f.tiles_m = static_cast<int>(f.n_pad / ssi::BM);
  f.vtiles_total = static_cast<int>(ceil_div(V, ssi::BN));
  f.combine_blocks = static_cast<int>(ceil_div(N, ssi::COMBINE_THREADS));
  f.cw = plan_compact(N, f.n_pad, D, 0);
  size_t o = f.cw.end;
  // per-split partials: the device plan keeps tiles_m * splits <= max(MAX_LSE_ITEMS, tiles_m)
  const size_t items = f.tiles_m > ssi::MAX_LSE_ITEMS ? f.tiles_m : ssi::MAX_LSE_ITEMS;
  const size_t part = align_up(items * ssi::BM * sizeof(float), 256);
  f.off_m = o, o += part;
  f.off_l = o, o += part;
  f.off_z = o, o += part;
  f.off_bl = o, o += align_up(static_cast<size_t>(f.combine_blocks) * sizeof(float), 256);
  f.off_bc = o, o += align_up(static_cast<size_t>(f.combine_blocks) * sizeof(int), 256);
  f.off_sync = o, o += 256;  // lockstep counter of the forward-only K1
  f.bytes = o;
  return f;
}

// Tuning knobs (environment, read once): for experiments from bench.py without rebuilding.
struct Tune {
  // logits / dlogits block scratch cap: 4.5 GiB of bf16 -- the 16,384 rows a rank holds when the 64 x 2048 batch
  // is split over 8 GPUs fit ONE row chunk at V = 133,258 (then the whole dX contraction hides the exchange)
  int64_t g_cap_elems = int64_t(4608) << 19;
  int64_t g_tiles = 0;                     // force the block width (256-column tiles); 0 = auto
  int group_m_grad = 16, group_m_dw = 1, group_m_dx = 1;  // m-tiles per raster group (1 = n-fastest)
  int store_hint = 0;  // 1 = evict-first L2 policy on the dlogits / dW / dX tile stores
  int k1_hints = 0, dw_hints = 0, dx_hints = 0;  // operand L2 policies: 10 * A + B; 1 evict-first, 2 evict-last
  int lockstep = 64;   // k-blocks between lockstep sync points of the dW / dX contractions (0 = off)
  int cta_group = 2;   // 2 = CTA pairs (tcgen05 cta_group::2, 256x256 tiles), 1 = single-CTA 128x256 tiles
  int lockstep_k1 = 1;  // single-pass K1: lockstep of the vocabulary sweep (one sync point per tile); 0 = off
  int shape_gate = 1;  // wide dW / dX: also enqueue the 256x256 kernel and let the device-side plan pick (small problems)
  int tail_split = 1;  // wide tiles: split the items of an at most half-full last wave into 256x256 halves
  int wide_dw = 2, wide_dx = 2;  // single-pass dW / dX: 2 = 256x512 tiles per CTA pair (needs cta_group 2), 1 = 256x256
  // single-pass: 1 = K1 stashes E = exp(logit - ref[row]) and the softmax normalisation is a per-row factor folded into
  // the dX epilogue and the X rows of dW (no pass over the block); 0 = stash logits + separate softmax-grad HBM pass
  int exp_stash = 1;
  int sparse_two_phase = -1;  // sparse (embedding) exchange: 1 = owner-reduce + apply, 0 = one phase, -1 = auto (world > 4)
  // single-pass K1: m-tiles (128 rows) per raster group, 0 = the whole row chunk (GemmParams::lse_group_m), and
  // evict-first L2 policy on its E-tile stores.  Together they keep the X rows of a group L2-resident next to the
  // 4.35 GB write stream: DRAM reads per 16k-row chunk 4.3 -> 2.0 GB, step -0.75 % (profiles/r02_ab_k1_l2hints*.txt)
  int k1_group_m = 64;
  int k1_store_first = 1;
  // data-parallel exchange: seconds a device-side wait for a peer may last before the step is declared lost (all
  // ranks poisoned); read when an exchange buffer is allocated or reset.  Default = torch.distributed's 10 minutes.
  int64_t dp_timeout_s = 600;
};
// The knobs are process-global and may be changed at run time (ssi_b200_set_option): every API entry takes a
// thread-local SNAPSHOT under the mutex (TuneScope), so one call never sees half of an update.
std::mutex g_tune_mu;
Tune& global_tune() {
  static Tune t = [] {
    Tune x;
    if (const char* e = getenv("SSI_B200_G_CAP_MIB")) x.g_cap_elems = atoll(e) * (int64_t(1) << 19);
    if (const char* e = getenv("SSI_B200_G_TILES")) x.g_tiles = atoll(e);
    if (const char* e = getenv("SSI_B200_GROUP_M_GRAD")) x.group_m_grad = atoi(e);
    if (const char* e = getenv("SSI_B200_GROUP_M_DW")) x.group_m_dw = atoi(e);
    if (const char* e = getenv("SSI_B200_GROUP_M_DX")) x.group_m_dx = atoi(e);
    if (const char* e = getenv("SSI_B200_STORE_HINT")) x.store_hint = atoi(e);
    if (const char* e = getenv("SSI_B200_CTA_GROUP")) x.cta_group = atoi(e) == 1 ? 1 : 2;
    if (const char* e = getenv("SSI_B200_K1_HINTS")) x.k1_hints = atoi(e);
    if (const char* e = getenv("SSI_B200_LOCKSTEP")) x.lockstep = atoi(e);
    if (const char* e = getenv("SSI_B200_DW_HINTS")) x.dw_hints = atoi(e);
    if (const char* e = getenv("SSI_B200_DX_HINTS")) x.dx_hints = atoi(e);
    if (const char* e = getenv("SSI_B200_LOCKSTEP_K1")) x.lockstep_k1 = atoi(e) ? 1 : 0;
    if (const char* e = getenv("SSI_B200_SHAPE_GATE")) x.shape_gate = atoi(e) ? 1 : 0;
    if (const char* e = getenv("SSI_B200_TAIL_SPLIT")) x.tail_split = atoi(e) ? 1 : 0;
    if (const char* e = getenv("SSI_B200_WIDE_DW")) x.wide_dw = atoi(e) == 1 ? 1 : 2;
    if (const char* e = getenv("SSI_B200_WIDE_DX")) x.wide_dx = atoi(e) == 1 ? 1 : 2;
    if (const char* e = getenv("SSI_B200_EXP_STASH")) x.exp_stash = atoi(e) ? 1 : 0;
    if (const char* e = getenv("SSI_B200_K1_GROUP_M")) x.k1_group_m = atoi(e);
    if (const char* e = getenv("SSI_B200_K1_STORE_FIRST")) x.k1_store_first = atoi(e) ? 1 : 0;
    if (const char* e = getenv("SSI_B200_SPARSE_TWO_PHASE")) x.sparse_two_phase = atoi(e);
    if (const char* e = getenv("SSI_B200_DP_TIMEOUT_S")) x.dp_timeout_s = atoll(e) > 0 ? atoll(e) : 600;
    return x;
  }();
  return t;
}
thread_local Tune tl_tune;
struct TuneScope {
  TuneScope() {
    std::lock_guard<std::mutex> lk(g_tune_mu);
    tl_tune = global_tune();
  }
};
Tune& tune() { return tl_tune; }  // the snapshot of the running API call

int cta_group() { return tune().cta_group; }
int tail_split_opt() { return tune().tail_split; }
int shape_gate_opt() { return tune().shape_gate; }
unsigned long long policy_of(int code) {
  return code == 1 ? ssi::kEvictFirst : code == 2 ? ssi::kEvictLast : code == 3 ? ssi::kEvictNormal : 0ull;
}
void set_hints(ssi::GemmParams& p, int hints) {
  p.a_policy = policy_of(hints / 10);
  p.b_policy = policy_of(hints % 10);
}

struct BwdPlan {
  int64_t n_pad;
  int tiles_m, vtiles_total, nblocks, vtiles_per_block;
  int64_t vc;  // columns of the dlogits block buffer
  CompactWs cw;
  size_t off_lse_c, off_g, off_dxacc, off_sync, bytes;
};

// SM count of the current device for host-side schedules (148 on B200; queried, not assumed)
int current_sms() {
  DeviceInfo* di = nullptr;
  if (device_info(&di) == 0 && di->sms > 0) return di->sms;
  return 148;  // no usable device in this process (size queries on a build box): the B200 figure
}

BwdPlan plan_bwd(int64_t N, int64_t V, int64_t D) {
  BwdPlan b;
  b.n_pad = ceil_div(N, ssi::BM) * ssi::BM;
  b.tiles_m = static_cast<int>(b.n_pad / ssi::BM);
  b.vtiles_total = static_cast<int>(ceil_div(V, ssi::BN));
  // Block width t (in 256-column tiles): as wide as the dlogits scratch cap allows, chosen so that the dW
  // contraction of a block (2t x ceil(D/256) output tiles on 148 persistent CTAs) wastes the fewest waves;
  // a small per-block charge prefers fewer, longer launches.
  const int64_t cap_elems = tune().g_cap_elems;
  int64_t t_max = cap_elems / b.n_pad / ssi::BN;
  if (t_max < 4) t_max = 4;
  if (t_max > b.vtiles_total) t_max = b.vtiles_total;
  const int d_tiles = static_cast<int>(ceil_div(D, ssi::BN));
  const int sms = current_sms();
  int64_t best_t = t_max;
  double best_cost = -1.0;
  for (int64_t t = (tune().g_tiles > 0 ? tune().g_tiles : 4); t <= t_max; ++t) {
    const int64_t full = b.vtiles_total / t, rem = b.vtiles_total % t;
    double waves = static_cast<double>(full) * ceil_div(2 * t * d_tiles, sms);
    if (rem) waves += ceil_div(2 * rem * d_tiles, sms);
    const double cost = waves + 0.25 * (full + (rem ? 1 : 0));
    if (best_cost < 0 || cost <= best_cost) best_cost = cost, best_t = t;
    if (tune().g_tiles > 0) break;
  }
  b.vtiles_per_block = static_cast<int>(best_t);
  b.nblocks = static_cast<int>(ceil_div(b.vtiles_total, best_t));
  b.vc = static_cast<int64_t>(b.vtiles_per_block) * ssi::BN;
  b.cw = plan_compact(N, b.n_pad, D, 0);
  size_t o = b.cw.end;
  b.off_lse_c = o, o += align_up(static_cast<size_t>(b.n_pad) * sizeof(float), 256);
  b.off_g = o, o += align_up(static_cast<size_t>(b.n_pad) * b.vc * 2, 1024);
  b.off_dxacc = o, o += align_up(static_cast<size_t>(b.n_pad) * D * sizeof(float), 1024);
  b.off_sync = o, o += 4096;  // lockstep counters of the dW / dX contractions (two per vocabulary block) + error word
  b.bytes = o;
  return b;
}

struct ScatPlan {
  int nwarps, passes, slabs;
  size_t off_k0, off_v0, off_k1, off_v1, off_urank, off_nuniq, off_ghist, off_partial, off_info, bytes;
};

ScatPlan plan_scatter(int64_t N, int64_t V, int64_t D) {
  ScatPlan s;
  s.nwarps = static_cast<int>(ceil_div(N, 32));
  s.slabs = static_cast<int>(ceil_div(D, ssi::SCAT_COLS));
  int bits = 1;
  while ((int64_t(1) << bits) <= V) ++bits;  // keys are in [0, V] (V = out-of-range sentinel)
  s.passes = (bits + ssi::SORT_BITS - 1) / ssi::SORT_BITS;
  size_t o = 0;
  const size_t arr = align_up(static_cast<size_t>(N) * 4, 256);
  s.off_k0 = o, o += arr;
  s.off_v0 = o, o += arr;
  s.off_k1 = o, o += arr;
  s.off_v1 = o, o += arr;
  s.off_urank = o, o += arr;
  s.off_nuniq = o, o += 256;  // [0] number of distinct ids, [16] arrival counter of the sort's grid barrier
  s.off_ghist = o, o += static_cast<size_t>(ssi::SORT_DIGITS) * ssi::SORT_MAX_BLOCKS * sizeof(uint32_t);
  s.off_partial = o, o += align_up(static_cast<size_t>(s.nwarps) * 2 * D * sizeof(float), 256);
  s.off_info = o, o += align_up(static_cast<size_t>(s.nwarps) * sizeof(ssi::ScatInfo), 256);
  s.bytes = o;
  return s;
}

template <int MODE>
int embedding_grad_scatter_impl(const void* dH, const int64_t* ids, int64_t N, int64_t V, int64_t D,
                                const ssi::ScatOut& out, void* ws, size_t ws_bytes, void* stream, int phases,
                                int cat = CAT_SCATTER);
ssi::ScatOut dense_out(void* dW);

// Single-pass ("stash") mode: forward also forms the gradients, one ROW chunk at a time, so the logits are
// computed once (6 N V D executed instead of 8 N V D).  Scratch: one bf16 block S/G [rows_cap, Vpad].
struct SpPlan {
  int64_t n_pad, vpad;
  int tiles_m, vtiles_total, combine_blocks, t_max, chunks_max;
  CompactWs cw;
  size_t off_m, off_l, off_z, off_lse_c, off_loss_c, off_bl, off_bc, off_sync, off_rstd, off_dsp, off_dxc, off_s, bytes;
  // exp-stash form: pilot reference, per-row factor, scaled X rows, workspace of the label-keyed onehot scatter
  size_t off_ref2, off_rowscale, off_labw, off_xs, off_lscat, lscat_bytes;
};
constexpr size_t SYNC_AREA_BYTES = 4096;
constexpr int NORM_BWD_BLOCKS = 148 * 2;            // grid of the fused norm-backward scatter
constexpr int NORM_BWD_PARTS = NORM_BWD_BLOCKS * 8;  // one dscale partial row per warp

int plan_single_pass(int64_t N, int64_t V, int64_t D, SpPlan* out) {
  SpPlan f;
  f.n_pad = ceil_div(N, ssi::BM) * ssi::BM;
  f.tiles_m = static_cast<int>(f.n_pad / ssi::BM);
  f.vtiles_total = static_cast<int>(ceil_div(V, ssi::BN));
  f.vpad = static_cast<int64_t>(f.vtiles_total) * ssi::BN;
  f.combine_blocks = static_cast<int>(ceil_div(N, ssi::COMBINE_THREADS));
  int64_t t_max = tune().g_cap_elems / f.vpad / ssi::BM;
  if (t_max < 1) t_max = 1;
  if (t_max > f.tiles_m) t_max = f.tiles_m;
  f.t_max = static_cast<int>(t_max);
  f.chunks_max = static_cast<int>(ceil_div(f.tiles_m, t_max)) + (t_max < f.tiles_m ? 1 : 0);
  if (f.chunks_max > ssi::MAX_ROW_CHUNKS)
    return fail(-1, "single-pass mode needs %d row chunks (> %d): raise SSI_B200_G_CAP_MIB", f.chunks_max,
                ssi::MAX_ROW_CHUNKS);
  f.cw = plan_compact(N, f.n_pad, D, 0);
  size_t o = f.cw.end;
  const size_t items = f.t_max > ssi::MAX_LSE_ITEMS ? f.t_max : ssi::MAX_LSE_ITEMS;
  const size_t part = align_up(items * ssi::BM * sizeof(float), 256);
  f.off_m = o, o += part;
  f.off_l = o, o += part;
  f.off_z = o, o += part;
  f.off_lse_c = o, o += align_up(static_cast<size_t>(f.n_pad) * sizeof(float), 256);
  f.off_loss_c = o, o += align_up(static_cast<size_t>(f.n_pad) * sizeof(float), 256);
  f.off_bl = o, o += align_up(static_cast<size_t>(f.combine_blocks) * sizeof(float), 256);
  f.off_bc = o, o += align_up(static_cast<size_t>(f.combine_blocks) * sizeof(int), 256);
  f.off_sync = o, o += SYNC_AREA_BYTES;
  // fused final RMSNorm (fwd_grad_norm): per-row rstd of the compact rows, per-warp partial rows of dscale
  f.off_rstd = o, o += align_up(static_cast<size_t>(f.n_pad) * sizeof(float), 256);
  f.off_dsp = o, o += align_up(static_cast<size_t>(NORM_BWD_PARTS) * D * sizeof(float), 256);
  f.off_dxc = o, o += align_up(static_cast<size_t>(f.n_pad) * D * 2, 1024);
  f.off_ref2 = o, o += align_up(items * ssi::BM * sizeof(float), 256);
  f.off_rowscale = o, o += align_up(static_cast<size_t>(f.n_pad) * sizeof(float), 256);
  f.off_labw = o, o += align_up(static_cast<size_t>(f.n_pad) * sizeof(float), 256);  // 1 - p_label per compact row
  f.off_xs = o, o += align_up(static_cast<size_t>(f.n_pad) * D * 2, 1024);
  f.lscat_bytes = plan_scatter(f.n_pad, V, D).bytes;
  f.off_lscat = o, o += align_up(f.lscat_bytes, 1024);
  f.off_s = o, o += align_up(static_cast<size_t>(f.t_max) * ssi::BM * f.vpad * 2, 1024);
  f.bytes = o;
  *out = f;
  return 0;
}

int check_common(const void* X, const void* W, int64_t N, int64_t V, int64_t D) {
  SSI_CHECK_ARG(X && W, "null X/W pointer");
  SSI_CHECK_ARG(N >= 1 && V >= 1 && D >= 64, "bad shape N=%lld V=%lld D=%lld", (long long)N, (long long)V,
                (long long)D);
  SSI_CHECK_ARG(D % 64 == 0, "D=%lld must be a multiple of 64", (long long)D);
  SSI_CHECK_ARG(N < (int64_t(1) << 30) && V < (int64_t(1) << 30), "N/V too large for 32-bit tile coordinates");
  return 0;
}

// scale_grads / grad-norm fold (aux_kernels.cuh GradFold): float partials in three segments -- [0, 4096) the dense
// writer's blocks (apply_dw / dp_reduce: the rank's own slice), [4096, 8192) dp_gather's blocks, [8192, end) the
// warps of the sparse read-modify-write site (embedding scatter / dp_sparse_add)
constexpr int FOLD_SEG = 4096;
struct FoldWs {
  float* base = nullptr;
  size_t floats = 0;
  float* seg(int i) const { return base ? base + static_cast<size_t>(i) * FOLD_SEG : nullptr; }
  size_t sparse_floats() const { return floats > 2 * FOLD_SEG ? floats - 2 * FOLD_SEG : 0; }
};
size_t fold_sparse_floats(int64_t max_rows, int64_t D) {
  const size_t scat = 2 * static_cast<size_t>(ceil_div(max_rows, 32)) * static_cast<size_t>(ceil_div(D, ssi::SCAT_COLS));
  return scat > 2 * FOLD_SEG ? scat : 2 * FOLD_SEG;
}
int fold_ws(void* ws, size_t bytes, FoldWs* out) {
  FoldWs f;
  if (ws) {
    SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 15) == 0, "fold workspace must be 16-byte aligned");
    SSI_CHECK_ARG(bytes >= (2 * FOLD_SEG + 2 * FOLD_SEG) * sizeof(float), "fold workspace too small: %zu", bytes);
    f.base = static_cast<float*>(ws), f.floats = bytes / sizeof(float);
  }
  *out = f;
  return 0;
}

}  // namespace

// ===================================================================================================
extern "C" {

int ssi_b200_version(void) { return SSI_B200_ABI_VERSION; }

const char* ssi_b200_last_error(void) { return g_err; }

int ssi_b200_set_option(const char* name, int64_t value) {
  SSI_CHECK_ARG(name, "null option name");
  std::lock_guard<std::mutex> lk(g_tune_mu);
  Tune& t = global_tune();
  if (!strcmp(name, "g_cap_mib")) {
    SSI_CHECK_ARG(value >= 1, "g_cap_mib must be >= 1");
    t.g_cap_elems = value * (int64_t(1) << 19);
  } else if (!strcmp(name, "g_tiles")) t.g_tiles = value;
  else if (!strcmp(name, "group_m_grad")) t.group_m_grad = static_cast<int>(value);
  else if (!strcmp(name, "group_m_dw")) t.group_m_dw = static_cast<int>(value);
  else if (!strcmp(name, "group_m_dx")) t.group_m_dx = static_cast<int>(value);
  else if (!strcmp(name, "store_hint")) t.store_hint = static_cast<int>(value);
  else if (!strcmp(name, "cta_group")) t.cta_group = (value == 1) ? 1 : 2;
  else if (!strcmp(name, "k1_hints")) t.k1_hints = static_cast<int>(value);
  else if (!strcmp(name, "lockstep")) t.lockstep = static_cast<int>(value);
  else if (!strcmp(name, "dw_hints")) t.dw_hints = static_cast<int>(value);
  else if (!strcmp(name, "dx_hints")) t.dx_hints = static_cast<int>(value);
  else if (!strcmp(name, "lockstep_k1")) t.lockstep_k1 = value ? 1 : 0;
  else if (!strcmp(name, "shape_gate")) t.shape_gate = value ? 1 : 0;
  else if (!strcmp(name, "tail_split")) t.tail_split = value ? 1 : 0;
  else if (!strcmp(name, "wide_dw")) t.wide_dw = (value == 1) ? 1 : 2;
  else if (!strcmp(name, "wide_dx")) t.wide_dx = (value == 1) ? 1 : 2;
  else if (!strcmp(name, "exp_stash")) t.exp_stash = value ? 1 : 0;
  else if (!strcmp(name, "k1_group_m")) t.k1_group_m = value > 0 ? static_cast<int>(value) : 0;
  else if (!strcmp(name, "k1_store_first")) t.k1_store_first = value ? 1 : 0;
  else if (!strcmp(name, "sparse_two_phase")) t.sparse_two_phase = value < 0 ? -1 : (value ? 1 : 0);
  else if (!strcmp(name, "dp_timeout_s")) t.dp_timeout_s = value > 0 ? value : 600;
  else return fail(-1, "unknown option '%s'", name);
  return 0;
}

int ssi_b200_get_option(const char* name, int64_t* value) {
  SSI_CHECK_ARG(name && value, "null pointer argument");
  std::lock_guard<std::mutex> lk(g_tune_mu);
  const Tune& t = global_tune();
  const struct { const char* n; int64_t v; } all[] = {
      {"g_cap_mib", t.g_cap_elems >> 19}, {"g_tiles", t.g_tiles}, {"group_m_grad", t.group_m_grad},
      {"group_m_dw", t.group_m_dw}, {"group_m_dx", t.group_m_dx}, {"store_hint", t.store_hint},
      {"cta_group", t.cta_group}, {"k1_hints", t.k1_hints}, {"lockstep", t.lockstep}, {"dw_hints", t.dw_hints},
      {"dx_hints", t.dx_hints}, {"lockstep_k1", t.lockstep_k1}, {"shape_gate", t.shape_gate},
      {"tail_split", t.tail_split}, {"wide_dw", t.wide_dw}, {"wide_dx", t.wide_dx}, {"exp_stash", t.exp_stash},
      {"k1_group_m", t.k1_group_m}, {"k1_store_first", t.k1_store_first}, {"sparse_two_phase", t.sparse_two_phase},
      {"dp_timeout_s", t.dp_timeout_s}};
  for (const auto& o : all)
    if (!strcmp(name, o.n)) {
      *value = o.v;
      return 0;
    }
  return fail(-1, "unknown option '%s'", name);
}

int64_t ssi_b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int ssi_b200_workspace_bytes(int64_t N, int64_t V, int64_t D, size_t* fwd_bytes, size_t* bwd_bytes,
                             size_t* scatter_bytes) {
  TuneScope tune_scope_;
  SSI_CHECK_ARG(N >= 1 && V >= 1 && D >= 8 && D % 8 == 0, "bad shape N=%lld V=%lld D=%lld", (long long)N,
                (long long)V, (long long)D);
  if (fwd_bytes || bwd_bytes)
    SSI_CHECK_ARG(D % 64 == 0, "D=%lld must be a multiple of 64 for the LM-head kernels", (long long)D);
  if (fwd_bytes) *fwd_bytes = plan_fwd(N, V, D).bytes;
  if (bwd_bytes) *bwd_bytes = plan_bwd(N, V, D).bytes;
  if (scatter_bytes) *scatter_bytes = plan_scatter(N, V, D).bytes;
  return 0;
}

int ssi_b200_lmhead_ce_fwd(const void* X, const void* W, const int64_t* labels, int64_t N, int64_t V, int64_t D,
                           int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* ws,
                           size_t ws_bytes, void* stream) {
  TuneScope tune_scope_;
  if (int r = check_common(X, W, N, V, D)) return r;
  SSI_CHECK_ARG(labels && lse && loss_out && n_valid && ws, "null pointer argument");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const FwdPlan f = plan_fwd(N, V, D);
  SSI_CHECK_ARG(ws_bytes >= f.bytes, "forward workspace too small: %zu < %zu", ws_bytes, f.bytes);
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "workspace must be 256-byte aligned");
  char* wsb = static_cast<char*>(ws);
  ssi::DevPlan* dplan = reinterpret_cast<ssi::DevPlan*>(wsb + f.cw.off_plan);
  int* pos = reinterpret_cast<int*>(wsb + f.cw.off_pos);
  int* idx = reinterpret_cast<int*>(wsb + f.cw.off_idx);
  int64_t* labels_c = reinterpret_cast<int64_t*>(wsb + f.cw.off_labels_c);
  void* Xc = wsb + f.cw.off_xc;

  // ---- compaction of the valid rows + device-side schedule
  {
    ProfScope ps(CAT_AUX, st);
    ssi::compact_scan_kernel<<<1, 1024, 0, st>>>(labels, static_cast<int>(N), static_cast<int>(f.n_pad),
                                                  ignore_index, pos, idx, labels_c, dplan, f.vtiles_total, di->sms,
                                                  0, 0, 0, nullptr, cta_group());
    SSI_LAUNCH_OK("compact_scan");
    ssi::gather_rows_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<const uint4*>(X), idx, dplan,
                                                         static_cast<int>(D / 8), static_cast<uint4*>(Xc), nullptr,
                                                         nullptr);
    SSI_LAUNCH_OK("gather_rows");
  }

  ssi::GemmParams p;
  memset(&p, 0, sizeof(p));
  if (int r = make_tmap(&p.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Xc, D, f.n_pad, D * 2, 64, ssi::BM)) return r;
  if (int r = make_tmap(&p.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W, D, V, D * 2, 64, ssi::BN / cta_group())) return r;
  p.dplan = dplan;
  p.m_from_plan = 1;
  p.tiles_m = f.tiles_m;  // upper bounds; the kernel takes tiles_m / splits from *dplan
  p.tiles_n = 1;
  p.k_blocks = static_cast<int>(D / ssi::BK);
  p.group_m = f.tiles_m;  // all m-tiles of one split run together: they stream the same W tiles through L2
  p.labels = labels_c;
  p.ignore_index = ignore_index;
  p.n_rows = static_cast<int>(f.n_pad);
  p.vocab = static_cast<int>(V);
  p.col0 = 0;
  p.vtiles_per_split = f.vtiles_total;
  p.vtiles_total = f.vtiles_total;
  p.part_m = reinterpret_cast<float*>(wsb + f.off_m);
  p.part_l = reinterpret_cast<float*>(wsb + f.off_l);
  p.part_z = reinterpret_cast<float*>(wsb + f.off_z);
  p.part_stride = static_cast<int>(f.n_pad);
  if (tune().lockstep_k1) {
    SSI_CUDA(cudaMemsetAsync(wsb + f.off_sync, 0, 256, st));
    p.sync_ctr = reinterpret_cast<unsigned int*>(wsb + f.off_sync), p.sync_every = 1;
    p.sync_err = p.sync_ctr + 63;
  }
  if (int r = launch_gemm<ssi::EPI_LSE, false, false>(p, di->sms, st, "lmhead_lse", CAT_LSE, di->sms)) return r;

  ProfScope ps_aux(CAT_AUX, st);
  float* blk_loss = reinterpret_cast<float*>(wsb + f.off_bl);
  int* blk_cnt = reinterpret_cast<int*>(wsb + f.off_bc);
  ssi::lse_combine_kernel<<<f.combine_blocks, ssi::COMBINE_THREADS, 0, st>>>(
      p.part_m, p.part_l, p.part_z, dplan, pos, static_cast<int>(N), lse, blk_loss, blk_cnt);
  SSI_LAUNCH_OK("lse_combine");
  ssi::loss_finalize_kernel<<<1, 256, 0, st>>>(blk_loss, blk_cnt, f.combine_blocks, loss_out, n_valid);
  SSI_LAUNCH_OK("loss_finalize");
  return 0;
}

int ssi_b200_lmhead_ce_bwd(const void* X, const void* W, const int64_t* labels, const float* lse,
                           const float* grad_out, const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                           int64_t ignore_index, void* dX, void* dW, int accumulate_dW, void* ws, size_t ws_bytes,
                           void* stream) {
  TuneScope tune_scope_;
  if (int r = check_common(X, W, N, V, D)) return r;
  SSI_CHECK_ARG(labels && lse && grad_out && n_valid_total && dX && dW && ws, "null pointer argument");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const BwdPlan b = plan_bwd(N, V, D);
  SSI_CHECK_ARG(ws_bytes >= b.bytes, "backward workspace too small: %zu < %zu", ws_bytes, b.bytes);
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "workspace must be 256-byte aligned");
  char* wsb = static_cast<char*>(ws);
  ssi::DevPlan* dplan = reinterpret_cast<ssi::DevPlan*>(wsb + b.cw.off_plan);
  int* pos = reinterpret_cast<int*>(wsb + b.cw.off_pos);
  int* idx = reinterpret_cast<int*>(wsb + b.cw.off_idx);
  int64_t* labels_c = reinterpret_cast<int64_t*>(wsb + b.cw.off_labels_c);
  void* Xc = wsb + b.cw.off_xc;
  float* lse_c = reinterpret_cast<float*>(wsb + b.off_lse_c);
  void* G = wsb + b.off_g;
  float* dxacc = reinterpret_cast<float*>(wsb + b.off_dxacc);
  const int d_tiles = static_cast<int>(ceil_div(D, ssi::BN));
  // lockstep throttle of the two long-K contractions, as in single-pass mode (one counter per launch)
  unsigned int* sync_area = reinterpret_cast<unsigned int*>(wsb + b.off_sync);
  const int lockstep = (b.nblocks <= 500) ? tune().lockstep : 0;
  if (lockstep > 0) SSI_CUDA(cudaMemsetAsync(sync_area, 0, 4096, st));

  // ---- the same compaction as forward (stateless ABI: recomputed from the labels, ~0.1 ms)
  {
    ProfScope ps(CAT_AUX, st);
    ssi::compact_scan_kernel<<<1, 1024, 0, st>>>(labels, static_cast<int>(N), static_cast<int>(b.n_pad),
                                                  ignore_index, pos, idx, labels_c, dplan, b.vtiles_total, di->sms,
                                                  0, 0, 0, nullptr, cta_group());
    SSI_LAUNCH_OK("compact_scan");
    ssi::gather_rows_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<const uint4*>(X), idx, dplan,
                                                         static_cast<int>(D / 8), static_cast<uint4*>(Xc), lse, lse_c);
    SSI_LAUNCH_OK("gather_rows");
  }

  // tensor maps that do not change across vocabulary blocks (X_c and G hold COMPACT rows)
  CUtensorMap tmX_k, tmX_mn, tmW_k, tmW_mn, tmG_st, tmG_k, tmG_mn, tmdW, tmdX;
  if (int r = make_tmap(&tmX_k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Xc, D, b.n_pad, D * 2, 64, ssi::BM)) return r;
  if (int r = make_tmap(&tmX_mn, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Xc, D, b.n_pad, D * 2, 64, 64)) return r;
  if (int r = make_tmap(&tmW_k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W, D, V, D * 2, 64, ssi::BN / cta_group())) return r;
  if (int r = make_tmap(&tmW_mn, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W, D, V, D * 2, 64, 64)) return r;
  if (int r = make_tmap(&tmG_st, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, G, b.vc, b.n_pad, b.vc * 2, 64, ssi::BM))
    return r;
  tmG_k = tmG_st;  // K-major A operand of dX = G W uses the same {64 v, 128 tok} box
  if (int r = make_tmap(&tmG_mn, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, G, b.vc, b.n_pad, b.vc * 2, 64, 64)) return r;
  if (int r = make_tmap(&tmdW, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dW, D, V, D * 2, 64, ssi::BM)) return r;
  if (int r = make_tmap(&tmdX, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dxacc, D, b.n_pad, D * 4, 32, ssi::BM)) return r;

  for (int blk = 0; blk < b.nblocks; ++blk) {
    const int vt0 = blk * b.vtiles_per_block;
    const int vtn = (b.vtiles_total - vt0) < b.vtiles_per_block ? (b.vtiles_total - vt0) : b.vtiles_per_block;
    const int v0 = vt0 * ssi::BN;  // first vocabulary row of this block

    // ---- K2a: recompute logits tiles, emit dlogits block G[:, v0 : v0 + vtn*256) in bf16
    {
      ssi::GemmParams p;
      memset(&p, 0, sizeof(p));
      p.tmA = tmX_k, p.tmB = tmW_k, p.tmC = tmG_st;
      p.dplan = dplan, p.m_from_plan = 1;
      p.tiles_m = b.tiles_m, p.tiles_n = vtn, p.k_blocks = static_cast<int>(D / ssi::BK);
      p.group_m = tune().group_m_grad;
      p.store_policy = tune().store_hint ? ssi::kEvictFirst : 0ull;
      p.b_r0 = v0;
      p.labels = labels_c, p.ignore_index = ignore_index;
      p.n_rows = static_cast<int>(b.n_pad), p.vocab = static_cast<int>(V), p.col0 = v0;
      p.lse = lse_c, p.grad_out = grad_out, p.n_valid = n_valid_total;
      if (int r = launch_gemm<ssi::EPI_GRAD, false, false>(p, di->sms, st, "lmhead_dlogits", CAT_DLOGITS)) return r;
    }
    // ---- K2b: dW[v0 : v0 + rows) (+)= G^T X_c   (M = vocabulary rows, N = D, K = valid tokens)
    {
      ssi::GemmParams p;
      memset(&p, 0, sizeof(p));
      p.tmA = tmG_mn, p.tmB = tmX_mn, p.tmC = tmdW;
      p.dplan = dplan, p.k_from_plan = 1;
      const int64_t vrows = (V - v0) < int64_t(vtn) * ssi::BN ? (V - v0) : int64_t(vtn) * ssi::BN;
      p.tiles_m = static_cast<int>(ceil_div(vrows, ssi::BM));
      p.tiles_n = d_tiles;
      p.k_blocks = static_cast<int>(b.n_pad / ssi::BK);
      p.group_m = tune().group_m_dw;
      p.store_policy = tune().store_hint ? ssi::kEvictFirst : 0ull;
      p.c_r0 = v0;
      p.c_reduce = accumulate_dW ? 1 : 0;
      if (lockstep > 0) p.sync_ctr = sync_area + 2 * blk, p.sync_err = sync_area + 1023, p.sync_every = lockstep;
      if (int r = launch_gemm_wide<ssi::EPI_BF16, true, true>(p, di->sms, st, "lmhead_dW", CAT_DW, tune().wide_dw))
        return r;
    }
    // ---- K2c: dXacc_c (+)= G W[v0 : v0 + vtn*256)   (M = valid tokens, N = D, K = vocabulary block)
    {
      ssi::GemmParams p;
      memset(&p, 0, sizeof(p));
      p.tmA = tmG_k, p.tmB = tmW_mn, p.tmC = tmdX;
      p.dplan = dplan, p.m_from_plan = 1;
      p.tiles_m = b.tiles_m, p.tiles_n = d_tiles;
      p.k_blocks = vtn * (ssi::BN / ssi::BK);
      p.group_m = tune().group_m_dx;
      p.b_k0 = v0;
      p.c_reduce = (blk > 0) ? 1 : 0;
      if (lockstep > 0) p.sync_ctr = sync_area + 2 * blk + 1, p.sync_err = sync_area + 1023, p.sync_every = lockstep;
      if (int r = launch_gemm_wide<ssi::EPI_F32, false, true>(p, di->sms, st, "lmhead_dX", CAT_DX, tune().wide_dx))
        return r;
    }
  }
  {
    // dX[r] = bf16(dXacc_c[pos[r]]) for valid rows, exact zeros for ignored rows
    ProfScope ps_aux(CAT_AUX, st);
    ssi::dx_finalize_kernel<<<di->sms * 8, 256, 0, st>>>(reinterpret_cast<const float4*>(dxacc), pos,
                                                         static_cast<int>(N), static_cast<int>(D / 4),
                                                         reinterpret_cast<uint2*>(dX));
    SSI_LAUNCH_OK("dx_finalize");
  }
  return 0;
}

int ssi_b200_single_pass_workspace_bytes(int64_t N, int64_t V, int64_t D, size_t* bytes) {
  TuneScope tune_scope_;
  SSI_CHECK_ARG(bytes, "null pointer argument");
  SSI_CHECK_ARG(N >= 1 && V >= 1 && D >= 64 && D % 64 == 0, "bad shape N=%lld V=%lld D=%lld", (long long)N,
                (long long)V, (long long)D);
  SpPlan f;
  if (int r = plan_single_pass(N, V, D, &f)) return r;
  *bytes = f.bytes;
  return 0;
}

int ssi_b200_lmhead_ce_fwd_grad(const void* X, const void* W, const int64_t* labels, const float* grad_hint,
                                const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                                int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* dX,
                                float* dW_acc, void* ws, size_t ws_bytes, void* stream) {
  return ssi_b200_lmhead_ce_fwd_grad_dp(X, W, labels, grad_hint, n_valid_total, N, V, D, ignore_index, lse, loss_out,
                                        n_valid, dX, dW_acc, ws, ws_bytes, stream, 0, nullptr, nullptr);
}

int ssi_b200_lmhead_ce_fwd_grad_dp(const void* X, const void* W, const int64_t* labels, const float* grad_hint,
                                   const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                                   int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* dX,
                                   float* dW_acc, void* ws, size_t ws_bytes, void* stream, int accumulate_dW_acc,
                                   void* dw_acc_init_event, void* dw_ready_event) {
  return ssi_b200_lmhead_ce_fwd_grad_norm(X, W, labels, grad_hint, n_valid_total, N, V, D, ignore_index, lse, loss_out,
                                          n_valid, dX, dW_acc, ws, ws_bytes, stream, accumulate_dW_acc,
                                          dw_acc_init_event, dw_ready_event, nullptr, 0.f, nullptr);
}

int ssi_b200_lmhead_ce_fwd_grad_norm(const void* X, const void* W, const int64_t* labels, const float* grad_hint,
                                     const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                                     int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* dX,
                                     float* dW_acc, void* ws, size_t ws_bytes, void* stream, int accumulate_dW_acc,
                                     void* dw_acc_init_event, void* dw_ready_event, const void* norm_scale,
                                     float norm_eps, float* dscale) {
  ssi_b200_fwd_grad_opts o;
  memset(&o, 0, sizeof(o));
  o.accumulate_dW_acc = accumulate_dW_acc, o.dw_acc_init_event = dw_acc_init_event, o.dw_ready_event = dw_ready_event;
  o.norm_scale = norm_scale, o.norm_eps = norm_eps, o.dscale = dscale;
  return ssi_b200_lmhead_ce_fwd_grad_ex(X, W, labels, grad_hint, n_valid_total, N, V, D, ignore_index, lse, loss_out,
                                        n_valid, dX, dW_acc, ws, ws_bytes, stream, &o);
}

int ssi_b200_lmhead_ce_fwd_grad_ex(const void* X, const void* W, const int64_t* labels, const float* grad_hint,
                                   const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                                   int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* dX,
                                   float* dW_acc, void* ws, size_t ws_bytes, void* stream,
                                   const ssi_b200_fwd_grad_opts* opts) {
  TuneScope tune_scope_;
  ssi_b200_fwd_grad_opts defaults;
  memset(&defaults, 0, sizeof(defaults));
  if (!opts) opts = &defaults;
  const int accumulate_dW_acc = opts->accumulate_dW_acc;
  void* const dw_acc_init_event = opts->dw_acc_init_event;
  void* const dw_ready_event = opts->dw_ready_event;
  void* const pre_grad_event = opts->pre_grad_event;
  const void* const norm_scale = opts->norm_scale;
  const float norm_eps = opts->norm_eps;
  float* const dscale = opts->dscale;
  if (int r = check_common(X, W, N, V, D)) return r;
  SSI_CHECK_ARG(!norm_scale || (dscale && (reinterpret_cast<uintptr_t>(norm_scale) & 15) == 0 && D <= 8192),
                "fused norm needs a 16-byte aligned scale, a dscale output and D <= 8192");
  SSI_CHECK_ARG(labels && grad_hint && lse && loss_out && n_valid && dX && dW_acc && ws, "null pointer argument");
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(dX) & 15) == 0 && (reinterpret_cast<uintptr_t>(dW_acc) & 15) == 0,
                "dX / dW_acc must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  SpPlan f;
  if (int r = plan_single_pass(N, V, D, &f)) return r;
  SSI_CHECK_ARG(ws_bytes >= f.bytes, "single-pass workspace too small: %zu < %zu", ws_bytes, f.bytes);
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "workspace must be 256-byte aligned");
  char* wsb = static_cast<char*>(ws);
  ssi::DevPlan* dplan = reinterpret_cast<ssi::DevPlan*>(wsb + f.cw.off_plan);
  int* pos = reinterpret_cast<int*>(wsb + f.cw.off_pos);
  int* idx = reinterpret_cast<int*>(wsb + f.cw.off_idx);
  int64_t* labels_c = reinterpret_cast<int64_t*>(wsb + f.cw.off_labels_c);
  void* Xc = wsb + f.cw.off_xc;
  float* part_m = reinterpret_cast<float*>(wsb + f.off_m);
  float* part_l = reinterpret_cast<float*>(wsb + f.off_l);
  float* part_z = reinterpret_cast<float*>(wsb + f.off_z);
  float* lse_c = reinterpret_cast<float*>(wsb + f.off_lse_c);
  float* loss_c = reinterpret_cast<float*>(wsb + f.off_loss_c);
  float* blk_loss = reinterpret_cast<float*>(wsb + f.off_bl);
  int* blk_cnt = reinterpret_cast<int*>(wsb + f.off_bc);
  void* dXc = wsb + f.off_dxc;
  void* S = wsb + f.off_s;
  float* ref2 = reinterpret_cast<float*>(wsb + f.off_ref2);
  float* rowscale_c = reinterpret_cast<float*>(wsb + f.off_rowscale);
  float* labw_c = reinterpret_cast<float*>(wsb + f.off_labw);
  void* Xs = wsb + f.off_xs;
  const bool estash = tune().exp_stash != 0;
  const int d_tiles = static_cast<int>(ceil_div(D, ssi::BN));
  const int64_t rows_cap = static_cast<int64_t>(f.t_max) * ssi::BM;
  const int64_t* nvt = n_valid_total ? n_valid_total : n_valid;
  unsigned int* sync_area = reinterpret_cast<unsigned int*>(wsb + f.off_sync);
  const int lockstep = tune().lockstep;
  // word 1022 = exp-stash overflow flag, word 1023 = lockstep gave up, the rest = lockstep counters
  if (lockstep > 0 || tune().lockstep_k1 || estash) SSI_CUDA(cudaMemsetAsync(sync_area, 0, SYNC_AREA_BYTES, st));

  {
    ProfScope ps(CAT_AUX, st);
    ssi::compact_scan_kernel<<<1, 1024, 0, st>>>(labels, static_cast<int>(N), static_cast<int>(f.n_pad),
                                                  ignore_index, pos, idx, labels_c, dplan, f.vtiles_total, di->sms,
                                                  f.t_max, f.chunks_max,
                                                  (cta_group() == 2 && tune().wide_dx == 2) ? (d_tiles + 1) / 2 : d_tiles,
                                                  n_valid, cta_group(),
                                                  (cta_group() == 2 && tune().wide_dx == 2) ? tune().tail_split : 0);
    SSI_LAUNCH_OK("compact_scan");
    if (norm_scale) {  // X holds PRE-norm hidden states: the final RMSNorm is applied while gathering the valid rows
      ssi::gather_rows_norm_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<const uint4*>(X), idx, dplan,
                                                                static_cast<int>(D / 8),
                                                                static_cast<const uint4*>(norm_scale), norm_eps,
                                                                static_cast<uint4*>(Xc),
                                                                reinterpret_cast<float*>(wsb + f.off_rstd));
    } else {
      ssi::gather_rows_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<const uint4*>(X), idx, dplan,
                                                           static_cast<int>(D / 8), static_cast<uint4*>(Xc), nullptr,
                                                           nullptr);
    }
    SSI_LAUNCH_OK("gather_rows");
  }

  CUtensorMap tmX_k, tmX_mn, tmW_k, tmW_mn, tmS_k, tmS_mn, tmdW, tmdXc;
  if (int r = make_tmap(&tmX_k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Xc, D, f.n_pad, D * 2, 64, ssi::BM)) return r;
  if (int r = make_tmap(&tmX_mn, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Xc, D, f.n_pad, D * 2, 64, 64)) return r;
  if (int r = make_tmap(&tmW_k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W, D, V, D * 2, 64, ssi::BN / cta_group())) return r;
  if (int r = make_tmap(&tmW_mn, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W, D, V, D * 2, 64, 64)) return r;
  if (int r = make_tmap(&tmS_k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, S, f.vpad, rows_cap, f.vpad * 2, 64, ssi::BM))
    return r;
  if (int r = make_tmap(&tmS_mn, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, S, f.vpad, rows_cap, f.vpad * 2, 64, 64))
    return r;
  if (int r = make_tmap(&tmdW, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dW_acc, D, V, D * 4, 32, ssi::BM)) return r;
  if (int r = make_tmap(&tmdXc, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dXc, D, f.n_pad, D * 2, 64, ssi::BM)) return r;
  CUtensorMap tmXs_mn = tmX_mn;  // dW's B operand: the scaled rows in the exp-stash form
  if (estash)
    if (int r = make_tmap(&tmXs_mn, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, Xs, D, f.n_pad, D * 2, 64, 64)) return r;

  for (int ch = 0; ch < f.chunks_max; ++ch) {
    const ssi::DevPlan* cplan = dplan + 1 + ch;
    // ---- K1: logits tiles of the chunk -> one bf16 block + per-split row partials
    auto k1_params = [&](ssi::GemmParams& p) {
      memset(&p, 0, sizeof(p));
      p.tmA = tmX_k, p.tmB = tmW_k, p.tmC = tmS_k;
      p.dplan = cplan, p.m_from_plan = 1, p.plan_rows = ssi::PR_A_ROWS | ssi::PR_LABEL_ROWS;
      p.tiles_m = f.t_max, p.tiles_n = 1, p.k_blocks = static_cast<int>(D / ssi::BK);
      p.group_m = f.t_max;
      p.lse_group_m = tune().k1_group_m;
      p.store_policy = (tune().store_hint || tune().k1_store_first) ? ssi::kEvictFirst : 0ull;
      p.labels = labels_c, p.ignore_index = ignore_index;
      p.n_rows = static_cast<int>(f.n_pad), p.vocab = static_cast<int>(V), p.col0 = 0;
      p.vtiles_per_split = f.vtiles_total, p.vtiles_total = f.vtiles_total;
      p.part_m = part_m, p.part_l = part_l, p.part_z = part_z;
      p.part_stride = static_cast<int>(rows_cap);
      set_hints(p, tune().k1_hints);
    };
    if (estash) {
      // pilot: the running max of the FIRST vocabulary tile, per row (log2 domain) = the reference of E below
      {
        ssi::GemmParams p;
        k1_params(p);
        p.vtiles_per_split = 1, p.vtiles_total = 1, p.lse_one_split = 1;
        p.part_m = ref2;
        if (int r = launch_gemm<ssi::EPI_LSE, false, false>(p, di->sms, st, "lmhead_pilot", CAT_LSE, di->sms)) return r;
      }
      ssi::GemmParams p;
      k1_params(p);
      p.row_ref2 = ref2;
      if (tune().lockstep_k1) p.sync_ctr = sync_area + 256 + ch, p.sync_err = sync_area + 1023, p.sync_every = 1;
      if (int r = launch_gemm<ssi::EPI_EXP_STASH, false, false>(p, di->sms, st, "lmhead_exp_stash", CAT_LSE, di->sms))
        return r;
      // data parallel: grad_hint / n_valid_total are filled by the counts exchange running on another stream beside
      // the first logits contraction; the combine below is their first consumer
      if (ch == 0 && pre_grad_event) SSI_CUDA(cudaStreamWaitEvent(st, static_cast<cudaEvent_t>(pre_grad_event), 0));
      ProfScope ps(CAT_AUX, st);
      ssi::exp_combine_kernel<<<64, 256, 0, st>>>(ref2, part_l, part_z, cplan, grad_hint, nvt, lse_c, loss_c,
                                                  rowscale_c, labw_c, sync_area + 1022);
      SSI_LAUNCH_OK("exp_combine");
      ssi::scale_rows_kernel<<<di->sms * 4, 256, 0, st>>>(static_cast<const uint4*>(Xc), rowscale_c, cplan,
                                                          static_cast<int>(D / 8), static_cast<uint4*>(Xs));
      SSI_LAUNCH_OK("scale_rows");
    } else {
      {
        ssi::GemmParams p;
        k1_params(p);
        if (tune().lockstep_k1) p.sync_ctr = sync_area + 256 + ch, p.sync_err = sync_area + 1023, p.sync_every = 1;
        if (int r = launch_gemm<ssi::EPI_LSE_STASH, false, false>(p, di->sms, st, "lmhead_lse_stash", CAT_LSE, di->sms))
          return r;
      }
      {
        ProfScope ps(CAT_AUX, st);
        ssi::chunk_combine_kernel<<<64, 256, 0, st>>>(part_m, part_l, part_z, cplan, lse_c, loss_c);
        SSI_LAUNCH_OK("chunk_combine");
      }
      // ---- S -> G in place (softmax - onehot) * grad_hint / n_valid_total
      if (ch == 0 && pre_grad_event) SSI_CUDA(cudaStreamWaitEvent(st, static_cast<cudaEvent_t>(pre_grad_event), 0));
      ProfScope ps(CAT_SOFTMAX, st);
      ssi::softmax_grad_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<uint4*>(S), static_cast<int>(f.vpad / 8),
                                                            cplan, lse_c, labels_c, static_cast<int>(V), grad_hint,
                                                            nvt);
      SSI_LAUNCH_OK("softmax_grad");
    }
    // ---- dW_acc (+)= G^T X_c[chunk]   (M = vocabulary rows, N = D, K = the chunk's token rows)
    {
      ssi::GemmParams p;
      memset(&p, 0, sizeof(p));
      p.tmA = tmS_mn, p.tmB = tmXs_mn, p.tmC = tmdW;
      p.dplan = cplan, p.k_from_plan = 1, p.plan_rows = ssi::PR_B_K;
      p.tiles_m = static_cast<int>(ceil_div(V, ssi::BM));
      p.tiles_n = d_tiles;
      p.k_blocks = static_cast<int>(rows_cap / ssi::BK);
      p.group_m = tune().group_m_dw;
      // the first NON-EMPTY round overwrites dW_acc and later rounds add -- or all add when the caller initialised it
      p.c_reduce = accumulate_dW_acc ? 1 : 0, p.c_reduce_plan = accumulate_dW_acc ? 0 : 1;
      if (ch == 0 && dw_acc_init_event)
        SSI_CUDA(cudaStreamWaitEvent(st, static_cast<cudaEvent_t>(dw_acc_init_event), 0));
      set_hints(p, tune().dw_hints);
      if (lockstep > 0) p.sync_ctr = sync_area + 2 * ch, p.sync_err = sync_area + 1023, p.sync_every = lockstep;
      if (int r = launch_gemm_wide<ssi::EPI_F32, true, true>(p, di->sms, st, "lmhead_dW", CAT_DW, tune().wide_dw))
        return r;
      if (ch == f.chunks_max - 1) {
        if (estash) {
          // the label column's term of dW in fp32: dW_acc[v, :] -= g * sum over the X_c rows whose label is v of
          // (1 - p_label[row]) * X_c[row, :]
          ssi::ScatOut o = dense_out(dW_acc);
          o.scale_num = grad_hint, o.scale_den = reinterpret_cast<const long long*>(nvt), o.scale_sign = -1.f;
          o.row_w = labw_c;
          if (int r = embedding_grad_scatter_impl<ssi::SCAT_DENSE_F32>(Xc, labels_c, f.n_pad, V, D, o,
                                                                       wsb + f.off_lscat, f.lscat_bytes, stream, 3,
                                                                       CAT_SOFTMAX))
            return r;
        }
        // dW_acc is final after the last round's contraction (the real chunks occupy the last rounds): let the
        // caller start the gradient exchange on another stream while this stream still runs the last dX
        if (dw_ready_event) SSI_CUDA(cudaEventRecord(static_cast<cudaEvent_t>(dw_ready_event), st));
      }
    }
    // ---- dXc[chunk] = G W   (M = the chunk's token rows, N = D, K = padded vocabulary)
    {
      ssi::GemmParams p;
      memset(&p, 0, sizeof(p));
      p.tmA = tmS_k, p.tmB = tmW_mn, p.tmC = tmdXc;
      p.dplan = cplan, p.m_from_plan = 1, p.plan_rows = ssi::PR_C_ROWS;
      p.tiles_m = f.t_max, p.tiles_n = d_tiles;
      p.k_blocks = f.vtiles_total * (ssi::BN / ssi::BK);
      p.group_m = tune().group_m_dx;
      if (estash) p.row_scale = rowscale_c;  // dlogits = rowscale[row] * E'
      set_hints(p, tune().dx_hints);
      if (lockstep > 0) p.sync_ctr = sync_area + 2 * ch + 1, p.sync_err = sync_area + 1023, p.sync_every = lockstep;
      if (int r = launch_gemm_wide<ssi::EPI_BF16, false, true>(p, di->sms, st, "lmhead_dX", CAT_DX, tune().wide_dx))
        return r;
    }
  }
  ProfScope ps_aux(CAT_AUX, st);
  ssi::stash_finish_kernel<<<f.combine_blocks, ssi::COMBINE_THREADS, 0, st>>>(lse_c, loss_c, pos,
                                                                              static_cast<int>(N), lse, blk_loss,
                                                                              blk_cnt);
  SSI_LAUNCH_OK("stash_finish");
  ssi::loss_finalize_kernel<<<1, 256, 0, st>>>(blk_loss, blk_cnt, f.combine_blocks, loss_out, n_valid);
  SSI_LAUNCH_OK("loss_finalize");
  if (norm_scale) {
    // backward of the fused norm while the compact dX rows go back to their positions: dX = d(pre-norm h)
    const int row_u4 = static_cast<int>(D / 8);
    if (estash) {
      ssi::dx_label_term_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<uint4*>(dXc), labels_c,
                                                             static_cast<int>(f.n_pad), row_u4,
                                                             static_cast<const uint4*>(W), static_cast<int>(V),
                                                             grad_hint, nvt, labw_c);
      SSI_LAUNCH_OK("dx_label_term");
    }
    const int maxj = (row_u4 + 31) / 32;
    float* partial = reinterpret_cast<float*>(wsb + f.off_dsp);
    const float* rstd_c = reinterpret_cast<const float*>(wsb + f.off_rstd);
#define SSI_NORM_SCATTER(J)                                                                                         \
  ssi::dx_scatter_norm_bwd_kernel<J><<<NORM_BWD_BLOCKS, 256, 0, st>>>(                                              \
      static_cast<const uint4*>(dXc), pos, static_cast<const uint4*>(X), static_cast<const uint4*>(norm_scale),     \
      rstd_c, static_cast<int>(N), row_u4, static_cast<uint4*>(dX), partial)
    if (maxj <= 2) SSI_NORM_SCATTER(2);
    else if (maxj <= 4) SSI_NORM_SCATTER(4);
    else if (maxj <= 8) SSI_NORM_SCATTER(8);
    else if (maxj <= 16) SSI_NORM_SCATTER(16);
    else SSI_NORM_SCATTER(32);
#undef SSI_NORM_SCATTER
    SSI_LAUNCH_OK("dx_scatter_norm_bwd");
    ssi::rmsnorm_dscale_reduce_kernel<<<static_cast<int>(ceil_div(D, 32)), 256, 0, st>>>(partial, NORM_BWD_PARTS,
                                                                                       static_cast<int>(D), dscale);
    SSI_LAUNCH_OK("rmsnorm_dscale_reduce");
    return 0;
  }
  ssi::dx_scatter_bf16_kernel<<<di->sms * 8, 256, 0, st>>>(
      static_cast<const uint4*>(dXc), pos, static_cast<int>(N), static_cast<int>(D / 8), static_cast<uint4*>(dX),
      estash ? static_cast<const uint4*>(W) : nullptr, labels_c, static_cast<int>(V), grad_hint, nvt, labw_c);
  SSI_LAUNCH_OK("dx_scatter");
  return 0;
}

int ssi_b200_lmhead_ce_apply_grad(void* dX, const float* dW_acc, const float* grad_out, const float* grad_hint,
                                  int64_t N, int64_t V, int64_t D, void* dW, int accumulate_dW, void* stream) {
  return ssi_b200_lmhead_ce_apply_grad_fold(dX, dW_acc, grad_out, grad_hint, N, V, D, dW, accumulate_dW, nullptr,
                                            nullptr, 0, stream);
}

int ssi_b200_lmhead_ce_apply_grad_fold(void* dX, const float* dW_acc, const float* grad_out, const float* grad_hint,
                                       int64_t N, int64_t V, int64_t D, void* dW, int accumulate_dW,
                                       const float* post_scale, void* fold_ws_ptr, size_t fold_ws_bytes,
                                       void* stream) {
  SSI_CHECK_ARG(grad_out && grad_hint && (dW_acc || !dW), "null pointer argument");
  SSI_CHECK_ARG(N >= 0 && V >= 1 && D >= 8 && D % 8 == 0, "bad shape N=%lld V=%lld D=%lld", (long long)N,
                (long long)V, (long long)D);
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(dX) & 15) == 0 && (reinterpret_cast<uintptr_t>(dW_acc) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(dW) & 15) == 0,
                "dX / dW_acc / dW must be 16-byte aligned");
  SSI_CHECK_ARG(dW || (!post_scale && !fold_ws_ptr), "the scale_grads fold needs dW");
  FoldWs fw;
  if (int r = fold_ws(fold_ws_ptr, fold_ws_bytes, &fw)) return r;
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (fw.base) SSI_CUDA(cudaMemsetAsync(fw.base, 0, fw.floats * sizeof(float), st));  // a fold begins here
  ProfScope ps(CAT_AUX, st);
  if (dX && N > 0) {
    ssi::apply_dx_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<uint4*>(dX), static_cast<size_t>(N) * D / 8,
                                                      grad_out, grad_hint);
    SSI_LAUNCH_OK("apply_dx");
  }
  if (dW) {
    const int grid = di->sms * 16 < FOLD_SEG ? di->sms * 16 : FOLD_SEG;
    ssi::GradFold gf;
    gf.post_scale = post_scale, gf.sq = fw.seg(0);
    ssi::apply_dw_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4*>(dW_acc),
                                               static_cast<size_t>(V) * D / 4, grad_out, grad_hint,
                                               accumulate_dW ? 1 : 0, static_cast<uint2*>(dW), gf);
    SSI_LAUNCH_OK("apply_dw");
  }
  return 0;
}

int ssi_b200_grad_fold_workspace_bytes(int64_t max_rows, int64_t D, size_t* bytes) {
  SSI_CHECK_ARG(bytes && max_rows >= 1 && D >= 8, "bad argument");
  *bytes = (2 * static_cast<size_t>(FOLD_SEG) + fold_sparse_floats(max_rows, D)) * sizeof(float);
  return 0;
}

int ssi_b200_grad_fold_sqnorm(const void* fold_ws_ptr, size_t fold_ws_bytes, float* sqnorm_out, void* stream) {
  SSI_CHECK_ARG(fold_ws_ptr && sqnorm_out, "null pointer argument");
  FoldWs fw;
  if (int r = fold_ws(const_cast<void*>(fold_ws_ptr), fold_ws_bytes, &fw)) return r;
  SSI_CHECK_ARG(fw.floats < (size_t(1) << 31), "fold workspace too large");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ProfScope ps(CAT_AUX, st);
  ssi::grad_fold_finalize_kernel<<<1, 256, 0, st>>>(fw.base, static_cast<int>(fw.floats), sqnorm_out);
  SSI_LAUNCH_OK("grad_fold_finalize");
  return 0;
}

}  // extern "C"
namespace {
// phases: bit 0 = sort the ids (leaves the order in ws), bit 1 = segmented reduction of the dH rows in that order
template <int MODE>
int embedding_grad_scatter_impl(const void* dH, const int64_t* ids, int64_t N, int64_t V, int64_t D,
                                const ssi::ScatOut& out, void* ws, size_t ws_bytes, void* stream, int phases, int cat) {
  SSI_CHECK_ARG(ws && (!(phases & 1) || ids) && (!(phases & 2) || dH), "null pointer argument");
  SSI_CHECK_ARG(N >= 1 && V >= 1 && D >= 8 && D % 8 == 0, "bad shape N=%lld V=%lld D=%lld", (long long)N,
                (long long)V, (long long)D);
  SSI_CHECK_ARG(N < (int64_t(1) << 31) && V < (int64_t(1) << 31) - 1, "N/V too large");
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(dH) & 15) == 0, "dH must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const ScatPlan s = plan_scatter(N, V, D);
  SSI_CHECK_ARG(ws_bytes >= s.bytes, "scatter workspace too small: %zu < %zu", ws_bytes, s.bytes);
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(ws) & 15) == 0, "scatter workspace must be 16-byte aligned");
  char* wsb = static_cast<char*>(ws);
  uint32_t* k[2] = {reinterpret_cast<uint32_t*>(wsb + s.off_k0), reinterpret_cast<uint32_t*>(wsb + s.off_k1)};
  uint32_t* v[2] = {reinterpret_cast<uint32_t*>(wsb + s.off_v0), reinterpret_cast<uint32_t*>(wsb + s.off_v1)};
  int* urank = reinterpret_cast<int*>(wsb + s.off_urank);
  int* nuniq = reinterpret_cast<int*>(wsb + s.off_nuniq);
  const int n = static_cast<int>(N);

  ProfScope ps_scat(cat, st);
  if (phases & 1) {
    static std::once_flag once[64];
    std::call_once(once[di->dev], [] {
      cudaFuncSetAttribute(ssi::sort_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ssi::SORT_SMEM_BYTES);
    });
    // one block per SORT_KEYS_PER_BLOCK ids; cooperative launch: the blocks synchronise through a counter in ws
    int G = static_cast<int>(ceil_div(N, ssi::SORT_KEYS_PER_BLOCK));
    const int g_max = di->sms < ssi::SORT_MAX_BLOCKS ? di->sms : ssi::SORT_MAX_BLOCKS;
    G = G < 1 ? 1 : (G > g_max ? g_max : G);
    SSI_CUDA(cudaMemsetAsync(nuniq, 0, 256, st));
    uint32_t* ghist = reinterpret_cast<uint32_t*>(wsb + s.off_ghist);
    unsigned int* bar = reinterpret_cast<unsigned int*>(nuniq) + 16;
    long long v_ll = V;
    int passes = s.passes;
    int n_arg = n;
    void* args[] = {&ids, &n_arg, &v_ll, &passes, &k[0], &v[0], &k[1], &v[1], &urank, &nuniq, &ghist, &bar};
    SSI_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(ssi::sort_grid_kernel), dim3(G),
                                         dim3(ssi::SORT_THREADS), args, ssi::SORT_SMEM_BYTES, st));
    SSI_LAUNCH_OK("sort_grid");
  }
  if (!(phases & 2)) return 0;
  const int cur = (s.passes - 1) & 1;
  float* partial = reinterpret_cast<float*>(wsb + s.off_partial);
  ssi::ScatInfo* info = reinterpret_cast<ssi::ScatInfo*>(wsb + s.off_info);
  // no shared memory, <= 72 registers: co-resides with a persistent tensor-core CTA
  const int64_t warps_total = static_cast<int64_t>(s.nwarps) * s.slabs;
  const unsigned grid = static_cast<unsigned>(ceil_div(warps_total, ssi::SCAT_THREADS / 32));
  if (out.fold.sq) SSI_CUDA(cudaMemsetAsync(out.fold.sq, 0, 2 * static_cast<size_t>(warps_total) * sizeof(float), st));
  ssi::scatter_reduce_kernel<MODE><<<grid, ssi::SCAT_THREADS, 0, st>>>(static_cast<const __nv_bfloat16*>(dH), k[cur],
                                                                      v[cur], urank, n, static_cast<uint32_t>(V),
                                                                      static_cast<int>(D), s.slabs, out, partial, info);
  SSI_LAUNCH_OK("scatter_reduce");
  ssi::scatter_fixup_kernel<MODE><<<grid, ssi::SCAT_THREADS, 0, st>>>(partial, info, s.nwarps,
                                                                     static_cast<uint32_t>(V), static_cast<int>(D),
                                                                     s.slabs, out);
  SSI_LAUNCH_OK("scatter_fixup");
  return 0;
}

ssi::ScatOut dense_out(void* dW) {
  ssi::ScatOut o;
  memset(&o, 0, sizeof(o));
  o.dense = dW;
  return o;
}
}  // namespace
extern "C" {

int ssi_b200_embedding_grad_scatter(const void* dH, const int64_t* ids, int64_t N, int64_t V, int64_t D, void* dW,
                                    void* ws, size_t ws_bytes, void* stream) {
  SSI_CHECK_ARG(dW && (reinterpret_cast<uintptr_t>(dW) & 15) == 0, "dW must be a 16-byte aligned device pointer");
  return embedding_grad_scatter_impl<ssi::SCAT_DENSE_BF16>(dH, ids, N, V, D, dense_out(dW), ws, ws_bytes, stream, 3);
}

int ssi_b200_embedding_grad_scatter_f32(const void* dH, const int64_t* ids, int64_t N, int64_t V, int64_t D,
                                        float* dW_acc, void* ws, size_t ws_bytes, void* stream) {
  SSI_CHECK_ARG(dW_acc && (reinterpret_cast<uintptr_t>(dW_acc) & 15) == 0, "dW_acc must be 16-byte aligned");
  return embedding_grad_scatter_impl<ssi::SCAT_DENSE_F32>(dH, ids, N, V, D, dense_out(dW_acc), ws, ws_bytes, stream,
                                                          3);
}

int ssi_b200_embedding_grad_sort(const int64_t* ids, int64_t N, int64_t V, int64_t D, void* ws, size_t ws_bytes,
                                 void* stream) {
  return embedding_grad_scatter_impl<ssi::SCAT_DENSE_F32>(nullptr, ids, N, V, D, dense_out(nullptr), ws, ws_bytes,
                                                          stream, 1);
}

int ssi_b200_embedding_grad_scatter_sorted(const void* dH, int64_t N, int64_t V, int64_t D, void* dW, void* ws,
                                           size_t ws_bytes, void* stream) {
  SSI_CHECK_ARG(dW && (reinterpret_cast<uintptr_t>(dW) & 15) == 0, "dW must be a 16-byte aligned device pointer");
  return embedding_grad_scatter_impl<ssi::SCAT_DENSE_BF16>(dH, nullptr, N, V, D, dense_out(dW), ws, ws_bytes, stream,
                                                           2);
}

int ssi_b200_embedding_grad_scatter_sorted_fold(const void* dH, int64_t N, int64_t V, int64_t D, void* dW, void* ws,
                                                size_t ws_bytes, const float* post_scale, void* fold_ws_ptr,
                                                size_t fold_ws_bytes, void* stream) {
  SSI_CHECK_ARG(dW && (reinterpret_cast<uintptr_t>(dW) & 15) == 0, "dW must be a 16-byte aligned device pointer");
  FoldWs fw;
  if (int r = fold_ws(fold_ws_ptr, fold_ws_bytes, &fw)) return r;
  SSI_CHECK_ARG(!fw.base || fw.sparse_floats() >= fold_sparse_floats(N, D), "fold workspace too small for N=%lld",
                (long long)N);
  ssi::ScatOut o = dense_out(dW);
  o.fold.post_scale = post_scale, o.fold.sq = fw.seg(2);
  return embedding_grad_scatter_impl<ssi::SCAT_DENSE_BF16>(dH, nullptr, N, V, D, o, ws, ws_bytes, stream, 2);
}

int ssi_b200_embedding_grad_scatter_sorted_f32(const void* dH, int64_t N, int64_t V, int64_t D, float* dW_acc,
                                               void* ws, size_t ws_bytes, void* stream) {
  SSI_CHECK_ARG(dW_acc && (reinterpret_cast<uintptr_t>(dW_acc) & 15) == 0, "dW_acc must be 16-byte aligned");
  return embedding_grad_scatter_impl<ssi::SCAT_DENSE_F32>(dH, nullptr, N, V, D, dense_out(dW_acc), ws, ws_bytes,
                                                          stream, 2);
}

// ---------------------------------------------------------------------------------------------------
// Data-parallel exchange of the tied gradient over NVLink peer memory (dp_exchange.cuh)
// ---------------------------------------------------------------------------------------------------
int ssi_b200_dp_buffer_bytes(int64_t V, int64_t D, int world, int64_t max_rows, size_t* bytes,
                             size_t* dw_acc_offset) {
  SSI_CHECK_ARG(bytes, "null pointer argument");
  SSI_CHECK_ARG(V >= 1 && D >= 8 && D % 8 == 0, "bad shape V=%lld D=%lld", (long long)V, (long long)D);
  SSI_CHECK_ARG(world >= 1 && world <= ssi::DP_MAX_WORLD, "world size %d not in [1, %d]", world, ssi::DP_MAX_WORLD);
  SSI_CHECK_ARG(max_rows >= 0 && max_rows < (int64_t(1) << 31), "bad max_rows %lld", (long long)max_rows);
  const ssi::DpLayout l = ssi::dp_layout(V, D, world, max_rows);
  *bytes = l.bytes;
  if (dw_acc_offset) *dw_acc_offset = l.off_acc;
  return 0;
}

namespace {
int dp_store_timeout(void* own_buf) {
  const unsigned long long ns = static_cast<unsigned long long>(tune().dp_timeout_s) * 1000000000ull;
  SSI_CUDA(cudaMemcpy(static_cast<char*>(own_buf) + offsetof(ssi::DpCtrl, timeout_ns), &ns, sizeof(ns),
                      cudaMemcpyHostToDevice));
  return 0;
}
}  // namespace

int ssi_b200_dp_alloc(size_t bytes, void** ptr, void* ipc_handle) {
  TuneScope tune_scope_;
  SSI_CHECK_ARG(ptr && ipc_handle && bytes >= ssi::DP_CTRL_BYTES, "bad argument");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  void* p = nullptr;
  SSI_CUDA(cudaMalloc(&p, bytes));
  SSI_CUDA(cudaMemset(p, 0, bytes));  // flags, and slot entries whose stale epoch tags could match
  if (int r = dp_store_timeout(p)) {
    cudaFree(p);
    return r;
  }
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail(static_cast<int>(e), "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(ipc_handle, &h, sizeof(h));
  *ptr = p;
  return 0;
}

int ssi_b200_dp_open(const void* ipc_handle, void** ptr) {
  SSI_CHECK_ARG(ipc_handle && ptr, "null pointer argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle, sizeof(h));
  SSI_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int ssi_b200_dp_close(void* ptr) {
  SSI_CHECK_ARG(ptr, "null pointer argument");
  SSI_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int ssi_b200_dp_free(void* ptr) {
  SSI_CHECK_ARG(ptr, "null pointer argument");
  SSI_CUDA(cudaFree(ptr));
  return 0;
}

int ssi_b200_dp_status(const void* own_buf, int* timed_out) {
  SSI_CHECK_ARG(own_buf && timed_out, "null pointer argument");
  unsigned int v = 0;
  SSI_CUDA(cudaMemcpy(&v, static_cast<const char*>(own_buf) + offsetof(ssi::DpCtrl, timed_out), sizeof(v),
                      cudaMemcpyDeviceToHost));
  *timed_out = static_cast<int>(v);
  return 0;
}

int ssi_b200_dp_reset(void* own_buf) {
  TuneScope tune_scope_;
  SSI_CHECK_ARG(own_buf, "null pointer argument");
  SSI_CUDA(cudaDeviceSynchronize());  // no exchange kernel of this rank may still be spinning or poisoning
  const unsigned int zero = 0;
  SSI_CUDA(cudaMemcpy(static_cast<char*>(own_buf) + offsetof(ssi::DpCtrl, timed_out), &zero, sizeof(zero),
                      cudaMemcpyHostToDevice));
  return dp_store_timeout(own_buf);
}

namespace {
int dp_peers(void* const* peer_bufs, int rank, int world, ssi::DpPeers* peers) {
  SSI_CHECK_ARG(peer_bufs, "null pointer argument");
  SSI_CHECK_ARG(world >= 1 && world <= ssi::DP_MAX_WORLD && rank >= 0 && rank < world, "bad rank %d / world %d", rank,
                world);
  memset(peers, 0, sizeof(*peers));
  for (int p = 0; p < world; ++p) {
    SSI_CHECK_ARG(peer_bufs[p], "null exchange buffer for rank %d", p);
    peers->buf[p] = static_cast<char*>(peer_bufs[p]);
  }
  return 0;
}
}  // namespace

int ssi_b200_dp_counts(void* const* peer_bufs, int rank, int world, uint32_t epoch, const int64_t* local_counts,
                       int64_t* global_counts, float* hint_out, void* host_flag, void* stream) {
  SSI_CHECK_ARG(local_counts && global_counts, "null pointer argument");
  ssi::DpPeers peers;
  if (int r = dp_peers(peer_bufs, rank, world, &peers)) return r;
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ProfScope ps(CAT_EXCHANGE, st);
  ssi::dp_counts_kernel<<<1, 32, 0, st>>>(peers, rank, world, epoch, local_counts, global_counts, hint_out,
                                          static_cast<unsigned int*>(host_flag));
  SSI_LAUNCH_OK("dp_counts");
  return 0;
}

int ssi_b200_dp_exchange(void* const* peer_bufs, int rank, int world, int64_t V, int64_t D, int64_t max_rows,
                         uint32_t epoch, const float* grad_out, const float* grad_hint, void* dW, int accumulate_dW,
                         void* host_flag, void* stream) {
  return ssi_b200_dp_exchange_fold(peer_bufs, rank, world, V, D, max_rows, epoch, grad_out, grad_hint, dW,
                                   accumulate_dW, host_flag, nullptr, nullptr, 0, stream);
}

int ssi_b200_dp_exchange_fold(void* const* peer_bufs, int rank, int world, int64_t V, int64_t D, int64_t max_rows,
                              uint32_t epoch, const float* grad_out, const float* grad_hint, void* dW,
                              int accumulate_dW, void* host_flag, const float* post_scale, void* fold_ws_ptr,
                              size_t fold_ws_bytes, void* stream) {
  return ssi_b200_dp_exchange_nvls(peer_bufs, nullptr, rank, world, V, D, max_rows, epoch, grad_out, grad_hint, dW,
                                   accumulate_dW, host_flag, post_scale, fold_ws_ptr, fold_ws_bytes, stream);
}

int ssi_b200_dp_exchange_nvls(void* const* peer_bufs, const void* multicast_buf, int rank, int world, int64_t V,
                              int64_t D, int64_t max_rows, uint32_t epoch, const float* grad_out,
                              const float* grad_hint, void* dW, int accumulate_dW, void* host_flag,
                              const float* post_scale, void* fold_ws_ptr, size_t fold_ws_bytes, void* stream) {
  SSI_CHECK_ARG(grad_out && grad_hint && dW, "null pointer argument");
  ssi::DpPeers peers;
  if (int r = dp_peers(peer_bufs, rank, world, &peers)) return r;
  SSI_CHECK_ARG(V >= 1 && D >= 8 && D % 8 == 0, "bad shape V=%lld D=%lld", (long long)V, (long long)D);
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(dW) & 15) == 0, "dW must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  FoldWs fw;
  if (int r = fold_ws(fold_ws_ptr, fold_ws_bytes, &fw)) return r;
  if (fw.base) SSI_CUDA(cudaMemsetAsync(fw.base, 0, fw.floats * sizeof(float), st));  // a fold begins here
  ssi::GradFold f_own, f_others;
  f_own.post_scale = post_scale, f_own.sq = fw.seg(0);
  f_others.post_scale = nullptr, f_others.sq = fw.seg(1);
  const ssi::DpLayout lay = ssi::dp_layout(V, D, world, max_rows);
  const long long total_f4 = static_cast<long long>(V) * D / 4;
  const long long slice_f4 = static_cast<long long>(ceil_div(V, world)) * D / 4;
  const int grid = di->sms * 2 < FOLD_SEG ? di->sms * 2 : FOLD_SEG;
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(multicast_buf) & 15) == 0, "multicast_buf must be 16-byte aligned");
  const float4* mc_acc = multicast_buf ? reinterpret_cast<const float4*>(static_cast<const char*>(multicast_buf) +
                                                                         lay.off_acc)
                                       : nullptr;
  uint2* out = static_cast<uint2*>(dW);
  unsigned int* hf = static_cast<unsigned int*>(host_flag);
  ProfScope ps(CAT_EXCHANGE, st);
#define SSI_DP_CASE(W_)                                                                                               \
  case W_:                                                                                                            \
    if (mc_acc && W_ > 1)                                                                                             \
      ssi::dp_reduce_kernel<W_, true><<<grid, ssi::DP_THREADS, 0, st>>>(peers, lay, rank, slice_f4, total_f4, epoch,  \
                                                                        grad_out, grad_hint, accumulate_dW ? 1 : 0,   \
                                                                        out, hf, f_own, mc_acc);                      \
    else                                                                                                              \
      ssi::dp_reduce_kernel<W_><<<grid, ssi::DP_THREADS, 0, st>>>(peers, lay, rank, slice_f4, total_f4, epoch,        \
                                                                  grad_out, grad_hint, accumulate_dW ? 1 : 0, out,    \
                                                                  hf, f_own);                                         \
    SSI_LAUNCH_OK("dp_reduce");                                                                                       \
    if (W_ > 1) {                                                                                                     \
      ssi::dp_gather_kernel<W_><<<grid, ssi::DP_THREADS, 0, st>>>(peers, lay, rank, slice_f4, total_f4, epoch, out,   \
                                                                  hf, f_others);                                      \
      SSI_LAUNCH_OK("dp_gather");                                                                                     \
    }                                                                                                                 \
    break;
  switch (world) {
    SSI_DP_CASE(1)
    SSI_DP_CASE(2)
    SSI_DP_CASE(3)
    SSI_DP_CASE(4)
    SSI_DP_CASE(5)
    SSI_DP_CASE(6)
    SSI_DP_CASE(7)
    SSI_DP_CASE(8)
    default:
      return fail(-1, "unsupported world size %d", world);
  }
#undef SSI_DP_CASE
  return 0;
}

int ssi_b200_dp_embedding_exchange(const void* dH, int64_t N, int64_t V, int64_t D, void* sort_ws,
                                   size_t sort_ws_bytes, void* const* peer_bufs, int rank, int world, int64_t max_rows,
                                   uint32_t epoch, void* dW, void* host_flag, void* stream) {
  return ssi_b200_dp_embedding_exchange_fold(dH, N, V, D, sort_ws, sort_ws_bytes, peer_bufs, rank, world, max_rows,
                                             epoch, dW, host_flag, nullptr, nullptr, 0, stream);
}

int ssi_b200_dp_embedding_exchange_fold(const void* dH, int64_t N, int64_t V, int64_t D, void* sort_ws,
                                        size_t sort_ws_bytes, void* const* peer_bufs, int rank, int world,
                                        int64_t max_rows, uint32_t epoch, void* dW, void* host_flag,
                                        const float* post_scale, void* fold_ws_ptr, size_t fold_ws_bytes,
                                        void* stream) {
  TuneScope tune_scope_;
  SSI_CHECK_ARG(dH && sort_ws && dW, "null pointer argument");
  ssi::DpPeers peers;
  if (int r = dp_peers(peer_bufs, rank, world, &peers)) return r;
  SSI_CHECK_ARG(N >= 1 && N <= max_rows, "N=%lld token rows do not fit the exchange buffer (max_rows=%lld)",
                (long long)N, (long long)max_rows);
  SSI_CHECK_ARG(V >= 1 && V < (int64_t(1) << 31) - 1 && D >= 8 && D % 8 == 0, "bad shape V=%lld D=%lld", (long long)V,
                (long long)D);
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(dW) & 15) == 0, "dW must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const ssi::DpLayout lay = ssi::dp_layout(V, D, world, max_rows);
  // 1. this rank's dH rows, summed per distinct id, into its peer-visible compact list
  ssi::ScatOut o;
  memset(&o, 0, sizeof(o));
  char* own = peers.buf[rank];
  o.rows = reinterpret_cast<__nv_bfloat16*>(own + lay.off_rows);
  o.ids = reinterpret_cast<int*>(own + lay.off_ids);
  o.slot = reinterpret_cast<unsigned long long*>(own + lay.off_slot);
  o.epoch = epoch;
  o.cap = static_cast<int>(max_rows);
  if (int r = embedding_grad_scatter_impl<ssi::SCAT_COMPACT>(dH, nullptr, N, V, D, o, sort_ws, sort_ws_bytes, stream, 2))
    return r;
  // 2. every id present on any rank: the ranks' rows added in rank order to the bf16 gradient row
  unsigned int* hf = static_cast<unsigned int*>(host_flag);
  const int grid = di->sms * 4;
  __nv_bfloat16* g = static_cast<__nv_bfloat16*>(dW);
  FoldWs fw;
  if (int r = fold_ws(fold_ws_ptr, fold_ws_bytes, &fw)) return r;
  const size_t sparse_warps = static_cast<size_t>(grid) * (ssi::DP_THREADS / 32);
  SSI_CHECK_ARG(!fw.base || fw.sparse_floats() >= sparse_warps, "fold workspace too small");
  if (fw.base) SSI_CUDA(cudaMemsetAsync(fw.seg(2), 0, fw.sparse_floats() * sizeof(float), st));
  ssi::GradFold gf;
  gf.post_scale = post_scale, gf.sq = fw.seg(2);
  ProfScope ps(CAT_EXCHANGE, st);
  const int tp_opt = tune().sparse_two_phase;
  const bool two_phase = tp_opt < 0 ? world > 4 : tp_opt != 0;
#define SSI_DP_CASE(W_)                                                                                            \
  case W_:                                                                                                         \
    if (two_phase) {                                                                                               \
      ssi::dp_sparse_reduce_kernel<W_><<<grid, ssi::DP_THREADS, 0, st>>>(peers, lay, rank, static_cast<int>(V),    \
                                                                         static_cast<int>(D), epoch, hf);          \
      ssi::dp_sparse_apply_kernel<W_><<<grid, ssi::DP_THREADS, 0, st>>>(peers, lay, rank, static_cast<int>(V),     \
                                                                        static_cast<int>(D), epoch, g, hf, gf);    \
    } else {                                                                                                       \
      ssi::dp_sparse_add_kernel<W_><<<grid, ssi::DP_THREADS, 0, st>>>(peers, lay, rank, static_cast<int>(V),       \
                                                                      static_cast<int>(D), epoch, g, hf, gf);      \
    }                                                                                                              \
    break;
  switch (world) {
    SSI_DP_CASE(1)
    SSI_DP_CASE(2)
    SSI_DP_CASE(3)
    SSI_DP_CASE(4)
    SSI_DP_CASE(5)
    SSI_DP_CASE(6)
    SSI_DP_CASE(7)
    SSI_DP_CASE(8)
    default:
      return fail(-1, "unsupported world size %d", world);
  }
#undef SSI_DP_CASE
  SSI_LAUNCH_OK("dp_sparse_add");
  return 0;
}

int ssi_b200_label_shift_count(const int64_t* labels, int64_t B, int64_t S, int64_t ignore_index,
                               int64_t* labels_out, int64_t* counts, void* stream) {
  SSI_CHECK_ARG(labels && labels_out && counts, "null pointer argument");
  SSI_CHECK_ARG(B >= 1 && S >= 1, "bad shape B=%lld S=%lld", (long long)B, (long long)S);
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  SSI_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(int64_t), st));
  const int64_t total = B * S;
  int grid = static_cast<int>(ceil_div(total, 256));
  if (grid > di->sms * 8) grid = di->sms * 8;
  ssi::label_shift_count_kernel<<<grid, 256, 0, st>>>(labels, B, S, ignore_index, labels_out,
                                                      reinterpret_cast<unsigned long long*>(counts));
  SSI_LAUNCH_OK("label_shift_count");
  return 0;
}

int ssi_b200_count_valid(const int64_t* labels, int64_t n, int64_t ignore_index, int64_t* counts, void* stream) {
  SSI_CHECK_ARG(labels && counts, "null pointer argument");
  SSI_CHECK_ARG(n >= 1, "bad length %lld", (long long)n);
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  SSI_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(int64_t), st));
  int grid = static_cast<int>(ceil_div(n, 1024));
  if (grid > di->sms * 4) grid = di->sms * 4;
  ProfScope ps(CAT_AUX, st);
  ssi::count_valid_kernel<<<grid, 256, 0, st>>>(labels, n, ignore_index, reinterpret_cast<unsigned long long*>(counts));
  SSI_LAUNCH_OK("count_valid");
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// SURVEY.md 8(f) "next" rows (next_rows.cuh)
// ---------------------------------------------------------------------------------------------------
int ssi_b200_rmsnorm_fwd(const void* x, const void* scale, float eps, int64_t N, int64_t D, void* y, float* rstd,
                         void* stream) {
  SSI_CHECK_ARG(x && scale && y, "null pointer argument");
  SSI_CHECK_ARG(N >= 1 && D >= 8 && D % 8 == 0 && N < (int64_t(1) << 31), "bad shape N=%lld D=%lld", (long long)N,
                (long long)D);
  SSI_CHECK_ARG(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(scale) | reinterpret_cast<uintptr_t>(y)) & 15) == 0,
                "x / scale / y must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ProfScope ps(CAT_AUX, st);
  ssi::rmsnorm_fwd_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<const uint4*>(x), static_cast<const uint4*>(scale),
                                                       eps, static_cast<int>(N), static_cast<int>(D / 8),
                                                       static_cast<uint4*>(y), rstd);
  SSI_LAUNCH_OK("rmsnorm_fwd");
  return 0;
}

namespace {
constexpr int RMS_BWD_BLOCKS_PER_SM = 2;
}

int ssi_b200_rmsnorm_bwd_workspace_bytes(int64_t N, int64_t D, size_t* bytes) {
  SSI_CHECK_ARG(bytes, "null pointer argument");
  SSI_CHECK_ARG(N >= 1 && D >= 8 && D % 8 == 0, "bad shape N=%lld D=%lld", (long long)N, (long long)D);
  // one fp32 partial row of dscale per warp of the backward grid (148 SMs x 2 blocks x 8 warps at most)
  *bytes = static_cast<size_t>(148) * RMS_BWD_BLOCKS_PER_SM * 8 * D * sizeof(float);
  return 0;
}

int ssi_b200_rmsnorm_bwd(const void* dy, const void* x, const void* scale, const float* rstd, int64_t N, int64_t D,
                         void* dx, float* dscale, void* ws, size_t ws_bytes, void* stream) {
  SSI_CHECK_ARG(dy && x && scale && rstd && dx && dscale && ws, "null pointer argument");
  SSI_CHECK_ARG(N >= 1 && D >= 8 && D % 8 == 0 && N < (int64_t(1) << 31), "bad shape N=%lld D=%lld", (long long)N,
                (long long)D);
  SSI_CHECK_ARG(D <= 8192, "D=%lld > 8192 is not supported by the RMSNorm backward", (long long)D);
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int sms = di->sms < 148 ? di->sms : 148;
  const int grid = sms * RMS_BWD_BLOCKS_PER_SM;
  const int nparts = grid * 8;
  SSI_CHECK_ARG(ws_bytes >= static_cast<size_t>(nparts) * D * sizeof(float), "RMSNorm backward workspace too small");
  const int row_u4 = static_cast<int>(D / 8);
  const int maxj = (row_u4 + 31) / 32;
  float* partial = static_cast<float*>(ws);
  ProfScope ps(CAT_AUX, st);
#define SSI_RMS_BWD(J)                                                                                              \
  ssi::rmsnorm_bwd_kernel<J><<<grid, 256, 0, st>>>(static_cast<const uint4*>(dy), static_cast<const uint4*>(x),     \
                                                   static_cast<const uint4*>(scale), rstd, static_cast<int>(N),    \
                                                   row_u4, static_cast<uint4*>(dx), partial)
  if (maxj <= 2) SSI_RMS_BWD(2);
  else if (maxj <= 4) SSI_RMS_BWD(4);
  else if (maxj <= 8) SSI_RMS_BWD(8);
  else if (maxj <= 16) SSI_RMS_BWD(16);
  else SSI_RMS_BWD(32);
#undef SSI_RMS_BWD
  SSI_LAUNCH_OK("rmsnorm_bwd");
  ssi::rmsnorm_dscale_reduce_kernel<<<static_cast<int>(ceil_div(D, 32)), 256, 0, st>>>(partial, nparts,
                                                                                      static_cast<int>(D), dscale);
  SSI_LAUNCH_OK("rmsnorm_dscale_reduce");
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// RMSNorm backward for callers that have no device memory to give us and may not ask the driver for any.
//
// ssi_b200_rmsnorm_bwd needs one fp32 partial row of dscale per warp of its grid, and expects the caller to bring
// that as `ws`.  A host that binds this C ABI without torch's caching allocator behind it does not always have
// such a buffer: a trainer or server that carved the whole GPU into its own arena at start-up has nothing left
// for cudaMalloc to hand out, and one that runs several streams (here: the tied-gradient exchange beside the
// decoder backward) cannot afford cudaMalloc / cudaFree in the middle of a step, because both stall every stream
// of the device while they run.  For those callers the partial rows live in the array below, which is part of
// the loaded library image: the driver materialises it once per GPU when the module is loaded there and keeps
// it until the process exits (18.5 MiB per GPU, used or not).
//
// `g_rms_bwd_static_scratch` as written in HOST code is only the host-side handle of that symbol, not a device
// address, so the launch below never passes it directly: rms_bwd_static_scratch() asks the runtime for the
// address the array has on the CURRENT device (cudaGetSymbolAddress) and remembers it per device ordinal.
// ---------------------------------------------------------------------------------------------------
namespace {
constexpr int RMS_BWD_STATIC_MAX_D = 2048;  // == SSI_B200_RMSNORM_BWD_STATIC_MAX_D in the header
constexpr size_t RMS_BWD_STATIC_FLOATS = static_cast<size_t>(148) * RMS_BWD_BLOCKS_PER_SM * 8 * RMS_BWD_STATIC_MAX_D;
}  // namespace

static __device__ __align__(256) float g_rms_bwd_static_scratch[RMS_BWD_STATIC_FLOATS];

namespace {
// Device address of the array above on the current device; looked up on the first call per device (set-up work:
// the runtime may load the module on that device at this point), cached afterwards.
int rms_bwd_static_scratch(int dev, float** out) {
  static float* ptrs[64];
  static std::mutex mu;
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    SSI_CUDA(cudaGetSymbolAddress(&p, g_rms_bwd_static_scratch));
    SSI_CHECK_ARG(p && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "static scratch of device %d is not usable", dev);
    ptrs[dev] = static_cast<float*>(p);
  }
  *out = ptrs[dev];
  return 0;
}
}  // namespace

int ssi_b200_rmsnorm_bwd_static(const void* dy, const void* x, const void* scale, const float* rstd, int64_t N,
                                int64_t D, void* dx, float* dscale, void* stream) {
  SSI_CHECK_ARG(dy && x && scale && rstd && dx && dscale, "null pointer argument");
  SSI_CHECK_ARG(N >= 1 && D >= 8 && D % 8 == 0 && N < (int64_t(1) << 31), "bad shape N=%lld D=%lld", (long long)N,
                (long long)D);
  SSI_CHECK_ARG(D <= RMS_BWD_STATIC_MAX_D,
                "D=%lld exceeds the %d columns the built-in scratch of ssi_b200_rmsnorm_bwd_static holds; use "
                "ssi_b200_rmsnorm_bwd with a caller-provided workspace",
                (long long)D, RMS_BWD_STATIC_MAX_D);
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  float* partial = nullptr;
  if (int r = rms_bwd_static_scratch(di->dev, &partial)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int sms = di->sms < 148 ? di->sms : 148;
  const int grid = sms * RMS_BWD_BLOCKS_PER_SM;
  const int nparts = grid * 8;  // nparts * D <= RMS_BWD_STATIC_FLOATS by the checks above
  const int row_u4 = static_cast<int>(D / 8);
  const int maxj = (row_u4 + 31) / 32;
  ProfScope ps(CAT_AUX, st);
  // Every warp of the grid overwrites its whole partial row, rows of an earlier call included, so nothing in the
  // shared array has to be cleared between calls.
#define SSI_RMS_BWD_STATIC(J)                                                                                        \
  ssi::rmsnorm_bwd_kernel<J><<<grid, 256, 0, st>>>(static_cast<const uint4*>(dy), static_cast<const uint4*>(x),     \
                                                   static_cast<const uint4*>(scale), rstd, static_cast<int>(N),    \
                                                   row_u4, static_cast<uint4*>(dx), partial)
  if (maxj <= 2) SSI_RMS_BWD_STATIC(2);
  else if (maxj <= 4) SSI_RMS_BWD_STATIC(4);
  else SSI_RMS_BWD_STATIC(8);
#undef SSI_RMS_BWD_STATIC
  SSI_LAUNCH_OK("rmsnorm_bwd_static");
  ssi::rmsnorm_dscale_reduce_kernel<<<static_cast<int>(ceil_div(D, 32)), 256, 0, st>>>(partial, nparts,
                                                                                      static_cast<int>(D), dscale);
  SSI_LAUNCH_OK("rmsnorm_dscale_reduce_static");
  return 0;
}

int ssi_b200_label_shift_count_types(const int64_t* labels, const int64_t* tokens, int64_t B, int64_t S,
                                     int64_t ignore_index, const int64_t* range_lo, const int64_t* range_hi,
                                     int n_ranges, int64_t pad_idx, int64_t* labels_out, int64_t* counts,
                                     int64_t* type_counts, void* stream) {
  SSI_CHECK_ARG(labels && tokens && labels_out && counts && type_counts, "null pointer argument");
  SSI_CHECK_ARG(B >= 1 && S >= 1, "bad shape B=%lld S=%lld", (long long)B, (long long)S);
  SSI_CHECK_ARG(n_ranges >= 0 && n_ranges <= ssi::MAX_TOKEN_RANGES && (n_ranges == 0 || (range_lo && range_hi)),
                "n_ranges=%d not in [0, %d] or null range arrays (HOST arrays)", n_ranges, ssi::MAX_TOKEN_RANGES);
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ssi::TokenRanges tr;
  memset(&tr, 0, sizeof(tr));
  tr.n = n_ranges;
  for (int k = 0; k < n_ranges; ++k) tr.lo[k] = range_lo[k], tr.hi[k] = range_hi[k];
  SSI_CUDA(cudaMemsetAsync(counts, 0, 2 * sizeof(int64_t), st));
  SSI_CUDA(cudaMemsetAsync(type_counts, 0, (n_ranges + 1) * sizeof(int64_t), st));
  const int64_t total = B * S;
  int grid = static_cast<int>(ceil_div(total, 256));
  if (grid > di->sms * 8) grid = di->sms * 8;
  ProfScope ps(CAT_AUX, st);
  ssi::label_shift_count_types_kernel<<<grid, 256, 0, st>>>(labels, tokens, B, S, ignore_index, tr, pad_idx, labels_out,
                                                            reinterpret_cast<unsigned long long*>(counts),
                                                            reinterpret_cast<unsigned long long*>(type_counts));
  SSI_LAUNCH_OK("label_shift_count_types");
  return 0;
}

namespace {
constexpr int SQNORM_BLOCKS = 148 * 8;
}

int ssi_b200_grad_scale_sqnorm_workspace_bytes(size_t* bytes) {
  SSI_CHECK_ARG(bytes, "null pointer argument");
  *bytes = SQNORM_BLOCKS * sizeof(float);
  return 0;
}

int ssi_b200_grad_scale_sqnorm(void* grad, int64_t n_elems, const float* scale, float* sqnorm_out, void* ws,
                               size_t ws_bytes, void* stream) {
  SSI_CHECK_ARG(grad && scale && sqnorm_out && ws, "null pointer argument");
  SSI_CHECK_ARG(n_elems >= 8 && n_elems % 8 == 0, "element count %lld must be a positive multiple of 8",
                (long long)n_elems);
  SSI_CHECK_ARG((reinterpret_cast<uintptr_t>(grad) & 15) == 0, "grad must be 16-byte aligned");
  SSI_CHECK_ARG(ws_bytes >= SQNORM_BLOCKS * sizeof(float), "workspace too small");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t n_u4 = static_cast<size_t>(n_elems) / 8;
  int grid = static_cast<int>(ceil_div(static_cast<int64_t>(n_u4), 256));
  if (grid > SQNORM_BLOCKS) grid = SQNORM_BLOCKS;
  ProfScope ps(CAT_AUX, st);
  ssi::grad_scale_sqnorm_kernel<<<grid, 256, 0, st>>>(static_cast<uint4*>(grad), n_u4, scale, static_cast<float*>(ws));
  SSI_LAUNCH_OK("grad_scale_sqnorm");
  ssi::sqnorm_finalize_kernel<<<1, 256, 0, st>>>(static_cast<const float*>(ws), grid, sqnorm_out);
  SSI_LAUNCH_OK("sqnorm_finalize");
  return 0;
}

int ssi_b200_grad_add(void* dst, const void* src, int64_t n_elems, void* stream) {
  SSI_CHECK_ARG(dst && src, "null pointer argument");
  SSI_CHECK_ARG(n_elems >= 8 && n_elems % 8 == 0, "element count %lld must be a positive multiple of 8",
                (long long)n_elems);
  SSI_CHECK_ARG(((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0,
                "dst / src must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ProfScope ps(CAT_AUX, st);
  ssi::grad_add_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<uint4*>(dst), static_cast<const uint4*>(src),
                                                    static_cast<size_t>(n_elems) / 8);
  SSI_LAUNCH_OK("grad_add");
  return 0;
}

int ssi_b200_single_pass_lockstep_status(const void* ws, int64_t N, int64_t V, int64_t D, int* gave_up) {
  TuneScope tune_scope_;
  SSI_CHECK_ARG(ws && gave_up, "null pointer argument");
  SpPlan f;
  if (int r = plan_single_pass(N, V, D, &f)) return r;
  unsigned int v[2] = {0, 0};  // {exp-stash overflow, lockstep gave up}
  SSI_CUDA(cudaMemcpy(v, static_cast<const char*>(ws) + f.off_sync + 1022 * sizeof(unsigned int), sizeof(v),
                      cudaMemcpyDeviceToHost));
  *gave_up = (v[1] ? 1 : 0) | (v[0] ? 2 : 0);
  return 0;
}

int ssi_b200_embedding_gather(const void* W, const int64_t* ids, int64_t N, int64_t V, int64_t D, void* out,
                              void* stream) {
  SSI_CHECK_ARG(W && ids && out, "null pointer argument");
  SSI_CHECK_ARG(N >= 1 && V >= 1 && D >= 8 && D % 8 == 0 && N < (int64_t(1) << 31), "bad shape N=%lld V=%lld D=%lld",
                (long long)N, (long long)V, (long long)D);
  SSI_CHECK_ARG(((reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                "W / out must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ProfScope ps(CAT_AUX, st);
  ssi::embedding_gather_kernel<<<di->sms * 8, 256, 0, st>>>(static_cast<const uint4*>(W), ids, static_cast<int>(N), V,
                                                            static_cast<int>(D / 8), static_cast<uint4*>(out));
  SSI_LAUNCH_OK("embedding_gather");
  return 0;
}

int ssi_b200_plan_info(int64_t N, int64_t V, int64_t D, int64_t* out) {
  TuneScope tune_scope_;
  SSI_CHECK_ARG(out, "null pointer argument");
  SSI_CHECK_ARG(N >= 1 && V >= 1 && D >= 64 && D % 64 == 0, "bad shape N=%lld V=%lld D=%lld", (long long)N,
                (long long)V, (long long)D);
  const FwdPlan f = plan_fwd(N, V, D);
  const BwdPlan b = plan_bwd(N, V, D);
  out[0] = 0, out[1] = 0, out[2] = f.tiles_m, out[3] = f.vtiles_total;  // splits are chosen on the device
  out[4] = b.nblocks, out[5] = b.vtiles_per_block, out[6] = b.vc, out[7] = b.n_pad;
  return 0;
}

int ssi_b200_profile_begin(void) {
  std::lock_guard<std::mutex> lk(g_prof.mu);
  g_prof.recs.clear();
  g_prof.used = 0;
  g_prof.on = true;
  return 0;
}

int ssi_b200_profile_end(double* ms_by_class, int64_t* launches_by_class) {
  std::lock_guard<std::mutex> lk(g_prof.mu);
  g_prof.on = false;
  for (int i = 0; i < CAT_COUNT; ++i) {
    if (ms_by_class) ms_by_class[i] = 0.0;
    if (launches_by_class) launches_by_class[i] = 0;
  }
  for (const ProfRec& r : g_prof.recs) {
    SSI_CUDA(cudaEventSynchronize(r.b));
    float ms = 0.f;
    SSI_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
    if (ms_by_class) ms_by_class[r.cat] += ms;
    if (launches_by_class) launches_by_class[r.cat] += 1;
  }
  g_prof.recs.clear();
  g_prof.used = 0;
  return 0;
}

int ssi_b200_debug_gemm(const void* A, const void* B, float* C, int64_t M, int64_t N, int64_t K, int a_mn_major,
                        int b_mn_major, int accumulate, void* stream) {
  TuneScope tune_scope_;
  SSI_CHECK_ARG(A && B && C, "null pointer argument");
  SSI_CHECK_ARG(M >= 1 && N >= 1 && K >= 1, "bad shape");
  SSI_CHECK_ARG(K % 8 == 0 && M % 8 == 0 && N % 8 == 0, "M, N, K must be multiples of 8 (16-byte TMA strides)");
  DeviceInfo* di = nullptr;
  if (int r = device_info(&di)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ssi::GemmParams p;
  memset(&p, 0, sizeof(p));
  if (!a_mn_major) {
    if (int r = make_tmap(&p.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, K, M, K * 2, 64, ssi::BM)) return r;
  } else {
    if (int r = make_tmap(&p.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, M, K, M * 2, 64, 64)) return r;
  }
  if (!b_mn_major) {
    if (int r = make_tmap(&p.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B, K, N, K * 2, 64, ssi::BN / cta_group())) return r;
  } else {
    if (int r = make_tmap(&p.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B, N, K, N * 2, 64, 64)) return r;
  }
  if (int r = make_tmap(&p.tmC, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, C, N, M, N * 4, 32, ssi::BM)) return r;
  p.tiles_m = static_cast<int>(ceil_div(M, ssi::BM));
  p.tiles_n = static_cast<int>(ceil_div(N, ssi::BN));
  p.k_blocks = static_cast<int>(ceil_div(K, ssi::BK));
  p.group_m = 8;
  p.c_reduce = accumulate ? 1 : 0;
  if (!a_mn_major && !b_mn_major) return launch_gemm<ssi::EPI_F32, false, false>(p, di->sms, st, "debug_gemm_kk", CAT_AUX);
  if (a_mn_major && b_mn_major)
    return launch_gemm_wide<ssi::EPI_F32, true, true>(p, di->sms, st, "debug_gemm_mm", CAT_AUX, tune().wide_dw);
  if (!a_mn_major && b_mn_major)
    return launch_gemm_wide<ssi::EPI_F32, false, true>(p, di->sms, st, "debug_gemm_km", CAT_AUX, tune().wide_dx);
  return fail(-1, "unsupported operand majors (a_mn=%d, b_mn=%d)", a_mn_major, b_mn_major);
}

}  // extern "C"
