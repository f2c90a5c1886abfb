/*
 * ssi_b200 — C ABI of the B200-native (sm_100a) fused tied-LM-head + chunked-cross-entropy path of
 * anilkeshwani/speech-integration.  This header is the drop-in boundary: plain pointers and sizes, no
 * torch types.  Every entry point only ENQUEUES work on `stream` (a cudaStream_t passed as void*), never
 * synchronises the device and never allocates; all pointers are device pointers owned by the caller.
 *
 * Return convention: 0 = OK; negative = argument / shape / alignment error detected on the host before
 * any launch (text via ssi_b200_last_error(), thread-local); positive = cudaError_t of a failed launch.
 * There is no CPU fallback and no multi-backend dispatch: a device that is not sm_100 is a hard error.
 *
 * Layouts (all row-major, contiguous):
 *   X  [N, D]  bf16  final (normed) hidden states, N = B*S token rows
 *   W  [V, D]  bf16  tied embedding / LM-head weight (tok_embeddings.weight)
 *   labels [N] int64 already shifted as ssi/loss.py:16 does; ignore_index rows contribute nothing
 *   dX [N, D]  bf16 ; dW [V, D] bf16 (the tied weight's .grad buffer)
 * Requirements: D % 64 == 0, all base pointers 16-byte aligned (TMA), N >= 1, V >= 1.
 *
 * Reference interfaces replaced (paths relative to the reference checkout; tt/ = third_party/torchtune/):
 *   ssi_b200_lmhead_ce_fwd   <- tt/torchtune/modules/transformer.py:499-502 (chunked_output: F.linear per chunk,
 *                               tt/torchtune/modules/tied_linear.py:25) + tt/torchtune/modules/loss/
 *                               ce_chunked_output_loss.py:45-83 (CEWithChunkedOutputLoss.forward)
 *   ssi_b200_lmhead_ce_bwd   <- autograd of the above as driven by ssi/trainer.py:393-394
 *   ssi_b200_embedding_grad_scatter <- backward of nn.Embedding built at tt/torchtune/models/llama3_2/
 *                               _component_builders.py:114 and called at tt/torchtune/modules/transformer.py:628
 *   ssi_b200_label_shift_count <- ssi/loss.py:16 (label shift) + ssi/trainer.py:391 (token count)
 *   ssi_b200_dp_*            <- no reference counterpart (the reference trains on one GPU); the exchange step of
 *                               SURVEY.md 8(e), semantics of tt/recipes/full_finetune_distributed.py:844-851,867
 */
#ifndef SSI_B200_H_
#define SSI_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSI_B200_ABI_VERSION 114

/* ABI version of the loaded library (== SSI_B200_ABI_VERSION it was built with). */
int ssi_b200_version(void);

/* Thread-local text of the last error returned on this thread ("" if none). */
const char* ssi_b200_last_error(void);

/* Number of tcgen05 / helper kernels launched by this process through the library so far
 * (used by bench.py for its gpu_launches claim). */
int64_t ssi_b200_launch_count(void);

/* Tuning knobs (also read once from the environment as SSI_B200_<NAME>): "g_cap_mib" = size cap of the bf16
 * logits/dlogits block scratch in MiB (default 4608; workspace sizes depend on it: query them again after changing
 * it), "cta_group" (2 = CTA pairs, default), "wide_dw" / "wide_dx" (2 = 256x512 tiles for the long-K contractions,
 * default), "tail_split" (1 = a last wave that is at most half full runs as 256x256 halves, default), "shape_gate" (1 = the 256x256 kernel is enqueued as well and the device-side plan picks, default), "lockstep"
 * (k-blocks between sync points of dW / dX, default 64, 0 = off), "lockstep_k1" (1 = K1 sweeps its vocabulary tiles
 * in lockstep, default), "exp_stash" (1 = single-pass K1 stashes exp(logit - ref) and the softmax normalisation is a
 * per-row factor, default; 0 = stash logits + a separate softmax-grad pass), "sparse_two_phase" (data-parallel
 * embedding exchange: -1 = two phases when world > 4, default; 0 / 1 force), "k1_group_m" (m-tiles per raster group
 * of the single-pass K1, default 64, 0 = all) and "k1_store_first" (1 = evict-first L2 policy on its E-tile stores,
 * default), "dp_timeout_s" (see ssi_b200_dp_status), "g_tiles", "group_m_grad", "group_m_dw", "group_m_dx", "store_hint", "k1_hints", "dw_hints",
 * "dx_hints" (experiments; see profiles/r01_tuning_notes.md). */
int ssi_b200_set_option(const char* name, int64_t value);
/* The value an option has now (the names of ssi_b200_set_option).  The data-parallel set-up compares the options
 * that every rank must agree on ("sparse_two_phase": ranks that disagree wait on different flags) across ranks. */
int ssi_b200_get_option(const char* name, int64_t* value);

/* Workspace sizes (bytes) for the given problem. Any out pointer may be NULL. Workspaces must be 256-byte
 * aligned. */
int ssi_b200_workspace_bytes(int64_t N, int64_t V, int64_t D, size_t* fwd_bytes, size_t* bwd_bytes,
                             size_t* scatter_bytes);

/* Tile schedule the library will use for this problem: out[8] = {forward vocabulary splits, 256-column tiles
 * per split, 128-row token tiles, 256-column vocabulary tiles, backward dlogits blocks, tiles per block,
 * block width in columns, padded rows}. Host only. */
int ssi_b200_plan_info(int64_t N, int64_t V, int64_t D, int64_t* out);

/*
 * Forward: logits = X * W^T are produced tile by tile in tensor memory and consumed by an online
 * log-sum-exp epilogue; they never reach HBM.
 *   lse      [N]  out  fp32 natural-log logsumexp per row (saved for backward); 0 for ignored rows, which are
 *                      compacted away before the tensor-core stages and never touch them
 *   loss_out [2]  out  fp32 {sum over valid rows of (lse - z_label), that sum / n_valid}
 *                      (n_valid == 0 -> loss_out[1] = NaN, as 0/0 in ce_chunked_output_loss.py:83)
 *   n_valid  [1]  out  int64 count of labels != ignore_index (ce_chunked_output_loss.py:66)
 */
int ssi_b200_lmhead_ce_fwd(const void* X, const void* W, const int64_t* labels, int64_t N, int64_t V, int64_t D,
                           int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* ws,
                           size_t ws_bytes, void* stream);

/*
 * Backward: recomputes the logits tiles, forms dlogits = (softmax - onehot) * grad_out / n_valid_total in
 * bf16 one vocabulary block at a time (bounded scratch inside `ws`), and contracts
 *   dX = dlogits * W      (overwritten)
 *   dW (+)= dlogits^T * X (accumulate_dW != 0 adds to the existing buffer, else overwrites)
 * on the tensor cores.
 *   grad_out      [1] fp32 device scalar: upstream dL/dloss (ssi/trainer.py:393 multiplies the loss by
 *                     num_tokens_iter, so this is that count)
 *   n_valid_total [1] int64 device scalar: divisor of the mean (the local n_valid from forward, or the
 *                     all-reduced global count under data parallelism)
 */
int ssi_b200_lmhead_ce_bwd(const void* X, const void* W, const int64_t* labels, const float* lse,
                           const float* grad_out, const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                           int64_t ignore_index, void* dX, void* dW, int accumulate_dW, void* ws, size_t ws_bytes,
                           void* stream);

/*
 * Single-pass mode (the default for training): forward ALSO forms the gradients, one bounded ROW chunk at a
 * time, so each logits tile is computed once (6 N V D flop executed instead of the 8 N V D of fwd + recompute
 * bwd).  Per chunk: the logits tiles are rounded to bf16 -- exactly the dtype the reference's logits have
 * (tied_linear.py:25) -- and stored into one reused block S [rows_cap, Vpad] inside `ws` (never the full
 * [N, V] matrix); the log-sum-exp runs on those rounded values (ce_chunked_output_loss.py:41); S is rewritten
 * in place into dlogits = (softmax - onehot) * grad_hint / n_valid_total (bf16, like autograd's); then
 *   dW_acc (+)= dlogits^T * X_chunk   fp32 [V, D], summed over ALL tokens before any rounding
 *   dX_chunk  = dlogits * W           bf16
 *   grad_hint     [1] fp32 device scalar: the upstream gradient the caller expects backward to receive
 *                     (ssi/trainer.py:393: num_tokens_iter); any value works, see apply_grad
 *   n_valid_total [1] int64 device scalar or NULL (= use this call's own n_valid)
 *   lse/loss_out/n_valid  as ssi_b200_lmhead_ce_fwd
 *   dX     [N, D] bf16 out (zeros for ignored rows);  dW_acc [V, D] fp32 out (overwritten)
 */
int ssi_b200_single_pass_workspace_bytes(int64_t N, int64_t V, int64_t D, size_t* bytes);
int ssi_b200_lmhead_ce_fwd_grad(const void* X, const void* W, const int64_t* labels, const float* grad_hint,
                                const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                                int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* dX,
                                float* dW_acc, void* ws, size_t ws_bytes, void* stream);

/* Data-parallel variant of the above (see ssi_b200_dp_* below), three extra arguments:
 *   accumulate_dW_acc  != 0: the caller initialised dW_acc (e.g. zeros + the embedding scatter-add) and every
 *                      dW contraction ADDS to it; 0: the first one overwrites it as above
 *   dw_acc_init_event  cudaEvent_t or NULL: `stream` waits for it right before the first dW contraction (the
 *                      initialisation may run on another stream beside the first logits contraction)
 *   dw_ready_event     cudaEvent_t or NULL: recorded on `stream` right after the LAST dW contraction, i.e. as soon
 *                      as dW_acc is final and while the last dX contraction is still to run, so that the exchange
 *                      of the tied gradient can start on another stream behind it. */
int ssi_b200_lmhead_ce_fwd_grad_dp(const void* X, const void* W, const int64_t* labels, const float* grad_hint,
                                   const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                                   int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* dX,
                                   float* dW_acc, void* ws, size_t ws_bytes, void* stream, int accumulate_dW_acc,
                                   void* dw_acc_init_event, void* dw_ready_event);

/* The same with the final RMSNorm (tt/torchtune/modules/rms_norm.py:38-43, applied at transformer.py:653) fused in:
 * when `norm_scale` (bf16 [D]) is not NULL, X holds the PRE-norm hidden states; the norm is applied while the valid
 * rows are gathered (no separate pass), its backward runs while the compact dX rows are scattered back, so dX is
 * d(pre-norm h) and `dscale` [D] fp32 receives the gradient of the scale (for the upstream gradient grad_hint;
 * scale it by grad_out / grad_hint like dW_acc). norm_scale == NULL: identical to ssi_b200_lmhead_ce_fwd_grad_dp. */
int ssi_b200_lmhead_ce_fwd_grad_norm(const void* X, const void* W, const int64_t* labels, const float* grad_hint,
                                     const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                                     int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* dX,
                                     float* dW_acc, void* ws, size_t ws_bytes, void* stream, int accumulate_dW_acc,
                                     void* dw_acc_init_event, void* dw_ready_event, const void* norm_scale,
                                     float norm_eps, float* dscale);

/* The general form: every optional argument of the variants above in one struct (zero-initialise it), plus
 *   pre_grad_event  cudaEvent_t or NULL: `stream` waits for it right before the FIRST softmax-grad pass, the first
 *                   consumer of *grad_hint and *n_valid_total -- under data parallelism both are produced by
 *                   ssi_b200_dp_counts on another stream beside the first logits contraction. */
typedef struct ssi_b200_fwd_grad_opts {
  int accumulate_dW_acc;
  void* dw_acc_init_event;
  void* dw_ready_event;
  void* pre_grad_event;
  const void* norm_scale;
  float norm_eps;
  float* dscale;
} ssi_b200_fwd_grad_opts;
int ssi_b200_lmhead_ce_fwd_grad_ex(const void* X, const void* W, const int64_t* labels, const float* grad_hint,
                                   const int64_t* n_valid_total, int64_t N, int64_t V, int64_t D,
                                   int64_t ignore_index, float* lse, float* loss_out, int64_t* n_valid, void* dX,
                                   float* dW_acc, void* ws, size_t ws_bytes, void* stream,
                                   const ssi_b200_fwd_grad_opts* opts);

/* Synchronising debug query on the last single-pass call that used `ws`: bit 0 of *gave_up is set if a lockstep wait
 * timed out (part of the persistent grid was not resident; the kernels then ran un-throttled -- slower, still
 * correct); bit 1 if the exp-stash form overflowed on some row -- a logit more than ~88 nats above the largest of the
 * row's first 256 vocabulary columns: that row's loss is inf and the gradients of the call are NaN (never silently
 * wrong); rerun with option "exp_stash" = 0, which has no such range limit. */
int ssi_b200_single_pass_lockstep_status(const void* ws, int64_t N, int64_t V, int64_t D, int* gave_up);

/*
 * Backward of single-pass mode: ratio = grad_out / grad_hint (exactly 1 when the hint was right);
 *   dX *= ratio in place (skipped on the device when ratio == 1; dX may be NULL)
 *   dW  = bf16((accumulate_dW ? dW : 0) + ratio * dW_acc)     -- one rounding of the whole-batch fp32 sum
 *        (dW may be NULL: the data-parallel exchange below then does this step for the summed gradient)
 */
int ssi_b200_lmhead_ce_apply_grad(void* dX, const float* dW_acc, const float* grad_out, const float* grad_hint,
                                  int64_t N, int64_t V, int64_t D, void* dW, int accumulate_dW, void* stream);

/*
 * Tied-embedding gradient: dW[ids[i], :] += dH[i, :] for i in [0, N), ids in [0, V) (ids < 0 or >= V are
 * skipped).  Rows are grouped by a stable radix sort of the ids (one launch: a single-block LSD sort with 9-bit
 * digits) and reduced in fp32 by a warp-level segmented reduction in sorted order (deterministic: no
 * floating-point atomics), one read-modify-write of each touched dW row.  Three launches in all.
 */
int ssi_b200_embedding_grad_scatter(const void* dH, const int64_t* ids, int64_t N, int64_t V, int64_t D, void* dW,
                                    void* ws, size_t ws_bytes, void* stream);

/* The same scatter-add into an fp32 [V, D] buffer (data-parallel mode: the embedding gradient joins the fp32
 * dW_acc of single-pass mode before the exchange, so the summed tied gradient is rounded to bf16 exactly once). */
int ssi_b200_embedding_grad_scatter_f32(const void* dH, const int64_t* ids, int64_t N, int64_t V, int64_t D,
                                        float* dW_acc, void* ws, size_t ws_bytes, void* stream);
/* The two halves of the above as separate calls sharing `ws`: the radix sort of the ids (needs only the token ids,
 * its kernels use shared memory) can run at the start of the step, so that the reduction (no shared memory) can
 * later co-run with a tensor-core contraction on another stream.  Same N, V, D and workspace in both calls. */
int ssi_b200_embedding_grad_sort(const int64_t* ids, int64_t N, int64_t V, int64_t D, void* ws, size_t ws_bytes,
                                 void* stream);
int ssi_b200_embedding_grad_scatter_sorted_f32(const void* dH, int64_t N, int64_t V, int64_t D, float* dW_acc,
                                               void* ws, size_t ws_bytes, void* stream);
/* ... and the reduction into a bf16 [V, D] buffer (the tied weight's .grad) from an earlier ssi_b200_embedding_grad_sort
 * (the sort needs only the token ids, so the embedding FORWARD can start it). */
int ssi_b200_embedding_grad_scatter_sorted(const void* dH, int64_t N, int64_t V, int64_t D, void* dW, void* ws,
                                           size_t ws_bytes, void* stream);

/*
 * Data-parallel exchange of the tied LM-head / embedding gradient over NVLink peer memory (one process per GPU,
 * one NVSwitch node; the reference is single-process, semantics follow tt/recipes/full_finetune_distributed.py:
 * 844-851,867: gradients of the global mean).  Each rank owns one exchange buffer
 *     [ctrl 4 KiB][dW_acc fp32 V x D][stage bf16 ceil(V / world) x D][slot u64 V][emb_ids i32 R][emb_rows bf16 R x D]
 * (R = max_rows = the largest number of token rows one call hands to dp_embedding_exchange) that its peers map through
 * CUDA IPC.  dp_buffer_bytes gives its size and the offset of dW_acc (pass that address as `dW_acc` of
 * ssi_b200_lmhead_ce_fwd_grad*).  dp_alloc / dp_open / dp_close / dp_free are SET-UP calls (they allocate and
 * synchronise); `ipc_handle` is 64 bytes to be carried to the peers by the host (torch.distributed all_gather).
 * The three exchanges of a step, in the order a training step produces their inputs (all enqueue-only; `epoch` is a
 * counter the caller increments by one per step, starting at 1, identical on all ranks; `host_flag` is NULL or a
 * pinned, device-mapped host word that a timed-out wait sets to 1 so the host can notice without synchronising --
 * after a time-out every exchange on every rank writes NaN instead of sums until ssi_b200_dp_reset):
 *
 * dp_counts             global_counts[0..1] = sum over ranks of local_counts[0..1] ({unshifted, shifted} valid-token
 *                       counts, ssi/trainer.py:391 and ce_chunked_output_loss.py:66) and, if hint_out != NULL,
 *                       hint_out[0] = float(global_counts[0]) (the multiplier of ssi/trainer.py:393).  Eight-byte
 *                       peer stores and flags, no collective library call.
 * dp_exchange           the LM-head part, as soon as dW_acc is final:
 *                         dW = bf16((accumulate_dW ? dW : 0) + (grad_out / grad_hint) * sum_ranks dW_acc_rank)
 *                       on every rank, bit-identical across ranks and runs: rank r reduces vocabulary slice r in fp32
 *                       from its peers' buffers (P2P loads) in rank order, rounds once, and all ranks then pull the
 *                       finished bf16 slices.  No shared memory: the kernels co-reside with the tensor-core CTAs of
 *                       the dX contraction still running on another stream.
 * dp_embedding_exchange the embedding part, once dH (the gradient reaching the embedding output) exists -- in a
 *                       training step that is AFTER the decoder backward, so it cannot join dW_acc before the dense
 *                       exchange.  `sort_ws` holds the order left by ssi_b200_embedding_grad_sort for the same
 *                       N, V, D.  This rank's dH rows are summed per distinct id (fp32, sorted order) into its compact
 *                       list (emb_rows bf16, slot[id] = epoch | index); then every rank adds, for every id present on
 *                       any rank, the ranks' rows in rank order in fp32 to its bf16 dW row (one rounding, identical on
 *                       all ranks).  Must follow this step's dp_exchange in stream order.  N <= max_rows.
 *                       world > 4 (option "sparse_two_phase": -1 auto, 0, 1): two phases -- the owner of an id
 *                       (id % world) sums the ranks' rows once (fp32, rank order, one rounding to bf16), every rank
 *                       then adds the reduced row: one row pulled per id instead of one per holder.
 */
int ssi_b200_dp_buffer_bytes(int64_t V, int64_t D, int world, int64_t max_rows, size_t* bytes,
                             size_t* dw_acc_offset);
int ssi_b200_dp_alloc(size_t bytes, void** ptr, void* ipc_handle);
int ssi_b200_dp_open(const void* ipc_handle, void** ptr);
int ssi_b200_dp_close(void* ptr);
int ssi_b200_dp_free(void* ptr);
/* Synchronising check (set-up / tear-down class): *timed_out != 0 if a wait inside an exchange -- on this rank or on
 * any peer, which then marks every rank -- gave up because a rank did not arrive within the time limit; every exchange
 * since writes NaN.  The limit is option "dp_timeout_s" / SSI_B200_DP_TIMEOUT_S (seconds, default 600 = the default
 * time-out of a torch.distributed process group), read by dp_alloc and dp_reset. */
int ssi_b200_dp_status(const void* own_buf, int* timed_out);
/* Collective recovery after a time-out (set-up class: synchronises the device): clears this rank's mark and
 * re-reads the time limit.  Every rank must call it between two host-side barriers (no exchange in flight anywhere),
 * and all ranks must continue with the same `epoch`. */
int ssi_b200_dp_reset(void* own_buf);
int ssi_b200_dp_counts(void* const* peer_bufs, int rank, int world, uint32_t epoch, const int64_t* local_counts,
                       int64_t* global_counts, float* hint_out, void* host_flag, void* stream);
int ssi_b200_dp_exchange(void* const* peer_bufs, int rank, int world, int64_t V, int64_t D, int64_t max_rows,
                         uint32_t epoch, const float* grad_out, const float* grad_hint, void* dW, int accumulate_dW,
                         void* host_flag, void* stream);
int ssi_b200_dp_embedding_exchange(const void* dH, int64_t N, int64_t V, int64_t D, void* sort_ws,
                                   size_t sort_ws_bytes, void* const* peer_bufs, int rank, int world, int64_t max_rows,
                                   uint32_t epoch, void* dW, void* host_flag, void* stream);

/* dst = bf16(float(dst) + float(src)) over n_elems bf16 values (NCCL exchange with gradient accumulation: only the
 * NEW micro-batch's contribution is summed over ranks, then joins the existing gradient with this call). */
int ssi_b200_grad_add(void* dst, const void* src, int64_t n_elems, void* stream);

/*
 * labels_out[b, s] = labels[b, s+1] for s < S-1, ignore_index for s == S-1 (ssi/loss.py:16);
 * counts[0] = #(labels != ignore_index) on the UNSHIFTED labels (ssi/trainer.py:391),
 * counts[1] = the same count on the shifted labels.
 */
int ssi_b200_label_shift_count(const int64_t* labels, int64_t B, int64_t S, int64_t ignore_index,
                               int64_t* labels_out, int64_t* counts, void* stream);

/* counts[0] = counts[1] = #(labels != ignore_index) over a flat array of n (already shifted) labels
 * (ce_chunked_output_loss.py:66): the local count a data-parallel loss call posts when the trainer announced none. */
int ssi_b200_count_valid(const int64_t* labels, int64_t n, int64_t ignore_index, int64_t* counts, void* stream);

/*
 * ---- Rows SURVEY.md 8(f) marks "next": the callers and data formats either side of the path (HBM-bound) ----
 *
 * Final RMSNorm (tt/torchtune/modules/rms_norm.py:38-43, applied at tt/torchtune/modules/transformer.py:653 right
 * before the projection): y = bf16((float(x) * rsqrt(mean(float(x)^2) + eps)) * float(scale)); x, y [N, D] bf16,
 * scale [D] bf16, rstd [N] fp32 (may be NULL in forward-only use).  Backward (fp32 math like autograd of the fp32
 * F.rms_norm): g = dy * scale, dx = rstd * (g - x * rstd^2 * mean(g * x)) -> bf16; dscale [D] fp32 =
 * sum over rows of dy * x * rstd (fixed-order partials in `ws`, deterministic).  D % 8 == 0, D <= 8192 (backward).
 */
int ssi_b200_rmsnorm_fwd(const void* x, const void* scale, float eps, int64_t N, int64_t D, void* y, float* rstd,
                         void* stream);
int ssi_b200_rmsnorm_bwd_workspace_bytes(int64_t N, int64_t D, size_t* bytes);
int ssi_b200_rmsnorm_bwd(const void* dy, const void* x, const void* scale, const float* rstd, int64_t N, int64_t D,
                         void* dx, float* dscale, void* ws, size_t ws_bytes, void* stream);

/*
 * ssi_b200_rmsnorm_bwd without the `ws` argument.  This one has no counterpart in the reference (which leaves all
 * memory management to PyTorch); it exists for hosts of this C ABI that cannot obtain device scratch at the point
 * where the backward runs:
 *   - a trainer or server that handed the whole GPU to its own arena at start-up, so that cudaMalloc has nothing
 *     left to give (and no torch caching allocator sits behind the binding to ask instead);
 *   - a step that keeps several streams busy (the tied-gradient exchange beside the decoder backward): cudaMalloc
 *     and cudaFree stall every stream of the device while they run, so allocating mid-step costs the overlap.
 * The per-warp partial rows of dscale therefore live in a fixed array inside the library image itself:
 * 148 x 2 x 8 x SSI_B200_RMSNORM_BWD_STATIC_MAX_D floats = 18.5 MiB, which every GPU that loads the library holds
 * until the process exits whether this entry point is used or not.  Consequences a caller must know:
 *   - D > SSI_B200_RMSNORM_BWD_STATIC_MAX_D is refused (negative return, text in ssi_b200_last_error()); use
 *     ssi_b200_rmsnorm_bwd with your own workspace for wider rows.  N is not limited by the array.
 *   - The FIRST call on each device (with that device current) also looks up where the array sits on that device.
 *     That is set-up work: make this first call before you begin any stream capture on that device.
 *   - All calls on one GPU share the one array, and it keeps whatever the previous call left in it.  The kernels
 *     overwrite every partial row they later read, so nothing needs clearing between calls -- but two calls whose
 *     kernels can be on the GPU at the same time (different streams without an ordering between them, or two host
 *     threads) would overwrite each other's partial sums.  Order them on one stream or with events.
 *   - The call itself allocates nothing and frees nothing: no cudaMalloc / cudaFree, no library routine that
 *     allocates.  dy, x, scale, rstd, dx and dscale are the caller's, exactly as for ssi_b200_rmsnorm_bwd, and the
 *     results are bit-identical to that entry point's.
 */
#define SSI_B200_RMSNORM_BWD_STATIC_MAX_D 2048
int ssi_b200_rmsnorm_bwd_static(const void* dy, const void* x, const void* scale, const float* rstd, int64_t N,
                                int64_t D, void* dx, float* dscale, void* stream);

/*
 * ssi_b200_label_shift_count plus the token-type histogram of ssi/train_utils.py:150-165 in the same pass, all
 * results left on the device (the reference does one host sync per range, ssi/trainer.py:388-391):
 *   range_lo / range_hi  HOST arrays [n_ranges] of inclusive token-id ranges (n_ranges <= 15)
 *   type_counts [n_ranges + 1] int64 out: tokens inside each range, then "total" = #(tokens != pad_idx)
 */
int ssi_b200_label_shift_count_types(const int64_t* labels, const int64_t* tokens, int64_t B, int64_t S,
                                     int64_t ignore_index, const int64_t* range_lo, const int64_t* range_hi,
                                     int n_ranges, int64_t pad_idx, int64_t* labels_out, int64_t* counts,
                                     int64_t* type_counts, void* stream);

/*
 * scale_grads (tt/torchtune/training/_grad_scaler.py:25-31: p.grad *= scaler) for one bf16 gradient buffer and, in
 * the same pass, the sum of squares of the SCALED values that clip_grad_norm_ (ssi/trainer.py:405-408) needs of
 * it: grad = bf16(float(grad) * scale[0]) in place; sqnorm_out[0] = sum(float(grad_new)^2), fp32, deterministic.
 */
int ssi_b200_grad_scale_sqnorm_workspace_bytes(size_t* bytes);
int ssi_b200_grad_scale_sqnorm(void* grad, int64_t n_elems, const float* scale, float* sqnorm_out, void* ws,
                               size_t ws_bytes, void* stream);

/*
 * The same two passes FOLDED into the kernels that write the tied gradient anyway (replaces
 * tt/torchtune/training/_grad_scaler.py:25-31 + the tied weight's share of clip_grad_norm_, ssi/trainer.py:404-408,
 * on the last micro-batch of an accumulation window): every "_fold" entry point below is its plain namesake plus
 *   post_scale    fp32 device scalar or NULL: the scaler of scale_grads, applied in fp32 BEFORE the one rounding to
 *                 bf16 (the dense site also rescales what the gradient buffer already held when it accumulates);
 *   fold_ws       device scratch (ssi_b200_grad_fold_workspace_bytes, 16-byte aligned) or NULL: receives each
 *                 kernel's share of sum(g^2) -- block partials of the new values at the dense site, warp partials of
 *                 new^2 - old^2 at the sparse site.
 * Order per optimizer step: the dense site first (lmhead_ce_apply_grad_fold or dp_exchange_fold: it zeroes
 * fold_ws), then the sparse site (embedding_grad_scatter_sorted_fold or dp_embedding_exchange_fold), then
 * ssi_b200_grad_fold_sqnorm: sqnorm_out[0] = sum over the whole [V, D] gradient of g^2 (partials added in a fixed
 * order in double: deterministic).  No separate pass over the gradient remains.
 */
int ssi_b200_grad_fold_workspace_bytes(int64_t max_rows, int64_t D, size_t* bytes);
int ssi_b200_grad_fold_sqnorm(const void* fold_ws, size_t fold_ws_bytes, float* sqnorm_out, void* stream);
int ssi_b200_lmhead_ce_apply_grad_fold(void* dX, const float* dW_acc, const float* grad_out, const float* grad_hint,
                                       int64_t N, int64_t V, int64_t D, void* dW, int accumulate_dW,
                                       const float* post_scale, void* fold_ws, size_t fold_ws_bytes, void* stream);
int ssi_b200_embedding_grad_scatter_sorted_fold(const void* dH, int64_t N, int64_t V, int64_t D, void* dW, void* ws,
                                                size_t ws_bytes, const float* post_scale, void* fold_ws,
                                                size_t fold_ws_bytes, void* stream);
int ssi_b200_dp_exchange_fold(void* const* peer_bufs, int rank, int world, int64_t V, int64_t D, int64_t max_rows,
                              uint32_t epoch, const float* grad_out, const float* grad_hint, void* dW,
                              int accumulate_dW, void* host_flag, const float* post_scale, void* fold_ws,
                              size_t fold_ws_bytes, void* stream);
/* ssi_b200_dp_exchange_fold with the in-switch reduction (NVLS): multicast_buf = the MULTICAST address of the
 * exchange buffers (all ranks' buffers bound to one multicast object, e.g. torch.distributed._symmetric_memory's
 * multicast_ptr) or NULL (= plain P2P loads).  The reduce then issues one multimem.ld_reduce per 16 bytes instead of
 * `world` peer loads: 1/world of the NVLink bytes into the GPU. */
int ssi_b200_dp_exchange_nvls(void* const* peer_bufs, const void* multicast_buf, int rank, int world, int64_t V,
                              int64_t D, int64_t max_rows, uint32_t epoch, const float* grad_out,
                              const float* grad_hint, void* dW, int accumulate_dW, void* host_flag,
                              const float* post_scale, void* fold_ws, size_t fold_ws_bytes, void* stream);
int ssi_b200_dp_embedding_exchange_fold(const void* dH, int64_t N, int64_t V, int64_t D, void* sort_ws,
                                        size_t sort_ws_bytes, void* const* peer_bufs, int rank, int world,
                                        int64_t max_rows, uint32_t epoch, void* dW, void* host_flag,
                                        const float* post_scale, void* fold_ws, size_t fold_ws_bytes, void* stream);

/* Forward embedding gather (tt/torchtune/modules/transformer.py:628): out[i, :] = W[ids[i], :], bf16, exact copy;
 * ids outside [0, V) give zero rows. */
int ssi_b200_embedding_gather(const void* W, const int64_t* ids, int64_t N, int64_t V, int64_t D, void* out,
                              void* stream);

/*
 * Optional per-kernel-class timing (used by bench.py for the roofline figure): between profile_begin and
 * profile_end every launch made through this library is bracketed by CUDA events on its own stream.
 * profile_end waits for those events and fills ms_by_class[8] / launches_by_class[8] (either may be NULL):
 *   0 = K1 logits + online-LSE (fwd), 1 = K2a dlogits recompute, 2 = K2b dW contraction, 3 = K2c dX contraction,
 *   4 = K3 embedding scatter (sort + segmented reduce, one entry per call), 5 = small helper kernels,
 *   6 = single-pass work on dlogits outside the contractions (exp-stash form: the label-keyed onehot scatter into
 *       dW_acc; classic form: the S -> dlogits pass, HBM-bound), 7 = data-parallel exchange.
 * This pair is the only part of the ABI that synchronises (profile_end) or allocates (events, lazily).
 */
int ssi_b200_profile_begin(void);
int ssi_b200_profile_end(double* ms_by_class, int64_t* launches_by_class);

/*
 * Test hook: C[M, N] (fp32, row-major, ldc = N) = A * B^T-style contraction on the same tcgen05 mainloop
 * the product kernels use.  a_mn_major == 0: A is [M, K] row-major (K contiguous); != 0: A is [K, M]
 * row-major (M contiguous).  Likewise B: [N, K] or [K, N].  accumulate != 0 adds into C (TMA reduce-add).
 * Supported (a_mn_major, b_mn_major): (0,0), (1,1), (0,1).
 */
int ssi_b200_debug_gemm(const void* A, const void* B, float* C, int64_t M, int64_t N, int64_t K, int a_mn_major,
                        int b_mn_major, int accumulate, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SSI_B200_H_ */
