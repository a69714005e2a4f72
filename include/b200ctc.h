/*
 * b200ctc.h - C ABI of the B200-native ESPnet CTC hot path (libb200ctc.so).
 *
 * The reference (DavidLBick/espnet-1) is pure Python and has NO native FFI for
 * this path: its boundary is the Python class API.  Each entry point below
 * states which reference call it replaces (file:line relative to the reference
 * tree); INTEGRATION.md shows the ctypes binding a maintainer would add.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller unless the name ends
 *    in `_host`; nothing is allocated or freed behind the caller's back except
 *    inside the `*_host` pipeline helpers (which own a cached staging arena).
 *  - every call is asynchronous on `stream` (a cudaStream_t passed as void*).
 *  - return value: 0 = OK, negative = error; b200ctc_last_error() gives a
 *    thread-local message.  There is NO CPU fallback: without a CUDA device
 *    (or with a non-sm_100 device) compute calls fail with B200CTC_ERR_DEVICE.
 *  - dtype codes: 0 = fp32, 1 = fp16, 2 = bf16 (arithmetic is always fp32).
 *  - thread-safe as long as distinct streams use distinct workspaces.
 */
#ifndef B200CTC_H_
#define B200CTC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200CTC_ABI_VERSION 6

#define B200CTC_OK 0
#define B200CTC_ERR_ARG (-1)     /* bad argument (null pointer, size, dtype) */
#define B200CTC_ERR_DEVICE (-2)  /* no usable CUDA device / launch failure   */
#define B200CTC_ERR_WORKSPACE (-3) /* workspace too small                    */

#define B200CTC_F32 0
#define B200CTC_F16 1
#define B200CTC_BF16 2

typedef void* b200ctc_stream_t; /* cudaStream_t */

/* ---- library -------------------------------------------------------------- */
int b200ctc_abi_version(void);
const char* b200ctc_last_error(void);
/* number of kernels launched by this library (process-wide) since the
 * last reset (bench.py's `gpu_launches`). */
int64_t b200ctc_launch_count(void);
void b200ctc_reset_launch_count(void);
/* optional per-phase timing with CUDA events recorded on the caller's stream
 * around the library's own launches (phase ids: 0 prep, 1 emission gather / row
 * pass A, 2 lattice, 3 grad_rows, 4 prefix_step, 5 softmax family, 6 rescale).  profile_read synchronises,
 * returns per-phase total milliseconds / launch counts and clears the log. */
void b200ctc_profile_enable(int on);
int b200ctc_profile_read(float* ms_total, int* counts, int n);
/* fills sm_count / compute capability of the current device */
int b200ctc_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- (a1) vocabulary soft-max family --------------------------------------
 * replaces Tensor.log_softmax / F.softmax / torch.argmax over the last dim:
 *   espnet2/asr/ctc.py:107,117,127 ; espnet/nets/pytorch_backend/ctc.py:142,152,161 ;
 *   espnet/nets/scorers/ctc.py:35,96,138.
 * `x` is (rows, V) row-major with row stride V, same dtype in and out.
 * mode: 0 = log_softmax, 1 = softmax.                                         */
int b200ctc_softmax_fwd(const void* x, void* y, int64_t rows, int V, int dtype,
                        int mode, b200ctc_stream_t stream);
/* The same over x + bias (bias: V values of x's dtype, added in fp32 before anything else):
 * `ctc_lo` is a Linear layer, and at the CTC projection's shape cuBLASLt adds its bias in a
 * separate pass over the (rows, V) output (0.38 ms of a 1.2 ms projection for 32 utterances);
 * the caller runs the GEMM without bias and the soft-max's one read of the row absorbs it
 * (espnet2/asr/ctc.py:107,117: F.softmax / F.log_softmax(self.ctc_lo(hs_pad))).  (ABI 6)      */
int b200ctc_softmax_fwd_bias(const void* x, const void* bias, void* y, int64_t rows,
                             int V, int dtype, int mode, b200ctc_stream_t stream);
/* autograd backward of the above: dx from y (the forward OUTPUT) and dy.      */
int b200ctc_softmax_bwd(const void* y, const void* dy, void* dx, int64_t rows,
                        int V, int dtype, int mode, b200ctc_stream_t stream);
/* first index of the row maximum (NaN counts as the maximum), int64 out.      */
int b200ctc_argmax(const void* x, int64_t* out, int64_t rows, int V, int dtype,
                   b200ctc_stream_t stream);
/* argmax of x + bias (bias as in b200ctc_softmax_fwd_bias: the projection's GEMM runs without
 * it; espnet2/asr/ctc.py:127 torch.argmax(self.ctc_lo(hs_pad), dim=2)).  (ABI 6)             */
int b200ctc_argmax_bias(const void* x, const void* bias, int64_t* out, int64_t rows,
                        int V, int dtype, b200ctc_stream_t stream);
/* best-path decode (espnet/nets/e2e_asr_common.py:159-191): argmax per frame,
 * collapse repeats, drop blanks.  x (B,T,V); out_ids (B,T) int64 padded with
 * -1; out_lens (B,) int64.  in_lens may be NULL (all T).                      */
int b200ctc_greedy_decode(const void* x, const int64_t* in_lens, int B, int T,
                          int V, int dtype, int blank, int64_t* out_ids,
                          int64_t* out_lens, b200ctc_stream_t stream);

/* ---- (a2,a3) CTC loss and its gradient w.r.t. the logits ------------------
 * replaces  th_pred.log_softmax(2) + torch.nn.CTCLoss(reduction="none",
 * zero_infinity=..) and the autograd backward of both:
 *   espnet2/asr/ctc.py:39-41,52-62 ; espnet/nets/pytorch_backend/ctc.py:38-40,56-65.
 * logits: (B,T,V) batch-major pre-softmax activations (what ctc_lo emits,
 *   espnet2/asr/ctc.py:82), row stride V.  targets: 1-D concatenated labels
 *   (espnet2/asr/ctc.py:91) with tgt_offsets (B+1,) prefix sums.
 * Two ways to get the gradient:
 *  - eager (training default): b200ctc_loss_fwd_grad computes nll AND
 *      grad[b,t,v] = grad_hint * d nll[b] / d logits[b,t,v]   (zeros for t >= in_lens[b])
 *    in one call.  The lattice runs on a sparse gather of the label columns, so
 *    the (B,T,V) logits are read in full exactly once and the gradient is
 *    written exactly once.  grad_hint is the upstream gradient the caller
 *    expects (1/B for loss.sum()/B); b200ctc_grad_rescale later multiplies by
 *    gscale[b*stride]/grad_hint and is a device-side no-op (every CTA reads one
 *    scalar and exits) when the real upstream gradient equals the hint.
 *  - deferred: b200ctc_loss_fwd(need_grad=1) keeps the lattices in `ws`
 *    (one full read of the logits for the row log-sum-exp), b200ctc_loss_bwd
 *    consumes `ws` and writes
 *      grad[b,t,v] = gscale_mul * gscale[b*gscale_stride] * d nll[b] / d logits[b,t,v];
 *    gscale may be NULL (=1), gscale_stride is 0 for a scalar upstream
 *    gradient, 1 for a per-utterance one.
 * fwd writes nll (B,) fp32 (0 where infeasible and zero_infinity).  tgt_offsets
 * may be NULL (the library prefix-sums tgt_lens itself).  loss_out may be NULL;
 * otherwise loss_out[0] = loss_scale * sum_b nll[b] is written by the same
 * launches (espnet2/asr/ctc.py:56-60 `loss.sum() / size` with loss_scale = 1/B). */
size_t b200ctc_loss_workspace_bytes(int B, int T, int V, int max_tgt_len);
int b200ctc_loss_fwd(const void* logits, int dtype, const int64_t* targets,
                     const int64_t* tgt_offsets, const int64_t* in_lens,
                     const int64_t* tgt_lens, int B, int T, int V,
                     int max_tgt_len, int blank, int zero_infinity, int need_grad,
                     float* nll, float* loss_out, float loss_scale, void* ws,
                     size_t ws_bytes, b200ctc_stream_t stream);
int b200ctc_loss_bwd(const void* logits, int dtype, const int64_t* in_lens,
                     int B, int T, int V, int max_tgt_len, int blank,
                     int zero_infinity, const float* gscale, int gscale_stride,
                     float gscale_mul, void* grad, void* ws, size_t ws_bytes,
                     b200ctc_stream_t stream);
int b200ctc_loss_fwd_grad(const void* logits, int dtype, const int64_t* targets,
                          const int64_t* tgt_offsets, const int64_t* in_lens,
                          const int64_t* tgt_lens, int B, int T, int V,
                          int max_tgt_len, int blank, int zero_infinity,
                          float grad_hint, float* nll, float* loss_out,
                          float loss_scale, void* grad, void* ws, size_t ws_bytes,
                          b200ctc_stream_t stream);
/* ---- (e) batch-sharded training: the loss all-reduce fused into the loss kernel
 * replaces the per-step all-reduce of {loss*weight, weight} over the data-
 * parallel ranks, espnet2/train/trainer.py:604-613 with
 * espnet2/torch_utils/recursive_op.py:8-47 (one process per GPU, one node).
 * A b200ctc_peer_t owns a small exchange buffer on the current device;
 * peer_create returns its 64-byte cudaIpc handle, the caller all-gathers the
 * handles of all ranks (any transport: torch.distributed, MPI, a file) and
 * peer_connect maps them.  b200ctc_loss_fwd_grad_sharded is loss_fwd_grad whose
 * final kernel ALSO pushes this rank's {sum_b nll_b, B} to every peer through
 * NVLink peer stores, waits for theirs and writes
 *   global_out[0] = sum_r sum_b nll / sum_r B_r ,  global_out[1] = sum_r B_r
 * (bit-identical on all ranks: summed in rank order).  Collective semantics:
 * every rank must make the same sequence of exchange calls; a rank with no
 * utterances calls b200ctc_peer_exchange({0,0}).  A peer that does not show up
 * within the time-out (default 30 s) yields NaN and sets the sticky flag read by
 * b200ctc_peer_error.
 * b200ctc_peer_configure(deferred = 1): loss_fwd_grad_sharded only PUSHES; the
 * wait for the peers' words - a barrier on the slowest rank that nothing on the
 * device needs inside the step - is done by the NEXT exchange call (for the
 * previous step: its words arrived a step ago) and by b200ctc_peer_wait, which
 * the caller issues on the stream where it reads global_out.  Ranks still stay
 * within one step of each other.  timeout_ms = 0 keeps the default.           */
typedef struct b200ctc_peer b200ctc_peer_t;
int b200ctc_peer_create(int world, int rank, b200ctc_peer_t** out,
                        unsigned char* handle_out /* 64 bytes */);
int b200ctc_peer_connect(b200ctc_peer_t* peer,
                         const unsigned char* all_handles /* world*64 bytes, rank order */);
int b200ctc_peer_destroy(b200ctc_peer_t* peer);
int b200ctc_peer_error(b200ctc_peer_t* peer); /* synchronises; 1 = a peer timed out */
int b200ctc_peer_configure(b200ctc_peer_t* peer, int deferred, int timeout_ms);
/* consume the latest pushed step if still owed: global_out = its result        */
int b200ctc_peer_wait(b200ctc_peer_t* peer, float* global_out, b200ctc_stream_t stream);
/* stand-alone exchange of a device pair {sum, weight} (same collective)       */
int b200ctc_peer_exchange(b200ctc_peer_t* peer, const float* local_pair,
                          float* global_out, b200ctc_stream_t stream);
int b200ctc_loss_fwd_grad_sharded(const void* logits, int dtype, const int64_t* targets,
                                  const int64_t* tgt_offsets, const int64_t* in_lens,
                                  const int64_t* tgt_lens, int B, int T, int V,
                                  int max_tgt_len, int blank, int zero_infinity,
                                  float grad_hint, float* nll, float* loss_out,
                                  float loss_scale, void* grad, void* ws,
                                  size_t ws_bytes, b200ctc_peer_t* peer,
                                  float* global_out, b200ctc_stream_t stream);
/* grad *= gscale[b*gscale_stride] * gscale_mul / grad_hint (no-op when equal)  */
int b200ctc_grad_rescale(void* grad, int dtype, const float* gscale,
                         int gscale_stride, float gscale_mul, float grad_hint,
                         int B, int T, int V, b200ctc_stream_t stream);
/* loss_fwd_grad / grad_rescale with ONE EXPECTED UPSTREAM GRADIENT PER UTTERANCE
 * (device vector utt_hints (B,), all non-zero) instead of a scalar: several CTC
 * heads that share labels but carry different loss weights - intermediate CTC,
 * espnet2/asr/espnet_model.py:243-286 - run as one batched call over
 * heads x utterances and backward stays a no-op check per utterance.  nll (B,)
 * only (no batch reduction: the caller sums per head).                         */
int b200ctc_loss_fwd_grad_hints(const void* logits, int dtype, const int64_t* targets,
                                const int64_t* tgt_offsets, const int64_t* in_lens,
                                const int64_t* tgt_lens, int B, int T, int V,
                                int max_tgt_len, int blank, int zero_infinity,
                                const float* utt_hints, float* nll, void* grad,
                                void* ws, size_t ws_bytes, b200ctc_stream_t stream);
int b200ctc_grad_rescale_hints(void* grad, int dtype, const float* gscale,
                               int gscale_stride, float gscale_mul,
                               const float* utt_hints, int B, int T, int V,
                               b200ctc_stream_t stream);
/* forced alignment (espnet/nets/pytorch_backend/ctc.py:163-226): max-semiring
 * lattice with back-pointers for ONE utterance; lp (T,V) fp32 log-probs,
 * out_ids (T,) int64 = label emitted at each frame (tgt_len >= 1).  `ws`: scratch of
 * b200ctc_align_workspace_bytes(T, tgt_len) bytes, 16-byte aligned (gathered emission columns
 * of the utterance, or packed back-pointers of inputs too long for shared memory); at most
 * ~50 k frames.                                                                              */
size_t b200ctc_align_workspace_bytes(int T, int tgt_len);
int b200ctc_forced_align(const float* lp, const int64_t* target, int T, int V,
                         int tgt_len, int blank, int64_t* out_ids, void* ws,
                         size_t ws_bytes, b200ctc_stream_t stream);
/* Allocation-free variant of b200ctc_forced_align.  NO counterpart in the
 * reference: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and result; the kernels' scratch (gathered
 * emissions or packed back-pointers) lives in a file-scope `static __device__` array of
 * B200CTC_ALIGN_STATIC_MAX_BYTES that every GPU loading the library holds until
 * exit.  Neither this call nor anything it calls allocates or frees memory.
 * T * (2*tgt_len+1) * 4 bytes beyond the maximum -> B200CTC_ERR_WORKSPACE with
 * a message naming the limit.  All calls on a device share the array (it keeps
 * the previous call's contents; every word that is read is written first by
 * the same call, so nothing needs clearing): calls that OVERLAP on one GPU
 * race - serialise them on one stream or with events.                        */
#define B200CTC_ALIGN_STATIC_MAX_BYTES (4u << 20) /* e.g. T=2000, tgt_len<=261 */
int b200ctc_forced_align_static(const float* lp, const int64_t* target, int T,
                                int V, int tgt_len, int blank, int64_t* out_ids,
                                b200ctc_stream_t stream);

/* ---- (a4-a7) batched CTC prefix scoring -----------------------------------
 * replaces CTCPrefixScoreTH (espnet/nets/ctc_prefix_score.py:11-269).        */
/* constructor padding (:47-50): frames t >= xlens[b] become logzero with
 * blank = 0, IN PLACE on the caller's (B,T,V) fp32 log-prob tensor.           */
int b200ctc_prefix_pad(float* lp, const int64_t* xlens, int B, int T, int V,
                       int blank, float logzero, b200ctc_stream_t stream);
/* initial state (:83-91): r (T,2,B*n_hyps) with r[:,0]=logzero and
 * r[:,1] = running sum over t of lp[b,t,blank].                               */
int b200ctc_prefix_init_state(const float* lp, int B, int T, int V, int n_hyps,
                              int blank, float logzero, float* r0,
                              b200ctc_stream_t stream);
/* attention window (:144-147): f_arg[h] = sum_t att_w[h,t]*t ; writes
 * minmax[0]=min_h f_arg, minmax[1]=max_h f_arg (fp32).                        */
int b200ctc_prefix_window(const float* att_w, int n_bh, int T, float* minmax,
                          b200ctc_stream_t stream);
/* time-contiguous copy of the padded posteriors: lp (B,T,V) -> lp_t (B,V,lp_t_ld),
 * frames [t0,T) only.  Made once per utterance batch (after prefix_pad, t0 = 0);
 * lets a pre-beam scoring step read one cache line per 32 frames of a candidate
 * instead of one each.  lp_t_ld >= T is the time stride of the copy: a streaming
 * caller (extend_prob, :222-242) allocates it with headroom and appends only
 * the new frames [T_old, T_new).                            (ABI 5: t0, lp_t_ld) */
int b200ctc_prefix_transpose(const float* lp, float* lp_t, int B, int T, int V,
                             int t0, int lp_t_ld, b200ctc_stream_t stream);
/* one scoring step (__call__, :68-187) over frames [start,end).
 *  lp (B,T,V) ; lp_t (B,V,T) from prefix_transpose or NULL ;
 *  r_prev (T,2,n_bh) ; s_prev (n_bh,V) or NULL (=0) ; last_ids (n_bh) ;
 *  scoring_ids (n_bh,snum) or NULL (then snum must equal V) ;
 *  outputs: r (T,2,n_bh,snum), log_psi (n_bh,V), scores (n_bh,V),
 *  idmap (n_bh,V) int64 (-1 = not scored; NULL iff scoring_ids is NULL).      */
int b200ctc_prefix_step(const float* lp, const float* lp_t, const float* r_prev,
                        const float* s_prev,
                        const int64_t* last_ids, const int64_t* scoring_ids,
                        const int64_t* end_frames, int B, int n_hyps, int T, int V,
                        int snum, int out_len, int start, int end, int blank,
                        int eos, float logzero, float* r, float* log_psi,
                        float* scores, int64_t* idmap, float* rsum_ws /*(T,n_bh) scratch*/,
                        int lp_t_ld /* time stride of lp_t (>= T; ABI 5) */,
                        const int* win_utt /* (B,2) {start,end} per utterance from
                                              prefix_window_utt, or NULL: [start,end) for all */,
                        b200ctc_stream_t stream);
/* Per-utterance attention windows - an opt-in EXTENSION: the reference
 * takes one window over the whole batch (:144-153); here the same formula runs over
 * the n_hyps hypotheses of each utterance alone, i.e. what the reference computes
 * when that utterance is decoded by itself.  Device-only (no read-back):
 * fprev (B,2) int32 {f_min,f_max} of the previous step or NULL ({0,1}, :95);
 * writes fnew (B,2) and win (B,2) {start,end} for b200ctc_prefix_step.          */
int b200ctc_prefix_window_utt(const float* att_w, int B, int n_hyps, int T, int margin,
                              int out_len, const int* fprev, int* fnew, int* win,
                              b200ctc_stream_t stream);
/* index_select_state (:189-220): best_ids (B,W) flat (hyp*V + label) ids.     */
int b200ctc_prefix_select(const float* r, const float* log_psi,
                          const int64_t* idmap, const int64_t* best_ids, int B,
                          int n_hyps, int W, int T, int V, int snum, float* r_new,
                          float* s_new, b200ctc_stream_t stream);

/* ---- (f1) beam pruning of the label-synchronous search --------------------
 * replaces the per-step glue of BatchBeamSearch.search
 * (espnet/nets/batch_beam_search.py:206-283) after the scorers have run.
 * Score matrices: K <= B200CTC_BEAM_MAX_SCORERS fp32 matrices of (n, V), given
 * as a HOST array of device pointers `score_ptrs[K]`, HOST `strides[2K]`
 * (row, column stride in elements; 0 = broadcast) and HOST `weights[K]`.
 *   w[h,v] = ((0 + weights[0]*s_0[h,v]) + weights[1]*s_1[h,v]) + ...   (:217-241,
 *   the reference's fp32 operation order; separate multiply and add)          */
#define B200CTC_BEAM_MAX_SCORERS 8
/* pre-beam (:228-234): out_ids (n_rows, k) int64 = top-k tokens of every row of w,
 * sorted by descending score (ties: lower token id first); k <= 64, V <= 55296. */
int b200ctc_beam_prebeam(const float* const* score_ptrs_host,
                         const int64_t* strides_host, const float* weights_host,
                         int K, int64_t n_rows, int V, int k, int64_t* out_ids,
                         b200ctc_stream_t stream);
/* one pruning step for U utterances x W beam slots (n = U*W rows), one CTA per
 * utterance: top-k over the utterance's W*V entries of w + run_score[h]
 * (batch_beam(), :98-118; k <= 64, W*V <= 55296), then for winner j of utterance u
 * (prev = flat / V, tok = flat % V, row = u*W + prev, o = u*k + j):
 *   out_val[o]  = w + run                  (the new hypothesis score, :258)
 *   out_flat[o] = prev*V + tok             (what index_select_state expects)
 *   out_run[o]  = out_val, or -inf if tok == eos, the utterance is inactive
 *                 (active[u] == 0) or this is its last position
 *                 (maxlen[u]-1 == step_i): the slot takes no further part (:318-347)
 *   cum_out[q][o] = cum_in[q][row] + s_q[row, tok]   (merge_scores, :260-266)
 *   yseq_out[o, :cur_len+1] = yseq_in[row, :cur_len] ++ tok   (append_token, :259)
 *   pack (3+K, U, k) fp32 = {out_val, prev, tok, cum_out[0..K)}: the one record the
 *                 host reads per step (ended hypotheses, end detection).
 * run_score, active, maxlen may be NULL.                                      */
int b200ctc_beam_select(const float* const* score_ptrs_host,
                        const int64_t* strides_host, const float* weights_host, int K,
                        const float* run_score, const float* cum_in,
                        const int64_t* yseq_in, int64_t yseq_in_stride, int cur_len,
                        const unsigned char* active, const int* maxlen, int step_i,
                        int U, int W, int V, int k, int eos, float* out_val,
                        int64_t* out_flat, float* out_run, float* cum_out,
                        int64_t* yseq_out, int64_t yseq_out_stride, float* pack,
                        b200ctc_stream_t stream);

/* One whole decoding position of the joint CTC/attention search (ABI 6) - everything of
 * BatchBeamSearch.search (espnet/nets/batch_beam_search.py:206-283) that follows the full
 * scorers, for U utterances x W beam slots (n = U*W rows), issued by ONE call on buffers the
 * caller allocated once per decode:
 *   beam_prebeam (if snum > 0)  ->  prefix_init_state (step_i == 0 only)  ->  prefix_step
 *   ->  beam_select (k = W)  ->  prefix_select  ->  pack copied to pinned host memory.
 * Same kernels, arguments and arithmetic as those entry points called one after the other (the
 * host-side mirror's generic path does exactly that; tests/test_beam_engine.py compares the
 * two), without the per-call argument marshalling and allocations of the host language: the
 * position is otherwise bound by the host, not by its kernels.  The CTC scorer is the LAST of
 * the K = K_full + 1 score matrices (the reference's accumulation order: full scorers, then
 * partial ones, :217-241).  The caller synchronises the stream before it reads pack_host. */
typedef struct b200ctc_beam_plan {
  int32_t U, W, V, T;
  int32_t snum;        /* pre-beam width (:228-234); 0 = the CTC scorer sees the full vocabulary */
  int32_t K_full;      /* number of full-scorer matrices handed in per position (<= 7) */
  int32_t prebeam_key; /* -1: pre-beam on the weighted sum of all full scorers ("full"),
                          else the index of the one full scorer it is taken from (weight 1) */
  int32_t blank, eos;
  float logzero;
  float weights[B200CTC_BEAM_MAX_SCORERS]; /* K_full full-scorer weights, then the CTC weight */
  /* CTC prefix scorer, device memory (shapes as in prefix_step / prefix_select) */
  const float* lp;           /* (U,T,V) padded log-probs */
  const float* lp_t;         /* (U,V,lp_t_ld) from prefix_transpose; required iff snum > 0 */
  int64_t lp_t_ld;
  const int64_t* end_frames; /* (U) */
  float* r_state[2];         /* (T,2,n)  state carried between positions, like the beam state */
  float* s_state[2];         /* (n,V)    below: [step_i & 1] read by prefix_step (position 0:
                                written by init_state first), [(step_i + 1) & 1] written by
                                prefix_select - so that position i + 1 can be in flight while
                                the host still copies position i's state */
  float* r;                  /* T*2*n*(snum or V) floats, private to the call (its layout may
                                differ from prefix_step's (T,2,n,snum)) */
  float* log_psi;            /* (n,V) */
  float* ctc_scores;         /* (n,V) */
  int64_t* idmap;            /* (n,V); unused when snum == 0 */
  float* rsum;               /* (T,n) scratch */
  int64_t* part_ids;         /* (n,snum) scratch: pre-beam candidates */
  /* beam state, device memory; [step_i & 1] is read, [(step_i + 1) & 1] is written */
  float* run_score[2];       /* (n) */
  float* cum[2];             /* (K,n) */
  int64_t* yseq[2];          /* (n, yseq_stride) prefixes */
  int64_t yseq_stride;
  const unsigned char* active; /* (U) */
  const int* maxlen;           /* (U) */
  float* out_val;            /* (n) */
  int64_t* out_flat;         /* (n) */
  float* pack;               /* (3+K,U,W) device */
  float* pack_host;          /* (3+K,U,W) PINNED host memory */
} b200ctc_beam_plan;
size_t b200ctc_beam_plan_bytes(void); /* sizeof(b200ctc_beam_plan): layout check for FFI mirrors */
/* full_ptrs_host[K_full] device pointers, full_strides_host[2*K_full] (row, column stride in
 * elements; 0 = broadcast) of this position's full-scorer matrices; cur_len = current prefix
 * length (primer included), i.e. prefix_step's out_len + 1.                                  */
int b200ctc_beam_position(const b200ctc_beam_plan* plan, const float* const* full_ptrs_host,
                          const int64_t* full_strides_host, int step_i, int cur_len,
                          b200ctc_stream_t stream);

/* ---- host-buffer pipeline (bench.py `e2e`) --------------------------------
 * Same arithmetic as loss_fwd + loss_bwd, but logits/targets/lengths are HOST
 * buffers (pinned for full speed) and nll (and optionally grad) are returned
 * to HOST buffers: utterance chunks are copied H2D on a copy stream while the
 * previous chunk computes.  Blocks until the results are in host memory.
 * grad_host may be NULL (loss only; the gradient stays in the device arena
 * returned through *grad_dev_out when that pointer is non-NULL).              */
int b200ctc_loss_fwd_bwd_host(const void* logits_host, int dtype,
                              const int64_t* targets_host,
                              const int64_t* in_lens_host,
                              const int64_t* tgt_lens_host, int B, int T, int V,
                              int blank, int zero_infinity, float gscale,
                              float* nll_host, void* grad_host,
                              void** grad_dev_out, int chunk_utts);
void b200ctc_host_pipeline_release(void);

#ifdef __cplusplus
}
#endif
#endif /* B200CTC_H_ */
