// One decoding position of the joint CTC/attention search behind ONE library call
// (include/b200ctc.h `b200ctc_beam_position`; reference: the part of BatchBeamSearch.search,
// /root/reference/espnet/nets/batch_beam_search.py:206-283, that follows the full scorers).
//
// Nothing new is computed here: the call issues the library's own pre-beam, prefix-scoring,
// pruning and state-selection entry points in the order the host-side mirror's generic path
// calls them, on buffers the caller allocated once per decode, and queues the copy of the
// step's record to pinned host memory behind them.  What it removes is the host work between
// those launches (argument marshalling and ~20 allocations per position in the host language),
// which is what bounded a position: ~0.2 ms of kernels inside 0.56 ms.
#include "common.cuh"

using namespace b200ctc;

extern "C" size_t b200ctc_beam_plan_bytes(void) { return sizeof(b200ctc_beam_plan); }

extern "C" int b200ctc_beam_position(const b200ctc_beam_plan* p, const float* const* full_ptrs_host,
                                     const int64_t* full_strides_host, int step_i, int cur_len,
                                     b200ctc_stream_t stream) {
  B2_REQUIRE(p != nullptr, "beam_position: null plan");
  const int U = p->U, W = p->W, V = p->V, T = p->T, Kf = p->K_full, K = p->K_full + 1;
  B2_REQUIRE(U > 0 && W > 0 && V > 0 && T > 0, "beam_position: bad shape U=%d W=%d V=%d T=%d", U, W, V, T);
  B2_REQUIRE(Kf >= 0 && K <= B200CTC_BEAM_MAX_SCORERS, "beam_position: K_full=%d (at most %d)", Kf,
             B200CTC_BEAM_MAX_SCORERS - 1);
  B2_REQUIRE(Kf == 0 || (full_ptrs_host && full_strides_host), "beam_position: null score-matrix arrays");
  B2_REQUIRE(p->snum >= 0 && p->snum < V, "beam_position: pre-beam width %d outside [0, V)", p->snum);
  B2_REQUIRE(p->snum == 0 || (Kf > 0 && p->lp_t && p->part_ids && p->idmap && p->prebeam_key < Kf),
             "beam_position: a pre-beam needs a full scorer, lp_t, part_ids and idmap");
  B2_REQUIRE(step_i >= 0 && cur_len >= 1 && cur_len < p->yseq_stride,
             "beam_position: cur_len=%d does not fit the prefix buffers (stride %lld)", cur_len,
             (long long)p->yseq_stride);
  B2_REQUIRE(p->lp && p->end_frames && p->r_state[0] && p->r_state[1] && p->s_state[0] && p->s_state[1] && p->r && p->log_psi && p->ctc_scores &&
                 p->rsum && p->run_score[0] && p->run_score[1] && p->cum[0] && p->cum[1] &&
                 p->yseq[0] && p->yseq[1] && p->out_val && p->out_flat && p->pack && p->pack_host,
             "beam_position: null buffer in the plan");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int n = U * W, in = step_i & 1, out = in ^ 1;
  const bool pre = p->snum > 0;
  int rc;

  // ---- pre-beam on the full scorers (batch_beam_search.py:228-234)
  if (pre) {
    if (p->prebeam_key < 0) {
      rc = b200ctc_beam_prebeam(full_ptrs_host, full_strides_host, p->weights, Kf, n, V, p->snum,
                                p->part_ids, stream);
    } else {
      const float one = 1.0f;
      rc = b200ctc_beam_prebeam(full_ptrs_host + p->prebeam_key, full_strides_host + 2 * p->prebeam_key,
                                &one, 1, n, V, p->snum, p->part_ids, stream);
    }
    if (rc != B200CTC_OK) return rc;
  }

  // ---- CTC prefix scores of the candidates (ctc_prefix_score.py:68-187; no attention window)
  if (step_i == 0) {
    rc = b200ctc_prefix_init_state(p->lp, U, T, V, W, p->blank, p->logzero, p->r_state[0], stream);
    if (rc != B200CTC_OK) return rc;
  }
  const int out_len = cur_len - 1;
  // `r` never leaves this call: where the scan kernel takes the step it is kept time-major
  // (n, snum, 2, T) - the kernel streams whole rows out and the selection below reads them back
  // coalesced - instead of the reference's (T, 2, n, snum) that the stand-alone entry points expose
  static const int tm_on = tuning_int("B200CTC_BEAM_TIME_MAJOR", 1);
  const int tm = (tm_on && pre && prefix_scan_applies(true, true, n, p->snum, T)) ? 1 : 0;
  // (the last token of every prefix is read in place: column cur_len - 1 of the prefix buffer)
  rc = prefix_step_impl(p->lp, pre ? p->lp_t : nullptr, p->r_state[in], step_i == 0 ? nullptr : p->s_state[in],
                           p->yseq[in] + (cur_len - 1), p->yseq_stride, pre ? p->part_ids : nullptr, p->end_frames, U, W, T, V,
                           pre ? p->snum : V, out_len, out_len > 1 ? out_len : 1, T, p->blank, p->eos,
                           p->logzero, p->r, p->log_psi, p->ctc_scores, pre ? p->idmap : nullptr, p->rsum,
                           (int)p->lp_t_ld, nullptr, tm, stream);
  if (rc != B200CTC_OK) return rc;

  // ---- pruning (batch_beam(), :98-118, and the gathers of :243-283)
  const float* ptrs[B200CTC_BEAM_MAX_SCORERS];
  int64_t strides[2 * B200CTC_BEAM_MAX_SCORERS];
  for (int q = 0; q < Kf; ++q) {
    ptrs[q] = full_ptrs_host[q];
    strides[2 * q] = full_strides_host[2 * q];
    strides[2 * q + 1] = full_strides_host[2 * q + 1];
  }
  ptrs[Kf] = p->ctc_scores;
  strides[2 * Kf] = V;
  strides[2 * Kf + 1] = 1;
  rc = b200ctc_beam_select(ptrs, strides, p->weights, K, p->run_score[in], p->cum[in], p->yseq[in],
                           p->yseq_stride, cur_len, p->active, p->maxlen, step_i, U, W, V, W, p->eos,
                           p->out_val, p->out_flat, p->run_score[out], p->cum[out], p->yseq[out],
                           p->yseq_stride, p->pack, stream);
  if (rc != B200CTC_OK) return rc;

  // ---- the surviving slots' CTC state (index_select_state, ctc_prefix_score.py:189-220)
  rc = prefix_select_impl(p->r, p->log_psi, pre ? p->idmap : nullptr, p->out_flat, U, W, W, T, V,
                          pre ? p->snum : V, p->r_state[out], p->s_state[out], tm, stream);
  if (rc != B200CTC_OK) return rc;

  // ---- the one record the host reads per position
  cudaError_t e = cudaMemcpyAsync(p->pack_host, p->pack, sizeof(float) * (size_t)(3 + K) * n,
                                  cudaMemcpyDeviceToHost, st);
  if (e != cudaSuccess) {
    set_error("beam_position: copy of the step record: %s", cudaGetErrorString(e));
    return B200CTC_ERR_DEVICE;
  }
  return check_launch("beam_position");
}
