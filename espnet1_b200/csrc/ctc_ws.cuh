// Workspace layout shared by the CTC forward (lse/gather + lattice recursion)
// and backward (fused gradient) kernels.
//
// All lattice quantities are in LOG2 units and NORMALISED so that no large
// magnitudes are ever added in fp32 (the reference's fp32 log-space lattice
// carries |alpha| ~ 4e3 at T=500,V=5000, i.e. ~3e-3 relative error in the
// occupancies; see DESIGN.md "numerics"):
//
//   lp2[t][v]   = logit*log2(e) - lse2[t]                  (log2-probability)
//   mt[t]       = max_v lp2[t][v] = -log2(sum_v 2^(y2-max)) (per-frame offset, <= 0)
//   em[t][k]    = lp2[t][label_k] - mt[t]   (k = 0: blank, k = 1..L: labels)  <= 0
//   alpha[t][s] = true log2-alpha - (sum_{tau<=t} mt[tau] + offA[t])
//   beta [t][s] = true log2-beta  - (sum_{tau>=t} mt[tau] + offB[t])
//
// offA/offB are the cumulative renormalisation shifts (fp64), so that
//   log2 gamma_t(s) = alpha + beta - em + (offA[t] + offB[t] - aend - lam)
// involves only O(1..100) magnitudes.
#pragma once
#include "common.cuh"

namespace b200ctc {

struct UttInfo {
  double lam;    // log2-sum of the two final normalised alphas
  double aend;   // offA at the last valid frame
  double msum;   // sum_t mt[t] over valid frames
  int T;         // clamped input length
  int L;         // clamped target length
  int zeroed;    // 1: loss was +inf and zero_infinity -> gradient is all zeros
  int infinite;  // 1: loss is +inf (kept when !zero_infinity -> NaN gradient)
  int blank_head;  // first label position (1-based k) whose label == blank, 0 = none
  int bad_label;   // 1: some label was outside [0,V) and got clamped
  int pad[2];
};

struct LossLayout {
  int B, T, V, Lmax, Lp, Sp;
  size_t lse2, mt, qt, em, alpha, beta, offA, offB, utt, lab, dup, total;
};

inline LossLayout make_loss_layout(int B, int T, int V, int Lmax) {
  LossLayout w;
  w.B = B; w.T = T; w.V = V; w.Lmax = Lmax;
  w.Lp = (int)align_up((size_t)Lmax + 1, 4);
  w.Sp = (int)align_up((size_t)2 * Lmax + 1, 4);
  size_t o = 0;
  const size_t BT = (size_t)B * T;
  auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes, 256); return r; };
  w.utt = take(sizeof(UttInfo) * (size_t)B);
  w.lab = take(4 * (size_t)B * w.Lp);
  w.dup = take(4 * (size_t)B * w.Lp);
  w.lse2 = take(4 * BT);
  w.mt = take(4 * BT);
  w.qt = take(4 * BT);
  w.offA = take(8 * BT);
  w.offB = take(8 * BT);
  w.em = take(4 * BT * w.Lp);
  w.alpha = take(4 * BT * w.Sp);
  w.beta = take(4 * BT * w.Sp);
  w.total = o;
  return w;
}

// dup[] packing: low 30 bits = (index of next label position with the same
// label) + 1, 0 = none; bit 30 = this position owns its vocabulary column.
constexpr int kDupFirst = 1 << 30;
constexpr int kDupMask = kDupFirst - 1;

}  // namespace b200ctc
