// Workspace layout shared by the CTC loss kernels (emission gather / row pass A,
// lattice recursion, gradient row pass B, nll finalize).
//
// All lattice quantities are in LOG2 units and NORMALISED so that no large
// magnitudes are ever added in fp32 (the reference's fp32 log-space lattice
// carries |alpha| ~ 4e3 at T=500,V=5000, i.e. ~3e-3 relative error in the
// occupancies; see DESIGN.md "numerics").  The recursion never needs the
// soft-max normaliser: any per-frame constant cancels in the occupancies
//   gamma_t(s) = alpha_t(s) beta_t(s) / (P * y_t(l_s)),
// so the lattice runs on RAW logits gathered at the label columns only,
//   y2[t][k]    = logit[t][label_k] * log2(e)           (k = 0: blank, 1..L: labels)
//   c[t]        = a per-frame constant (max_k y2[t][k] from the gather kernel,
//                 or the row maximum from row pass A)
//   em[t][k]    = y2[t][k] - c[t]
//   alpha[t][sp] = true log2-alpha~ - (sum_{tau<=t} c[tau] + FA_w[t])
//   beta [t][sp] = true log2-beta~  - (sum_{tau>=t} c[tau] + FB_w[t])
// where alpha~/beta~ are the recursions on the unnormalised scores 2^y2 and
// FA_w/FB_w are INTEGER re-centring offsets kept per (block of kRenormEvery
// recursion steps, warp of the lattice CTA; the recursion step of frame t is t
// for alpha and T_b-1-t for beta): every warp re-centres its own slice on its own maximum, so no CTA-wide
// reduction sits on the serial chain, and integer offsets combine exactly:
//   log2 gamma_t(s) = alpha + beta - em + (FA_wA[t] + FB_wB[t] - aend) - lam
//   log2 P          = lam + aend + sum_t (c[t] - lse2[t])
// with lse2[t] the row's log2-sum-exp (row pass A, or row pass B in the eager
// gradient mode).  beta is stored in recursion order: beta[t][sp], sp = S-1-s.
#pragma once
#include <stddef.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200ctc {

struct UttInfo {
  double lam;      // log2-sum of the two final normalised alphas
  int T;           // clamped input length          (T, L are read as one int2)
  int L;           // clamped target length
  int aend;        // FA of the warp that owns position S-1 at the last valid frame
  int zeroed;      // 1: loss was +inf and zero_infinity -> gradient is all zeros
  int infinite;    // 1: loss is +inf (kept when !zero_infinity -> NaN gradient)
  int blank_head;  // first label position (1-based k) whose label == blank, 0 = none
  int bad_label;   // 1: some label was outside [0,V) and got clamped
  int pad;
};

static_assert(offsetof(UttInfo, T) % 8 == 0 && offsetof(UttInfo, L) == offsetof(UttInfo, T) + 4,
              "UttInfo::{T,L} must form an aligned int2");

constexpr int kRenormEvery = 16;  // lattice re-centring period (steps); one frame offset per block
constexpr int kEmLabel0 = 4;  // em row: [blank, -, -, -, label_1 .. label_L]

struct LossLayout {
  int B, T, V, Lmax, Lp, Sp, R, NW, NB;
  int NG;  // frame groups of kRenormEvery frames per utterance (fused forward: gather -> lattice flags)
  // wide lattices are split over CL CTAs of NWc warps each (ctc_lattice.cuh "SPLIT"); the
  // CTAs of one (utterance, direction) hand their boundary values on through global memory
  int CL, NWc, Tb1;  // Tb1 = padded T+1: boundary values of one CTA boundary
  size_t lse2, cmax, em, alpha, beta, offA, offB, utt, lab, dup, counter, total;
  // sync words: {role ticket, sticky error} then the per-(utterance, group) gather flags -
  // adjacent, so that one memset clears them (ctc_fused.cuh)
  size_t ticket, err, flags;
  size_t split_tickets;  // two ints: dispatch-order tickets of the split lattice, per direction
  size_t bnd_val, bnd_f, bnd_ready;  // [B][2][CL-1] x {Tb1 floats, NB ints, 8 ints}
};
// sticky error word of a loss workspace (LossLayout::err; 0 = none).  A kernel that gives up
// waiting for another CTA sets it; nll_kernel then returns NaN losses (never silent garbage)
// and b200ctc_loss_status() reports it.
constexpr int kWsErrHandoff = 1;  // split lattice: an upstream CTA's boundary values never arrived
constexpr int kWsErrGather = 2;   // fused forward: a frame group's emissions never arrived
constexpr int kSplitWarps = 5;   // compute warps per CTA of a split lattice (R = 4)
constexpr int kBndReadyStride = 8;  // ints per ready counter (32 B apart)

// positions per thread of the lattice CTA (even); one warp covers 32*R positions.
// Warp w+1 of the systolic lattice trails warp w by one block of K >= 4 steps, and
// all warps read one shared ring of emission chunks (K frames of Lp floats): the
// ring must span NW+1 blocks.  R is the smallest value (<= 16 warps) that fits.
constexpr size_t kLatticeSmemBudget = 222 * 1024;  // of the 227 KB an SM can hand to one CTA
inline int lattice_R(int Smax, int Lp) {
  if (const char* e = tuning_env("B200CTC_LAT_R")) {  // tuning builds only
    const int r = atoi(e);
    if ((r == 2 || r == 4 || r == 8 || r == 16 || r == 32) && Smax <= 16 * 32 * r) return r;
  }
  for (int r = 2; r <= 16; r *= 2) {
    const int nw = (Smax + 32 * r - 1) / (32 * r);
    if (nw > 16) continue;
    const size_t fixed = (size_t)nw * 300 + 256;
    if ((size_t)(nw + 1) * 4 * Lp * 4 + fixed <= kLatticeSmemBudget) return r;
  }
  return 32;
}

inline LossLayout make_loss_layout(int B, int T, int V, int Lmax) {
  LossLayout w;
  w.B = B; w.T = T; w.V = V; w.Lmax = Lmax;
  const int S = 2 * Lmax + 1;
  w.Lp = kEmLabel0 + (int)align_up((size_t)(Lmax > 0 ? Lmax : 1), 4);
  w.Sp = (int)align_up((size_t)S, 32);  // rows of alpha / beta own their 128-byte lines (L2 discard)
  w.R = lattice_R(S, w.Lp);
  w.NW = (S + 32 * w.R - 1) / (32 * w.R);
  w.CL = 1;
  w.NWc = w.NW;
  // Long-form lattices (more than 16 warps of R <= 4, or an emission ring that does not fit):
  // one SM is then MUFU-bound (~3 MUFU per position and step) and stuck with 4-step blocks.
  // Split the positions over CL CTAs instead: R = 4, kSplitWarps warps per CTA, each CTA
  // stages only ITS label columns of the emission rows (16-step blocks fit again).
  static const int split = tuning_int("B200CTC_LAT_SPLIT", 1);
  if (split && (w.R > 4 || (w.R == 4 && w.NW > 8))) {  // (R = 4 with > 8 warps: 4/8-step blocks)
    const int nw4 = (S + 32 * 4 - 1) / (32 * 4);
    const int cl = (nw4 + kSplitWarps - 1) / kSplitWarps;
    // one CTA per SM (shared memory): only worth it while all CTAs of both directions are
    // resident together; a second wave would start when the first has finished
    if ((long long)B * 2 * cl <= sm_count()) {
      w.R = 4;
      w.NW = nw4;
      w.NWc = kSplitWarps;
      w.CL = cl;
    }
  }
  w.Tb1 = (int)align_up((size_t)T + 1 + 16, 4);  // + one block: a partial block's loads overshoot
  w.NB = (T + kRenormEvery - 1) / kRenormEvery;  // renorm blocks per utterance
  size_t o = 0;
  const size_t BT = (size_t)B * T;
  auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes, 256); return r; };
  w.counter = take(sizeof(int));  // "last CTA reduces" ticket of nll_kernel (zeroed by prep)
  w.NG = (T + kRenormEvery - 1) / kRenormEvery;
  if (w.NG < 1) w.NG = 1;
  w.ticket = take(256);
  w.err = w.ticket + 4;
  w.split_tickets = w.ticket + 8;
  w.flags = take(4 * (size_t)B * w.NG);
  w.utt = take(sizeof(UttInfo) * (size_t)B);
  w.lab = take(4 * (size_t)B * w.Lp);
  w.dup = take(4 * (size_t)B * w.Lp);
  w.lse2 = take(4 * BT);
  w.cmax = take(4 * BT);
  w.offA = take(4 * (size_t)B * w.NB * w.NW);
  w.offB = take(4 * (size_t)B * w.NB * w.NW);
  w.em = take(4 * BT * w.Lp);
  w.alpha = take(4 * BT * w.Sp);
  w.beta = take(4 * BT * w.Sp);
  const size_t nb = (size_t)B * 2 * (size_t)(w.CL - 1);
  w.bnd_val = take(4 * nb * w.Tb1);
  w.bnd_f = take(4 * nb * w.NB);
  w.bnd_ready = take(4 * nb * kBndReadyStride);
  w.total = o;
  return w;
}

// dup[] packing: low 30 bits = (index of next label position with the same
// label) + 1, 0 = none; bit 30 = this position owns its vocabulary column.
constexpr int kDupFirst = 1 << 30;
constexpr int kDupMask = kDupFirst - 1;

}  // namespace b200ctc
