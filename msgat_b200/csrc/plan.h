// plan.h -- host-side derivation of buffer layouts and tile shapes from a msgat_meam_desc.
// Shared by the forward and the backward so that `saved` written by one is read by the other.
#pragma once
#include <stdint.h>
#include "../../include/msgat_sm100.h"

#define MSGAT_SMEM_BUDGET (200 * 1024)

struct Plan {
  int B, C, N, O, Oc, Ot, Os, flags, n_conv, dil[MSGAT_MAX_CONV], nnz, precision;
  int M;                            // components of a grouped call (>= 1)
  int64_t x_sb, x_sc, x_sm;
  int64_t scratch_sm;               // floats between the scratch buffers of consecutive components (filled by api.cu)
  int s_pre;  // spatial branch aggregates Z = W h (Os channels) instead of h (C channels)
  int Cz;     // channels of the Z buffer the spatial gather reads
  // forward tiles
  int NTp, ntile_p;  // prep kernel: nodes per tile, tiles per sample
  int NTa, ntile_a;  // apply kernel
  int NTb, ntile_b;  // backward main kernel
  // saved layout (float offsets)
  int64_t o_qs, o_qt, o_Z, o_m, o_l, o_E, o_U, o_Y1, o_A2, o_pw, o_qc, o_Pc, o_Mc, o_K, o_Pt, saved_total;
  // forward scratch
  int64_t o_pqc, o_pK, fwd_scratch;
};

static inline int rup4(int v) { return (v + 3) & ~3; }
static inline int64_t rup4l(int64_t v) { return (v + 3) & ~(int64_t)3; }

// shared memory (floats) of the apply kernel for NT nodes per tile
static inline int64_t apply_smem_floats(const Plan& p, int NT) {
  int PS = NT * 12 + 4;
  int64_t f = 0;
  f += (int64_t)p.C * PS;                              // X
  f += (int64_t)p.C * PS;                              // H
  if (p.Ot) f += (int64_t)p.C * PS;                    // OT
  if (p.n_conv >= 2) f += (int64_t)p.Ot * PS;          // YA
  if (p.n_conv >= 3) f += (int64_t)p.Ot * PS;          // YB
  if (p.Os && !p.s_pre) f += (int64_t)p.C * PS;        // OS (aggregated h)
  if (p.flags & MSGAT_F_RES) f += (int64_t)p.C * rup4(p.O) + rup4(p.O);
  if (p.Oc) f += (int64_t)p.C * rup4(p.Oc) + rup4(p.Oc);
  for (int i = 0; i < p.n_conv; ++i) f += 2 * (int64_t)(i == 0 ? p.C : p.Ot) * rup4(p.Ot) + rup4(p.Ot);
  if (p.Os && (p.flags & MSGAT_F_SMIX) && !p.s_pre) f += (int64_t)p.C * rup4(p.Os);
  f += 144 + 24;
  return f;
}

static inline int make_plan(const msgat_meam_desc* d, int nnz, Plan* p) {
  p->B = d->B; p->C = d->C; p->N = d->N; p->Oc = d->Oc; p->Ot = d->Ot; p->Os = d->Os;
  p->O = d->Oc + d->Ot + d->Os; p->flags = d->flags; p->n_conv = d->n_conv; p->nnz = nnz;
  p->precision = d->precision; p->x_sb = d->x_sb; p->x_sc = d->x_sc;
  p->M = d->M > 1 ? d->M : 1; p->x_sm = d->x_sm; p->scratch_sm = 0;
  if (p->M > MSGAT_MAX_GROUP) return -8;
  for (int i = 0; i < MSGAT_MAX_CONV; ++i) p->dil[i] = d->dil[i];
  if (d->B <= 0 || d->C <= 0 || d->N <= 0 || p->O <= 0) return -1;
  if (d->n_conv < 0 || d->n_conv > MSGAT_MAX_CONV) return -2;
  if (d->Oc && !(d->flags & MSGAT_F_CMIX) && d->Oc != d->C) return -3;
  if (d->Ot && d->n_conv == 0 && d->Ot != d->C) return -4;
  if (d->Os && !(d->flags & MSGAT_F_SMIX) && d->Os != d->C) return -5;
  if (d->C > 128 || p->O > 384) return -6;
  p->s_pre = (d->Os && (d->flags & MSGAT_F_SMIX) && d->C > d->Os) ? 1 : 0;
  p->Cz = d->Os ? (p->s_pre ? d->Os : d->C) : 0;
  // tiles
  p->NTp = d->C <= 4 ? 64 : 16;   // narrow (first-layer) inputs: 16-node tiles leave most of a block idle (C * 16 rows on 256 threads)
  p->ntile_p = (d->N + p->NTp - 1) / p->NTp;
  int NT = 16;
  while (NT > 2 && apply_smem_floats(*p, NT) * 4 > MSGAT_SMEM_BUDGET) NT >>= 1;
  if (apply_smem_floats(*p, NT) * 4 > MSGAT_SMEM_BUDGET) return -7;
  p->NTa = NT;
  p->ntile_a = (d->N + NT - 1) / NT;
  p->NTb = 0; p->ntile_b = 0;  // filled by the backward
  int64_t o = 0;
  const int64_t BN = (int64_t)d->B * d->N;
  p->o_qs = o; o += rup4l(BN * 12);
  p->o_qt = o; o += rup4l(BN * 12);
  p->o_Z = o; o += rup4l(BN * 12 * p->Cz);
  p->o_m = o; o += rup4l(BN);
  p->o_l = o; o += rup4l(2 * BN);        // fp64: the softmax denominator
  p->o_E = o; o += rup4l(2 * BN * 12);   // fp64: E = softmax(S) q
  p->o_U = o; o += d->Os ? rup4l(BN * 12) : 0;
  // output of the first conv of a two-conv chain, kept for the tensor-core weight-gradient kernel
  p->o_Y1 = o; o += (d->n_conv == 2) ? rup4l(BN * 12 * d->Ot) : 0;
  // first-conv taps a_now | a_past = W1 h before the temporal mix, kept for dPt in the tensor-core data-gradient kernel
  p->o_A2 = o; o += (d->n_conv == 2 && d->C >= 8) ? rup4l(BN * 12 * 2 * d->Ot) : 0;
  p->o_pw = o; o += rup4l((int64_t)d->B * nnz);
  p->o_qc = o; o += rup4l((int64_t)d->B * d->C * 12);
  p->o_Pc = o; o += rup4l((int64_t)d->B * d->C * d->C);
  p->o_Mc = o; o += rup4l((int64_t)d->B * d->Oc * d->C);
  p->o_K = o; o += rup4l((int64_t)d->B * 240);
  p->o_Pt = o; o += rup4l((int64_t)d->B * 144);
  o = (o + 3) & ~(int64_t)3;
  p->saved_total = o;
  int64_t s = 0;
  p->o_pqc = s; s += rup4l(2 * (int64_t)d->B * p->ntile_p * d->C * 12);  // fp64 partials
  p->o_pK = s; s += rup4l(2 * (int64_t)d->B * p->ntile_p * 240);
  p->fwd_scratch = s;
  return 0;
}
