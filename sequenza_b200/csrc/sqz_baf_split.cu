// sqz_baf_split.cu -- split form of the baf.model.fit fast path (see sqz_baf_grid.cu for
// the mathematics and for the fused fallback).
//
// The factor  wub[c] = prior_c^(-1/3) * min_B (1 + xb^2/5)  of a segment depends on the
// cellularity and the copy number, not on the ploidy.  A grid column (one cellularity)
// has nP/PT ploidy tiles, so the fused kernel recomputes it nP/PT times.  Here:
//
//   baf_colterm_kernel   one thread per (cellularity, segment): the min over B for every
//                        copy number, stored as sqrt(wub) (8 bytes; wub is its square, one
//                        DMUL per copy number where it is needed) in a column-major scratch
//                        [column][CNt][segment] (coalesced over segments);
//   baf_grid_split_kernel one CTA per (sample, cellularity, ploidy tile, segment chunk):
//                        the scratch rows of the CTA's next 128 segments are staged in
//                        shared memory by TMA bulk copies (cp.async.bulk + mbarrier, one
//                        row per copy number, two stages); every thread reads its own
//                        entry of each row and evaluates
//                            q = (xr' * sqrt(wub))^2 + wub            (2 DFMA)
//                            qmin = min(qmin, q)                      (DSETP + 2 FSEL)
//                        for PT ploidies per copy number.  Running products stay in
//                        registers, their binary exponents in shared memory.
//
// On B200 every warp-wide math instruction (fp64 or the selects of the running minimum)
// holds the SMSP math dispatch for ~2.2 cycles and they do not overlap
// (tools/microbench*.cu), so the kernel is built to minimise math instructions per
// (copy number, ploidy): 5 here against 6.8 in the fused kernel.
#include "sqz_async.cuh"
#include "sqz_baf_grid.cuh"
#include "../../include/sequenza_b200.h"

namespace sqz {

// ---- pre-pass: BAF factor per (column, copy number, segment) ---------------------------
__global__ void __launch_bounds__(256)
baf_colterm_kernel(BafGridArgs a, CnModel m, int ic0, int64_t N_total,
                   double* __restrict__ cw) {
  extern __shared__ __align__(16) double sm[];
  double* mb = sm;            // [T] model BAF of this column (R/model.R:9-14)
  double* b_a = sm + m.T;     // [nc] model BAF at B = 0
  double* b_invd = b_a + m.nc;  // [nc] 1 / (model BAF step per unit of B)
  const double cell = a.cell[ic0 + blockIdx.y];
  for (int t = threadIdx.x; t < m.T; t += blockDim.x) {
    int c = 0;
    while (m.toff[c + 1] <= t) ++c;
    mb[t] = theoretical_baf((double)(m.CNt_min + c), (double)(t - m.toff[c]), cell,
                            (double)m.CNn);
  }
  for (int c = threadIdx.x; c < m.nc; c += blockDim.x) {
    const double den = ((double)(m.CNt_min + c) * cell) + (double)m.CNn * (1 - cell);
    b_a[c] = (1 - cell) / den;
    b_invd[c] = den / cell;
  }
  __syncthreads();
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= N_total) return;
  const double2 bk = a.bk[s];
  const double Bf = a.Bf[s];
  const int64_t Npad = (N_total + 1) & ~(int64_t)1;  // even row pitch (16-byte aligned rows)
  double* out = cw + (int64_t)blockIdx.y * m.nc * Npad + s;
  for (int c = 0; c < m.nc; ++c) {
    // min over B of |xb| with xb = (mb_B - Bf) kb: NaN never wins (R/bayes.R:70-72,
    // which.max skips NaN).  The model BAF is linear in B (R/model.R:10-11), so for a
    // copy number with many rows only the two rows that bracket Bf are read: the nearest
    // row is always one of them (also when rounding puts floor() one off, because then Bf
    // sits on a row and that row is in both brackets).  Short rows are scanned.
    const int t0 = m.toff[c], nb = m.toff[c + 1] - t0;
    double m0 = INF;
    if (nb > 3) {
      int b0 = __double2int_rd((Bf - b_a[c]) * b_invd[c]);  // NaN -> 0
      b0 = max(0, min(b0, nb - 2));
      const double x0 = fma(mb[t0 + b0], bk.x, bk.y), x1 = fma(mb[t0 + b0 + 1], bk.x, bk.y);
      if (fabs(x0) < fabs(m0)) m0 = x0;
      if (fabs(x1) < fabs(m0)) m0 = x1;
    } else {
      for (int t = t0; t < t0 + nb; ++t) {
        const double x0 = fma(mb[t], bk.x, bk.y);
        if (fabs(x0) < fabs(m0)) m0 = x0;
      }
    }
    const double u = fma(m0, m0, 1.0);  // 1 + xb^2/5
    // ratio.priority keeps the BAF factor apart from the ratio term (R/bayes.R:81-88)
    out[(int64_t)c * Npad] = a.rp ? u : sqrt(u * m.w[c]);
  }
}

// ---- grid kernel ------------------------------------------------------------------------
// Staging of the pre-pass rows: one elected thread issues a bulk async copy (TMA, 1-D) per
// copy number for the CTA's next 128 segments -- each row of the scratch is contiguous over
// segments -- into the other half of a two-stage shared-memory buffer; an mbarrier per
// stage counts the bytes; every thread then reads its own 16-byte entry of each row.
template <int PT, bool RP, int NCT>
__global__ void __launch_bounds__(128, 2)
baf_grid_split_kernel(BafGridArgs a, CnModel m, const double* __restrict__ cw, int ic0,
                      int ub0, int nUb, int64_t N_total, double* __restrict__ partM,
                      long long* __restrict__ partE) {
  const int nc = NCT > 0 ? NCT : m.nc;  // compile-time for the two usual ranges (0..7, 0..20)
  extern __shared__ __align__(16) double sm[];
  const int nthr = blockDim.x, tid = threadIdx.x;
  const int nthr2 = nthr + 2;  // stage row: a run may start one segment early (alignment)
  const int64_t Npad = (N_total + 1) & ~(int64_t)1;  // row pitch of the scratch (even)
  double* lmr = sm;                                                // [nc][PT]
  double* stage = sm + m.nc * PT;                                  // [2][nc][nthr + 2]
  int* Ec = reinterpret_cast<int*>(stage + 2 * m.nc * nthr2);     // [PT][nthr]
  __shared__ double redM[4][PT];
  __shared__ long long redE[4][PT];
  __shared__ unsigned redBad[4];
  __shared__ __align__(8) unsigned long long bar[2];

  int64_t b = blockIdx.x;
  const int ub = ub0 + (int)(b % nUb);  // unit (column, tile), counted from the batch's first column
  b /= nUb;
  const int chunk = (int)(b % a.nChunk);
  const int64_t smp = b / a.nChunk;
  const int icl = ub / a.nPT, pt = ub - icl * a.nPT;
  const int ic = ic0 + icl;
  const int p0 = pt * PT;
  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    mbar_fence_init();
  }

  // log model depth ratio of this (cellularity, ploidy tile): R/model.R:1-7, :49-56
  const double cell = a.cell[ic];
  const double adr = a.adr[smp];
  for (int i = tid; i < m.nc * PT; i += nthr) {
    const int c = i / PT, j = i - c * PT;
    const int ip = min(p0 + j, a.nP - 1);  // ragged last tile repeats the last ploidy
    lmr[i] = log(theoretical_depth_ratio((double)(m.CNt_min + c), cell, a.ploidy[ip],
                                         (double)m.CNn, adr));
  }
#pragma unroll
  for (int j = 0; j < PT; ++j) Ec[j * nthr + tid] = 0;
  __syncthreads();

  const int64_t s0 = a.seg_off[smp], s1 = a.seg_off[smp + 1];
  const int64_t per = (s1 - s0 + a.nChunk - 1) / a.nChunk;
  const int64_t cb = s0 + (int64_t)chunk * per;
  const int64_t ce = min(cb + per, s1);
  const double* cwc = cw + (int64_t)icl * m.nc * Npad;  // this column's scratch rows

  double P[PT];
#pragma unroll
  for (int j = 0; j < PT; ++j) P[j] = 1.0;
  unsigned bad = 0;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);

  const int n_it = ce > cb ? (int)((ce - cb + nthr - 1) / nthr) : 0;
  auto issue = [&](int buf, int itn) {  // elected thread: rows of the CTA's iteration itn
    const int64_t base = cb + (int64_t)itn * nthr;
    const unsigned cnt = (unsigned)min((int64_t)nthr, ce - base);
    // bulk copies move multiples of 16 bytes from 16-byte aligned addresses: the run starts
    // at the even segment at or below `base` and ends at the next even count
    const int64_t b0 = base & ~(int64_t)1;
    const unsigned n8 = (unsigned)(((base - b0) + cnt + 1) & ~1u);
    mbar_expect_tx(&bar[buf], (unsigned)m.nc * n8 * (unsigned)sizeof(double));
    double* dst = stage + (buf * m.nc) * nthr2;
    const double* src = cwc + b0;
    for (int c = 0; c < nc; ++c, dst += nthr2, src += Npad)
      tma_bulk_g2s(dst, src, n8 * (unsigned)sizeof(double), &bar[buf]);
  };

  int64_t s = cb + tid;
  double2 rk_n = make_double2(0., 0.);
  if (s < ce) rk_n = a.rk[s];
  if (tid == 0 && n_it > 0) issue(0, 0);
  int it = 0;
  for (int itn = 0; itn < n_it; ++itn) {
    const int buf = itn & 1;
    const bool act = s < ce;
    const double2 rk = rk_n;
    const int64_t sn = s + nthr;
    if (itn + 1 < n_it) {
      __syncthreads();  // every thread has finished reading the other stage (iteration itn-1)
      if (tid == 0) issue(buf ^ 1, itn + 1);
      if (sn < ce) rk_n = a.rk[sn];  // ratio terms of the next segment: register prefetch
    }
    mbar_wait(&bar[buf], (unsigned)((itn >> 1) & 1));
    if (!act) {
      s = sn;
      continue;
    }
    // the stage holds the run that starts at the even segment at or below the CTA's first
    const double* col = stage + (buf * m.nc) * nthr2 + (int)(cb & 1) + tid;

    double qmin[PT];
    double sel[RP ? PT : 1], q2[RP ? PT : 1];  // q2: runner-up ratio score (ratio.priority)
#pragma unroll
    for (int j = 0; j < PT; ++j) qmin[j] = INF;
    if (RP) {
#pragma unroll
      for (int j = 0; j < (RP ? PT : 1); ++j) {
        sel[j] = 1.0;
        q2[j] = INF;
      }
    }
#pragma unroll 2
    for (int c = 0; c < nc; ++c) {
      const double cwv = col[c * nthr2];  // sqrt(wub), or the BAF factor u with ratio.priority
      const double* l = lmr + c * PT;
      if (!RP) {
        const double krw = rk.x * cwv, erw = rk.y * cwv, wub = cwv * cwv;
#pragma unroll
        for (int j = 0; j < PT; ++j) {
          const double x = fma(l[j], krw, erw);
          const double q = fma(x, x, wub);
          if (q < qmin[j]) qmin[j] = q;  // strict <: first minimum, NaN never wins
        }
      } else {
        // ratio.priority: CNt chosen on the ratio score alone (R/bayes.R:81-88)
        const double wc = m.w[c];
#pragma unroll
        for (int j = 0; j < PT; ++j) {
          const double x = fma(l[j], rk.x, rk.y);
          const double ar = fma(x, x, 1.0) * wc;
          if (ar < qmin[j]) {
            q2[RP ? j : 0] = qmin[j];
            qmin[j] = ar;
            sel[RP ? j : 0] = cwv;
          } else if (ar < q2[RP ? j : 0]) {
            q2[RP ? j : 0] = ar;
          }
        }
      }
    }
    if (RP) {
      // The reference picks the copy number on dt() + log(prior) (R/bayes.R:81); here it is
      // picked on (1 + x^2) prior^(-1/3).  When the two best are closer than rounding can
      // tell apart, the two orders may disagree and the BAF factor of the other copy number
      // would change the result by more than rounding: hand the grid point to the faithful
      // kernel instead.
#pragma unroll
      for (int j = 0; j < (RP ? PT : 1); ++j)
        if (q2[j] <= qmin[j] * (1.0 + 1e-12)) bad |= 1u << j;
    }
#pragma unroll
    for (int j = 0; j < PT; ++j) P[j] *= RP ? qmin[j] * sel[RP ? j : 0] : qmin[j];
    if ((++it & 15) == 0) {
      // every 16th segment: binary exponent -> integer column (q >= 1; sixteen factors cannot
      // overflow unless one is already beyond 2^63, which is caught as "bad" and repaired)
#pragma unroll
      for (int j = 0; j < PT; ++j) {
        const int hi = __double2hiint(P[j]), lo = __double2loint(P[j]);
        if ((unsigned)(hi - 0x00100000) >= 0x7FE00000u) bad |= 1u << j;
        Ec[j * nthr + tid] += (hi >> 20) - 1023;
        P[j] = __hiloint2double((hi & 0x000FFFFF) | 0x3FF00000, lo);
      }
    }
    s = sn;
  }

  // CTA reduction: sum of log(mantissa) (each in [0, log 2)), exact integer exponents
  const int lane = tid & 31, warp = tid >> 5, nwarp = (nthr + 31) >> 5;
#pragma unroll
  for (int j = 0; j < PT; ++j) {
    const int hi = __double2hiint(P[j]), lo = __double2loint(P[j]);
    if ((unsigned)(hi - 0x00100000) >= 0x7FE00000u) bad |= 1u << j;
    double v = log(__hiloint2double((hi & 0x000FFFFF) | 0x3FF00000, lo));
    long long e = (long long)Ec[j * nthr + tid] + ((hi >> 20) - 1023);
    for (int off = 16; off; off >>= 1) {
      v += __shfl_xor_sync(0xffffffffu, v, off);
      e += __shfl_xor_sync(0xffffffffu, e, off);
    }
    if (lane == 0) {
      redM[warp][j] = v;
      redE[warp][j] = e;
    }
  }
  for (int off = 16; off; off >>= 1) bad |= __shfl_xor_sync(0xffffffffu, bad, off);
  if (lane == 0) redBad[warp] = bad;
  __syncthreads();
  if (tid < PT && p0 + tid < a.nP) {
    double v = 0.;
    long long e = 0;
    unsigned bd = 0;
    for (int w = 0; w < nwarp; ++w) {
      v += redM[w][tid];
      e += redE[w][tid];
      bd |= redBad[w];
    }
    if ((bd >> tid) & 1u) v = __longlong_as_double(0x7FF8000000000000LL);
    const int64_t Gs = (int64_t)a.nCs * a.nP;
    const int64_t o = ((int64_t)chunk * a.S + smp) * Gs + (int64_t)ic * a.nP + p0 + tid;
    partM[o] = v;
    partE[o] = e;
  }
}

// ---- grid kernel with exact candidate pruning (plain mode, few copy numbers) -------------
// For one segment and one ploidy tile only a few copy numbers can be the minimiser.  Over a
// tile of sorted ploidies (checked on the host) the scaled ratio residual x_cj of copy
// number c is monotone in j, so it sweeps [xm_c - h_c, xm_c + h_c] (midpoint and half-width
// from the first and last table entries of the row), and
//     q_cj = wub_c (1 + x_cj^2)  >=  LB_c := wub_c (1 + max(|xm_c| - h_c, 0)^2).
// Each thread (a) computes LB_c for every copy number, (b) evaluates the one with the
// smallest bound over the whole tile, (c) takes Qmax >= max_j qmin_j, and (d) evaluates only
// the copy numbers with LB_c < Qmax (1 + 1e-12).  A copy number that is skipped has
// q_cj >= LB_c >= qmin_j for every j, so it cannot lower any minimum: the result equals
// evaluating all of them.  Candidates differ per thread; the rows of the model table are
// padded by two doubles so that threads reading different rows hit different banks.
// NCT > 0: the number of copy numbers is a compile-time constant (8 = sequenza.fit's CNt
// 0..7), which lets the bound pass unroll into independent chains held in registers.
template <int PT, int NCT>
__global__ void __launch_bounds__(128, 2)
baf_grid_pruned_kernel(BafGridArgs a, CnModel m, const double* __restrict__ cw, int ic0,
                       int ub0, int nUb, int64_t N_total, double* __restrict__ partM,
                       long long* __restrict__ partE, unsigned long long* __restrict__ stats) {
  constexpr int LS = PT + 2;  // padded row stride of the log-ratio table
  constexpr bool FIX = NCT > 0;
  const int nc = FIX ? NCT : m.nc;
  extern __shared__ __align__(16) double sm[];
  const int nthr = blockDim.x, tid = threadIdx.x;
  double* lmr = sm;                                                  // [nc][LS]
  double* lmid = sm + m.nc * LS;                                     // [nc] row midpoint
  double* lhw = lmid + m.nc;                                         // [nc] row half-width
  double* stage = sm + m.nc * LS + 2 * ((m.nc + 1) & ~1);           // [3][nc][nthr]
  double* LBc = stage + 3 * m.nc * nthr;                             // [nc][nthr]
  int* Ec = reinterpret_cast<int*>(LBc + m.nc * nthr);              // [PT][nthr]
  __shared__ double redM[4][PT];
  __shared__ long long redE[4][PT];
  __shared__ unsigned redBad[4];

  int64_t b = blockIdx.x;
  const int ub = ub0 + (int)(b % nUb);  // unit (column, tile), counted from the batch's first column
  b /= nUb;
  const int chunk = (int)(b % a.nChunk);
  const int64_t smp = b / a.nChunk;
  const int icl = ub / a.nPT, pt = ub - icl * a.nPT;
  const int ic = ic0 + icl;
  const int p0 = pt * PT;

  const double cell = a.cell[ic];
  const double adr = a.adr[smp];
  for (int i = tid; i < m.nc * PT; i += nthr) {
    const int c = i / PT, j = i - c * PT;
    const int ip = min(p0 + j, a.nP - 1);  // ragged last tile repeats the last ploidy
    lmr[c * LS + j] = log(theoretical_depth_ratio((double)(m.CNt_min + c), cell, a.ploidy[ip],
                                                  (double)m.CNn, adr));
  }
#pragma unroll
  for (int j = 0; j < PT; ++j) Ec[j * nthr + tid] = 0;
  __syncthreads();
  for (int c = tid; c < m.nc; c += nthr) {
    const double f = lmr[c * LS], l = lmr[c * LS + PT - 1];
    lmid[c] = 0.5 * (f + l);
    lhw[c] = fabs(0.5 * (f - l));
  }
  __syncthreads();

  const int64_t s0 = a.seg_off[smp], s1 = a.seg_off[smp + 1];
  const int64_t per = (s1 - s0 + a.nChunk - 1) / a.nChunk;
  const int64_t cb = s0 + (int64_t)chunk * per;
  const int64_t ce = min(cb + per, s1);
  const int64_t Npad = (N_total + 1) & ~(int64_t)1;  // row pitch of the scratch
  const double* cwc = cw + (int64_t)icl * m.nc * Npad;

  double P[PT];
#pragma unroll
  for (int j = 0; j < PT; ++j) P[j] = 1.0;
  unsigned bad = 0;
  unsigned n_eval = 0;  // copy numbers this thread evaluated (reported when profiling)
  const double INF = __longlong_as_double(0x7FF0000000000000LL);

  auto issue = [&](int buf, int64_t seg) {
    double* dst = stage + (buf * m.nc) * nthr + tid;
    const double* src = cwc + seg;
#pragma unroll
    for (int c = 0; c < nc; ++c, dst += nthr, src += Npad) cp_async8(dst, src);
    cp_async_commit();
  };

  // three-stage pipeline: the rows of a segment are requested two iterations before use
  int64_t s = cb + tid;
  double2 rk_n = make_double2(0., 0.);
  if (s < ce) {
    rk_n = a.rk[s];
    issue(0, s);
  } else {
    cp_async_commit();
  }
  if (s + nthr < ce) issue(1, s + nthr);
  else cp_async_commit();
  int it = 0, buf = 0;
  while (s < ce) {
    const double2 rk = rk_n;
    const int64_t sn = s + nthr;
    if (sn < ce) rk_n = a.rk[sn];
    if (sn + nthr < ce) issue(buf >= 1 ? buf - 1 : 2, sn + nthr);  // stage (buf + 2) % 3
    else cp_async_commit();                                        // keep the group count even
    cp_async_wait<2>();
    const double* col = stage + (buf * m.nc) * nthr + tid;

    // (a) lower bounds; the argmin only has to be a good guess, so it is taken on the
    // high words with the copy-number index packed into the low six bits
    const double akr = fabs(rk.x);
    unsigned kbest = 0xFFFFFFFFu;
    double lbr[FIX ? NCT : 1];
#pragma unroll
    for (int c = 0; c < nc; ++c) {
      double t = fabs(fma(lmid[c], rk.x, rk.y)) - lhw[c] * akr;
      if (t < 0.0) t = 0.0;
      const double rv = col[c * nthr], wv = rv * rv;
      const double lb = fma(wv, t * t, wv);
      if (FIX) lbr[FIX ? c : 0] = lb;
      else LBc[c * nthr + tid] = lb;
      kbest = min(kbest, ((unsigned)__double2hiint(lb) & 0xFFFFFFC0u) | (unsigned)c);
    }
    const int first = (int)(kbest & 63u) < nc ? (int)(kbest & 63u) : 0;

    double qmin[PT];
    auto eval = [&](int c) {
      const double cwv = col[c * nthr];
      const double* l = lmr + c * LS;
      const double krw = rk.x * cwv, erw = rk.y * cwv, wub = cwv * cwv;
#pragma unroll
      for (int j = 0; j < PT; ++j) {
        const double x = fma(l[j], krw, erw);
        const double q = fma(x, x, wub);
        if (q < qmin[j]) qmin[j] = q;  // strict <, NaN never wins
      }
    };
    // (b) most promising copy number: its q is the running minimum as it stands (a NaN
    // stays, reaches the product and is flagged for repair).  (c) Qmax: x is monotone over
    // the tile, so q = x^2 + wub peaks at one end; the larger high word plus one step of
    // 2^-20 also covers a last-place wobble of log() in the table.
    {
      const double cwv = col[first * nthr];
      const double* l = lmr + first * LS;
      const double krw = rk.x * cwv, erw = rk.y * cwv, wub = cwv * cwv;
#pragma unroll
      for (int j = 0; j < PT; ++j) {
        const double x = fma(l[j], krw, erw);
        qmin[j] = fma(x, x, wub);
      }
    }
    const unsigned qh = max((unsigned)__double2hiint(qmin[0]) & 0x7FFFFFFFu,
                            (unsigned)__double2hiint(qmin[PT - 1]) & 0x7FFFFFFFu);
    const double qmax = qh >= 0x7FF00000u ? INF : __hiloint2double((int)(qh + 1u), 0);
    // (d) the copy numbers that can still lower a minimum
    unsigned long long need = 0;
#pragma unroll
    for (int c = 0; c < nc; ++c)
      if ((FIX ? lbr[FIX ? c : 0] : LBc[c * nthr + tid]) < qmax) need |= 1ull << c;
    need &= ~(1ull << first);
    n_eval += 1u + (unsigned)__popcll(need);
    while (need) {  // per-thread trip count; the warp runs as long as its slowest lane
      const int c = __ffsll((long long)need) - 1;
      need &= need - 1;
      eval(c);
    }

#pragma unroll
    for (int j = 0; j < PT; ++j) P[j] *= qmin[j];
    if ((++it & 15) == 0) {  // q < 2^63 is ample; an overflow is flagged and repaired
#pragma unroll
      for (int j = 0; j < PT; ++j) {
        const int hi = __double2hiint(P[j]), lo = __double2loint(P[j]);
        if ((unsigned)(hi - 0x00100000) >= 0x7FE00000u) bad |= 1u << j;
        Ec[j * nthr + tid] += (hi >> 20) - 1023;
        P[j] = __hiloint2double((hi & 0x000FFFFF) | 0x3FF00000, lo);
      }
    }
    s = sn;
    buf = buf == 2 ? 0 : buf + 1;
  }

  const int lane = tid & 31, warp = tid >> 5, nwarp = (nthr + 31) >> 5;
#pragma unroll
  for (int j = 0; j < PT; ++j) {
    const int hi = __double2hiint(P[j]), lo = __double2loint(P[j]);
    if ((unsigned)(hi - 0x00100000) >= 0x7FE00000u) bad |= 1u << j;
    double v = log(__hiloint2double((hi & 0x000FFFFF) | 0x3FF00000, lo));
    long long e = (long long)Ec[j * nthr + tid] + ((hi >> 20) - 1023);
    for (int off = 16; off; off >>= 1) {
      v += __shfl_xor_sync(0xffffffffu, v, off);
      e += __shfl_xor_sync(0xffffffffu, e, off);
    }
    if (lane == 0) {
      redM[warp][j] = v;
      redE[warp][j] = e;
    }
  }
  for (int off = 16; off; off >>= 1) bad |= __shfl_xor_sync(0xffffffffu, bad, off);
  if (lane == 0) redBad[warp] = bad;
  if (stats) {
    for (int off = 16; off; off >>= 1) n_eval += __shfl_xor_sync(0xffffffffu, n_eval, off);
    if (lane == 0) atomicAdd(stats, (unsigned long long)n_eval);
  }
  __syncthreads();
  if (tid < PT && p0 + tid < a.nP) {
    double v = 0.;
    long long e = 0;
    unsigned bd = 0;
    for (int w = 0; w < nwarp; ++w) {
      v += redM[w][tid];
      e += redE[w][tid];
      bd |= redBad[w];
    }
    if ((bd >> tid) & 1u) v = __longlong_as_double(0x7FF8000000000000LL);
    const int64_t Gs = (int64_t)a.nCs * a.nP;
    const int64_t o = ((int64_t)chunk * a.S + smp) * Gs + (int64_t)ic * a.nP + p0 + tid;
    partM[o] = v;
    partE[o] = e;
  }
}

// ---- host side ---------------------------------------------------------------------------
// one column of pre-pass output: [copy number][segment], rows padded to an even length
size_t baf_split_column_bytes(int nc, int64_t N_total) {
  const size_t npad = ((size_t)(N_total > 0 ? N_total : 1) + 1) & ~(size_t)1;
  return (size_t)nc * npad * sizeof(double);
}

static size_t split_smem_bytes(int nc, int PT, int threads) {
  return (size_t)nc * PT * sizeof(double) + (size_t)2 * nc * (threads + 2) * sizeof(double) +
         (size_t)PT * threads * sizeof(int);
}

static size_t pruned_smem_bytes(int nc, int PT, int threads) {
  return (size_t)(nc * (PT + 2) + 2 * ((nc + 1) & ~1)) * sizeof(double) +
         (size_t)3 * nc * threads * sizeof(double) + (size_t)nc * threads * sizeof(double) +
         (size_t)PT * threads * sizeof(int);
}

template <int PT>
static int launch_pruned(const BafGridArgs& a, const CnModel& m, int threads, int64_t N_total,
                         const double* cw, int ic0, int ub0, int nUb, double* partM,
                         long long* partE, unsigned long long* stats, cudaStream_t st) {
  const size_t smb = pruned_smem_bytes(m.nc, PT, threads);
  const int64_t ctas = a.S * (int64_t)nUb * a.nChunk;
  if (ctas > 0x7fffffffLL) return fail(SQZ_ERR_SIZE, "grid too large for one launch");
  if (m.nc == 8) {  // CNt 0..7, what sequenza.fit always uses (R/bayes.R:40)
    SQZ_CUDA(cudaFuncSetAttribute(baf_grid_pruned_kernel<PT, 8>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    baf_grid_pruned_kernel<PT, 8><<<(unsigned)ctas, threads, smb, st>>>(
        a, m, cw, ic0, ub0, nUb, N_total, partM, partE, stats);
    SQZ_CUDA(cudaGetLastError());
    return SQZ_OK;
  }
  SQZ_CUDA(cudaFuncSetAttribute(baf_grid_pruned_kernel<PT, 0>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  baf_grid_pruned_kernel<PT, 0><<<(unsigned)ctas, threads, smb, st>>>(a, m, cw, ic0, ub0, nUb,
                                                                     N_total, partM, partE, stats);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

bool baf_split_supported(int nc, int PT, int threads) {
  return threads <= 128 && split_smem_bytes(nc, PT, threads) <= 200 * 1024;
}

template <int PT, bool RP>
static int launch_split(const BafGridArgs& a, const CnModel& m, int threads, int64_t N_total,
                        const double* cw, int ic0, int ub0, int nUb, double* partM,
                        long long* partE, cudaStream_t st) {
  const size_t smb = split_smem_bytes(m.nc, PT, threads);
  const int64_t ctas = a.S * (int64_t)nUb * a.nChunk;
  if (ctas > 0x7fffffffLL) return fail(SQZ_ERR_SIZE, "grid too large for one launch");
#define SQZ_SPLIT_NC(NCV)                                                                   \
  do {                                                                                      \
    SQZ_CUDA(cudaFuncSetAttribute(baf_grid_split_kernel<PT, RP, NCV>,                       \
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); \
    baf_grid_split_kernel<PT, RP, NCV><<<(unsigned)ctas, threads, smb, st>>>(               \
        a, m, cw, ic0, ub0, nUb, N_total, partM, partE);                                    \
  } while (0)
  if (!RP && m.nc == 21) SQZ_SPLIT_NC(21);
  else if (!RP && m.nc == 8) SQZ_SPLIT_NC(8);
  else SQZ_SPLIT_NC(0);
#undef SQZ_SPLIT_NC
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

// Runs pre-pass + grid kernel over the span's columns that hold units of this launch, in
// batches of columns that fit cw_bytes.
int baf_split_run(const BafGridArgs& a, const CnModel& m, int PT, int threads, int64_t N_total,
                  void* cw_buf, size_t cw_bytes, double* partM, long long* partE,
                  cudaStream_t st, bool prune, unsigned long long* stats) {
  if (prune && (a.rp || pruned_smem_bytes(m.nc, PT, threads) > 200 * 1024)) prune = false;
  const size_t per_col = baf_split_column_bytes(m.nc, N_total);
  int64_t nCb_max = (int64_t)(cw_bytes / per_col);
  if (nCb_max < 1) return fail(SQZ_ERR_SIZE, "split path needs room for one column");
  if (nCb_max > a.nCs) nCb_max = a.nCs;
  if (nCb_max > 65535) nCb_max = 65535;
  double* cw = (double*)cw_buf;
  const bool rp = a.rp != 0;
  const int64_t u_end = (int64_t)a.u_first + a.nU;  // units (column, tile) of this launch
  const int64_t c_lo = a.u_first / a.nPT, c_hi = (u_end + a.nPT - 1) / a.nPT;  // columns touched
  for (int64_t ic0 = c_lo; ic0 < c_hi; ic0 += nCb_max) {
    const int64_t ic1 = ic0 + nCb_max < c_hi ? ic0 + nCb_max : c_hi;
    const int nCb = (int)(ic1 - ic0);
    const int64_t ua = (a.u_first > ic0 * a.nPT ? a.u_first : ic0 * a.nPT) - ic0 * a.nPT;
    const int64_t ub = (u_end < ic1 * a.nPT ? u_end : ic1 * a.nPT) - ic0 * a.nPT;
    if (ub <= ua) continue;
    const int ub0 = (int)ua, nUb = (int)(ub - ua);
    dim3 grid((unsigned)((N_total + 255) / 256), (unsigned)nCb);
    baf_colterm_kernel<<<grid, 256, (size_t)(m.T + 2 * m.nc) * sizeof(double), st>>>(
        a, m, (int)ic0, N_total, cw);
    SQZ_CUDA(cudaGetLastError());
    int rc;
    if (prune) {
      if (PT == 8) rc = launch_pruned<8>(a, m, threads, N_total, cw, (int)ic0, ub0, nUb, partM, partE, stats, st);
      else if (PT == 16) rc = launch_pruned<16>(a, m, threads, N_total, cw, (int)ic0, ub0, nUb, partM, partE, stats, st);
      else rc = launch_pruned<32>(a, m, threads, N_total, cw, (int)ic0, ub0, nUb, partM, partE, stats, st);
      if (rc) return rc;
      continue;
    }
#define SQZ_SPLIT(PTV)                                                                          \
  rc = rp ? launch_split<PTV, true>(a, m, threads, N_total, cw, (int)ic0, ub0, nUb, partM, partE, st) \
          : launch_split<PTV, false>(a, m, threads, N_total, cw, (int)ic0, ub0, nUb, partM, partE, st)
    if (PT == 8) SQZ_SPLIT(8);
    else if (PT == 16) SQZ_SPLIT(16);
    else SQZ_SPLIT(32);
#undef SQZ_SPLIT
    if (rc) return rc;
  }
  return SQZ_OK;
}

}  // namespace sqz
