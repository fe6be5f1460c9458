// sqz_baf_colwarp.cu -- pruned grid kernel with one cellularity column per lane.
//
// baf_grid_pruned_kernel (sqz_baf_split.cu) gives every lane its own segment.  The set of
// copy numbers that survive the bound differs from segment to segment, and a warp runs as
// long as its slowest lane: 1.2 evaluations per thread become 2.6 per warp.  For ONE
// segment the surviving set changes slowly with the cellularity, so here the 32 lanes of a
// warp take 32 neighbouring columns of the same segment and ploidy tile: 1.33 evaluations
// per warp on the same data (tools/sim/lane_mapping.py).
//
//   pre-pass  baf_colterm_cg_kernel: lanes = the 32 columns of a group; writes ONE number per
//             (segment, copy number, column) as [group][segment][copy number][32 columns],
//             so the rows of one segment are contiguous (2 KB for CNt 0..7).  For 8 copy
//             numbers it is sqrt(wub) and wub is the square (one DMUL where needed); for
//             21 it is wub and the square root is taken for the one or two copy numbers
//             that get evaluated.  Half the scratch and traffic of the (wub, sqrt) pairs.
//   CTA       = ONE warp: 32 columns x one ploidy tile x one segment chunk.  Warps of
//             different tiles cost different amounts (the candidate counts differ), so
//             nothing is shared between them and the hardware scheduler balances them;
//             the tiles of one (group, chunk) are neighbours in the grid and find each
//             other's rows in L2.
//   ring      two stages of 4 (CNt 0..7) or 1 (CNt 0..20) segments; lane 0 fills a stage
//             with one TMA bulk copy (plus one for the segments' ratio terms) and
//             completion is counted on an mbarrier; the stage is refilled as soon as the
//             warp has left it.
//   batch     the segments of a stage are analysed together (bounds, argmin, dynamic loads,
//             warp vote: a serial chain per segment, independent between segments; results
//             parked in shared memory), then evaluated one after the other by ONE copy of
//             the evaluation code (long runs of independent DFMA/DMUL; the kernel has to
//             stay small: with every warp at a different place, a larger body stalls on
//             instruction fetch).
//   thread    = one (column, tile): the tile's -log(ploidy term) centred on its midpoint in
//             registers, running products in registers, exponents in shared memory
//
// log mr of (copy number c, ploidy j) is split as  la_c - ld_j  (R/model.R:1-7 has
// mr = adr * A_c / D_j), so x = (la_c - ld_j - log r) k sqrt(wub) = fma(nld_j, krw, base_c):
// the per-thread table is PT + nc numbers, not PT * nc.
//
// Candidate rule as in baf_grid_pruned_kernel.  The warp-uniform number of extra candidates
// (0 in 3/4 of the iterations, 1 in 1/5) selects the loop: the minimum over candidates is
// taken ploidy by ploidy, so no array of minima is kept next to the products.
#include <stdlib.h>

#include <type_traits>

#include "sqz_async.cuh"
#include "sqz_baf_grid.cuh"
#include "../../include/sequenza_b200.h"

namespace sqz {

namespace {
// Segments per stage of the ring = segments whose analyses run together (their serial
// chains overlap in one basic block).  Two stages: one is worked on, the other in flight.
// (4 for CNt 0..7; 1 or 2 for CNt 0..20, see baf_colwarp_run)
constexpr int kPreSeg = 256;  // segments per pre-pass CTA
}  // namespace

// sqrt(wub) as the grid kernel stores it for CNt 0..7.  The grid kernel turns the high word
// of this number into a float with one integer instruction, which is right for [1, 2^128)
// and for +Inf / NaN (they wrap to a small float: a bound that is too low, which is safe);
// wub >= 1 always (the priors are normalised, R/bayes.R:59).  A factor of 2^128 and more
// (a BAF residual of 1e38 standard deviations) is stored as +Inf: the copy number can then
// only be chosen if every other is infinite too, and that grid point is flagged and repaired.
__device__ __forceinline__ double cw_sqrt_for_float(double wub) {
  const double r = sqrt(wub);
  return r < 0x1p128 ? r : __longlong_as_double(0x7FF0000000000000LL);
}

// ---- pre-pass: BAF factor per (column group, segment, copy number, column) --------------
// Same arithmetic as baf_colterm_kernel (sqz_baf_split.cu); a short row scan replaces the
// bracket search (at most 4 rows per copy number for CNt <= 7).
__global__ void __launch_bounds__(256)
baf_colterm_cg_kernel(BafGridArgs a, CnModel m, int ic0, int nCb, int64_t N_total, int store_sqrt,
                      double* __restrict__ cwg) {
  extern __shared__ __align__(16) double sm[];
  double* mb = sm;                   // [T][32] model BAF of the group's columns (R/model.R:9-14)
  double* b_a = sm + m.T * 32;       // [nc][32] model BAF at B = 0
  double* b_invd = b_a + m.nc * 32;  // [nc][32] 1 / (model BAF step per unit of B)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int cg = blockIdx.y;
  const int icl = min(cg * 32 + lane, nCb - 1);  // ragged last group repeats a column
  const double cell = a.cell[ic0 + icl];
  for (int t = warp; t < m.T; t += nwarp) {
    int c = 0;
    while (m.toff[c + 1] <= t) ++c;
    mb[t * 32 + lane] = theoretical_baf((double)(m.CNt_min + c), (double)(t - m.toff[c]), cell,
                                        (double)m.CNn);
  }
  for (int c = warp; c < m.nc; c += nwarp) {
    const double den = ((double)(m.CNt_min + c) * cell) + (double)m.CNn * (1 - cell);
    b_a[c * 32 + lane] = (1 - cell) / den;
    b_invd[c * 32 + lane] = den / cell;
  }
  __syncthreads();
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const int64_t sb = (int64_t)blockIdx.x * kPreSeg;
  const int64_t se = min(sb + kPreSeg, N_total);
  for (int64_t s = sb + warp; s < se; s += nwarp) {
    const double2 bk = a.bk[s];
    const double Bf = a.Bf[s];
    double* out = cwg + (((int64_t)cg * N_total + s) * m.nc) * 32 + lane;
    for (int c = 0; c < m.nc; ++c) {
      // min over B of |xb|: NaN never wins (R/bayes.R:70-72); long rows read only the two
      // rows that bracket Bf (see baf_colterm_kernel)
      const int t0 = m.toff[c], nb = m.toff[c + 1] - t0;
      double m0 = INF;
      if (nb > 3) {
        int b0 = __double2int_rd((Bf - b_a[c * 32 + lane]) * b_invd[c * 32 + lane]);  // NaN -> 0
        b0 = max(0, min(b0, nb - 2));
        const double x0 = fma(mb[(t0 + b0) * 32 + lane], bk.x, bk.y);
        const double x1 = fma(mb[(t0 + b0 + 1) * 32 + lane], bk.x, bk.y);
        if (fabs(x0) < fabs(m0)) m0 = x0;
        if (fabs(x1) < fabs(m0)) m0 = x1;
      } else {
        for (int t = t0; t < t0 + nb; ++t) {
          const double x0 = fma(mb[t * 32 + lane], bk.x, bk.y);
          if (fabs(x0) < fabs(m0)) m0 = x0;
        }
      }
      const double wub = fma(m0, m0, 1.0) * m.w[c];
      out[c * 32] = store_sqrt ? cw_sqrt_for_float(wub) : wub;
    }
  }
}

// The same pre-pass for CNt 0..7 with two or more normal copies (what sequenza.fit always
// runs, R/bayes.R:40): copy number c has c/2 + 1 BAF rows, 20 in all, known at compile time.
// A lane keeps its column's 20 model BAFs and 8 prior weights in registers and the loops
// unroll completely: ~300 instructions per segment and warp instead of ~2,600 (the generic
// kernel walks the type table with run-time trip counts), which leaves the kernel bound by
// the 8 bytes per (segment, copy number, column) it writes.  Same arithmetic, same bits:
// scanning all rows finds the same nearest row as the bracket search above.
__global__ void __launch_bounds__(256)
baf_colterm_cg8_kernel(BafGridArgs a, CnModel m, int ic0, int nCb, int64_t N_total,
                       double* __restrict__ cwg) {
  constexpr int NC = 8, T = 20;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int cg = blockIdx.y;
  const int icl = min(cg * 32 + lane, nCb - 1);  // ragged last group repeats a column
  const double cell = a.cell[ic0 + icl];
  double mb[T], w[NC];
  {
    int t = 0;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      w[c] = m.w[c];
#pragma unroll
      for (int b = 0; b <= c / 2; ++b, ++t)
        mb[t] = theoretical_baf((double)c, (double)b, cell, (double)m.CNn);  // R/model.R:9-14
    }
  }
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const int64_t sb = (int64_t)blockIdx.x * kPreSeg;
  const int64_t se = min(sb + kPreSeg, N_total);
  for (int64_t s = sb + warp; s < se; s += nwarp) {
    const double2 bk = a.bk[s];
    double* out = cwg + (((int64_t)cg * N_total + s) * NC) * 32 + lane;
    int t = 0;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      double m0 = INF;  // min over B of |xb|: NaN never wins (R/bayes.R:70-72)
#pragma unroll
      for (int b = 0; b <= c / 2; ++b, ++t) {
        const double x0 = fma(mb[t], bk.x, bk.y);
        if (fabs(x0) < fabs(m0)) m0 = x0;
      }
      out[c * 32] = cw_sqrt_for_float(fma(m0, m0, 1.0) * w[c]);
    }
  }
}

// ---- grid kernel -------------------------------------------------------------------------
// What the analysis of one segment leaves for its evaluation, per lane, in shared memory:
//   three doubles  k1, b1, w1: first candidate, x_j = fma(nld_j, k1, b1), q_j = fma(x_j, x_j, w1)
//   word 0         further copy numbers this lane cannot exclude | first candidate << 24
//                  (a lane without a further one repeats its first)
//   word 1         upper bound of the binary exponent the products gain | (the longest
//                  candidate list in the warp, warp-uniform) << 16
// (the two words share an 8-byte slot next to the doubles: a record is [4][32] doubles)

// (Measured and dropped: two warps sharing a CTA's tile, 16 ploidies and half of a stage's
// analyses each, handed over in shared memory -- 128 registers, four warps per scheduler:
// 31.4 ms on config 3 against 31.3 for one warp; with three per scheduler at 168 registers
// the same.  More warps do not help: the kernel is bound by issue, not by latency.
// Also measured and dropped: the two phases on two warps of a CTA -- an analyst that requests
// and analyses the stages one ahead and an evaluator that only streams the fp64 work, meeting
// on mbarriers over a ring of three stages, four such CTAs per SM with the evaluators spread
// over the schedulers by %warpid: 37.2 ms.  A lone evaluator warp per scheduler reaches 2.7
// cycles per DFMA in the one-candidate loop but loses the time between segments and in the
// two-candidate loop that a second warp of the same kind fills here.)
template <int PT, int NC, int SB, int OCC, bool F32>
__global__ void __launch_bounds__(32, OCC)
baf_grid_colwarp_kernel(BafGridArgs a, CnModel m, const double* __restrict__ cwg, int ic0,
                        int nCb, int ub0, int nUb, int64_t N_total, double* __restrict__ partM,
                        long long* __restrict__ partE, unsigned long long* __restrict__ stats) {
  constexpr int kSegDoubles = NC * 32;  // one segment of a column group
  constexpr int kSegB = SB;             // segments per stage = segments analysed together
  constexpr int kRing = 2;              // one stage is worked on while the other is in flight
  constexpr bool SQ = NC <= 8;          // the scratch holds sqrt(wub) (else wub)
  constexpr int nthr = 32;
  extern __shared__ __align__(16) double sm[];
  const int lane = threadIdx.x, tid = threadIdx.x;
  double* stage = sm;                                                           // [kRing][kSegB][NC][32]
  RkRec* rxS = reinterpret_cast<RkRec*>(stage + kRing * kSegB * kSegDoubles);   // [kRing][kSegB]
  double* lm = reinterpret_cast<double*>(rxS + kRing * kSegB) + lane;          // [NC][32]
  float* lmf = reinterpret_cast<float*>(lm - lane + NC * 32) + lane;            // [NC][32] (F32)
  int* Ec = reinterpret_cast<int*>(lmf - lane + (F32 ? NC * 32 : 0));           // [PT][32]
  double* pk = reinterpret_cast<double*>(Ec + PT * 32) + lane;                  // [kSegB][4][32]
  __shared__ __align__(8) unsigned long long full[kRing];

  // CTA -> (unit of this launch, chunk, sample); units run tile-fastest inside a column
  // group, so the tiles of one (group, chunk) are neighbours in the grid and share L2 rows
  int64_t b = blockIdx.x;
  const int ub = ub0 + (int)(b % nUb);  // unit, counted from the first group of this batch
  b /= nUb;
  const int chunk = (int)(b % a.nChunk);
  const int64_t smp = b / a.nChunk;
  const int cg = ub / a.nPT, tile = ub - cg * a.nPT;

  if (lane == 0) {
    for (int s = 0; s < kRing; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncwarp();
  const int p0 = tile * PT;
  const int icl_raw = cg * 32 + lane;
  const int icl = min(icl_raw, nCb - 1);  // ragged last group: the lane repeats a column
  const double cell = a.cell[ic0 + icl];
  const double adr = a.adr[smp];

  // ---- per-thread model: nld_j = -log D_j centred on the tile midpoint -----------------
  // (the logarithms run in rolled loops through the not yet used stage buffer: one copy of
  // log() in the kernel instead of PT + NC inlined ones keeps the code small enough for the
  // instruction cache when every warp of an SM is at a different place)
  double nld[PT];
  static_assert(PT * 32 <= kRing * kSegB * kSegDoubles, "stage buffer too small for the set-up");
#pragma unroll 1
  for (int j = 0; j < PT; ++j) {
    const int ip = min(p0 + j, a.nP - 1);  // ragged last tile repeats the last ploidy
    stage[j * 32 + lane] = -log((a.ploidy[ip] / 2.0 * cell) + 1 - cell);
  }
#pragma unroll
  for (int j = 0; j < PT; ++j) nld[j] = stage[j * 32 + lane];
  const double nmid = 0.5 * (nld[0] + nld[PT - 1]);
#pragma unroll
  for (int j = 0; j < PT; ++j) nld[j] -= nmid;
#pragma unroll 1
  for (int c = 0; c < NC; ++c) {
    const double A = (1 - cell) + ((double)(m.CNt_min + c) / (double)m.CNn * cell);
    const double l = log(adr * A) + nmid;
    lm[c * 32] = l;
    if (F32) lmf[c * 32] = __double2float_rn(l);
  }
#pragma unroll
  for (int j = 0; j < PT; ++j) Ec[j * nthr + tid] = 0;
  // The bounds of the analysis are taken in fp32, rounded so that they never exceed the
  // fp64 value (see analyse).  hcf: half-width of the tile in -log D plus the rounding the
  // fp32 midpoint residual can carry, per unit of |k|: 2^-22 (|lm| + half-width) covers the
  // conversions of lm, k, b and the fused multiply-add (2^-24 each) with room to spare; the
  // part that scales with |b| comes with the segment (rkx.f.z).
  const double hw = fmax(fabs(nld[0]), fabs(nld[PT - 1]));
  float hcf = 0.f;
  if (F32) {
    float lmax = 0.f;
#pragma unroll 1
    for (int c = 0; c < NC; ++c) {
      const float v = fabsf(lmf[c * 32]);
      // (-Inf, cellularity 1 with CNt 0, gives an infinite residual whatever the margin)
      if (v < __int_as_float(0x7F800000)) lmax = fmaxf(lmax, v);
    }
    hcf = __double2float_ru(hw * (1.0 + 0x1p-22) + 0x1p-22 * (double)lmax);
  }

  const int64_t s0 = a.seg_off[smp], s1 = a.seg_off[smp + 1];
  const int64_t per = (s1 - s0 + a.nChunk - 1) / a.nChunk;
  const int64_t cb = s0 + (int64_t)chunk * per;
  const int64_t ce = min(cb + per, s1);
  const double* cwc = cwg + (int64_t)cg * N_total * kSegDoubles;  // this group's rows

  // Requests run one stage ahead of the consumption; the state below is warp-uniform, lane
  // 0 issues.  A stage is requested again only after every lane has left it (the
  // __syncwarp before the call).
  const double* f_src = cwc + cb * kSegDoubles;
  const RkRec* f_rx = a.rkx + cb;
  int f_left = (int)(ce > cb ? ce - cb : 0);  // segments not yet requested
  unsigned f_s = 0;
  // (shared-memory addresses in the 32-bit window, converted once)
  const unsigned stage_a = smem_u32(stage), rx_a = smem_u32(rxS), full_a = smem_u32(full);
  constexpr unsigned kStageBytes = kSegB * kSegDoubles * sizeof(double);
  auto fetch = [&]() {
    if (f_left <= 0) return;
    const unsigned cnt = f_left < kSegB ? (unsigned)f_left : (unsigned)kSegB;
    if (lane == 0) {
      const unsigned bar = full_a + f_s * 8u;
      mbar_expect_tx_a(bar, cnt * (unsigned)(kSegDoubles * sizeof(double) + sizeof(RkRec)));
      tma_bulk_g2s_a(stage_a + f_s * kStageBytes, f_src,
                     cnt * (unsigned)(kSegDoubles * sizeof(double)), bar);
      tma_bulk_g2s_a(rx_a + f_s * (unsigned)(kSegB * sizeof(RkRec)), f_rx,
                     cnt * (unsigned)sizeof(RkRec), bar);
    }
    f_src += kSegB * kSegDoubles;
    f_rx += kSegB;
    f_left -= kSegB;
    f_s ^= 1u;
  };

  double P[PT];
#pragma unroll
  for (int j = 0; j < PT; ++j) P[j] = 1.0;
  unsigned bad = 0, n_eval = 0;
  // Upper bound of the binary exponent the products may have gained since the last
  // renormalisation: every factor is at most the first candidate's q at a tile end.
  int grown = 0;

  // ---- analysis of one segment: which copy numbers have to be evaluated ------------------
  // A serial chain (bounds -> argmin -> dynamic loads -> Qmax -> candidate mask -> warp
  // vote).  The segments of a stage are analysed together, in one basic block, so that
  // their chains overlap; the evaluations (long runs of independent DFMA/DMUL) follow.
  // B segments at once, written phase by phase ACROSS the segments (all bound passes, all
  // argmins, all dynamic loads, ...): the hardware issues a warp's instructions in order, so
  // independent chains overlap only if they are interleaved in the instruction stream.  The
  // warp votes and the stores come last: a warp-synchronous operation or a shared-memory
  // store in the middle would pin the order.
  // (base: the first segment's slot in the ring; slot0: its slot in the stage, where the
  // results are parked)
  auto analyse = [&](int base, int slot0, auto bt) {
    constexpr int B = decltype(bt)::value;
    double2 rk[B];
    const double* col[B];
    unsigned lbh[B][NC], kbest[B];
    float kf[B], bf[B], sf[B];
#pragma unroll
    for (int g = 0; g < B; ++g) {
      rk[g] = rxS[base + g].rk;
      if (F32) {
        const float4 rf = rxS[base + g].f;
        kf[g] = rf.x;
        bf[g] = rf.y;
        // what |x_mid| may lose: the half-width of the tile and the rounding
        sf[g] = __fmaf_ru(fabsf(rf.x), hcf, rf.z);
      }
      col[g] = stage + (base + g) * kSegDoubles + lane;  // stored value of c: [c * 32]
      kbest[g] = 0xFFFFFFFFu;
    }
    // (a) lower bound of every copy number over the tile, wub (1 + t^2) with t = max(|x_mid|
    // - half-width, 0); argmin with the copy number packed into the low bits.
    if (F32) {
      // CNt 0..7: in fp32 and never above the fp64 value.  The midpoint residual loses its
      // rounding margin (sf), the stored factor is cut to the 20 mantissa bits of its high
      // word (one multiply-add on the integer pattern; the pre-pass keeps it in the range
      // where that is right), products and sums round down.  A bound that is a little low
      // costs an evaluation, never a result.  Bit patterns of non-negative floats order as
      // integers.
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const float lmc = lmf[c * 32];
#pragma unroll
        for (int g = 0; g < B; ++g) {
          const float t = fmaxf(__fsub_rd(fabsf(fmaf(lmc, kf[g], bf[g])), sf[g]), 0.f);
          const unsigned hi = reinterpret_cast<const unsigned*>(col[g])[c * 64 + 1];
          const float rv = __uint_as_float(hi * 8u + 0x40000000u);  // see cw_sqrt_for_float
          float lb;
          if (SQ) {
            const float u = __fmul_rd(rv, t);
            lb = __fmaf_rd(u, u, __fmul_rd(rv, rv));
          } else {
            lb = __fmul_rd(rv, __fmaf_rd(t, t, 1.f));
          }
          lbh[g][c] = __float_as_uint(lb);
          kbest[g] = min(kbest[g], (lbh[g][c] & 0xFFFFFFC0u) | (unsigned)c);
        }
      }
    } else {
      // CNt 0..20: in fp64, compared on the high words (the float copy of lm would cost two
      // of the twelve warps per SM).  max(t0, 0) = t0/2 + |t0/2| with exact halves: one DADD.
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const double lmc = lm[c * 32];
#pragma unroll
        for (int g = 0; g < B; ++g) {
          const double hx = 0.5 * rk[g].x, hy = 0.5 * rk[g].y;
          const double t0 = fabs(fma(lmc, hx, hy)) - hw * fabs(hx);
          const double t = t0 + fabs(t0);
          const double rv = col[g][c * 32], u = SQ ? rv * t : t * t;
          lbh[g][c] = (unsigned)__double2hiint(SQ ? fma(u, u, rv * rv) : fma(rv, u, rv));
          kbest[g] = min(kbest[g], (lbh[g][c] & 0xFFFFFFC0u) | (unsigned)c);
        }
      }
    }
    // (b) the most promising copy number
    int first[B];
    double v1[B], lf[B];
#pragma unroll
    for (int g = 0; g < B; ++g) {
      first[g] = (int)(kbest[g] & 63u) < NC ? (int)(kbest[g] & 63u) : 0;
      v1[g] = col[g][first[g] * 32];
      lf[g] = lm[first[g] * 32];
    }
    double k1[B], b1[B], w1[B];
#pragma unroll
    for (int g = 0; g < B; ++g) {
      const double r1 = SQ ? v1[g] : sqrt(v1[g]);
      w1[g] = SQ ? v1[g] * v1[g] : v1[g];
      k1[g] = rk[g].x * r1;
      b1[g] = fma(lf[g], k1[g], rk[g].y * r1);
    }
    // (c) Qmax from the two ends of the tile
    unsigned qh[B];
#pragma unroll
    for (int g = 0; g < B; ++g) {
      const double xa = fma(nld[0], k1[g], b1[g]), xb = fma(nld[PT - 1], k1[g], b1[g]);
      qh[g] = max((unsigned)__double2hiint(fma(xa, xa, w1[g])) & 0x7FFFFFFFu,
                  (unsigned)__double2hiint(fma(xb, xb, w1[g])) & 0x7FFFFFFFu);
    }
    // (d) copy numbers that can still lower a minimum: bound < Qmax, the double (qh + 1, 0),
    // which is exact as a float (20 mantissa bits; beyond the float range: +Inf, which
    // keeps every finite bound)
    unsigned need[B];
#pragma unroll
    for (int g = 0; g < B; ++g) {
      // (fp64 bounds: high word of the bound at most qh; Qmax = +Inf keeps every finite one)
      const unsigned qcap = F32 ? min(qh[g] + 1u, 0x47F00000u) * 8u + 0x40000000u
                                : (qh[g] >= 0x7FF00000u ? 0x7FF00000u : qh[g] + 1u);
      need[g] = 0;
#pragma unroll
      for (int c = 0; c < NC; ++c)
        if (lbh[g][c] < qcap) need[g] |= 1u << c;
      need[g] &= ~(1u << first[g]);
    }
    unsigned nmax[B];
#pragma unroll
    for (int g = 0; g < B; ++g) nmax[g] = __reduce_max_sync(0xffffffffu, (unsigned)__popc(need[g]));
#pragma unroll
    for (int g = 0; g < B; ++g) {
      double* pr = pk + (slot0 + g) * 4 * 32;
      pr[0 * 32] = k1[g];
      pr[1 * 32] = b1[g];
      pr[2 * 32] = w1[g];
      *reinterpret_cast<uint2*>(pr + 3 * 32) =
          make_uint2(need[g] | ((unsigned)first[g] << 24),
                     (qh[g] >> 20) | (nmax[g] << 16));  // exponent field: at most 0x7FF
    }
  };

  // ---- evaluation of one analysed segment: the products take in its minimum ----------------
  struct CwPar {
    double k1, b1, w1;
    unsigned need, nmax;
    int first, gr;
  };
  auto load_par = [&](const double* pr) {
    CwPar sg;
    sg.k1 = pr[0 * 32];
    sg.b1 = pr[1 * 32];
    sg.w1 = pr[2 * 32];
    const uint2 w = *reinterpret_cast<const uint2*>(pr + 3 * 32);
    sg.need = w.x & 0x00FFFFFFu;
    sg.first = (int)(w.x >> 24);
    sg.gr = (int)(w.y & 0xFFFFu) - 1021;
    sg.nmax = w.y >> 16;
    return sg;
  };
  auto evaluate = [&](int slot, const CwPar& sg) {
    if (sg.nmax == 0) {
      // common case (3 of 4): the first candidate alone; no minimum to take
#pragma unroll
      for (int j = 0; j < PT; ++j) {
        const double x = fma(nld[j], sg.k1, sg.b1);
        P[j] *= fma(x, x, sg.w1);
      }
    } else {
      const double2 rk = rxS[slot].rk;
      const double* col = stage + slot * kSegDoubles + lane;
      unsigned need = sg.need;
      // a lane with no further candidate repeats its first one: the minimum is unchanged
      const int c2 = need ? __ffs((int)need) - 1 : sg.first;
      need &= need - 1;
      const double v2 = col[c2 * 32];
      const double r2 = SQ ? v2 : sqrt(v2), w2 = SQ ? v2 * v2 : v2;
      const double k2 = rk.x * r2;
      const double b2 = fma(lm[c2 * 32], k2, rk.y * r2);
      if (sg.nmax == 1) {
#pragma unroll
        for (int j = 0; j < PT; ++j) {
          const double x = fma(nld[j], sg.k1, sg.b1), y = fma(nld[j], k2, b2);
          double q = fma(x, x, sg.w1);
          const double q2 = fma(y, y, w2);
          if (q2 < q) q = q2;  // strict <: a NaN in the first stays and is flagged
          P[j] *= q;
        }
      } else {
        // rare (1 in 20): several further candidates.  Eight ploidies at a time keep the
        // minima in registers; the candidate list is walked again for every block.
#pragma unroll
        for (int jb = 0; jb < PT; jb += 8) {
          double q[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const double x = fma(nld[jb + j], sg.k1, sg.b1), y = fma(nld[jb + j], k2, b2);
            q[j] = fma(x, x, sg.w1);
            const double q2 = fma(y, y, w2);
            if (q2 < q[j]) q[j] = q2;
          }
          unsigned nd = need;
          for (unsigned k = 2; k <= sg.nmax; ++k) {
            const int c3 = nd ? __ffs((int)nd) - 1 : sg.first;
            nd &= nd - 1;
            const double v3 = col[c3 * 32];
            const double r3 = SQ ? v3 : sqrt(v3), w3 = SQ ? v3 * v3 : v3;
            const double k3 = rk.x * r3;
            const double b3 = fma(lm[c3 * 32], k3, rk.y * r3);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const double y = fma(nld[jb + j], k3, b3);
              const double q3 = fma(y, y, w3);
              if (q3 < q[j]) q[j] = q3;
            }
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) P[jb + j] *= q[j];
        }
      }
    }
    n_eval += 1u + sg.nmax;  // what the warp executes: every lane follows the longest list
    grown += sg.gr;
  };

  fetch();
  int s = 0;
  unsigned phase = 0;
#pragma unroll 1
  for (int left = (int)(ce > cb ? ce - cb : 0); left > 0; left -= kSegB) {
    __syncwarp();
    fetch();  // the stage left in the previous round is requested again
    mbar_wait_a(full_a + (unsigned)s * 8u, phase);
    const int base = s * kSegB;
    const int nsg = left < kSegB ? left : kSegB;
    if (nsg == kSegB) {
      analyse(base, 0, std::integral_constant<int, kSegB>());
    } else {  // ragged last stage of the chunk
      int k = 0;
      if (kSegB > 2 && nsg >= 2) {
        analyse(base, 0, std::integral_constant<int, (kSegB > 2 ? 2 : 1)>());
        k = 2;
      }
      if (nsg > k) analyse(base + k, k, std::integral_constant<int, 1>());
    }
    // (each lane reads back only what it wrote itself: no warp synchronisation needed; the
    // record is loaded where it is used -- fetching the next one during the evaluation cost
    // thirteen instructions per segment in moves and index arithmetic, 3 % of the fit)
    {
      const double* pr = pk;
#pragma unroll 1
      for (int sg = 0; sg < nsg; ++sg, pr += 4 * 32) evaluate(base + sg, load_par(pr));
    }
    // Once per stage: the products are renormalised (binary exponent -> integer column)
    // before they can overflow; q >= 1, so they never underflow.  `grown` bounds the
    // exponent gained since the last time; a stage that gains more than 2^300 on its own
    // overflows to Inf, which is flagged and repaired like any other non-finite product.
    if (__any_sync(0xffffffffu, grown > 700)) {
      grown = 0;
#pragma unroll
      for (int j = 0; j < PT; ++j) {
        const int hi = __double2hiint(P[j]), lo = __double2loint(P[j]);
        if ((unsigned)(hi - 0x00100000) >= 0x7FE00000u) bad |= 1u << j;
        Ec[j * nthr + tid] += (hi >> 20) - 1023;
        P[j] = __hiloint2double((hi & 0x000FFFFF) | 0x3FF00000, lo);
      }
    }
    if (++s == kRing) {
      s = 0;
      phase ^= 1u;
    }
  }

  if (stats) {
    if (icl_raw >= nCb) n_eval = 0;
    for (int off = 16; off; off >>= 1) n_eval += __shfl_xor_sync(0xffffffffu, n_eval, off);
    if (lane == 0 && n_eval) atomicAdd(stats, (unsigned long long)n_eval);
  }
  if (icl_raw >= nCb) return;
  const int64_t Gs = (int64_t)a.nCs * a.nP;
  const int64_t o = ((int64_t)chunk * a.S + smp) * Gs + (int64_t)(ic0 + icl) * a.nP + p0;
  // the products leave through the (drained) stage buffer so that the loop with the
  // logarithm can stay rolled; a lane reads back only its own values
#pragma unroll
  for (int j = 0; j < PT; ++j) stage[j * 32 + lane] = P[j];
#pragma unroll 1
  for (int j = 0; j < PT; ++j) {
    if (p0 + j < a.nP) {
      const double pj = stage[j * 32 + lane];
      const int hi = __double2hiint(pj), lo = __double2loint(pj);
      const bool bd = ((bad >> j) & 1u) || (unsigned)(hi - 0x00100000) >= 0x7FE00000u;
      const double v = log(__hiloint2double((hi & 0x000FFFFF) | 0x3FF00000, lo));
      partM[o + j] = bd ? __longlong_as_double(0x7FF8000000000000LL) : v;
      partE[o + j] = (long long)Ec[j * nthr + tid] + ((hi >> 20) - 1023);
    }
  }
}

// ---- host side ---------------------------------------------------------------------------
static size_t colwarp_smem_bytes(int NC, int PT, int segb, bool f32) {
  const int ring = 2;
  return (size_t)ring * segb * NC * 32 * sizeof(double) +
         ring * segb * sizeof(RkRec) +
         (size_t)NC * 32 * (sizeof(double) + (f32 ? sizeof(float) : 0)) +
         (size_t)PT * 32 * sizeof(int) +
         (size_t)segb * 32 * 4 * sizeof(double);
}

static size_t colterm_cg_smem_bytes(int nc, int T) {
  return (size_t)(T + 2 * nc) * 32 * sizeof(double);
}

// CNt 0..7 (sequenza.fit, R/bayes.R:40) and CNt 0..20 (sequenza.results, R/results.R:128).
// The pre-pass keeps the model BAF of 32 columns in shared memory: a shifted range with
// many rows per copy number (e.g. CNt 3..23) would not fit the default 48 KB and takes the
// split kernel instead.
bool baf_colwarp_supported(int nc, int CNt_min, int T, bool rp) {
  (void)CNt_min;
  return (nc == 8 || nc == 21) && !rp && colterm_cg_smem_bytes(nc, T) <= 48 * 1024;
}

size_t baf_colwarp_group_bytes(int nc, int64_t N_total) {
  return (size_t)nc * 32 * sizeof(double) * (size_t)(N_total > 0 ? N_total : 1);
}

template <int PT, int NC, int SB, int OCC, bool F32>
static int launch_colwarp(const BafGridArgs& a, const CnModel& m, int64_t N_total,
                          const double* cwg, int ic0, int nCb, int ub0, int nUb, double* partM,
                          long long* partE, unsigned long long* stats, cudaStream_t st) {
  const int64_t ctas = a.S * (int64_t)nUb * a.nChunk;
  if (ctas > 0x7fffffffLL) return fail(SQZ_ERR_SIZE, "grid too large for one launch");
  const size_t smb = colwarp_smem_bytes(NC, PT, SB, F32);
  SQZ_CUDA(cudaFuncSetAttribute(baf_grid_colwarp_kernel<PT, NC, SB, OCC, F32>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024));
  SQZ_CUDA(cudaFuncSetAttribute(baf_grid_colwarp_kernel<PT, NC, SB, OCC, F32>,
                                cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  baf_grid_colwarp_kernel<PT, NC, SB, OCC, F32><<<(unsigned)ctas, 32, smb, st>>>(
      a, m, cwg, ic0, nCb, ub0, nUb, N_total, partM, partE, stats);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

// Pre-pass + grid kernel over the span's column groups that hold units of this launch, as
// many 32-column groups per launch as fit cw_bytes.
int baf_colwarp_run(const BafGridArgs& a, const CnModel& m, int PT, int64_t N_total,
                    void* cw_buf, size_t cw_bytes, double* partM, long long* partE,
                    cudaStream_t st, unsigned long long* stats) {
  if (!baf_colwarp_supported(m.nc, m.CNt_min, m.T, a.rp != 0))
    return fail(SQZ_ERR_ARG, "column-per-lane kernel: unsupported model");
  if (PT != 8 && PT != 16 && PT != 32)
    return fail(SQZ_ERR_ARG, "column-per-lane kernel: ploidy tile must be 8, 16 or 32");
  int64_t groups = (int64_t)(cw_bytes / baf_colwarp_group_bytes(m.nc, N_total));
  if (groups < 1) return fail(SQZ_ERR_SIZE, "column-per-lane kernel needs room for 32 columns");
  if (groups > 2047) groups = 2047;
  double* cwg = (double*)cw_buf;
  const size_t pre_smem = colterm_cg_smem_bytes(m.nc, m.T);
  const int64_t u_end = (int64_t)a.u_first + a.nU;       // units of this launch: [u_first, u_end)
  const int64_t g_lo = a.u_first / a.nPT, g_hi = (u_end + a.nPT - 1) / a.nPT;  // groups touched
  for (int64_t g0 = g_lo; g0 < g_hi; g0 += groups) {
    const int64_t g1 = g0 + groups < g_hi ? g0 + groups : g_hi;
    const int64_t ic0 = g0 * 32;
    const int nCb = (int)((a.nCs - ic0 < (g1 - g0) * 32) ? (a.nCs - ic0) : (g1 - g0) * 32);
    // units of the batch, counted from its first group
    const int64_t ua = (a.u_first > g0 * a.nPT ? a.u_first : g0 * a.nPT) - g0 * a.nPT;
    const int64_t ub = (u_end < g1 * a.nPT ? u_end : g1 * a.nPT) - g0 * a.nPT;
    if (nCb <= 0 || ub <= ua) continue;
    dim3 grid((unsigned)((N_total + kPreSeg - 1) / kPreSeg), (unsigned)((nCb + 31) / 32));
    if (m.nc == 8 && m.CNt_min == 0 && m.CNn >= 2 && m.T == 20)
      baf_colterm_cg8_kernel<<<grid, 256, 0, st>>>(a, m, (int)ic0, nCb, N_total, cwg);
    else
      baf_colterm_cg_kernel<<<grid, 256, pre_smem, st>>>(a, m, (int)ic0, nCb, N_total,
                                                         m.nc <= 8 ? 1 : 0, cwg);
    SQZ_CUDA(cudaGetLastError());
    int rc;
#define SQZ_CW(PTV, NCV, SBV, OCCV)                                                     \
  rc = launch_colwarp<PTV, NCV, SBV, OCCV, (NCV <= 8)>(                                 \
      a, m, N_total, cwg, (int)ic0, nCb, (int)ua, (int)(ub - ua), partM, partE, stats, st)
    // (measured and dropped: two segments per stage, 36.4 ms against 34.4; twelve warps
    // per SM at 168 registers with two per stage, 51 ms -- the spills cost more than the
    // occupancy gains)
    if (m.nc == 8) {
      if (PT == 8) SQZ_CW(8, 8, 4, 8);
      else if (PT == 16) SQZ_CW(16, 8, 4, 8);
      else SQZ_CW(32, 8, 4, 8);
    } else {
      // 21 copy numbers: a stage is 5.25 KB per segment, one segment per stage; the narrow
      // tile needs 158 registers, so twelve warps fit an SM (measured: two segments per
      // stage and seven warps, 1.24-1.27 ms on config 2 against 1.08)
      if (PT == 8) SQZ_CW(8, 21, 1, 12);
      else if (PT == 16) SQZ_CW(16, 21, 1, 8);
      else SQZ_CW(32, 21, 1, 8);
    }
#undef SQZ_CW
    if (rc) return rc;
  }
  return SQZ_OK;
}

}  // namespace sqz
