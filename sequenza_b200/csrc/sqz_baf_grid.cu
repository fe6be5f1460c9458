// sqz_baf_grid.cu -- baf.model.fit grid search (R/bayes.R:103-116 over R/bayes.R:38-101)
// and baf.bayes at one grid point, hand-written for sm_100a.
//
// What the reference does per grid point g = (cellularity c, ploidy p) and segment s:
//   post[t] = dt((log mr[t] - log r_s)/sr_s, 5, log) + log prior[t]
//           + dt((mb[t] - Bf_s)/sb_s, 5, log)                     t = (CNt, B) rows
//   LPP[g] += max_t post[t]                                       (R/bayes.R:61-79, :112)
//
// Fast path (SQZ_MODE_FAST).  With df = 5,  dt(x, 5, log) = k5 - 3 log(1 + x^2/5), so
//   max_t post[t] = 2 k5 - 3 log( min_t (1 + xr_t^2/5) (1 + xb_t^2/5) prior_t^(-1/3) ).
// Three exact factorisations make this cheap:
//   (1) xr depends on the type only through CNt and xb does not depend on ploidy, so
//       min_t = min_CNt [ (1 + xr^2/5) * w_CNt * min_B (1 + xb^2/5) ] and the inner
//       min over B is shared by every ploidy of a cellularity column;
//   (2) sum_s log q_s = log prod_s q_s: the per-segment minima are multiplied into a
//       running product whose binary exponent is peeled off into an integer every four
//       segments, so the main loop has no log at all;
//   (3) the segment-size weights enter through sr_s = sd/sqrt(w) only.
// One CTA owns one cellularity, a tile of PT ploidies and one chunk of a sample's
// segments; every thread streams segments (two coalesced 16-byte loads each,
// prefetched one iteration ahead) and keeps PT running products in registers.
// Anything non-finite (NaN model points at cellularity 1, zero sd, overflow) makes the
// product Inf/NaN, which flags that grid point; flagged points are recomputed by the
// faithful kernel below, which walks the nmath dt() sequence type by type.
#include <stdlib.h>

#include <mutex>
#include <vector>

#include "sqz_common.cuh"
#include "sqz_baf_grid.cuh"
#include "../../include/sequenza_b200.h"

namespace sqz {

// ---- per-segment preparation for the fast path ------------------------------------
__global__ void baf_prep_kernel(BafGridArgs a, double2* rk, RkRec* rkx, double2* bk, int* nB) {
  const int64_t smp = blockIdx.y;
  const int64_t s0 = a.seg_off[smp], s1 = a.seg_off[smp + 1];
  const double inv_sqrt5 = 0.44721359549995793928;
  int cnt = 0;
  for (int64_t s = s0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < s1;
       s += (int64_t)gridDim.x * blockDim.x) {
    double sr = a.sdr[s] / sqrt(a.wr[s]);
    double kr = (1.0 / sr) * inv_sqrt5;
    double lr = log(a.ratio[s]);
    double Bf = a.Bf[s];
    double sb = a.sdBf[s] / sqrt(a.wBf[s]);
    bool hasB = !isnan(Bf) && !isnan(sb);  // R/bayes.R:68-69
    double kb = 0., eb = 0.;
    if (hasB) {
      kb = (1.0 / sb) * inv_sqrt5;
      eb = -(Bf * kb);
    }
    rk[s] = make_double2(kr, -(lr * kr));
    {
      const float kf = __double2float_rn(kr), ef = __double2float_rn(-(lr * kr));
      RkRec r;
      r.rk = make_double2(kr, -(lr * kr));
      r.f = make_float4(kf, ef, __fmul_ru(fabsf(ef), 0x1p-22f), 0.f);
      rkx[s] = r;
    }
    bk[s] = make_double2(kb, eb);
    cnt += hasB ? 1 : 0;
  }
  // block-wide count, one atomic per CTA
  __shared__ int sh;
  if (threadIdx.x == 0) sh = 0;
  __syncthreads();
  for (int off = 16; off; off >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(&sh, cnt);
  __syncthreads();
  if (threadIdx.x == 0 && sh) atomicAdd(&nB[smp], sh);
}

// ---- fast grid kernel ---------------------------------------------------------------
template <int PT, bool RP>
__global__ void __launch_bounds__(256, ((PT <= 16 && !RP) ? 2 : 1))
baf_grid_fast_kernel(BafGridArgs a, CnModel m, double* __restrict__ partM,
                     long long* __restrict__ partE) {
  extern __shared__ double sm[];
  double* lmr = sm;                 // [nc][PT] log model depth ratio
  double* mb = sm + m.nc * PT;      // [T] model BAF, type order
  __shared__ double redM[8][PT];
  __shared__ long long redE[8][PT];
  __shared__ unsigned redBad[8];

  int64_t b = blockIdx.x;
  const int u = a.u_first + (int)(b % a.nU);  // unit (column, tile) of the span
  b /= a.nU;
  const int chunk = (int)(b % a.nChunk);
  const int64_t smp = b / a.nChunk;
  const int ic = u / a.nPT, pt = u - ic * a.nPT;
  const int p0 = pt * PT;
  const int tid = threadIdx.x;

  // model tables for this (cellularity, ploidy tile): R/model.R:1-14 via R/model.R:49-56
  const double cell = a.cell[ic];
  const double adr = a.adr[smp];
  for (int i = tid; i < m.nc * PT; i += blockDim.x) {
    int c = i / PT, j = i - c * PT;
    int ip = min(p0 + j, a.nP - 1);  // ragged last tile repeats the last ploidy
    lmr[i] = log(theoretical_depth_ratio((double)(m.CNt_min + c), cell, a.ploidy[ip],
                                         (double)m.CNn, adr));
  }
  for (int t = tid; t < m.T; t += blockDim.x) {
    int c = 0;
    while (m.toff[c + 1] <= t) ++c;
    mb[t] = theoretical_baf((double)(m.CNt_min + c), (double)(t - m.toff[c]), cell,
                            (double)m.CNn);
  }
  __syncthreads();

  const int64_t s0 = a.seg_off[smp], s1 = a.seg_off[smp + 1];
  const int64_t per = (s1 - s0 + a.nChunk - 1) / a.nChunk;
  const int64_t cb = s0 + (int64_t)chunk * per;
  const int64_t ce = min(cb + per, s1);

  double P[PT];
  int E[PT];
#pragma unroll
  for (int j = 0; j < PT; ++j) {
    P[j] = 1.0;
    E[j] = 0;
  }
  unsigned bad = 0;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);

  auto renorm = [&]() {
#pragma unroll
    for (int j = 0; j < PT; ++j) {
      int hi = __double2hiint(P[j]), lo = __double2loint(P[j]);
      // not a positive normal number (Inf, NaN, zero, denormal, negative): flag
      if ((unsigned)(hi - 0x00100000) >= 0x7FE00000u) bad |= 1u << j;
      E[j] += (hi >> 20) - 1023;
      P[j] = __hiloint2double((hi & 0x000FFFFF) | 0x3FF00000, lo);
    }
  };

  int64_t s = cb + tid;
  double2 rk_n = make_double2(0., 0.), bk_n = rk_n;
  if (s < ce) {
    rk_n = a.rk[s];
    bk_n = a.bk[s];
  }
  int it = 0;
  while (s < ce) {
    const double2 rk = rk_n, bk = bk_n;
    const int64_t sn = s + blockDim.x;
    if (sn < ce) {  // prefetch the next segment of this thread
      rk_n = a.rk[sn];
      bk_n = a.bk[sn];
    }
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
    int t = 0;
    for (int c = 0; c < m.nc; ++c) {
      // min over B of |xb| (NaN never wins): R/bayes.R:70-72.  Four independent running
      // minima break the compare/select dependency chain; the signed value is kept and
      // |.| is folded into the compare, so no fp64 slot is spent on fabs.
      const int tend = m.toff[c + 1];
      double m0 = INF, m1 = INF, m2 = INF, m3 = INF;
      for (; t + 4 <= tend; t += 4) {
        const double x0 = fma(mb[t], bk.x, bk.y), x1 = fma(mb[t + 1], bk.x, bk.y);
        const double x2 = fma(mb[t + 2], bk.x, bk.y), x3 = fma(mb[t + 3], bk.x, bk.y);
        if (fabs(x0) < fabs(m0)) m0 = x0;
        if (fabs(x1) < fabs(m1)) m1 = x1;
        if (fabs(x2) < fabs(m2)) m2 = x2;
        if (fabs(x3) < fabs(m3)) m3 = x3;
      }
      if (t + 2 <= tend) {
        const double x0 = fma(mb[t], bk.x, bk.y), x1 = fma(mb[t + 1], bk.x, bk.y);
        if (fabs(x0) < fabs(m0)) m0 = x0;
        if (fabs(x1) < fabs(m1)) m1 = x1;
        t += 2;
      }
      if (t < tend) {
        const double x2 = fma(mb[t], bk.x, bk.y);
        if (fabs(x2) < fabs(m2)) m2 = x2;
        ++t;
      }
      if (fabs(m1) < fabs(m0)) m0 = m1;
      if (fabs(m3) < fabs(m2)) m2 = m3;
      if (fabs(m2) < fabs(m0)) m0 = m2;
      const double mabs = m0;
      const double u = fma(mabs, mabs, 1.0);  // 1 + xb^2/5
      const double wc = m.w[c];
      const double wub = u * wc;
      const double* l = lmr + c * PT;
#pragma unroll
      for (int j = 0; j < PT; ++j) {
        double x = fma(l[j], rk.x, rk.y);
        double tt = fma(x, x, 1.0);  // 1 + xr^2/5
        if (!RP) {
          double q = tt * wub;
          if (q < qmin[j]) qmin[j] = q;
        } else {
          // ratio.priority: CNt chosen on the ratio score alone (R/bayes.R:81-88)
          double ar = tt * wc;
          if (ar < qmin[j]) {
            q2[RP ? j : 0] = qmin[j];
            qmin[j] = ar;
            sel[RP ? j : 0] = u;
          } else if (ar < q2[RP ? j : 0]) {
            q2[RP ? j : 0] = ar;
          }
        }
      }
    }
    if (RP) {
      // near-tie between the two best ratio scores: the reference ranks them on
      // dt() + log(prior) (R/bayes.R:81), which may order them the other way -> repair
#pragma unroll
      for (int j = 0; j < (RP ? PT : 1); ++j)
        if (q2[j] <= qmin[j] * (1.0 + 1e-12)) bad |= 1u << j;
    }
#pragma unroll
    for (int j = 0; j < PT; ++j) P[j] *= RP ? qmin[j] * sel[RP ? j : 0] : qmin[j];
    if ((++it & 3) == 0) renorm();
    s = sn;
  }
  renorm();

  // CTA reduction: sum of log(mantissa) (each in [0, log 2)), exact integer exponents
  const int lane = tid & 31, warp = tid >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int j = 0; j < PT; ++j) {
    double v = log(P[j]);
    long long e = E[j];
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

// ---- finalize: LPP[g] = (n + nB) k5 - 3 (E ln2 + M) in double-double ----------------
// One thread per (sample, grid point of the span); points outside the launch's units are
// left alone.  The result goes to every destination in a.outs (this GPU's buffer and, for
// a multi-GPU fit, the same offset of each peer's buffer over NVLink).
__global__ void baf_grid_finalize_kernel(BafGridArgs a, const double* __restrict__ partM,
                                         const long long* __restrict__ partE,
                                         int* __restrict__ fix_list,
                                         int* __restrict__ fix_count) {
  const int64_t Gs = (int64_t)a.nCs * a.nP;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.S * Gs) return;
  const int64_t smp = i / Gs;
  const int64_t g = i - smp * Gs;
  const int ic = (int)(g / a.nP), ip = (int)(g - (int64_t)ic * a.nP);
  if (!in_units(a, ic, ip)) return;
  double M = 0.;
  long long E = 0;
  for (int ch = 0; ch < a.nChunk; ++ch) {
    M += partM[(int64_t)ch * a.S * Gs + i];
    E += partE[(int64_t)ch * a.S * Gs + i];
  }
  const double n = (double)(a.seg_off[smp + 1] - a.seg_off[smp]);
  const double Kn = n + (double)a.nB[smp];
  const double k5 = a.t5 - SQZ_LN_SQRT_2PI;
  const double ln2_hi = 0.693147180559945286, ln2_lo = 2.319046813846299558e-17;
  const double Ed = (double)E;
  // p1 + e1 = Kn * k5
  double p1 = Kn * k5, e1 = fma(Kn, k5, -p1);
  // L + eL = E * ln2
  double L = Ed * ln2_hi, eL = fma(Ed, ln2_hi, -L) + Ed * ln2_lo;
  // s + es = L + M
  double sL = L + M, bb = sL - L;
  double es = ((L - (sL - bb)) + (M - bb)) + eL;
  // p3 + e3 = 3 (s + es)
  double p3 = 3.0 * sL, e3 = fma(3.0, sL, -p3) + 3.0 * es;
  double r = p1 - p3, b2 = r - p1;
  double er = ((p1 - (r - b2)) + (-p3 - b2)) + (e1 - e3);
  double res = r + er;
  if (n == 0) res = 0.;
  store_result(a, smp, ic, ip, res);
  if (isnan(res)) {
    int k = atomicAdd(fix_count, 1);
    fix_list[k] = (int)i;
  }
}

// ---- faithful evaluation of one segment at one grid point (R/bayes.R:61-94) -------
// lratio[c] = log(model depth.ratio) per copy number, baf[t] per type row.
// Returns the winning copy-number index kc (-1: every posterior NaN) and B.
__device__ __forceinline__ void baf_eval_faithful(const CnModel& m, const double* lratio,
                                                  const double* baf, double Bf, double lr,
                                                  double sdBf, double sdr, double wBf,
                                                  double wr, int rp, double t5, int& kc,
                                                  int& kb, double& lpp) {
  const double sr = sdr / sqrt(wr);
  const double sb = sdBf / sqrt(wBf);
  const bool hasB = !isnan(Bf) && !isnan(sb);
  kc = -1;
  kb = 0;
  double best = 0.;
  if (!rp) {
    for (int c = 0; c < m.nc; ++c) {
      const double scr = dt5_log((lratio[c] - lr) / sr, t5) + m.lprior[c];
      for (int t = m.toff[c]; t < m.toff[c + 1]; ++t) {
        const double post = hasB ? scr + dt5_log((baf[t] - Bf) / sb, t5) : scr;
        if (!isnan(post) && (kc < 0 || post > best)) {  // which.max: first, NaN skipped
          best = post;
          kc = c;
          kb = t - m.toff[c];
        }
      }
    }
  } else {
    int kr = -1;
    double bests = 0.;
    for (int c = 0; c < m.nc; ++c) {
      const double scr = dt5_log((lratio[c] - lr) / sr, t5) + m.lprior[c];
      if (!isnan(scr) && (kr < 0 || scr > bests)) {
        bests = scr;
        kr = c;
      }
    }
    if (kr >= 0) {
      for (int t = m.toff[kr]; t < m.toff[kr + 1]; ++t) {
        const double post = hasB ? bests + dt5_log((baf[t] - Bf) / sb, t5) : bests;
        if (!isnan(post) && (kc < 0 || post > best)) {
          best = post;
          kc = kr;
          kb = t - m.toff[kr];
        }
      }
    }
  }
  lpp = (kc < 0) ? __longlong_as_double(0x7FF8000000000000LL) : best;
}

__device__ __forceinline__ void build_point_model(const CnModel& m, double cell,
                                                  double ploidy, double adr,
                                                  double* lratio, double* baf) {
  for (int c = threadIdx.x; c < m.nc; c += blockDim.x)
    lratio[c] = log(theoretical_depth_ratio((double)(m.CNt_min + c), cell, ploidy,
                                            (double)m.CNn, adr));
  for (int t = threadIdx.x; t < m.T; t += blockDim.x) {
    int c = 0;
    while (m.toff[c + 1] <= t) ++c;
    baf[t] = theoretical_baf((double)(m.CNt_min + c), (double)(t - m.toff[c]), cell,
                             (double)m.CNn);
  }
}

// Faithful grid kernel: persistent CTAs walk a list of (sample, grid point) items --
// either the points the fast path flagged (fix_list / fix_count) or, when fix_list is
// null, every point (SQZ_MODE_FAITHFUL).
__global__ void __launch_bounds__(256)
baf_grid_faithful_kernel(BafGridArgs a, CnModel m, const int* __restrict__ fix_list,
                         const int* __restrict__ fix_count) {
  extern __shared__ double sm[];
  double* lratio = sm;       // [nc]
  double* baf = sm + m.nc;   // [T]
  __shared__ double red[3][8];
  const int64_t Gs = (int64_t)a.nCs * a.nP;
  const int64_t n_items = fix_list ? (int64_t)*fix_count : a.S * Gs;
  for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int64_t i = fix_list ? (int64_t)fix_list[item] : item;
    const int64_t smp = i / Gs, g = i - smp * Gs;
    const int ic = (int)(g / a.nP), ip = (int)(g - (int64_t)ic * a.nP);
    if (!fix_list && !in_units(a, ic, ip)) continue;  // block-uniform
    __syncthreads();
    build_point_model(m, a.cell[ic], a.ploidy[ip], a.adr[smp], lratio, baf);
    __syncthreads();
    DD acc = dd_zero();
    for (int64_t s = a.seg_off[smp] + threadIdx.x; s < a.seg_off[smp + 1];
         s += blockDim.x) {
      int kc, kb;
      double lpp;
      baf_eval_faithful(m, lratio, baf, a.Bf[s], log(a.ratio[s]), a.sdBf[s], a.sdr[s],
                        a.wBf[s], a.wr[s], a.rp, a.t5, kc, kb, lpp);
      dd_add(acc, lpp);  // R sum(): no na.rm (R/bayes.R:112)
    }
    // CTA reduction in double-double, fixed order
    for (int off = 16; off; off >>= 1) {
      DD o;
      o.hi = __shfl_xor_sync(0xffffffffu, acc.hi, off);
      o.lo = __shfl_xor_sync(0xffffffffu, acc.lo, off);
      o.special = __shfl_xor_sync(0xffffffffu, acc.special, off);
      dd_merge(acc, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
      red[0][warp] = acc.hi;
      red[1][warp] = acc.lo;
      red[2][warp] = acc.special;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      DD tot = dd_zero();
      for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) {
        DD o{red[0][w], red[1][w], red[2][w]};
        dd_merge(tot, o);
      }
      store_result(a, smp, ic, ip, dd_value(tot));
    }
  }
}

// ---- baf.bayes at one grid point: one thread per segment, N x 4 column-major -----
__global__ void __launch_bounds__(256)
baf_point_kernel(int64_t N, const double* Bf, const double* ratio, const double* sdBf,
                 const double* sdr, const double* wBf, const double* wr, double cell,
                 double ploidy, double adr, CnModel m, int rp, double t5,
                 double* __restrict__ out) {
  extern __shared__ double sm[];
  double* lratio = sm;
  double* baf = sm + m.nc;
  build_point_model(m, cell, ploidy, adr, lratio, baf);
  __syncthreads();
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < N;
       s += (int64_t)gridDim.x * blockDim.x) {
    int kc, kb;
    double lpp;
    const double bf = Bf[s];
    baf_eval_faithful(m, lratio, baf, bf, log(ratio[s]), sdBf[s], sdr[s], wBf[s], wr[s],
                      rp, t5, kc, kb, lpp);
    double cn = na_real(), A = na_real(), B = na_real();
    if (kc >= 0) {
      cn = (double)(m.CNt_min + kc);
      B = (double)kb;
      A = cn - B;
      if (isnan(bf)) A = B = na_real();  // R/bayes.R:90-92
    }
    out[s] = cn;
    out[N + s] = A;
    out[2 * N + s] = B;
    out[3 * N + s] = lpp;
  }
}

// ---- host side ---------------------------------------------------------------------
int build_cn_model(int kind, int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                   const double* prior_value, int n_prior, CnModel* m, bool* fast_ok) {
  if (CNt_min < 0 || CNt_max < CNt_min) return fail(SQZ_ERR_ARG, "need 0 <= CNt.min <= CNt.max");
  if (CNt_max - CNt_min + 1 > kMaxCN)
    return fail(SQZ_ERR_SIZE, "CNt.max - CNt.min + 1 exceeds 64 copy numbers");
  if (CNn < 1) return fail(SQZ_ERR_ARG, "CNn must be >= 1");
  if (n_prior < 0 || (n_prior > 0 && (!prior_CN || !prior_value)))
    return fail(SQZ_ERR_ARG, "priors.table pointers missing");
  m->CNt_min = CNt_min;
  m->nc = CNt_max - CNt_min + 1;
  m->CNn = CNn;
  int T = 0;
  double v[kMaxCN];
  long double psum = 0.0L;
  bool ok = true;
  for (int c = 0; c < m->nc; ++c) {
    int cn = CNt_min + c;
    int rows = kind == 0 ? (CNn < 2 ? 1 : cn / 2 + 1) : cn + 1;
    m->toff[c] = T;
    T += rows;
    v[c] = 1.0;
    for (int i = 0; i < n_prior; ++i)
      if ((double)cn == prior_CN[i]) v[c] = prior_value[i];  // R/bayes.R:56-58
    psum += (long double)v[c] * rows;
  }
  m->toff[m->nc] = T;
  m->T = T;
  for (int c = m->nc + 1; c <= kMaxCN; ++c) m->toff[c] = T;
  for (int c = 0; c < kMaxCN; ++c) {
    if (c < m->nc) {
      m->prior[c] = v[c] / (double)psum;  // R/bayes.R:59
      m->lprior[c] = log(m->prior[c]);
      m->w[c] = cbrt(1.0 / m->prior[c]);
      if (!(m->prior[c] >= 0.0) || !isfinite(m->prior[c])) ok = false;
    } else {
      m->prior[c] = m->lprior[c] = m->w[c] = 0.;
    }
  }
  if (fast_ok) *fast_ok = ok && CNn >= 2;
  return SQZ_OK;
}

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// ---- optional per-kernel timing (sqz_profile_*): CUDA events recorded on the launch
// stream around the three phases of a fast-mode grid call, read back after the caller
// has synchronised.  Off by default; bench.py turns it on for the roofline numbers.
// All of this state is guarded by g_prof_mu: grid calls may come from several host
// threads (one per device in the host entry points).
struct ProfSet {
  cudaEvent_t ev[4];
  int path = 0;                    // 0 fused, 1 split, 2 pruned (segment lanes), 3 pruned (column lanes)
  unsigned long long n_eval = 0;   // (segment, tile, copy number) evaluations, pruned path
  int launches = 0;                // kernels launched by this call
};
static std::mutex g_prof_mu;
static bool g_prof_on = false;
static int g_prof_dev = -1;
static int g_prof_n = 0;
static const int kProfMax = 4096;
static std::vector<ProfSet>* g_prof = nullptr;

int profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = on != 0;
  g_prof_n = 0;
  return SQZ_OK;
}

static ProfSet* prof_next() {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (!g_prof_on) return nullptr;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  if (!g_prof) {
    g_prof = new std::vector<ProfSet>();
    g_prof->reserve(kProfMax);  // pointers handed out below stay valid
  }
  if (dev != g_prof_dev) {  // events belong to a device: start over on a device change
    for (auto& p : *g_prof)
      for (int k = 0; k < 4; ++k) cudaEventDestroy(p.ev[k]);
    g_prof->clear();
    g_prof_dev = dev;
    g_prof_n = 0;
  }
  if (g_prof_n >= kProfMax) return nullptr;
  if ((int)g_prof->size() <= g_prof_n) {
    ProfSet p;
    for (int k = 0; k < 4; ++k)
      if (cudaEventCreate(&p.ev[k]) != cudaSuccess) return nullptr;
    g_prof->push_back(p);
  }
  ProfSet* r = &(*g_prof)[g_prof_n++];
  r->path = 0;
  r->n_eval = 0;
  r->launches = 0;
  return r;
}

// kernel path, candidate-evaluation count and kernel-launch count of the calls recorded
// since profile_enable(1) (read before profile_collect resets them; the caller has
// synchronised)
int profile_counters(int* path, unsigned long long* n_eval, long long* launches) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  int p = 0;
  unsigned long long n = 0;
  long long l = 0;
  if (g_prof)
    for (int i = 0; i < g_prof_n; ++i) {
      p = (*g_prof)[i].path;
      n += (*g_prof)[i].n_eval;
      l += (*g_prof)[i].launches;
    }
  if (path) *path = p;
  if (n_eval) *n_eval = n;
  if (launches) *launches = l;
  return SQZ_OK;
}

// sums (ms) of prep / fast kernel / finalize+repair over the calls recorded since
// profile_enable(1); the caller must have synchronised the stream(s).
int profile_collect(double* ms3, int* n_calls) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  double acc[3] = {0, 0, 0};
  int n = 0;
  if (g_prof)
    for (int i = 0; i < g_prof_n; ++i) {
      float a = 0, b = 0, c = 0;
      ProfSet& p = (*g_prof)[i];
      if (cudaEventElapsedTime(&a, p.ev[0], p.ev[1]) != cudaSuccess ||
          cudaEventElapsedTime(&b, p.ev[1], p.ev[2]) != cudaSuccess ||
          cudaEventElapsedTime(&c, p.ev[2], p.ev[3]) != cudaSuccess) {
        cudaGetLastError();
        continue;
      }
      acc[0] += a;
      acc[1] += b;
      acc[2] += c;
      ++n;
    }
  g_prof_n = 0;
  if (ms3) {
    ms3[0] = acc[0];
    ms3[1] = acc[1];
    ms3[2] = acc[2];
  }
  if (n_calls) *n_calls = n;
  return SQZ_OK;
}

// ---- planning -----------------------------------------------------------------------
// Everything here is a function of the WHOLE grid (nC x nP), the segment table (N_total,
// S) and the model -- never of the shard a launch works on -- so that a grid point gets
// the same bits from any shard and any GPU count (see sqz_baf_grid.cuh).
struct GridPlan {
  int family;  // 0 fused, 1 split, 2 pruned (segment per lane), 3 pruned (column per lane)
  int PT, nPT, nChunk, threads;
  int W;       // columns per unit: 32 for family 3, else 1
};

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

// Fraction of SM-time idle in the last wave when `ctas` CTAs run `per_sm` at a time.
static double tail_waste(int64_t ctas, int per_sm) {
  const int64_t slots = 148LL * per_sm;
  const int64_t waves = (ctas + slots - 1) / slots;
  return 1.0 - (double)ctas / (double)(waves * slots);
}

// Candidate pruning (baf_grid_pruned_kernel) is used for few copy numbers and a sorted
// ploidy axis.  Measured on B200: 8 copy numbers -> 1.5-2.4x over the split kernel on fine
// and medium grids for any sd scale (tools/worstcase.py); 21 copy numbers -> no gain.
static int prune_mode(int nc, bool monotone, bool rp) {
  const int e = env_int("SQZ_PRUNE", -1);  // 0/1 override (experiments only)
  if (!monotone || rp) return 0;
  return e < 0 ? (nc <= 12 ? 1 : 0) : (e != 0);
}

// Chunk count of the column-per-lane kernel.  It follows the segment table only.  Every
// warp of that kernel walks all segments of its chunk: at least 64 per chunk, at most 64
// chunks (128 for one long table), fewer for a cohort (the samples already fill the machine, and
// every chunk costs a slab of partial results).
constexpr int64_t kColwarpMaxChunks = 128;
static int64_t colwarp_chunks(int64_t N_total, int64_t S) {
  const int64_t avg = S > 0 ? (N_total + S - 1) / S : 0;
  int64_t top = S > 0 ? 2048 / S : 64;
  if (top > 64) top = 64;
  int64_t k = avg / 64;
  if (k > top) k = top;
  // One long table: up to 128 chunks of 512 segments or more.  An eighth of the fine grid
  // (one GPU of eight) is then 7.5 waves of one-warp CTAs instead of 3.7, and the idle tail
  // at the end of the kernel shrinks with the CTAs: measured on config 3, 64 -> 128 chunks
  // costs 1.8 % on one GPU (32.9 -> 33.5 ms) and gains 4.4 % on the slowest of eight
  // shards (4.96 -> 4.74 ms).  (SQZ_CW_MAXCHUNKS: experiments only.)
  const int64_t cap = env_int("SQZ_CW_MAXCHUNKS", (int)kColwarpMaxChunks);
  if (S == 1 && avg / 512 > 64) k = avg / 512 < cap ? avg / 512 : cap;
  return k < 1 ? 1 : k;
}

constexpr int64_t kNarrowMaxChunks = 128;
// Chunk count of the column-per-lane kernel with 8-wide tiles (twelve one-warp CTAs per SM):
// the units of the WHOLE grid times the chunks should fill the machine in whole waves.
static int64_t narrow_chunks(int64_t N_total, int64_t S, int64_t units_total) {
  const int64_t avg = S > 0 ? (N_total + S - 1) / S : 0;
  int64_t cap = avg / 128;
  if (cap > kNarrowMaxChunks) cap = kNarrowMaxChunks;
  if (S > 1 && cap > 2048 / S) cap = 2048 / S;
  if (cap < 1) cap = 1;
  const int64_t slots = 148 * 12, ctas = (units_total > 0 ? units_total : 1) * (S > 0 ? S : 1);
  int64_t best = slots / ctas;  // one wave
  for (int64_t k = 2; k <= 4; ++k)
    if (k * slots / ctas <= cap) best = k * slots / ctas;
  if (best > cap) best = cap;
  return best < 1 ? 1 : best;
}

// Largest chunk count either planner may pick; the partial buffers are sized for it.
// Segment-per-lane kernels keep >= 8 segments per thread of a 64-thread CTA.
static int64_t max_chunks(int64_t N_total, int64_t S) {
  const int64_t avg = S > 0 ? (N_total + S - 1) / S : 0;
  int64_t a = avg / (64 * 8);
  if (a > 64) a = 64;
  const int64_t b = colwarp_chunks(N_total, S);
  if (b > a) a = b;
  // the narrow-tile column-per-lane plan (CNt 0..20 on a coarse ploidy axis) fills the
  // machine with chunks of 128 segments or more, at most kNarrowMaxChunks of them
  int64_t c = avg / 128;
  if (c > kNarrowMaxChunks) c = kNarrowMaxChunks;
  if (S > 1 && c > 2048 / S) c = 2048 / S;
  if (c > a) a = c;
  const int64_t e = env_int("SQZ_CW_MAXCHUNKS", 0);  // experiments: room for SQZ_CHUNKS
  if (e > a && e <= 512 && S == 1 && avg / e >= 32) a = e;
  return a < 1 ? 1 : a;
}

// Ploidy tile of the column-per-lane kernel, or 0 when a segment-per-lane kernel is the
// better choice for this grid.  The column-per-lane kernel runs one warp per (32 columns,
// tile, chunk) and there are 8 warp slots per SM: it wins when those units fill the
// machine for several waves (measured: tools/colwarp_vs_pruned.py), and when few lanes of
// the last 32-column group idle.  32-wide tiles when there are enough units, else 16-wide.
static int colwarp_tile(const CnModel& m, bool rp, int64_t N_total, int64_t S, int64_t nP,
                        int64_t nC) {
  const int e = env_int("SQZ_COLWARP", -1);  // 0/1 override (experiments and tests only)
  if (e == 0 || !baf_colwarp_supported(m.nc, m.CNt_min, m.T, rp) || N_total <= 0) return 0;
  const int pt_env = env_int("SQZ_PT", 0);
  if (e == 1) return (pt_env == 8 || pt_env == 16 || pt_env == 32) ? pt_env : (nP <= 96 ? 16 : 32);
  const int64_t groups = (nC + 31) / 32;
  if (nC * 10 < groups * 32 * 6) return 0;
  const int64_t slots = 148 * 8, ch = colwarp_chunks(N_total, S);
  const int64_t u32 = S * groups * ((nP + 31) / 32) * ch, u16 = S * groups * ((nP + 15) / 16) * ch;
  if (m.nc > 8) {
    // CNt 0..20.  Fine ploidy axes: 16-wide tiles.  Coarse ones (the default 91 x 61 grid):
    // 8-wide tiles keep about three copy numbers alive per warp, and with one segment per
    // stage twelve warps fit an SM -- 1.08 ms against the split kernel's 1.40 on config 2
    // (1.30 / 1.40 with every sd x 0.1, where most candidates survive); short tables stay
    // with the split kernel.
    if (nP >= 200 && u16 >= 2 * slots) return 16;
    const int64_t avg = S > 0 ? (N_total + S - 1) / S : 0;
    return (avg >= 4096 && nP >= 16) ? 8 : 0;
  }
  // 32-wide tiles need fewer instructions per evaluation; 16-wide ones give twice the units
  // and a cheaper per-chunk prologue.  Measured (tools/colwarp_vs_pruned.py,
  // tools/shard_tile_width.py): 32 wins from 6 waves of units on, and from 2 waves on when
  // the chunks are long.
  const int64_t per_chunk = (S > 0 ? (N_total + S - 1) / S : 0) / ch;
  if (nP > 96 && (u32 >= 6 * slots || (u32 >= 2 * slots && per_chunk >= 500))) return 32;
  if (u16 >= 3 * slots) return 16;
  return 0;
}

static GridPlan plan_fit(int64_t N_total, int64_t S, int64_t nP, int64_t nC, const CnModel& m,
                         bool rp, bool monotone, bool allow_colwarp = true) {
  GridPlan g;
  const int cw_pt = (allow_colwarp && monotone && env_int("SQZ_PRUNE", 1) != 0)
                        ? colwarp_tile(m, rp, N_total, S, nP, nC) : 0;
  if (cw_pt) {
    g.family = 3;
    g.W = 32;
    g.PT = cw_pt;
    g.nPT = (int)((nP + g.PT - 1) / g.PT);
    g.threads = 32;  // one warp per CTA
    g.nChunk = (int)(g.PT == 8 && m.nc > 8
                         ? narrow_chunks(N_total, S, ((nC + 31) / 32) * g.nPT)
                         : colwarp_chunks(N_total, S));
    const int ch_env = env_int("SQZ_CHUNKS", 0);
    if (ch_env >= 1 && ch_env <= max_chunks(N_total, S)) g.nChunk = ch_env;
    return g;
  }
  const bool pruned = prune_mode(m.nc, monotone, rp) != 0;
  g.family = pruned ? 2 : 1;
  g.W = 1;
  const int cand[3] = {8, 16, 32};
  double bestc = 0;
  g.PT = 32;
  for (int k = 0; k < (rp ? 2 : 3); ++k) {  // PT = 32 spills registers with ratio.priority
    int pt = cand[k];
    double cost = (double)((nP + pt - 1) / pt) * (pt * (4.0 * m.nc + 1) + 2.0 * m.T + 2.0 * m.nc);
    if (k == 0 || cost < bestc) {
      bestc = cost;
      g.PT = pt;
    }
  }
  // with pruning and a coarse ploidy axis (a 32-wide tile would span most of it and keep
  // every copy number alive) 16-wide tiles are faster: default 91 x 61 grid, 0.57 vs 0.65 ms
  if (pruned && nP <= 96 && g.PT == 32) g.PT = 16;
  const int pt_env = env_int("SQZ_PT", 0);  // tuning overrides (experiments only)
  if (pt_env == 8 || pt_env == 16 || (pt_env == 32 && !rp)) g.PT = pt_env;
  g.nPT = (int)((nP + g.PT - 1) / g.PT);
  const int64_t avg = S > 0 ? (N_total + S - 1) / S : 0;
  // 128-thread CTAs: two resident per SM even at 255 registers, finer tail granularity
  g.threads = avg > 64 ? 128 : 64;
  const int th_env = env_int("SQZ_THREADS", 0);
  if (th_env == 64 || th_env == 128 || th_env == 256) g.threads = th_env;
  // resident CTAs per SM from the register budget of each instantiation
  const int regs = (g.PT == 32 || rp) ? 256 : 128;
  int per_sm = 65536 / (regs * g.threads);
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  // Split each sample's segments into chunks until the whole grid fills the machine for a
  // few waves, then pick the chunk count with the smallest idle tail in that range.
  const int64_t base = S * nC * g.nPT;
  int64_t cap = avg / ((int64_t)g.threads * 8);  // keep >= 8 segments per thread
  if (cap < 1) cap = 1;
  if (cap > 64) cap = 64;
  int64_t lo = base > 0 ? (148LL * per_sm * 3 + base - 1) / base : 1;
  if (lo > cap) lo = cap;
  if (lo < 1) lo = 1;
  int64_t hi = lo * 2 + 2;
  if (hi > cap) hi = cap;
  int64_t best = lo;
  double bw = 2.0;
  for (int64_t k = lo; k <= hi; ++k) {
    double wst = tail_waste(base * k, per_sm) + 0.002 * (double)(k - lo);  // mild bias to fewer
    if (wst < bw) {
      bw = wst;
      best = k;
    }
  }
  g.nChunk = (int)best;
  const int ch_env = env_int("SQZ_CHUNKS", 0);
  if (ch_env >= 1 && ch_env <= max_chunks(N_total, S)) g.nChunk = ch_env;
  return g;
}

static bool ploidy_monotone(const double* h_ploidy, int64_t nP) {
  if (!h_ploidy) return true;
  bool up = true, dn = true;
  for (int64_t i = 1; i < nP; ++i) {
    if (!(h_ploidy[i] >= h_ploidy[i - 1])) up = false;
    if (!(h_ploidy[i] <= h_ploidy[i - 1])) dn = false;
  }
  return up || dn;
}

// the most samples one batch of launches may hold: (sample, grid point) items are indexed
// with 32-bit integers in the repair list
static int64_t samples_per_batch(int64_t S, int64_t nP, int64_t nCs) {
  const int64_t per = nCs * nP > 0 ? nCs * nP : 1;
  int64_t sb = 0x7fffffffLL / per;
  if (sb > 65535) sb = 65535;  // gridDim.y of the preparation kernel
  return sb < S ? sb : S;
}

struct Workspace {
  double2 *rk, *bk;
  RkRec* rkx;
  int* nB;
  double *ploidy, *cell;
  double* partM;
  long long* partE;
  int *fix_list, *fix_count;
  unsigned long long* stats;  // candidate evaluations of the pruned kernels (profiling only)
  size_t bytes;
};

// nCs = columns of the span a launch works on
static Workspace carve(void* base, int64_t N_total, int64_t S, int64_t nP, int64_t nCs) {
  // partials are sized for the largest chunk count the planner can pick
  Workspace w;
  size_t off = 0;
  char* p = (char*)base;
  auto take = [&](size_t bytes) {
    void* r = p ? (void*)(p + off) : nullptr;
    off += align256(bytes);
    return r;
  };
  const size_t n = (size_t)(N_total > 0 ? N_total : 1);
  const int64_t Sb = samples_per_batch(S, nP, nCs);
  const size_t items = (size_t)(Sb * nCs * nP > 0 ? Sb * nCs * nP : 1);
  const int64_t want = max_chunks(N_total, S);
  w.rk = (double2*)take(n * sizeof(double2));
  w.bk = (double2*)take(n * sizeof(double2));
  w.rkx = (RkRec*)take(n * sizeof(RkRec));
  w.nB = (int*)take((size_t)(S > 0 ? S : 1) * sizeof(int));
  w.ploidy = (double*)take((size_t)(nP > 0 ? nP : 1) * sizeof(double));
  w.cell = (double*)take((size_t)(nCs > 0 ? nCs : 1) * sizeof(double));
  w.partM = (double*)take(items * (size_t)want * sizeof(double));
  w.partE = (long long*)take(items * (size_t)want * sizeof(long long));
  w.fix_list = (int*)take(items * sizeof(int));
  w.fix_count = (int*)take(sizeof(int));
  w.stats = (unsigned long long*)take(sizeof(unsigned long long));
  w.bytes = off;
  return w;
}

// plan of a fast-mode call, for reporting (bench.py roofline) and for sharding by units:
// out8 = {PT, nPT, chunks, threads, T, kernel family, W, units of the whole grid}
int baf_grid_plan(int64_t N_total, int64_t S, const double* h_ploidy, int64_t nP, int64_t nC,
                  int CNt_min, int CNt_max, int CNn, int ratio_priority, int64_t* out8) {
  CnModel m;
  bool ok;
  int rc = build_cn_model(0, CNt_min, CNt_max, CNn, nullptr, nullptr, 0, &m, &ok);
  if (rc) return rc;
  const GridPlan g = plan_fit(N_total, S, nP, nC, m, ratio_priority != 0,
                              ploidy_monotone(h_ploidy, nP));
  out8[0] = g.PT;
  out8[1] = g.nPT;
  out8[2] = g.nChunk;
  out8[3] = g.threads;
  out8[4] = m.T;
  out8[5] = g.family;
  out8[6] = g.W;
  out8[7] = ((nC + g.W - 1) / g.W) * g.nPT;
  return SQZ_OK;
}

// columns [span0, span1) that the units [u0, u1) of the whole grid touch
static void unit_span(const GridPlan& g, int64_t nC, int64_t u0, int64_t u1, int64_t* span0,
                      int64_t* span1) {
  const int64_t g0 = u0 / g.nPT, g1 = (u1 + g.nPT - 1) / g.nPT;
  *span0 = g0 * g.W;
  *span1 = g1 * g.W < nC ? g1 * g.W : nC;
}

size_t baf_grid_workspace_bytes(int64_t N_total, int64_t S, int64_t nP, int64_t nCs) {
  return carve(nullptr, N_total, S, nP, nCs).bytes;
}

// minimum + room for the pre-pass output of as many of the span's columns as fit under
// `cap_bytes`; the kernel family follows the whole grid (nC columns)
size_t baf_grid_workspace_preferred_bytes(int64_t N_total, int64_t S, int64_t nP, int64_t nC,
                                          int64_t nCs, int n_copy_numbers, size_t cap_bytes) {
  const size_t base = carve(nullptr, N_total, S, nP, nCs).bytes;
  if (n_copy_numbers < 1 || n_copy_numbers > kMaxCN || nCs < 1) return base;
  if (nC < nCs) nC = nCs;
  const size_t per_col = baf_split_column_bytes(n_copy_numbers, N_total);
  size_t cols = (size_t)nCs;
  CnModel m;
  bool ok;
  if (build_cn_model(0, 0, n_copy_numbers - 1, 2, nullptr, nullptr, 0, &m, &ok) == SQZ_OK &&
      colwarp_tile(m, false, N_total, S, nP, nC)) {
    // the column-per-lane kernel would take this fit: 8 bytes per (column, copy number,
    // segment) in whole 32-column groups.  (A call that cannot use it after all --
    // ratio.priority, unsorted ploidies -- runs the split kernel in batches.)
    size_t want = ((cols + 31) / 32) * baf_colwarp_group_bytes(n_copy_numbers, N_total);
    if (want < per_col) want = per_col;
    if (cap_bytes > 0 && want > cap_bytes) want = cap_bytes;
    return base + align256(want);
  }
  if (cap_bytes > 0 && cols * per_col > cap_bytes) cols = cap_bytes / per_col;
  if (cols < 1) cols = 1;
  return base + align256(cols * per_col);
}

// the same for a shard given as units [u0, u1) of the whole grid
size_t baf_grid_units_workspace_bytes(int64_t N_total, int64_t S, const double* h_ploidy,
                                      int64_t nP, int64_t nC, int CNt_min, int CNt_max, int CNn,
                                      int ratio_priority, int64_t u0, int64_t u1,
                                      size_t cap_bytes) {
  CnModel m;
  bool ok;
  if (build_cn_model(0, CNt_min, CNt_max, CNn, nullptr, nullptr, 0, &m, &ok) != SQZ_OK) return 0;
  const GridPlan g = plan_fit(N_total, S, nP, nC, m, ratio_priority != 0,
                              ploidy_monotone(h_ploidy, nP));
  if (u0 < 0 || u1 < u0 || u1 > ((nC + g.W - 1) / g.W) * g.nPT) return 0;
  if (u0 == u1) return carve(nullptr, N_total, S, nP, 1).bytes;
  int64_t span0, span1;
  unit_span(g, nC, u0, u1, &span0, &span1);
  return baf_grid_workspace_preferred_bytes(N_total, S, nP, nC, span1 - span0, m.nc, cap_bytes);
}

// ---- launch ---------------------------------------------------------------------------
// What part of the grid a call computes and where the results go.
struct GridShard {
  bool by_units;     // false: columns [a0, a1) -> out[S][(a1 - a0) * nP], shard-relative
                     // true : units   [a0, a1) -> out[S][nC * nP], whole-grid layout
  int64_t a0, a1;
  int n_out;
  double* outs[kMaxOut];
};

static int baf_grid_core(int64_t S, const int64_t* d_seg_off, int64_t N_total,
                         const double* d_Bf, const double* d_ratio, const double* d_sdBf,
                         const double* d_sdr, const double* d_wBf, const double* d_wr,
                         const double* h_ploidy, int64_t nP, const double* h_cell, int64_t nC,
                         const GridShard& sh, const double* d_adr, int CNt_min, int CNt_max,
                         int CNn, const double* prior_CN, const double* prior_value,
                         int n_prior, int ratio_priority, int mode, void* workspace,
                         size_t workspace_bytes, cudaStream_t st, bool allow_profile) {
  if (S < 0 || N_total < 0 || nP < 0 || nC < 0)
    return fail(SQZ_ERR_ARG, "negative size");
  if (mode != SQZ_MODE_FAST && mode != SQZ_MODE_FAITHFUL)
    return fail(SQZ_ERR_ARG, "mode must be SQZ_MODE_FAST or SQZ_MODE_FAITHFUL");
  if (sh.n_out < 1 || sh.n_out > kMaxOut)
    return fail(SQZ_ERR_ARG, "between 1 and 8 output buffers");
  CnModel m;
  bool fast_ok;
  int rc = build_cn_model(0, CNt_min, CNt_max, CNn, prior_CN, prior_value, n_prior, &m,
                          &fast_ok);
  if (rc) return rc;
  const bool rp = ratio_priority != 0;
  // candidate pruning relies on the residual being monotone over a ploidy tile
  const bool mono = ploidy_monotone(h_ploidy, nP);
  GridPlan g = plan_fit(N_total, S, nP, nC, m, rp, mono);

  // the span of columns this call works on, and its units
  int64_t span0, span1, u_first, nU;
  if (!sh.by_units) {
    if (sh.a0 < 0 || sh.a1 > nC || sh.a0 > sh.a1)
      return fail(SQZ_ERR_ARG, "bad cellularity shard");
    span0 = sh.a0;
    span1 = sh.a1;
    u_first = 0;
    nU = ((span1 - span0 + g.W - 1) / g.W) * g.nPT;
  } else {
    const int64_t n_units = ((nC + g.W - 1) / g.W) * g.nPT;
    if (sh.a0 < 0 || sh.a1 > n_units || sh.a0 > sh.a1)
      return fail(SQZ_ERR_ARG, "bad unit shard (see sqz_baf_grid_plan for the unit count)");
    unit_span(g, nC, sh.a0, sh.a1, &span0, &span1);
    u_first = sh.a0 - (sh.a0 / g.nPT) * g.nPT;
    nU = sh.a1 - sh.a0;
  }
  const int64_t nCs = span1 - span0;
  if (S == 0 || nCs <= 0 || nP == 0 || nU == 0) return SQZ_OK;
  if (nCs * nP > 0x7fffffffLL)
    return fail(SQZ_ERR_SIZE, "more than 2^31-1 grid points in one shard");
  if (S > 65535 * 64LL) return fail(SQZ_ERR_SIZE, "more than 4,194,240 samples in one call");
  for (int k = 0; k < sh.n_out; ++k)
    if (!sh.outs[k]) return fail(SQZ_ERR_ARG, "an output buffer is null");
  if (!workspace || ((uintptr_t)workspace & 255))
    return fail(SQZ_ERR_ARG, "workspace must be non-null and 256-byte aligned");
  Workspace w = carve(workspace, N_total, S, nP, nCs);
  if (w.bytes > workspace_bytes) return fail(SQZ_ERR_SIZE, "workspace too small");

  SQZ_CUDA(cudaMemcpyAsync(w.ploidy, h_ploidy, nP * sizeof(double), cudaMemcpyHostToDevice, st));
  SQZ_CUDA(cudaMemcpyAsync(w.cell, h_cell + span0, nCs * sizeof(double), cudaMemcpyHostToDevice, st));

  // pre-pass scratch behind the fixed part of the workspace
  char* cw_buf = (char*)workspace + w.bytes;
  const size_t cw_bytes = workspace_bytes > w.bytes ? workspace_bytes - w.bytes : 0;
  if (g.family == 3 && cw_bytes < baf_colwarp_group_bytes(m.nc, N_total)) {
    // no room for one 32-column group of pre-pass output: segment-per-lane kernels
    g = plan_fit(N_total, S, nP, nC, m, rp, mono, false);
    if (sh.by_units) return fail(SQZ_ERR_SIZE, "workspace too small for a unit shard of the "
                                               "column-per-lane kernel");
    nU = nCs * g.nPT;
  }
  const bool use_split = g.family == 3 ||
                         (env_int("SQZ_SPLIT", 1) != 0 && N_total > 0 &&
                          baf_split_supported(m.nc, g.PT, g.threads) &&
                          cw_bytes >= baf_split_column_bytes(m.nc, N_total));
  if (!use_split) g.family = 0;

  BafGridArgs a;
  a.seg_off = d_seg_off;
  a.Bf = d_Bf;
  a.ratio = d_ratio;
  a.sdBf = d_sdBf;
  a.sdr = d_sdr;
  a.wBf = d_wBf;
  a.wr = d_wr;
  a.rk = w.rk;
  a.rkx = w.rkx;
  a.bk = w.bk;
  a.ploidy = w.ploidy;
  a.cell = w.cell;
  a.nP = (int)nP;
  a.nCs = (int)nCs;
  a.nPT = g.nPT;
  a.nChunk = g.nChunk;
  a.PT = g.PT;
  a.W = g.W;
  a.u_first = (int)u_first;
  a.nU = (int)nU;
  a.rp = rp ? 1 : 0;
  a.t5 = dt5_const();
  a.out_sS = sh.by_units ? nC * nP : nCs * nP;
  a.out_col0 = sh.by_units ? span0 : 0;
  a.n_out = sh.n_out;

  const size_t sm_faith = (size_t)(m.nc + m.T) * sizeof(double);
  const int faithful_ctas = 148 * 4;
  const bool fast = mode == SQZ_MODE_FAST && fast_ok;
  ProfSet* prof = (fast && allow_profile) ? prof_next() : nullptr;
  int launches = 0;

  if (fast) {
    SQZ_CUDA(cudaMemsetAsync(w.nB, 0, S * sizeof(int), st));
    if (prof) {
      SQZ_CUDA(cudaMemsetAsync(w.stats, 0, sizeof(unsigned long long), st));
      SQZ_CUDA(cudaEventRecord(prof->ev[0], st));
    }
    // per-segment preparation, all samples (gridDim.y holds at most 65535 of them)
    const int64_t avg = (N_total + S - 1) / S;
    int bx = (int)((avg + 255) / 256);
    if (bx < 1) bx = 1;
    if (bx > 1024) bx = 1024;
    for (int64_t sa = 0; sa < S; sa += 65535) {
      const int64_t sb = sa + 65535 < S ? sa + 65535 : S;
      BafGridArgs ap = a;
      ap.S = sb - sa;
      ap.seg_off = d_seg_off + sa;
      dim3 grid(bx, (unsigned)(sb - sa));
      baf_prep_kernel<<<grid, 256, 0, st>>>(ap, w.rk, w.rkx, w.bk, w.nB + sa);
      SQZ_CUDA(cudaGetLastError());
      ++launches;
    }
    if (prof) SQZ_CUDA(cudaEventRecord(prof->ev[1], st));
  }

  // batches of samples: (sample, grid point) items are indexed with 32-bit integers
  const int64_t Sb = samples_per_batch(S, nP, nCs);
  for (int64_t sa = 0; sa < S; sa += Sb) {
    const int64_t sb = sa + Sb < S ? sa + Sb : S;
    a.S = sb - sa;
    a.seg_off = d_seg_off + sa;
    a.adr = d_adr + sa;
    a.nB = w.nB + sa;
    for (int k = 0; k < sh.n_out; ++k) a.outs[k] = sh.outs[k] + sa * a.out_sS;
    if (!fast) {
      baf_grid_faithful_kernel<<<faithful_ctas, 256, sm_faith, st>>>(a, m, nullptr, nullptr);
      SQZ_CUDA(cudaGetLastError());
      continue;
    }
    SQZ_CUDA(cudaMemsetAsync(w.fix_count, 0, sizeof(int), st));
    if (use_split) {
      rc = g.family == 3
               ? baf_colwarp_run(a, m, g.PT, N_total, cw_buf, cw_bytes, w.partM, w.partE, st,
                                 prof ? w.stats : nullptr)
               : baf_split_run(a, m, g.PT, g.threads, N_total, cw_buf, cw_bytes, w.partM,
                               w.partE, st, g.family == 2, prof ? w.stats : nullptr);
      if (rc) return rc;
      launches += 2;  // pre-pass + grid kernel (per batch of columns; one batch when the
                      // preferred workspace is given)
    } else {
      const int64_t ctas = a.S * nU * g.nChunk;
      if (ctas > 0x7fffffffLL) return fail(SQZ_ERR_SIZE, "grid too large for one launch");
      const size_t smb = (size_t)(m.nc * g.PT + m.T) * sizeof(double);
#define SQZ_LAUNCH_FAST(PTV)                                                              \
  do {                                                                                    \
    if (rp)                                                                               \
      baf_grid_fast_kernel<PTV, true><<<(unsigned)ctas, g.threads, smb, st>>>(a, m, w.partM, w.partE); \
    else                                                                                  \
      baf_grid_fast_kernel<PTV, false><<<(unsigned)ctas, g.threads, smb, st>>>(a, m, w.partM, w.partE); \
  } while (0)
      if (g.PT == 8) SQZ_LAUNCH_FAST(8);
      else if (g.PT == 16) SQZ_LAUNCH_FAST(16);
      else SQZ_LAUNCH_FAST(32);
#undef SQZ_LAUNCH_FAST
      SQZ_CUDA(cudaGetLastError());
      ++launches;
    }
    if (prof && sb == S) SQZ_CUDA(cudaEventRecord(prof->ev[2], st));
    const int64_t items = a.S * nCs * nP;
    baf_grid_finalize_kernel<<<(unsigned)((items + 255) / 256), 256, 0, st>>>(
        a, w.partM, w.partE, w.fix_list, w.fix_count);
    SQZ_CUDA(cudaGetLastError());
    baf_grid_faithful_kernel<<<faithful_ctas, 256, sm_faith, st>>>(a, m, w.fix_list,
                                                                  w.fix_count);
    SQZ_CUDA(cudaGetLastError());
    launches += 2;
  }
  if (prof) {
    prof->path = g.family;
    prof->launches = launches;
    SQZ_CUDA(cudaMemcpyAsync(&prof->n_eval, w.stats, sizeof(unsigned long long),
                             cudaMemcpyDeviceToHost, st));
    SQZ_CUDA(cudaEventRecord(prof->ev[3], st));
  }
  return SQZ_OK;
}

// column shard [ic0, ic1): results in d_out[S][(ic1 - ic0) * nP]
int baf_grid_launch(int64_t S, const int64_t* d_seg_off, int64_t N_total, const double* d_Bf,
                    const double* d_ratio, const double* d_sdBf, const double* d_sdr,
                    const double* d_wBf, const double* d_wr, const double* h_ploidy,
                    int64_t nP, const double* h_cell, int64_t nC, int64_t ic0, int64_t ic1,
                    const double* d_adr, int CNt_min, int CNt_max, int CNn,
                    const double* prior_CN, const double* prior_value, int n_prior,
                    int ratio_priority, int mode, void* workspace, size_t workspace_bytes,
                    cudaStream_t st, double* d_out, bool allow_profile) {
  GridShard sh;
  sh.by_units = false;
  sh.a0 = ic0;
  sh.a1 = ic1;
  sh.n_out = 1;
  sh.outs[0] = d_out;
  if (S > 0 && nP > 0 && ic1 > ic0 && !d_out) return fail(SQZ_ERR_ARG, "output buffer is null");
  if (!d_out) sh.outs[0] = (double*)workspace;  // nothing will be written
  return baf_grid_core(S, d_seg_off, N_total, d_Bf, d_ratio, d_sdBf, d_sdr, d_wBf, d_wr,
                       h_ploidy, nP, h_cell, nC, sh, d_adr, CNt_min, CNt_max, CNn, prior_CN,
                       prior_value, n_prior, ratio_priority, mode, workspace, workspace_bytes,
                       st, allow_profile);
}

// unit shard [u0, u1) of the whole grid: results at their final place of every buffer in
// d_outs (each S x nC x nP, ploidy fastest) -- this GPU's and, over NVLink, its peers'
int baf_grid_launch_units(int64_t S, const int64_t* d_seg_off, int64_t N_total,
                          const double* d_Bf, const double* d_ratio, const double* d_sdBf,
                          const double* d_sdr, const double* d_wBf, const double* d_wr,
                          const double* h_ploidy, int64_t nP, const double* h_cell, int64_t nC,
                          int64_t u0, int64_t u1, const double* d_adr, int CNt_min,
                          int CNt_max, int CNn, const double* prior_CN,
                          const double* prior_value, int n_prior, int ratio_priority, int mode,
                          void* workspace, size_t workspace_bytes, cudaStream_t st,
                          double* const* d_outs, int n_out) {
  if (n_out < 1 || n_out > kMaxOut || !d_outs)
    return fail(SQZ_ERR_ARG, "between 1 and 8 output buffers");
  GridShard sh;
  sh.by_units = true;
  sh.a0 = u0;
  sh.a1 = u1;
  sh.n_out = n_out;
  for (int k = 0; k < n_out; ++k) sh.outs[k] = d_outs[k];
  return baf_grid_core(S, d_seg_off, N_total, d_Bf, d_ratio, d_sdBf, d_sdr, d_wBf, d_wr,
                       h_ploidy, nP, h_cell, nC, sh, d_adr, CNt_min, CNt_max, CNn, prior_CN,
                       prior_value, n_prior, ratio_priority, mode, workspace, workspace_bytes,
                       st, true);
}

int baf_point_launch(const double* d_Bf, const double* d_ratio, const double* d_sdBf,
                     const double* d_sdr, const double* d_wBf, const double* d_wr, int64_t N,
                     double cellularity, double ploidy, double adr, int CNt_min, int CNt_max,
                     int CNn, const double* prior_CN, const double* prior_value, int n_prior,
                     int ratio_priority, cudaStream_t st, double* d_out) {
  if (N < 0) return fail(SQZ_ERR_ARG, "negative N");
  CnModel m;
  int rc = build_cn_model(0, CNt_min, CNt_max, CNn, prior_CN, prior_value, n_prior, &m,
                          nullptr);
  if (rc) return rc;
  if (N == 0) return SQZ_OK;
  int64_t blocks = (N + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  baf_point_kernel<<<(unsigned)blocks, 256, (size_t)(m.nc + m.T) * sizeof(double), st>>>(
      N, d_Bf, d_ratio, d_sdBf, d_sdr, d_wBf, d_wr, cellularity, ploidy, adr, m,
      ratio_priority ? 1 : 0, dt5_const(), d_out);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

// ---- grid fit for callers that must not allocate ------------------------------------
// Same kernels as sqz_baf_grid_dev; the only difference is where the scratch lives.
// Instead of a caller-provided workspace it is this fixed array inside the library
// image, so the call path contains no cudaMalloc/cudaFree and no allocating routine.
// Every GPU that loads the library carries these bytes until the process exits.
constexpr size_t kNoAllocScratchBytes = 16u << 20;  // 16 MiB
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not an address a
// kernel can use; the device address is looked up once per device and kept here.
constexpr int kMaxDevices = 64;
static void* g_noalloc_ptr[kMaxDevices] = {nullptr};
static std::mutex g_noalloc_mu;

static int noalloc_scratch_for_current_device(void** out) {
  int dev = -1;
  SQZ_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices)
    return fail(SQZ_ERR_SIZE, "device ordinal beyond the 64 this entry point tracks");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    SQZ_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));  // with `dev` current
    if (!p) return fail(SQZ_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return SQZ_OK;
}

}  // namespace sqz

extern "C" int sqz_baf_grid_dev_noalloc(
    int64_t S, const int64_t* d_seg_offsets, int64_t N_total, const double* d_Bf,
    const double* d_depth_ratio, const double* d_sd_Bf, const double* d_sd_ratio,
    const double* d_weight_Bf, const double* d_weight_ratio, const double* ploidy, int64_t nP,
    const double* cellularity, int64_t nC, int64_t ic0, int64_t ic1,
    const double* d_avg_depth_ratio, int CNt_min, int CNt_max, int CNn, const double* prior_CN,
    const double* prior_value, int n_prior, int ratio_priority, int mode, void* stream,
    double* d_lpp_sum_out) {
  using namespace sqz;
  {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1) {
      cudaGetLastError();
      return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
    }
  }
  if (S < 0 || N_total < 0 || nP < 0 || nC < 0 || ic0 < 0 || ic1 > nC || ic0 > ic1)
    return fail(SQZ_ERR_ARG, "negative size or bad cellularity shard");
  const size_t need = baf_grid_workspace_bytes(N_total, S, nP, ic1 - ic0);
  if (need > kNoAllocScratchBytes) {
    char msg[256];
    snprintf(msg, sizeof msg,
             "sqz_baf_grid_dev_noalloc: this problem needs %zu bytes of scratch but the "
             "built-in scratch is %zu bytes (16 MiB); use sqz_baf_grid_dev with a "
             "caller-owned workspace instead",
             need, kNoAllocScratchBytes);
    return fail(SQZ_ERR_SIZE, msg);
  }
  void* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc) return rc;
  // the optional per-kernel timing creates CUDA events on first use: it is switched off for
  // this call (allow_profile = false), a path that promises not to allocate
  rc = baf_grid_launch(S, d_seg_offsets, N_total, d_Bf, d_depth_ratio, d_sd_Bf, d_sd_ratio,
                       d_weight_Bf, d_weight_ratio, ploidy, nP, cellularity, nC, ic0, ic1,
                       d_avg_depth_ratio, CNt_min, CNt_max, CNn, prior_CN, prior_value,
                       n_prior, ratio_priority, mode, scratch, kNoAllocScratchBytes,
                       (cudaStream_t)stream, d_lpp_sum_out, false);
  return rc;
}
