// sqz_mufreq.cu -- mufreq.bayes (R/bayes.R:126-173) and the mufreq.model.fit grid
// (R/bayes.R:175-186), hand-written for sm_100a.
//
// Per position the reference evaluates, for every (CNt, Mt) row t,
//   score.r[t] = dbinom(round(w.r * r/(1+r)), w.r, mr_t/(1+mr_t)) * prior_t   (R/bayes.R:25-29, :153)
//   score.m[t] = dbinom(round(F * w.m), w.m, mufreq_t or 0.01)                 (R/bayes.R:1-6)
//   post[t]    = score.r[t] * score.m[t],  exact zeros -> 1e-323               (R/bayes.R:155-156)
// then takes CN* = CNt[which.max(score.r)] and the first maximum of log(post) among the
// rows with CNt == CN* (R/bayes.R:157-164).
//
// The kernels keep R's nmath operation sequence (dbinom_raw, bd0, stirlerr) and only
// remove repetition that cannot change a bit of the result:
//   - score.r is the same number for all rows of one CNt, so it is evaluated once per
//     copy number;
//   - the three stirlerr terms and the lf term of dbinom_raw depend on (x, n) only, so
//     they are evaluated once per position (and, for the grid, once per fit in a
//     preparation pass) instead of once per row;
//   - rows outside CN* never reach the output, so only the CN* rows get a score.m.
#include <stdlib.h>

#include "sqz_common.cuh"
#include "../../include/sequenza_b200.h"

namespace sqz {

int build_cn_model(int kind, int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                   const double* prior_value, int n_prior, CnModel* m, bool* fast_ok);

// model tables for one grid point: prob_r per copy number, model mufreq per row
__device__ __forceinline__ void build_muf_model(const CnModel& m, double cell, double ploidy,
                                                double adr, double* probr, double* mm) {
  for (int c = threadIdx.x; c < m.nc; c += blockDim.x) {
    double mr = theoretical_depth_ratio((double)(m.CNt_min + c), cell, ploidy,
                                        (double)m.CNn, adr);
    probr[c] = mr / (1 + mr);  // R/bayes.R:27
  }
  for (int t = threadIdx.x; t < m.T; t += blockDim.x) {
    int c = 0;
    while (m.toff[c + 1] <= t) ++c;
    double v = theoretical_mufreq((double)(m.CNt_min + c), (double)(t - m.toff[c]), cell,
                                  (double)m.CNn);
    mm[t] = (v == 0) ? 0.01 : v;  // R/bayes.R:3 (seq.errors)
  }
}

__device__ __forceinline__ void mufreq_eval(const CnModel& m, const double* probr,
                                            const double* mm, const BinomHoist& hr,
                                            const BinomHoist& hm, int& kc, int& kt,
                                            double& lpp) {
  kc = -1;
  kt = 0;
  double bests = 0.;
  for (int c = 0; c < m.nc; ++c) {
    const double sr = binom_finish(hr, probr[c]) * m.prior[c];
    if (!isnan(sr) && (kc < 0 || sr > bests)) {  // which.max(score.r)
      bests = sr;
      kc = c;
    }
  }
  if (kc < 0) {
    lpp = __longlong_as_double(0x7FF8000000000000LL);
    return;
  }
  int k = -1;
  double best = 0.;
  for (int t = m.toff[kc]; t < m.toff[kc + 1]; ++t) {
    double post = bests * binom_finish(hm, mm[t]);
    if (post == 0) post = __longlong_as_double(2LL);  // 1e-323, R/bayes.R:148, :156
    const double lp = log(post);
    if (!isnan(lp) && (k < 0 || lp > best)) {
      best = lp;
      k = t;
    }
  }
  if (k < 0) {
    kc = -1;
    lpp = __longlong_as_double(0x7FF8000000000000LL);
    return;
  }
  kt = k - m.toff[kc];
  lpp = best;
}

__device__ __forceinline__ void hoist_position(double F, double ratio, double wm, double wr,
                                               BinomHoist& hr, BinomHoist& hm) {
  hr = binom_hoist(nearbyint(wr * (ratio / (1 + ratio))), wr);  // R/bayes.R:26
  hm = binom_hoist(nearbyint(F * wm), wm);                      // R/bayes.R:4
}

__global__ void __launch_bounds__(256)
mufreq_point_kernel(int64_t N, const double* F, const double* ratio, const double* wm,
                    const double* wr, double cell, double ploidy, double adr, CnModel m,
                    double* __restrict__ out) {
  extern __shared__ double sm[];
  double* probr = sm;
  double* mm = sm + m.nc;
  build_muf_model(m, cell, ploidy, adr, probr, mm);
  __syncthreads();
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < N;
       s += (int64_t)gridDim.x * blockDim.x) {
    BinomHoist hr, hm;
    hoist_position(F[s], ratio[s], wm[s], wr[s], hr, hm);
    int kc, kt;
    double lpp;
    mufreq_eval(m, probr, mm, hr, hm, kc, kt, lpp);
    out[s] = kc >= 0 ? (double)m.CNn : na_real();
    out[N + s] = kc >= 0 ? (double)(m.CNt_min + kc) : na_real();
    out[2 * N + s] = kc >= 0 ? (double)kt : na_real();
    out[3 * N + s] = lpp;
  }
}

__global__ void mufreq_prep_kernel(int64_t N, const double* F, const double* ratio,
                                   const double* wm, const double* wr, BinomHoist* hr,
                                   BinomHoist* hm) {
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < N;
       s += (int64_t)gridDim.x * blockDim.x)
    hoist_position(F[s], ratio[s], wm[s], wr[s], hr[s], hm[s]);
}

// grid: persistent CTAs over this shard's grid points; threads stride over positions
__global__ void __launch_bounds__(256)
mufreq_grid_kernel(int64_t N, const BinomHoist* __restrict__ hr,
                   const BinomHoist* __restrict__ hm, const double* __restrict__ ploidy,
                   int nP, const double* __restrict__ cell, int nCs, double adr, CnModel m,
                   double* __restrict__ out) {
  extern __shared__ double sm[];
  double* probr = sm;
  double* mm = sm + m.nc;
  __shared__ double red[3][8];
  const int64_t items = (int64_t)nP * nCs;
  for (int64_t g = blockIdx.x; g < items; g += gridDim.x) {
    const int ic = (int)(g / nP), ip = (int)(g - (int64_t)ic * nP);
    __syncthreads();
    build_muf_model(m, cell[ic], ploidy[ip], adr, probr, mm);
    __syncthreads();
    DD acc = dd_zero();
    for (int64_t s = threadIdx.x; s < N; s += blockDim.x) {
      const BinomHoist a = hr[s], b = hm[s];
      int kc, kt;
      double lpp;
      mufreq_eval(m, probr, mm, a, b, kc, kt, lpp);
      dd_add(acc, lpp);
    }
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
      out[g] = dd_value(tot);
    }
  }
}

__global__ void model_points_kernel(int kind, double cell, double ploidy, double adr,
                                    CnModel m, double* out) {
  const int T = m.T;
  for (int t = threadIdx.x; t < T; t += blockDim.x) {
    int c = 0;
    while (m.toff[c + 1] <= t) ++c;
    const double cn = (double)(m.CNt_min + c), second = (double)(t - m.toff[c]);
    out[t] = cn;
    out[T + t] = second;
    out[2 * T + t] = theoretical_depth_ratio(cn, cell, ploidy, (double)m.CNn, adr);
    out[3 * T + t] = kind == 0 ? theoretical_baf(cn, second, cell, (double)m.CNn)
                               : theoretical_mufreq(cn, second, cell, (double)m.CNn);
    out[4 * T + t] = m.prior[c];
  }
}

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

size_t mufreq_grid_workspace_bytes(int64_t N, int64_t nP, int64_t nCs) {
  const size_t n = (size_t)(N > 0 ? N : 1);
  return 2 * align256(n * sizeof(BinomHoist)) + align256((size_t)(nP > 0 ? nP : 1) * 8) +
         align256((size_t)(nCs > 0 ? nCs : 1) * 8);
}

int mufreq_point_launch(const double* d_F, const double* d_ratio, const double* d_wm,
                        const double* d_wr, int64_t N, double cellularity, double ploidy,
                        double adr, int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                        const double* prior_value, int n_prior, cudaStream_t st,
                        double* d_out) {
  if (N < 0) return fail(SQZ_ERR_ARG, "negative N");
  CnModel m;
  int rc = build_cn_model(1, CNt_min, CNt_max, CNn, prior_CN, prior_value, n_prior, &m,
                          nullptr);
  if (rc) return rc;
  if (N == 0) return SQZ_OK;
  int64_t blocks = (N + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  mufreq_point_kernel<<<(unsigned)blocks, 256, (size_t)(m.nc + m.T) * sizeof(double), st>>>(
      N, d_F, d_ratio, d_wm, d_wr, cellularity, ploidy, adr, m, d_out);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

int mufreq_grid_launch(const double* d_F, const double* d_ratio, const double* d_wm,
                       const double* d_wr, int64_t N, const double* h_ploidy, int64_t nP,
                       const double* h_cell, int64_t nC, int64_t ic0, int64_t ic1, double adr,
                       int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                       const double* prior_value, int n_prior, void* workspace,
                       size_t workspace_bytes, cudaStream_t st, double* d_out) {
  if (N < 0 || nP < 0 || nC < 0 || ic0 < 0 || ic1 > nC || ic0 > ic1)
    return fail(SQZ_ERR_ARG, "negative size or bad cellularity shard");
  const int64_t nCs = ic1 - ic0;
  if (nCs == 0 || nP == 0) return SQZ_OK;
  CnModel m;
  int rc = build_cn_model(1, CNt_min, CNt_max, CNn, prior_CN, prior_value, n_prior, &m,
                          nullptr);
  if (rc) return rc;
  if (!workspace || ((uintptr_t)workspace & 255))
    return fail(SQZ_ERR_ARG, "workspace must be non-null and 256-byte aligned");
  if (mufreq_grid_workspace_bytes(N, nP, nCs) > workspace_bytes)
    return fail(SQZ_ERR_SIZE, "workspace too small");
  char* p = (char*)workspace;
  const size_t n = (size_t)(N > 0 ? N : 1);
  BinomHoist* hr = (BinomHoist*)p;
  p += align256(n * sizeof(BinomHoist));
  BinomHoist* hm = (BinomHoist*)p;
  p += align256(n * sizeof(BinomHoist));
  double* d_ploidy = (double*)p;
  p += align256((size_t)nP * 8);
  double* d_cell = (double*)p;
  SQZ_CUDA(cudaMemcpyAsync(d_ploidy, h_ploidy, nP * sizeof(double), cudaMemcpyHostToDevice, st));
  SQZ_CUDA(cudaMemcpyAsync(d_cell, h_cell + ic0, nCs * sizeof(double), cudaMemcpyHostToDevice, st));
  if (N > 0) {
    int64_t blocks = (N + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    mufreq_prep_kernel<<<(unsigned)blocks, 256, 0, st>>>(N, d_F, d_ratio, d_wm, d_wr, hr, hm);
    SQZ_CUDA(cudaGetLastError());
  }
  int64_t ctas = nP * nCs;
  if (ctas > 148 * 8) ctas = 148 * 8;
  const int threads = N > 2048 ? 256 : 128;
  mufreq_grid_kernel<<<(unsigned)ctas, threads, (size_t)(m.nc + m.T) * sizeof(double), st>>>(
      N, hr, hm, d_ploidy, (int)nP, d_cell, (int)nCs, adr, m, d_out);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

int model_points_launch(int kind, double cellularity, double ploidy, double adr, int CNt_min,
                        int CNt_max, int CNn, const double* prior_CN,
                        const double* prior_value, int n_prior, int* T_out, double* h_out) {
  if (kind != 0 && kind != 1) return fail(SQZ_ERR_ARG, "kind must be 0 (baf) or 1 (mufreq)");
  CnModel m;
  int rc = build_cn_model(kind, CNt_min, CNt_max, CNn, prior_CN, prior_value, n_prior, &m,
                          nullptr);
  if (rc) return rc;
  if (T_out) *T_out = m.T;
  if (!h_out) return SQZ_OK;
  double* d = nullptr;
  SQZ_CUDA(cudaMalloc(&d, (size_t)m.T * 5 * sizeof(double)));
  model_points_kernel<<<1, 128>>>(kind, cellularity, ploidy, adr, m, d);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess)
    e = cudaMemcpy(h_out, d, (size_t)m.T * 5 * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return fail(SQZ_ERR_CUDA, cudaGetErrorString(e));
  return SQZ_OK;
}

// ---- scalar weights: the answer depends on two integers only ---------------------------
// sequenza.results calls mufreq.bayes with the default scalar weights 100 / 100
// (R/results.R:153-157).  Then the row of a position is a function of
//   k_r = round(w.r * r/(1+r)) in 0..w.r   and   k_m = round(F * w.m) in 0..w.m
// alone, so the (w.r+1) x (w.m+1) answers are computed once with the same nmath sequence
// (mufreq_scores_kernel: one thread per binomial density; mufreq_table_kernel: one thread
// per pair of counts) and the per-position kernel becomes a
// streaming lookup: 16 bytes in, 32 bytes out, HBM-bound.  Positions whose counts fall
// outside the table (NA, negative ratio, F outside [0,1]) take the direct path.
// Stage 1: every binomial density the table needs, one thread each:
//   srp[kr][c] = dbinom(kr, w.r, prob_c) * prior_c     (score.r, R/bayes.R:149-153)
//   smv[km][t] = dbinom(km, w.m, mufreq_t)             (score.m, R/bayes.R:151-152)
__global__ void __launch_bounds__(128)
mufreq_scores_kernel(int nr, int nm, double wr, double wm, double cell, double ploidy,
                     double adr, CnModel m, double* __restrict__ srp, double* __restrict__ smv) {
  extern __shared__ double sm[];
  double* probr = sm;
  double* mm = sm + m.nc;
  build_muf_model(m, cell, ploidy, adr, probr, mm);
  __syncthreads();
  const int64_t n1 = (int64_t)nr * m.nc, n = n1 + (int64_t)nm * m.T;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    if (i < n1) {
      const int kr = (int)(i / m.nc), c = (int)(i - (int64_t)kr * m.nc);
      srp[i] = binom_finish(binom_hoist((double)kr, wr), probr[c]) * m.prior[c];
    } else {
      const int64_t k = i - n1;
      const int km = (int)(k / m.T), t = (int)(k - (int64_t)km * m.T);
      smv[k] = binom_finish(binom_hoist((double)km, wm), mm[t]);
    }
  }
}

// Stage 2: which.max(score.r) per kr (first maximum, NaN skipped), then for every (kr, km)
// the first maximum of log(post) among the rows of that copy number (R/bayes.R:155-164).
// One WARP per pair of counts: the lanes take the copy numbers, then the rows, and a
// shuffle reduction keeps the first maximum (a thread per pair walked up to 21 logarithms
// one after the other: 15 us of latency for a 40 KB table).  Besides the double4 answers
// (what the global-memory lookup reads) the table is written in a compact form, 8 bytes of
// LPP and a 16-bit code (CNt << 8 | Mt, 0xFFFF = NA) per pair, which fits shared memory.
struct FirstMax {
  double v;
  int k;  // -1: none yet
};
__device__ __forceinline__ FirstMax first_max(FirstMax a, FirstMax b) {
  if (b.k < 0) return a;
  if (a.k < 0) return b;
  if (b.v > a.v || (b.v == a.v && b.k < a.k)) return b;
  return a;
}
__device__ __forceinline__ FirstMax warp_first_max(FirstMax a) {
#pragma unroll
  for (int off = 16; off; off >>= 1) {
    FirstMax b;
    b.v = __shfl_xor_sync(0xffffffffu, a.v, off);
    b.k = __shfl_xor_sync(0xffffffffu, a.k, off);
    a = first_max(a, b);
  }
  return a;
}

constexpr unsigned kMufNaCode = 0xFFFFu;

__global__ void __launch_bounds__(256)
mufreq_table_kernel(int nr, int nm, CnModel m, const double* __restrict__ srp,
                    const double* __restrict__ smv, double4* __restrict__ table,
                    double* __restrict__ lpp8, unsigned short* __restrict__ code16) {
  const int lane = threadIdx.x & 31;
  const int64_t n = (int64_t)nr * nm;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
    const int kr = (int)(i / nm), km = (int)(i - (int64_t)kr * nm);
    FirstMax bc = {0., -1};
    for (int c = lane; c < m.nc; c += 32) {
      const double sr = srp[(int64_t)kr * m.nc + c];
      if (!isnan(sr) && (bc.k < 0 || sr > bc.v)) bc = {sr, c};
    }
    bc = warp_first_max(bc);
    const int kc = bc.k;
    FirstMax bt = {0., -1};
    if (kc >= 0)
      for (int t = m.toff[kc] + lane; t < m.toff[kc + 1]; t += 32) {
        double post = bc.v * smv[(int64_t)km * m.T + t];
        if (post == 0) post = __longlong_as_double(2LL);  // 1e-323
        const double lp = log(post);
        if (!isnan(lp) && (bt.k < 0 || lp > bt.v)) bt = {lp, t};
      }
    bt = warp_first_max(bt);
    if (lane == 0) {
      const bool ok = kc >= 0 && bt.k >= 0;
      const int cnt = ok ? m.CNt_min + kc : 0, mt = ok ? bt.k - m.toff[kc] : 0;
      const double lpp = ok ? bt.v : __longlong_as_double(0x7FF8000000000000LL);
      table[i] = make_double4(ok ? (double)m.CNn : na_real(), ok ? (double)cnt : na_real(),
                              ok ? (double)mt : na_real(), lpp);
      lpp8[i] = lpp;
      code16[i] = (unsigned short)(ok ? ((unsigned)cnt << 8 | (unsigned)mt) : kMufNaCode);
    }
  }
}

__global__ void __launch_bounds__(256)
mufreq_lookup_kernel(int64_t N, const double* __restrict__ F, const double* __restrict__ ratio,
                     int nr, int nm, double wr, double wm, double cell, double ploidy,
                     double adr, CnModel m, const double4* __restrict__ table,
                     double* __restrict__ out) {
  extern __shared__ double sm[];
  double* probr = sm;
  double* mm = sm + m.nc;
  build_muf_model(m, cell, ploidy, adr, probr, mm);  // for the rare direct path
  __syncthreads();
  // four positions per thread and iteration: all input loads, then all table lookups,
  // then the stores, so that several memory requests are in flight per thread
  constexpr int U = 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t s0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s0 < N; s0 += U * stride) {
    double r[U], f[U], kr[U], km[U];
    double4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t s = s0 + u * stride;
      r[u] = s < N ? ratio[s] : 1.0;
      f[u] = s < N ? F[s] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      kr[u] = nearbyint(wr * (r[u] / (1 + r[u])));  // R/bayes.R:26
      km[u] = nearbyint(f[u] * wm);                  // R/bayes.R:4
      const bool in = kr[u] >= 0.0 && kr[u] < (double)nr && km[u] >= 0.0 && km[u] < (double)nm;
      v[u] = table[in ? (int64_t)kr[u] * nm + (int64_t)km[u] : 0];
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t s = s0 + u * stride;
      if (s >= N) continue;
      const bool in = kr[u] >= 0.0 && kr[u] < (double)nr && km[u] >= 0.0 && km[u] < (double)nm;
      if (!in) {  // outside the table (NA, ratio < 0, F outside [0, 1]): direct evaluation
        const BinomHoist hr = binom_hoist(kr[u], wr), hm = binom_hoist(km[u], wm);
        int kc, kt;
        double lpp;
        mufreq_eval(m, probr, mm, hr, hm, kc, kt, lpp);
        v[u] = make_double4(kc >= 0 ? (double)m.CNn : na_real(),
                            kc >= 0 ? (double)(m.CNt_min + kc) : na_real(),
                            kc >= 0 ? (double)kt : na_real(), lpp);
      }
      out[s] = v[u].x;
      out[N + s] = v[u].y;
      out[2 * N + s] = v[u].z;
      out[3 * N + s] = v[u].w;
    }
  }
}

// The same lookup with 128-bit accesses: a thread takes two neighbouring positions (one
// double2 load per input vector, one double2 store per output column) and two such pairs per
// iteration; inputs and outputs are touched once, so they are loaded / stored with the
// streaming hints and leave the L2 to the 326 KB answer table.  Needs an even N and 16-byte
// aligned vectors (the launcher checks); positions outside the table take the direct
// evaluation, as above.
__global__ void __launch_bounds__(256)
mufreq_lookup2_kernel(int64_t N, const double2* __restrict__ F2, const double2* __restrict__ R2,
                      int nr, int nm, double wr, double wm, double cell, double ploidy,
                      double adr, CnModel m, const double4* __restrict__ table,
                      double* __restrict__ out) {
  extern __shared__ double sm[];
  double* probr = sm;
  double* mm = sm + m.nc;
  build_muf_model(m, cell, ploidy, adr, probr, mm);  // for the rare direct path
  __syncthreads();
  constexpr int U = 2;  // pairs per thread and iteration: four positions in flight
  const int64_t NP = N >> 1, stride = (int64_t)gridDim.x * blockDim.x;
  double2* o0 = reinterpret_cast<double2*>(out);
  double2* o1 = reinterpret_cast<double2*>(out + N);
  double2* o2 = reinterpret_cast<double2*>(out + 2 * N);
  double2* o3 = reinterpret_cast<double2*>(out + 3 * N);
  for (int64_t p0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p0 < NP; p0 += U * stride) {
    double2 r[U], f[U];
    double4 v[2 * U];
    double kr[2 * U], km[2 * U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t p = p0 + u * stride;
      r[u] = p < NP ? __ldcs(R2 + p) : make_double2(1.0, 1.0);
      f[u] = p < NP ? __ldcs(F2 + p) : make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int u = 0; u < 2 * U; ++u) {
      const double rr = (u & 1) ? r[u >> 1].y : r[u >> 1].x;
      const double ff = (u & 1) ? f[u >> 1].y : f[u >> 1].x;
      kr[u] = nearbyint(wr * (rr / (1 + rr)));  // R/bayes.R:26
      km[u] = nearbyint(ff * wm);               // R/bayes.R:4
      const bool in = kr[u] >= 0.0 && kr[u] < (double)nr && km[u] >= 0.0 && km[u] < (double)nm;
      v[u] = table[in ? (int64_t)kr[u] * nm + (int64_t)km[u] : 0];
    }
#pragma unroll
    for (int u = 0; u < 2 * U; ++u) {
      const bool in = kr[u] >= 0.0 && kr[u] < (double)nr && km[u] >= 0.0 && km[u] < (double)nm;
      if (!in && p0 + (u >> 1) * stride < NP) {  // NA, ratio < 0, F outside [0, 1]
        const BinomHoist hr = binom_hoist(kr[u], wr), hm = binom_hoist(km[u], wm);
        int kc, kt;
        double lpp;
        mufreq_eval(m, probr, mm, hr, hm, kc, kt, lpp);
        v[u] = make_double4(kc >= 0 ? (double)m.CNn : na_real(),
                            kc >= 0 ? (double)(m.CNt_min + kc) : na_real(),
                            kc >= 0 ? (double)kt : na_real(), lpp);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t p = p0 + u * stride;
      if (p >= NP) continue;
      __stcs(o0 + p, make_double2(v[2 * u].x, v[2 * u + 1].x));
      __stcs(o1 + p, make_double2(v[2 * u].y, v[2 * u + 1].y));
      __stcs(o2 + p, make_double2(v[2 * u].z, v[2 * u + 1].z));
      __stcs(o3 + p, make_double2(v[2 * u].w, v[2 * u + 1].w));
    }
  }
}

// The lookup with the compact table in SHARED memory (101 x 101 pairs: 100 KB, two CTAs per
// SM): every position costs one 8-byte and one 2-byte shared-memory load instead of a
// 32-byte gather through L1/L2, whose tag stage -- 32 different lines per warp request --
// took as long as the HBM streams themselves.  Inputs and outputs as in
// mufreq_lookup2_kernel (128-bit streaming accesses, two pairs of positions per thread and
// iteration).  Copy numbers and row numbers come back as doubles by the 2^52 trick.
__device__ __forceinline__ double small_uint_to_double(unsigned k) {
  return __hiloint2double(0x43300000, (int)k) - 4503599627370496.0;
}

// positions outside the table (NA, ratio < 0, F outside [0, 1]): the direct evaluation, kept
// out of line so that the streaming loop stays within 64 registers
__device__ __noinline__ double4 muf_direct(const CnModel& m, const double* probr,
                                           const double* mm, double kr, double wr, double km,
                                           double wm) {
  const BinomHoist hr = binom_hoist(kr, wr), hm = binom_hoist(km, wm);
  int kc, kt;
  double lpp;
  mufreq_eval(m, probr, mm, hr, hm, kc, kt, lpp);
  return make_double4(kc >= 0 ? (double)m.CNn : na_real(),
                      kc >= 0 ? (double)(m.CNt_min + kc) : na_real(),
                      kc >= 0 ? (double)kt : na_real(), lpp);
}

__global__ void __launch_bounds__(512, 2)
mufreq_lookup_smem_kernel(int64_t N, const double2* __restrict__ F2,
                          const double2* __restrict__ R2, int nr, int nm, double wr, double wm,
                          double cell, double ploidy, double adr,
                          const __grid_constant__ CnModel m, const double* __restrict__ lpp8,
                          const unsigned short* __restrict__ code16, double* __restrict__ out) {
  extern __shared__ __align__(16) double sm[];
  const int n = nr * nm, n8 = (n + 1) & ~1, n16 = (n + 7) & ~7;  // padded to 16 bytes
  double* s_lpp = sm;
  unsigned short* s_code = reinterpret_cast<unsigned short*>(sm + n8);
  double* probr = reinterpret_cast<double*>(s_code + n16);
  double* mm = probr + m.nc;
  {
    const int4* g = reinterpret_cast<const int4*>(lpp8);
    int4* d = reinterpret_cast<int4*>(s_lpp);
    for (int i = threadIdx.x; i < n8 / 2; i += blockDim.x) d[i] = g[i];
    g = reinterpret_cast<const int4*>(code16);
    d = reinterpret_cast<int4*>(s_code);
    for (int i = threadIdx.x; i < n16 / 8; i += blockDim.x) d[i] = g[i];
  }
  build_muf_model(m, cell, ploidy, adr, probr, mm);  // for the rare direct path
  __syncthreads();
  constexpr int U = 2;  // pairs per thread and iteration: four positions in flight
  const int64_t NP = N >> 1, stride = (int64_t)gridDim.x * blockDim.x;
  const double cnn = (double)m.CNn;
  double2* o0 = reinterpret_cast<double2*>(out);
  double2* o1 = reinterpret_cast<double2*>(out + N);
  double2* o2 = reinterpret_cast<double2*>(out + 2 * N);
  double2* o3 = reinterpret_cast<double2*>(out + 3 * N);
  for (int64_t p0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p0 < NP; p0 += U * stride) {
    double2 r[U], f[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t p = p0 + u * stride;
      r[u] = p < NP ? __ldcs(R2 + p) : make_double2(1.0, 1.0);
      f[u] = p < NP ? __ldcs(F2 + p) : make_double2(0.0, 0.0);
    }
    double c1[2 * U], c2[2 * U], c3[2 * U], lp[2 * U];
    unsigned rare = 0;
#pragma unroll
    for (int u = 0; u < 2 * U; ++u) {
      const double rr = (u & 1) ? r[u >> 1].y : r[u >> 1].x;
      const double ff = (u & 1) ? f[u >> 1].y : f[u >> 1].x;
      const double kr = nearbyint(wr * (rr / (1 + rr)));  // R/bayes.R:26
      const double km = nearbyint(ff * wm);               // R/bayes.R:4
      const bool in = kr >= 0.0 && kr < (double)nr && km >= 0.0 && km < (double)nm;
      const int idx = in ? (int)kr * nm + (int)km : 0;
      const unsigned cd = s_code[idx];
      const bool ok = cd != kMufNaCode;
      c1[u] = ok ? cnn : na_real();
      c2[u] = ok ? small_uint_to_double(cd >> 8) : na_real();
      c3[u] = ok ? small_uint_to_double(cd & 255u) : na_real();
      lp[u] = s_lpp[idx];
      if (!in) rare |= 1u << u;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t p = p0 + u * stride;
      if (p >= NP) continue;
      __stcs(o0 + p, make_double2(c1[2 * u], c1[2 * u + 1]));
      __stcs(o1 + p, make_double2(c2[2 * u], c2[2 * u + 1]));
      __stcs(o2 + p, make_double2(c3[2 * u], c3[2 * u + 1]));
      __stcs(o3 + p, make_double2(lp[2 * u], lp[2 * u + 1]));
    }
    // positions outside the table (NA, ratio < 0, F outside [0, 1]; rare): evaluated directly
    // and written over the placeholder this thread has just stored
    while (rare) {
      const int u = __ffs((int)rare) - 1;
      rare &= rare - 1;
      const int64_t p = p0 + (u >> 1) * stride;
      if (p >= NP) continue;
      const int64_t sidx = 2 * p + (u & 1);
      const double rr = reinterpret_cast<const double*>(R2)[sidx];
      const double ff = reinterpret_cast<const double*>(F2)[sidx];
      const double4 d = muf_direct(m, probr, mm, nearbyint(wr * (rr / (1 + rr))), wr,
                                   nearbyint(ff * wm), wm);
      out[sidx] = d.x;
      out[N + sidx] = d.y;
      out[2 * N + sidx] = d.z;
      out[3 * N + sidx] = d.w;
    }
  }
}

static size_t muf_compact_bytes(int64_t n) {  // lpp8 + code16, each padded to 16 bytes
  return (size_t)((n + 1) & ~(int64_t)1) * sizeof(double) +
         (size_t)((n + 7) & ~(int64_t)7) * sizeof(unsigned short);
}

// number of table bytes sqz_mufreq_point_scalar_dev needs (0: weights not usable for the
// table, the caller should use the vector-weight entry point)
size_t mufreq_scalar_table_bytes(double wm, double wr) {
  if (!(wm >= 0) || !(wr >= 0) || wm != nearbyint(wm) || wr != nearbyint(wr)) return 0;
  if (wm > 4096 || wr > 4096 || (wm + 1) * (wr + 1) > 4.0e6) return 0;
  // answer table + the two score tables of stage 1, sized for any copy-number range
  // + the compact copy of the answers for the shared-memory lookup
  return (size_t)((wm + 1) * (wr + 1)) * sizeof(double4) +
         (size_t)(wr + 1) * kMaxCN * sizeof(double) +
         (size_t)(wm + 1) * kMaxTypes * sizeof(double) +
         muf_compact_bytes((int64_t)((wm + 1) * (wr + 1)));
}

int mufreq_point_scalar_launch(const double* d_F, const double* d_ratio, int64_t N, double wm,
                               double wr, double cellularity, double ploidy, double adr,
                               int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                               const double* prior_value, int n_prior, void* table,
                               size_t table_bytes, cudaStream_t st, double* d_out) {
  if (N < 0) return fail(SQZ_ERR_ARG, "negative N");
  const size_t need = mufreq_scalar_table_bytes(wm, wr);
  if (need == 0)
    return fail(SQZ_ERR_ARG, "scalar weights must be non-negative integers <= 4096 (use the "
                             "vector-weight entry point otherwise)");
  if (!table || table_bytes < need || ((uintptr_t)table & 31))
    return fail(SQZ_ERR_SIZE, "table buffer missing, misaligned or too small");
  CnModel m;
  int rc = build_cn_model(1, CNt_min, CNt_max, CNn, prior_CN, prior_value, n_prior, &m,
                          nullptr);
  if (rc) return rc;
  if (N == 0) return SQZ_OK;
  const int nr = (int)wr + 1, nm = (int)wm + 1;
  const size_t smb = (size_t)(m.nc + m.T) * sizeof(double);
  double* srp = (double*)((char*)table + (size_t)nr * nm * sizeof(double4));
  double* smv = srp + (size_t)nr * kMaxCN;
  double* lpp8 = smv + (size_t)nm * kMaxTypes;
  const int64_t npair = (int64_t)nr * nm;
  unsigned short* code16 = reinterpret_cast<unsigned short*>(lpp8 + ((npair + 1) & ~(int64_t)1));
  int64_t sb = ((int64_t)nr * m.nc + (int64_t)nm * m.T + 127) / 128;
  if (sb > 148 * 16) sb = 148 * 16;
  mufreq_scores_kernel<<<(unsigned)sb, 128, smb, st>>>(nr, nm, wr, wm, cellularity, ploidy, adr,
                                                      m, srp, smv);
  SQZ_CUDA(cudaGetLastError());
  int64_t tb = (npair * 32 + 255) / 256;  // one warp per pair of counts
  if (tb > 148 * 16) tb = 148 * 16;
  mufreq_table_kernel<<<(unsigned)tb, 256, 0, st>>>(nr, nm, m, srp, smv, (double4*)table, lpp8,
                                                   code16);
  SQZ_CUDA(cudaGetLastError());
  const bool vec2 = (N % 2 == 0) &&
                    !(((uintptr_t)d_F | (uintptr_t)d_ratio | (uintptr_t)d_out) & 15);
  // compact table in shared memory: two CTAs of 512 threads per SM when it fits half an SM
  const size_t smem_tab = muf_compact_bytes(npair) + smb;
  static const bool no_smem_tab = getenv("SQZ_MUFREQ_NO_SMEM") != nullptr;
  if (vec2 && !no_smem_tab && N >= (1 << 16) && smem_tab <= 112 * 1024 &&
      CNt_max <= 255) {
    SQZ_CUDA(cudaFuncSetAttribute(mufreq_lookup_smem_kernel,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tab));
    mufreq_lookup_smem_kernel<<<148 * 2, 512, smem_tab, st>>>(
        N, (const double2*)d_F, (const double2*)d_ratio, nr, nm, wr, wm, cellularity, ploidy,
        adr, m, lpp8, code16, d_out);
    SQZ_CUDA(cudaGetLastError());
    return SQZ_OK;
  }
  int64_t blocks = ((vec2 ? N / 2 : N) + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  if (vec2)
    mufreq_lookup2_kernel<<<(unsigned)blocks, 256, smb, st>>>(
        N, (const double2*)d_F, (const double2*)d_ratio, nr, nm, wr, wm, cellularity, ploidy,
        adr, m, (const double4*)table, d_out);
  else
    mufreq_lookup_kernel<<<(unsigned)blocks, 256, smb, st>>>(N, d_F, d_ratio, nr, nm, wr, wm,
                                                            cellularity, ploidy, adr, m,
                                                            (const double4*)table, d_out);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

}  // namespace sqz
