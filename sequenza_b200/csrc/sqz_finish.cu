// sqz_finish.cu -- what happens to result$LPP after the grid search, on the device:
//   R/bayes.R:120-122   LogSumLik = log(sum(exp(LPP - max(LPP, na.rm = TRUE)))) + max;
//                       lpp = exp(z - LogSumLik)
//   R/solutions.R:1-15  get.ci: sort(lpp, decreasing), cumsum, the value at which the
//                       cumulative mass first reaches `level`, and which marginal maxima
//                       (apply(lpp, 1, max), apply(lpp, 2, max)) are at or above it.
// SURVEY 8f ranks this as the next row (N1): at 901 x 601 the host-side sort and the
// tapply/normalise steps cost more than the 8-GPU fit itself.
//
// No sort is needed.  get.ci only asks, for each marginal maximum v, whether v >= the
// threshold element, and that holds iff the total mass of the elements strictly greater
// than v is below `level` (v's first occurrence in the descending order then sits at or
// before the crossing).  So the kernel computes, for every marginal maximum, the mass
// above it: the nP + nC maxima are sorted in shared memory, every element finds its rank
// among them by binary search and adds its value to that rank's bin, and a suffix sum over
// the bins gives the masses.  Bins are 64-bit fixed point (2^-62 units) added with integer
// atomics, so the result does not depend on the order of the additions.  R's cumsum runs
// in long double; the two can disagree on a confidence limit only if a cumulative mass
// lies within ~1e-13 of `level`.
//
// One CTA per sample (a cohort has hundreds of small grids; a single fine grid keeps
// 1024 threads busy for ~0.1 ms).  Requires nP + nC <= 4096.
#include "sqz_common.cuh"
#include "../../include/sequenza_b200.h"

namespace sqz {

constexpr int kFinishThreads = 1024;
constexpr int kMaxMarginals = 4096;

__device__ __forceinline__ double block_max(double v, double* red) {
  for (int off = 16; off; off >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, off));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double r = red[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r = fmax(r, red[w]);
  __syncthreads();  // every thread has read the partials before the caller reuses `red`
  return r;
}

__global__ void __launch_bounds__(kFinishThreads)
fit_finish_kernel(const double* __restrict__ lpp_sum, int nP, int nC,
                  double* __restrict__ znorm, double* __restrict__ row_max,
                  double* __restrict__ col_max, double* __restrict__ row_above,
                  double* __restrict__ col_above, double* __restrict__ log_sum_lik) {
  __shared__ double red[3][32];
  __shared__ double s_lsl;
  extern __shared__ __align__(16) unsigned char dyn[];
  int Mp = 1;
  while (Mp < nP + nC) Mp <<= 1;
  double* thr = reinterpret_cast<double*>(dyn);                       // [Mp] sorted maxima
  unsigned long long* bins = reinterpret_cast<unsigned long long*>(thr + Mp);  // [Mp + 1]
  unsigned short* perm = reinterpret_cast<unsigned short*>(bins + Mp + 1);     // [Mp]
  const int64_t G = (int64_t)nP * nC;
  const int64_t smp = blockIdx.x;
  const double* z = lpp_sum + smp * G;
  double* zn = znorm + smp * G;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const double NEG_INF = __longlong_as_double(0xFFF0000000000000LL);

  // max(LPP, na.rm = TRUE)
  double mx = NEG_INF;
  for (int64_t i = tid; i < G; i += nthr) {
    const double v = z[i];
    if (!isnan(v) && v > mx) mx = v;
  }
  mx = block_max(mx, red[0]);
  // sum(exp(LPP - max)) without na.rm (a NaN poisons everything, as in R), double-double
  DD acc = dd_zero();
  for (int64_t i = tid; i < G; i += nthr) dd_add(acc, exp(z[i] - mx));
  for (int off = 16; off; off >>= 1) {
    DD o;
    o.hi = __shfl_xor_sync(0xffffffffu, acc.hi, off);
    o.lo = __shfl_xor_sync(0xffffffffu, acc.lo, off);
    o.special = __shfl_xor_sync(0xffffffffu, acc.special, off);
    dd_merge(acc, o);
  }
  {
    const int lane = tid & 31, warp = tid >> 5;
    if (lane == 0) {
      red[0][warp] = acc.hi;
      red[1][warp] = acc.lo;
      red[2][warp] = acc.special;
    }
    __syncthreads();
    if (tid == 0) {
      DD tot = dd_zero();
      for (int w = 0; w < (nthr >> 5); ++w) {
        DD o{red[0][w], red[1][w], red[2][w]};
        dd_merge(tot, o);
      }
      s_lsl = log(dd_value(tot)) + mx;
      log_sum_lik[smp] = s_lsl;
    }
    __syncthreads();
  }
  const double lsl = s_lsl;
  for (int64_t i = tid; i < G; i += nthr) zn[i] = exp(z[i] - lsl);
  __syncthreads();

  // marginal maxima.  apply(lpp, 1, max): over cellularities for each ploidy (row);
  // apply(lpp, 2, max): over ploidies for each cellularity (column).  max() keeps NaN.
  const int M = nP + nC;
  for (int k = tid; k < M; k += nthr) {
    double v = NEG_INF;
    bool nan = false;
    if (k < nP) {
      for (int c = 0; c < nC; ++c) {
        const double x = zn[(int64_t)c * nP + k];
        nan |= isnan(x);
        v = fmax(v, x);
      }
      if (nan) v = __longlong_as_double(0x7FF8000000000000LL);
      row_max[smp * nP + k] = v;
    } else {
      const int c = k - nP;
      for (int p = 0; p < nP; ++p) {
        const double x = zn[(int64_t)c * nP + p];
        nan |= isnan(x);
        v = fmax(v, x);
      }
      if (nan) v = __longlong_as_double(0x7FF8000000000000LL);
      col_max[smp * nC + c] = v;
    }
    thr[k] = nan ? __longlong_as_double(0x7FF0000000000000LL) : v;  // NaN sorts last
    perm[k] = (unsigned short)k;
  }
  for (int k = M + tid; k < Mp; k += nthr) {
    thr[k] = __longlong_as_double(0x7FF0000000000000LL);
    perm[k] = (unsigned short)0xFFFF;
  }
  __syncthreads();
  // bitonic sort of (thr, perm), ascending
  for (int size = 2; size <= Mp; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < Mp; i += nthr) {
        const int j = i ^ stride;
        if (j > i) {
          const bool up = (i & size) == 0;
          const double a = thr[i], b = thr[j];
          if ((a > b) == up) {
            thr[i] = b;
            thr[j] = a;
            const unsigned short t = perm[i];
            perm[i] = perm[j];
            perm[j] = t;
          }
        }
      }
      __syncthreads();
    }
  for (int k = tid; k <= Mp; k += nthr) bins[k] = 0ull;
  __syncthreads();
  // every element joins the bin "greater than exactly k of the sorted maxima"
  const double SCALE = 4611686018427387904.0;  // 2^62
  for (int64_t i = tid; i < G; i += nthr) {
    const double x = zn[i];
    if (!(x > 0.0)) continue;  // zeros and NaN carry no mass
    int lo = 0, hi = M;        // number of thr[] strictly below x
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (thr[mid] < x) lo = mid + 1;
      else hi = mid;
    }
    atomicAdd(&bins[lo], (unsigned long long)(x * SCALE + 0.5));
  }
  __syncthreads();
  // mass strictly above sorted maximum i = sum of bins[k], k > i  (suffix sum, one thread)
  if (tid == 0) {
    unsigned long long run = 0ull;
    for (int k = M; k >= 1; --k) {
      run += bins[k];
      bins[k] = run;
    }
  }
  __syncthreads();
  for (int i = tid; i < Mp; i += nthr) {
    const int k = perm[i];
    if (k == 0xFFFF) continue;  // padding of the sort
    // equal maxima share the mass above the LAST of them in sorted order
    int e = i;
    while (e + 1 < Mp && thr[e + 1] == thr[i]) ++e;
    const double above = (double)bins[e + 1] / SCALE;
    if (k < nP) row_above[smp * nP + k] = above;
    else col_above[smp * nC + (k - nP)] = above;
  }
}

int fit_finish_launch(const double* d_lpp_sum, int64_t S, int64_t nP, int64_t nC,
                      double* d_znorm, double* d_row_max, double* d_col_max,
                      double* d_row_above, double* d_col_above, double* d_lsl,
                      cudaStream_t st) {
  if (S < 0 || nP < 1 || nC < 1) return fail(SQZ_ERR_ARG, "need S >= 0, nP >= 1, nC >= 1");
  if (nP + nC > kMaxMarginals)
    return fail(SQZ_ERR_SIZE, "sqz_fit_finish handles at most 4096 ploidy + cellularity values");
  if (nP * nC > 0x7fffffffLL) return fail(SQZ_ERR_SIZE, "grid too large");
  if (S == 0) return SQZ_OK;
  int Mp = 1;
  while (Mp < nP + nC) Mp <<= 1;
  const size_t smb = (size_t)Mp * 8 + (size_t)(Mp + 1) * 8 + (size_t)Mp * 2 + 16;
  SQZ_CUDA(cudaFuncSetAttribute(fit_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                100 * 1024));
  fit_finish_kernel<<<(unsigned)S, kFinishThreads, smb, st>>>(
      d_lpp_sum, (int)nP, (int)nC, d_znorm, d_row_max, d_col_max, d_row_above, d_col_above,
      d_lsl);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

}  // namespace sqz
