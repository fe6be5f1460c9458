// sqz_common.cuh -- shared host/device pieces of libsequenza_b200 (sm_100a).
//
// Everything here is fp64.  The translation units are compiled with -fmad=false so
// that the "faithful" code below rounds after every multiply and add, the way R's
// nmath does on x86-64; the fast grid kernel asks for fused operations explicitly
// with fma().
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include <string>

namespace sqz {

// ---- error plumbing (host) ------------------------------------------------------
int fail(int code, const std::string& msg);

#define SQZ_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      char b__[512];                                                                \
      snprintf(b__, sizeof b__, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__, \
               cudaGetErrorString(e__));                                            \
      return ::sqz::fail(SQZ_ERR_CUDA, b__);                                        \
    }                                                                               \
  } while (0)

constexpr int kMaxCN = 64;       // distinct copy numbers (CNt.max - CNt.min + 1)
constexpr int kMaxTypes = 2112;  // (CNt, B) or (CNt, Mt) rows for 64 copy numbers

// R's NA_real_ is the NaN with low word 1954; R prints any other NaN as "NaN".
__host__ __device__ inline double na_real() {
#ifdef __CUDA_ARCH__
  return __longlong_as_double(0x7FF00000000007A2LL);
#else
  union { double d; uint64_t u; } v;
  v.u = 0x7FF00000000007A2ULL;
  return v.d;
#endif
}

#define SQZ_LN_SQRT_2PI 0.918938533204672741780329736406
#define SQZ_LN_2PI 1.837877066409345483560659472811

// ---- R nmath, restated (stats::dt and stats::dbinom are what R/misc.R:3,
// R/bayes.R:5 and R/bayes.R:28 call; nmath is not part of the reference tree) ----

// stirlerr(n) for 2n = 0..30 (n <= 15), 40-digit values rounded by the compiler.
__host__ __device__ inline double sferr_half(int k) {
  switch (k) {
    case 1: return 0.1534264097200273452913839393;
    case 2: return 0.08106146679532725821967026359;
    case 3: return 0.05481412105191765389613870235;
    case 4: return 0.04134069595540929409382208141;
    case 5: return 0.03316287351993628748511050974;
    case 6: return 0.02767792568499833914878929275;
    case 7: return 0.02374616365629749597133027909;
    case 8: return 0.02079067210376509311152277177;
    case 9: return 0.01848845053267318523077935748;
    case 10: return 0.01664469118982119216319486537;
    case 11: return 0.01513497322191737887351383688;
    case 12: return 0.01387612882307074799874572702;
    case 13: return 0.01281046524292022692425065528;
    case 14: return 0.01189670994589177009505572412;
    case 15: return 0.0111045597582069173266307552;
    case 16: return 0.01041126526197209649747856713;
    case 17: return 0.009799416126158803298390373402;
    case 18: return 0.009255462182712732917728636633;
    case 19: return 0.008768700134139385462955047269;
    case 20: return 0.00833056343336287125646931866;
    case 21: return 0.007934114564314020547249562491;
    case 22: return 0.007573675487951840794972024212;
    case 23: return 0.007244554301320383179546196602;
    case 24: return 0.006942840107209529865664152663;
    case 25: return 0.006665247032707682442356180895;
    case 26: return 0.006408994188004207068439631083;
    case 27: return 0.006171712263039457647534604798;
    case 28: return 0.005951370112758847735624416046;
    case 29: return 0.005746216513010115682026102477;
    case 30: return 0.00555473355196280137103868996;
    default: return 0.0;
  }
}

__host__ __device__ inline double stirlerr(double n) {
  const double S0 = 0.083333333333333333333;
  const double S1 = 0.00277777777777777777778;
  const double S2 = 0.00079365079365079365079365;
  const double S3 = 0.000595238095238095238095238;
  const double S4 = 0.0008417508417508417508417508;
  if (n <= 15.0) {
    double nn = n + n;
    if (nn == (double)(int)nn) return sferr_half((int)nn);
    return lgamma(n + 1.) - (n + 0.5) * log(n) + n - SQZ_LN_SQRT_2PI;
  }
  double nn = n * n;
  if (n > 500) return (S0 - S1 / nn) / n;
  if (n > 80) return (S0 - (S1 - S2 / nn) / nn) / n;
  if (n > 35) return (S0 - (S1 - (S2 - S3 / nn) / nn) / nn) / n;
  return (S0 - (S1 - (S2 - (S3 - S4 / nn) / nn) / nn) / nn) / n;
}

__host__ __device__ inline double bd0(double x, double np) {
  if (!isfinite(x) || !isfinite(np) || np == 0.0) return NAN;
  if (fabs(x - np) < 0.1 * (x + np)) {
    double v = (x - np) / (x + np);
    double s = (x - np) * v;
    if (fabs(s) < 2.2250738585072014e-308) return s;
    double ej = 2 * x * v;
    v = v * v;
    for (int j = 1; j < 1000; j++) {
      ej *= v;
      double s1 = s + ej / ((j << 1) + 1);
      if (s1 == s) return s1;
      s = s1;
    }
  }
  return x * log(x / np) + np - x;
}

// The x-independent part of dt(x, 5, log = TRUE):
//   t = -bd0(n/2, (n+1)/2) + stirlerr((n+1)/2) - stirlerr(n/2),  n = 5.
__host__ __device__ inline double dt5_const() {
  return -bd0(2.5, 3.0) + stirlerr(3.0) - stirlerr(2.5);
}

// dt(x, df = 5, log = TRUE), central t, operation order of nmath dt.c.
__host__ __device__ inline double dt5_log(double x, double t5) {
  const double n = 5.0;
  if (isnan(x)) return x + n;
  if (!isfinite(x)) return -INFINITY;
  double x2n = x * x / n, l_x2n, u;
  if (x2n > 1. / 2.220446049250313e-16) {
    double ax = fabs(x);
    l_x2n = log(ax) - log(n) / 2.;
    u = n * l_x2n;
  } else if (x2n > 0.2) {
    l_x2n = log(1 + x2n) / 2.;
    u = n * l_x2n;
  } else {
    l_x2n = log1p(x2n) / 2.;
    u = -bd0(n / 2., (n + x * x) / 2.) + x * x / 2.;
  }
  return t5 - u - (SQZ_LN_SQRT_2PI + l_x2n);
}

// dbinom_raw(x, n, p, q, log = FALSE) with the three stirlerr terms and lf hoisted by
// the caller (they depend on x and n only, not on the model type).
struct BinomHoist {
  double x, n;     // forced to integers
  double lc0;      // stirlerr(n) - stirlerr(x) - stirlerr(n-x)
  double half_lf;  // 0.5 * (log(2 pi) + log(x) + log1p(-x/n))
  int kind;        // 0: general 0 < x < n; 1: x == 0; 2: x == n; 3: density 0; 4: NaN
  double special;  // value for kind 3/4
};

__host__ __device__ inline bool r_nonint(double x) {
  return fabs(x - nearbyint(x)) > 1e-7 * fmax(1., fabs(x));
}

// The x/n-only prologue of dbinom(x, n, p): argument checks of dbinom.c, then the
// pieces of dbinom_raw that do not involve p.
__host__ __device__ inline BinomHoist binom_hoist(double x, double n) {
  BinomHoist h;
  h.x = x;
  h.n = n;
  h.lc0 = 0;
  h.half_lf = 0;
  h.special = 0;
  if (isnan(x) || isnan(n)) {
    h.kind = 4;
    h.special = x + n;
    return h;
  }
  if (n < 0 || r_nonint(n)) {
    h.kind = 4;
    h.special = NAN;
    return h;
  }
  if (r_nonint(x) || x < 0 || !isfinite(x)) {
    h.kind = 3;
    return h;
  }
  n = nearbyint(n);
  x = nearbyint(x);
  h.x = x;
  h.n = n;
  if (x == 0) {
    h.kind = 1;
  } else if (x == n) {
    h.kind = 2;
  } else if (x < 0 || x > n) {
    h.kind = 3;
  } else {
    h.kind = 0;
    h.lc0 = stirlerr(n) - stirlerr(x) - stirlerr(n - x);
    h.half_lf = 0.5 * (SQZ_LN_2PI + log(x) + log1p(-x / n));
  }
  return h;
}

// dbinom(x, n, p, log = FALSE) given the hoisted part.
__host__ __device__ inline double binom_finish(const BinomHoist& h, double p) {
  if (h.kind == 4) return isnan(p) ? h.special + p : h.special;
  if (isnan(p)) return h.x + h.n + p;
  if (p < 0 || p > 1) return NAN;
  if (h.kind == 3) return 0.;
  double q = 1 - p, x = h.x, n = h.n;
  if (p == 0) return (x == 0) ? 1. : 0.;
  if (q == 0) return (x == n) ? 1. : 0.;
  if (h.kind == 1) {
    if (n == 0) return 1.;
    double lc = (p < 0.1) ? -bd0(n, n * q) - n * p : n * log(q);
    return exp(lc);
  }
  if (h.kind == 2) {
    double lc = (q < 0.1) ? -bd0(n, n * p) - n * q : n * log(p);
    return exp(lc);
  }
  double lc = h.lc0 - bd0(x, n * p) - bd0(n - x, n * q);
  return exp(lc - h.half_lf);
}

// ---- model points (R/model.R:1-20), operation order as written there ------------
__host__ __device__ inline double theoretical_depth_ratio(double CNt, double cellularity,
                                                          double ploidy, double CNn,
                                                          double avg_depth_ratio) {
  double cellu_copy_term = (1 - cellularity) + (CNt / CNn * cellularity);
  double ploidy_cellu_term = (ploidy / 2.0 * cellularity) + 1 - cellularity;
  return avg_depth_ratio * cellu_copy_term / ploidy_cellu_term;
}

__host__ __device__ inline double theoretical_baf(double CNt, double B, double cellularity,
                                                  double CNn) {
  if (CNn <= 1) return na_real();
  return ((B * cellularity) + (1 - cellularity)) /
         ((CNt * cellularity) + CNn * (1 - cellularity));
}

__host__ __device__ inline double theoretical_mufreq(double CNt, double Mt,
                                                     double cellularity, double CNn) {
  double normal_alleles = (CNt - Mt) * cellularity + CNn * (1 - cellularity);
  double all_alleles = (CNt * cellularity) + CNn * (1 - cellularity);
  return 1 - (normal_alleles / all_alleles);
}

// ---- double-double accumulation (stands in for R's long double sum()) -----------
struct DD {
  double hi, lo, special;  // special collects +-Inf / NaN terms, added plainly
};
__host__ __device__ inline DD dd_zero() { return DD{0., 0., 0.}; }
__host__ __device__ inline void dd_add(DD& a, double x) {
  if (!isfinite(x)) {
    a.special += x;
    return;
  }
  double s = a.hi + x;
  double bb = s - a.hi;
  double err = (a.hi - (s - bb)) + (x - bb);
  a.hi = s;
  a.lo += err;
}
__host__ __device__ inline void dd_merge(DD& a, const DD& b) {
  dd_add(a, b.hi);
  a.lo += b.lo;
  a.special += b.special;
}
__host__ __device__ inline double dd_value(const DD& a) {
  double v = a.hi + a.lo;
  return (a.special != 0.0) ? v + a.special : v;  // NaN != 0 is true
}

// ---- model description shared by the kernels (passed by value; lives in the
// kernel-parameter constant bank) --------------------------------------------------
struct CnModel {
  int CNt_min;
  int nc;              // number of distinct copy numbers
  int CNn;
  int T;               // number of types (rows of the types matrix)
  int toff[kMaxCN + 1];  // first type row of each copy number
  double prior[kMaxCN];   // normalised prior of each copy number's rows
  double lprior[kMaxCN];  // log(prior)
  double w[kMaxCN];       // prior^(-1/3): fast-path weight (baf only)
};

}  // namespace sqz
