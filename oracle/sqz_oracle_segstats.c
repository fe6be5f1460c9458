/*
 * sqz_oracle_segstats.c -- CPU restatement of the per-segment statistics of
 * segment.breaks (R/segments.R:113-198) and b_allele_freq (R/model.R:69-98).
 *
 * TEST INFRASTRUCTURE ONLY (see the header of sqz_oracle.c).  PARITY STATUS: "parity
 * unpinned": the reference's tests never call segment.breaks and R is not installed; the
 * restatement is cross-checked against an independent numpy/FFT restatement of R's
 * density.default (tests/test_oracle_segstats.py).
 *
 * Third-party arithmetic not under /root/reference (GNU R stats, version unpinned):
 *   cut.default / .bincode (right = TRUE, include.lowest = FALSE): interval k holds
 *       b[k] < x <= b[k+1];
 *   sum (long double), sd = sqrt(var): cov.c two-pass with a refined long-double mean,
 *       denominator n - 1, NA for n < 2;
 *   quantile type 7, IQR, bw.nrd0: 0.9 * min(sd, IQR / 1.34) * n^-0.2 with its fall-backs;
 *   density.default (gaussian, n = 512, cut = 3): linear binning onto 2n points (BinDist),
 *       convolution with dnorm at kords = seq(0, 2 (up - lo), length.out = 2n) (R does it
 *       by FFT; the circular correlation is summed directly here), pmax(0, .), approx() onto
 *       seq(from, to, length.out = 512);
 *   dnorm (R >= 3.1: split exponent beyond |x| = 5).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SEG_NA (sqzo_na_real_seg())
static double sqzo_na_real_seg(void) {
  union { double d; uint64_t u; } v;
  v.u = 0x7FF00000000007A2ULL;
  return v.d;
}

/* stats::dnorm(x, 0, sigma), not log (nmath/dnorm.c) */
static double r_dnorm(double x, double sigma) {
  const double M_1_SQRT_2PI_ = 0.398942280401432677939946059934;
  if (isnan(x) || isnan(sigma)) return x + sigma;
  if (sigma < 0) return NAN;
  if (!isfinite(sigma)) return 0.;
  if (!isfinite(x)) return 0.;
  if (sigma == 0) return (x == 0) ? INFINITY : 0.;
  x = fabs(x / sigma);
  if (x >= 2 * sqrt(DBL_MAX)) return 0.;
  if (x < 5) return M_1_SQRT_2PI_ * exp(-0.5 * x * x) / sigma;
  if (x > sqrt(-2 * M_LN2 * (DBL_MIN_EXP + 1 - DBL_MANT_DIG))) return 0.;
  double x1 = ldexp(nearbyint(ldexp(x, 16)), -16);
  double x2 = x - x1;
  return M_1_SQRT_2PI_ / sigma * (exp(-0.5 * x1 * x1) * exp((-0.5 * x2 - x1) * x2));
}

/* stats::sd over the non-NA values of v[0..n): cov.c's two-pass variance */
static double r_sd(const double* v, int64_t n) {
  long double sum = 0.0L;
  int64_t m = 0;
  for (int64_t i = 0; i < n; i++)
    if (!isnan(v[i])) { sum += v[i]; m++; }
  if (m < 2) return SEG_NA;
  long double mean = sum / m, t = 0.0L;
  for (int64_t i = 0; i < n; i++)
    if (!isnan(v[i])) t += (v[i] - mean);
  mean += t / m;
  long double ss = 0.0L;
  for (int64_t i = 0; i < n; i++)
    if (!isnan(v[i])) { long double d = v[i] - mean; ss += d * d; }
  return sqrt((double)(ss / (m - 1)));
}

/* base::mean over the non-NA values (summary.c: long double sum, one refinement pass) */
static double r_mean(const double* v, int64_t n) {
  long double sum = 0.0L;
  int64_t m = 0;
  for (int64_t i = 0; i < n; i++)
    if (!isnan(v[i])) { sum += v[i]; m++; }
  if (m == 0) return NAN;
  long double mean = sum / m, t = 0.0L;
  for (int64_t i = 0; i < n; i++)
    if (!isnan(v[i])) t += (v[i] - mean);
  return (double)(mean + t / m);
}

static int cmp_double(const void* a, const void* b) {
  double x = *(const double*)a, y = *(const double*)b;
  return (x > y) - (x < y);
}

/* quantile(x, p, type = 7) on sorted x */
static double r_quantile7(const double* xs, int64_t n, double p) {
  double index = 1 + (double)(n - 1 > 0 ? n - 1 : 0) * p;
  double lo = floor(index), hi = ceil(index);
  double qs = xs[(int64_t)lo - 1];
  double xh = xs[(int64_t)hi - 1];
  if (index > lo && xh != qs) {
    double h = index - lo;
    qs = (1 - h) * qs + h * xh;
  }
  return qs;
}

/* bw.nrd0 (stats/R/bandwidths.R) */
static double r_bw_nrd0(const double* x, const double* xs, int64_t n) {
  double hi = r_sd(x, n);
  double iqr = r_quantile7(xs, n, 0.75) - r_quantile7(xs, n, 0.25);
  double lo = fmin(hi, iqr / 1.34);
  if (!(lo != 0)) {  /* if(!(lo <- min(hi, IQR/1.34))) */
    lo = hi;
    if (!(lo != 0)) {
      lo = fabs(x[0]);
      if (!(lo != 0)) lo = 1.;
    }
  }
  return 0.9 * lo * pow((double)n, -0.2);
}

/* Mode of the weighted kernel density of c(Bf, Af) (R/model.R:69-85), folded to <= 0.5
 * (R/model.R:86-88).  m het positions; returns NA for m = 0, Bf for m = 1.
 * If dens512 is non-null it receives density()'s 512 y values (for cross-checks). */
double sqzo_b_allele_freq_mode(const double* Af, const double* Bf, const double* good_reads,
                               int64_t m, double* dens512) {
  if (m == 0) return SEG_NA;
  if (m == 1) return Bf[0];
  const int n = 512;
  const int64_t N = 2 * m;
  double* x = (double*)malloc(sizeof(double) * N * 3);
  double* xs = x + N;
  double* w = xs + N;
  long double gsum = 0.0L;
  for (int64_t i = 0; i < m; i++) gsum += good_reads[i];
  const double two_sum = 2 * (double)gsum;
  for (int64_t i = 0; i < m; i++) {
    x[i] = Bf[i];
    x[m + i] = Af[i];
    w[i] = w[m + i] = good_reads[i] / two_sum;
  }
  memcpy(xs, x, sizeof(double) * N);
  qsort(xs, N, sizeof(double), cmp_double);
  const double bw = r_bw_nrd0(x, xs, N);
  const double from = xs[0] - 3 * bw, to = xs[N - 1] + 3 * bw;
  const double lo = from - 4 * bw, up = to + 4 * bw;
  /* BinDist (stats/src/massdist.c) */
  double y[1024];
  memset(y, 0, sizeof y);
  {
    const int ixmin = 0, ixmax = n - 2;
    const double xdelta = (up - lo) / (n - 1);
    for (int64_t i = 0; i < N; i++) {
      if (isfinite(x[i])) {
        double xpos = (x[i] - lo) / xdelta;
        int ix = (int)floor(xpos);
        double fx = xpos - ix, wi = w[i];
        if (ixmin <= ix && ix <= ixmax) {
          y[ix] += wi * (1 - fx);
          y[ix + 1] += wi * fx;
        } else if (ix == -1) {
          y[0] += wi * fx;
        } else if (ix == ixmax + 1) {
          y[ix] += wi * (1 - fx);
        }
      }
    }
  }
  /* kords <- seq.int(0, 2*(up-lo), length.out = 2n); kords[(n+2):(2n)] <- -kords[n:2];
   * kords <- dnorm(kords, sd = bw); then the circular correlation with y, / length(y) is
   * the FFT's own normalisation and cancels */
  double K[1024];
  {
    const double by = 2 * (up - lo) / (2 * n - 1);
    for (int d = 0; d <= n; d++) K[d] = r_dnorm(d == 2 * n - 1 ? 2 * (up - lo) : d * by, bw);
    for (int d = n + 1; d < 2 * n; d++) K[d] = K[2 * n - d];
  }
  double kords[512], xords[512];
  for (int j = 0; j < n; j++) {
    long double acc = 0.0L;
    for (int mm = 0; mm < 2 * n; mm++)
      if (y[mm] != 0) acc += (long double)y[mm] * K[((mm - j) % (2 * n) + 2 * n) % (2 * n)];
    double v = (double)acc;
    kords[j] = v > 0 ? v : 0;
  }
  {
    const double by = (up - lo) / (n - 1);
    for (int j = 0; j < n; j++) xords[j] = j == n - 1 ? up : lo + j * by;
  }
  /* x <- seq.int(from, to, length.out = 512); y <- approx(xords, kords, x)$y */
  double gx[512], gy[512];
  {
    const double by = (to - from) / (n - 1);
    for (int j = 0; j < n; j++) {
      const double v = j == n - 1 ? to : from + j * by;
      gx[j] = v;
      int i = 0, jj = n - 1;
      if (v < xords[i] || v > xords[jj]) { gy[j] = NAN; continue; }
      while (i < jj - 1) {
        int ij = (i + jj) / 2;
        if (v < xords[ij]) jj = ij; else i = ij;
      }
      if (v == xords[jj]) gy[j] = kords[jj];
      else if (v == xords[i]) gy[j] = kords[i];
      else gy[j] = kords[i] + (kords[jj] - kords[i]) * ((v - xords[i]) / (xords[jj] - xords[i]));
    }
  }
  if (dens512) memcpy(dens512, gy, sizeof gy);
  /* points.max <- which(diff(sign(diff(dd$y))) == -2) + 1, else == -1 (R/model.R:73-76);
   * b.val <- l.max[which.max(dd$y[dd$x %in% l.max])]: the highest local maximum, first on
   * ties.  One departure, stated in DESIGN.md: a maximum whose top is a run of exactly equal
   * values (c(Bf, Af) is symmetric about 0.5 when Af = 1 - Bf, and the 512-point grid has no
   * point at 0.5, so the two central values are equal) counts as a maximum at the first
   * point of the run.  R's FFT leaves rounding noise of 1e-16 on the two values, so its -2
   * rule finds one of them -- after folding the same number to rounding -- whereas the
   * noise-free sum here would fall through to a spurious maximum in the tail. */
  int best = -1, cand = -1;
  for (int j = 0; j + 1 < n; j++) {
    const double d = gy[j + 1] - gy[j];
    if (d > 0) {
      cand = j + 1;
    } else if (d < 0) {
      if (cand >= 0 && (best < 0 || gy[cand] > gy[best])) best = cand;
      cand = -1;
    }
  }
  if (best < 0) {  /* R/model.R:74-76: the == -1 rule (a shoulder at either end) */
    for (int j = 1; j < n - 1; j++) {
      const double d0 = gy[j] - gy[j - 1], d1 = gy[j + 1] - gy[j];
      const int s0 = (d0 > 0) - (d0 < 0), s1 = (d1 > 0) - (d1 < 0);
      if (s1 - s0 == -1 && (best < 0 || gy[j] > gy[best])) best = j;
    }
  }
  double b_val = best >= 0 ? gx[best] : SEG_NA;
  free(x);
  if (best >= 0 && b_val > 0.5) b_val = 1 - b_val;
  return b_val;
}

/* Per-segment statistics of one genome (R/segments.R:113-198).
 * Rows of chromosome c are [row_off[c], row_off[c+1]), its breaks.vect (2 values per
 * segment: the output of unique.breaks, sorted) is brk[2 * seg_off[c] .. 2 * seg_off[c+1]).
 * out is n_seg_total x 6 column-major: Bf, N.BAF, sd.BAF, depth.ratio, N.ratio, sd.ratio
 * (R/segments.R:164-170).  Rows need not be sorted by position. */
int sqzo_segment_stats(int64_t n_chr, const int64_t* row_off, const int64_t* seg_off,
                       const double* position, const double* adjusted_ratio,
                       const double* depth_normal, const double* good_reads, const double* Af,
                       const double* Bf, const unsigned char* is_het, const double* brk,
                       double min_reads_baf, int weighted_mean, double* out) {
  const int64_t S = seg_off[n_chr];
  for (int64_t c = 0; c < n_chr; c++) {
    const int64_t r0 = row_off[c], r1 = row_off[c + 1], n = r1 - r0;
    double* v = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1) * 5);
    double *vb = v + n, *hA = vb + n, *hB = hA + n, *hG = hB + n;
    for (int64_t s = seg_off[c]; s < seg_off[c + 1]; s++) {
      const double b0 = brk[2 * s], b1 = brk[2 * s + 1];  /* (b0, b1]: cut(), right = TRUE */
      long double srw = 0.0L, swr = 0.0L;
      int64_t nr = 0, nb = 0, nv = 0;
      for (int64_t i = r0; i < r1; i++) {
        if (!(position[i] > b0 && position[i] <= b1)) continue;
        nr++;
        const double a = adjusted_ratio[i];
        if (weighted_mean) {
          const double wr = sqrt(depth_normal[i]), rw = a * wr;
          if (!isnan(rw)) srw += rw;
          if (!isnan(wr)) swr += wr;
          v[nv++] = rw / wr;
        } else {
          v[nv++] = a;
        }
        if (is_het[i] && good_reads[i] >= min_reads_baf) {
          if (weighted_mean) {
            const double wb = sqrt(good_reads[i]);
            vb[nb] = (Bf[i] * wb) / wb;
          } else {
            vb[nb] = Bf[i];
          }
          hA[nb] = Af[i];
          hB[nb] = Bf[i];
          hG[nb] = good_reads[i];
          nb++;
        }
      }
      out[0 * S + s] = sqzo_b_allele_freq_mode(hA, hB, hG, nb, NULL);
      out[1 * S + s] = (double)nb;
      out[2 * S + s] = r_sd(vb, nb);
      out[3 * S + s] = weighted_mean ? (double)srw / (double)swr : r_mean(v, nv);
      out[4 * S + s] = (double)nr;
      out[5 * S + s] = r_sd(v, nv);
    }
    free(v);
  }
  return 0;
}
