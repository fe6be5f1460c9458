/*
 * sqz_oracle.c -- CPU restatement of sequenza 3.0.0's purity/ploidy likelihood path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in sequenza_b200/ (the product) may link, import
 * or call this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and only as the checker or as the timed CPU baseline.
 *
 * PARITY STATUS: "parity unpinned" against a live R run.  The reference is pure R and R
 * is not installed in this image, so oracle/_ref cannot be built.  What this oracle IS
 * pinned against (tests/test_oracle_golden.py):
 *   - tests/testthat/test.model.R:4-55      theoretical.* values, both type matrices
 *   - inst/doc/sequenza.html                six rendered baf.bayes rows (CNt/A/B exact,
 *                                           LPP at the 3-dp input tolerance, 5e-3)
 *   - data/CP.example.RData                 layout of list(ploidy, cellularity, lpp) and
 *                                           the get.ci point estimate (ploidy 2, cell. 0.88)
 *   - scipy.stats.t / scipy.stats.binom / mpmath at 40 digits for dt and dbinom
 *
 * Third-party arithmetic that is NOT under /root/reference: GNU R's nmath (stats::dt,
 * stats::dbinom, round, sum), version unpinned by the reference (DESCRIPTION:24,
 * "R (>= 3.2.0)"; package built 2019-05).  The published algorithm of R 3.4 .. 4.3 is
 * restated here from its documentation and Loader's "Fast and accurate computation of
 * binomial probabilities" (2000): dt.c (central t, with the large-x^2/n branch),
 * dbinom.c (dbinom_raw), bd0.c (series + log form), stirlerr.c (half-integer table to
 * n = 15, then the 5-term asymptotic series), fround.c (digits = 0 -> rint, half-even),
 * summary.c rsum (long double accumulator, no na.rm).
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <stdatomic.h>
#include <unistd.h>

#define SQZO_LN_SQRT_2PI 0.918938533204672741780329736406
#define SQZO_LN_2PI 1.837877066409345483560659472811
#define SQZO_1_SQRT_2PI 0.398942280401432677939946059934

/* R's NA_real_: a NaN whose low word is 1954 (arithmetic.c R_NaReal). */
static double na_real(void) {
  union { double d; uint64_t u; } v;
  v.u = 0x7FF00000000007A2ULL;
  return v.d;
}
double sqzo_na_real(void) { return na_real(); }

/* ---- nmath: stirlerr(n) = log(n!) - log(sqrt(2 pi n) (n/e)^n)  (stirlerr.c) ---- */
static const double sferr_halves[31] = {
    0.0, /* n = 0: placeholder, never read */
    0.1534264097200273452913839393,   0.08106146679532725821967026359,
    0.05481412105191765389613870235,  0.04134069595540929409382208141,
    0.03316287351993628748511050974,  0.02767792568499833914878929275,
    0.02374616365629749597133027909,  0.02079067210376509311152277177,
    0.01848845053267318523077935748,  0.01664469118982119216319486537,
    0.01513497322191737887351383688,  0.01387612882307074799874572702,
    0.01281046524292022692425065528,  0.01189670994589177009505572412,
    0.0111045597582069173266307552,   0.01041126526197209649747856713,
    0.009799416126158803298390373402, 0.009255462182712732917728636633,
    0.008768700134139385462955047269, 0.00833056343336287125646931866,
    0.007934114564314020547249562491, 0.007573675487951840794972024212,
    0.007244554301320383179546196602, 0.006942840107209529865664152663,
    0.006665247032707682442356180895, 0.006408994188004207068439631083,
    0.006171712263039457647534604798, 0.005951370112758847735624416046,
    0.005746216513010115682026102477, 0.00555473355196280137103868996};

double sqzo_stirlerr(double n) {
  const double S0 = 0.083333333333333333333;        /* 1/12 */
  const double S1 = 0.00277777777777777777778;      /* 1/360 */
  const double S2 = 0.00079365079365079365079365;   /* 1/1260 */
  const double S3 = 0.000595238095238095238095238;  /* 1/1680 */
  const double S4 = 0.0008417508417508417508417508; /* 1/1188 */
  double nn;
  if (n <= 15.0) {
    nn = n + n;
    if (nn == (int)nn) return sferr_halves[(int)nn];
    return lgamma(n + 1.) - (n + 0.5) * log(n) + n - SQZO_LN_SQRT_2PI;
  }
  nn = n * n;
  if (n > 500) return (S0 - S1 / nn) / n;
  if (n > 80) return (S0 - (S1 - S2 / nn) / nn) / n;
  if (n > 35) return (S0 - (S1 - (S2 - S3 / nn) / nn) / nn) / n;
  return (S0 - (S1 - (S2 - (S3 - S4 / nn) / nn) / nn) / nn) / n;
}

/* ---- nmath: bd0(x, np) = x log(x/np) + np - x, series when x ~ np  (bd0.c) ---- */
double sqzo_bd0(double x, double np) {
  double ej, s, s1, v;
  int j;
  if (!isfinite(x) || !isfinite(np) || np == 0.0) return NAN;
  if (fabs(x - np) < 0.1 * (x + np)) {
    v = (x - np) / (x + np);
    s = (x - np) * v;
    if (fabs(s) < DBL_MIN) return s;
    ej = 2 * x * v;
    v = v * v;
    for (j = 1; j < 1000; j++) {
      ej *= v;
      s1 = s + ej / ((j << 1) + 1);
      if (s1 == s) return s1;
      s = s1;
    }
  }
  return x * log(x / np) + np - x;
}

/* ---- nmath: central Student t density (dt.c); this is what R/misc.R:3 reaches
 * because dt2's `ncp` is missing and stats::dt then takes the C_dt branch. ---- */
double sqzo_dt(double x, double n, int give_log) {
  if (isnan(x) || isnan(n)) return x + n;
  if (n <= 0) return NAN;
  if (!isfinite(x)) return give_log ? -INFINITY : 0.0;
  if (!isfinite(n)) {
    /* dnorm(x, 0, 1): not reachable from sequenza (df = 5) */
    double ax = fabs(x);
    return give_log ? -(SQZO_LN_SQRT_2PI + 0.5 * ax * ax)
                    : SQZO_1_SQRT_2PI * exp(-0.5 * ax * ax);
  }
  double u, t = -sqzo_bd0(n / 2., (n + 1) / 2.) + sqzo_stirlerr((n + 1) / 2.) -
                sqzo_stirlerr(n / 2.);
  double x2n = x * x / n, ax = 0., l_x2n;
  int lrg_x2n = (x2n > 1. / DBL_EPSILON);
  if (lrg_x2n) {
    ax = fabs(x);
    l_x2n = log(ax) - log(n) / 2.;
    u = n * l_x2n;
  } else if (x2n > 0.2) {
    l_x2n = log(1 + x2n) / 2.;
    u = n * l_x2n;
  } else {
    l_x2n = log1p(x2n) / 2.;
    u = -sqzo_bd0(n / 2., (n + x * x) / 2.) + x * x / 2.;
  }
  if (give_log) return t - u - (SQZO_LN_SQRT_2PI + l_x2n);
  double I_sqrt_ = (lrg_x2n ? sqrt(n) / ax : exp(-l_x2n));
  return exp(t - u) * SQZO_1_SQRT_2PI * I_sqrt_;
}

/* ---- nmath: dbinom_raw / dbinom (dbinom.c) ---- */
static double dbinom_raw(double x, double n, double p, double q, int give_log) {
  double lf, lc;
  if (p == 0) return (x == 0) ? (give_log ? 0. : 1.) : (give_log ? -INFINITY : 0.);
  if (q == 0) return (x == n) ? (give_log ? 0. : 1.) : (give_log ? -INFINITY : 0.);
  if (x == 0) {
    if (n == 0) return give_log ? 0. : 1.;
    lc = (p < 0.1) ? -sqzo_bd0(n, n * q) - n * p : n * log(q);
    return give_log ? lc : exp(lc);
  }
  if (x == n) {
    lc = (q < 0.1) ? -sqzo_bd0(n, n * p) - n * q : n * log(p);
    return give_log ? lc : exp(lc);
  }
  if (x < 0 || x > n) return give_log ? -INFINITY : 0.;
  lc = sqzo_stirlerr(n) - sqzo_stirlerr(x) - sqzo_stirlerr(n - x) - sqzo_bd0(x, n * p) -
       sqzo_bd0(n - x, n * q);
  lf = SQZO_LN_2PI + log(x) + log1p(-x / n);
  return give_log ? (lc - 0.5 * lf) : exp(lc - 0.5 * lf);
}

static int r_nonint(double x) {
  return fabs(x - nearbyint(x)) > 1e-7 * fmax(1., fabs(x));
}

double sqzo_dbinom(double x, double n, double p, int give_log) {
  if (isnan(x) || isnan(n) || isnan(p)) return x + n + p;
  if (p < 0 || p > 1 || n < 0 || r_nonint(n)) return NAN;
  if (r_nonint(x)) return give_log ? -INFINITY : 0.; /* R warns, returns 0 */
  if (x < 0 || !isfinite(x)) return give_log ? -INFINITY : 0.;
  n = nearbyint(n);
  x = nearbyint(x);
  return dbinom_raw(x, n, p, 1 - p, give_log);
}

/* R's round(x, 0): fround.c digits == 0 -> private_rint -> half-to-even. */
double sqzo_round0(double x) { return nearbyint(x); }

/* ---- R/model.R:1-7 ---- */
double sqzo_theoretical_depth_ratio(double CNt, double cellularity, double ploidy,
                                    double CNn, double normal_ploidy,
                                    double avg_depth_ratio) {
  double cellu_copy_term = (1 - cellularity) + (CNt / CNn * cellularity);
  double ploidy_cellu_term = (ploidy / normal_ploidy * cellularity) + 1 - cellularity;
  return avg_depth_ratio * cellu_copy_term / ploidy_cellu_term;
}

/* ---- R/model.R:9-14 ---- */
double sqzo_theoretical_baf(double CNt, double B, double cellularity, double CNn) {
  double baf = ((B * cellularity) + (1 - cellularity)) /
               ((CNt * cellularity) + CNn * (1 - cellularity));
  if (CNn <= 1) baf = na_real();
  return baf;
}

/* ---- R/model.R:16-20 ---- */
double sqzo_theoretical_mufreq(double CNt, double Mt, double cellularity, double CNn) {
  double normal_alleles = (CNt - Mt) * cellularity + CNn * (1 - cellularity);
  double all_alleles = (CNt * cellularity) + CNn * (1 - cellularity);
  return 1 - (normal_alleles / all_alleles);
}

/* ---- R/model.R:22-37: rows ordered CNt ascending, B = 0..trunc(CNt/2) (B = 0 only when
 * CNn < 2).  out is T x 3 column-major (CNn, CNt, B); returns T.  out may be NULL. ---- */
int sqzo_baf_types_matrix(int CNt_min, int CNt_max, int CNn, double* out) {
  int T = 0;
  for (int c = CNt_min; c <= CNt_max; c++) T += (CNn < 2) ? 1 : (c / 2 + 1);
  if (!out) return T;
  int t = 0;
  for (int c = CNt_min; c <= CNt_max; c++) {
    int nb = (CNn < 2) ? 1 : (c / 2 + 1);
    for (int b = 0; b < nb; b++, t++) {
      out[t] = CNn;
      out[T + t] = c;
      out[2 * T + t] = b;
    }
  }
  return T;
}

/* ---- R/model.R:39-47: rows CNt ascending, Mt = 0..CNt ---- */
int sqzo_mufreq_types_matrix(int CNt_min, int CNt_max, int CNn, double* out) {
  int T = 0;
  for (int c = CNt_min; c <= CNt_max; c++) T += c + 1;
  if (!out) return T;
  int t = 0;
  for (int c = CNt_min; c <= CNt_max; c++)
    for (int m = 0; m <= c; m++, t++) {
      out[t] = CNn;
      out[T + t] = c;
      out[2 * T + t] = m;
    }
  return T;
}

/* R's which.max: first maximum, NaN skipped; -1 when nothing is comparable. */
static int which_max(const double* v, int n) {
  int idx = -1;
  double s = 0;
  for (int i = 0; i < n; i++)
    if (!isnan(v[i]) && (idx < 0 || v[i] > s)) {
      s = v[i];
      idx = i;
    }
  return idx;
}

typedef struct {
  int T;
  double *CNt, *A, *B, *lratio, *baf, *lprior, *prior;
} baf_model_t;

/* R/bayes.R:46-59 + :62-63: types, model points, priors (value per CNt, normalised over
 * the T types), test.ratio = log(model depth.ratio). */
static int baf_model_build(baf_model_t* m, double cellularity, double ploidy, double adr,
                           int CNt_min, int CNt_max, int CNn, int n_prior,
                           const double* prior_CN, const double* prior_value) {
  int T = sqzo_baf_types_matrix(CNt_min, CNt_max, CNn, NULL);
  double* ty = (double*)malloc(sizeof(double) * 3 * T);
  double* buf = (double*)malloc(sizeof(double) * 7 * T);
  if (!ty || !buf) return -1;
  sqzo_baf_types_matrix(CNt_min, CNt_max, CNn, ty);
  m->T = T;
  m->CNt = buf;
  m->A = buf + T;
  m->B = buf + 2 * T;
  m->lratio = buf + 3 * T;
  m->baf = buf + 4 * T;
  m->lprior = buf + 5 * T;
  m->prior = buf + 6 * T;
  double psum = 0;
  for (int t = 0; t < T; t++) {
    double c = ty[T + t], b = ty[2 * T + t];
    m->CNt[t] = c;
    m->A[t] = c - b;
    m->B[t] = b;
    m->lratio[t] =
        log(sqzo_theoretical_depth_ratio(c, cellularity, ploidy, CNn, 2., adr));
    m->baf[t] = sqzo_theoretical_baf(c, b, cellularity, CNn);
    double pr = 1.;
    for (int i = 0; i < n_prior; i++)
      if (c == prior_CN[i]) pr = prior_value[i];
    m->prior[t] = pr;
    psum += pr; /* R sum(): long double, but T <= a few hundred small values */
  }
  for (int t = 0; t < T; t++) {
    m->prior[t] = m->prior[t] / psum;
    m->lprior[t] = log(m->prior[t]);
  }
  free(ty);
  return 0;
}

/* R/bayes.R:61-94 for one segment.  out4 = CNt, A, B, LPP.  margin (optional) = gap
 * between the winning posterior and the best posterior of any OTHER (CNt,B) row the
 * rule could have picked; used by the tests to tell a real mismatch from a near-tie. */
static void baf_bayes_one(const baf_model_t* m, double Bf, double lr, double sdBf,
                          double sdr, double wBf, double wr, int ratio_priority,
                          double* score_r, double* post, double* out4, double* margin) {
  int T = m->T;
  double sr = sdr / sqrt(wr);
  for (int t = 0; t < T; t++)
    score_r[t] = sqzo_dt((m->lratio[t] - lr) / sr, 5., 1) + m->lprior[t];
  double sb = sdBf / sqrt(wBf);
  if (!isnan(Bf) && !isnan(sb)) {
    for (int t = 0; t < T; t++)
      post[t] = score_r[t] + sqzo_dt((m->baf[t] - Bf) / sb, 5., 1);
  } else {
    for (int t = 0; t < T; t++) post[t] = score_r[t];
  }
  int k = -1;
  double gap = INFINITY;
  if (!ratio_priority) {
    k = which_max(post, T);
    if (k >= 0 && margin)
      for (int t = 0; t < T; t++)
        if (t != k && !isnan(post[t]) && post[k] - post[t] < gap) gap = post[k] - post[t];
  } else {
    int kr = which_max(score_r, T);
    if (kr >= 0) {
      double cn = m->CNt[kr];
      double s = 0;
      for (int t = 0; t < T; t++)
        if (m->CNt[t] == cn && !isnan(post[t]) && (k < 0 || post[t] > s)) {
          s = post[t];
          k = t;
        }
      if (k >= 0 && margin) {
        for (int t = 0; t < T; t++) {
          if (m->CNt[t] == cn && t != k && !isnan(post[t]) && post[k] - post[t] < gap)
            gap = post[k] - post[t];
          if (m->CNt[t] != cn && !isnan(score_r[t]) && score_r[kr] - score_r[t] < gap)
            gap = score_r[kr] - score_r[t];
        }
      }
    }
  }
  if (k < 0) {
    /* the reference stops with an R error here (as.numeric of a 0-row data.frame);
     * the restatement reports NA so a caller can see which segment was degenerate */
    out4[0] = out4[1] = out4[2] = na_real();
    out4[3] = NAN;
    if (margin) *margin = NAN;
    return;
  }
  out4[0] = m->CNt[k];
  out4[1] = m->A[k];
  out4[2] = m->B[k];
  out4[3] = post[k];
  if (isnan(Bf)) out4[1] = out4[2] = na_real();
  if (margin) *margin = gap;
}

/* ---- R/bayes.R:38-101 baf.bayes.  All per-segment vectors have length N (the caller
 * does R's scalar recycling).  out is N x 4 column-major (CNt, A, B, LPP). ---- */
int sqzo_baf_bayes(int64_t N, const double* Bf, const double* depth_ratio,
                   const double* sd_Bf, const double* sd_ratio, const double* w_Bf,
                   const double* w_ratio, double cellularity, double ploidy,
                   double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                   int n_prior, const double* prior_CN, const double* prior_value,
                   int ratio_priority, double* out, double* margin) {
  baf_model_t m;
  if (baf_model_build(&m, cellularity, ploidy, avg_depth_ratio, CNt_min, CNt_max, CNn,
                      n_prior, prior_CN, prior_value))
    return -1;
  double* tmp = (double*)malloc(sizeof(double) * 2 * m.T);
  for (int64_t s = 0; s < N; s++) {
    double o4[4];
    baf_bayes_one(&m, Bf[s], log(depth_ratio[s]), sd_Bf[s], sd_ratio[s], w_Bf[s],
                  w_ratio[s], ratio_priority, tmp, tmp + m.T, o4,
                  margin ? margin + s : NULL);
    out[s] = o4[0];
    out[N + s] = o4[1];
    out[2 * N + s] = o4[2];
    out[3 * N + s] = o4[3];
  }
  free(tmp);
  free(m.CNt);
  return 0;
}

/* Worker pool over grid points: stands in for the reference's mclapply fork pool
 * (R/bayes.R:114-115, :184-185).  One grid point per task, dynamic pick-up. */
typedef struct {
  int is_mufreq;
  int64_t N;
  const double *a, *b, *c, *d, *e, *f;
  const double *ploidy, *cellularity;
  int64_t nP;
  double adr;
  int CNt_min, CNt_max, CNn, n_prior;
  const double *prior_CN, *prior_value;
  int ratio_priority;
  int64_t g0, g1;
  double* out;
  atomic_llong next;
  atomic_int err;
} grid_job_t;

static void mufreq_bayes_one_fwd(const void* m, double CNn, double F, double ratio,
                                 double wm, double wr, double* t1, double* t2,
                                 double* o4);
static int muf_model_build_fwd(void* m, double c, double p, double adr, int lo, int hi,
                               int CNn, int np, const double* pc, const double* pv,
                               int* T, double** to_free);

static void* grid_worker(void* arg) {
  grid_job_t* j = (grid_job_t*)arg;
  for (;;) {
    int64_t g = (int64_t)atomic_fetch_add(&j->next, 1);
    if (g >= j->g1) break;
    int64_t ip = g % j->nP, ic = g / j->nP;
    long double acc = 0.0L;
    if (!j->is_mufreq) {
      baf_model_t m;
      if (baf_model_build(&m, j->cellularity[ic], j->ploidy[ip], j->adr, j->CNt_min,
                          j->CNt_max, j->CNn, j->n_prior, j->prior_CN, j->prior_value)) {
        atomic_store(&j->err, 1);
        continue;
      }
      double* tmp = (double*)malloc(sizeof(double) * 2 * m.T);
      for (int64_t s = 0; s < j->N; s++) {
        double o4[4];
        baf_bayes_one(&m, j->a[s], j->b[s], j->c[s], j->d[s], j->e[s], j->f[s],
                      j->ratio_priority, tmp, tmp + m.T, o4, NULL);
        acc += o4[3]; /* R sum(): long double accumulator, NaN kept (no na.rm) */
      }
      free(tmp);
      free(m.CNt);
    } else {
      char mbuf[128];
      int T;
      double* to_free;
      if (muf_model_build_fwd(mbuf, j->cellularity[ic], j->ploidy[ip], j->adr,
                              j->CNt_min, j->CNt_max, j->CNn, j->n_prior, j->prior_CN,
                              j->prior_value, &T, &to_free)) {
        atomic_store(&j->err, 1);
        continue;
      }
      double* tmp = (double*)malloc(sizeof(double) * 2 * T);
      for (int64_t s = 0; s < j->N; s++) {
        double o4[4];
        mufreq_bayes_one_fwd(mbuf, j->CNn, j->a[s], j->b[s], j->c[s], j->d[s], tmp,
                             tmp + T, o4);
        acc += o4[3];
      }
      free(tmp);
      free(to_free);
    }
    j->out[g - j->g0] = (double)acc;
  }
  return NULL;
}

static int grid_run(grid_job_t* j, int nthreads) {
  if (nthreads <= 0) nthreads = (int)sysconf(_SC_NPROCESSORS_ONLN);
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  atomic_store(&j->next, (long long)j->g0);
  atomic_store(&j->err, 0);
  pthread_t th[256];
  int started = 0;
  for (int i = 1; i < nthreads; i++)
    if (pthread_create(&th[started], NULL, grid_worker, j) == 0) started++;
  grid_worker(j);
  for (int i = 0; i < started; i++) pthread_join(th[i], NULL);
  return atomic_load(&j->err) ? -1 : 0;
}

/* ---- R/bayes.R:103-116: result$LPP, one sum(baf.bayes(...)[,4]) per grid point in
 * expand.grid(ploidy, cellularity) order (ploidy fastest).  R's sum() accumulates in
 * long double and does not drop NaN.  Grid points g0 <= g < g1 only (for timing a
 * bounded sample); lpp_sum has g1-g0 entries.  A pthread pool over grid points
 * (grid_run above) stands in for the reference's mclapply fork pool. ---- */
int sqzo_baf_model_fit_sums(int64_t N, const double* Bf, const double* depth_ratio,
                            const double* sd_Bf, const double* sd_ratio,
                            const double* w_Bf, const double* w_ratio,
                            const double* ploidy, int64_t nP, const double* cellularity,
                            int64_t nC, double avg_depth_ratio, int CNt_min, int CNt_max,
                            int CNn, int n_prior, const double* prior_CN,
                            const double* prior_value, int ratio_priority, int64_t g0,
                            int64_t g1, int nthreads, double* lpp_sum) {
  double* lr = (double*)malloc(sizeof(double) * (N > 0 ? N : 1));
  if (!lr) return -1;
  for (int64_t s = 0; s < N; s++) lr[s] = log(depth_ratio[s]);
  (void)nC;
  grid_job_t job = {0};
  job.is_mufreq = 0;
  job.N = N;
  job.a = Bf;
  job.b = lr;
  job.c = sd_Bf;
  job.d = sd_ratio;
  job.e = w_Bf;
  job.f = w_ratio;
  job.ploidy = ploidy;
  job.nP = nP;
  job.cellularity = cellularity;
  job.adr = avg_depth_ratio;
  job.CNt_min = CNt_min;
  job.CNt_max = CNt_max;
  job.CNn = CNn;
  job.n_prior = n_prior;
  job.prior_CN = prior_CN;
  job.prior_value = prior_value;
  job.ratio_priority = ratio_priority;
  job.g0 = g0;
  job.g1 = g1;
  job.out = lpp_sum;
  int rc = grid_run(&job, nthreads);
  free(lr);
  return rc;
}

typedef struct {
  int T;
  double *CNt, *Mt, *ratio, *mufreq, *prior;
} muf_model_t;

/* R/bayes.R:132-144 */
static int muf_model_build(muf_model_t* m, double cellularity, double ploidy, double adr,
                           int CNt_min, int CNt_max, int CNn, int n_prior,
                           const double* prior_CN, const double* prior_value) {
  int T = sqzo_mufreq_types_matrix(CNt_min, CNt_max, CNn, NULL);
  double* ty = (double*)malloc(sizeof(double) * 3 * T);
  double* buf = (double*)malloc(sizeof(double) * 5 * T);
  if (!ty || !buf) return -1;
  sqzo_mufreq_types_matrix(CNt_min, CNt_max, CNn, ty);
  m->T = T;
  m->CNt = buf;
  m->Mt = buf + T;
  m->ratio = buf + 2 * T;
  m->mufreq = buf + 3 * T;
  m->prior = buf + 4 * T;
  double psum = 0;
  for (int t = 0; t < T; t++) {
    double c = ty[T + t], mt = ty[2 * T + t];
    m->CNt[t] = c;
    m->Mt[t] = mt;
    m->mufreq[t] = sqzo_theoretical_mufreq(c, mt, cellularity, CNn);
    m->ratio[t] = sqzo_theoretical_depth_ratio(c, cellularity, ploidy, CNn, 2., adr);
    double pr = 1.;
    for (int i = 0; i < n_prior; i++)
      if (c == prior_CN[i]) pr = prior_value[i];
    m->prior[t] = pr;
    psum += pr;
  }
  for (int t = 0; t < T; t++) m->prior[t] /= psum;
  free(ty);
  return 0;
}

/* R/bayes.R:145-166 for one position; out4 = CNn, CNt, Mt, LPP */
static void mufreq_bayes_one(const muf_model_t* m, double CNn, double F, double ratio,
                             double w_mufreq, double w_ratio, double* score_r,
                             double* lpost, double* out4) {
  int T = m->T;
  /* depth.ratio.dbinom, R/bayes.R:25-29 */
  double kr = sqzo_round0(w_ratio * (ratio / (1 + ratio)));
  /* mufreq.dbinom, R/bayes.R:1-6 */
  double km = sqzo_round0(F * w_mufreq);
  const double min_offset = 1e-323;
  for (int t = 0; t < T; t++) {
    double prob = m->ratio[t] / (1 + m->ratio[t]);
    double sr = sqzo_dbinom(kr, w_ratio, prob, 0);
    double mm = m->mufreq[t];
    if (mm == 0) mm = 0.01;
    double sm = sqzo_dbinom(km, w_mufreq, mm, 0);
    sr = sr * m->prior[t];
    double post = sr * sm;
    if (post == 0) post = min_offset;
    score_r[t] = sr;
    lpost[t] = log(post);
  }
  int kr_idx = which_max(score_r, T);
  int k = -1;
  if (kr_idx >= 0) {
    double cn = m->CNt[kr_idx], s = 0;
    for (int t = 0; t < T; t++)
      if (m->CNt[t] == cn && !isnan(lpost[t]) && (k < 0 || lpost[t] > s)) {
        s = lpost[t];
        k = t;
      }
  }
  if (k < 0) {
    out4[0] = out4[1] = out4[2] = na_real();
    out4[3] = NAN;
    return;
  }
  out4[0] = CNn;
  out4[1] = m->CNt[k];
  out4[2] = m->Mt[k];
  out4[3] = lpost[k];
}

static void mufreq_bayes_one_fwd(const void* m, double CNn, double F, double ratio,
                                 double wm, double wr, double* t1, double* t2,
                                 double* o4) {
  mufreq_bayes_one((const muf_model_t*)m, CNn, F, ratio, wm, wr, t1, t2, o4);
}
static int muf_model_build_fwd(void* m, double c, double p, double adr, int lo, int hi,
                               int CNn, int np, const double* pc, const double* pv,
                               int* T, double** to_free) {
  muf_model_t* mm = (muf_model_t*)m;
  int rc = muf_model_build(mm, c, p, adr, lo, hi, CNn, np, pc, pv);
  if (!rc) {
    *T = mm->T;
    *to_free = mm->CNt;
  }
  return rc;
}

/* ---- R/bayes.R:126-173 mufreq.bayes; out N x 4 column-major (CNn, CNt, Mt, LPP) ---- */
int sqzo_mufreq_bayes(int64_t N, const double* mufreq, const double* depth_ratio,
                      const double* w_mufreq, const double* w_ratio, double cellularity,
                      double ploidy, double avg_depth_ratio, int CNt_min, int CNt_max,
                      int CNn, int n_prior, const double* prior_CN,
                      const double* prior_value, double* out) {
  muf_model_t m;
  if (muf_model_build(&m, cellularity, ploidy, avg_depth_ratio, CNt_min, CNt_max, CNn,
                      n_prior, prior_CN, prior_value))
    return -1;
  double* tmp = (double*)malloc(sizeof(double) * 2 * m.T);
  for (int64_t s = 0; s < N; s++) {
    double o4[4];
    mufreq_bayes_one(&m, CNn, mufreq[s], depth_ratio[s], w_mufreq[s], w_ratio[s], tmp,
                     tmp + m.T, o4);
    out[s] = o4[0];
    out[N + s] = o4[1];
    out[2 * N + s] = o4[2];
    out[3 * N + s] = o4[3];
  }
  free(tmp);
  free(m.CNt);
  return 0;
}

/* ---- R/bayes.R:175-186: per-grid-point sum(mufreq.bayes(...)[,4]) ---- */
int sqzo_mufreq_model_fit_sums(int64_t N, const double* mufreq, const double* depth_ratio,
                               const double* w_mufreq, const double* w_ratio,
                               const double* ploidy, int64_t nP,
                               const double* cellularity, int64_t nC,
                               double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                               int n_prior, const double* prior_CN,
                               const double* prior_value, int64_t g0, int64_t g1,
                               int nthreads, double* lpp_sum) {
  (void)nC;
  grid_job_t job = {0};
  job.is_mufreq = 1;
  job.N = N;
  job.a = mufreq;
  job.b = depth_ratio;
  job.c = w_mufreq;
  job.d = w_ratio;
  job.ploidy = ploidy;
  job.nP = nP;
  job.cellularity = cellularity;
  job.adr = avg_depth_ratio;
  job.CNt_min = CNt_min;
  job.CNt_max = CNt_max;
  job.CNn = CNn;
  job.n_prior = n_prior;
  job.prior_CN = prior_CN;
  job.prior_value = prior_value;
  job.g0 = g0;
  job.g1 = g1;
  job.out = lpp_sum;
  return grid_run(&job, nthreads);
}

/* ---- R/bayes.R:120-122: LogSumLik and znorm = exp(z - LogSumLik), in place over the
 * length-G vector.  max(na.rm = TRUE) but sum() without, exactly as the reference. ---- */
void sqzo_normalise_lpp(int64_t G, const double* lpp_sum, double* znorm) {
  double mx = -INFINITY;
  int any = 0;
  for (int64_t g = 0; g < G; g++)
    if (!isnan(lpp_sum[g]) && (!any || lpp_sum[g] > mx)) {
      mx = lpp_sum[g];
      any = 1;
    }
  long double acc = 0.0L;
  for (int64_t g = 0; g < G; g++) acc += exp(lpp_sum[g] - mx);
  double lsl = log((double)acc) + mx;
  for (int64_t g = 0; g < G; g++) znorm[g] = exp(lpp_sum[g] - lsl);
}

int sqzo_max_threads(void) { return (int)sysconf(_SC_NPROCESSORS_ONLN); }
