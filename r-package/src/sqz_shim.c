/*
 * sqz_shim.c -- .Call shim between sequenza's R code and libsequenza_b200.so.
 *
 * UNTESTED IN THIS REPOSITORY: the build image has no R (no R.h / Rinternals.h), so
 * this file is compiled only where R exists (see Makevars).  It contains no arithmetic:
 * every routine checks its SEXP arguments, calls one C-ABI entry point of
 * include/sequenza_b200.h and turns a non-zero status into an R error.
 *
 * Ownership: inputs are R-owned REALSXPs, read-only.  Outputs are allocated with
 * Rf_allocVector / Rf_allocMatrix and PROTECTed until return.  Device buffers belong to
 * the library (cached per device, released by sqzr_release), so nothing needs freeing
 * before Rf_error() longjmps.  .Call runs on R's main thread; the library is entered
 * from that thread only.
 */
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>
#include <stdint.h>

#include "sequenza_b200.h"

static void need_real(SEXP x, R_xlen_t n, const char* what) {
  if (TYPEOF(x) != REALSXP) Rf_error("%s must be a double vector (use as.double)", what);
  if (n >= 0 && XLENGTH(x) != n)
    Rf_error("%s has length %lld, expected %lld", what, (long long)XLENGTH(x), (long long)n);
}

static void raise(int rc) {
  if (rc != SQZ_OK) Rf_error("sequenza_b200: %s", sqz_last_error());
}

/* replaces fit.cp + pblapply in baf.model.fit (R/bayes.R:109-115); returns result$LPP */
SEXP sqzr_baf_grid(SEXP Bf, SEXP ratio, SEXP sdBf, SEXP sdr, SEXP wBf, SEXP wr, SEXP ploidy,
                   SEXP cellularity, SEXP adr, SEXP cn_min, SEXP cn_max, SEXP cnn, SEXP pCN,
                   SEXP pval, SEXP rp, SEXP n_gpus, SEXP mode) {
  R_xlen_t n = XLENGTH(ratio);
  need_real(Bf, n, "Bf");
  need_real(ratio, n, "depth.ratio");
  need_real(sdBf, n, "sd.Bf");
  need_real(sdr, n, "sd.ratio");
  need_real(wBf, n, "weight.Bf");
  need_real(wr, n, "weight.ratio");
  need_real(ploidy, -1, "ploidy");
  need_real(cellularity, -1, "cellularity");
  need_real(pCN, -1, "priors.table$CN");
  need_real(pval, XLENGTH(pCN), "priors.table$value");
  R_xlen_t nP = XLENGTH(ploidy), nC = XLENGTH(cellularity);
  SEXP out = PROTECT(Rf_allocVector(REALSXP, nP * nC));
  int rc = sqz_baf_grid(REAL(Bf), REAL(ratio), REAL(sdBf), REAL(sdr), REAL(wBf), REAL(wr),
                        (int64_t)n, REAL(ploidy), (int64_t)nP, REAL(cellularity), (int64_t)nC,
                        Rf_asReal(adr), Rf_asInteger(cn_min), Rf_asInteger(cn_max),
                        Rf_asInteger(cnn), REAL(pCN), REAL(pval), (int)XLENGTH(pCN),
                        Rf_asLogical(rp) == TRUE, Rf_asInteger(n_gpus), Rf_asInteger(mode),
                        REAL(out));
  UNPROTECT(1);
  raise(rc);
  return out;
}

/* replaces the body of baf.bayes (R/bayes.R:46-98); returns the N x 4 matrix */
SEXP sqzr_baf_point(SEXP Bf, SEXP ratio, SEXP sdBf, SEXP sdr, SEXP wBf, SEXP wr,
                    SEXP cellularity, SEXP ploidy, SEXP adr, SEXP cn_min, SEXP cn_max,
                    SEXP cnn, SEXP pCN, SEXP pval, SEXP rp) {
  R_xlen_t n = XLENGTH(ratio);
  need_real(Bf, n, "Bf");
  need_real(ratio, n, "depth.ratio");
  need_real(sdBf, n, "sd.Bf");
  need_real(sdr, n, "sd.ratio");
  need_real(wBf, n, "weight.Bf");
  need_real(wr, n, "weight.ratio");
  need_real(pCN, -1, "priors.table$CN");
  need_real(pval, XLENGTH(pCN), "priors.table$value");
  SEXP out = PROTECT(Rf_allocMatrix(REALSXP, (int)n, 4));
  int rc = sqz_baf_point(REAL(Bf), REAL(ratio), REAL(sdBf), REAL(sdr), REAL(wBf), REAL(wr),
                         (int64_t)n, Rf_asReal(cellularity), Rf_asReal(ploidy), Rf_asReal(adr),
                         Rf_asInteger(cn_min), Rf_asInteger(cn_max), Rf_asInteger(cnn),
                         REAL(pCN), REAL(pval), (int)XLENGTH(pCN), Rf_asLogical(rp) == TRUE,
                         REAL(out));
  UNPROTECT(1);
  raise(rc);
  return out;
}

/* replaces the body of mufreq.bayes (R/bayes.R:132-170); returns the N x 4 matrix */
SEXP sqzr_mufreq_point(SEXP F, SEXP ratio, SEXP wm, SEXP wr, SEXP cellularity, SEXP ploidy,
                       SEXP adr, SEXP cn_min, SEXP cn_max, SEXP cnn, SEXP pCN, SEXP pval) {
  R_xlen_t n = XLENGTH(F);
  need_real(F, n, "mufreq");
  need_real(ratio, n, "depth.ratio");
  need_real(wm, n, "weight.mufreq");
  need_real(wr, n, "weight.ratio");
  need_real(pCN, -1, "priors.table$CN");
  need_real(pval, XLENGTH(pCN), "priors.table$value");
  SEXP out = PROTECT(Rf_allocMatrix(REALSXP, (int)n, 4));
  int rc = sqz_mufreq_point(REAL(F), REAL(ratio), REAL(wm), REAL(wr), (int64_t)n,
                            Rf_asReal(cellularity), Rf_asReal(ploidy), Rf_asReal(adr),
                            Rf_asInteger(cn_min), Rf_asInteger(cn_max), Rf_asInteger(cnn),
                            REAL(pCN), REAL(pval), (int)XLENGTH(pCN), REAL(out));
  UNPROTECT(1);
  raise(rc);
  return out;
}

/* mufreq.bayes with scalar weights (the reference's defaults, R/results.R:153-157) */
SEXP sqzr_mufreq_point_scalar(SEXP F, SEXP ratio, SEXP wm, SEXP wr, SEXP cellularity,
                              SEXP ploidy, SEXP adr, SEXP cn_min, SEXP cn_max, SEXP cnn,
                              SEXP pCN, SEXP pval) {
  R_xlen_t n = XLENGTH(F);
  need_real(F, n, "mufreq");
  need_real(ratio, n, "depth.ratio");
  need_real(pCN, -1, "priors.table$CN");
  need_real(pval, XLENGTH(pCN), "priors.table$value");
  SEXP out = PROTECT(Rf_allocMatrix(REALSXP, (int)n, 4));
  int rc = sqz_mufreq_point_scalar(REAL(F), REAL(ratio), (int64_t)n, Rf_asReal(wm),
                                   Rf_asReal(wr), Rf_asReal(cellularity), Rf_asReal(ploidy),
                                   Rf_asReal(adr), Rf_asInteger(cn_min), Rf_asInteger(cn_max),
                                   Rf_asInteger(cnn), REAL(pCN), REAL(pval),
                                   (int)XLENGTH(pCN), REAL(out));
  UNPROTECT(1);
  raise(rc);
  return out;
}

/* replaces fit.cp + pblapply in mufreq.model.fit (R/bayes.R:179-185) */
SEXP sqzr_mufreq_grid(SEXP F, SEXP ratio, SEXP wm, SEXP wr, SEXP ploidy, SEXP cellularity,
                      SEXP adr, SEXP cn_min, SEXP cn_max, SEXP cnn, SEXP pCN, SEXP pval,
                      SEXP n_gpus) {
  R_xlen_t n = XLENGTH(F);
  need_real(F, n, "mufreq");
  need_real(ratio, n, "depth.ratio");
  need_real(wm, n, "weight.mufreq");
  need_real(wr, n, "weight.ratio");
  need_real(ploidy, -1, "ploidy");
  need_real(cellularity, -1, "cellularity");
  need_real(pCN, -1, "priors.table$CN");
  need_real(pval, XLENGTH(pCN), "priors.table$value");
  R_xlen_t nP = XLENGTH(ploidy), nC = XLENGTH(cellularity);
  SEXP out = PROTECT(Rf_allocVector(REALSXP, nP * nC));
  int rc = sqz_mufreq_grid(REAL(F), REAL(ratio), REAL(wm), REAL(wr), (int64_t)n, REAL(ploidy),
                           (int64_t)nP, REAL(cellularity), (int64_t)nC, Rf_asReal(adr),
                           Rf_asInteger(cn_min), Rf_asInteger(cn_max), Rf_asInteger(cnn),
                           REAL(pCN), REAL(pval), (int)XLENGTH(pCN), Rf_asInteger(n_gpus),
                           REAL(out));
  UNPROTECT(1);
  raise(rc);
  return out;
}

SEXP sqzr_device_count(void) { return Rf_ScalarInteger(sqz_device_count()); }

SEXP sqzr_release(void) {
  sqz_release();
  return R_NilValue;
}

static const R_CallMethodDef call_methods[] = {
    {"sqzr_baf_grid", (DL_FUNC)&sqzr_baf_grid, 17},
    {"sqzr_baf_point", (DL_FUNC)&sqzr_baf_point, 15},
    {"sqzr_mufreq_point", (DL_FUNC)&sqzr_mufreq_point, 12},
    {"sqzr_mufreq_point_scalar", (DL_FUNC)&sqzr_mufreq_point_scalar, 12},
    {"sqzr_mufreq_grid", (DL_FUNC)&sqzr_mufreq_grid, 13},
    {"sqzr_device_count", (DL_FUNC)&sqzr_device_count, 0},
    {"sqzr_release", (DL_FUNC)&sqzr_release, 0},
    {NULL, NULL, 0}};

void R_init_sequenza(DllInfo* dll) {
  R_registerRoutines(dll, NULL, call_methods, NULL, NULL);
  R_useDynamicSymbols(dll, FALSE);
}
