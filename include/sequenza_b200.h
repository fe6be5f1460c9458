/*
 * sequenza_b200.h -- C ABI of libsequenza_b200.so, the B200 (sm_100a) implementation of
 * sequenza 3.0.0's cellularity x ploidy likelihood path.
 *
 * The reference is pure R and has no FFI boundary of its own; these entry points are
 * the narrowest cuts through R/bayes.R that keep every R-visible behaviour (see
 * INTEGRATION.md for the .Call shim and the R-side edits).  Paths below are relative to
 * the reference tree.
 *
 * Conventions
 *   - Every function returns 0 (SQZ_OK) or a negative SQZ_ERR_* code; the message is
 *     available from sqz_last_error() (thread-local).  Nothing aborts the process.
 *   - All vectors are fp64.  Inputs are read-only.  Per-segment / per-position vectors
 *     all have length N (the R caller recycles scalars, as data.frame() does at
 *     R/bayes.R:43-45 and :130-131).  NA_real_ and NaN are both "NA" on input
 *     (is.na()); NA on output carries R's NA_real_ bit pattern 0x7FF00000000007A2.
 *   - Matrices are column-major, as in R.
 *   - priors.table is passed as two parallel vectors prior_CN / prior_value of length
 *     n_prior (R/bayes.R:55-59: later rows override earlier ones; values are divided by
 *     their sum over all type rows).
 *   - "host" entry points take host pointers and do their own transfers; "_dev" entry
 *     points take device pointers valid on the CURRENT device plus a caller-owned
 *     workspace and a cudaStream_t (passed as void*), and are asynchronous.
 *   - CNt_max - CNt_min + 1 <= 64.
 */
#ifndef SEQUENZA_B200_H
#define SEQUENZA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SQZ_OK 0
#define SQZ_ERR_ARG (-1)    /* bad argument (R: stop()) */
#define SQZ_ERR_CUDA (-2)   /* CUDA runtime failure; see sqz_last_error() */
#define SQZ_ERR_NOGPU (-3)  /* no usable device */
#define SQZ_ERR_SIZE (-4)   /* size beyond a stated limit */

/* grid evaluation mode */
#define SQZ_MODE_FAST 0     /* factorised product kernel + faithful repair of flagged points */
#define SQZ_MODE_FAITHFUL 1 /* every (grid point, segment, type) through the nmath sequence */

const char* sqz_version(void);
const char* sqz_last_error(void);
int sqz_device_count(void);
/* Frees every cached per-device buffer and stream held by the host entry points. */
int sqz_release(void);

/* ---------------------------------------------------------------------------------
 * B1  baf.model.fit grid -- replaces fit.cp + pblapply, R/bayes.R:109-115.
 * lpp_sum_out[ic * nP + ip] = sum(baf.bayes(cellularity[ic], ploidy[ip], ...)[, 4]),
 * i.e. result$LPP in expand.grid(ploidy, cellularity) order (R/bayes.R:106-107, :116).
 * The tapply / LogSumLik / exp steps (R/bayes.R:117-123) stay in R.
 * n_gpus >= 1: the grid is dealt to that many devices of this process in whole units
 * (tiles of columns x ploidies planned from the whole grid, so the result has the same
 * bits for any n_gpus); the segment table reaches device 0 once and the others over NVLink,
 * and every device stores its results into device 0's buffer.
 * --------------------------------------------------------------------------------- */
int sqz_baf_grid(const double* Bf, const double* depth_ratio, const double* sd_Bf,
                 const double* sd_ratio, const double* weight_Bf,
                 const double* weight_ratio, int64_t N, const double* ploidy, int64_t nP,
                 const double* cellularity, int64_t nC, double avg_depth_ratio,
                 int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                 const double* prior_value, int n_prior, int ratio_priority, int n_gpus,
                 int mode, double* lpp_sum_out);

/* Cohort form of B1 (BASELINE config 5): S samples share one grid.  Sample k owns
 * segments seg_offsets[k] .. seg_offsets[k+1]-1 of the concatenated vectors and
 * avg_depth_ratio[k]; lpp_sum_out is S x (nC*nP), sample-major.  One launch per device;
 * with n_gpus > 1 and S >= n_gpus the cohort is split by sample, not by unit. */
int sqz_baf_grid_batch(int64_t S, const int64_t* seg_offsets, const double* Bf,
                       const double* depth_ratio, const double* sd_Bf,
                       const double* sd_ratio, const double* weight_Bf,
                       const double* weight_ratio, const double* ploidy, int64_t nP,
                       const double* cellularity, int64_t nC,
                       const double* avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                       const double* prior_CN, const double* prior_value, int n_prior,
                       int ratio_priority, int n_gpus, int mode, double* lpp_sum_out);

/* Device-resident form of B1 for callers that own device memory (one process per GPU,
 * torch.distributed): segment vectors, seg_offsets (S+1 int64) and avg_depth_ratio (S)
 * are DEVICE pointers; ploidy / cellularity / priors are HOST pointers (they
 * parameterise the launch).  Evaluates cellularity columns ic0 <= ic < ic1 only and
 * writes S x ((ic1-ic0)*nP) doubles to d_lpp_sum_out (device).  workspace must hold
 * at least sqz_baf_grid_workspace_bytes(...) bytes, 256-byte aligned. */
size_t sqz_baf_grid_workspace_bytes(int64_t N_total, int64_t S, int64_t nP, int64_t nC_shard);
/* The minimum above selects the fused kernel.  With more room the library runs the faster
 * split forms: a per-column pre-pass whose output is streamed back by the grid kernel, 8 *
 * n_copy_numbers * N_total bytes per cellularity column, in whole groups of 32 columns for
 * the kernel that large fits with CNt 0..7 or 0..20 get (one cellularity per lane); columns
 * are processed in batches that fit.  This returns minimum + room for all nC_shard columns
 * of the kernel the planner picks for the WHOLE grid of nC columns (the kernel, its tile
 * width and its segment chunks never depend on the shard, so that a grid point gets the
 * same bits from any shard), at most cap_bytes of it (0 = no cap; at least one column).
 * The host entry points use a cap of 16 GiB.  n_copy_numbers = CNt_max - CNt_min + 1. */
size_t sqz_baf_grid_workspace_preferred_bytes(int64_t N_total, int64_t S, int64_t nP,
                                              int64_t nC, int64_t nC_shard,
                                              int n_copy_numbers, size_t cap_bytes);
int sqz_baf_grid_dev(int64_t S, const int64_t* d_seg_offsets, int64_t N_total,
                     const double* d_Bf, const double* d_depth_ratio,
                     const double* d_sd_Bf, const double* d_sd_ratio,
                     const double* d_weight_Bf, const double* d_weight_ratio,
                     const double* ploidy, int64_t nP, const double* cellularity,
                     int64_t nC, int64_t ic0, int64_t ic1, const double* d_avg_depth_ratio,
                     int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                     const double* prior_value, int n_prior, int ratio_priority, int mode,
                     void* workspace, size_t workspace_bytes, void* stream,
                     double* d_lpp_sum_out);

/* B1 sharded by UNITS instead of whole columns, for an even split of the grid over GPUs
 * (the loop being split is still fit.cp over expand.grid, R/bayes.R:106-115).  The grid is
 * cut into units of W columns x PT ploidies counted from its origin, unit = group * nPT +
 * tile; sqz_baf_grid_plan reports W, PT, nPT and the unit count of the whole grid.  This
 * call evaluates units u0 <= u < u1 and writes every result at its FINAL place of a
 * whole-grid buffer (S x nC*nP doubles, ploidy fastest) -- of each of the n_out buffers in
 * d_outs (a HOST array of device pointers): the caller's own buffer and, for a multi-GPU
 * fit, the peers' buffers mapped into this process (peer access or CUDA IPC), so that the
 * last kernel of the fit is also its all-gather over NVLink.  Grid points outside the
 * unit range are not touched.  The caller orders the peers' reads after this call
 * (sqz_peer_signal / sqz_peer_wait or a stream/host synchronisation).
 * Workspace: sqz_baf_grid_units_workspace_bytes(...) (cap_bytes as above). */
size_t sqz_baf_grid_units_workspace_bytes(int64_t N_total, int64_t S, const double* ploidy,
                                          int64_t nP, int64_t nC, int CNt_min, int CNt_max,
                                          int CNn, int ratio_priority, int64_t u0, int64_t u1,
                                          size_t cap_bytes);
int sqz_baf_grid_dev_units(int64_t S, const int64_t* d_seg_offsets, int64_t N_total,
                           const double* d_Bf, const double* d_depth_ratio,
                           const double* d_sd_Bf, const double* d_sd_ratio,
                           const double* d_weight_Bf, const double* d_weight_ratio,
                           const double* ploidy, int64_t nP, const double* cellularity,
                           int64_t nC, int64_t u0, int64_t u1,
                           const double* d_avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                           const double* prior_CN, const double* prior_value, int n_prior,
                           int ratio_priority, int mode, void* workspace,
                           size_t workspace_bytes, void* stream, double* const* d_outs,
                           int n_out);

/* Result buffers shared by the GPUs of one box, one process per GPU -- the assembly of
 * result$LPP that the reference gets from its mclapply workers over pipes
 * (R/bayes.R:114-116).  A peer buffer is sqz_peer_data_offset() bytes of header (one
 * 8-byte flag per rank) followed by data_bytes of data.  sqz_peer_alloc creates one on the
 * current device and exports its 64-byte CUDA IPC handle; the other processes map it with
 * sqz_peer_open (addresses valid in the opening process; NVLink peer mapping) and pass
 * `base + sqz_peer_data_offset()` of every rank to sqz_baf_grid_dev_units as d_outs.
 * After the fit, on the same stream: sqz_peer_signal (all earlier stores of this GPU made
 * visible, then flag[my_rank] = epoch in every buffer of d_bases, a host array of the n
 * ranks' bases in rank order) and sqz_peer_wait (until all n flags of the own buffer show
 * at least epoch; a peer missing for 10 s traps the kernel).  Epochs must grow from call
 * to call; alternate between two data areas so that nobody overwrites what a slower rank
 * still reads.  At most 16 ranks. */
size_t sqz_peer_data_offset(void);
int sqz_peer_alloc(size_t data_bytes, void** d_base, unsigned char* handle64,
                   size_t* total_bytes);
int sqz_peer_open(const unsigned char* handle64, void** d_base);
int sqz_peer_close(void* d_base);
int sqz_peer_free(void* d_base);
int sqz_peer_signal(void* const* d_bases, int n, int my_rank, uint64_t epoch, void* stream);
int sqz_peer_wait(void* d_base, int n, uint64_t epoch, void* stream);

/* ---------------------------------------------------------------------------------
 * Grid fit for callers that cannot allocate device memory.  This entry point has no
 * counterpart in the reference.
 *
 * Who needs it: code that runs where cudaMalloc is off limits -- a host thread inside a
 * real-time or latency-bounded loop (cudaMalloc/cudaFree take a driver lock and may
 * synchronise the whole device), a plugin whose embedding application owns the memory
 * budget and forbids further allocations after start-up, or a process that has already
 * handed all free device memory to another pool and would simply get an out-of-memory
 * error.  Such callers still own their input and output buffers, but have nowhere to
 * get the workspace that sqz_baf_grid_dev asks for.
 *
 * What it does: exactly what sqz_baf_grid_dev does (same arguments minus the workspace,
 * same kernels, same results), with the scratch taken from a fixed 16 MiB array that is
 * part of the library image on every GPU that loads it.  The call performs no
 * cudaMalloc, cudaFree or other allocating call.
 *
 * Limits and rules:
 *   - the problem must fit: sqz_baf_grid_workspace_bytes(N_total, S, nP, ic1 - ic0) must
 *     be <= 16 MiB (16,777,216 bytes); larger problems are refused with SQZ_ERR_SIZE and
 *     a message naming both numbers.  BASELINE config 2 (3e4 segments, 91 x 61 grid)
 *     needs about 6.4 MiB;
 *   - the first call on each device looks up the device address of the built-in scratch
 *     (one cudaGetSymbolAddress); make that first call before starting any stream
 *     capture on that device;
 *   - all calls on a device share the one scratch array, and it keeps whatever the
 *     previous call left in it (each call re-initialises the parts it needs zeroed).
 *     Two calls whose GPU work overlaps in time -- different streams, or different host
 *     threads -- race on it and corrupt each other's results.  Issue the calls on one
 *     stream, or order them with events;
 *   - per-kernel timing (sqz_profile_enable) is ignored by this entry point.
 * --------------------------------------------------------------------------------- */
int sqz_baf_grid_dev_noalloc(int64_t S, const int64_t* d_seg_offsets, int64_t N_total,
                             const double* d_Bf, const double* d_depth_ratio,
                             const double* d_sd_Bf, const double* d_sd_ratio,
                             const double* d_weight_Bf, const double* d_weight_ratio,
                             const double* ploidy, int64_t nP, const double* cellularity,
                             int64_t nC, int64_t ic0, int64_t ic1,
                             const double* d_avg_depth_ratio, int CNt_min, int CNt_max,
                             int CNn, const double* prior_CN, const double* prior_value,
                             int n_prior, int ratio_priority, int mode, void* stream,
                             double* d_lpp_sum_out);

/* ---------------------------------------------------------------------------------
 * B2  baf.bayes at one (cellularity, ploidy) -- replaces R/bayes.R:46-98.
 * out is N x 4 column-major: CNt, A, B, LPP (R/bayes.R:99).  Callers: R/results.R:128,
 * :139, R/graphics.R:419.  Rows whose posterior is NaN for every type (the reference
 * stops with an R error there) come back as NA, NA, NA, NaN.
 * --------------------------------------------------------------------------------- */
int sqz_baf_point(const double* Bf, const double* depth_ratio, const double* sd_Bf,
                  const double* sd_ratio, const double* weight_Bf,
                  const double* weight_ratio, int64_t N, double cellularity, double ploidy,
                  double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                  const double* prior_CN, const double* prior_value, int n_prior,
                  int ratio_priority, double* out);
int sqz_baf_point_dev(const double* d_Bf, const double* d_depth_ratio,
                      const double* d_sd_Bf, const double* d_sd_ratio,
                      const double* d_weight_Bf, const double* d_weight_ratio, int64_t N,
                      double cellularity, double ploidy, double avg_depth_ratio,
                      int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                      const double* prior_value, int n_prior, int ratio_priority,
                      void* stream, double* d_out);

/* ---------------------------------------------------------------------------------
 * B3  mufreq.bayes at one (cellularity, ploidy) -- replaces R/bayes.R:132-170.
 * out is N x 4 column-major: CNn, CNt, Mt, LPP (R/bayes.R:171).  Callers:
 * R/results.R:153, :161.  weight_* have length N.
 * --------------------------------------------------------------------------------- */
int sqz_mufreq_point(const double* mufreq, const double* depth_ratio,
                     const double* weight_mufreq, const double* weight_ratio, int64_t N,
                     double cellularity, double ploidy, double avg_depth_ratio,
                     int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                     const double* prior_value, int n_prior, double* out);
int sqz_mufreq_point_dev(const double* d_mufreq, const double* d_depth_ratio,
                         const double* d_weight_mufreq, const double* d_weight_ratio,
                         int64_t N, double cellularity, double ploidy,
                         double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                         const double* prior_CN, const double* prior_value, int n_prior,
                         void* stream, double* d_out);

/* B3 with scalar weights, as sequenza.results calls it (weight.mufreq = weight.ratio = 100,
 * R/results.R:153-157): the N x 4 result depends on round(F * w.m) and
 * round(w.r * r/(1+r)) only, so the (w.r+1) x (w.m+1) answers are computed once with the
 * same arithmetic and every position becomes a lookup (16 bytes in, 32 bytes out).
 * Weights must be non-negative integers <= 4096 with (w.m+1)(w.r+1) <= 4e6, otherwise
 * SQZ_ERR_ARG (use sqz_mufreq_point).  The _dev form needs a device buffer of
 * sqz_mufreq_scalar_table_bytes(...) bytes, 32-byte aligned. */
size_t sqz_mufreq_scalar_table_bytes(double weight_mufreq, double weight_ratio);
int sqz_mufreq_point_scalar(const double* mufreq, const double* depth_ratio, int64_t N,
                            double weight_mufreq, double weight_ratio, double cellularity,
                            double ploidy, double avg_depth_ratio, int CNt_min, int CNt_max,
                            int CNn, const double* prior_CN, const double* prior_value,
                            int n_prior, double* out);
int sqz_mufreq_point_scalar_dev(const double* d_mufreq, const double* d_depth_ratio, int64_t N,
                                double weight_mufreq, double weight_ratio, double cellularity,
                                double ploidy, double avg_depth_ratio, int CNt_min, int CNt_max,
                                int CNn, const double* prior_CN, const double* prior_value,
                                int n_prior, void* d_table, size_t table_bytes, void* stream,
                                double* d_out);

/* ---------------------------------------------------------------------------------
 * B4  mufreq.model.fit grid -- replaces fit.cp + pblapply, R/bayes.R:179-185.
 * Same output convention as sqz_baf_grid.
 * --------------------------------------------------------------------------------- */
int sqz_mufreq_grid(const double* mufreq, const double* depth_ratio,
                    const double* weight_mufreq, const double* weight_ratio, int64_t N,
                    const double* ploidy, int64_t nP, const double* cellularity,
                    int64_t nC, double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                    const double* prior_CN, const double* prior_value, int n_prior,
                    int n_gpus, double* lpp_sum_out);
int sqz_mufreq_grid_dev(const double* d_mufreq, const double* d_depth_ratio,
                        const double* d_weight_mufreq, const double* d_weight_ratio,
                        int64_t N, const double* ploidy, int64_t nP,
                        const double* cellularity, int64_t nC, int64_t ic0, int64_t ic1,
                        double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                        const double* prior_CN, const double* prior_value, int n_prior,
                        void* workspace, size_t workspace_bytes, void* stream,
                        double* d_lpp_sum_out);
size_t sqz_mufreq_grid_workspace_bytes(int64_t N, int64_t nP, int64_t nC_shard);

/* ---------------------------------------------------------------------------------
 * After the grid: normalisation (R/bayes.R:120-122) and the reductions get.ci needs
 * (R/solutions.R:1-15), for S result$LPP vectors of one nP x nC grid (S > 1: a cohort).
 *   lpp            S x (nP*nC): exp(LPP - LogSumLik); as an R matrix [ploidy x cellularity]
 *                  (column-major) when the grid axes are already sorted and unique
 *   row_max        S x nP: apply(lpp, 1, max)        col_max        S x nC: apply(lpp, 2, max)
 *   row_mass_above S x nP, col_mass_above S x nC: total lpp mass of the grid points whose
 *                  lpp is strictly greater than that marginal maximum.  A marginal value is
 *                  inside get.ci's `level` region iff its mass_above < level -- the same
 *                  set the reference gets from sort + cumsum (R/solutions.R:2-5, :10-14),
 *                  without a sort.  Masses are accumulated in 2^-62 fixed point.
 *   log_sum_lik    S: LogSumLik
 * max(na.rm = TRUE) / sum() without na.rm as in the reference (any NaN in LPP makes the
 * whole lpp NaN).  Requires nP + nC <= 4096.  The tapply step (R/bayes.R:117) is the
 * identity for sorted unique axes; callers with duplicated or unsorted axis values keep
 * the R code.
 * --------------------------------------------------------------------------------- */
int sqz_fit_finish(const double* lpp_sum, int64_t S, int64_t nP, int64_t nC, double* lpp,
                   double* row_max, double* col_max, double* row_mass_above,
                   double* col_mass_above, double* log_sum_lik);
int sqz_fit_finish_dev(const double* d_lpp_sum, int64_t S, int64_t nP, int64_t nC,
                       double* d_lpp, double* d_row_max, double* d_col_max,
                       double* d_row_mass_above, double* d_col_mass_above,
                       double* d_log_sum_lik, void* stream);

/* ---------------------------------------------------------------------------------
 * Before the grid: the per-segment statistics of segment.breaks, R/segments.R:113-198
 * (weighted.mean = TRUE: :146-170, FALSE: :171-190), with b_allele_freq's kernel-density
 * mode, R/model.R:69-98 -- the step that produces the segment table sequenza.fit filters
 * and fits (R/extract.R:175-185 calls it per chromosome).  The breakpoints themselves
 * (copynumber::aspcf, R/segments.R:1-111) stay with the caller.
 *   rows       all chromosomes concatenated, chromosome c = rows row_off[c] ..
 *              row_off[c+1]-1, sorted by position inside a chromosome; fp64 seqz columns
 *              position, adjusted.ratio, depth.normal, good.reads, Af, Bf, and is_het (one
 *              byte per row: zygosity.normal == "het").  NaN/NA_real_ are NA (na.rm = TRUE)
 *   breaks     breaks_vect holds (start, end] of segment s at [2s], [2s+1]: the output of
 *              unique.breaks (R/segments.R:133-139), increasing inside a chromosome;
 *              chromosome c owns segments seg_off[c] .. seg_off[c+1]-1.  As cut() does, a
 *              row belongs to a segment if start < position <= end
 *   out        n_seg x 6 column-major: Bf, N.BAF, sd.BAF, depth.ratio, N.ratio, sd.ratio
 *              (R/segments.R:164-170); NA where the reference has NA (sd of fewer than two
 *              values, Bf without a het position)
 * The final filter (N.ratio per Mb >= 2, R/segments.R:196-197) and the chromosome labels
 * stay with the caller.  The density is density.default's algorithm (bw.nrd0, linear
 * binning onto 512 points, gaussian kernel, approx) with the convolution summed directly
 * instead of by FFT; one stated departure: a maximum whose top is a run of exactly equal
 * values counts at its first point.
 * The _dev form needs sqz_segment_stats_workspace_bytes(n_rows, n_seg) bytes, 256-byte
 * aligned (a hand-over record per segment and a compact copy of the het rows).
 * --------------------------------------------------------------------------------- */
int sqz_segment_stats(int64_t n_chr, const int64_t* row_off, const int64_t* seg_off,
                      const double* position, const double* adjusted_ratio,
                      const double* depth_normal, const double* good_reads, const double* Af,
                      const double* Bf, const unsigned char* is_het, const double* breaks_vect,
                      double min_reads_baf, int weighted_mean, double* out);
size_t sqz_segment_stats_workspace_bytes(int64_t n_rows, int64_t n_seg);
int sqz_segment_stats_dev(int64_t n_chr, const int64_t* d_row_off, const int64_t* d_seg_off,
                          int64_t n_rows, int64_t n_seg, const double* d_position,
                          const double* d_adjusted_ratio, const double* d_depth_normal,
                          const double* d_good_reads, const double* d_Af, const double* d_Bf,
                          const unsigned char* d_is_het, const double* d_breaks_vect,
                          double min_reads_baf, int weighted_mean, void* workspace,
                          size_t workspace_bytes, void* stream, double* d_out);

/* ---------------------------------------------------------------------------------
 * Model points as the kernels compute them (R/model.R:49-67), for checking the device
 * arithmetic against tests/testthat/test.model.R.  out is T x 5 column-major:
 * CNt, second index (B or Mt), model depth.ratio, model BAF or mufreq, normalised prior.
 * kind 0 = baf types (R/model.R:22-37), 1 = mufreq types (R/model.R:39-47).
 * Returns T through *T_out; out may be NULL to query T.
 * --------------------------------------------------------------------------------- */
int sqz_model_points(int kind, double cellularity, double ploidy, double avg_depth_ratio,
                     int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                     const double* prior_value, int n_prior, int* T_out, double* out);

/* ---------------------------------------------------------------------------------
 * Measured fp64 issue rate of the current device, for the roofline denominator that
 * MEASURED_PEAKS.json does not carry: dependent-free DFMA chains on every SM.
 * tflops_out = 2 * DFMA / s / 1e12.  No counterpart in the reference.
 * --------------------------------------------------------------------------------- */
int sqz_fp64_peak(int device, int repeats, double* tflops_out);

/* Launch geometry of the fast grid path for a grid of nC x nP points, for reporting and for
 * sharding by units: out8 = ploidies per tile (PT), tiles per column (nPT), segment chunks,
 * threads per CTA, number of (CNt, B) types, kernel family (0 fused, 1 split, 2 pruned with
 * one segment per lane, 3 pruned with one cellularity column per lane -- what the preferred
 * workspace gives), columns per unit (W), units of the whole grid (ceil(nC / W) * nPT).
 * `ploidy` (nP host values, or NULL = sorted) matters because candidate pruning needs a
 * monotone ploidy axis.  No counterpart in the reference. */
int sqz_baf_grid_plan(int64_t N_total, int64_t S, const double* ploidy, int64_t nP, int64_t nC,
                      int CNt_min, int CNt_max, int CNn, int ratio_priority, int64_t* out8);

/* Per-kernel timing of fast-mode grid calls on the current device: after
 * sqz_profile_enable(1), every sqz_baf_grid_dev / sqz_baf_grid call records CUDA events
 * on its launch stream around (preparation | grid kernel | finalize + repair);
 * sqz_profile_collect sums them in ms3[0..2] over *n_calls calls and resets.  The caller
 * synchronises the stream before collecting.  No counterpart in the reference. */
int sqz_profile_enable(int on);
int sqz_profile_collect(double* ms3, int* n_calls);
/* Call before sqz_profile_collect: *path = kernel path of the recorded calls (0 fused, 1
 * split, 2 split with candidate pruning, one segment per lane, 3 candidate pruning with
 * one cellularity column per lane); *n_eval = (segment, ploidy tile, copy number)
 * evaluations the pruned kernel actually made, summed over the recorded calls (path 3:
 * per column, counting what the column's warp executed). */
/* *launches = kernels the recorded calls launched (counted at the launch sites). */
int sqz_profile_counters(int* path, unsigned long long* n_eval, long long* launches);

#ifdef __cplusplus
}
#endif
#endif /* SEQUENZA_B200_H */
