/* Plain C99 consumer of include/sequenza_b200.h: what an R .Call shim (or any other C
 * caller) links against.  Takes the address of every entry point, prints the library
 * version and device count, and -- when a GPU is present -- fits a tiny grid.
 *   gcc -std=c99 -Iinclude examples/abi_check.c -Lsequenza_b200 -lsequenza_b200 \
 *       -Wl,-rpath,$PWD/sequenza_b200 -o abi_check */
#include <stdio.h>
#include "sequenza_b200.h"

typedef void (*fn_t)(void);

int main(void) {
  const fn_t syms[] = {
      (fn_t)sqz_version, (fn_t)sqz_last_error, (fn_t)sqz_device_count,
      (fn_t)sqz_release, (fn_t)sqz_baf_grid, (fn_t)sqz_baf_grid_batch,
      (fn_t)sqz_baf_grid_workspace_bytes,
      (fn_t)sqz_baf_grid_workspace_preferred_bytes, (fn_t)sqz_baf_grid_dev,
      (fn_t)sqz_baf_grid_dev_noalloc, (fn_t)sqz_baf_point,
      (fn_t)sqz_baf_point_dev, (fn_t)sqz_mufreq_point,
      (fn_t)sqz_mufreq_point_dev, (fn_t)sqz_mufreq_scalar_table_bytes,
      (fn_t)sqz_mufreq_point_scalar, (fn_t)sqz_mufreq_point_scalar_dev,
      (fn_t)sqz_mufreq_grid, (fn_t)sqz_mufreq_grid_dev,
      (fn_t)sqz_mufreq_grid_workspace_bytes, (fn_t)sqz_fit_finish,
      (fn_t)sqz_fit_finish_dev, (fn_t)sqz_model_points,
      (fn_t)sqz_fp64_peak, (fn_t)sqz_baf_grid_plan,
      (fn_t)sqz_profile_enable, (fn_t)sqz_profile_collect,
      (fn_t)sqz_profile_counters, (fn_t)sqz_baf_grid_units_workspace_bytes,
      (fn_t)sqz_baf_grid_dev_units, (fn_t)sqz_peer_data_offset,
      (fn_t)sqz_peer_alloc, (fn_t)sqz_peer_open, (fn_t)sqz_peer_close,
      (fn_t)sqz_peer_free, (fn_t)sqz_peer_signal, (fn_t)sqz_peer_wait,
      (fn_t)sqz_segment_stats, (fn_t)sqz_segment_stats_workspace_bytes,
      (fn_t)sqz_segment_stats_dev};
  printf("%s: %d entry points, %d device(s)\n", sqz_version(),
         (int)(sizeof syms / sizeof syms[0]), sqz_device_count());
  {
    /* three segments, 2 x 2 grid, sequenza.fit's default priors (CN = 2 -> 2) */
    const double Bf[3] = {0.45, 0.30, 0.10}, ratio[3] = {1.0, 1.4, 0.7};
    const double sdB[3] = {0.05, 0.05, 0.05}, sdr[3] = {0.5, 0.6, 0.7};
    const double w[3] = {10.0, 20.0, 5.0};
    const double ploidy[2] = {2.0, 3.0}, cell[2] = {0.5, 0.9};
    const double pcn[1] = {2.0}, pval[1] = {2.0};
    double lpp[4];
    int rc = sqz_baf_grid(Bf, ratio, sdB, sdr, w, w, 3, ploidy, 2, cell, 2, 1.0, 0, 7, 2, pcn,
                          pval, 1, 0, 1, SQZ_MODE_FAST, lpp);
    if (rc == SQZ_OK)
      printf("LPP sums: %.6f %.6f %.6f %.6f\n", lpp[0], lpp[1], lpp[2], lpp[3]);
    else
      printf("sqz_baf_grid -> %d (%s)\n", rc, sqz_last_error());
    return (rc == SQZ_OK || rc == SQZ_ERR_NOGPU) ? 0 : 1;
  }
}
