// sqz_baf_grid.cuh -- types shared by the baf grid translation units.
#pragma once
#include "sqz_common.cuh"

namespace sqz {

struct BafGridArgs {
  int64_t S;
  const int64_t* seg_off;  // [S+1]
  const double *Bf, *ratio, *sdBf, *sdr, *wBf, *wr;  // raw per-segment vectors
  const double2* rk;  // fast path: (kr, er)   xr/sqrt5 = fma(log mr, kr, er)
  const double2* bk;  // fast path: (kb, eb)   xb/sqrt5 = fma(mb, kb, eb)
  const int* nB;      // [S] segments that carry a BAF term
  const double* adr;  // [S] avg.depth.ratio
  const double* ploidy;  // [nP]
  const double* cell;    // [nCs] this shard's cellularities
  int nP, nCs, nPT, nChunk;
  int rp;
  double t5;
};


// Split form of the fast path (sqz_baf_split.cu): the per-(cellularity, segment, CNt)
// BAF factor is computed once per column by a pre-pass and streamed back by the grid
// kernel, instead of being recomputed by every ploidy tile.
size_t baf_split_column_bytes(int nc, int64_t N_total);
bool baf_split_supported(int nc, int PT, int threads);
int baf_split_run(const BafGridArgs& a, const CnModel& m, int PT, int threads, int64_t N_total,
                  void* cw_buf, size_t cw_bytes, double* partM, long long* partE,
                  cudaStream_t st, bool prune, unsigned long long* stats);

// Pruned grid kernel with one cellularity column per lane (sqz_baf_colwarp.cu), with its
// own pre-pass and scratch layout: baf_colwarp_group_bytes per group of 32 columns.
// One warp per CTA.
bool baf_colwarp_supported(int nc, bool rp);
size_t baf_colwarp_group_bytes(int nc, int64_t N_total);
int baf_colwarp_run(const BafGridArgs& a, const CnModel& m, int PT, int64_t N_total,
                    void* cw_buf, size_t cw_bytes, double* partM, long long* partE,
                    cudaStream_t st, unsigned long long* stats);

}  // namespace sqz
