// sqz_baf_grid.cuh -- types shared by the baf grid translation units.
//
// Geometry of a grid-fit call.  The cellularity x ploidy grid (nC x nP, ploidy fastest,
// R/bayes.R:106-107) is cut into UNITS of W columns x PT ploidies: W = 32 for the
// column-per-lane kernel, 1 for the segment-per-lane kernels.  Unit u = group * nPT + tile
// with groups and tiles counted from the grid's origin.  The kernel family, W, PT and the
// number of segment chunks are planned from the WHOLE grid and the segment table, never
// from the shard, so a grid point gets the same bits whichever shard (column block or
// unit range, any GPU count) computes it.
//
// A launch works on a SPAN of neighbouring columns [span0, span1) (all indices in the
// kernels are relative to span0: a.cell[0] is column span0) and on the units
// [u_first, u_first + nU) counted from the span's first group.  A column shard is the span
// itself with all of its units; a unit shard's span is the whole groups it touches.
#pragma once
#include "sqz_common.cuh"

namespace sqz {

constexpr int kMaxOut = 8;  // destinations of one result (this GPU and its peers)

struct alignas(32) RkRec {
  double2 rk;
  float4 f;  // (float)kr, (float)er, 2^-22 |er| rounded up, unused
};

struct BafGridArgs {
  int64_t S;
  const int64_t* seg_off;  // [S+1]
  const double *Bf, *ratio, *sdBf, *sdr, *wBf, *wr;  // raw per-segment vectors
  const double2* rk;  // fast path: (kr, er)   xr/sqrt5 = fma(log mr, kr, er)
  // the same with (kr, er) as floats and 2^-22 |er| rounded up behind it (the fp32 bounds of
  // the column-per-lane kernel): one 32-byte record per segment, one bulk copy per stage
  const RkRec* rkx;
  const double2* bk;  // fast path: (kb, eb)   xb/sqrt5 = fma(mb, kb, eb)
  const int* nB;      // [S] segments that carry a BAF term
  const double* adr;  // [S] avg.depth.ratio
  const double* ploidy;  // [nP]
  const double* cell;    // [nCs] cellularities of the span
  int nP, nCs, nPT, nChunk;
  int PT, W;            // unit shape
  int u_first, nU;      // units of this launch, counted from the span's first group
  int rp;
  double t5;
  // where results go: outs[k][smp * out_sS + (column in span + out_col0) * nP + ploidy]
  int64_t out_sS;
  int64_t out_col0;
  int n_out;
  double* outs[kMaxOut];
};

// does the (span column, ploidy index) belong to the units of this launch?
__device__ __forceinline__ bool in_units(const BafGridArgs& a, int ic, int ip) {
  const int u = (ic / a.W) * a.nPT + ip / a.PT - a.u_first;
  return (unsigned)u < (unsigned)a.nU;
}

__device__ __forceinline__ void store_result(const BafGridArgs& a, int64_t smp, int ic, int ip,
                                             double v) {
  const int64_t o = smp * a.out_sS + ((int64_t)ic + a.out_col0) * a.nP + ip;
  for (int k = 0; k < a.n_out; ++k) a.outs[k][o] = v;
}

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
bool baf_colwarp_supported(int nc, int CNt_min, int T, bool rp);
size_t baf_colwarp_group_bytes(int nc, int64_t N_total);
int baf_colwarp_run(const BafGridArgs& a, const CnModel& m, int PT, int64_t N_total,
                    void* cw_buf, size_t cw_bytes, double* partM, long long* partE,
                    cudaStream_t st, unsigned long long* stats);

}  // namespace sqz
