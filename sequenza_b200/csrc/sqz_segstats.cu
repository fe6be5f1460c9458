// sqz_segstats.cu -- the per-segment statistics of segment.breaks (R/segments.R:113-198) and
// the kernel-density mode of b_allele_freq (R/model.R:69-98): the step that turns the seqz
// rows of a genome and its breakpoints into the segment table sequenza.fit consumes.
//
// Layout: the rows of all chromosomes concatenated (row_off[n_chr + 1]), sorted by position
// inside each chromosome, one fp64 vector per seqz column plus a byte per row for
// zygosity.normal == "het"; breaks_vect holds two values per segment, (start, end], the
// output of unique.breaks (R/segments.R:133-139), increasing inside a chromosome.  A segment
// is then a contiguous run of rows, found by two binary searches.
//
//   segstat_reduce_kernel  one CTA per segment, ONE pass over its rows: N.ratio, the weighted
//                          sums of the depth ratio, and sum / sum of squares in double-double
//                          (106 bits), from which var = (S2 - S1^2 / n) / (n - 1) comes out
//                          to ~1e-30 (mean^2 / var + 1) relative -- R's two-pass var() to far
//                          below the 1e-12 the tests ask; identical values give exactly 0, as
//                          two passes do.  At most 49 bytes per row: HBM-bound.
//                          The het rows' (Bf, Af, good.reads) are copied next to each other
//                          into the workspace (warp-aggregated slots, any order: everything
//                          downstream is order-independent), so that what concerns the B
//                          allele runs on full warps in the next kernel instead of on the 30 %
//                          of the lanes that hold a het row.
//   segstat_kde_kernel     one CTA per segment with at least two het positions: N.BAF's
//                          companions (sd.BAF; n, sum of good.reads, sd / min / max of
//                          c(Bf, Af)) from the compact rows, then
//                          quantile(type 7) by radix selection on the ordered bit patterns
//                          (the four order statistics of the two quartiles in the same eight
//                          passes),
//                          bw.nrd0, density.default's linear binning onto 512 points (64-bit
//                          fixed-point shared-memory atomics: order-independent), the
//                          correlation with dnorm summed directly (R does it by FFT), approx()
//                          onto the 512 output points, and the highest local maximum.
#include "sqz_common.cuh"
#include "../../include/sequenza_b200.h"

namespace sqz {

namespace {
constexpr int kSegThreads = 256;
constexpr int kDensN = 512;  // density.default's n

struct SegArgs {
  int64_t n_chr;
  const int64_t* row_off;  // [n_chr + 1]
  const int64_t* seg_off;  // [n_chr + 1]
  const double *pos, *ratio, *dnorm, *good, *Af, *Bf;
  const unsigned char* het;
  const double* brk;  // [2 * S]
  double min_reads;
  int weighted;
  int64_t S;
};

// per segment, handed from the reduction to the density kernel
struct SegMid {
  int64_t lo, hi;  // its rows
  int64_t nb;      // het rows with enough reads
  double x0;       // Bf of the first of them: x[1] of c(Bf, Af), bw.nrd0's last resort
};

// ---- double-double helpers -------------------------------------------------------------
struct D2 {
  double hi, lo;
};
__device__ __forceinline__ D2 d2_add(D2 a, double x) {  // a + x, x finite
  const double s = a.hi + x, bb = s - a.hi;
  const double e = (a.hi - (s - bb)) + (x - bb);
  const double lo = a.lo + e;
  const double h = s + lo;
  return D2{h, lo - (h - s)};
}
// streaming form: the high word takes the value, the rounding error of that one addition
// goes to the low word (two-sum), no renormalisation -- the chain from row to row is a
// single addition per accumulator; d2_merge at the end brings the pair back to normal form
__device__ __forceinline__ void d2_acc(D2& a, double x) {
  const double s = a.hi + x, bb = s - a.hi;
  a.lo += (a.hi - (s - bb)) + (x - bb);
  a.hi = s;
}
__device__ __forceinline__ void d2_acc_sq(D2& a, double x) {  // a += x^2
  const double p = x * x;
  d2_acc(a, p);
  a.lo += fma(x, x, -p);
}
__device__ __forceinline__ D2 d2_merge(D2 a, D2 b) { return d2_add(d2_add(a, b.hi), b.lo); }
__device__ __forceinline__ D2 d2_sq(D2 a) {  // a^2
  const double p = a.hi * a.hi;
  const double e = fma(a.hi, a.hi, -p) + 2.0 * (a.hi * a.lo);
  const double h = p + e;
  return D2{h, e - (h - p)};
}
__device__ __forceinline__ D2 d2_div(D2 a, double n) {  // a / n
  const double q = a.hi / n;
  const double r = fma(-q, n, a.hi) + a.lo;
  const double q2 = r / n;
  const double h = q + q2;
  return D2{h, q2 - (h - q)};
}

// stats::sd = sqrt(var) from the sums, n - 1 in the denominator, NA below two values
__device__ __forceinline__ double d2_sd(const D2& s1, const D2& s2, long long n, double mn,
                                        double mx) {
  if (n < 2) return na_real();
  if (mn == mx) return 0.0;  // identical values: exactly zero, as the two-pass var gives
  const D2 q = d2_div(d2_sq(s1), (double)n);
  const D2 r = d2_add(d2_add(s2, -q.hi), -q.lo);
  const double var = (r.hi + r.lo) / (double)(n - 1);
  return sqrt(var > 0 ? var : 0.0);
}

// first index in [a, b) whose position is > x (positions sorted; b if none), by the whole
// CTA: every round samples the range at kSegThreads points, so a chromosome of a million rows
// takes three dependent loads instead of twenty
__device__ __forceinline__ int64_t block_upper_bound(const double* p, int64_t a, int64_t b,
                                                     double x, int* sh) {
  while (b > a) {  // block-uniform
    const int64_t step = (b - a + kSegThreads - 1) / kSegThreads;
    const int64_t idx = a + (int64_t)threadIdx.x * step;
    if (threadIdx.x == 0) *sh = kSegThreads;
    __syncthreads();
    if (idx >= b || p[idx] > x) atomicMin(sh, (int)threadIdx.x);
    __syncthreads();
    const int first = *sh;  // first sample beyond x (or beyond the range)
    __syncthreads();
    const int64_t na = first > 0 ? a + (int64_t)(first - 1) * step + 1 : a;
    const int64_t nb = first < kSegThreads ? (a + (int64_t)first * step < b ? a + (int64_t)first * step : b) : b;
    a = na;
    b = nb;
  }
  return a;
}
}  // namespace

// ---- reductions of one segment (R/segments.R:140-170) -----------------------------------
__global__ void __launch_bounds__(kSegThreads, 3)
segstat_reduce_kernel(SegArgs a, SegMid* __restrict__ mid, double* __restrict__ cx,
                      double* __restrict__ cg, double* __restrict__ out) {
  __shared__ unsigned s_slot;
  __shared__ D2 red[4 * 8];
  __shared__ long long redl[3 * 8];
  __shared__ double redd[2 * 8];
  __shared__ int s_first;
  const int64_t s = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int64_t c = 0;
  while (c + 1 < a.n_chr && a.seg_off[c + 1] <= s) ++c;
  const int64_t r0 = a.row_off[c], r1 = a.row_off[c + 1];
  if (threadIdx.x == 0) s_slot = 0;
  // cut(position, breaks.vect): right-closed intervals (b0, b1] (R/segments.R:140-141)
  const int64_t lo = block_upper_bound(a.pos, r0, r1, a.brk[2 * s], &s_first);
  const int64_t hi = block_upper_bound(a.pos, lo, r1, a.brk[2 * s + 1], &s_first);
  __syncthreads();
  const bool W = a.weighted != 0;
  const double INF = __longlong_as_double(0x7FF0000000000000LL);
  const double NaN = __longlong_as_double(0x7FF8000000000000LL);

  // one pass: counts, sums and sums of squares of the depth ratio; het rows set aside
  D2 sum[4] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};  // 0 rw, 1 w.r, 2 v, 3 v^2
  long long nv = 0, nb = 0, first = 0x7fffffffffffffffLL;
  double vmin = INF, vmax = -INF;
  double2* cx2 = reinterpret_cast<double2*>(cx);  // the segment owns cx2[lo, hi) and cg[lo, hi)
  auto row = [&](int64_t i, double r, double dn, double g, bool het, double bf, double af) {
    double v;
    if (W) {
      const double wr = sqrt(dn), rw = r * wr;  // R/segments.R:116-117
      if (!isnan(rw)) d2_acc(sum[0], rw);
      if (!isnan(wr)) d2_acc(sum[1], wr);
      v = rw / wr;
    } else {
      v = r;
    }
    if (!isnan(v)) {
      d2_acc(sum[2], v);
      d2_acc_sq(sum[3], v);
      ++nv;
      vmin = fmin(vmin, v);
      vmax = fmax(vmax, v);
    }
    // R/segments.R:128-129: het rows with enough reads go to the compact copy, one slot
    // range per warp and batch row (every lane of the warp is here: uniform trip counts)
    const bool hb = het && g >= a.min_reads;
    const unsigned hm = __ballot_sync(0xffffffffu, hb);
    if (hm) {
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(&s_slot, (unsigned)__popc(hm));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (hb) {
        const int64_t slot = lo + base + __popc(hm & ((1u << lane) - 1u));
        cx2[slot] = make_double2(bf, af);
        cg[slot] = g;
        ++nb;
        if (i < first) first = i;
      }
    }
  };
  // four rows per thread in flight: the loads of a batch are issued before any of it is
  // consumed.  Rows past the end are read as NA / not het and change nothing.
  constexpr int kBatch = 4;
  for (int64_t b0 = lo; b0 < hi; b0 += kBatch * (int64_t)kSegThreads) {  // block-uniform
    double r[kBatch], dn[kBatch], g[kBatch], bf[kBatch], af[kBatch];
    bool het[kBatch];
#pragma unroll
    for (int k = 0; k < kBatch; ++k) {
      const int64_t ii = b0 + k * (int64_t)kSegThreads + threadIdx.x;
      const bool in = ii < hi;
      r[k] = in ? a.ratio[ii] : NaN;
      dn[k] = in && W ? a.dnorm[ii] : NaN;
      g[k] = in ? a.good[ii] : 0.;
      het[k] = in && a.het[ii] != 0;
      bf[k] = in ? a.Bf[ii] : 0.;
      af[k] = in ? a.Af[ii] : 0.;
    }
#pragma unroll
    for (int k = 0; k < kBatch; ++k)
      row(b0 + k * (int64_t)kSegThreads + threadIdx.x, r[k], dn[k], g[k], het[k], bf[k], af[k]);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) sum[k] = d2_add(D2{sum[k].hi, 0.}, sum[k].lo);  // normal form
  // Block totals, for thread 0 alone (it writes the segment's row): the chains go through
  // the shuffle levels side by side, one barrier, thread 0 folds the eight warps' partials.
#pragma unroll
  for (int off = 16; off; off >>= 1) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      D2 o;
      o.hi = __shfl_xor_sync(0xffffffffu, sum[k].hi, off);
      o.lo = __shfl_xor_sync(0xffffffffu, sum[k].lo, off);
      sum[k] = d2_merge(sum[k], o);
    }
    nv += __shfl_xor_sync(0xffffffffu, nv, off);
    nb += __shfl_xor_sync(0xffffffffu, nb, off);
    first = min(first, __shfl_xor_sync(0xffffffffu, first, off));
    vmin = fmin(vmin, __shfl_xor_sync(0xffffffffu, vmin, off));
    vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, off));
  }
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < 4; ++k) red[k * 8 + warp] = sum[k];
    redl[0 * 8 + warp] = nv;
    redl[1 * 8 + warp] = nb;
    redl[2 * 8 + warp] = first;
    redd[0 * 8 + warp] = vmin;
    redd[1 * 8 + warp] = vmax;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
#pragma unroll
  for (int w = 1; w < kSegThreads / 32; ++w) {
#pragma unroll
    for (int k = 0; k < 4; ++k) sum[k] = d2_merge(sum[k], red[k * 8 + w]);
    nv += redl[0 * 8 + w];
    nb += redl[1 * 8 + w];
    first = min(first, redl[2 * 8 + w]);
    vmin = fmin(vmin, redd[0 * 8 + w]);
    vmax = fmax(vmax, redd[1 * 8 + w]);
  }
  const D2 mv = nv > 0 ? d2_div(sum[2], (double)nv) : D2{0, 0};
  const int64_t S = a.S;
  const double x0 = nb >= 1 ? a.Bf[first] : 0.0;
  out[0 * S + s] = nb == 1 ? x0 : na_real();  // length(Bf) == 1 (R/model.R:93-94); nb >= 2:
  out[1 * S + s] = (double)nb;                // the density kernel fills Bf and sd.BAF in
  out[2 * S + s] = na_real();
  out[3 * S + s] = W ? (sum[0].hi + sum[0].lo) / (sum[1].hi + sum[1].lo)
                     : (nv > 0 ? mv.hi + mv.lo : NaN);
  out[4 * S + s] = (double)(hi - lo);
  out[5 * S + s] = d2_sd(sum[2], sum[3], nv, vmin, vmax);
  SegMid m;
  m.lo = lo;
  m.hi = hi;
  m.nb = nb;
  m.x0 = x0;
  mid[s] = m;
}

// ---- the mode of the kernel density of c(Bf, Af) (R/model.R:69-88) ------------------------
namespace {
// doubles as unsigned keys with the same order
__device__ __forceinline__ unsigned long long dkey(double x) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double dunkey(unsigned long long k) {
  const unsigned long long b = (k >> 63) ? (k & 0x7FFFFFFFFFFFFFFFULL) : ~k;
  return __longlong_as_double((long long)b);
}

// stats::dnorm(x, 0, sigma) (nmath/dnorm.c, R >= 3.1)
__device__ __forceinline__ double r_dnorm(double x, double sigma) {
  const double k = 0.398942280401432677939946059934;
  x = fabs(x / sigma);
  if (x < 5) return k * exp(-0.5 * x * x) / sigma;
  if (x > 38.56804181549334) return 0.;  // sqrt(-2 ln2 (DBL_MIN_EXP + 1 - DBL_MANT_DIG))
  const double x1 = ldexp(nearbyint(ldexp(x, 16)), -16);
  const double x2 = x - x1;
  return k / sigma * (exp(-0.5 * x1 * x1) * exp((-0.5 * x2 - x1) * x2));
}
}  // namespace

__global__ void __launch_bounds__(kSegThreads, 4)
segstat_kde_kernel(SegArgs a, const SegMid* __restrict__ mid, const double* __restrict__ cx,
                   const double* __restrict__ cg, double* __restrict__ out) {
  __shared__ unsigned hist[4][256];
  __shared__ unsigned long long s_prefix[4];
  __shared__ long long s_rank[4];
  __shared__ unsigned s_cnt[4], s_ncand[4];
  __shared__ unsigned long long ybin[kDensN + 1];  // 2^-62 fixed point
  __shared__ double K[kDensN + 1], kords[kDensN], gy[kDensN];  // gy doubles as the bins in fp64
  __shared__ double s_q[4];
  __shared__ D2 red[5 * 8];
  __shared__ long long redl[8];
  __shared__ double redd[4 * 8];
  __shared__ double s_stat[4];  // sum(good.reads), sd / min / max of c(Bf, Af)
  const int64_t s = blockIdx.x;
  const SegMid m = mid[s];
  if (m.nb < 2) return;  // 0: NA, 1: Bf itself -- written by the reduction kernel
  const int tid = threadIdx.x;
  const int64_t N = 2 * m.nb;

  const double* x = cx + 2 * m.lo;  // c(Bf, Af) of the het rows, interleaved: N values
  const double* gw = cg + m.lo;     // good.reads of the same rows

  // sd.BAF (R/segments.R:152-154) and what the density needs, from the compact rows: every
  // lane holds a het row here.  Sums in double-double as in the reduction kernel.
  {
    const double INF = __longlong_as_double(0x7FF0000000000000LL);
    const bool W = a.weighted != 0;
    const double2* x2 = reinterpret_cast<const double2*>(x);
    D2 sum[5] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}};  // vb, vb^2, good, x, x^2
    long long nvb = 0;
    double bmin = INF, bmax = -INF, xmin = INF, xmax = -INF;
    for (int64_t i = tid; i < m.nb; i += kSegThreads) {
      const double2 v = x2[i];
      const double g = gw[i];
      double vb = v.x;
      if (W) {
        const double wb = sqrt(g);  // R/segments.R:118-119
        vb = (v.x * wb) / wb;
      }
      if (!isnan(vb)) {
        d2_acc(sum[0], vb);
        d2_acc_sq(sum[1], vb);
        ++nvb;
        bmin = fmin(bmin, vb);
        bmax = fmax(bmax, vb);
      }
      d2_acc(sum[2], g);
      d2_acc(sum[3], v.x);
      d2_acc(sum[3], v.y);
      d2_acc_sq(sum[4], v.x);
      d2_acc_sq(sum[4], v.y);
      xmin = fmin(xmin, fmin(v.x, v.y));
      xmax = fmax(xmax, fmax(v.x, v.y));
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) sum[k] = d2_add(D2{sum[k].hi, 0.}, sum[k].lo);  // normal form
#pragma unroll
    for (int off = 16; off; off >>= 1) {
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        D2 o;
        o.hi = __shfl_xor_sync(0xffffffffu, sum[k].hi, off);
        o.lo = __shfl_xor_sync(0xffffffffu, sum[k].lo, off);
        sum[k] = d2_merge(sum[k], o);
      }
      nvb += __shfl_xor_sync(0xffffffffu, nvb, off);
      bmin = fmin(bmin, __shfl_xor_sync(0xffffffffu, bmin, off));
      bmax = fmax(bmax, __shfl_xor_sync(0xffffffffu, bmax, off));
      xmin = fmin(xmin, __shfl_xor_sync(0xffffffffu, xmin, off));
      xmax = fmax(xmax, __shfl_xor_sync(0xffffffffu, xmax, off));
    }
    const int lane = tid & 31, warp = tid >> 5;
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < 5; ++k) red[k * 8 + warp] = sum[k];
      redl[warp] = nvb;
      redd[0 * 8 + warp] = bmin;
      redd[1 * 8 + warp] = bmax;
      redd[2 * 8 + warp] = xmin;
      redd[3 * 8 + warp] = xmax;
    }
    __syncthreads();
    if (tid == 0) {
#pragma unroll
      for (int w = 1; w < kSegThreads / 32; ++w) {
#pragma unroll
        for (int k = 0; k < 5; ++k) sum[k] = d2_merge(sum[k], red[k * 8 + w]);
        nvb += redl[w];
        bmin = fmin(bmin, redd[0 * 8 + w]);
        bmax = fmax(bmax, redd[1 * 8 + w]);
        xmin = fmin(xmin, redd[2 * 8 + w]);
        xmax = fmax(xmax, redd[3 * 8 + w]);
      }
      out[2 * a.S + s] = d2_sd(sum[0], sum[1], nvb, bmin, bmax);  // column 2: sd.BAF
      s_stat[0] = sum[2].hi + sum[2].lo;
      s_stat[1] = d2_sd(sum[3], sum[4], N, xmin, xmax);
      s_stat[2] = xmin;
      s_stat[3] = xmax;
    }
    __syncthreads();
  }
  const double m_gsum = s_stat[0], m_sdx = s_stat[1], m_xmin = s_stat[2], m_xmax = s_stat[3];

  // quantile(x, c(.25, .75), type = 7): index = 1 + (n - 1) p; (1 - h) x[lo] + h x[hi].
  // The four order statistics by radix selection on the ordered bit patterns, 8 bits a
  // pass, most significant first, all four in the same passes.
  double qidx[2], qfl[2];
  long long want[4];
  for (int q = 0; q < 2; ++q) {
    qidx[q] = 1 + (double)(N - 1) * (q == 0 ? 0.25 : 0.75);
    qfl[q] = floor(qidx[q]);
    want[2 * q] = (long long)qfl[q] - 1;
    want[2 * q + 1] = (long long)ceil(qidx[q]) - 1;
  }
  if (tid < 4) {
    s_prefix[tid] = 0;
    s_rank[tid] = want[tid];
  }
  for (int shift = 56; shift >= 0; shift -= 8) {
    for (int t = tid; t < 4 * 256; t += kSegThreads) (&hist[0][0])[t] = 0;
    __syncthreads();
    const unsigned long long mask = shift == 56 ? 0ULL : (~0ULL << (shift + 8));
    const unsigned long long p0 = s_prefix[0], p1 = s_prefix[1], p2 = s_prefix[2],
                             p3 = s_prefix[3];
    for (int64_t i = tid; i < N; i += kSegThreads) {
      const unsigned long long k = dkey(x[i]);
      const unsigned long long hi_bits = k & mask;
      const unsigned d = (unsigned)(k >> shift) & 255u;
      if (hi_bits == p0) atomicAdd(&hist[0][d], 1u);
      if (hi_bits == p1) atomicAdd(&hist[1][d], 1u);
      if (hi_bits == p2) atomicAdd(&hist[2][d], 1u);
      if (hi_bits == p3) atomicAdd(&hist[3][d], 1u);
    }
    __syncthreads();
    if (tid < 4 * 32) {
      // warp w finds the digit of order statistic w: the bin where the running count passes
      // the rank.  A lane sums its eight bins, a shuffle scan gives the count below the
      // lane's bins, and the one lane whose bins hold the rank walks them.
      const int w = tid >> 5, lane = tid & 31;
      const long long r = s_rank[w];
      unsigned c8[8], tot = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        c8[k] = hist[w][lane * 8 + k];
        tot += c8[k];
      }
      unsigned incl = tot;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const unsigned o = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += o;
      }
      const long long below = (long long)(incl - tot);
      // the rank is below the total of the candidates, so exactly one lane qualifies; a
      // rank past the end (cannot happen) would leave the last bin, as the serial walk did
      const bool mine = (r >= below && r < below + (long long)tot) ||
                        (lane == 31 && r >= (long long)incl);
      if (mine) {
        long long rr = r - below;
        int b = 0;
        unsigned cb = c8[0];
#pragma unroll
        for (int k = 0; k < 7; ++k)
          if (b == k && rr >= (long long)c8[k]) {
            rr -= c8[k];
            ++b;
            cb = c8[k + 1];
          }
        s_rank[w] = rr;
        s_prefix[w] |= (unsigned long long)(lane * 8 + b) << shift;
        s_cnt[w] = cb;  // values that share the digits found so far
      }
    }
    __syncthreads();
    // Once few values are left around each of the four order statistics (three passes
    // usually: 24 leading bits), they are gathered and ranked directly instead of walking
    // all values five more times.  Many equal values keep a bin full: the passes go on.
    constexpr unsigned kCand = 128;  // per order statistic, in the histogram's memory
    if (shift <= 40 && shift > 0 &&
        max(max(s_cnt[0], s_cnt[1]), max(s_cnt[2], s_cnt[3])) <= kCand) {
      unsigned long long* cand = reinterpret_cast<unsigned long long*>(&hist[0][0]);
      static_assert(sizeof(hist) >= 4 * kCand * sizeof(unsigned long long), "candidate lists");
      if (tid < 4) s_ncand[tid] = 0;
      __syncthreads();
      const unsigned long long fixed = ~0ULL << shift;
      const unsigned long long q0 = s_prefix[0], q1 = s_prefix[1], q2 = s_prefix[2],
                               q3 = s_prefix[3];
      for (int64_t i = tid; i < N; i += kSegThreads) {
        const unsigned long long k = dkey(x[i]), hb = k & fixed;
        if (hb == q0) cand[0 * kCand + atomicAdd(&s_ncand[0], 1u)] = k;
        if (hb == q1) cand[1 * kCand + atomicAdd(&s_ncand[1], 1u)] = k;
        if (hb == q2) cand[2 * kCand + atomicAdd(&s_ncand[2], 1u)] = k;
        if (hb == q3) cand[3 * kCand + atomicAdd(&s_ncand[3], 1u)] = k;
      }
      __syncthreads();
      if (tid < 4 * 32) {
        const int w = tid >> 5, lane = tid & 31;
        const int nc = (int)s_ncand[w];
        const long long r = s_rank[w];
        const unsigned long long* cw = cand + w * kCand;
        for (int c = lane; c < nc; c += 32) {
          const unsigned long long kc = cw[c];
          long long rank = 0;
          for (int j = 0; j < nc; ++j) {
            const unsigned long long kj = cw[j];
            rank += (kj < kc || (kj == kc && j < c)) ? 1 : 0;
          }
          if (rank == r) s_prefix[w] = kc;
        }
      }
      __syncthreads();
      break;
    }
  }
  if (tid < 2) {
    const double xl = dunkey(s_prefix[2 * tid]), xh = dunkey(s_prefix[2 * tid + 1]);
    double v = xl;
    if (qidx[tid] > qfl[tid] && xh != xl) {
      const double h = qidx[tid] - qfl[tid];
      v = (1 - h) * xl + h * xh;
    }
    s_q[tid] = v;
  }
  __syncthreads();
  // bw.nrd0: 0.9 min(sd, IQR / 1.34) n^-0.2, with its fall-backs for a zero spread
  double lo_ = fmin(m_sdx, (s_q[1] - s_q[0]) / 1.34);
  if (!(lo_ != 0)) {
    lo_ = m_sdx;
    if (!(lo_ != 0)) {
      lo_ = fabs(m.x0);
      if (!(lo_ != 0)) lo_ = 1.;
    }
  }
  const double bw = 0.9 * lo_ * pow((double)N, -0.2);
  const double from = m_xmin - 3 * bw, to = m_xmax + 3 * bw;  // cut = 3
  const double lo = from - 4 * bw, up = to + 4 * bw;
  const int n = kDensN;

  // BinDist: linear binning of the weights good.reads / (2 sum(good.reads)) onto n points
  for (int j = tid; j <= n; j += kSegThreads) ybin[j] = 0;
  for (int d = tid; d <= n; d += kSegThreads) K[d] = r_dnorm(d * (2 * (up - lo) / (2 * n - 1)), bw);
  __syncthreads();
  {
    const double xdelta = (up - lo) / (n - 1), two_sum = 2 * m_gsum;
    const double SC = 4611686018427387904.0;  // 2^62
    const double2* x2 = reinterpret_cast<const double2*>(x);  // (Bf, Af) of a het row
    for (int64_t i = tid; i < m.nb; i += kSegThreads) {
      const double w = gw[i] / two_sum;  // the same weight for both values of the row
      const double2 xv2 = x2[i];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const double xv = h ? xv2.y : xv2.x;
        if (!isfinite(xv)) continue;
        const double xpos = (xv - lo) / xdelta;
        const int ix = (int)floor(xpos);
        const double fx = xpos - ix;
        if (0 <= ix && ix <= n - 2) {
          atomicAdd(&ybin[ix], (unsigned long long)(w * (1 - fx) * SC + 0.5));
          atomicAdd(&ybin[ix + 1], (unsigned long long)(w * fx * SC + 0.5));
        } else if (ix == -1) {
          atomicAdd(&ybin[0], (unsigned long long)(w * fx * SC + 0.5));
        } else if (ix == n - 1) {
          atomicAdd(&ybin[ix], (unsigned long long)(w * (1 - fx) * SC + 0.5));
        }
      }
    }
  }
  __syncthreads();
  // the bins as doubles, once (the same value the conversion inside the sum gave: the
  // 64-bit integer rounded to 53 bits, times 2^-62 exactly); gy is free until approx() below
  for (int j = tid; j < n; j += kSegThreads)
    gy[j] = (double)ybin[j] * (1.0 / 4611686018427387904.0);
  __syncthreads();
  // the correlation of the bins with dnorm(kords) (R: fft(fft(y) * Conj(fft(kords)))): the
  // kernel is symmetric and only the first n bins are occupied, so K[|m - j|] is all of it
  // A thread takes two neighbouring outputs: their kernel values are one apart, so a step of
  // the sum loads one bin and one kernel value for two products.  Every output still adds
  // its terms in increasing bin order, empty bins skipped, as before.
  static_assert(kDensN == 2 * kSegThreads, "two outputs per thread");
  {
    const int j0 = 2 * tid;  // outputs j0 and j0 + 1
    double acc0 = 0., acc1 = 0.;
    double kprev = K[j0 + 1];  // K[|mm - (j0 + 1)|] at mm = 0
#pragma unroll 4
    for (int mm = 0; mm < n; ++mm) {
      const int d0 = mm > j0 ? mm - j0 : j0 - mm;
      const double k0 = K[d0];  // = K[|mm - j0|]; the next step's K[|mm + 1 - (j0 + 1)|]
      const double y = gy[mm];
      if (y != 0.) {
        acc0 += y * k0;
        acc1 += y * kprev;
      }
      kprev = k0;
    }
    kords[j0] = acc0 > 0 ? acc0 : 0;
    kords[j0 + 1] = acc1 > 0 ? acc1 : 0;
  }
  __syncthreads();
  // approx(xords, kords, seq(from, to, length.out = n))
  {
    const double byx = (up - lo) / (n - 1), byg = (to - from) / (n - 1);
    for (int j = tid; j < n; j += kSegThreads) {
      const double v = j == n - 1 ? to : from + j * byg;
      int i = 0, jj = n - 1;
      auto xo = [&](int t) { return t == n - 1 ? up : lo + t * byx; };
      while (i < jj - 1) {
        const int ij = (i + jj) / 2;
        if (v < xo(ij)) jj = ij;
        else i = ij;
      }
      double y;
      if (v == xo(jj)) y = kords[jj];
      else if (v == xo(i)) y = kords[i];
      else y = kords[i] + (kords[jj] - kords[i]) * ((v - xo(i)) / (xo(jj) - xo(i)));
      gy[j] = y;
    }
  }
  __syncthreads();
  if (tid != 0) return;
  // the highest local maximum, first on ties (R/model.R:73-85); a top that is a run of equal
  // values counts at its first point (R's FFT noise makes one of them the maximum; after the
  // fold about 0.5 that is the same number to rounding)
  int best = -1, cand = -1;
  for (int j = 0; j + 1 < n; ++j) {
    const double d = gy[j + 1] - gy[j];
    if (d > 0) {
      cand = j + 1;
    } else if (d < 0) {
      if (cand >= 0 && (best < 0 || gy[cand] > gy[best])) best = cand;
      cand = -1;
    }
  }
  if (best < 0) {
    for (int j = 1; j < n - 1; ++j) {
      const double d0 = gy[j] - gy[j - 1], d1 = gy[j + 1] - gy[j];
      const int s0 = (d0 > 0) - (d0 < 0), s1 = (d1 > 0) - (d1 < 0);
      if (s1 - s0 == -1 && (best < 0 || gy[j] > gy[best])) best = j;
    }
  }
  double b_val = na_real();
  if (best >= 0) {
    b_val = best == n - 1 ? to : from + best * ((to - from) / (n - 1));
    if (b_val > 0.5) b_val = 1 - b_val;  // R/model.R:86-88
  }
  out[s] = b_val;  // column 0: Bf
}

// ---- host side ---------------------------------------------------------------------------
static size_t seg_a256(size_t x) { return (x + 255) & ~(size_t)255; }
// per segment hand-over + the compact copy of the het rows (24 bytes per row at most)
size_t segment_stats_workspace_bytes(int64_t n_rows, int64_t n_seg) {
  const size_t n = (size_t)(n_rows > 0 ? n_rows : 1);
  return seg_a256((size_t)(n_seg > 0 ? n_seg : 1) * sizeof(SegMid)) + seg_a256(16 * n) +
         seg_a256(8 * n);
}

int segment_stats_launch(int64_t n_chr, const int64_t* d_row_off, const int64_t* d_seg_off,
                         int64_t n_rows, int64_t n_seg, const double* d_pos, const double* d_ratio,
                         const double* d_dnorm, const double* d_good, const double* d_Af,
                         const double* d_Bf, const unsigned char* d_het, const double* d_brk,
                         double min_reads_baf, int weighted_mean, void* workspace,
                         size_t workspace_bytes, cudaStream_t st, double* d_out) {
  if (n_chr < 0 || n_seg < 0) return fail(SQZ_ERR_ARG, "negative size");
  if (n_seg == 0) return SQZ_OK;
  if (n_seg > 0x7fffffffLL) return fail(SQZ_ERR_SIZE, "more than 2^31-1 segments");
  if (!workspace || ((uintptr_t)workspace & 255) ||
      workspace_bytes < segment_stats_workspace_bytes(n_rows, n_seg))
    return fail(SQZ_ERR_SIZE, "segment-statistics workspace too small or not 256-byte aligned");
  SegArgs a;
  a.n_chr = n_chr;
  a.row_off = d_row_off;
  a.seg_off = d_seg_off;
  a.pos = d_pos;
  a.ratio = d_ratio;
  a.dnorm = d_dnorm;
  a.good = d_good;
  a.Af = d_Af;
  a.Bf = d_Bf;
  a.het = d_het;
  a.brk = d_brk;
  a.min_reads = min_reads_baf;
  a.weighted = weighted_mean ? 1 : 0;
  a.S = n_seg;
  SegMid* mid = (SegMid*)workspace;
  double* cx = (double*)((char*)workspace + seg_a256((size_t)n_seg * sizeof(SegMid)));
  double* cg = (double*)((char*)cx + seg_a256(16 * (size_t)(n_rows > 0 ? n_rows : 1)));
  segstat_reduce_kernel<<<(unsigned)n_seg, kSegThreads, 0, st>>>(a, mid, cx, cg, d_out);
  SQZ_CUDA(cudaGetLastError());
  segstat_kde_kernel<<<(unsigned)n_seg, kSegThreads, 0, st>>>(a, mid, cx, cg, d_out);
  SQZ_CUDA(cudaGetLastError());
  return SQZ_OK;
}

}  // namespace sqz
