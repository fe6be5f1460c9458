// sqz_api.cu -- extern "C" surface of libsequenza_b200.so (see include/sequenza_b200.h).
//
// Host entry points stage R-owned host vectors into cached per-device buffers, run the
// kernels on a private stream per device and copy each device's slice of the result
// straight into the caller's output (no device-to-device gather: the consumer is the
// R process).  The _dev entry points are thin argument checks over the launchers.
#include <algorithm>
#include <mutex>
#include <thread>
#include <vector>

#include "sqz_common.cuh"
#include "../../include/sequenza_b200.h"

namespace sqz {

static thread_local std::string g_err;
int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}

// launchers (sqz_baf_grid.cu, sqz_mufreq.cu)
size_t baf_grid_workspace_bytes(int64_t N_total, int64_t S, int64_t nP, int64_t nCs);
size_t baf_grid_workspace_preferred_bytes(int64_t N_total, int64_t S, int64_t nP, int64_t nC,
                                          int64_t nCs, int n_copy_numbers, size_t cap_bytes);
size_t baf_grid_units_workspace_bytes(int64_t N_total, int64_t S, const double* h_ploidy,
                                      int64_t nP, int64_t nC, int CNt_min, int CNt_max, int CNn,
                                      int ratio_priority, int64_t u0, int64_t u1,
                                      size_t cap_bytes);
int baf_grid_launch_units(int64_t S, const int64_t* d_seg_off, int64_t N_total,
                          const double* d_Bf, const double* d_ratio, const double* d_sdBf,
                          const double* d_sdr, const double* d_wBf, const double* d_wr,
                          const double* h_ploidy, int64_t nP, const double* h_cell, int64_t nC,
                          int64_t u0, int64_t u1, const double* d_adr, int CNt_min,
                          int CNt_max, int CNn, const double* prior_CN,
                          const double* prior_value, int n_prior, int ratio_priority, int mode,
                          void* workspace, size_t workspace_bytes, cudaStream_t st,
                          double* const* d_outs, int n_out);
int baf_grid_launch(int64_t S, const int64_t* d_seg_off, int64_t N_total, const double* d_Bf,
                    const double* d_ratio, const double* d_sdBf, const double* d_sdr,
                    const double* d_wBf, const double* d_wr, const double* h_ploidy,
                    int64_t nP, const double* h_cell, int64_t nC, int64_t ic0, int64_t ic1,
                    const double* d_adr, int CNt_min, int CNt_max, int CNn,
                    const double* prior_CN, const double* prior_value, int n_prior,
                    int ratio_priority, int mode, void* workspace, size_t workspace_bytes,
                    cudaStream_t st, double* d_out, bool allow_profile = true);
int baf_point_launch(const double* d_Bf, const double* d_ratio, const double* d_sdBf,
                     const double* d_sdr, const double* d_wBf, const double* d_wr, int64_t N,
                     double cellularity, double ploidy, double adr, int CNt_min, int CNt_max,
                     int CNn, const double* prior_CN, const double* prior_value, int n_prior,
                     int ratio_priority, cudaStream_t st, double* d_out);
int baf_grid_plan(int64_t N_total, int64_t S, const double* h_ploidy, int64_t nP, int64_t nC,
                  int CNt_min, int CNt_max, int CNn, int ratio_priority, int64_t* out8);
int profile_enable(int on);
int profile_collect(double* ms3, int* n_calls);
int profile_counters(int* path, unsigned long long* n_eval, long long* launches);
size_t mufreq_grid_workspace_bytes(int64_t N, int64_t nP, int64_t nCs);
int mufreq_point_launch(const double* d_F, const double* d_ratio, const double* d_wm,
                        const double* d_wr, int64_t N, double cellularity, double ploidy,
                        double adr, int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                        const double* prior_value, int n_prior, cudaStream_t st,
                        double* d_out);
int mufreq_grid_launch(const double* d_F, const double* d_ratio, const double* d_wm,
                       const double* d_wr, int64_t N, const double* h_ploidy, int64_t nP,
                       const double* h_cell, int64_t nC, int64_t ic0, int64_t ic1, double adr,
                       int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                       const double* prior_value, int n_prior, void* workspace,
                       size_t workspace_bytes, cudaStream_t st, double* d_out);
size_t mufreq_scalar_table_bytes(double wm, double wr);
int mufreq_point_scalar_launch(const double* d_F, const double* d_ratio, int64_t N, double wm,
                               double wr, double cellularity, double ploidy, double adr,
                               int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                               const double* prior_value, int n_prior, void* table,
                               size_t table_bytes, cudaStream_t st, double* d_out);
int fit_finish_launch(const double* d_lpp_sum, int64_t S, int64_t nP, int64_t nC,
                      double* d_znorm, double* d_row_max, double* d_col_max,
                      double* d_row_above, double* d_col_above, double* d_lsl,
                      cudaStream_t st);
size_t segment_stats_workspace_bytes(int64_t n_rows, int64_t n_seg);
int segment_stats_launch(int64_t n_chr, const int64_t* d_row_off, const int64_t* d_seg_off,
                         int64_t n_rows, int64_t n_seg, const double* d_pos, const double* d_ratio,
                         const double* d_dnorm, const double* d_good, const double* d_Af,
                         const double* d_Bf, const unsigned char* d_het, const double* d_brk,
                         double min_reads_baf, int weighted_mean, void* workspace,
                         size_t workspace_bytes, cudaStream_t st, double* d_out);
int model_points_launch(int kind, double cellularity, double ploidy, double adr, int CNt_min,
                        int CNt_max, int CNn, const double* prior_CN,
                        const double* prior_value, int n_prior, int* T_out, double* h_out);

// ---- per-device cache used by the host entry points --------------------------------
struct DevCtx {
  bool live = false;
  cudaStream_t st = nullptr;
  cudaEvent_t ev_in = nullptr, ev_done = nullptr;  // inputs staged / work finished
  void* in = nullptr;   // staged inputs
  size_t in_cap = 0;
  void* ws = nullptr;   // kernel workspace
  size_t ws_cap = 0;
  void* out = nullptr;  // result slice
  size_t out_cap = 0;
  unsigned long long peers = 0;  // devices whose memory this device may address
};
static std::mutex g_mu;
static std::vector<DevCtx> g_ctx;

static int grow(void** p, size_t* cap, size_t need) {
  if (need <= *cap) return SQZ_OK;
  if (*p) SQZ_CUDA(cudaFree(*p));
  *p = nullptr;
  *cap = 0;
  // a little headroom for small buffers that creep up; none for multi-GiB requests
  size_t want = need + (need < ((size_t)256 << 20) ? need / 4 : 0) + 256;
  SQZ_CUDA(cudaMalloc(p, want));
  *cap = want;
  return SQZ_OK;
}

// Workspace of a grid fit: the preferred size if the device has it, else halved towards
// the minimum (fewer pre-pass columns per batch; the results do not depend on it).
static int grow_ws(DevCtx* c, size_t need_min, size_t need_pref, size_t* got) {
  size_t want = need_pref > need_min ? need_pref : need_min;
  if (want <= c->ws_cap) {
    *got = want;
    return SQZ_OK;
  }
  for (;;) {
    if (c->ws) cudaFree(c->ws);
    c->ws = nullptr;
    c->ws_cap = 0;
    cudaError_t e = cudaMalloc(&c->ws, want + 256);
    if (e == cudaSuccess) {
      c->ws_cap = want + 256;
      *got = want;
      return SQZ_OK;
    }
    cudaGetLastError();  // out of memory: clear it and ask for less
    if (e != cudaErrorMemoryAllocation || want == need_min)
      return fail(SQZ_ERR_CUDA, std::string("cudaMalloc of the grid workspace failed: ") +
                                    cudaGetErrorString(e));
    want = need_min + (want - need_min) / 2;
    if (want - need_min < ((size_t)1 << 20)) want = need_min;
  }
}

static int ctx_for(int dev, DevCtx** out) {
  if ((int)g_ctx.size() <= dev) g_ctx.resize(dev + 1);
  DevCtx& c = g_ctx[dev];
  SQZ_CUDA(cudaSetDevice(dev));
  if (!c.live) {
    SQZ_CUDA(cudaStreamCreateWithFlags(&c.st, cudaStreamNonBlocking));
    SQZ_CUDA(cudaEventCreateWithFlags(&c.ev_in, cudaEventDisableTiming));
    SQZ_CUDA(cudaEventCreateWithFlags(&c.ev_done, cudaEventDisableTiming));
    c.live = true;
  }
  *out = &c;
  return SQZ_OK;
}

// may device `dev` (current) address memory of device `peer`?  Enabled on first use.
static bool peer_ok(int dev, int peer) {
  if (dev == peer) return true;
  if (dev >= 64 || peer >= 64) return false;
  DevCtx& c = g_ctx[dev];
  if (c.peers & (1ull << peer)) return true;
  int can = 0;
  if (cudaDeviceCanAccessPeer(&can, dev, peer) != cudaSuccess || !can) {
    cudaGetLastError();
    return false;
  }
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return false;
  }
  cudaGetLastError();
  c.peers |= 1ull << peer;
  return true;
}

static int usable_devices(int n_gpus, int* ndev) {
  int have = 0;
  cudaError_t e = cudaGetDeviceCount(&have);
  if (e != cudaSuccess || have < 1) {
    cudaGetLastError();
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  }
  if (n_gpus < 1) return fail(SQZ_ERR_ARG, "n_gpus must be >= 1");
  *ndev = n_gpus < have ? n_gpus : have;
  return SQZ_OK;
}

static inline size_t a256(size_t x) { return (x + 255) & ~(size_t)255; }

struct DeviceGuard {
  int prev = -1;
  DeviceGuard() { cudaGetDevice(&prev); }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// wait for everything queued on the first ndev devices (also on error paths: nothing may
// still be running on caller-visible buffers when a host entry point returns)
static cudaError_t sync_devices(int ndev) {
  cudaError_t first = cudaSuccess;
  for (int d = 0; d < ndev && d < (int)g_ctx.size(); ++d) {
    if (!g_ctx[d].live) continue;
    cudaError_t e = cudaSetDevice(d);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g_ctx[d].st);
    if (e != cudaSuccess && first == cudaSuccess) first = e;
    cudaGetLastError();
  }
  return first;
}

struct GridCall {  // arguments every device of a host grid call shares
  int64_t S;
  const int64_t* seg_off;
  const double* const* vec;
  const double* ploidy;
  int64_t nP;
  const double* cell;
  int64_t nC;
  const double* adr;
  int CNt_min, CNt_max, CNn;
  const double* prior_CN;
  const double* prior_value;
  int n_prior, ratio_priority, mode;
};

// One sample table, several devices: the grid is cut into units (sqz_baf_grid.cuh) and
// every device writes its results straight into device 0's whole-grid buffer over NVLink
// (peer access).  The segment table crosses PCIe once, to device 0, and reaches the other
// devices by peer copies; each device is driven by its own host thread; the assembled
// LPP vector leaves device 0 in one copy.
static int baf_grid_host_units(const GridCall& g, int ndev, double* out) {
  const int64_t N = g.seg_off[g.S], G = g.nP * g.nC;
  int64_t plan[8];
  int rc = baf_grid_plan(N, g.S, g.ploidy, g.nP, g.nC, g.CNt_min, g.CNt_max, g.CNn,
                         g.ratio_priority, plan);
  if (rc) return rc;
  const int64_t n_units = plan[7];
  if (ndev > n_units) ndev = (int)n_units;
  const size_t vbytes = a256((size_t)(N > 0 ? N : 1) * 8);
  const size_t off_b = a256((size_t)(g.S + 1) * 8), adr_b = a256((size_t)g.S * 8);
  const size_t in_need = 6 * vbytes + off_b + adr_b;
  for (int d = 0; d < ndev; ++d) {
    DevCtx* c;
    if ((rc = ctx_for(d, &c))) return rc;
    if ((rc = grow(&c->in, &c->in_cap, in_need))) return rc;
  }
  DevCtx* root = &g_ctx[0];
  SQZ_CUDA(cudaSetDevice(0));
  if ((rc = grow(&root->out, &root->out_cap, (size_t)(g.S * G) * 8))) return rc;
  {  // host -> device 0, once
    char* p = (char*)root->in;
    for (int k = 0; k < 6; ++k, p += vbytes)
      if (N > 0)
        SQZ_CUDA(cudaMemcpyAsync(p, g.vec[k], (size_t)N * 8, cudaMemcpyHostToDevice, root->st));
    SQZ_CUDA(cudaMemcpyAsync(p, g.seg_off, (size_t)(g.S + 1) * 8, cudaMemcpyHostToDevice, root->st));
    SQZ_CUDA(cudaMemcpyAsync(p + off_b, g.adr, (size_t)g.S * 8, cudaMemcpyHostToDevice, root->st));
    SQZ_CUDA(cudaEventRecord(root->ev_in, root->st));
  }
  std::vector<int> rcs(ndev, SQZ_OK);
  std::vector<std::string> msgs(ndev);
  auto run_dev = [&](int d) {
    auto body = [&]() -> int {
      DevCtx* c = &g_ctx[d];
      SQZ_CUDA(cudaSetDevice(d));
      if (d != 0) {
        if (!peer_ok(d, 0)) return fail(SQZ_ERR_CUDA, "no peer access to device 0");
        SQZ_CUDA(cudaStreamWaitEvent(c->st, root->ev_in, 0));
        SQZ_CUDA(cudaMemcpyPeerAsync(c->in, d, root->in, 0, in_need, c->st));
      }
      const int64_t u0 = n_units * d / ndev, u1 = n_units * (d + 1) / ndev;
      const size_t ws_min = baf_grid_units_workspace_bytes(N, g.S, g.ploidy, g.nP, g.nC, g.CNt_min,
                                                           g.CNt_max, g.CNn, g.ratio_priority, u0,
                                                           u1, 1);
      const size_t ws_pref = baf_grid_units_workspace_bytes(N, g.S, g.ploidy, g.nP, g.nC,
                                                            g.CNt_min, g.CNt_max, g.CNn,
                                                            g.ratio_priority, u0, u1,
                                                            (size_t)16 << 30);
      size_t ws_got = 0;
      int r = grow_ws(c, ws_min, ws_pref, &ws_got);
      if (r) return r;
      const char* p = (const char*)c->in;
      const double* dv[6];
      for (int k = 0; k < 6; ++k, p += vbytes) dv[k] = (const double*)p;
      double* outs[1] = {(double*)root->out};
      r = baf_grid_launch_units(g.S, (const int64_t*)p, N, dv[0], dv[1], dv[2], dv[3], dv[4],
                                dv[5], g.ploidy, g.nP, g.cell, g.nC, u0, u1,
                                (const double*)(p + off_b), g.CNt_min, g.CNt_max, g.CNn,
                                g.prior_CN, g.prior_value, g.n_prior, g.ratio_priority, g.mode,
                                c->ws, ws_got, c->st, outs, 1);
      if (r) return r;
      SQZ_CUDA(cudaEventRecord(c->ev_done, c->st));
      return SQZ_OK;
    };
    rcs[d] = body();
    if (rcs[d]) msgs[d] = g_err;
  };
  {
    std::vector<std::thread> th;
    for (int d = 1; d < ndev; ++d) th.emplace_back(run_dev, d);
    run_dev(0);
    for (auto& t : th) t.join();
  }
  for (int d = 0; d < ndev; ++d)
    if (rcs[d]) {
      sync_devices(ndev);
      return fail(rcs[d], msgs[d]);
    }
  SQZ_CUDA(cudaSetDevice(0));
  for (int d = 1; d < ndev; ++d) SQZ_CUDA(cudaStreamWaitEvent(root->st, g_ctx[d].ev_done, 0));
  SQZ_CUDA(cudaMemcpyAsync(out, root->out, (size_t)(g.S * G) * 8, cudaMemcpyDeviceToHost,
                           root->st));
  cudaError_t e = cudaStreamSynchronize(root->st);
  if (e != cudaSuccess) {
    sync_devices(ndev);
    return fail(SQZ_ERR_CUDA, std::string("grid fit failed: ") + cudaGetErrorString(e));
  }
  return SQZ_OK;
}

static int baf_grid_host(int64_t S, const int64_t* seg_off, const double* const vec[6],
                         const double* ploidy, int64_t nP, const double* cell, int64_t nC,
                         const double* adr, int CNt_min, int CNt_max, int CNn,
                         const double* prior_CN, const double* prior_value, int n_prior,
                         int ratio_priority, int n_gpus, int mode, double* out) {
  if (S < 0 || nP < 0 || nC < 0) return fail(SQZ_ERR_ARG, "negative size");
  if (S > 0 && !seg_off) return fail(SQZ_ERR_ARG, "seg_offsets is null");
  const int64_t N = S > 0 ? seg_off[S] : 0;
  if (S > 0 && seg_off[0] != 0) return fail(SQZ_ERR_ARG, "seg_offsets[0] must be 0");
  for (int64_t k = 0; k < S; ++k)
    if (seg_off[k + 1] < seg_off[k]) return fail(SQZ_ERR_ARG, "seg_offsets must be non-decreasing");
  for (int k = 0; k < 6; ++k)
    if (N > 0 && !vec[k]) return fail(SQZ_ERR_ARG, "a per-segment vector is null");
  if ((nP > 0 && !ploidy) || (nC > 0 && !cell) || (S > 0 && !adr) || (S * nP * nC > 0 && !out))
    return fail(SQZ_ERR_ARG, "null pointer argument");
  if (S == 0 || nP == 0 || nC == 0) return SQZ_OK;
  int ndev = 0;
  int rc = usable_devices(n_gpus, &ndev);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  const int64_t G = nP * nC;
  // Shards.  A cohort (at least one sample per device) is cut by sample at
  // segment-balanced boundaries, so that each device receives only its own samples'
  // segments.  Otherwise every device works on all segments and the grid is cut into
  // units (baf_grid_host_units) or, without peer access between the devices, into
  // contiguous blocks of cellularity columns.
  const bool by_sample = ndev > 1 && S >= ndev;
  if (!by_sample && ndev > 1) {
    bool peers = true;
    for (int d = 1; d < ndev && peers; ++d) {
      DevCtx* c;
      if ((rc = ctx_for(d, &c))) return rc;
      peers = peer_ok(d, 0);
    }
    if (peers) {
      GridCall g{S, seg_off, vec, ploidy, nP, cell, nC, adr, CNt_min, CNt_max, CNn,
                 prior_CN, prior_value, n_prior, ratio_priority, mode};
      return baf_grid_host_units(g, ndev, out);
    }
  }
  if (!by_sample && ndev > nC) ndev = (int)nC;
  std::vector<int64_t> c0(ndev + 1), s0(ndev + 1);
  for (int d = 0; d <= ndev; ++d) {
    c0[d] = by_sample ? (d == 0 ? 0 : nC) : nC * d / ndev;
    if (!by_sample) {
      s0[d] = d == 0 ? 0 : S;
    } else if (N == 0) {
      s0[d] = S * d / ndev;
    } else {
      s0[d] = std::lower_bound(seg_off, seg_off + S + 1, N / ndev * d + N % ndev * d / ndev) -
              seg_off;
    }
  }
  s0[0] = 0;
  s0[ndev] = S;
  std::vector<int64_t> off_local(by_sample ? S + ndev : 0);
  auto queue_device = [&](int d) -> int {
    // device d: samples [sa, sb) x columns [ca, cb)
    const int64_t sa = by_sample ? s0[d] : 0, sb = by_sample ? s0[d + 1] : S;
    const int64_t ca = by_sample ? 0 : c0[d], cb = by_sample ? nC : c0[d + 1];
    const int64_t Sd = sb - sa, nCs = cb - ca;
    if (Sd <= 0 || nCs <= 0) return SQZ_OK;
    const int64_t n_lo = seg_off[sa], Nd = seg_off[sb] - n_lo;
    const int64_t* h_off = seg_off;
    if (by_sample) {
      int64_t* o = off_local.data() + sa + d;
      for (int64_t k = 0; k <= Sd; ++k) o[k] = seg_off[sa + k] - n_lo;
      h_off = o;
    }
    DevCtx* c;
    int r;
    if ((r = ctx_for(d, &c))) return r;
    const size_t vbytes = a256((size_t)(Nd > 0 ? Nd : 1) * 8);
    const size_t in_need = 6 * vbytes + a256((size_t)(Sd + 1) * 8) + a256((size_t)Sd * 8);
    if ((r = grow(&c->in, &c->in_cap, in_need))) return r;
    const int ncn = CNt_max - CNt_min + 1;
    size_t ws_got = 0;
    if ((r = grow_ws(c, baf_grid_workspace_preferred_bytes(Nd, Sd, nP, nC, nCs, ncn, 1),
                     baf_grid_workspace_preferred_bytes(Nd, Sd, nP, nC, nCs, ncn,
                                                        (size_t)16 << 30), &ws_got)))
      return r;
    if ((r = grow(&c->out, &c->out_cap, (size_t)(Sd * nCs * nP) * 8))) return r;
    char* p = (char*)c->in;
    const double* dv[6];
    for (int k = 0; k < 6; ++k) {
      dv[k] = (const double*)p;
      if (Nd > 0)
        SQZ_CUDA(cudaMemcpyAsync(p, vec[k] + n_lo, (size_t)Nd * 8, cudaMemcpyHostToDevice, c->st));
      p += vbytes;
    }
    int64_t* d_off = (int64_t*)p;
    SQZ_CUDA(cudaMemcpyAsync(d_off, h_off, (size_t)(Sd + 1) * 8, cudaMemcpyHostToDevice, c->st));
    p += a256((size_t)(Sd + 1) * 8);
    double* d_adr = (double*)p;
    SQZ_CUDA(cudaMemcpyAsync(d_adr, adr + sa, (size_t)Sd * 8, cudaMemcpyHostToDevice, c->st));
    return baf_grid_launch(Sd, d_off, Nd, dv[0], dv[1], dv[2], dv[3], dv[4], dv[5], ploidy, nP,
                           cell, nC, ca, cb, d_adr, CNt_min, CNt_max, CNn, prior_CN,
                           prior_value, n_prior, ratio_priority, mode, c->ws, ws_got, c->st,
                           (double*)c->out);
  };
  auto read_back = [&](int d) -> int {
    const int64_t sa = by_sample ? s0[d] : 0, sb = by_sample ? s0[d + 1] : S;
    const int64_t ca = by_sample ? 0 : c0[d], cb = by_sample ? nC : c0[d + 1];
    const int64_t Sd = sb - sa, nCs = cb - ca;
    if (Sd <= 0 || nCs <= 0) return SQZ_OK;
    DevCtx* c = &g_ctx[d];
    SQZ_CUDA(cudaSetDevice(d));
    // device slice [Sd][nCs*nP] -> host rows sa.. of [S][G] at column offset ca*nP
    SQZ_CUDA(cudaMemcpy2DAsync(out + sa * G + ca * nP, (size_t)G * 8, c->out,
                               (size_t)(nCs * nP) * 8, (size_t)(nCs * nP) * 8, (size_t)Sd,
                               cudaMemcpyDeviceToHost, c->st));
    return SQZ_OK;
  };
  for (int d = 0; d < ndev && !rc; ++d) rc = queue_device(d);
  // Results are copied only after every device has its work queued: a device-to-host
  // copy into pageable memory blocks the calling thread until that device is done.
  for (int d = 0; d < ndev && !rc; ++d) rc = read_back(d);
  std::string msg = rc ? g_err : std::string();
  // also after an error: nothing stays queued behind the caller's back
  const cudaError_t e = sync_devices(ndev);
  if (rc) return fail(rc, msg);
  if (e != cudaSuccess)
    return fail(SQZ_ERR_CUDA, std::string("grid fit failed: ") + cudaGetErrorString(e));
  return SQZ_OK;
}

// ---- fp64 issue-rate probe -----------------------------------------------------------
__global__ void __launch_bounds__(256) fp64_peak_kernel(double* out, int iters, double seed) {
  double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4,
         a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 0.999999, c = 1e-9;
  for (int i = 0; i < iters; ++i) {
    a0 = fma(a0, m, c);
    a1 = fma(a1, m, c);
    a2 = fma(a2, m, c);
    a3 = fma(a3, m, c);
    a4 = fma(a4, m, c);
    a5 = fma(a5, m, c);
    a6 = fma(a6, m, c);
    a7 = fma(a7, m, c);
  }
  double r = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (r == 12345.678) out[0] = r;  // keeps the chains alive
}

// single-device host wrapper shared by the point entry points: stage k vectors of
// length N, run `launch`, copy N x 4 back.
template <class F>
static int point_host(int k, const double* const* vec, int64_t N, double* out, F launch) {
  if (N < 0) return fail(SQZ_ERR_ARG, "negative N");
  for (int i = 0; i < k; ++i)
    if (N > 0 && !vec[i]) return fail(SQZ_ERR_ARG, "a per-row vector is null");
  if (N > 0 && !out) return fail(SQZ_ERR_ARG, "out is null");
  int ndev = 0;
  int rc = usable_devices(1, &ndev);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  int cur = 0;
  SQZ_CUDA(cudaGetDevice(&cur));
  DevCtx* c;
  if ((rc = ctx_for(cur, &c))) return rc;
  const size_t vbytes = a256((size_t)(N > 0 ? N : 1) * 8);
  if ((rc = grow(&c->in, &c->in_cap, k * vbytes))) return rc;
  if ((rc = grow(&c->out, &c->out_cap, (size_t)(N > 0 ? N : 1) * 32))) return rc;
  const double* dv[8];
  char* p = (char*)c->in;
  for (int i = 0; i < k; ++i) {
    dv[i] = (const double*)p;
    if (N > 0) SQZ_CUDA(cudaMemcpyAsync(p, vec[i], (size_t)N * 8, cudaMemcpyHostToDevice, c->st));
    p += vbytes;
  }
  rc = launch(dv, c->st, (double*)c->out);
  if (rc) return rc;
  if (N > 0)
    SQZ_CUDA(cudaMemcpyAsync(out, c->out, (size_t)N * 32, cudaMemcpyDeviceToHost, c->st));
  SQZ_CUDA(cudaStreamSynchronize(c->st));
  return SQZ_OK;
}

}  // namespace sqz

using namespace sqz;

extern "C" {

const char* sqz_version(void) { return "sequenza_b200 0.1.0 (sm_100a)"; }
const char* sqz_last_error(void) { return g_err.c_str(); }

int sqz_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int sqz_release(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  for (size_t d = 0; d < g_ctx.size(); ++d) {
    DevCtx& c = g_ctx[d];
    if (!c.live) continue;
    cudaSetDevice((int)d);
    cudaStreamSynchronize(c.st);
    if (c.in) cudaFree(c.in);
    if (c.ws) cudaFree(c.ws);
    if (c.out) cudaFree(c.out);
    cudaStreamDestroy(c.st);
    if (c.ev_in) cudaEventDestroy(c.ev_in);
    if (c.ev_done) cudaEventDestroy(c.ev_done);
    c = DevCtx();
  }
  return SQZ_OK;
}

int sqz_baf_grid(const double* Bf, const double* depth_ratio, const double* sd_Bf,
                 const double* sd_ratio, const double* weight_Bf, const double* weight_ratio,
                 int64_t N, const double* ploidy, int64_t nP, const double* cellularity,
                 int64_t nC, double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                 const double* prior_CN, const double* prior_value, int n_prior,
                 int ratio_priority, int n_gpus, int mode, double* lpp_sum_out) {
  if (N < 0) return fail(SQZ_ERR_ARG, "negative N");
  const int64_t off[2] = {0, N};
  const double* vec[6] = {Bf, depth_ratio, sd_Bf, sd_ratio, weight_Bf, weight_ratio};
  return baf_grid_host(1, off, vec, ploidy, nP, cellularity, nC, &avg_depth_ratio, CNt_min,
                       CNt_max, CNn, prior_CN, prior_value, n_prior, ratio_priority, n_gpus,
                       mode, lpp_sum_out);
}

int sqz_baf_grid_batch(int64_t S, const int64_t* seg_offsets, const double* Bf,
                       const double* depth_ratio, const double* sd_Bf, const double* sd_ratio,
                       const double* weight_Bf, const double* weight_ratio,
                       const double* ploidy, int64_t nP, const double* cellularity,
                       int64_t nC, const double* avg_depth_ratio, int CNt_min, int CNt_max,
                       int CNn, const double* prior_CN, const double* prior_value,
                       int n_prior, int ratio_priority, int n_gpus, int mode,
                       double* lpp_sum_out) {
  const double* vec[6] = {Bf, depth_ratio, sd_Bf, sd_ratio, weight_Bf, weight_ratio};
  return baf_grid_host(S, seg_offsets, vec, ploidy, nP, cellularity, nC, avg_depth_ratio,
                       CNt_min, CNt_max, CNn, prior_CN, prior_value, n_prior, ratio_priority,
                       n_gpus, mode, lpp_sum_out);
}

size_t sqz_baf_grid_workspace_bytes(int64_t N_total, int64_t S, int64_t nP, int64_t nC_shard) {
  return baf_grid_workspace_bytes(N_total, S, nP, nC_shard);
}

size_t sqz_baf_grid_workspace_preferred_bytes(int64_t N_total, int64_t S, int64_t nP,
                                              int64_t nC, int64_t nC_shard,
                                              int n_copy_numbers, size_t cap_bytes) {
  return baf_grid_workspace_preferred_bytes(N_total, S, nP, nC, nC_shard, n_copy_numbers,
                                            cap_bytes);
}

size_t sqz_baf_grid_units_workspace_bytes(int64_t N_total, int64_t S, const double* ploidy,
                                          int64_t nP, int64_t nC, int CNt_min, int CNt_max,
                                          int CNn, int ratio_priority, int64_t u0, int64_t u1,
                                          size_t cap_bytes) {
  return baf_grid_units_workspace_bytes(N_total, S, ploidy, nP, nC, CNt_min, CNt_max, CNn,
                                        ratio_priority, u0, u1, cap_bytes);
}

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
                           int n_out) {
  if (sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return baf_grid_launch_units(S, d_seg_offsets, N_total, d_Bf, d_depth_ratio, d_sd_Bf,
                               d_sd_ratio, d_weight_Bf, d_weight_ratio, ploidy, nP, cellularity,
                               nC, u0, u1, d_avg_depth_ratio, CNt_min, CNt_max, CNn, prior_CN,
                               prior_value, n_prior, ratio_priority, mode, workspace,
                               workspace_bytes, (cudaStream_t)stream, d_outs, n_out);
}

int sqz_baf_grid_dev(int64_t S, const int64_t* d_seg_offsets, int64_t N_total,
                     const double* d_Bf, const double* d_depth_ratio, const double* d_sd_Bf,
                     const double* d_sd_ratio, const double* d_weight_Bf,
                     const double* d_weight_ratio, const double* ploidy, int64_t nP,
                     const double* cellularity, int64_t nC, int64_t ic0, int64_t ic1,
                     const double* d_avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                     const double* prior_CN, const double* prior_value, int n_prior,
                     int ratio_priority, int mode, void* workspace, size_t workspace_bytes,
                     void* stream, double* d_lpp_sum_out) {
  if (sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return baf_grid_launch(S, d_seg_offsets, N_total, d_Bf, d_depth_ratio, d_sd_Bf, d_sd_ratio,
                         d_weight_Bf, d_weight_ratio, ploidy, nP, cellularity, nC, ic0, ic1,
                         d_avg_depth_ratio, CNt_min, CNt_max, CNn, prior_CN, prior_value,
                         n_prior, ratio_priority, mode, workspace, workspace_bytes,
                         (cudaStream_t)stream, d_lpp_sum_out);
}

int sqz_baf_point(const double* Bf, const double* depth_ratio, const double* sd_Bf,
                  const double* sd_ratio, const double* weight_Bf, const double* weight_ratio,
                  int64_t N, double cellularity, double ploidy, double avg_depth_ratio,
                  int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                  const double* prior_value, int n_prior, int ratio_priority, double* out) {
  const double* vec[6] = {Bf, depth_ratio, sd_Bf, sd_ratio, weight_Bf, weight_ratio};
  return point_host(6, vec, N, out, [&](const double** dv, cudaStream_t st, double* d_out) {
    return baf_point_launch(dv[0], dv[1], dv[2], dv[3], dv[4], dv[5], N, cellularity, ploidy,
                            avg_depth_ratio, CNt_min, CNt_max, CNn, prior_CN, prior_value,
                            n_prior, ratio_priority, st, d_out);
  });
}

int sqz_baf_point_dev(const double* d_Bf, const double* d_depth_ratio, const double* d_sd_Bf,
                      const double* d_sd_ratio, const double* d_weight_Bf,
                      const double* d_weight_ratio, int64_t N, double cellularity,
                      double ploidy, double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                      const double* prior_CN, const double* prior_value, int n_prior,
                      int ratio_priority, void* stream, double* d_out) {
  if (sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return baf_point_launch(d_Bf, d_depth_ratio, d_sd_Bf, d_sd_ratio, d_weight_Bf,
                          d_weight_ratio, N, cellularity, ploidy, avg_depth_ratio, CNt_min,
                          CNt_max, CNn, prior_CN, prior_value, n_prior, ratio_priority,
                          (cudaStream_t)stream, d_out);
}

int sqz_mufreq_point(const double* mufreq, const double* depth_ratio,
                     const double* weight_mufreq, const double* weight_ratio, int64_t N,
                     double cellularity, double ploidy, double avg_depth_ratio, int CNt_min,
                     int CNt_max, int CNn, const double* prior_CN, const double* prior_value,
                     int n_prior, double* out) {
  const double* vec[4] = {mufreq, depth_ratio, weight_mufreq, weight_ratio};
  return point_host(4, vec, N, out, [&](const double** dv, cudaStream_t st, double* d_out) {
    return mufreq_point_launch(dv[0], dv[1], dv[2], dv[3], N, cellularity, ploidy,
                               avg_depth_ratio, CNt_min, CNt_max, CNn, prior_CN, prior_value,
                               n_prior, st, d_out);
  });
}

size_t sqz_mufreq_scalar_table_bytes(double weight_mufreq, double weight_ratio) {
  return mufreq_scalar_table_bytes(weight_mufreq, weight_ratio);
}

int sqz_mufreq_point_scalar_dev(const double* d_mufreq, const double* d_depth_ratio, int64_t N,
                                double weight_mufreq, double weight_ratio, double cellularity,
                                double ploidy, double avg_depth_ratio, int CNt_min, int CNt_max,
                                int CNn, const double* prior_CN, const double* prior_value,
                                int n_prior, void* d_table, size_t table_bytes, void* stream,
                                double* d_out) {
  if (sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return mufreq_point_scalar_launch(d_mufreq, d_depth_ratio, N, weight_mufreq, weight_ratio,
                                    cellularity, ploidy, avg_depth_ratio, CNt_min, CNt_max, CNn,
                                    prior_CN, prior_value, n_prior, d_table, table_bytes,
                                    (cudaStream_t)stream, d_out);
}

int sqz_mufreq_point_scalar(const double* mufreq, const double* depth_ratio, int64_t N,
                            double weight_mufreq, double weight_ratio, double cellularity,
                            double ploidy, double avg_depth_ratio, int CNt_min, int CNt_max,
                            int CNn, const double* prior_CN, const double* prior_value,
                            int n_prior, double* out) {
  const size_t tbytes = mufreq_scalar_table_bytes(weight_mufreq, weight_ratio);
  if (tbytes == 0)
    return fail(SQZ_ERR_ARG, "scalar weights must be non-negative integers <= 4096 (use "
                             "sqz_mufreq_point otherwise)");
  const double* vec[2] = {mufreq, depth_ratio};
  return point_host(2, vec, N, out, [&](const double** dv, cudaStream_t st, double* d_out) {
    int cur = 0;
    SQZ_CUDA(cudaGetDevice(&cur));
    DevCtx& c = g_ctx[cur];
    int rc = grow(&c.ws, &c.ws_cap, tbytes);
    if (rc) return rc;
    return mufreq_point_scalar_launch(dv[0], dv[1], N, weight_mufreq, weight_ratio, cellularity,
                                      ploidy, avg_depth_ratio, CNt_min, CNt_max, CNn, prior_CN,
                                      prior_value, n_prior, c.ws, c.ws_cap, st, d_out);
  });
}

int sqz_mufreq_point_dev(const double* d_mufreq, const double* d_depth_ratio,
                         const double* d_weight_mufreq, const double* d_weight_ratio,
                         int64_t N, double cellularity, double ploidy, double avg_depth_ratio,
                         int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                         const double* prior_value, int n_prior, void* stream, double* d_out) {
  if (sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return mufreq_point_launch(d_mufreq, d_depth_ratio, d_weight_mufreq, d_weight_ratio, N,
                             cellularity, ploidy, avg_depth_ratio, CNt_min, CNt_max, CNn,
                             prior_CN, prior_value, n_prior, (cudaStream_t)stream, d_out);
}

size_t sqz_mufreq_grid_workspace_bytes(int64_t N, int64_t nP, int64_t nC_shard) {
  return mufreq_grid_workspace_bytes(N, nP, nC_shard);
}

int sqz_mufreq_grid_dev(const double* d_mufreq, const double* d_depth_ratio,
                        const double* d_weight_mufreq, const double* d_weight_ratio, int64_t N,
                        const double* ploidy, int64_t nP, const double* cellularity,
                        int64_t nC, int64_t ic0, int64_t ic1, double avg_depth_ratio,
                        int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                        const double* prior_value, int n_prior, void* workspace,
                        size_t workspace_bytes, void* stream, double* d_lpp_sum_out) {
  if (sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return mufreq_grid_launch(d_mufreq, d_depth_ratio, d_weight_mufreq, d_weight_ratio, N,
                            ploidy, nP, cellularity, nC, ic0, ic1, avg_depth_ratio, CNt_min,
                            CNt_max, CNn, prior_CN, prior_value, n_prior, workspace,
                            workspace_bytes, (cudaStream_t)stream, d_lpp_sum_out);
}

int sqz_mufreq_grid(const double* mufreq, const double* depth_ratio,
                    const double* weight_mufreq, const double* weight_ratio, int64_t N,
                    const double* ploidy, int64_t nP, const double* cellularity, int64_t nC,
                    double avg_depth_ratio, int CNt_min, int CNt_max, int CNn,
                    const double* prior_CN, const double* prior_value, int n_prior,
                    int n_gpus, double* lpp_sum_out) {
  if (N < 0 || nP < 0 || nC < 0) return fail(SQZ_ERR_ARG, "negative size");
  const double* vec[4] = {mufreq, depth_ratio, weight_mufreq, weight_ratio};
  for (int k = 0; k < 4; ++k)
    if (N > 0 && !vec[k]) return fail(SQZ_ERR_ARG, "a per-position vector is null");
  if ((nP > 0 && !ploidy) || (nC > 0 && !cellularity) || (nP * nC > 0 && !lpp_sum_out))
    return fail(SQZ_ERR_ARG, "null pointer argument");
  if (nP == 0 || nC == 0) return SQZ_OK;
  int ndev = 0;
  int rc = usable_devices(n_gpus, &ndev);
  if (rc) return rc;
  if (ndev > nC) ndev = (int)nC;
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  std::vector<int64_t> c0(ndev + 1);
  for (int d = 0; d <= ndev; ++d) c0[d] = nC * d / ndev;
  for (int d = 0; d < ndev; ++d) {
    DevCtx* c;
    if ((rc = ctx_for(d, &c))) return rc;
    const int64_t nCs = c0[d + 1] - c0[d];
    const size_t vbytes = a256((size_t)(N > 0 ? N : 1) * 8);
    if ((rc = grow(&c->in, &c->in_cap, 4 * vbytes))) return rc;
    if ((rc = grow(&c->ws, &c->ws_cap, mufreq_grid_workspace_bytes(N, nP, nCs)))) return rc;
    if ((rc = grow(&c->out, &c->out_cap, (size_t)(nCs * nP) * 8))) return rc;
    char* p = (char*)c->in;
    const double* dv[4];
    for (int k = 0; k < 4; ++k) {
      dv[k] = (const double*)p;
      if (N > 0) SQZ_CUDA(cudaMemcpyAsync(p, vec[k], (size_t)N * 8, cudaMemcpyHostToDevice, c->st));
      p += vbytes;
    }
    rc = mufreq_grid_launch(dv[0], dv[1], dv[2], dv[3], N, ploidy, nP, cellularity, nC, c0[d],
                            c0[d + 1], avg_depth_ratio, CNt_min, CNt_max, CNn, prior_CN,
                            prior_value, n_prior, c->ws, c->ws_cap, c->st, (double*)c->out);
    if (rc) return rc;
  }
  for (int d = 0; d < ndev; ++d) {  // copy back only after all devices have been queued
    const int64_t nCs = c0[d + 1] - c0[d];
    SQZ_CUDA(cudaSetDevice(d));
    SQZ_CUDA(cudaMemcpyAsync(lpp_sum_out + c0[d] * nP, g_ctx[d].out, (size_t)(nCs * nP) * 8,
                             cudaMemcpyDeviceToHost, g_ctx[d].st));
  }
  for (int d = 0; d < ndev; ++d) {
    SQZ_CUDA(cudaSetDevice(d));
    SQZ_CUDA(cudaStreamSynchronize(g_ctx[d].st));
  }
  return SQZ_OK;
}

int sqz_fit_finish_dev(const double* d_lpp_sum, int64_t S, int64_t nP, int64_t nC,
                       double* d_lpp, double* d_row_max, double* d_col_max,
                       double* d_row_mass_above, double* d_col_mass_above,
                       double* d_log_sum_lik, void* stream) {
  if (sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return fit_finish_launch(d_lpp_sum, S, nP, nC, d_lpp, d_row_max, d_col_max, d_row_mass_above,
                           d_col_mass_above, d_log_sum_lik, (cudaStream_t)stream);
}

int sqz_fit_finish(const double* lpp_sum, int64_t S, int64_t nP, int64_t nC, double* lpp,
                   double* row_max, double* col_max, double* row_mass_above,
                   double* col_mass_above, double* log_sum_lik) {
  if (S < 0 || nP < 1 || nC < 1) return fail(SQZ_ERR_ARG, "need S >= 0, nP >= 1, nC >= 1");
  if (S > 0 && (!lpp_sum || !lpp || !row_max || !col_max || !row_mass_above ||
                !col_mass_above || !log_sum_lik))
    return fail(SQZ_ERR_ARG, "null pointer argument");
  if (S == 0) return SQZ_OK;
  int ndev = 0;
  int rc = usable_devices(1, &ndev);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  int cur = 0;
  SQZ_CUDA(cudaGetDevice(&cur));
  DevCtx* c;
  if ((rc = ctx_for(cur, &c))) return rc;
  const size_t G = (size_t)(nP * nC), nS = (size_t)S;
  const size_t in_b = a256(nS * G * 8);
  const size_t out_b = a256(nS * G * 8) + 2 * a256(nS * nP * 8) + 2 * a256(nS * nC * 8) +
                       a256(nS * 8);
  if ((rc = grow(&c->in, &c->in_cap, in_b))) return rc;
  if ((rc = grow(&c->out, &c->out_cap, out_b))) return rc;
  char* p = (char*)c->out;
  double* d_zn = (double*)p;
  p += a256(nS * G * 8);
  double* d_rm = (double*)p;
  p += a256(nS * nP * 8);
  double* d_ra = (double*)p;
  p += a256(nS * nP * 8);
  double* d_cm = (double*)p;
  p += a256(nS * nC * 8);
  double* d_ca = (double*)p;
  p += a256(nS * nC * 8);
  double* d_l = (double*)p;
  SQZ_CUDA(cudaMemcpyAsync(c->in, lpp_sum, nS * G * 8, cudaMemcpyHostToDevice, c->st));
  rc = fit_finish_launch((const double*)c->in, S, nP, nC, d_zn, d_rm, d_cm, d_ra, d_ca, d_l,
                         c->st);
  if (rc) return rc;
  SQZ_CUDA(cudaMemcpyAsync(lpp, d_zn, nS * G * 8, cudaMemcpyDeviceToHost, c->st));
  SQZ_CUDA(cudaMemcpyAsync(row_max, d_rm, nS * nP * 8, cudaMemcpyDeviceToHost, c->st));
  SQZ_CUDA(cudaMemcpyAsync(row_mass_above, d_ra, nS * nP * 8, cudaMemcpyDeviceToHost, c->st));
  SQZ_CUDA(cudaMemcpyAsync(col_max, d_cm, nS * nC * 8, cudaMemcpyDeviceToHost, c->st));
  SQZ_CUDA(cudaMemcpyAsync(col_mass_above, d_ca, nS * nC * 8, cudaMemcpyDeviceToHost, c->st));
  SQZ_CUDA(cudaMemcpyAsync(log_sum_lik, d_l, nS * 8, cudaMemcpyDeviceToHost, c->st));
  SQZ_CUDA(cudaStreamSynchronize(c->st));
  return SQZ_OK;
}

size_t sqz_segment_stats_workspace_bytes(int64_t n_rows, int64_t n_seg) {
  return segment_stats_workspace_bytes(n_rows, n_seg);
}

int sqz_segment_stats_dev(int64_t n_chr, const int64_t* d_row_off, const int64_t* d_seg_off,
                          int64_t n_rows, int64_t n_seg, const double* d_position,
                          const double* d_adjusted_ratio, const double* d_depth_normal,
                          const double* d_good_reads, const double* d_Af, const double* d_Bf,
                          const unsigned char* d_is_het, const double* d_breaks_vect,
                          double min_reads_baf, int weighted_mean, void* workspace,
                          size_t workspace_bytes, void* stream, double* d_out) {
  if (sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return segment_stats_launch(n_chr, d_row_off, d_seg_off, n_rows, n_seg, d_position,
                              d_adjusted_ratio,
                              d_depth_normal, d_good_reads, d_Af, d_Bf, d_is_het,
                              d_breaks_vect, min_reads_baf, weighted_mean, workspace,
                              workspace_bytes, (cudaStream_t)stream, d_out);
}

int sqz_segment_stats(int64_t n_chr, const int64_t* row_off, const int64_t* seg_off,
                      const double* position, const double* adjusted_ratio,
                      const double* depth_normal, const double* good_reads, const double* Af,
                      const double* Bf, const unsigned char* is_het, const double* breaks_vect,
                      double min_reads_baf, int weighted_mean, double* out) {
  if (n_chr < 0) return fail(SQZ_ERR_ARG, "negative n_chr");
  if (n_chr == 0) return SQZ_OK;
  if (!row_off || !seg_off) return fail(SQZ_ERR_ARG, "offset arrays are null");
  const int64_t N = row_off[n_chr], S = seg_off[n_chr];
  if (row_off[0] != 0 || seg_off[0] != 0) return fail(SQZ_ERR_ARG, "offsets must start at 0");
  for (int64_t c = 0; c < n_chr; ++c) {
    if (row_off[c + 1] < row_off[c] || seg_off[c + 1] < seg_off[c])
      return fail(SQZ_ERR_ARG, "offsets must be non-decreasing");
    for (int64_t i = row_off[c] + 1; i < row_off[c + 1]; ++i)
      if (position && position[i] < position[i - 1])
        return fail(SQZ_ERR_ARG, "positions must be sorted inside each chromosome");
    for (int64_t k = 2 * seg_off[c] + 1; k < 2 * seg_off[c + 1]; ++k)
      if (breaks_vect && !(breaks_vect[k] > breaks_vect[k - 1]))
        return fail(SQZ_ERR_ARG, "breaks must increase inside each chromosome (unique.breaks)");
  }
  if (S == 0) return SQZ_OK;
  if ((N > 0 && (!position || !adjusted_ratio || !depth_normal || !good_reads || !Af || !Bf ||
                 !is_het)) || !breaks_vect || !out)
    return fail(SQZ_ERR_ARG, "null pointer argument");
  int ndev = 0;
  int rc = usable_devices(1, &ndev);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  int cur = 0;
  SQZ_CUDA(cudaGetDevice(&cur));
  DevCtx* c;
  if ((rc = ctx_for(cur, &c))) return rc;
  const size_t vb = a256((size_t)(N > 0 ? N : 1) * 8), hb = a256((size_t)(N > 0 ? N : 1));
  const size_t ob = a256((size_t)(n_chr + 1) * 8), bb = a256((size_t)S * 16);
  if ((rc = grow(&c->in, &c->in_cap, 6 * vb + hb + 2 * ob + bb))) return rc;
  if ((rc = grow(&c->ws, &c->ws_cap, segment_stats_workspace_bytes(N, S)))) return rc;
  if ((rc = grow(&c->out, &c->out_cap, (size_t)S * 48))) return rc;
  char* p = (char*)c->in;
  const double* src[6] = {position, adjusted_ratio, depth_normal, good_reads, Af, Bf};
  const double* dv[6];
  for (int k = 0; k < 6; ++k, p += vb) {
    dv[k] = (const double*)p;
    if (N > 0) SQZ_CUDA(cudaMemcpyAsync(p, src[k], (size_t)N * 8, cudaMemcpyHostToDevice, c->st));
  }
  const unsigned char* d_het = (const unsigned char*)p;
  if (N > 0) SQZ_CUDA(cudaMemcpyAsync(p, is_het, (size_t)N, cudaMemcpyHostToDevice, c->st));
  p += hb;
  const int64_t* d_ro = (const int64_t*)p;
  SQZ_CUDA(cudaMemcpyAsync(p, row_off, (size_t)(n_chr + 1) * 8, cudaMemcpyHostToDevice, c->st));
  p += ob;
  const int64_t* d_so = (const int64_t*)p;
  SQZ_CUDA(cudaMemcpyAsync(p, seg_off, (size_t)(n_chr + 1) * 8, cudaMemcpyHostToDevice, c->st));
  p += ob;
  const double* d_brk = (const double*)p;
  SQZ_CUDA(cudaMemcpyAsync(p, breaks_vect, (size_t)S * 16, cudaMemcpyHostToDevice, c->st));
  rc = segment_stats_launch(n_chr, d_ro, d_so, N, S, dv[0], dv[1], dv[2], dv[3], dv[4], dv[5],
                            d_het, d_brk, min_reads_baf, weighted_mean, c->ws, c->ws_cap, c->st,
                            (double*)c->out);
  if (rc) {
    cudaStreamSynchronize(c->st);
    return rc;
  }
  SQZ_CUDA(cudaMemcpyAsync(out, c->out, (size_t)S * 48, cudaMemcpyDeviceToHost, c->st));
  SQZ_CUDA(cudaStreamSynchronize(c->st));
  return SQZ_OK;
}

int sqz_model_points(int kind, double cellularity, double ploidy, double avg_depth_ratio,
                     int CNt_min, int CNt_max, int CNn, const double* prior_CN,
                     const double* prior_value, int n_prior, int* T_out, double* out) {
  if (out && sqz_device_count() < 1)
    return fail(SQZ_ERR_NOGPU, "no CUDA device available: libsequenza_b200 has no CPU path");
  return model_points_launch(kind, cellularity, ploidy, avg_depth_ratio, CNt_min, CNt_max,
                             CNn, prior_CN, prior_value, n_prior, T_out, out);
}

int sqz_baf_grid_plan(int64_t N_total, int64_t S, const double* ploidy, int64_t nP, int64_t nC,
                      int CNt_min, int CNt_max, int CNn, int ratio_priority, int64_t* out8) {
  if (!out8) return fail(SQZ_ERR_ARG, "out8 is null");
  return baf_grid_plan(N_total, S, ploidy, nP, nC, CNt_min, CNt_max, CNn, ratio_priority, out8);
}

int sqz_profile_enable(int on) { return profile_enable(on); }
int sqz_profile_collect(double* ms3, int* n_calls) { return profile_collect(ms3, n_calls); }
int sqz_profile_counters(int* path, unsigned long long* n_eval, long long* launches) {
  return profile_counters(path, n_eval, launches);
}

int sqz_fp64_peak(int device, int repeats, double* tflops_out) {
  if (!tflops_out) return fail(SQZ_ERR_ARG, "tflops_out is null");
  int ndev = 0;
  int rc = usable_devices(1, &ndev);
  if (rc) return rc;
  DeviceGuard guard;
  SQZ_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  SQZ_CUDA(cudaGetDeviceProperties(&prop, device));
  double* d = nullptr;
  SQZ_CUDA(cudaMalloc(&d, 8));
  cudaEvent_t e0, e1;
  SQZ_CUDA(cudaEventCreate(&e0));
  SQZ_CUDA(cudaEventCreate(&e1));
  const int iters = 1 << 15, blocks = prop.multiProcessorCount * 8, threads = 256;
  if (repeats < 1) repeats = 1;
  double best = 0;
  for (int r = 0; r < repeats + 1; ++r) {
    SQZ_CUDA(cudaEventRecord(e0));
    fp64_peak_kernel<<<blocks, threads>>>(d, iters, 1.0);
    SQZ_CUDA(cudaEventRecord(e1));
    SQZ_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    SQZ_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    double tf = 2.0 * 8 * (double)iters * blocks * threads / (ms * 1e-3) / 1e12;
    if (r > 0 && tf > best) best = tf;  // first launch is warm-up
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *tflops_out = best;
  return SQZ_OK;
}

}  // extern "C"
