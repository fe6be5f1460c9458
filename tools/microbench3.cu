// Third calibration: is occupancy or instruction-level parallelism the lever on the fp64 pipe?
//  (A) DFMA throughput for 1, 2, 4, 8 warps per SM sub-partition x 1..8 independent chains
//      per thread (a dependent DFMA chain gives the pipe latency);
//  (B) the column-per-lane main loop (x = fma(nld_j, k, b); q = fma(x, x, w); P_j *= q for
//      32 ploidies, registers only) at one warp per CTA, 4 and 8 CTAs per SM, with a serial
//      "analysis" chain of CH dependent fp64 operations in front of it, either in the same
//      iteration (as the round-1 kernel) or computed one iteration ahead (software pipeline);
//  (C) DFMA and FSEL/IMNMX issued together: do the fp64 and the ALU pipe overlap?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o microbench3 microbench3.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void k_dfma(double* out, int iters, double m, double c) {
  double a[ILP];
#pragma unroll
  for (int i = 0; i < ILP; i++) a[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = fma(a[i], m, c);
  }
  double r = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) r += a[i];
  if (r == 1.2345) out[0] = r;
}

// (B) CH = length of the dependent chain, PIPE = chain result used one iteration later
template <int CH, bool PIPE>
__global__ void __launch_bounds__(32, 8) k_colwarp(double* out, int iters, double k0, double b0) {
  double nld[32], P[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    nld[j] = 1e-3 * (j - 16) + 1e-6 * threadIdx.x;
    P[j] = 1.0;
  }
  double seed = 1.0 + 1e-9 * threadIdx.x;
  double k = k0, b = b0, w = 1.0;  // parameters of the current segment
  for (int it = 0; it < iters; ++it) {
    // "analysis": a serial chain that yields the next segment's (k, b, w)
    double t = seed;
#pragma unroll
    for (int i = 0; i < CH; ++i) t = fma(t, 0.999999, 1e-7);
    seed = t;
    const double kn = k0 + 1e-9 * t, bn = b0 - 1e-9 * t, wn = 1.0 + 1e-12 * t;
    if (!PIPE) {
      k = kn;
      b = bn;
      w = wn;
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const double x = fma(nld[j], k, b);
      P[j] *= fma(x, x, w);
    }
    if (PIPE) {
      k = kn;
      b = bn;
      w = wn;
    }
    if ((it & 63) == 63) {
#pragma unroll
      for (int j = 0; j < 32; ++j) P[j] = 1.0 + 1e-30 * P[j];
    }
  }
  double r = 0;
#pragma unroll
  for (int j = 0; j < 32; ++j) r += P[j];
  if (r == 1.2345) out[0] = r;
}

// (C) MODE 0: 32 DFMA; 1: 32 FSEL pairs (fp64 min on integer compare result); 2: both
template <int MODE>
__global__ void k_overlap(double* out, int iters, double m, double c) {
  double a[16], q[16];
#pragma unroll
  for (int i = 0; i < 16; i++) {
    a[i] = threadIdx.x + i;
    q[i] = 1e300 - i;
  }
  int pred = threadIdx.x;
  for (int it = 0; it < iters; ++it) {
    pred = pred * 5 + 1;
#pragma unroll
    for (int i = 0; i < 16; i++) {
      if (MODE != 1) a[i] = fma(a[i], m, c);
      if (MODE != 0) {  // two FSEL on a predicate that costs no fp64 instruction
        const bool p = (pred >> i) & 1;
        q[i] = p ? q[(i + 1) & 15] : q[i];
      }
    }
  }
  double r = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) r += a[i] + q[i];
  if (r == 1.2345) out[0] = r;
}

template <class F>
float timeit(F f) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  f();
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  f();
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  return ms;
}

int main() {
  double* d;
  cudaMalloc(&d, 64);
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double ghz = clk * 1e-6;
  printf("SMs %d, max clock %.3f GHz\n", sms, ghz);
  const int iters = 1 << 15;
  printf("(A) DFMA: SMSP cycles per warp instruction (2.0 = pipe peak)\n");
#define RUN_A(ILP)                                                                          \
  for (int wps : {1, 2, 4, 8}) {                                                            \
    const int threads = 128, blocks = sms * wps; /* wps CTAs of 4 warps = wps warps/SMSP */ \
    float ms = timeit([&] { k_dfma<ILP><<<blocks, threads>>>(d, iters, 0.999, 1e-9); });    \
    double inst = (double)ILP * iters * wps;     /* warp instructions per SMSP */           \
    printf("  chains/thread %d, warps/SMSP %d: %.3f ms, %.2f cycles per DFMA\n", ILP, wps,  \
           ms, ms * 1e-3 * ghz * 1e9 / inst);                                               \
  }
  RUN_A(1) RUN_A(2) RUN_A(4) RUN_A(8)
  printf("(B) column-per-lane main loop, one warp per CTA: SMSP cycles per (segment, tile); "
         "96 fp64 instructions = 192 cycles at pipe peak\n");
#define RUN_B(CH, PIPE)                                                                       \
  for (int cps : {4, 8}) {                                                                    \
    float ms = timeit([&] { k_colwarp<CH, PIPE><<<sms * cps, 32>>>(d, 4096, 1.1, 0.3); });    \
    printf("  chain %3d %s, %d warps/SMSP: %.3f ms, %.0f SMSP cycles per (iteration, warp)\n", \
           CH, PIPE ? "pipelined " : "same iter.", cps / 4, ms,                               \
           ms * 1e-3 * ghz * 1e9 / (4096.0 * cps / 4));                                       \
  }
  RUN_B(0, false) RUN_B(16, false) RUN_B(16, true) RUN_B(40, false) RUN_B(40, true)
  RUN_B(80, false) RUN_B(80, true)
  printf("(C) fp64 pipe and ALU pipe together, 4 warps/SMSP, 16 per iteration each\n");
  for (int mode = 0; mode < 3; ++mode) {
    float ms = 0;
    if (mode == 0) ms = timeit([&] { k_overlap<0><<<sms * 4, 128>>>(d, iters, 0.999, 1e-9); });
    if (mode == 1) ms = timeit([&] { k_overlap<1><<<sms * 4, 128>>>(d, iters, 0.999, 1e-9); });
    if (mode == 2) ms = timeit([&] { k_overlap<2><<<sms * 4, 128>>>(d, iters, 0.999, 1e-9); });
    printf("  %s: %.3f ms, %.2f SMSP cycles per loop element\n",
           mode == 0 ? "DFMA only      " : mode == 1 ? "2 FSEL only    " : "DFMA + 2 FSEL  ", ms,
           ms * 1e-3 * ghz * 1e9 / (16.0 * iters * 4));
  }
  return 0;
}
