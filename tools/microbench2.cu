// Second calibration round: cost of the fp64 compare+select and of integer-compare minima.
#include <cstdio>
#include <cuda_runtime.h>
template<int PT, int MODE>
__global__ void k(double* out, int iters, double kr, double er, double w0) {
  __shared__ double l[64*32];
  for (int i=threadIdx.x;i<64*32;i+=blockDim.x) l[i]=1.0+1e-3*i;
  __syncthreads();
  double qm[PT];
  for (int j=0;j<PT;j++) qm[j]=1e300;
  double krr = kr + 1e-9*threadIdx.x, err = er, w = w0;
  for (int it=0; it<iters; ++it) {
    const double* lp = l + (it & 63)*32;
#pragma unroll
    for (int j=0;j<PT;j++) {
      double x=fma(lp[j],krr,err);
      if (MODE==0) {            // 3 fp64, no min: DFMA DFMA DMUL, accumulate by DADD-free trick
        double t=fma(x,x,1.0); qm[j]=t*w + 0.0*qm[j];   // (compiler keeps DMUL + DFMA)
      } else if (MODE==1) {     // full mix with fp64 compare
        double t=fma(x,x,1.0); double q=t*w; if (q<qm[j]) qm[j]=q;
      } else if (MODE==2) {     // full mix with 64-bit unsigned integer compare (q > 0)
        double t=fma(x,x,1.0); double q=t*w;
        unsigned long long a=__double_as_longlong(q), b=__double_as_longlong(qm[j]);
        qm[j]=__longlong_as_double(a<b?a:b);
      } else if (MODE==3) {     // only compare+select on precomputed x (1 fp64 DFMA + DSETP)
        if (x<qm[j]) qm[j]=x;
      } else if (MODE==4) {     // DFMA + integer min
        unsigned long long a=__double_as_longlong(x), b=__double_as_longlong(qm[j]);
        qm[j]=__longlong_as_double(a<b?a:b);
      } else if (MODE==5) {     // predicated recompute: DFMA DFMA DMUL DSETP @P DMUL (no FSEL)
        double t=fma(x,x,1.0); double q=t*w; if (q<qm[j]) qm[j]=t*w*1.0000000001;
      }
    }
    w += 1e-7; err -= 1e-9;
  }
  double r=0; for(int j=0;j<PT;j++) r+=qm[j];
  if (r==1.2345) out[0]=r;
}
template<class F> float timeit(F f) {
  cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms,a,b); return ms;
}
int main() {
  double* d; cudaMalloc(&d, 64);
  const int iters=20000, sms=148;
  auto rep=[&](const char* name, float ms, int PT, int blocks, int threads){
    double elts = (double)PT*iters*blocks*threads/32.0;
    double cyc = ms*1e-3*1.965e9*sms*4.0/elts;
    printf("%-44s bps %d thr %d: %.3f ms  SMSP-cycles per (warp,element) %.2f\n", name, blocks/sms, threads, ms, cyc);
  };
  for (int bps : {1,2}) for (int th : {256}) {
    int blocks=sms*bps;
    rep("M0 DFMA DFMA DMUL(+DFMA acc)", timeit([&]{k<16,0><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}),16,blocks,th);
    rep("M1 DFMA DFMA DMUL DSETP 2FSEL  PT16", timeit([&]{k<16,1><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}),16,blocks,th);
    rep("M1 ... PT32", timeit([&]{k<32,1><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}),32,blocks,th);
    rep("M2 DFMA DFMA DMUL + u64 int min  PT16", timeit([&]{k<16,2><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}),16,blocks,th);
    rep("M2 ... PT32", timeit([&]{k<32,2><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}),32,blocks,th);
    rep("M3 DFMA DSETP 2FSEL", timeit([&]{k<16,3><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}),16,blocks,th);
    rep("M4 DFMA + u64 int min", timeit([&]{k<16,4><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}),16,blocks,th);
    rep("M5 predicated recompute", timeit([&]{k<16,5><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}),16,blocks,th);
  }
  return 0;
}
