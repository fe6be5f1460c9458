// Microbenchmarks that calibrate the fp64 pipe on B200 for the instruction mix of
// baf_grid_fast_kernel.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false
#include <cstdio>
#include <cuda_runtime.h>

#define CHK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("%s: %s\n",#x,cudaGetErrorString(e)); return 1;}}while(0)

// (1) pure DFMA, all-register operands, ILP chains
template<int ILP>
__global__ void k_dfma(double* out, int iters, double m, double c) {
  double a[ILP];
  for (int i=0;i<ILP;i++) a[i]=threadIdx.x+i;
  for (int it=0; it<iters; ++it) {
#pragma unroll
    for (int i=0;i<ILP;i++) a[i]=fma(a[i],m,c);
  }
  double r=0; for(int i=0;i<ILP;i++) r+=a[i];
  if (r==1.2345) out[0]=r;
}
// (2) the j-loop mix: x=fma(l,kr,er); t=fma(x,x,1); q=t*w; if(q<qm) qm=q     (l from smem broadcast)
template<int PT, bool USE_SMEM>
__global__ void k_mix(double* out, int iters, double kr, double er, double w0) {
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
      double lv = USE_SMEM ? lp[j] : (double)(1.0+j);
      double x=fma(lv,krr,err);
      double t=fma(x,x,1.0);
      double q=t*w;
      if (q<qm[j]) qm[j]=q;
    }
    w += 1e-7; err -= 1e-9;
  }
  double r=0; for(int j=0;j<PT;j++) r+=qm[j];
  if (r==1.2345) out[0]=r;
}
// (3) variant: 5 fp64 ops, no FSEL: predicated recompute
template<int PT>
__global__ void k_mix_pred(double* out, int iters, double kr, double er, double w0) {
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
      double t=fma(x,x,w);     // 3-op variant: q = x'^2 + w  (x' pre-scaled)
      if (t<qm[j]) qm[j]=t;
    }
    w += 1e-7; err -= 1e-9;
  }
  double r=0; for(int j=0;j<PT;j++) r+=qm[j];
  if (r==1.2345) out[0]=r;
}
// (4) DSETP+FSEL only
template<int PT>
__global__ void k_min(double* out, int iters, double w0) {
  double qm[PT], v[PT];
  for (int j=0;j<PT;j++) {qm[j]=1e300; v[j]=threadIdx.x+j;}
  for (int it=0; it<iters; ++it) {
#pragma unroll
    for (int j=0;j<PT;j++) { if (v[(j+it)%PT] < qm[j]) qm[j]=v[(j+it)%PT]; }
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
  double* d; CHK(cudaMalloc(&d, 64));
  cudaDeviceProp p; cudaGetDeviceProperties(&p,0);
  int sms=p.multiProcessorCount; int clk=0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("SMs %d clock %d kHz\n", sms, clk);
  const int iters=20000;
  auto rep=[&](const char* name, float ms, double fp64_instr_per_thread, int blocks, int threads){
    double inst = fp64_instr_per_thread*blocks*threads/32.0; // warp-instr
    double per_smsp_cycle = inst/(sms*4.0)/(ms*1e-3*1.965e9);
    printf("%-34s blocks/SM %d thr %d: %.3f ms  fp64 warp-inst/SMSP/cycle(at 1.965GHz) %.3f\n", name, blocks/sms, threads, ms, per_smsp_cycle);
  };
  for (int bps : {1,2,4}) for (int th : {128,256}) {
    int blocks=sms*bps;
    rep("dfma ILP8", timeit([&]{k_dfma<8><<<blocks,th>>>(d,iters,0.999,1e-9);}), 8.0*iters, blocks, th);
  }
  for (int bps : {1,2}) for (int th : {128,256}) {
    int blocks=sms*bps;
    rep("mix PT16 smem (4 fp64/elt)", timeit([&]{k_mix<16,true><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}), 64.0*iters, blocks, th);
    rep("mix PT16 nosmem", timeit([&]{k_mix<16,false><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}), 64.0*iters, blocks, th);
    rep("mix PT32 smem", timeit([&]{k_mix<32,true><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}), 128.0*iters, blocks, th);
    rep("mix3 PT16 (3 fp64/elt)", timeit([&]{k_mix_pred<16><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}), 48.0*iters, blocks, th);
    rep("mix3 PT32 (3 fp64/elt)", timeit([&]{k_mix_pred<32><<<blocks,th>>>(d,iters,1.1,0.3,1.5);}), 96.0*iters, blocks, th);
    rep("min only PT16 (1 fp64/elt)", timeit([&]{k_min<16><<<blocks,th>>>(d,iters,1.5);}), 16.0*iters, blocks, th);
  }
  return 0;
}
