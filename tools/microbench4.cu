// microbench4.cu -- issue rate of the packed fp32 instructions (FFMA2) and of the 3-input
// FMNMX3 against plain FFMA on sm_100a: SMSP cycles per warp instruction with 8 independent
// chains per thread.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb4 microbench4.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITER = 4096, CH = 8;

template <int MODE>
__global__ void kern(float* out, float seed) {
  float a[CH], b[CH];
#pragma unroll
  for (int i = 0; i < CH; ++i) {
    a[i] = seed + i + threadIdx.x;
    b[i] = seed * 0.5f + i;
  }
  const float m = 1.0000001f, c = 1e-9f;
#pragma unroll 1
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (MODE == 0) {  // 2 x FFMA.RM
        a[i] = __fmaf_rd(a[i], m, c);
        b[i] = __fmaf_rd(b[i], m, c);
      } else if (MODE == 1) {  // 1 x FFMA2.RM on the pair
        unsigned long long v, mm, cc;
        asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[i]), "f"(b[i]));
        asm("mov.b64 %0, {%1, %1};" : "=l"(mm) : "f"(m));
        asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
        asm("fma.rm.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(mm), "l"(cc));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(b[i]) : "l"(v));
      } else if (MODE == 2) {  // 2 x FMNMX + 2 x FADD (t = max(|x| - s, 0) the plain way)
        a[i] = fmaxf(__fsub_rd(fabsf(a[i]), c), 0.f);
        b[i] = fmaxf(__fsub_rd(fabsf(b[i]), c), 0.f);
      } else if (MODE == 3) {  // 2 x FMNMX3
        float r, s;
        asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a[i]), "f"(b[i]), "f"(c));
        asm("max.f32 %0, %1, %2, %3;" : "=f"(s) : "f"(b[i]), "f"(r), "f"(m));
        a[i] = r;
        b[i] = s;
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < CH; ++i) s += a[i] + b[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, int inst_per_pair, int sms, double ghz) {
  float* d;
  cudaMalloc(&d, 148 * 1024 * sizeof(float) * 4);
  for (int thr : {128, 256, 512}) {
    kern<MODE><<<sms, thr>>>(d, 1.f);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<MODE><<<sms, thr>>>(d, 1.f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double warps_per_smsp = thr / 32 / 4.0;
    const double inst = (double)ITER * CH * inst_per_pair * warps_per_smsp;
    printf("  %-34s warps/SMSP %.0f: %.3f ms, %.2f SMSP cycles per warp instruction\n", name,
           warps_per_smsp, ms, ms * 1e-3 * ghz * 1e9 / inst);
  }
  cudaFree(d);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int khz;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ghz = khz * 1e-6;
  printf("SMs %d, max clock %.3f GHz\n(D) packed fp32 and 3-input min/max, 8 chains per thread\n",
         p.multiProcessorCount, ghz);
  run<0>("2 x FFMA.RM", 2, p.multiProcessorCount, ghz);
  run<1>("1 x FFMA2.RM (two floats)", 1, p.multiProcessorCount, ghz);
  run<2>("2 x (FADD.RM + FMNMX)", 4, p.multiProcessorCount, ghz);
  run<3>("2 x FMNMX3", 2, p.multiProcessorCount, ghz);
  return 0;
}
