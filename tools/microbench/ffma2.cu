// Microbenchmark: issue-slot cost of packed FFMA2 vs scalar FFMA on sm_100a, alone and mixed
// with alu-pipe work (LOP3).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITER = 4096;

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float a, float b, unsigned m) {
  float x[16];
  unsigned u[8];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x * 1e-3f + i;
#pragma unroll
  for (int i = 0; i < 8; ++i) u[i] = threadIdx.x + i;
  for (int it = 0; it < ITER; ++it) {
    if (MODE >= 4) {
    } else if (MODE == 0 || MODE == 2) {
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], a, b);
    } else {
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        float2 v = __ffma2_rn(make_float2(x[i], x[i + 1]), make_float2(a, a), make_float2(b, b));
        x[i] = v.x; x[i + 1] = v.y;
      }
    }
    if (MODE == 4) {  // 3-register FFMA: every operand a distinct live register
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], x[(i + 5) & 15], x[(i + 11) & 15]);
    }
    if (MODE == 5) {  // 3-register FFMA2
      float2 v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        v[i] = __ffma2_rn(make_float2(x[2 * i], x[2 * i + 1]), make_float2(x[(2 * i + 4) & 15], x[(2 * i + 5) & 15]),
                          make_float2(x[(2 * i + 10) & 15], x[(2 * i + 11) & 15]));
#pragma unroll
      for (int i = 0; i < 8; ++i) { x[2 * i] = v[i].x; x[2 * i + 1] = v[i].y; }
    }
    if (MODE == 6) {  // FMUL 2-register + FADD
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] = x[i] * x[(i + 5) & 15];
    }
    if (MODE >= 2 && MODE < 4) {
#pragma unroll
      for (int i = 0; i < 8; ++i) u[i] = (u[i] ^ m) & (u[(i + 1) & 7] | m);
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
  unsigned w = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) w ^= u[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)w;
}

template <int MODE>
void run(const char* name, float* d) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = 148 * 8;
  k<MODE><<<grid, 256>>>(d, 1.0001f, 0.5f, 0x5a5a5a5au);
  cudaEventRecord(e0);
  k<MODE><<<grid, 256>>>(d, 1.0001f, 0.5f, 0x5a5a5a5au);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double fma = (double)grid * 256 * ITER * 16;
  printf("%-28s %8.3f ms  %7.2f TFMA/s (%.1f TFLOP/s)\n", name, ms, fma / ms / 1e9, 2 * fma / ms / 1e9);
}

int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  run<0>("16 FFMA", d);
  run<1>("8 FFMA2", d);
  run<2>("16 FFMA + 8 LOP3", d);
  run<3>("8 FFMA2 + 8 LOP3", d);
  run<4>("16 FFMA 3-reg", d);
  run<5>("8 FFMA2 3-reg", d);
  run<6>("16 FMUL 2-reg", d);
  return 0;
}
