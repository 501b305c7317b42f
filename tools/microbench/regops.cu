// Microbenchmark: issue cost of scalar FFMA on sm_100a as a function of how many DISTINCT register-file
// operands it reads and whether consecutive instructions share one (operand reuse cache).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o regops regops.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITER = 4096;

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float a, float b) {
  float x[16], y[4], z[4];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x * 1e-3f + i;
#pragma unroll
  for (int i = 0; i < 4; ++i) { y[i] = 1.0f + threadIdx.x * 1e-6f * (i + 1); z[i] = threadIdx.x * 1e-7f * (i + 1); }
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) x[i] = fmaf(x[i], a, b);                           // 1 register + 2 constants
      if (MODE == 1) x[i] = fmaf(x[i], y[0], b);                        // 2 registers (one shared by all) + constant
      if (MODE == 2) x[i] = fmaf(x[i], y[0], z[0]);                     // 3 registers, two shared by all instructions
      if (MODE == 3) x[i] = fmaf(x[i], y[i & 3], z[0]);                 // 3 registers, one shared
      if (MODE == 4) x[i] = fmaf(x[i], x[(i + 5) & 15], x[(i + 11) & 15]);  // 3 registers, none shared
      if (MODE == 5) x[i] = fmaf(x[i], x[(i + 5) & 15], b);             // 2 registers + constant, none shared
      if (MODE == 6) x[i] = fmaf(y[i & 3], z[(i >> 2) & 3], x[i]);      // accumulate form, 3 registers, b-slot shared by 4 in a row
      if (MODE == 7) x[i] = fmaf(x[i], x[i], x[(i + 5) & 15]);          // 2 distinct registers in 3 slots
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + y[1] + y[2] + y[3] + z[1] + z[2] + z[3];
}

template <int MODE>
void run(const char* name, float* d) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = 148 * 8;
  k<MODE><<<grid, 256>>>(d, 1.0001f, 0.5f);
  cudaEventRecord(e0);
  k<MODE><<<grid, 256>>>(d, 1.0001f, 0.5f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double fma = (double)grid * 256 * ITER * 16;
  int dev = 0, khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  // warp-instructions per cycle per SM sub-partition at the nominal boost clock
  double ipc = fma / 32.0 / (ms * 1e-3) / (khz * 1e3) / (148 * 4);
  printf("%-62s %7.3f ms %6.2f TFMA/s  %.2f FFMA/clk/SMSP (at %d MHz)\n", name, ms, fma / ms / 1e9, ipc, khz / 1000);
}

int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  run<0>("x = fma(x, const, const)", d);
  run<1>("x = fma(x, y0, const)        2 regs, y0 shared", d);
  run<2>("x = fma(x, y0, z0)           3 regs, y0 z0 shared", d);
  run<3>("x = fma(x, y[i&3], z0)       3 regs, z0 shared", d);
  run<4>("x = fma(x, x', x'')          3 regs, none shared", d);
  run<5>("x = fma(x, x', const)        2 regs, none shared", d);
  run<6>("x = fma(y[i&3], z[i>>2], x)  3 regs, z shared by runs of 4", d);
  run<7>("x = fma(x, x, x')            2 distinct regs in 3 slots", d);
  return 0;
}
