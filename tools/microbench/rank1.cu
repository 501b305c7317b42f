// Microbenchmark: rank-1 update A[7][6] -= v s^T (the inner operation of the Householder solve) as 42 scalar
// FFMAs vs 21 packed FFMA2 (+7 register-pair packs), to see whether FFMA2 lifts the 3-register-operand issue limit.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int ITER = 2048;

template <int MODE>
__global__ void __launch_bounds__(128) k(float* out, float eps) {
  float A[7][6], v[7], s[6];
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    v[r] = 1e-3f * (threadIdx.x + r);
#pragma unroll
    for (int j = 0; j < 6; ++j) A[r][j] = 0.01f * (r * 6 + j) + threadIdx.x * 1e-4f;
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) s[j] = eps * (j + 1);
  for (int it = 0; it < ITER; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int r = 0; r < 7; ++r)
#pragma unroll
        for (int j = 0; j < 6; ++j) A[r][j] = fmaf(-s[j], v[r], A[r][j]);
    } else {
#pragma unroll
      for (int r = 0; r < 7; ++r) {
        const float2 vv = make_float2(v[r], v[r]);
#pragma unroll
        for (int j = 0; j < 6; j += 2) {
          const float2 a = __ffma2_rn(make_float2(-s[j], -s[j + 1]), vv, make_float2(A[r][j], A[r][j + 1]));
          A[r][j] = a.x; A[r][j + 1] = a.y;
        }
      }
    }
    // perturb v and s a little from A so nothing is loop-invariant (7 + 6 cheap ops)
#pragma unroll
    for (int r = 0; r < 7; ++r) v[r] = fmaf(A[r][0], eps, v[r]);
#pragma unroll
    for (int j = 0; j < 6; ++j) s[j] = fmaf(A[0][j], eps, s[j]);
  }
  float acc = 0;
#pragma unroll
  for (int r = 0; r < 7; ++r)
#pragma unroll
    for (int j = 0; j < 6; ++j) acc += A[r][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE>
void run(const char* name, float* d) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = 148 * 6;
  k<MODE><<<grid, 128>>>(d, 1e-6f);
  cudaEventRecord(e0);
  k<MODE><<<grid, 128>>>(d, 1e-6f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double warps = (double)grid * 4, cyc = ms * 1e-3 * 1.965e9;
  // 6 CTAs x 4 warps = 24 warps per SM = 6 per SMSP
  printf("%-34s %7.3f ms   %.1f cycles per rank-1 update per warp-slot (6 warps/SMSP)\n", name, ms, cyc / ITER / 6);
  (void)warps;
}

int main() {
  float* d; cudaMalloc(&d, 148 * 6 * 128 * 4);
  run<0>("42 FFMA (+13)", d);
  run<1>("21 FFMA2 + 7 packs (+13)", d);
  return 0;
}
