/*
 * Plain C host for libmsik.so: no Python, no torch.  Shows the drop-in boundary of include/msik.h the way a
 * native binding would use it:
 *
 *   chain description -> msik_chain_create -> device buffers in the PhysX layout -> msik_delta_step
 *   (path A: q + pinv(J(q)) a, mani_skill/agents/controllers/utils/kinematics.py:156-172) -> joint targets
 *
 * and checks the result against the fp64 oracle's answers stored in the fixture
 * (tests/golden/c_host_panda.bin, written by tools/gen_c_host_fixture.py).
 *
 *   gcc -std=c99 -Iinclude -I/usr/local/cuda/include examples/c_host/msik_host_example.c \
 *       -Lmaniskill_b200 -lmsik -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,$PWD/maniskill_b200 -o msik_host_example
 *   ./msik_host_example tests/golden/c_host_panda.bin
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "msik.h"

#define CHECK_CUDA(x)                                                                  \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));     \
      return 2;                                                                        \
    }                                                                                  \
  } while (0)
#define CHECK_MSIK(x)                                                                  \
  do {                                                                                 \
    int rc_ = (x);                                                                     \
    if (rc_ != MSIK_OK) {                                                              \
      fprintf(stderr, "%s:%d: msik error %d: %s\n", __FILE__, __LINE__, rc_, msik_last_error()); \
      return 2;                                                                        \
    }                                                                                  \
  } while (0)

/* fixture header (little endian): magic, n_envs, qpos_stride, action_stride, n_ctrl, then
 * msik_chain_desc, msik_step_cfg, qpos[n_envs*qs], action[n_envs*as], expected[n_envs*n_ctrl] (fp32) */
typedef struct fixture_header {
  uint32_t magic;
  int32_t n_envs, qpos_stride, action_stride, n_ctrl;
} fixture_header;

static int read_exact(FILE* f, void* dst, size_t n) { return fread(dst, 1, n, f) == n ? 0 : -1; }

int main(int argc, char** argv) {
  const char* path = argc > 1 ? argv[1] : "tests/golden/c_host_panda.bin";
  FILE* f = fopen(path, "rb");
  if (!f) {
    fprintf(stderr, "cannot open %s\n", path);
    return 2;
  }
  fixture_header h;
  msik_chain_desc desc;
  msik_step_cfg cfg;
  if (read_exact(f, &h, sizeof h) || h.magic != 0x4b49534du /* "MSIK" */ || read_exact(f, &desc, sizeof desc) ||
      read_exact(f, &cfg, sizeof cfg)) {
    fprintf(stderr, "bad fixture %s\n", path);
    return 2;
  }
  const size_t nq = (size_t)h.n_envs * h.qpos_stride, na = (size_t)h.n_envs * h.action_stride, nt = (size_t)h.n_envs * h.n_ctrl;
  float* h_q = (float*)malloc(nq * sizeof(float));
  float* h_a = (float*)malloc(na * sizeof(float));
  float* h_want = (float*)malloc(nt * sizeof(float));
  float* h_got = (float*)malloc(nt * sizeof(float));
  if (!h_q || !h_a || !h_want || !h_got || read_exact(f, h_q, nq * sizeof(float)) || read_exact(f, h_a, na * sizeof(float)) ||
      read_exact(f, h_want, nt * sizeof(float))) {
    fprintf(stderr, "short fixture %s\n", path);
    return 2;
  }
  fclose(f);

  int sms = 0, major = 0, minor = 0;
  CHECK_MSIK(msik_device_check(&sms, &major, &minor));
  msik_chain_t chain = NULL;
  CHECK_MSIK(msik_chain_create(&desc, &chain));
  char kname[64] = "?";
  CHECK_MSIK(msik_chain_kernel_name(chain, kname, (int)sizeof kname));
  printf("libmsik %d on sm_%d%d (%d SMs), kernel for this chain: %s\n", msik_version(), major, minor, sms, kname);

  float *d_q = NULL, *d_a = NULL, *d_t = NULL;
  cudaStream_t stream;
  CHECK_CUDA(cudaStreamCreate(&stream));
  CHECK_CUDA(cudaMalloc((void**)&d_q, nq * sizeof(float)));
  CHECK_CUDA(cudaMalloc((void**)&d_a, na * sizeof(float)));
  CHECK_CUDA(cudaMalloc((void**)&d_t, nt * sizeof(float)));
  CHECK_CUDA(cudaMemcpyAsync(d_q, h_q, nq * sizeof(float), cudaMemcpyHostToDevice, stream));
  CHECK_CUDA(cudaMemcpyAsync(d_a, h_a, na * sizeof(float), cudaMemcpyHostToDevice, stream));
  /* the fixture's cfg has MSIK_F_NO_DRIVE: dense (n_envs, n_ctrl) joint targets, no drive-target buffer */
  CHECK_MSIK(msik_delta_step(chain, &cfg, d_q, NULL, d_a, NULL, d_t, NULL, (void*)stream));
  CHECK_CUDA(cudaMemcpyAsync(h_got, d_t, nt * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CHECK_CUDA(cudaStreamSynchronize(stream));

  double worst = 0.0;
  for (size_t i = 0; i < nt; ++i) {
    const double e = fabs((double)h_got[i] - (double)h_want[i]);
    if (!(e <= worst)) worst = e; /* also catches NaN */
  }
  printf("%d envs, max |joint target - oracle| = %.3g rad (bar 1e-5), %lld launches\n", h.n_envs, worst, (long long)msik_launch_count());
  CHECK_MSIK(msik_chain_destroy(chain));
  cudaFree(d_q); cudaFree(d_a); cudaFree(d_t);
  cudaStreamDestroy(stream);
  free(h_q); free(h_a); free(h_want); free(h_got);
  return worst < 1e-5 ? 0 : 1;
}
