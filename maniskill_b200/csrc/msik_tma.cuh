// Persistent, TMA-staged variant of the path-A kernel (sm_100a).
//
// One CTA per (SM x resident-CTA slot) walks the env tiles  blockIdx.x, +gridDim.x, ...
// Each tile's PhysX rows (qpos, flat action, old drive targets) are contiguous spans, so they
// are moved by the bulk-copy engine (cp.async.bulk, 1-D TMA) straight into shared memory and
// signalled on an mbarrier: no per-thread load instructions, no address arithmetic, and the
// loads of tile k+2 are in flight while tile k is computed (2-stage ring for qpos/action,
// 4 buffers for the target tile because it is also the source of the asynchronous bulk
// store of the results).  Threads then pick their own row from shared memory (odd row
// strides -> conflict-free), do the whole FK / Jacobian / min-norm solve in registers and
// drop the new joint targets into the staged target tile, which one thread hands back to
// the bulk-copy engine (cp.async.bulk.global.shared::cta).
#pragma once
#include "msik_kernels.cuh"

namespace msik {

constexpr int QA_STAGES = 2;
constexpr int T_BUFS = 4;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared -> global, tracked by the bulk async-group of the issuing thread
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// bytes of dynamic shared memory of delta_step_tma_kernel for a step configuration
__host__ __device__ inline size_t tma_smem_bytes(const msik_step_cfg& c, bool drive) {
  size_t fl = (size_t)QA_STAGES * (round4(TPB * c.qpos_stride) + round4(TPB * c.action_stride));
  if (drive) fl += (size_t)T_BUFS * round4(TPB * c.target_stride);
  return fl * sizeof(float);
}

template <class Chain, int C0, int NR, int NC, bool MASKED>
__global__ void __launch_bounds__(TPB, 6)
delta_step_tma_kernel(const __grid_constant__ StepParams P) {
  constexpr int N = Chain::N;
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t full[QA_STAGES];

  const int tid = threadIdx.x;
  const int qs = P.cfg.qpos_stride, as = P.cfg.action_stride, ts = P.cfg.target_stride;
  const bool drive = !(P.cfg.flags & MSIK_F_NO_DRIVE);
  const uint32_t qbytes = (uint32_t)(TPB * qs) * 4u, abytes = (uint32_t)(TPB * as) * 4u, tbytes = (uint32_t)(TPB * ts) * 4u;

  // shared-memory carve-up: [q stage 0..1][a stage 0..1][t buffer 0..3], each 16-byte aligned
  const int qfl = round4(TPB * qs), afl = round4(TPB * as), tfl = round4(TPB * ts);
  float* const q_base = smem;
  float* const a_base = q_base + QA_STAGES * qfl;
  float* const t_base = a_base + QA_STAGES * afl;

  const int nfull = P.cfg.n_envs / TPB;  // full tiles; the ragged tail is handled after the loop
  const int first = blockIdx.x, step = gridDim.x;
  const int mine = first < nfull ? (nfull - first + step - 1) / step : 0;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < QA_STAGES; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto issue = [&](int k) {  // thread 0: start the loads of this CTA's k-th tile
    const int s = k % QA_STAGES;
    const size_t env0 = (size_t)(first + k * step) * TPB;
    mbar_expect_tx(&full[s], qbytes + abytes + (drive ? tbytes : 0u));
    bulk_g2s(q_base + s * qfl, P.qpos + env0 * qs, qbytes, &full[s]);
    bulk_g2s(a_base + s * afl, P.action + env0 * as, abytes, &full[s]);
    if (drive) bulk_g2s(t_base + (k % T_BUFS) * tfl, P.drive + env0 * ts, tbytes, &full[s]);
  };
  if (tid == 0) {
    if (mine > 0) issue(0);
    if (mine > 1) issue(1);
  }

  // per-thread shared-memory offsets of this thread's row
  const int qrow = tid * qs, arow = tid * as + P.cfg.action_offset, trow = tid * ts;

  for (int k = 0; k < mine; ++k) {
    const int s = k % QA_STAGES;
    mbar_wait(&full[s], (uint32_t)(k / QA_STAGES) & 1u);
    float q[N], a[NC];
#pragma unroll
    for (int i = 0; i < N; ++i) q[i] = q_base[s * qfl + qrow + P.cfg.anc_col[i]];
#pragma unroll
    for (int j = 0; j < NC; ++j) a[j] = a_base[s * afl + arow + j];
    __syncthreads();  // stage s (qpos / action) has been consumed by every thread
    if (tid == 0 && k + QA_STAGES < mine) {
      bulk_wait_read<1>();  // the store that last read target buffer (k+2)%4 (tile k-2) is done
      issue(k + QA_STAGES);
    }

    float b[NC];
    preprocess_action<NC>(P, a, b);
    float sn[N], cs[N];
    joint_sincos<Chain>(P, q, sn, cs);
    float Jl[N][3], Ja[N][3], pe[3], Re[9];
    Chain::fkjac(P.chain, q, sn, cs, Jl, Ja, pe, Re);
    float A[NR][NC];
    build_At<Chain, C0, NR, NC, MASKED>(P, Jl, Ja, A);
    float dq[NR], qn[NR];
    min_norm_solve<NR, NC>(A, b, dq, P.rank_rtol);
#pragma unroll
    for (int r = 0; r < NR; ++r) qn[r] = q[C0 + r] + dq[r];

    const size_t env = (size_t)(first + k * step) * TPB + tid;
    if (P.tq_out || P.sq_out) {
      int j = 0;
#pragma unroll
      for (int r = 0; r < NR; ++r) {
        const bool on = !MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u);
        if (on) {
          if (P.tq_out) P.tq_out[env * P.n_ctrl + j] = qn[r];
          if (P.sq_out) P.sq_out[env * P.n_ctrl + j] = q[C0 + r];
          ++j;
        }
      }
    }
    if (drive) {
      float* tt = t_base + (k % T_BUFS) * tfl;
#pragma unroll
      for (int r = 0; r < NR; ++r) {
        const bool on = !MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u);
        if (on) tt[trow + P.cfg.anc_col[C0 + r]] = qn[r];
      }
      fence_proxy_async();  // generic-proxy writes -> visible to the bulk-copy engine
      __syncthreads();
      if (tid == 0) {
        bulk_s2g(P.drive + (size_t)(first + k * step) * TPB * ts, tt, tbytes);
        bulk_commit();
      }
    }
  }
  if (tid == 0 && drive) bulk_wait_all<0>();

  // ragged tail (n_envs % TPB rows): generic loads / stores, done by one CTA
  const int tail = P.cfg.n_envs - nfull * TPB;
  if (tail > 0 && blockIdx.x == (unsigned)(nfull % step)) {
    __syncthreads();
    const int env0 = nfull * TPB;
    Tiles Tt;
    Tt.q = q_base; Tt.a = a_base; Tt.t = t_base;
    float q[N], a[NC];
    int row;
    StepParams const& Pr = P;
    load_env<Chain, NC, false>(Pr, Tt, env0, tail, drive, q, a, row);
    float b[NC];
    preprocess_action<NC>(P, a, b);
    float sn[N], cs[N];
    joint_sincos<Chain>(P, q, sn, cs);
    float Jl[N][3], Ja[N][3], pe[3], Re[9];
    Chain::fkjac(P.chain, q, sn, cs, Jl, Ja, pe, Re);
    float A[NR][NC];
    build_At<Chain, C0, NR, NC, MASKED>(P, Jl, Ja, A);
    float dq[NR], qn[NR];
    min_norm_solve<NR, NC>(A, b, dq, P.rank_rtol);
#pragma unroll
    for (int r = 0; r < NR; ++r) qn[r] = q[C0 + r] + dq[r];
    store_env<Chain, C0, NR, MASKED, false>(P, Tt, env0, tail, row, q, qn);
  }
}

}  // namespace msik
