// Persistent, TMA-staged variant of the path-A kernel (sm_100a).
//
// One warp per (SM x resident-warp slot) walks the env tiles  w, w + W, ...
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
#include <type_traits>

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
// shared-state-space accessors on 32-bit byte addresses: the loop keeps three per-lane row addresses
// in registers and everything else is an immediate or a constant-bank operand
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_u32(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_u32(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s_u32(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar)
               : "memory");
}
// L2 prefetch of a global range (no value is consumed: legal before griddepcontrol.wait, L2 is the coherence point)
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_s2g_u32(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
// one lane of the (converged) warp; the same lane for the same mask every time, so the bulk-group bookkeeping
// (commit / wait_group are per thread) stays with one thread.  ptxas issues UBLKCP under an elected predicate
// directly; under `lane == 0` it wraps every copy in a per-active-lane loop.
__device__ __forceinline__ bool elect_one() {
#ifdef MSIK_NO_ELECT
  return (threadIdx.x & 31) == 0;
#else
  uint32_t p;
  asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}\n" : "=r"(p));
  return p != 0;
#endif
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Warp-granular pipeline: every warp owns its own 2-stage (qpos, action) ring, 4 target
// buffers and mbarriers, and walks warp-tiles of 32 envs  w, w + W, w + 2W, ...  (W = resident
// warps of the grid).  Nothing in the loop synchronises across warps: lane 0 arms the mbarrier
// and issues the bulk copies, the warp waits on its own barrier, __syncwarp() orders the
// shared-memory hand-over.  A 32-row span is 128 * stride bytes: always a multiple of 16.
#ifndef MSIK_TMA_TPB
#define MSIK_TMA_TPB 128
#endif
constexpr int TMA_TPB = MSIK_TMA_TPB;  // threads per CTA of the persistent kernel
constexpr int WARPS = TMA_TPB / 32;

// bytes of dynamic shared memory of delta_step_tma_kernel for a step configuration
// (warps = warps per CTA: 4 for the usual rows; wide rows -- Realman's 41 floats -- run 2 or 1 warps per CTA so that the
// per-warp rings still fit, msik.cu choose_tma)
__host__ __device__ inline size_t tma_smem_bytes(const msik_step_cfg& c, bool drive, int warps = WARPS) {
  size_t fl = (size_t)QA_STAGES * 32 * (c.qpos_stride + c.action_stride);
  if (drive) fl += (size_t)T_BUFS * 32 * c.target_stride;
  return fl * warps * sizeof(float);
}

// Sibling joint-position controllers fused into the launch (msik_aux_cfg): elementwise, rounded
// exactly like the torch expressions they replace (separate multiply and add, no FMA contraction).
__device__ __forceinline__ void aux_targets(const StepParams& P, const float* __restrict__ qrow,
                                            const float* __restrict__ arow0, float* __restrict__ trow) {
  if (P.aux_uniform) {
    // every sibling column takes the same value (mimic joints, no use_delta): compute it once
    float v = arow0[P.aux.action_col[0]];
    if (P.aux.flags[0] & MSIK_AUX_NORMALIZE)
      v = __fadd_rn(P.aux.mid[0], __fmul_rn(P.aux.half[0], fminf(fmaxf(v, -1.f), 1.f)));
    for (int c = 0; c < P.aux.n_cols; ++c) trow[P.aux.col[c]] = v;
    return;
  }
  for (int c = 0; c < P.aux.n_cols; ++c) {
    float v = arow0[P.aux.action_col[c]];
    const int f = P.aux.flags[c];
    if (f & MSIK_AUX_NORMALIZE) v = __fadd_rn(P.aux.mid[c], __fmul_rn(P.aux.half[c], fminf(fmaxf(v, -1.f), 1.f)));
    if (f & MSIK_AUX_DELTA) v = __fadd_rn(qrow[P.aux.col[c]], v);
    trow[P.aux.col[c]] = v;
  }
}

// the same on shared-memory byte addresses: qa / ta = this lane's qpos / target row, aa = its action row at action_offset
__device__ __forceinline__ void aux_targets_smem(const StepParams& P, uint32_t qa, uint32_t aa, uint32_t ta) {
  if (P.aux_uniform) {
    float v = lds_f32(aa + 4u * (uint32_t)P.aux_acol_rel);
    if (P.aux.flags[0] & MSIK_AUX_NORMALIZE)
      v = __fadd_rn(P.aux.mid[0], __fmul_rn(P.aux.half[0], fminf(fmaxf(v, -1.f), 1.f)));
    sts_f32(ta + 4u * (uint32_t)P.aux.col[0], v);
    if (P.aux.n_cols > 1) sts_f32(ta + 4u * (uint32_t)P.aux.col[1], v);
    for (int c = 2; c < P.aux.n_cols; ++c) sts_f32(ta + 4u * (uint32_t)P.aux.col[c], v);
    return;
  }
  for (int c = 0; c < P.aux.n_cols; ++c) {
    float v = lds_f32(aa + 4u * (uint32_t)(P.aux.action_col[c] - P.cfg.action_offset));
    const int f = P.aux.flags[c];
    if (f & MSIK_AUX_NORMALIZE) v = __fadd_rn(P.aux.mid[c], __fmul_rn(P.aux.half[c], fminf(fmaxf(v, -1.f), 1.f)));
    if (f & MSIK_AUX_DELTA) v = __fadd_rn(lds_f32(qa + 4u * (uint32_t)P.aux.col[c]), v);
    sts_f32(ta + 4u * (uint32_t)P.aux.col[c], v);
  }
}

// IDENT: chain joint i sits in qpos / target column i (stock Panda, xArm): shared-memory
// accesses get immediate offsets instead of one address add each.
#ifndef MSIK_TMA_MIN_CTAS
#define MSIK_TMA_MIN_CTAS 5
#endif
#ifndef MSIK_TMA_PREFETCH
#define MSIK_TMA_PREFETCH 1  // tiles per warp prefetched into L2 ahead of griddepcontrol.wait (0: none; 1 / 2 / 3 / 4 measured: 19.35 / 19.44 / 19.65 / 20.18 us)
#endif
// LEAN: the stock stepping configuration, decided on the host (tma_lean): drive targets written, no tq / sq
// copies, no debug stage.  The cold branches leave the loop, which shortens it and its instruction-cache footprint.
// LEAN values: 0 = general; 1 = stock Panda-like dict: arm + ONE normalised action column driving one or two sibling
// columns (mimic gripper) which together cover the whole target row, no velocity controller -- neither the column count,
// the flag nor the old-target reload is tested per tile;
// 2 = the same hoisting (drive targets written, no tq / sq copies, no debug stage) with ANY fused sibling columns
// (none included) and the fused velocity controller kept inside the loop -- the arm alone, the stock Fetch dict
// (arm + gripper + body + base).
// Chain2 (with its stock controlled range C02, NR2) = the second EE controller of the same CombinedController, solved from
// the same staged rows in the same iteration (msik_delta_step_dual; the fork's Realman robot drives both arms from one
// flat action, mobile_realman.py:436-443): one pass over the 164-byte rows instead of two.  void = none.
struct NoChain { static constexpr int N = 1; };
template <class Chain, int C0, int NR, int NC, bool MASKED, bool IDENT, int LEAN = 0, class Chain2 = NoChain, int C02 = 0, int NR2 = 1>
__global__ void __launch_bounds__(TMA_TPB, MSIK_TMA_MIN_CTAS)
delta_step_tma_kernel(const __grid_constant__ StepParams P) {
  constexpr int N = Chain::N;
  constexpr bool DUAL = !std::is_same<Chain2, NoChain>::value;
  constexpr int N2 = Chain2::N;
  static_assert(!DUAL || (LEAN != 0 && !MASKED), "dual launches serve the stock stepping configuration only");
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t full[WARPS][QA_STAGES];

  const int lane = threadIdx.x & 31;
  // warp-uniform (lets ptxas use uniform registers); a one-warp CTA makes every address below CTA-uniform
  const int warp = WARPS == 1 ? 0 : __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int qs = P.cfg.qpos_stride, as = P.cfg.action_stride, ts = P.cfg.target_stride;
  const bool drive = LEAN || !(P.cfg.flags & MSIK_F_NO_DRIVE);
  // old targets are needed only to preserve foreign columns; LEAN = 1 is launched for whole-row writers only (tma_lean)
  const bool load_t = LEAN != 1 && drive && !P.full_row;
  const uint32_t qbytes = (uint32_t)qs * 128u, abytes = (uint32_t)as * 128u, tbytes = (uint32_t)ts * 128u;
  uint32_t txbytes = qbytes + abytes + (load_t ? tbytes : 0u);
  asm volatile("" : "+r"(txbytes));  // kept in a register (the arrive takes one): ptxas otherwise re-derives it per tile

  // this warp's slice of shared memory: [q0 q1][a0 a1][t0 t1 t2 t3], as shared-space byte addresses
  const uint32_t per_warp = QA_STAGES * (qbytes + abytes) + (drive ? T_BUFS * tbytes : 0u);
  const uint32_t wq = smem_u32(smem) + (uint32_t)warp * per_warp;  // warp-uniform bases
  const uint32_t wa = wq + QA_STAGES * qbytes;
  const uint32_t wt = wa + QA_STAGES * abytes;
  const uint32_t bar0 = smem_u32(full[warp]);

  const int nfull = P.cfg.n_envs >> 5;  // full warp-tiles; the ragged tail is handled after the loop
  const int nw = (int)(blockDim.x >> 5);  // warps per CTA of THIS launch (<= WARPS)
  const int first = blockIdx.x * nw + warp, step = gridDim.x * nw;
  const int mine = first < nfull ? (nfull - first + step - 1) / step : 0;

  if (lane == 0) {  /*init*/
#pragma unroll
    for (int s = 0; s < QA_STAGES; ++s) mbar_init(&full[warp][s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();

  // running global addresses of the next tile to load / to store (bytes advance by a constant).
  // They are kept warp-uniform (every lane advances them), so the bulk-copy operands sit in uniform registers.
  const char* gq = reinterpret_cast<const char*>(P.qpos) + (size_t)(uint32_t)first * qbytes;
  const char* ga = reinterpret_cast<const char*>(P.action) + (size_t)(uint32_t)first * abytes;
  char* gs = reinterpret_cast<char*>(P.drive) + (size_t)(uint32_t)first * tbytes;
  // (unsigned: a signed `step` costs a sign-extension + 64-bit multiply-add per pointer per tile where ptxas rematerialises these)
  const size_t dq_b = (size_t)(uint32_t)step * qbytes, da_b = (size_t)(uint32_t)step * abytes, dt_b = (size_t)(uint32_t)step * tbytes;

  // lane 0 starts the loads of one tile into stage s / target buffer tb; `gt` = its old-target rows
  // (sq / sa / bar = shared-memory addresses of the stage's qpos rows, action rows and mbarrier)
  auto issue = [&](uint32_t sq, uint32_t sa, uint32_t bar, uint32_t tb, const char* q_src, const char* a_src, const char* gt) {
    mbar_expect_tx_u32(bar, txbytes);
    bulk_g2s_u32(sq, q_src, qbytes, bar);
    bulk_g2s_u32(sa, a_src, abytes, bar);
    if (load_t) bulk_g2s_u32(wt + tb * tbytes, gt, tbytes, bar);
  };
  // programmatic dependent launch: everything above overlapped the previous kernel's tail; wait for
  // its memory before the first global access, then let OUR successor start scheduling early
#if MSIK_TMA_PREFETCH > 0
  // While the previous kernel of the stream drains (this grid was scheduled early: programmatic dependent launch), pull this
  // warp's first tiles towards L2.  No value is consumed before the wait below, and L2 is the coherence point of every
  // writer, so this is legal ahead of griddepcontrol.wait; it takes the DRAM latency of the first loads (about 1 us per
  // launch, the larger part of the per-launch ramp) off the critical path.
  if (elect_one()) {
#pragma unroll
    for (int t = 0; t < MSIK_TMA_PREFETCH; ++t)
      if (t < mine) {
        bulk_prefetch_l2(gq + t * dq_b, qbytes);
        bulk_prefetch_l2(ga + t * da_b, abytes);
        if (load_t) bulk_prefetch_l2(gs + t * dt_b, tbytes);
      }
  }
#endif
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (elect_one()) {
    if (mine > 0) issue(wq, wa, bar0, 0, gq, ga, gs);
    if (mine > 1) issue(wq + qbytes, wa + abytes, bar0 + 8u, 1, gq + dq_b, ga + da_b, gs + dt_b);
  }
  gq += 2 * dq_b; ga += 2 * da_b;  // tile k + 2 is loaded while tile k is solved

  // the loop state of a lane: the byte offsets of its row inside a qpos / action stage, the address of its row in
  // target buffer 0.  The empty asm keeps them in registers (ptxas otherwise rebuilds them from the constant bank
  // every iteration).
  uint32_t lq = (uint32_t)lane * (uint32_t)qs * 4u;
  uint32_t la = ((uint32_t)lane * (uint32_t)as + (uint32_t)P.cfg.action_offset) * 4u;
  uint32_t lt = wt + (uint32_t)lane * (uint32_t)ts * 4u;
  asm volatile("" : "+r"(lq), "+r"(la), "+r"(lt));
  // LEAN == 1: the (at most two) sibling columns as ready-made row addresses and constants, for the same reason
  uint32_t la_aux = la + 4u * (uint32_t)P.aux_acol_rel, lt_aux0 = lt + 4u * (uint32_t)P.aux.col[0];
  uint32_t lt_aux1 = lt + 4u * (uint32_t)P.aux.col[P.aux.n_cols > 1 ? 1 : 0];  // one column: the same word is stored twice
  float aux_mid = P.aux.mid[0], aux_half = P.aux.half[0];
  if constexpr (LEAN == 1) asm volatile("" : "+r"(la_aux), "+r"(lt_aux0), "+r"(lt_aux1), "+f"(aux_mid), "+f"(aux_half));

  // The stage of tile k alternates (QA_STAGES = 2): the shared-memory addresses of its rows and of its mbarrier, and the
  // phase parity to wait for, are carried as XOR toggles between the two stages' values (one uniform instruction each)
  // instead of being derived from k (shift, mask, multiply, add: the kernel is issue-bound).
  static_assert(QA_STAGES == 2, "stage addresses are toggled");
  const uint32_t xq = wq ^ (wq + qbytes), xa = wa ^ (wa + abytes), xb = bar0 ^ (bar0 + 8u);
  uint32_t sq = wq, sa = wa, sbar = bar0, phase = 0, odd = 0;
  const int last_refill = mine - QA_STAGES;  // tiles k < last_refill have a successor two tiles on to prefetch
  for (int k = 0; k < mine; ++k) {
    const uint32_t tb = (uint32_t)k & (T_BUFS - 1);
    mbar_wait_u32(sbar, phase);
    const uint32_t qa = sq + lq, aa = sa + la, ta = lt + tb * tbytes;
    float q[N], a[NC];
#pragma unroll
    for (int i = 0; i < N; ++i) q[i] = lds_f32(qa + 4u * (uint32_t)(IDENT ? i : P.cfg.anc_col[i]));
#pragma unroll
    for (int j = 0; j < NC; ++j) a[j] = lds_f32(aa + 4u * j);
    float q2[N2], a2[NC];
    if constexpr (DUAL) {
#pragma unroll
      for (int i = 0; i < N2; ++i) q2[i] = lds_f32(qa + 4u * (uint32_t)P.cfg2.anc_col[i]);
      const uint32_t aa2 = aa + 4u * (uint32_t)(P.cfg2.action_offset - P.cfg.action_offset);
#pragma unroll
      for (int j = 0; j < NC; ++j) a2[j] = lds_f32(aa2 + 4u * j);
    }
    if constexpr (LEAN == 1) {  // one normalised action column -> one or two sibling columns (tma_lean)
      const float v = __fadd_rn(aux_mid, __fmul_rn(aux_half, fminf(fmaxf(lds_f32(sa + la_aux), -1.f), 1.f)));
      sts_f32(lt_aux0 + tb * tbytes, v);
      sts_f32(lt_aux1 + tb * tbytes, v);
    } else {
      if (drive && P.aux.n_cols) aux_targets_smem(P, qa, aa, ta);
      if (P.vel.n_cols) {  // fused velocity controller: reads this env's staged rows, per-lane stores into the qvel row
        const float* srow = reinterpret_cast<const float*>(reinterpret_cast<const char*>(smem) + (aa - smem_u32(smem)));
        const float ori = P.vel.ego == MSIK_VEL_JOINT ? 0.f : lds_f32(qa + 4u * (uint32_t)P.vel.ori_col);
        const size_t env = (size_t)(first + k * step) * 32 + lane;
        vel_apply(P.vel, srow + (P.vel.action_offset - P.cfg.action_offset), ori, P.qvel + env * P.vel.qvel_stride);
      }
    }
    // Stage s is refilled by the bulk-copy engine (async proxy) right here, while the loads above went through the generic
    // proxy: without a cross-proxy fence nothing orders them (ptxas drops the converged __syncwarp and does not wait for the
    // LDS results before UBLKCP), and an L2-hot refill can land before a late LDS executes -- wrong rows in the first
    // 128-byte chunks of a tile at 2^20 envs, round 1.  (Issuing the refill after the store instead,
    // under the ONE fence at the end of the iteration: measured 22.2 vs 21.9 us per 2^20 envs -- the lost iteration of
    // prefetch lead costs more than the fence saves, profiles/r02g_tma_variants.txt.)
    fence_proxy_async();
    __syncwarp();
    if (k < last_refill) {
      if (elect_one()) {
        bulk_wait_read<1>();
        issue(sq, sa, sbar, ((uint32_t)k + QA_STAGES) & (T_BUFS - 1), gq, ga, gs + 2 * dt_b);
      }
      gq += dq_b; ga += da_b;
    }
    phase ^= odd;  // both stages of a phase consumed -> the other parity
    odd ^= 1u; sq ^= xq; sa ^= xa; sbar ^= xb;

    float qn[NR];
    if (!LEAN && (P.debug & 1)) {
#pragma unroll
      for (int r = 0; r < NR; ++r) qn[r] = q[C0 + r] + a[r % NC];
    } else {
      solve_env<Chain, C0, NR, NC, MASKED>(P, q, a, qn);
    }

    if (!LEAN && (P.tq_out || P.sq_out)) {
      const size_t env = (size_t)(first + k * step) * 32 + lane;
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
#pragma unroll
      for (int r = 0; r < NR; ++r) {
        const bool on = !MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u);
        if (on) sts_f32(ta + 4u * (uint32_t)(IDENT ? C0 + r : P.cfg.anc_col[C0 + r]), qn[r]);
      }
    }
    if constexpr (DUAL) {  // second arm: same rows, its own chain / action slice / columns
      float qn2[NR2];
      solve_env<Chain2, C02, NR2, NC, false, 1>(P, q2, a2, qn2);
#pragma unroll
      for (int r = 0; r < NR2; ++r) sts_f32(ta + 4u * (uint32_t)P.cfg2.anc_col[C02 + r], qn2[r]);
    }
    fence_proxy_async();  // generic-proxy writes of the target tile -> visible to the bulk-copy engine
    __syncwarp();
    if (elect_one()) {
      if (drive) {
        bulk_s2g_u32(gs, wt + tb * tbytes, tbytes);
        bulk_commit();
      }
    }
    if (drive) gs += dt_b;
  }
  if (drive && elect_one()) bulk_wait_all<0>();  // shared memory must outlive the bulk stores

  // ragged tail (n_envs % 32 rows): one warp, plain per-lane loads / stores of its own row
  const int tail = P.cfg.n_envs - (nfull << 5);
  if (tail > 0 && first == nfull % step && lane < tail) {
    const size_t env = (size_t)(nfull << 5) + lane;
    float q[N], a[NC];
#pragma unroll
    for (int i = 0; i < N; ++i) q[i] = __ldg(P.qpos + env * qs + P.cfg.anc_col[i]);
#pragma unroll
    for (int j = 0; j < NC; ++j) a[j] = __ldg(P.action + env * as + P.cfg.action_offset + j);
    if (drive && P.aux.n_cols) aux_targets(P, P.qpos + env * qs, P.action + env * as, P.drive + env * ts);
    if (P.vel.n_cols)
      vel_apply(P.vel, P.action + env * as + P.vel.action_offset,
                P.vel.ego == MSIK_VEL_JOINT ? 0.f : __ldg(P.qpos + env * qs + P.vel.ori_col), P.qvel + env * P.vel.qvel_stride);
    float qn[NR];
    solve_env<Chain, C0, NR, NC, MASKED>(P, q, a, qn);
    int j = 0;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      const bool on = !MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u);
      if (on) {
        if (P.tq_out) P.tq_out[env * P.n_ctrl + j] = qn[r];
        if (P.sq_out) P.sq_out[env * P.n_ctrl + j] = q[C0 + r];
        if (drive) P.drive[env * ts + P.cfg.anc_col[C0 + r]] = qn[r];
        ++j;
      }
    }
    if constexpr (DUAL) {
      float q2[N2], a2[NC], qn2[NR2];
#pragma unroll
      for (int i = 0; i < N2; ++i) q2[i] = __ldg(P.qpos + env * qs + P.cfg2.anc_col[i]);
#pragma unroll
      for (int jj = 0; jj < NC; ++jj) a2[jj] = __ldg(P.action + env * as + P.cfg2.action_offset + jj);
      solve_env<Chain2, C02, NR2, NC, false, 1>(P, q2, a2, qn2);
#pragma unroll
      for (int r = 0; r < NR2; ++r) P.drive[env * ts + P.cfg2.anc_col[C02 + r]] = qn2[r];
    }
  }
}

}  // namespace msik
