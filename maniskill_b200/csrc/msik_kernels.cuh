// Device code of the batched end-effector controller (sm_100a).
//
// Mapping: ONE THREAD PER ENV.  The whole per-env problem (7-11 joint values, the
// 6 x n Jacobian, the Householder factors) lives in registers, there is no
// inter-thread communication in the math, and global memory is touched only
// through block-wide tiles: each CTA of TPB envs moves the contiguous span of
// TPB PhysX rows with 128-bit coalesced accesses into shared memory, where every
// thread picks / deposits its own row (row strides 9, 7, 15 ... are odd => the
// stride-`row` shared-memory accesses are bank-conflict free).
// The warp-cooperative mapping (lane per joint, shuffle prefix products) is msik_warp.cuh: built for the
// iterative path B, where few envs leave the SMs idle; DESIGN.md section 4 has the measured A/B.
#pragma once
#include "../../include/msik.h"
#include "msik_math.cuh"
#include "generated/chains.cuh"
#include "msik_ik.cuh"

namespace msik {

constexpr int TPB = 128;

struct StepParams {
  msik_chain_desc chain;  // padded to the kernel's joint bucket (generic kernels only)
  msik_step_cfg cfg;
  msik_iter_cfg it;
  const float* qpos;
  const int32_t* row_index;
  const float* action;
  float* drive;
  float* tq_out;
  float* sq_out;
  const float* prev_pose;
  float* pose_out;
  int32_t* iters_out;
  float* err_out;
  float* jac_out;
  float mid[3];   // 0.5 (high + low)
  float half[3];  // 0.5 (high - low)
  unsigned long long colmask;  // controlled columns of a target row (stride <= 64)
  float rank_rtol;
  int32_t vec_ok;  // all tile bases 16-byte aligned
  int32_t merge;   // 1: read old target rows, merge, store whole rows as float4
  int32_t n_ctrl;
  int32_t debug;     // bit 0: skip the FK / solve (memory-pipeline floor measurement, MSIK_DEBUG_NOMATH=1)
  int32_t aux_uniform;  // all aux columns share action column / flags / scale and none is a delta
  int32_t full_row;  // arm + aux columns cover the whole target row: old targets are not read
  int32_t aux_acol_rel;  // aux.action_col[0] - cfg.action_offset (uniform aux: one action column for all)
  int32_t na;        // target_ik_kernel: action dims (3 | 6)
  int32_t gather;    // target_ik_kernel: per-lane row access (row_index / rows too wide for CTA tiles)
  float ik_guard;    // target_ik_kernel: smallest pivot the fast step is trusted with (-1: damping == 0)
  float ik_lambda;   // sqrt(damping)
  int32_t ik_xoff;   // target_ik_kernel: float offset of the compaction exchange buffer in dynamic smem (-1: no compaction)
  int32_t ik_compact_from;  // first iteration with a compaction checkpoint
  int32_t ik_sort;   // target_ik_kernel: 1 = regroup the CTA's envs by remaining work after iteration 0 (needs ik_xoff >= 0)
  // second EE controller of the same CombinedController solved in the same launch (delta_step_tma_kernel<..., Chain2>;
  // msik_delta_step_dual): only the fields its solve and its row access need
  int32_t tma_warps;  // warps per CTA of the persistent launch (host side only)
  int32_t dual;       // 1: cfg2 & co. are valid
  msik_step_cfg cfg2;
  float mid2[3], half2[3];
  float rank_rtol2;
  msik_vel_cfg vel;  // velocity controller fused into the launch (n_cols = 0: none), see msik_delta_step_fused_vel
  float* qvel;       // cuda_articulation_target_qvel
  msik_aux_cfg aux;  // sibling joint-position controllers fused into the launch (n_cols = 0: none)
};

// One env of the velocity controllers (pd_joint_vel.py:38-42, pd_base_vel.py:19-33,50-72; include/msik.h msik_vel_targets):
// arow = this controller's slice of the env's action row, ori = qpos of the orientation joint, vrow = the env's row of
// cuda_articulation_target_qvel.
__device__ __forceinline__ void vel_apply(const msik_vel_cfg& c, const float* __restrict__ arow, float ori, float* __restrict__ vrow) {
  const int na = c.ego == MSIK_VEL_FORWARD ? c.n_cols - 1 : c.n_cols;
  auto act = [&](int k) {
    float v = arow[k];
    if (c.flags & MSIK_AUX_NORMALIZE) v = __fadd_rn(c.mid[k], __fmul_rn(c.half[k], fminf(fmaxf(v, -1.f), 1.f)));
    return v;
  };
  if (c.ego == MSIK_VEL_JOINT) {
    for (int k = 0; k < na; ++k) vrow[c.col[k]] = act(k);
    return;
  }
  float sn, cs;
  sincos_joint(ori, sn, cs);
  const float a0 = act(0);
  if (c.ego == MSIK_VEL_BASE) {
    const float a1 = act(1);
    vrow[c.col[0]] = fmaf(cs, a0, -sn * a1);
    vrow[c.col[1]] = fmaf(sn, a0, cs * a1);
    for (int k = 2; k < na; ++k) vrow[c.col[k]] = act(k);
  } else {  // forward only: the y velocity fed to the rotation is zero, action entries 1.. are joints 2..
    vrow[c.col[0]] = cs * a0;
    vrow[c.col[1]] = sn * a0;
    for (int k = 1; k < na; ++k) vrow[c.col[k + 1]] = act(k);
  }
}

// ----------------------------------------------------------------------------
// generic table-driven chain: N is a padded joint bucket, joint_type 0 = padding
// ----------------------------------------------------------------------------
template <int NN>
struct ChainGeneric {
  static constexpr int N = NN;
  static constexpr bool kGeneric = true;
  static constexpr bool kLastLinZero = false;
  static constexpr uint32_t kRevoluteMask = 0xffffffffu;
  static constexpr bool kHasReduced = false;  // the reduced solve exists for generated chains only
  static constexpr int kRedC0 = 0, kRedNR = 1, kRedNU = 4, kRedM = 0, kRedSkip = 0;
  static constexpr uint32_t kRedZB = 0u;
  static __device__ __forceinline__ void reduced(const float (&)[N], const float (&)[N], const float (&)[N], const float (&)[6],
                                                 float (&)[kRedNU][4], float (&)[2][kRedNU], float (&)[4], float (&)[2]) {}
  static __device__ __forceinline__ void reduced_scatter(const float (&)[kRedNU], const float (&)[2], float (&)[kRedNR]) {}
  static __device__ __forceinline__ void reduced_ik_sys(const float (&)[N], const float (&)[N], const float (&)[N], const float (&)[3],
                                                        const float (&)[9], float (&)[kRedNU][4], float (&)[2][kRedNU], float (&)[3],
                                                        float (&)[9], float (&)[9], float (&)[3]) {}
  static __device__ __forceinline__ void reduced_ik_rhs(const float (&)[3], const float (&)[3], const float (&)[3], float (&)[4],
                                                        float (&)[2]) {}
  static constexpr bool kHasLinear = false;  // nor the frame-optimised 3-dof system
  static constexpr int kLinC0 = 0, kLinNR = 1;
  static constexpr bool kLinLastZero = false;
  static __device__ __forceinline__ void linear(const float (&)[N], const float (&)[N], const float (&)[N], const float (&)[3],
                                                float (&)[kLinNR][3], float (&)[3]) {}
  static __device__ __forceinline__ void fkjac(const msik_chain_desc& d, const float (&q)[N],
                                               const float (&sn)[N], const float (&cs)[N],
                                               float (&Jl)[N][3], float (&Ja)[N][3], float (&pe)[3],
                                               float (&Re)[9]) {
    float R[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f};
    // position = base + p: `base` takes every translation that precedes the first revolute
    // joint (e.g. Fetch's +-20 m base prismatics), so the lever arms pe - o_i below are
    // formed from `p` alone and do not cancel against a large base offset in fp32.
    float base[3] = {0.f, 0.f, 0.f};
    float p[3] = {0.f, 0.f, 0.f};
    float o[N][3];
    bool seen = false;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const float* g = d.pre[i];
      float T[9];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float c = fmaf(R[3 * r], g[9], fmaf(R[3 * r + 1], g[10], R[3 * r + 2] * g[11]));
        p[r] += seen ? c : 0.f;
        base[r] += seen ? 0.f : c;
#pragma unroll
        for (int j = 0; j < 3; ++j)
          T[3 * r + j] = fmaf(R[3 * r], g[j], fmaf(R[3 * r + 1], g[3 + j], R[3 * r + 2] * g[6 + j]));
      }
      const int ty = d.joint_type[i];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        Ja[i][r] = T[3 * r + 2];  // joint axis z_i (angular column; fixed up below)
        o[i][r] = p[r];
      }
      if (ty == MSIK_JOINT_REVOLUTE) {
        seen = true;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          R[3 * r] = fmaf(cs[i], T[3 * r], sn[i] * T[3 * r + 1]);
          R[3 * r + 1] = fmaf(cs[i], T[3 * r + 1], -sn[i] * T[3 * r]);
          R[3 * r + 2] = T[3 * r + 2];
        }
      } else {
#pragma unroll
        for (int k = 0; k < 9; ++k) R[k] = T[k];
        if (ty == MSIK_JOINT_PRISMATIC) {
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const float c = T[3 * r + 2] * q[i];
            p[r] += seen ? c : 0.f;
            base[r] += seen ? 0.f : c;
          }
        }
      }
    }
    const float* g = d.pre[N];
    float pr[3];  // EE position relative to `base`
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      pr[r] = fmaf(R[3 * r], g[9], fmaf(R[3 * r + 1], g[10], fmaf(R[3 * r + 2], g[11], p[r])));
      pe[r] = base[r] + pr[r];
#pragma unroll
      for (int j = 0; j < 3; ++j)
        Re[3 * r + j] = fmaf(R[3 * r], g[j], fmaf(R[3 * r + 1], g[3 + j], R[3 * r + 2] * g[6 + j]));
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const int ty = d.joint_type[i];
      const float zx = Ja[i][0], zy = Ja[i][1], zz = Ja[i][2];
      if (ty == MSIK_JOINT_REVOLUTE) {
        const float dx = pr[0] - o[i][0], dy = pr[1] - o[i][1], dz = pr[2] - o[i][2];
        Jl[i][0] = zy * dz - zz * dy;
        Jl[i][1] = zz * dx - zx * dz;
        Jl[i][2] = zx * dy - zy * dx;
      } else if (ty == MSIK_JOINT_PRISMATIC) {
        Jl[i][0] = zx; Jl[i][1] = zy; Jl[i][2] = zz;
        Ja[i][0] = 0.f; Ja[i][1] = 0.f; Ja[i][2] = 0.f;
      } else {
        Jl[i][0] = 0.f; Jl[i][1] = 0.f; Jl[i][2] = 0.f;
        Ja[i][0] = 0.f; Ja[i][1] = 0.f; Ja[i][2] = 0.f;
      }
    }
  }
};

// Table-driven chain of EXACTLY NN revolute joints (the common case of the robots without generated code: 6- / 7-axis arms
// such as the reference's xArm6 and xMate3, whose URDFs are downloaded at run time).  Same table as ChainGeneric, but no
// joint-type dispatch, no prismatic terms and no base / relative split of the position (there are no base prismatics whose
// +-20 m offsets the split keeps out of the lever arms): the straight-line code is a third shorter.
template <int NN>
struct ChainGenericRev : ChainGeneric<NN> {
  static constexpr int N = NN;
  static constexpr bool kGeneric = false;  // (joint types are not read from the table: every joint is revolute)
  static __device__ __forceinline__ void fkjac(const msik_chain_desc& d, const float (&q)[N], const float (&sn)[N], const float (&cs)[N],
                                               float (&Jl)[N][3], float (&Ja)[N][3], float (&pe)[3], float (&Re)[9]) {
    (void)q;
    float R[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f};
    float p[3] = {0.f, 0.f, 0.f};
    float o[N][3];
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const float* g = d.pre[i];
      float T[9];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        p[r] = fmaf(R[3 * r], g[9], fmaf(R[3 * r + 1], g[10], fmaf(R[3 * r + 2], g[11], p[r])));
#pragma unroll
        for (int j = 0; j < 3; ++j)
          T[3 * r + j] = fmaf(R[3 * r], g[j], fmaf(R[3 * r + 1], g[3 + j], R[3 * r + 2] * g[6 + j]));
      }
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        Ja[i][r] = T[3 * r + 2];  // joint axis z_i = angular column
        o[i][r] = p[r];
        R[3 * r] = fmaf(cs[i], T[3 * r], sn[i] * T[3 * r + 1]);
        R[3 * r + 1] = fmaf(cs[i], T[3 * r + 1], -sn[i] * T[3 * r]);
        R[3 * r + 2] = T[3 * r + 2];
      }
    }
    const float* g = d.pre[N];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      pe[r] = fmaf(R[3 * r], g[9], fmaf(R[3 * r + 1], g[10], fmaf(R[3 * r + 2], g[11], p[r])));
#pragma unroll
      for (int j = 0; j < 3; ++j)
        Re[3 * r + j] = fmaf(R[3 * r], g[j], fmaf(R[3 * r + 1], g[3 + j], R[3 * r + 2] * g[6 + j]));
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const float zx = Ja[i][0], zy = Ja[i][1], zz = Ja[i][2];
      const float dx = pe[0] - o[i][0], dy = pe[1] - o[i][1], dz = pe[2] - o[i][2];
      Jl[i][0] = zy * dz - zz * dy;
      Jl[i][1] = zz * dx - zx * dz;
      Jl[i][2] = zx * dy - zy * dx;
    }
  }
};

// ----------------------------------------------------------------------------
// tile movers: contiguous span of rows <-> shared memory
// ----------------------------------------------------------------------------
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}

__device__ __forceinline__ void tile_load(float* __restrict__ tile, const float* __restrict__ src,
                                          int nfloats, bool vec) {
  const int tid = threadIdx.x;
  if (vec) {
    const int nv = nfloats >> 2;
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* t4 = reinterpret_cast<float4*>(tile);
#pragma unroll 4
    for (int i = tid; i < nv; i += TPB) t4[i] = ldg_stream(s4 + i);
    const int done = nv << 2;
    if (tid < nfloats - done) tile[done + tid] = __ldg(src + done + tid);
  } else {
#pragma unroll 4
    for (int i = tid; i < nfloats; i += TPB) tile[i] = __ldg(src + i);
  }
}

__device__ __forceinline__ void tile_store(float* __restrict__ dst, const float* __restrict__ tile,
                                           int nfloats, bool vec) {
  const int tid = threadIdx.x;
  if (vec) {
    const int nv = nfloats >> 2;
    float4* d4 = reinterpret_cast<float4*>(dst);
    const float4* t4 = reinterpret_cast<const float4*>(tile);
#pragma unroll 4
    for (int i = tid; i < nv; i += TPB) d4[i] = t4[i];
    const int done = nv << 2;
    if (tid < nfloats - done) dst[done + tid] = tile[done + tid];
  } else {
#pragma unroll 4
    for (int i = tid; i < nfloats; i += TPB) dst[i] = tile[i];
  }
}

// coalesced store of the controlled columns only (colmask bit = column index)
__device__ __forceinline__ void tile_store_masked(float* __restrict__ dst,
                                                  const float* __restrict__ tile, int nfloats,
                                                  int stride, unsigned long long colmask) {
  const int tid = threadIdx.x;
  int col = tid % stride;
  const int step = TPB % stride;
  for (int i = tid; i < nfloats; i += TPB) {
    if ((colmask >> col) & 1ull) dst[i] = tile[i];
    col += step;
    if (col >= stride) col -= stride;
  }
}

__host__ __device__ __forceinline__ int round4(int x) { return (x + 3) & ~3; }

// ----------------------------------------------------------------------------
// action pre-processing: BaseController._preprocess_action (base_controller.py:120-133),
// PDEEPoseController._clip_and_scale_action (pd_ee_pose.py:197-210),
// gym_utils.clip_and_scale_action (gym_utils.py:102-105)
// ----------------------------------------------------------------------------
// ARM selects which EE controller's constants a function reads: 0 = P.cfg / P.mid / P.half / P.rank_rtol (every kernel),
// 1 = the second controller of a dual launch (P.cfg2 ...).  Compile-time, so ARM = 0 code is what it always was.
template <int ARM>
__device__ __forceinline__ const msik_step_cfg& arm_cfg(const StepParams& P) {
  if constexpr (ARM == 0) return P.cfg; else return P.cfg2;
}
template <int ARM>
__device__ __forceinline__ float arm_mid(const StepParams& P, int k) { if constexpr (ARM == 0) return P.mid[k]; else return P.mid2[k]; }
template <int ARM>
__device__ __forceinline__ float arm_half(const StepParams& P, int k) { if constexpr (ARM == 0) return P.half[k]; else return P.half2[k]; }
template <int ARM>
__device__ __forceinline__ float arm_rtol(const StepParams& P) { if constexpr (ARM == 0) return P.rank_rtol; else return P.rank_rtol2; }

template <int NC, int ARM = 0>
__device__ __forceinline__ void preprocess_action(const StepParams& P, const float (&a)[NC],
                                                  float (&b)[NC]) {
  if (arm_cfg<ARM>(P).flags & MSIK_F_NORMALIZE) {
#pragma unroll
    for (int k = 0; k < 3; ++k)
      b[k] = fmaf(arm_half<ARM>(P, k), fminf(fmaxf(a[k], -1.f), 1.f), arm_mid<ARM>(P, k));
    if (NC == 6) {
      const float rx = a[3], ry = a[4], rz = a[5];
      const float n2 = fmaf(rx, rx, fmaf(ry, ry, rz * rz));
      // rot * (1 / |rot|) where |rot| > 1, then * rot_lower (pd_ee_pose.py:205-209): one combined scale factor
      float inv = rsqrt_approx(fmaxf(n2, 1.f));
      inv = inv * fmaf(-0.5f * fmaxf(n2, 1.f), inv * inv, 1.5f);  // one Newton step; exactly 1 for n2 <= 1
      const float sc = arm_cfg<ARM>(P).rot_scale * inv;
      b[3] = rx * sc; b[4] = ry * sc; b[5] = rz * sc;
    }
  } else {
#pragma unroll
    for (int k = 0; k < NC; ++k) b[k] = a[k];
  }
}

template <class Chain>
__device__ __forceinline__ void joint_sincos(const StepParams& P, const float (&q)[Chain::N],
                                             float (&sn)[Chain::N], float (&cs)[Chain::N]) {
  if constexpr (!Chain::kGeneric) {  // joint types known at compile time: revolute joints two at a time (packed fp32x2)
    sincos_revolute_pairs<Chain::N, Chain::kRevoluteMask>(q, sn, cs);
    return;
  }
#pragma unroll
  for (int i = 0; i < Chain::N; ++i) {
    if (Chain::kGeneric) {
      if (P.chain.joint_type[i] == MSIK_JOINT_REVOLUTE) sincos_joint(q[i], sn[i], cs[i]);
      else { sn[i] = 0.f; cs[i] = 1.f; }
    } else if ((Chain::kRevoluteMask >> i) & 1u) {
      sincos_joint(q[i], sn[i], cs[i]);
    } else {
      sn[i] = 0.f; cs[i] = 1.f;
    }
  }
}

// J^T rows of the controlled joint range [C0, C0+NR) -> A, uncontrolled rows zeroed
template <class Chain, int C0, int NR, int NC, bool MASKED>
__device__ __forceinline__ void build_At(const StepParams& P, const float (&Jl)[Chain::N][3],
                                         const float (&Ja)[Chain::N][3], float (&A)[NR][NC]) {
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const bool on = !MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      A[r][k] = on ? Jl[C0 + r][k] : 0.f;
      if (NC == 6) A[r][3 + k] = on ? Ja[C0 + r][k] : 0.f;
    }
  }
}

// shared-memory carve-up of a CTA (non-gather mode)
struct Tiles {
  float* q;
  float* a;
  float* t;
  float* x;  // target_ik_kernel: exchange buffer of the lane compaction (NULL when the launch has none)
};
// words per env of that exchange buffer: q[N], tp[3], Rt[9], err_min, epos2, erot2, packed flags, iters
#define IK_XWORDS(N) ((N) + 17)
// the same record at an odd stride (bank-conflict free) + one sort key per thread in front: the regrouping buffer
#define IK_SORT_STRIDE(N) (((N) + 17) | 1)
#define IK_SORT_WORDS(N) (TPB + 4 + TPB * IK_SORT_STRIDE(N))
#ifndef MSIK_IK_COMPACTION
#define MSIK_IK_COMPACTION 0  // lane compaction of target_ik_kernel: built, measured slower, compiled out (see the kernel)
#endif
__device__ __forceinline__ Tiles carve(float* smem, const msik_step_cfg& c) {
  Tiles t;
  t.q = smem;
  t.a = t.q + round4(TPB * c.qpos_stride);
  t.t = t.a + round4(TPB * c.action_stride);
  t.x = nullptr;
  return t;
}

// common input stage: q (all chain joints) and raw action slice of this thread's env
template <class Chain, int NA, bool GATHER>
__device__ __forceinline__ void load_env(const StepParams& P, const Tiles& T, int env0, int nrows,
                                         bool need_old_target, float (&q)[Chain::N], float (&a)[NA],
                                         int& row) {
  const int tid = threadIdx.x;
  const int qs = P.cfg.qpos_stride, as = P.cfg.action_stride, ts = P.cfg.target_stride;
  const bool active = tid < nrows;
  row = env0 + tid;
  if (!GATHER) {
    tile_load(T.q, P.qpos + (size_t)env0 * qs, nrows * qs, P.vec_ok);
    if (P.action) tile_load(T.a, P.action + (size_t)env0 * as, nrows * as, P.vec_ok);
    if (need_old_target) tile_load(T.t, P.drive + (size_t)env0 * ts, nrows * ts, P.vec_ok);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < Chain::N; ++i) q[i] = active ? T.q[tid * qs + P.cfg.anc_col[i]] : 0.f;
#pragma unroll
    for (int k = 0; k < NA; ++k) a[k] = (active && P.action) ? T.a[tid * as + P.cfg.action_offset + k] : 0.f;
  } else {
    if (active && P.row_index) row = P.row_index[env0 + tid];
#pragma unroll
    for (int i = 0; i < Chain::N; ++i)
      q[i] = active ? __ldg(P.qpos + (size_t)row * qs + P.cfg.anc_col[i]) : 0.f;
#pragma unroll
    for (int k = 0; k < NA; ++k)
      a[k] = (active && P.action) ? __ldg(P.action + (size_t)(env0 + tid) * as + P.cfg.action_offset + k) : 0.f;
  }
}

// common output stage: new joint targets of the controlled range -> drive buffer (+ dense copies)
template <class Chain, int C0, int NR, bool MASKED, bool GATHER>
__device__ __forceinline__ void store_env(const StepParams& P, const Tiles& T, int env0, int nrows,
                                          int row, const float (&q)[Chain::N],
                                          const float (&qn)[NR]) {
  const int tid = threadIdx.x;
  const int qs = P.cfg.qpos_stride, ts = P.cfg.target_stride;
  const bool active = tid < nrows;
  const int env = env0 + tid;
  if (active && (P.tq_out || P.sq_out)) {
    int j = 0;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      const bool on = !MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u);
      if (on) {
        if (P.tq_out) P.tq_out[(size_t)env * P.n_ctrl + j] = qn[r];
        if (P.sq_out) P.sq_out[(size_t)env * P.n_ctrl + j] = q[C0 + r];
        ++j;
      }
    }
  }
  if (P.cfg.flags & MSIK_F_NO_DRIVE) return;
  if (!GATHER) {
    // deposit into the staging tile: the merged old-target tile, or (masked-store mode) this
    // thread's own qpos row, which nobody else reads
    float* stage = P.merge ? T.t : T.q;
    const int ss = P.merge ? ts : qs;
    if (active) {
#pragma unroll
      for (int r = 0; r < NR; ++r) {
        const bool on = !MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u);
        if (on) stage[tid * ss + P.cfg.anc_col[C0 + r]] = qn[r];
      }
    }
    __syncthreads();
    if (P.merge) tile_store(P.drive + (size_t)env0 * ts, stage, nrows * ts, P.vec_ok);
    else tile_store_masked(P.drive + (size_t)env0 * ts, stage, nrows * ts, ts, P.colmask);
  } else if (active) {
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      const bool on = !MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u);
      if (on) P.drive[(size_t)row * ts + P.cfg.anc_col[C0 + r]] = qn[r];
    }
  }
}

// One env of path A: pre-processed action -> FK -> Jacobian -> min-norm solve -> q0 + dq (kinematics.py:156-172).
// Generated chains whose stock controlled range this instantiation serves take the reduced solve
// (chaingen.py gen_reduced, msik_math.cuh reduced_min_norm); -DMSIK_NO_REDUCED keeps the dense one for A/B runs.
template <class Chain, int C0, int NR, int NC, bool MASKED, int ARM = 0>
__device__ __forceinline__ void solve_env(const StepParams& P, const float (&q)[Chain::N], const float (&a)[NC],
                                          float (&qn)[NR]) {
  static_assert(ARM == 0 || (!MASKED && !Chain::kGeneric), "the second arm of a dual launch is a generated chain on its stock range");
  constexpr int N = Chain::N;
  float b[NC];
  preprocess_action<NC, ARM>(P, a, b);
  float sn[N], cs[N];
#ifdef MSIK_DEBUG_STAGES
  if (P.debug & 4) {  // timing experiment: no sincos
#pragma unroll
    for (int i = 0; i < N; ++i) { sn[i] = q[i]; cs[i] = 1.f - q[i]; }
  } else
#endif
  joint_sincos<Chain>(P, q, sn, cs);
  float dq[NR];
#ifndef MSIK_NO_REDUCED
  if constexpr (Chain::kHasReduced && NC == 6 && !MASKED && C0 == Chain::kRedC0 && NR == Chain::kRedNR) {
    constexpr int NU = Chain::kRedNU;
    float At[NU][4], Bm[2][NU], a1[4], a2[2], u[NU], v[2];
    Chain::reduced(q, sn, cs, b, At, Bm, a1, a2);
#ifdef MSIK_DEBUG_STAGES
    if (P.debug & 2) {  // timing experiment: a few FMAs instead of the reduced solve
#pragma unroll
      for (int i = 0; i < NU; ++i)
        u[i] = fmaf(At[i][0], a1[0], fmaf(At[i][1], a1[1], fmaf(At[i][2], a1[2], fmaf(At[i][3], a1[3], Bm[0][i] * a2[0] + Bm[1][i] * a2[1]))));
      v[0] = a2[0]; v[1] = a2[1];
    } else
#endif
    reduced_min_norm<NU, Chain::kRedSkip, Chain::kRedZB>(At, Bm, a1, a2, u, v, arm_rtol<ARM>(P));
    Chain::reduced_scatter(u, v, dq);
  } else if constexpr (Chain::kHasLinear && NC == 3 && !MASKED && C0 == Chain::kLinC0 && NR == Chain::kLinNR) {
    // pd_ee_delta_pos: linear Jacobian generated in the cheapest joint frame (chaingen.py gen_linear), dense NR x 3 solve
    float A[NR][3], a3[3];
    Chain::linear(q, sn, cs, b, A, a3);
    min_norm_solve<NR, 3, Chain::kLinLastZero ? 3 : 0>(A, a3, dq, arm_rtol<ARM>(P));
  } else
#endif
  {
    float Jl[N][3], Ja[N][3], pe[3], Re[9];
    Chain::fkjac(P.chain, q, sn, cs, Jl, Ja, pe, Re);
    float A[NR][NC];
    build_At<Chain, C0, NR, NC, MASKED>(P, Jl, Ja, A);
#ifdef MSIK_DEBUG_STAGES
    if (P.debug & 2) {  // timing experiment: J^T b instead of the min-norm solve
#pragma unroll
      for (int r = 0; r < NR; ++r) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < NC; ++k) acc = fmaf(A[r][k], b[k], acc);
        dq[r] = acc;
      }
    } else
#endif
    min_norm_solve<NR, NC, (Chain::kLastLinZero && C0 + NR == Chain::N) ? 3 : 0>(A, b, dq, arm_rtol<ARM>(P));
  }
#pragma unroll
  for (int r = 0; r < NR; ++r) qn[r] = q[C0 + r] + dq[r];
}

// ----------------------------------------------------------------------------
// Path A: q0 + pinv(J(q0)) @ action   (kinematics.py:156-172), fused with the
// action pre-processing and the drive-target scatter.
// ----------------------------------------------------------------------------
template <class Chain, int C0, int NR, int NC, bool MASKED, bool GATHER>
__global__ void __launch_bounds__(TPB)
delta_step_kernel(const __grid_constant__ StepParams P) {
  constexpr int N = Chain::N;
  extern __shared__ __align__(16) float smem[];
  const int env0 = blockIdx.x * TPB;
  const int nrows = min(TPB, P.cfg.n_envs - env0);
  const Tiles T = carve(smem, P.cfg);

  float q[N], a[NC];
  int row;
  load_env<Chain, NC, GATHER>(P, T, env0, nrows, P.merge && !(P.cfg.flags & MSIK_F_NO_DRIVE), q, a, row);

  float qn[NR];
  solve_env<Chain, C0, NR, NC, MASKED>(P, q, a, qn);

  store_env<Chain, C0, NR, MASKED, GATHER>(P, T, env0, nrows, row, q, qn);
}

// ----------------------------------------------------------------------------
// target pose: PDEEPosController.compute_target_pose (pd_ee_pose.py:76-89) and
// PDEEPoseController.compute_target_pose (pd_ee_pose.py:212-238)
// b = pre-processed action; prev = previous pose (raw 7) when use_delta.
// ----------------------------------------------------------------------------
template <int NC>
__device__ __forceinline__ void compose_target(int flags, const float (&b)[NC], const float* prev,
                                               float (&tp)[3], Quat& tq) {
  const bool use_delta = flags & MSIK_F_USE_DELTA;
  Quat dq{1.f, 0.f, 0.f, 0.f};
  if (NC == 6) {
    float m[9];
    euler_xyz_to_mat(b[3], b[4], b[5], m);
    dq = mat_to_quat(m);
  }
  if (!use_delta) {
    tp[0] = b[0]; tp[1] = b[1]; tp[2] = b[2];
    tq = dq;
    return;
  }
  const Quat pq{prev[3], prev[4], prev[5], prev[6]};
  const float d[3] = {b[0], b[1], b[2]};
  if (NC == 6) {
    tq = (flags & MSIK_F_BODY_ROT) ? quat_mul(pq, dq) : quat_mul(dq, pq);
    if (flags & MSIK_F_BODY_TRANS) {
      float r[3];
      quat_apply(pq, d, r);
      tp[0] = prev[0] + r[0]; tp[1] = prev[1] + r[1]; tp[2] = prev[2] + r[2];
    } else {
      tp[0] = prev[0] + d[0]; tp[1] = prev[1] + d[1]; tp[2] = prev[2] + d[2];
    }
  } else {
    // Pose.__mul__ with an identity-rotation delta (pose.py:187-201)
    if (flags & MSIK_F_BODY_TRANS) {
      float r[3];
      quat_apply(pq, d, r);
      tp[0] = prev[0] + r[0]; tp[1] = prev[1] + r[1]; tp[2] = prev[2] + r[2];
    } else {
      // delta * prev: p = d + apply(identity, prev.p)
      tp[0] = d[0] + prev[0]; tp[1] = d[1] + prev[1]; tp[2] = d[2] + prev[2];
    }
    tq = quat_standardize(pq);  // quaternion_multiply standardises
  }
}

// ----------------------------------------------------------------------------
// Path B: iterative damped-least-squares IK to an absolute target pose, the whole
// <= max_iterations loop in one launch, per-env convergence, no host sync
// (replaces pk.PseudoInverseIK.solve called at kinematics.py:144-155).  The per-env
// arithmetic is msik_ik.cuh; this kernel is its warp-level control:
//   main loop : every env that is still being solved takes the fast (undamped, reduced) step; an env whose
//               smallest pivot drops under the guard parks (its damping is no longer below fp32 resolution);
//   tail loop : parked envs continue with the exact damped step (augmented QR).  Warps without parked envs
//               skip it; with damping == 0 nothing ever parks.
// P.na = action dims (3 or 6); the solve is always 6-d (pos_only is ignored by the reference on this path,
// kinematics.py:144-155).  P.gather: per-lane row access instead of CTA tiles (Articulation._data_index).
// ----------------------------------------------------------------------------
#ifndef MSIK_IK_MIN_CTAS
#define MSIK_IK_MIN_CTAS 4
#endif
template <class Chain, int C0, int NR, bool MASKED, bool SORT = false>
__global__ void __launch_bounds__(TPB, MSIK_IK_MIN_CTAS)
target_ik_kernel(const __grid_constant__ StepParams P) {
  constexpr int N = Chain::N;
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x;
  const int env0 = blockIdx.x * TPB;
  const int nrows = min(TPB, P.cfg.n_envs - env0);
  const bool active = tid < nrows;
  const int env = env0 + tid;
  Tiles T = carve(smem, P.cfg);
  if (P.ik_xoff >= 0) T.x = smem + P.ik_xoff;
  const int qs = P.cfg.qpos_stride, as = P.cfg.action_stride, ts = P.cfg.target_stride;
  const bool gather = P.gather != 0;
  const bool drive = !(P.cfg.flags & MSIK_F_NO_DRIVE);
  const int na = P.na;

  float q[N], a[6];
  int row = env;
  if (!gather) {
    tile_load(T.q, P.qpos + (size_t)env0 * qs, nrows * qs, P.vec_ok);
    if (P.action) tile_load(T.a, P.action + (size_t)env0 * as, nrows * as, P.vec_ok);
    if (P.merge && drive) tile_load(T.t, P.drive + (size_t)env0 * ts, nrows * ts, P.vec_ok);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < N; ++i) q[i] = active ? T.q[tid * qs + P.cfg.anc_col[i]] : 0.f;
#pragma unroll
    for (int k = 0; k < 6; ++k) a[k] = (active && P.action && k < na) ? T.a[tid * as + P.cfg.action_offset + k] : 0.f;
  } else {
    if (active && P.row_index) row = P.row_index[env];
#pragma unroll
    for (int i = 0; i < N; ++i) q[i] = active ? __ldg(P.qpos + (size_t)row * qs + P.cfg.anc_col[i]) : 0.f;
#pragma unroll
    for (int k = 0; k < 6; ++k)
      a[k] = (active && P.action && k < na) ? __ldg(P.action + (size_t)env * as + P.cfg.action_offset + k) : 0.f;
  }
  if (active && P.sq_out) {  // controller._start_qpos (pd_ee_pose.py:95)
    int j = 0;
#pragma unroll
    for (int r = 0; r < NR; ++r)
      if (!MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u)) P.sq_out[(size_t)env * P.n_ctrl + j++] = q[C0 + r];
  }

  // target pose: compute_target_pose (pd_ee_pose.py:76-89,212-238) or given (compute_ik called directly)
  float tp[3];
  Quat tq;
  {
    float prev[7] = {0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f};
    if (active && (P.cfg.flags & (MSIK_F_USE_DELTA | MSIK_F_TARGET_GIVEN))) {
#pragma unroll
      for (int k = 0; k < 7; ++k) prev[k] = P.prev_pose[(size_t)env * 7 + k];
    }
    if (P.cfg.flags & MSIK_F_TARGET_GIVEN) {
      tp[0] = prev[0]; tp[1] = prev[1]; tp[2] = prev[2];
      tq = Quat{prev[3], prev[4], prev[5], prev[6]};
    } else if (na == 6) {
      float b[6];
      preprocess_action<6>(P, a, b);
      compose_target<6>(P.cfg.flags, b, prev, tp, tq);
    } else {
      const float a3[3] = {a[0], a[1], a[2]};
      float b[3];
      preprocess_action<3>(P, a3, b);
      compose_target<3>(P.cfg.flags, b, prev, tp, tq);
    }
  }
  if (active && P.pose_out) {
    float* o = P.pose_out + (size_t)env * 7;
    o[0] = tp[0]; o[1] = tp[1]; o[2] = tp[2];
    o[3] = tq.w; o[4] = tq.x; o[5] = tq.y; o[6] = tq.z;
  }
  // pk converts the target through a rotation matrix (Transform3d): unit norm
  float Rt[9];
  {
    const float n2 = tq.w * tq.w + tq.x * tq.x + tq.y * tq.y + tq.z * tq.z;
    const float inv = rcp_nr(sqrt_nr(n2));
    quat_to_mat(Quat{tq.w * inv, tq.x * inv, tq.y * inv, tq.z * inv}, Rt);
  }

  IkLoop s;
  ik_loop_init(s, active);
  bool parked = false;
  // Lane compaction (-DMSIK_IK_COMPACTION=1, off in the shipped build: measured slower, profiles/r02e_ik_variants.txt -- the
  // compiled-in but disabled code alone cost 2 us per 65 536 envs in registers and per-iteration tests).  An env keeps its
  // identity (slot = its thread index at load time) while it may MOVE between lanes: every kCompactEvery iterations the CTA
  // counts the envs still being solved and, when at least one whole warp's worth has finished since the last time, packs the
  // live ones into the lowest lanes through shared memory (finished envs write their result first).  Iteration counts differ
  // per env (Panda near rest: 12..26, mean 21.6), so without this a warp runs to its slowest lane and ~14 % of the issued
  // lane-iterations are idle.
#if MSIK_IK_COMPACTION
  int slot = tid;
  bool valid = active;          // this thread holds an env whose result has not been written yet
  int warps_used = (nrows + 31) >> 5;
  float* xbuf = T.x;            // TPB x IK_XWORDS exchange buffer (non-NULL when the launch provides it)
  __shared__ int wcount[TPB / 32];
#else
  int slot = tid;       // the env this thread holds (row of the CTA's tile); changes once when the CTA regroups (below)
  bool valid = active;  // ... and whether it is a real env
#endif

  auto emit = [&]() {
    const int e = env0 + slot;
    if (P.iters_out) P.iters_out[e] = parked ? -s.iters : s.iters;  // negative: finished by the damped step
    if (P.err_out) {
      P.err_out[(size_t)e * 2] = sqrt_nr(s.epos2);
#ifdef MSIK_IK_SORT_DEBUG
      P.err_out[(size_t)e * 2 + 1] = (float)tid;  // which thread finished this env
#else
      P.err_out[(size_t)e * 2 + 1] = sqrt_nr(s.erot2);
#endif
    }
    if (P.tq_out) {
      int j = 0;
#pragma unroll
      for (int r = 0; r < NR; ++r)
        if (!MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u)) P.tq_out[(size_t)e * P.n_ctrl + j++] = q[C0 + r];
    }
    if (!drive) return;
    if (!gather) {
      float* stage = P.merge ? T.t : T.q;  // merged old-target tile, or the env's own qpos row (masked-store mode)
      const int ss = P.merge ? ts : qs;
#pragma unroll
      for (int r = 0; r < NR; ++r)
        if (!MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u)) stage[slot * ss + P.cfg.anc_col[C0 + r]] = q[C0 + r];
    } else {
      const int rw = P.row_index ? P.row_index[e] : e;
#pragma unroll
      for (int r = 0; r < NR; ++r)
        if (!MASKED || ((P.cfg.ctrl_mask >> (C0 + r)) & 1u)) P.drive[(size_t)rw * ts + P.cfg.anc_col[C0 + r]] = q[C0 + r];
    }
  };

  // Regrouping by remaining work (P.ik_sort; one exchange through shared memory, after iteration 0).  A warp runs until its
  // slowest lane is done (Panda near rest: 12..26 iterations per env, mean 21.6, warp maxima 24..27) and the iterations an env
  // still needs follow from its error after the first step (the error shrinks by the step size every iteration), so the CTA
  // ranks its envs by that error once and deals them out in rank order: each warp then holds envs that finish together, and
  // the warps of the CTA need ~20 / 22 / 23 / 25.4 iterations instead of ~25 each (-9 % warp-iterations).  The ranking is a
  // counting sort on the predicted number of iterations (one shared-memory atomic per env); an all-pairs rank cost as many
  // instructions as the regrouping saves.  Per-env arithmetic is untouched: results are bit-identical, only the lane an
  // env runs in changes (`slot` keeps its row).
  float* const sortbuf = (MSIK_IK_COMPACTION == 0 && SORT) ? T.x : nullptr;  // SORT = false: all of this folds away
  for (int it = 0; it < P.it.max_iterations; ++it) {
    if (SORT && sortbuf != nullptr && it == 1) {  // CTA-uniform; every warp gets here (no warp leaves the loop at it == 0, below)
      constexpr int XS = IK_SORT_STRIDE(N);
      constexpr int NB = 64;  // buckets = predicted iterations still to go (clamped)
      int* hist = reinterpret_cast<int*>(sortbuf);  // [NB] counts, then [NB] exclusive prefix sums
      float* rec = sortbuf + TPB + 4;
      if (tid < NB) hist[tid] = 0;
      __syncthreads();
      // the error shrinks by (1 - lr) per iteration, so an env needs log(err / tol) / -log(1 - lr) more of them: that number
      // (the larger of the position's and the rotation's) is its bucket.  Finished / idle / parked lanes: bucket 0.
      const float pt2 = P.it.pos_tol * P.it.pos_tol, rt2 = P.it.rot_tol * P.it.rot_tol;
      const float ratio2 = fmaxf(s.epos2 * rcp_approx(pt2 + 1e-37f), s.erot2 * rcp_approx(rt2 + 1e-37f));  // (err / tol)^2
      const float lr = P.it.lr > 0.f && P.it.lr < 1.f ? P.it.lr : 0.2f;
      const float per_log2 = -0.5f * rcp_approx(__log2f(1.f - lr));  // iterations per log2 of the squared ratio
      int bkt = 0;
      if (s.remaining && !parked) bkt = min(NB - 1, max(1, 1 + (int)(__log2f(fmaxf(ratio2, 1.f)) * per_log2)));
      const int pos = atomicAdd(&hist[bkt], 1);  // place inside the bucket (any order: the envs of a bucket finish together)
      __syncthreads();
      if (tid < 32) {  // exclusive prefix sums of the 64 counts, two per lane
        const int v0 = hist[2 * tid], v1 = hist[2 * tid + 1];
        int incl = v0 + v1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int up = __shfl_up_sync(0xffffffffu, incl, d);
          if (tid >= d) incl += up;
        }
        const int excl = incl - (v0 + v1);
        hist[NB + 2 * tid] = excl;
        hist[NB + 2 * tid + 1] = excl + v0;
      }
      __syncthreads();
      const int rank = hist[NB + bkt] + pos;
      // rank order straight onto the threads: warp w gets ranks 32 w .. 32 w + 31.  No rotation between CTAs is needed (one
      // by a per-SM counter was measured: 53 vs 49 us) -- the hardware already spreads the warps of co-resident CTAs over
      // the sub-partitions differently.
      const int dest = rank;
      {
        float* x = rec + dest * XS;
#pragma unroll
        for (int i = 0; i < N; ++i) x[i] = q[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) x[N + i] = tp[i];
#pragma unroll
        for (int i = 0; i < 9; ++i) x[N + 3 + i] = Rt[i];
        x[N + 12] = s.err_min; x[N + 13] = s.epos2; x[N + 14] = s.erot2;
        x[N + 15] = __int_as_float(slot | (parked ? 0x10000 : 0) | (s.remaining ? 0x20000 : 0) | (valid ? 0x40000 : 0) | (s.noimp << 19));
        x[N + 16] = __int_as_float(s.iters);
      }
      __syncthreads();
      {
        const float* x = rec + tid * XS;
#pragma unroll
        for (int i = 0; i < N; ++i) q[i] = x[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) tp[i] = x[N + i];
#pragma unroll
        for (int i = 0; i < 9; ++i) Rt[i] = x[N + 3 + i];
        s.err_min = x[N + 12]; s.epos2 = x[N + 13]; s.erot2 = x[N + 14];
        const int pk = __float_as_int(x[N + 15]);
        slot = pk & 0xffff; parked = (pk & 0x10000) != 0; s.remaining = (pk & 0x20000) != 0; valid = (pk & 0x40000) != 0;
        s.noimp = pk >> 19;
        s.iters = __float_as_int(x[N + 16]);
      }
    }
#if MSIK_IK_COMPACTION
    constexpr int kCompactEvery = 2;
    if (xbuf != nullptr && (it % kCompactEvery) == 0 && it >= P.ik_compact_from) {  // CTA-uniform
      const int n_run = __syncthreads_count(s.remaining && !parked);
      if (n_run == 0) break;
      if (n_run <= (warps_used - 1) * 32) {
        const bool live = valid && s.remaining;  // running or parked: still needs iterations
        if (valid && !live) { emit(); valid = false; }
        const unsigned bal = __ballot_sync(0xffffffffu, live);
        if ((tid & 31) == 0) wcount[tid >> 5] = __popc(bal);
        __syncthreads();
        int base = 0, total = 0;
#pragma unroll
        for (int w = 0; w < TPB / 32; ++w) {
          const int c = wcount[w];
          base += w < (tid >> 5) ? c : 0;
          total += c;
        }
        if (live) {
          float* x = xbuf + (base + __popc(bal & ((1u << (tid & 31)) - 1u))) * IK_XWORDS(N);
#pragma unroll
          for (int i = 0; i < N; ++i) x[i] = q[i];
#pragma unroll
          for (int i = 0; i < 3; ++i) x[N + i] = tp[i];
#pragma unroll
          for (int i = 0; i < 9; ++i) x[N + 3 + i] = Rt[i];
          x[N + 12] = s.err_min; x[N + 13] = s.epos2; x[N + 14] = s.erot2;
          x[N + 15] = __int_as_float(slot | (parked ? 0x10000 : 0) | (s.noimp << 17));
          x[N + 16] = __int_as_float(s.iters);
        }
        __syncthreads();
        valid = tid < total;
        if (valid) {
          const float* x = xbuf + tid * IK_XWORDS(N);
#pragma unroll
          for (int i = 0; i < N; ++i) q[i] = x[i];
#pragma unroll
          for (int i = 0; i < 3; ++i) tp[i] = x[N + i];
#pragma unroll
          for (int i = 0; i < 9; ++i) Rt[i] = x[N + 3 + i];
          s.err_min = x[N + 12]; s.epos2 = x[N + 13]; s.erot2 = x[N + 14];
          const int pk = __float_as_int(x[N + 15]);
          slot = pk & 0xffff; parked = (pk & 0x10000) != 0; s.noimp = pk >> 17;
          s.iters = __float_as_int(x[N + 16]);
          s.remaining = true;
        } else {
          s.remaining = false; parked = false;
        }
        warps_used = (total + 31) >> 5;
        __syncthreads();  // xbuf / wcount may be rewritten at the next checkpoint
      }
    }
#endif
    const bool run = s.remaining && !parked;
    if (!__any_sync(0xffffffffu, run)) {
#if MSIK_IK_COMPACTION
      if (xbuf != nullptr) continue;  // keep in step with the CTA's checkpoints
#endif
      if (SORT && sortbuf != nullptr && it == 0) continue;  // an idle warp still takes part in the regrouping (it may get work)
      break;
    }
    float dq[NR], ep2, er2, pmin;
    ik_iter_fast<Chain, C0, NR, MASKED>(P.chain, P.cfg.ctrl_mask, P.rank_rtol, q, tp, Rt, dq, ep2, er2, pmin);
    if (run) {
      if (pmin < P.ik_guard) parked = true;
      else ik_advance<Chain, C0, NR, MASKED>(s, P.it, P.cfg.ctrl_mask, q, dq, ep2, er2);
    }
  }
  if (__any_sync(0xffffffffu, parked)) {
    for (;;) {
      const bool run = parked && s.remaining && s.iters < P.it.max_iterations;
      if (!__any_sync(0xffffffffu, run)) break;
      float dq[NR], ep2, er2;
      ik_iter_dls<Chain, C0, NR, MASKED>(P.chain, P.cfg.ctrl_mask, P.ik_lambda, q, tp, Rt, dq, ep2, er2);
      if (run) ik_advance<Chain, C0, NR, MASKED>(s, P.it, P.cfg.ctrl_mask, q, dq, ep2, er2);
    }
  }
  if (valid) emit();
  if (!drive || gather) return;
  __syncthreads();
  {
    float* stage = P.merge ? T.t : T.q;
    if (P.merge) tile_store(P.drive + (size_t)env0 * ts, stage, nrows * ts, P.vec_ok);
    else tile_store_masked(P.drive + (size_t)env0 * ts, stage, nrows * ts, ts, P.colmask);
  }
}

// ----------------------------------------------------------------------------
// FK (+ optional Jacobian): pk chain.forward_kinematics / chain.jacobian (kinematics.py:157)
// ----------------------------------------------------------------------------
template <class Chain>
__global__ void __launch_bounds__(TPB) fk_kernel(const __grid_constant__ StepParams P, int n_real) {
  constexpr int N = Chain::N;
  const int env = blockIdx.x * TPB + threadIdx.x;
  if (env >= P.cfg.n_envs) return;
  const int row = P.row_index ? P.row_index[env] : env;
  float q[N];
#pragma unroll
  for (int i = 0; i < N; ++i) q[i] = __ldg(P.qpos + (size_t)row * P.cfg.qpos_stride + P.cfg.anc_col[i]);
  float sn[N], cs[N];
  joint_sincos<Chain>(P, q, sn, cs);
  float Jl[N][3], Ja[N][3], pe[3], Re[9];
  Chain::fkjac(P.chain, q, sn, cs, Jl, Ja, pe, Re);
  if (P.pose_out) {
    Quat qq = mat_to_quat(Re);
    if (P.prev_pose) {
      // world pose = root pose * FK (Pose.__mul__, pose.py:187-201); prev_pose carries the (n_envs,7) root poses
      const float* rp = P.prev_pose + (size_t)env * 7;
      const Quat rq{__ldg(rp + 3), __ldg(rp + 4), __ldg(rp + 5), __ldg(rp + 6)};
      float w[3];
      quat_apply(rq, pe, w);
      pe[0] = __ldg(rp + 0) + w[0]; pe[1] = __ldg(rp + 1) + w[1]; pe[2] = __ldg(rp + 2) + w[2];
      qq = quat_mul(rq, qq);
    }
    float* o = P.pose_out + (size_t)env * 7;
    o[0] = pe[0]; o[1] = pe[1]; o[2] = pe[2];
    o[3] = qq.w; o[4] = qq.x; o[5] = qq.y; o[6] = qq.z;
  }
  if (P.jac_out) {
    float* J = P.jac_out + (size_t)env * 6 * n_real;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      if (i < n_real) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          J[k * n_real + i] = Jl[i][k];
          J[(3 + k) * n_real + i] = Ja[i][k];
        }
      }
    }
  }
}

}  // namespace msik
