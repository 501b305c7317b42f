// Small controller-state kernels (reset / interpolation / pose bookkeeping).
// Included by msik.cu only (non-template __global__ functions: one definition).
#pragma once
#include "msik_kernels.cuh"

namespace msik {

// ----------------------------------------------------------------------------
// small controller-state kernels
// ----------------------------------------------------------------------------
// ee_pose_at_base = root.inv() * ee  (pd_ee_pose.py:61-64, pose.py:187-220)
__global__ void __launch_bounds__(TPB)
ee_pose_at_base_kernel(const float* __restrict__ body, int stride, const int32_t* __restrict__ root_idx,
                       const int32_t* __restrict__ ee_idx, int n, float* __restrict__ out) {
  const int e = blockIdx.x * TPB + threadIdx.x;
  if (e >= n) return;
  const float* r = body + (size_t)root_idx[e] * stride;
  const float* l = body + (size_t)ee_idx[e] * stride;
  const Quat rq{r[3], r[4], r[5], r[6]};
  const Quat ri = quat_conj(rq);
  const float np[3] = {-r[0], -r[1], -r[2]};
  float ip[3];
  quat_apply(ri, np, ip);  // inverse translation
  const Quat lq{l[3], l[4], l[5], l[6]};
  const Quat oq = quat_mul(ri, lq);
  const float lp[3] = {l[0], l[1], l[2]};
  float rp[3];
  quat_apply(ri, lp, rp);
  float* o = out + (size_t)e * 7;
  o[0] = ip[0] + rp[0]; o[1] = ip[1] + rp[1]; o[2] = ip[2] + rp[2];
  o[3] = oq.w; o[4] = oq.x; o[5] = oq.y; o[6] = oq.z;
}

// sibling joint-position controllers (msik_aux_cfg) when the path-A launch could not fuse them
__global__ void __launch_bounds__(TPB) aux_targets_kernel(const __grid_constant__ StepParams P) {
  const int e = blockIdx.x * TPB + threadIdx.x;
  if (e >= P.cfg.n_envs) return;
  const size_t row = P.row_index ? P.row_index[e] : e;
  const float* qrow = P.qpos + row * P.cfg.qpos_stride;
  const float* arow = P.action + (size_t)e * P.cfg.action_stride;
  float* trow = P.drive + row * P.cfg.target_stride;
  for (int c = 0; c < P.aux.n_cols; ++c) {
    float v = arow[P.aux.action_col[c]];
    const int f = P.aux.flags[c];
    if (f & MSIK_AUX_NORMALIZE) v = __fadd_rn(P.aux.mid[c], __fmul_rn(P.aux.half[c], fminf(fmaxf(v, -1.f), 1.f)));
    if (f & MSIK_AUX_DELTA) v = __fadd_rn(qrow[P.aux.col[c]], v);
    trow[P.aux.col[c]] = v;
  }
}

// pd_joint_vel.py:38-42, pd_base_vel.py:19-33,50-72 + articulation.py:743-761 (see include/msik.h msik_vel_targets)
__global__ void __launch_bounds__(TPB) vel_targets_kernel(const __grid_constant__ msik_vel_cfg c, const float* __restrict__ qpos,
                                                          const int32_t* __restrict__ row_index,
                                                          const float* __restrict__ action, float* __restrict__ qvel) {
  const int e = blockIdx.x * TPB + threadIdx.x;
  if (e >= c.n_envs) return;
  const size_t row = row_index ? row_index[e] : e;
  const float ori = c.ego == MSIK_VEL_JOINT ? 0.f : qpos[row * c.qpos_stride + c.ori_col];
  vel_apply(c, action + (size_t)e * c.action_stride + c.action_offset, ori, qvel + row * c.qvel_stride);
}

template <int NA>
__global__ void __launch_bounds__(TPB) target_pose_kernel(const __grid_constant__ StepParams P) {
  const int e = blockIdx.x * TPB + threadIdx.x;
  if (e >= P.cfg.n_envs) return;
  float a[NA], b[NA];
#pragma unroll
  for (int k = 0; k < NA; ++k) a[k] = P.action[(size_t)e * P.cfg.action_stride + P.cfg.action_offset + k];
  preprocess_action<NA>(P, a, b);
  float prev[7] = {0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f};
  if (P.cfg.flags & MSIK_F_USE_DELTA) {
#pragma unroll
    for (int k = 0; k < 7; ++k) prev[k] = P.prev_pose[(size_t)e * 7 + k];
  }
  float tp[3];
  Quat tq;
  compose_target<NA>(P.cfg.flags, b, prev, tp, tq);
  float* o = P.pose_out + (size_t)e * 7;
  o[0] = tp[0]; o[1] = tp[1]; o[2] = tp[2];
  o[3] = tq.w; o[4] = tq.x; o[5] = tq.y; o[6] = tq.z;
}

// trajectory/utils/actions/conversion.py:52-81,144-183: target / current EE pose -> normalised delta action
struct DeltaActionParams {
  float mid[6], inv_half[6];
  int32_t n, nd;
};
__global__ void __launch_bounds__(TPB)
pose_to_delta_action_kernel(const DeltaActionParams P, const float* __restrict__ target, const float* __restrict__ ee,
                            float* __restrict__ action, uint8_t* __restrict__ exact) {
  const int e = blockIdx.x * TPB + threadIdx.x;
  if (e >= P.n) return;
  const float* t = target + (size_t)e * 7;
  const float* c = ee + (size_t)e * 7;
  float a[6] = {t[0] - c[0], t[1] - c[1], t[2] - c[2], 0.f, 0.f, 0.f};
  if (P.nd == 6) {
    // sapien Pose.__mul__: plain Hamilton product, no sign standardisation
    const Quat d = quat_raw_mul(Quat{c[3], c[4], c[5], c[6]}, Quat{t[3], -t[4], -t[5], -t[6]});
    // transforms3d.quaternions.quat2axangle + the wrap of conversion.py:55-56
    const float nq = d.w * d.w + d.x * d.x + d.y * d.y + d.z * d.z;
    const float inv = nq > 0.f ? rsqrtf(nq) : 0.f;
    const float w = fminf(fmaxf(d.w * inv, -1.f), 1.f);
    const float x = d.x * inv, y = d.y * inv, z = d.z * inv;
    const float len2 = x * x + y * y + z * z;
    if (len2 >= 1.2789769e-13f) {  // (3 eps_fp32)^2: below it the rotation is the identity, axis (1,0,0), angle 0
      // 2 acos(w) of the unit quaternion, taken as 2 atan2(|v|, w): acos loses half the digits near w = 1 in fp32
      const float vn = sqrtf(len2);
      float ang = 2.f * atan2f(vn, w);
      if (ang > 3.14159265358979f) ang -= 6.28318530717959f;
      const float k = ang / vn;
      a[3] = x * k; a[4] = y * k; a[5] = z * k;
    }
  }
  bool ok = true;
#pragma unroll
  for (int k = 0; k < 6; ++k) a[k] = (a[k] - P.mid[k]) * P.inv_half[k];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if (fabsf(a[k]) > 1.f) ok = false;
    a[k] = fminf(fmaxf(a[k], -1.f), 1.f);
  }
  if (P.nd == 6) {
    const float n2 = a[3] * a[3] + a[4] * a[4] + a[5] * a[5];
    if (n2 > 1.f) {
      const float s = rsqrtf(n2);
      a[3] *= s; a[4] *= s; a[5] *= s;
      ok = false;
    }
  }
  for (int k = 0; k < P.nd; ++k) action[(size_t)e * P.nd + k] = a[k];
  if (exact) exact[e] = ok ? 1 : 0;
}

// before_simulation_step with interpolate (pd_joint_pos.py:92-98)
__global__ void __launch_bounds__(TPB)
interp_targets_kernel(const __grid_constant__ StepParams P, const float* __restrict__ start,
                      const float* __restrict__ step_size, float step) {
  const int idx = blockIdx.x * TPB + threadIdx.x;
  const int c = P.n_ctrl;
  if (idx >= P.cfg.n_envs * c) return;
  const int e = idx / c, j = idx - e * c;
  // j-th controlled chain joint -> column
  int col = 0, seen = 0;
  for (int i = 0; i < MSIK_MAX_JOINTS; ++i) {
    if ((P.cfg.ctrl_mask >> i) & 1u) {
      if (seen == j) { col = P.cfg.anc_col[i]; break; }
      ++seen;
    }
  }
  const int row = P.row_index ? P.row_index[e] : e;
  P.drive[(size_t)row * P.cfg.target_stride + col] = fmaf(step_size[idx], step, start[idx]);
}

// masked reset (pd_joint_pos.py:51-66, pd_ee_pose.py:66-74)
__global__ void __launch_bounds__(TPB)
reset_masked_kernel(const __grid_constant__ StepParams P, const uint8_t* __restrict__ mask,
                    const float* __restrict__ ee_pose, float* __restrict__ start_qpos,
                    float* __restrict__ target_qpos, float* __restrict__ target_pose) {
  const int e = blockIdx.x * TPB + threadIdx.x;
  if (e >= P.cfg.n_envs || !mask[e]) return;
  const int row = P.row_index ? P.row_index[e] : e;
  int j = 0;
  for (int i = 0; i < MSIK_MAX_JOINTS; ++i) {
    if ((P.cfg.ctrl_mask >> i) & 1u) {
      const float v = P.qpos[(size_t)row * P.cfg.qpos_stride + P.cfg.anc_col[i]];
      if (start_qpos) start_qpos[(size_t)e * P.n_ctrl + j] = v;
      if (target_qpos) target_qpos[(size_t)e * P.n_ctrl + j] = v;
      ++j;
    }
  }
  if (ee_pose && target_pose) {
#pragma unroll
    for (int k = 0; k < 7; ++k) target_pose[(size_t)e * 7 + k] = ee_pose[(size_t)e * 7 + k];
  }
}

}  // namespace msik
