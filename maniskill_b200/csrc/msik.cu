// C ABI of the EE-controller hot path (include/msik.h).  Host-side glue only:
// argument validation, chain matching, kernel-variant selection, launches.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "msik_launch.cuh"
#include "msik_state_kernels.cuh"
#include "msik_warp.cuh"

namespace msik {

std::atomic<long long> g_launches{0};

#define MSIK_DECL_OPS(name, n) extern const ChainOps kOps_##name;
MSIK_FOREACH_SPECIALIZED_CHAIN(MSIK_DECL_OPS)
extern const ChainOps kOps_generic7;
extern const ChainOps kOps_generic8;
extern const ChainOps kOps_generic12;
extern const ChainOps kOps_generic16;
extern const ChainOps kOps_generic6r;
extern const ChainOps kOps_generic7r;

#define MSIK_OPS_PTR(name, n) &kOps_##name,
static const ChainOps* const kSpecialised[] = {MSIK_FOREACH_SPECIALIZED_CHAIN(MSIK_OPS_PTR)};
static const ChainOps* const kGeneric[] = {&kOps_generic7, &kOps_generic8, &kOps_generic12, &kOps_generic16};

// chain specialisations compiled at run time (maniskill_b200/jit.py) and registered by msik_register_chain_ops
static std::mutex g_plugin_mu;
static std::vector<const ChainOps*> g_plugins;

static thread_local std::string t_err;

static int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  t_err = buf;
  return code;
}

static int cuda_fail(cudaError_t e, const char* what) {
  return fail(MSIK_ECUDA, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
}

// env count up to which msik_target_ik (mapping = auto) takes the lane-per-joint kernel; set from the measured crossover
constexpr int kIkWarpMaxEnvs = 0;

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int store_mode_merge() {
  static int mode = [] {
    const char* e = getenv("MSIK_STORE_MODE");
    return (e && !strcmp(e, "masked")) ? 0 : 1;
  }();
  return mode;
}

}  // namespace msik

using namespace msik;

struct msik_chain {
  msik_chain_desc desc;    // as given
  msik_chain_desc padded;  // padded to ops->n joints (type 0, identity pre), tool moved to [ops->n]
  const ChainOps* ops;
};

// Host-buffer pipeline: one stream per engine (host->device copies, kernels, device->host copies) so that
// each copy engine sees its transfers back to back, chunks rotate over kPipeSlots device buffers, and
// events carry the chunk-to-chunk dependencies between the three streams.
constexpr int kPipeSlots = 8;
struct msik_pipeline {
  int chunk, qs, as, nslots;
  bool trace;
  cudaEvent_t ev_t0;
  cudaStream_t s_in, s_k, s_out;
  cudaEvent_t ev_in[kPipeSlots];   // inputs of the slot are on the device
  cudaEvent_t ev_k[kPipeSlots];    // kernel of the slot is done (inputs may be overwritten, targets may be read)
  cudaEvent_t ev_out[kPipeSlots];  // targets of the slot are on the host (the target buffer may be overwritten)
  float* d_q[kPipeSlots];
  float* d_a[kPipeSlots];
  float* d_t[kPipeSlots];
};

extern "C" {

int msik_version(void) { return MSIK_VERSION; }
const char* msik_last_error(void) { return t_err.c_str(); }
int64_t msik_launch_count(void) { return (int64_t)g_launches.load(); }

int msik_device_check(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(MSIK_ENODEV, "no CUDA device: %s", cudaGetErrorString(e));
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceProperties");
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (prop.major != 10)
    return fail(MSIK_ENODEV, "device %s is sm_%d%d; this library is built for sm_100a only and has no fallback",
                prop.name, prop.major, prop.minor);
  return MSIK_OK;
}

int msik_chain_create(const msik_chain_desc* desc, msik_chain_t* out) {
  if (!desc || !out) return fail(MSIK_EINVAL, "msik_chain_create: NULL argument");
  const int n = desc->n_joints;
  if (n < 1 || n > MSIK_MAX_JOINTS) return fail(MSIK_EINVAL, "n_joints=%d out of range [1,%d]", n, MSIK_MAX_JOINTS);
  for (int i = 0; i < n; ++i)
    if (desc->joint_type[i] != MSIK_JOINT_REVOLUTE && desc->joint_type[i] != MSIK_JOINT_PRISMATIC)
      return fail(MSIK_EINVAL, "joint %d has unsupported type %d", i, desc->joint_type[i]);
  msik_chain* c = new (std::nothrow) msik_chain;
  if (!c) return fail(MSIK_ENOMEM, "out of host memory");
  c->desc = *desc;
  c->ops = nullptr;
  const bool no_spec = getenv("MSIK_FORCE_GENERIC") != nullptr;
  auto matches = [&](const ChainOps* ops) {
    if (no_spec || ops->n != n) return false;
    bool same = memcmp(ops->pre, &desc->pre[0][0], sizeof(float) * 12 * (n + 1)) == 0;
    for (int i = 0; same && i < n; ++i) same = ops->types[i] == desc->joint_type[i];
    return same;
  };
  for (const ChainOps* ops : kSpecialised)
    if (matches(ops)) { c->ops = ops; break; }
  if (!c->ops) {
    std::lock_guard<std::mutex> lk(g_plugin_mu);
    for (const ChainOps* ops : g_plugins)
      if (matches(ops)) { c->ops = ops; break; }
  }
  if (!c->ops && !getenv("MSIK_NO_GENERIC_REV")) {  // all-revolute 6- / 7-axis arm: the leaner table-driven kernels
    bool allrev = true;
    for (int i = 0; i < n; ++i) allrev = allrev && desc->joint_type[i] == MSIK_JOINT_REVOLUTE;
    if (allrev && n == 6) c->ops = &kOps_generic6r;
    if (allrev && n == 7) c->ops = &kOps_generic7r;
  }
  if (!c->ops) {
    for (const ChainOps* ops : kGeneric)
      if (ops->n >= n) { c->ops = ops; break; }
  }
  // padded table for the generic kernels
  const int N = c->ops->n;
  memset(&c->padded, 0, sizeof c->padded);
  c->padded.n_joints = N;
  for (int i = 0; i < N; ++i) {
    if (i < n) {
      c->padded.joint_type[i] = desc->joint_type[i];
      memcpy(c->padded.pre[i], desc->pre[i], sizeof(float) * 12);
      c->padded.lower[i] = desc->lower[i];
      c->padded.upper[i] = desc->upper[i];
    } else {
      c->padded.pre[i][0] = c->padded.pre[i][4] = c->padded.pre[i][8] = 1.f;
    }
  }
  memcpy(c->padded.pre[N], desc->pre[n], sizeof(float) * 12);
  *out = c;
  return MSIK_OK;
}

int msik_register_chain_ops(const void* ops_) {
  const ChainOps* ops = static_cast<const ChainOps*>(ops_);
  if (!ops || !ops->name || ops->generic || !ops->pre || !ops->types || ops->n < 1 || ops->n > MSIK_MAX_JOINTS || !ops->delta ||
      !ops->target_ik || !ops->delta_tma || !ops->fk)
    return fail(MSIK_EINVAL, "not a chain-specialisation table of this library version");
  std::lock_guard<std::mutex> lk(g_plugin_mu);
  for (const ChainOps* o : g_plugins)
    if (o == ops) return MSIK_OK;
  g_plugins.push_back(ops);
  return MSIK_OK;
}

int msik_chain_destroy(msik_chain_t chain) {
  delete chain;
  return MSIK_OK;
}

int msik_chain_kernel_name(msik_chain_t chain, char* buf, int cap) {
  if (!chain || !buf || cap < 1) return fail(MSIK_EINVAL, "msik_chain_kernel_name: bad argument");
  snprintf(buf, cap, "%s", chain->ops->name);
  return MSIK_OK;
}

}  // extern "C"

namespace msik {

// fills the launch parameters shared by every kernel; returns the ctrl mode
static int prepare(msik_chain_t chain, const msik_step_cfg* cfg, StepParams& P, int* mode, int* ndim,
                   bool need_action) {
  if (!chain || !cfg) return fail(MSIK_EINVAL, "NULL chain / cfg");
  const int n = chain->desc.n_joints;
  if (cfg->n_envs < 0) return fail(MSIK_EINVAL, "n_envs=%d", cfg->n_envs);
  if (cfg->qpos_stride < 1 || cfg->target_stride < 1) return fail(MSIK_EINVAL, "bad row stride");
  const uint32_t all = n >= 32 ? 0xffffffffu : ((1u << n) - 1u);
  if (cfg->ctrl_mask == 0 || (cfg->ctrl_mask & ~all))
    return fail(MSIK_EINVAL, "ctrl_mask=0x%x does not fit a chain of %d joints", cfg->ctrl_mask, n);
  unsigned long long colmask = 0;
  for (int i = 0; i < n; ++i) {
    const int c = cfg->anc_col[i];
    if (c < 0 || c >= cfg->qpos_stride || c >= cfg->target_stride)
      return fail(MSIK_EINVAL, "anc_col[%d]=%d outside row stride", i, c);
    if ((cfg->ctrl_mask >> i) & 1u && c < 64) colmask |= 1ull << c;
  }
  const int nd = (cfg->flags & MSIK_F_POS_ONLY) ? 3 : 6;
  if (need_action) {
    if (cfg->action_offset < 0 || cfg->action_offset + nd > cfg->action_stride)
      return fail(MSIK_EINVAL, "action slice [%d,%d) outside action stride %d", cfg->action_offset,
                  cfg->action_offset + nd, cfg->action_stride);
  }
  memset(&P, 0, sizeof P);
  P.chain = chain->padded;
  P.cfg = *cfg;
  for (int i = n; i < MSIK_MAX_JOINTS; ++i) P.cfg.anc_col[i] = cfg->anc_col[n - 1];  // padding joints read a valid column
  for (int k = 0; k < 3; ++k) {
    P.mid[k] = 0.5f * (cfg->pos_high[k] + cfg->pos_low[k]);
    P.half[k] = 0.5f * (cfg->pos_high[k] - cfg->pos_low[k]);
  }
  P.colmask = colmask;
  P.n_ctrl = __builtin_popcount(cfg->ctrl_mask);
  // torch.linalg.pinv's default rtol = max(rows, cols) * eps_fp32 of the (nd x n_ctrl) Jacobian it is handed
  P.rank_rtol = (float)(nd > P.n_ctrl ? nd : P.n_ctrl) * 1.1920929e-07f;
  if (mode) {
    if (cfg->ctrl_mask == all) *mode = CTRL_FULL;
    else if (chain->ops->range_mask && cfg->ctrl_mask == chain->ops->range_mask) *mode = CTRL_RANGE;
    else *mode = CTRL_MASKED;
    if (chain->ops->n != n) *mode = CTRL_MASKED;  // padded bucket: padding joints must stay masked
  }
  if (ndim) *ndim = nd;
  return MSIK_OK;
}

// tiled (shared-memory staged) or gather I/O; fills smem size, vec_ok, merge
static bool choose_tiled(StepParams& P, const int32_t* row_index, size_t* smem) {
  const msik_step_cfg& c = P.cfg;
  const bool drive = !(c.flags & MSIK_F_NO_DRIVE);
  size_t fl = (size_t)round4(TPB * c.qpos_stride) + round4(TPB * c.action_stride) +
              (drive ? round4(TPB * c.target_stride) : 0);
  *smem = fl * sizeof(float);
  if (row_index || *smem > 160 * 1024) return false;
  P.vec_ok = aligned16(P.qpos) && aligned16(P.action) && (!drive || aligned16(P.drive)) &&
             (c.action_stride > 0 || !P.action);
  P.merge = store_mode_merge();
  if (!P.merge && (c.target_stride > 64 || c.target_stride != c.qpos_stride)) P.merge = 1;
  if (!drive) P.merge = 0;
  return true;
}

// persistent TMA-staged kernel: identity rows and every tile base 16-byte aligned
// (tile sizes are TPB * stride * 4 bytes, always a multiple of 16)
static bool choose_tma(StepParams& P, const int32_t* row_index, size_t* smem) {
  static const bool off = [] {
    const char* e = getenv("MSIK_DELTA_KERNEL");
    return e && !strcmp(e, "tiled");
  }();
  if (off || row_index) return false;
  const msik_step_cfg& c = P.cfg;
  const bool drive = !(c.flags & MSIK_F_NO_DRIVE);
  if (!aligned16(P.qpos) || !aligned16(P.action) || (drive && !aligned16(P.drive))) return false;
  // warps per CTA: 4 while the per-warp rings fit, 2 or 1 for wide rows (Realman: 41-float rows = 37 KB per warp)
  static const int warps_env = [] { const char* e = getenv("MSIK_TMA_WARPS"); return e ? atoi(e) : 0; }();  // experiment knob
  int warps = (warps_env >= 1 && warps_env <= WARPS) ? warps_env : WARPS;
  while (warps > 1 && tma_smem_bytes(c, drive, warps) > (size_t)kTmaMaxSmem) warps >>= 1;
  *smem = tma_smem_bytes(c, drive, warps);
  if (*smem > (size_t)kTmaMaxSmem) return false;  // wider rows than this: tiled kernel
  P.tma_warps = warps;
  static const int dbg = [] { const char* e = getenv("MSIK_DEBUG_NOMATH"); return e ? atoi(e) : 0; }();
  P.debug = dbg;
  P.vec_ok = 1;
  P.merge = drive ? 1 : 0;
  return true;
}

}  // namespace msik

static int check_vel_cfg(const msik_vel_cfg* cfg) {
  if (!cfg) return fail(MSIK_EINVAL, "NULL cfg");
  if (cfg->n_envs < 0) return fail(MSIK_EINVAL, "n_envs=%d", cfg->n_envs);
  if (cfg->ego < MSIK_VEL_JOINT || cfg->ego > MSIK_VEL_FORWARD) return fail(MSIK_EINVAL, "ego=%d", cfg->ego);
  const int min_cols = cfg->ego == MSIK_VEL_JOINT ? 1 : 3;  // pd_base_vel.py:15: x, y and the orientation joint
  if (cfg->n_cols < min_cols || cfg->n_cols > MSIK_MAX_AUX)
    return fail(MSIK_EINVAL, "n_cols=%d outside [%d, %d]", cfg->n_cols, min_cols, MSIK_MAX_AUX);
  const int na = cfg->ego == MSIK_VEL_FORWARD ? cfg->n_cols - 1 : cfg->n_cols;
  if (cfg->action_offset < 0 || cfg->action_offset + na > cfg->action_stride)
    return fail(MSIK_EINVAL, "action slice [%d, %d) outside the action row (%d)", cfg->action_offset, cfg->action_offset + na, cfg->action_stride);
  unsigned long long seen = 0;
  for (int k = 0; k < cfg->n_cols; ++k) {
    const int c = cfg->col[k];
    if (c < 0 || c >= cfg->qvel_stride || c >= 64) return fail(MSIK_EINVAL, "col[%d]=%d outside the row", k, c);
    if ((seen >> c) & 1ull) return fail(MSIK_EINVAL, "col[%d]=%d is written twice", k, c);
    seen |= 1ull << c;
  }
  if (cfg->ego != MSIK_VEL_JOINT && (cfg->ori_col < 0 || cfg->ori_col >= cfg->qpos_stride))
    return fail(MSIK_EINVAL, "ori_col=%d outside the qpos row", cfg->ori_col);
  return MSIK_OK;
}

extern "C" {

int msik_delta_step(msik_chain_t chain, const msik_step_cfg* cfg, const float* qpos,
                    const int32_t* row_index, const float* action, float* drive_target,
                    float* target_qpos_out, float* start_qpos_out, void* stream) {
  return msik_delta_step_fused(chain, cfg, nullptr, qpos, row_index, action, drive_target, target_qpos_out,
                               start_qpos_out, stream);
}

int msik_delta_step_fused(msik_chain_t chain, const msik_step_cfg* cfg, const msik_aux_cfg* aux,
                          const float* qpos, const int32_t* row_index, const float* action,
                          float* drive_target, float* target_qpos_out, float* start_qpos_out,
                          void* stream) {
  return msik_delta_step_fused_vel(chain, cfg, aux, nullptr, qpos, row_index, action, drive_target, nullptr,
                                   target_qpos_out, start_qpos_out, stream);
}

// everything msik_delta_step_fused_vel decides before it launches: validated parameters + the kernel variant
struct DeltaPlan {
  StepParams P;
  int mode, nd;
  size_t smem;
  bool tma, tiled;
  bool vel_separate;  // tiled / gather variants: the velocity targets in their own small launch
  msik_vel_cfg vel;
  // second EE controller (msik_delta_step_dual): solved inside the same persistent launch (P.dual) or, where that does not
  // apply, by a launch of its own right after the first arm's
  bool second_separate;
  msik_chain_t chain2;
  msik_step_cfg cfg2;
};

static int plan_delta(msik_chain_t chain, const msik_step_cfg* cfg, const msik_aux_cfg* aux, const msik_vel_cfg* vel,
                      const float* qpos, const int32_t* row_index, const float* action, float* drive_target,
                      float* target_qvel, float* target_qpos_out, float* start_qpos_out, DeltaPlan& D,
                      const msik_dual_cfg* second = nullptr) {
  StepParams& P = D.P;
  D.second_separate = false;
  D.chain2 = nullptr;
  int rc = prepare(chain, cfg, P, &D.mode, &D.nd, true);
  if (rc) return rc;
  const int nd = D.nd;
  if (!(cfg->flags & MSIK_F_USE_DELTA) || (cfg->flags & MSIK_F_USE_TARGET))
    return fail(MSIK_EINVAL, "msik_delta_step serves use_delta && !use_target only (kinematics.py:156-172)");
  if (!qpos || !action) return fail(MSIK_EINVAL, "NULL qpos / action");
  if (!drive_target && !(cfg->flags & MSIK_F_NO_DRIVE)) return fail(MSIK_EINVAL, "NULL drive_target");
  if (P.n_ctrl < nd)
    return fail(MSIK_EINVAL, "%d controlled joints < %d task dimensions: least-squares case not supported", P.n_ctrl, nd);
  if (aux && aux->n_cols) {
    if (cfg->flags & MSIK_F_NO_DRIVE) return fail(MSIK_EINVAL, "fused sibling columns need the drive-target buffer");
    if (aux->n_cols < 0 || aux->n_cols > MSIK_MAX_AUX) return fail(MSIK_EINVAL, "aux n_cols=%d out of range", aux->n_cols);
    unsigned long long seen = P.colmask;
    for (int k = 0; k < aux->n_cols; ++k) {
      const int c = aux->col[k];
      if (c < 0 || c >= cfg->target_stride || c >= cfg->qpos_stride || c >= 64)
        return fail(MSIK_EINVAL, "aux col[%d]=%d outside the row", k, c);
      if ((seen >> c) & 1ull) return fail(MSIK_EINVAL, "aux col[%d]=%d is written twice", k, c);
      seen |= 1ull << c;
      if (aux->action_col[k] < 0 || aux->action_col[k] >= cfg->action_stride)
        return fail(MSIK_EINVAL, "aux action_col[%d]=%d outside the action row", k, aux->action_col[k]);
    }
    P.aux = *aux;
    P.aux_uniform = 1;
    P.aux_acol_rel = aux->n_cols ? aux->action_col[0] - cfg->action_offset : 0;
    for (int k = 0; k < aux->n_cols; ++k)
      if (aux->action_col[k] != aux->action_col[0] || aux->flags[k] != aux->flags[0] || aux->mid[k] != aux->mid[0] ||
          aux->half[k] != aux->half[0] || (aux->flags[k] & MSIK_AUX_DELTA))
        P.aux_uniform = 0;
    const unsigned long long all_cols = cfg->target_stride >= 64 ? ~0ull : ((1ull << cfg->target_stride) - 1ull);
    P.full_row = cfg->target_stride <= 64 && seen == all_cols;
  }
  const bool has_vel = vel && vel->n_cols;
  if (has_vel) {
    rc = check_vel_cfg(vel);
    if (rc) return rc;
    if (vel->n_envs != cfg->n_envs || vel->qpos_stride != cfg->qpos_stride || vel->action_stride != cfg->action_stride)
      return fail(MSIK_EINVAL, "fused velocity controller: n_envs / qpos_stride / action_stride differ from the step's");
    if (!target_qvel) return fail(MSIK_EINVAL, "NULL target_qvel");
    D.vel = *vel;
  }
  P.qpos = qpos; P.row_index = row_index; P.action = action; P.drive = drive_target;
  P.tq_out = target_qpos_out; P.sq_out = start_qpos_out;
  P.qvel = target_qvel;
  D.smem = 0;
  D.vel_separate = false;
  D.tma = choose_tma(P, row_index, &D.smem);
  if (D.tma && has_vel) P.vel = *vel;  // written inside the persistent kernel's loop
  if (second && second->chain) {
    const msik_step_cfg& c2 = second->cfg;
    msik_chain_t ch2 = second->chain;
    const int n2 = ch2->desc.n_joints;
    const int nd2 = (c2.flags & MSIK_F_POS_ONLY) ? 3 : 6;
    if (c2.n_envs != cfg->n_envs || c2.qpos_stride != cfg->qpos_stride || c2.target_stride != cfg->target_stride ||
        c2.action_stride != cfg->action_stride)
      return fail(MSIK_EINVAL, "dual step: the second controller's n_envs / strides differ from the first one's");
    if (!(c2.flags & MSIK_F_USE_DELTA) || (c2.flags & (MSIK_F_USE_TARGET | MSIK_F_NO_DRIVE)))
      return fail(MSIK_EINVAL, "dual step: the second controller must be on the one-step delta path and write drive targets");
    if (c2.action_offset < 0 || c2.action_offset + nd2 > c2.action_stride)
      return fail(MSIK_EINVAL, "dual step: second action slice outside the action row");
    unsigned long long colmask2 = 0;
    for (int i = 0; i < n2; ++i) {
      const int col = c2.anc_col[i];
      if (col < 0 || col >= c2.qpos_stride) return fail(MSIK_EINVAL, "dual step: second anc_col[%d]=%d outside the row", i, col);
      if (((c2.ctrl_mask >> i) & 1u) && col < 64) colmask2 |= 1ull << col;
    }
    if (colmask2 & P.colmask) return fail(MSIK_EINVAL, "dual step: the two controllers drive a common column");
    for (int k = 0; k < P.aux.n_cols; ++k)
      if ((colmask2 >> P.aux.col[k]) & 1ull)
        return fail(MSIK_EINVAL, "dual step: sibling column %d is driven by the second controller too", P.aux.col[k]);
    D.chain2 = ch2;
    D.cfg2 = c2;
    const ChainOps* o1 = chain->ops;
    const bool fused = D.tma && tma_lean(P) != 0 && o1->delta_tma_dual && o1->dual_partner && !strcmp(o1->dual_partner, ch2->ops->name) &&
                       D.mode == CTRL_RANGE && ch2->ops->n == n2 && ch2->ops->range_mask && c2.ctrl_mask == ch2->ops->range_mask &&
                       nd2 == D.nd && cfg->target_stride <= 64;
    if (fused) {
      P.dual = 1;
      P.cfg2 = c2;
      for (int i = n2; i < MSIK_MAX_JOINTS; ++i) P.cfg2.anc_col[i] = c2.anc_col[n2 - 1];
      for (int k = 0; k < 3; ++k) {
        P.mid2[k] = 0.5f * (c2.pos_high[k] + c2.pos_low[k]);
        P.half2[k] = 0.5f * (c2.pos_high[k] - c2.pos_low[k]);
      }
      const int nc2 = __builtin_popcount(c2.ctrl_mask);
      P.rank_rtol2 = (float)(nd2 > nc2 ? nd2 : nc2) * 1.1920929e-07f;
      unsigned long long seen = P.colmask | colmask2;
      for (int k = 0; k < P.aux.n_cols; ++k) seen |= 1ull << P.aux.col[k];
      const unsigned long long all_cols = cfg->target_stride >= 64 ? ~0ull : ((1ull << cfg->target_stride) - 1ull);
      P.full_row = seen == all_cols;
    } else {
      D.second_separate = true;
    }
  }
  if (D.tma) {
    D.tiled = false;
    return MSIK_OK;
  }
  D.vel_separate = has_vel;
  D.tiled = choose_tiled(P, row_index, &D.smem);
  return MSIK_OK;
}

static int launch_first_arm(msik_chain_t chain, const DeltaPlan& D, void* stream);

static int launch_delta_plan(msik_chain_t chain, const DeltaPlan& D, void* stream) {
  int rc = launch_first_arm(chain, D, stream);
  if (rc == MSIK_OK && D.second_separate && D.P.cfg.n_envs > 0)  // no fused kernel for this pair / layout: the second arm on its own
    rc = msik_delta_step(D.chain2, &D.cfg2, D.P.qpos, D.P.row_index, D.P.action, D.P.drive, nullptr, nullptr, stream);
  return rc;
}

static int launch_first_arm(msik_chain_t chain, const DeltaPlan& D, void* stream) {
  const StepParams& P = D.P;
  if (P.cfg.n_envs == 0) return MSIK_OK;
  if (D.tma) {
    cudaError_t e = P.dual ? chain->ops->delta_tma_dual(P, D.nd, D.smem, (cudaStream_t)stream)
                           : chain->ops->delta_tma(P, D.nd, D.mode, D.smem, (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_fail(e, "delta_step_tma_kernel launch");
    return MSIK_OK;
  }
  if (D.vel_separate) {
    int rc = msik_vel_targets(&D.vel, P.qpos, P.row_index, P.action, P.qvel, stream);
    if (rc) return rc;
  }
  cudaError_t e = chain->ops->delta(P, D.nd, !D.tiled, D.mode, D.smem, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "delta_step_kernel launch");
  if (P.aux.n_cols) {  // tiled / gather variants: sibling columns in a second small launch
    aux_targets_kernel<<<(P.cfg.n_envs + TPB - 1) / TPB, TPB, 0, (cudaStream_t)stream>>>(P);
    g_launches.fetch_add(1);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "aux_targets_kernel launch");
  }
  return MSIK_OK;
}

int msik_delta_step_fused_vel(msik_chain_t chain, const msik_step_cfg* cfg, const msik_aux_cfg* aux, const msik_vel_cfg* vel,
                              const float* qpos, const int32_t* row_index, const float* action, float* drive_target,
                              float* target_qvel, float* target_qpos_out, float* start_qpos_out, void* stream) {
  DeltaPlan D;
  int rc = plan_delta(chain, cfg, aux, vel, qpos, row_index, action, drive_target, target_qvel, target_qpos_out, start_qpos_out, D);
  if (rc) return rc;
  return launch_delta_plan(chain, D, stream);
}

int msik_delta_step_dual(msik_chain_t chain, const msik_step_cfg* cfg, const msik_dual_cfg* second, const msik_aux_cfg* aux,
                         const msik_vel_cfg* vel, const float* qpos, const int32_t* row_index, const float* action,
                         float* drive_target, float* target_qvel, void* stream) {
  DeltaPlan D;
  int rc = plan_delta(chain, cfg, aux, vel, qpos, row_index, action, drive_target, target_qvel, nullptr, nullptr, D, second);
  if (rc) return rc;
  return launch_delta_plan(chain, D, stream);
}

// ------------------------------------------------------------------ prepared step
struct msik_step {
  msik_chain_t chain;
  msik_dual_cfg second;
  bool has_second;
  msik_step_cfg cfg;
  msik_aux_cfg aux;
  msik_vel_cfg vel;
  bool has_aux, has_vel;
  const float* qpos;
  const int32_t* row_index;
  float* drive;
  float* qvel;
  DeltaPlan plan;  // built for a 16-byte aligned action pointer
};

int msik_step_create(msik_chain_t chain, const msik_step_cfg* cfg, const msik_aux_cfg* aux, const msik_vel_cfg* vel,
                     const float* qpos, const int32_t* row_index, float* drive_target, float* target_qvel, msik_step_t* out) {
  return msik_step_create_dual(chain, cfg, nullptr, aux, vel, qpos, row_index, drive_target, target_qvel, out);
}

int msik_step_create_dual(msik_chain_t chain, const msik_step_cfg* cfg, const msik_dual_cfg* second, const msik_aux_cfg* aux,
                          const msik_vel_cfg* vel, const float* qpos, const int32_t* row_index, float* drive_target,
                          float* target_qvel, msik_step_t* out) {
  if (!out) return fail(MSIK_EINVAL, "NULL out");
  msik_step* s = new (std::nothrow) msik_step;
  if (!s) return fail(MSIK_ENOMEM, "out of host memory");
  const float* aligned_dummy = reinterpret_cast<const float*>(uintptr_t(256));
  int rc = plan_delta(chain, cfg, aux, vel, qpos, row_index, aligned_dummy, drive_target, target_qvel, nullptr, nullptr, s->plan, second);
  if (rc) { delete s; return rc; }
  s->has_second = second && second->chain;
  if (s->has_second) s->second = *second;
  s->chain = chain; s->cfg = *cfg;
  s->has_aux = aux && aux->n_cols; s->has_vel = vel && vel->n_cols;
  if (s->has_aux) s->aux = *aux;
  if (s->has_vel) s->vel = *vel;
  s->qpos = qpos; s->row_index = row_index; s->drive = drive_target; s->qvel = target_qvel;
  *out = s;
  return MSIK_OK;
}

int msik_step_launch(msik_step_t s, const float* action, void* stream) {
  if (!s || !action) return fail(MSIK_EINVAL, "NULL step / action");
  if (!aligned16(action))  // the plan was made for an aligned action buffer: decide again for this one
    return msik_delta_step_dual(s->chain, &s->cfg, s->has_second ? &s->second : nullptr, s->has_aux ? &s->aux : nullptr,
                                s->has_vel ? &s->vel : nullptr, s->qpos, s->row_index, action, s->drive, s->qvel, stream);
  s->plan.P.action = action;
  return launch_delta_plan(s->chain, s->plan, stream);
}

int msik_step_destroy(msik_step_t s) {
  delete s;
  return MSIK_OK;
}

int msik_target_ik(msik_chain_t chain, const msik_step_cfg* cfg, const msik_iter_cfg* it,
                   const float* qpos, const int32_t* row_index, const float* action,
                   const float* prev_pose, float* target_pose_out, float* drive_target,
                   float* target_qpos_out, float* start_qpos_out, int32_t* iters_out,
                   float* err_out, void* stream) {
  StepParams P;
  int mode, nd;
  int rc = prepare(chain, cfg, P, &mode, &nd, cfg && !(cfg->flags & MSIK_F_TARGET_GIVEN));
  if (rc) return rc;
  if (!it) return fail(MSIK_EINVAL, "NULL iter cfg");
  const bool given = cfg->flags & MSIK_F_TARGET_GIVEN;
  if (!qpos || (!action && !given)) return fail(MSIK_EINVAL, "NULL qpos / action");
  if (((cfg->flags & MSIK_F_USE_DELTA) || given) && !prev_pose) return fail(MSIK_EINVAL, "use_delta / target_given needs prev_pose");
  if (!drive_target && !(cfg->flags & MSIK_F_NO_DRIVE)) return fail(MSIK_EINVAL, "NULL drive_target");
  if (P.n_ctrl < 6) return fail(MSIK_EINVAL, "%d controlled joints < 6 task dimensions", P.n_ctrl);
  if (it->max_iterations < 0) return fail(MSIK_EINVAL, "max_iterations < 0");
  if (cfg->n_envs == 0) return MSIK_OK;
  if (!(it->damping >= 0.f)) return fail(MSIK_EINVAL, "damping=%g must be >= 0", (double)it->damping);
  P.it = *it;
  if (it->rank_rtol > 0.f) P.rank_rtol = it->rank_rtol;
  // damping == 0: the fast (undamped) step everywhere.  damping > 0: the fast step while an env's smallest pivot stays
  // above the guard (there the damping is below fp32 resolution), the exact damped step from then on (msik_ik.cuh).
  P.ik_lambda = it->damping > 0.f ? sqrtf(it->damping) : 0.f;
  P.ik_guard = it->damping > 0.f ? (it->guard_pivot > 0.f ? it->guard_pivot : ik_pivot_guard(it->damping)) : -1.f;
  P.na = nd;
  P.qpos = qpos; P.row_index = row_index; P.action = action; P.drive = drive_target;
  P.tq_out = target_qpos_out; P.sq_out = start_qpos_out;
  P.prev_pose = prev_pose; P.pose_out = target_pose_out; P.iters_out = iters_out; P.err_out = err_out;
  // mapping: lane-per-joint groups for small batches (the step is then the serial latency of the iterations and a group
  // of lanes walks an iteration faster than one lane), one thread per env otherwise (profiles/r02f_ik_warp_ab.txt)
  {
    static const int force = [] { const char* v = getenv("MSIK_IK_WARP"); return v ? atoi(v) : -1; }();
    static const int max_envs = [] { const char* v = getenv("MSIK_IK_WARP_MAX_ENVS"); return v ? atoi(v) : kIkWarpMaxEnvs; }();
    const bool can = chain->ops->n <= WLPE;
    bool warp = it->mapping == MSIK_IK_LANE_PER_JOINT || (it->mapping == MSIK_IK_AUTO && cfg->n_envs <= max_envs);
    if (it->mapping == MSIK_IK_AUTO && force >= 0) warp = force != 0;
    if (it->mapping == MSIK_IK_LANE_PER_JOINT && !can)
      return fail(MSIK_EINVAL, "lane-per-joint mapping serves chains of <= %d joints (this one has %d)", WLPE, chain->ops->n);
    if (warp && can) {
      const long long threads = (long long)cfg->n_envs * WLPE;
      (void)cudaGetLastError();
      target_ik_warp_kernel<<<(unsigned)((threads + TPB - 1) / TPB), TPB, 0, (cudaStream_t)stream>>>(P);
      g_launches.fetch_add(1);
      cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) return cuda_fail(e, "target_ik_warp_kernel launch");
      return MSIK_OK;
    }
  }
  size_t smem = 0;
  const bool tiled = choose_tiled(P, row_index, &smem);
  P.gather = tiled ? 0 : 1;
  if (!tiled) smem = 0;
  // Lane compaction (msik_kernels.cuh target_ik_kernel): exchange buffer behind the tiles.  Built and measured on B200
  // (profiles/r02e_ik_variants.txt): 52.1 vs 51.3 us at 65536 envs, 565 vs 539 us at 2^20 -- the kernel is bound by the
  // latency of each warp's dependent chain, not by issue slots, so freeing lanes buys nothing and the barriers cost.
  // Compiled out of the shipped library (the disabled code alone cost 2 us per 65 536 envs); a build with
  // -DMSIK_IK_COMPACTION=1 has it, and MSIK_IK_COMPACT=1 then turns it on for A/B runs.
  P.ik_xoff = -1;
  P.ik_compact_from = 4;
  P.ik_sort = 0;
#if !MSIK_IK_COMPACTION
  {
    // Regrouping by remaining work (target_ik_kernel): pays once a sub-partition holds several warps, i.e. from about two
    // CTAs per SM; below that a step is the latency of its slowest warp whatever the grouping.  MSIK_IK_SORT=0 / 1 overrides.
    static const int sort_env = [] { const char* v = getenv("MSIK_IK_SORT"); return v ? atoi(v) : -1; }();
    static thread_local int sms_dev = -1, sms = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev != sms_dev) {
      if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms = 0;
      sms_dev = dev;
    }
    const int ctas = (cfg->n_envs + TPB - 1) / TPB;
    const bool want = sort_env >= 0 ? sort_env != 0 : (sms > 0 && ctas >= 2 * sms);
    const size_t sbytes = sizeof(float) * (size_t)IK_SORT_WORDS(chain->ops->n);
    if (want && mode != CTRL_MASKED && cfg->n_envs > 32 && P.it.max_iterations >= 2 && smem + sbytes <= 160 * 1024) {
      P.ik_sort = 1;
      P.ik_xoff = (int)(smem / sizeof(float));
      smem += sbytes;
    }
  }
#endif
#if MSIK_IK_COMPACTION
  static const int compact = [] { const char* v = getenv("MSIK_IK_COMPACT"); return v ? atoi(v) : 0; }();
  const size_t xbytes = (size_t)TPB * IK_XWORDS(chain->ops->n) * sizeof(float);
  if (compact && cfg->n_envs > 32 && smem + xbytes <= 160 * 1024) {
    P.ik_xoff = (int)(smem / sizeof(float));
    smem += xbytes;
  }
#endif
  cudaError_t e = chain->ops->target_ik(P, mode, smem, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "target_ik_kernel launch");
  return MSIK_OK;
}

int msik_fk(msik_chain_t chain, const msik_step_cfg* cfg, const float* qpos, const int32_t* row_index,
            float* pose_out, float* jac_out, void* stream) {
  StepParams P;
  int rc = prepare(chain, cfg, P, nullptr, nullptr, false);
  if (rc) return rc;
  if (!qpos) return fail(MSIK_EINVAL, "NULL qpos");
  if (cfg->n_envs == 0) return MSIK_OK;
  P.qpos = qpos; P.row_index = row_index; P.pose_out = pose_out; P.jac_out = jac_out;
  cudaError_t e = chain->ops->fk(P, chain->desc.n_joints, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "fk_kernel launch");
  return MSIK_OK;
}

int msik_fk_world(msik_chain_t chain, const msik_step_cfg* cfg, const float* qpos, const int32_t* row_index,
                  const float* root_pose, float* pose_out, void* stream) {
  StepParams P;
  int rc = prepare(chain, cfg, P, nullptr, nullptr, false);
  if (rc) return rc;
  if (!qpos || !root_pose || !pose_out) return fail(MSIK_EINVAL, "NULL argument");
  if (cfg->n_envs == 0) return MSIK_OK;
  P.qpos = qpos; P.row_index = row_index; P.pose_out = pose_out; P.prev_pose = root_pose;
  cudaError_t e = chain->ops->fk(P, chain->desc.n_joints, (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "fk_kernel launch");
  return MSIK_OK;
}

int msik_ee_pose_at_base(const float* body_data, int32_t body_stride, const int32_t* root_index,
                         const int32_t* ee_index, int32_t n_envs, float* pose_out, void* stream) {
  if (!body_data || !root_index || !ee_index || !pose_out) return fail(MSIK_EINVAL, "NULL argument");
  if (body_stride < 7) return fail(MSIK_EINVAL, "body_stride=%d < 7", body_stride);
  if (n_envs <= 0) return n_envs == 0 ? MSIK_OK : fail(MSIK_EINVAL, "n_envs < 0");
  ee_pose_at_base_kernel<<<(n_envs + TPB - 1) / TPB, TPB, 0, (cudaStream_t)stream>>>(
      body_data, body_stride, root_index, ee_index, n_envs, pose_out);
  g_launches.fetch_add(1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "ee_pose_at_base_kernel launch");
  return MSIK_OK;
}

// cfg-only parameter block for the kernels that need no chain
static int prepare_nochain(const msik_step_cfg* cfg, StepParams& P) {
  if (!cfg) return fail(MSIK_EINVAL, "NULL cfg");
  if (cfg->n_envs < 0) return fail(MSIK_EINVAL, "n_envs < 0");
  memset(&P, 0, sizeof P);
  P.cfg = *cfg;
  for (int k = 0; k < 3; ++k) {
    P.mid[k] = 0.5f * (cfg->pos_high[k] + cfg->pos_low[k]);
    P.half[k] = 0.5f * (cfg->pos_high[k] - cfg->pos_low[k]);
  }
  P.n_ctrl = __builtin_popcount(cfg->ctrl_mask);
  return MSIK_OK;
}

int msik_target_pose(const msik_step_cfg* cfg, const float* action, const float* prev_pose,
                     float* target_pose_out, void* stream) {
  StepParams P;
  int rc = prepare_nochain(cfg, P);
  if (rc) return rc;
  if (!action || !target_pose_out) return fail(MSIK_EINVAL, "NULL argument");
  if ((cfg->flags & MSIK_F_USE_DELTA) && !prev_pose) return fail(MSIK_EINVAL, "use_delta needs prev_pose");
  const int nd = (cfg->flags & MSIK_F_POS_ONLY) ? 3 : 6;
  if (cfg->action_offset < 0 || cfg->action_offset + nd > cfg->action_stride)
    return fail(MSIK_EINVAL, "action slice outside action stride");
  if (cfg->n_envs == 0) return MSIK_OK;
  P.action = action; P.prev_pose = prev_pose; P.pose_out = target_pose_out;
  const int grid = (cfg->n_envs + TPB - 1) / TPB;
  if (nd == 6) target_pose_kernel<6><<<grid, TPB, 0, (cudaStream_t)stream>>>(P);
  else target_pose_kernel<3><<<grid, TPB, 0, (cudaStream_t)stream>>>(P);
  g_launches.fetch_add(1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "target_pose_kernel launch");
  return MSIK_OK;
}

int msik_pose_to_delta_action(const float pos_low[3], const float pos_high[3], float rot_low, float rot_high,
                              int32_t pos_only, const float* target_pose, const float* ee_pose, int32_t n_envs,
                              float* action_out, uint8_t* exact_out, void* stream) {
  if (!pos_low || !pos_high || !target_pose || !ee_pose || !action_out) return fail(MSIK_EINVAL, "NULL argument");
  if (n_envs <= 0) return n_envs == 0 ? MSIK_OK : fail(MSIK_EINVAL, "n_envs < 0");
  DeltaActionParams P;
  P.n = n_envs;
  P.nd = pos_only ? 3 : 6;
  for (int k = 0; k < 6; ++k) {
    const float lo = k < 3 ? pos_low[k] : rot_low, hi = k < 3 ? pos_high[k] : rot_high;
    const float half = 0.5f * (hi - lo);
    if (k < P.nd && !(half != 0.f)) return fail(MSIK_EINVAL, "empty action range in dimension %d", k);
    P.mid[k] = 0.5f * (hi + lo);
    P.inv_half[k] = half != 0.f ? 1.f / half : 0.f;
  }
  pose_to_delta_action_kernel<<<(n_envs + TPB - 1) / TPB, TPB, 0, (cudaStream_t)stream>>>(P, target_pose, ee_pose, action_out, exact_out);
  g_launches.fetch_add(1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "pose_to_delta_action_kernel launch");
  return MSIK_OK;
}

int msik_interp_targets(const msik_step_cfg* cfg, const float* start, const float* step_size,
                        float step, const int32_t* row_index, float* drive_target, void* stream) {
  StepParams P;
  int rc = prepare_nochain(cfg, P);
  if (rc) return rc;
  if (!start || !step_size || !drive_target) return fail(MSIK_EINVAL, "NULL argument");
  if (P.n_ctrl == 0 || cfg->n_envs == 0) return MSIK_OK;
  P.row_index = row_index; P.drive = drive_target;
  const long long total = (long long)cfg->n_envs * P.n_ctrl;
  interp_targets_kernel<<<(unsigned)((total + TPB - 1) / TPB), TPB, 0, (cudaStream_t)stream>>>(P, start, step_size, step);
  g_launches.fetch_add(1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "interp_targets_kernel launch");
  return MSIK_OK;
}

int msik_vel_targets(const msik_vel_cfg* cfg, const float* qpos, const int32_t* row_index, const float* action,
                     float* target_qvel, void* stream) {
  int rc = check_vel_cfg(cfg);
  if (rc) return rc;
  if (!action || !target_qvel || (cfg->ego != MSIK_VEL_JOINT && !qpos)) return fail(MSIK_EINVAL, "NULL argument");
  if (cfg->n_envs == 0) return MSIK_OK;
  vel_targets_kernel<<<(cfg->n_envs + TPB - 1) / TPB, TPB, 0, (cudaStream_t)stream>>>(*cfg, qpos, row_index, action, target_qvel);
  g_launches.fetch_add(1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "vel_targets_kernel launch");
  return MSIK_OK;
}

int msik_reset_masked(const msik_step_cfg* cfg, const uint8_t* reset_mask, const float* qpos,
                      const int32_t* row_index, const float* ee_pose, float* start_qpos,
                      float* target_qpos, float* target_pose, void* stream) {
  StepParams P;
  int rc = prepare_nochain(cfg, P);
  if (rc) return rc;
  if (!reset_mask || !qpos) return fail(MSIK_EINVAL, "NULL argument");
  if (cfg->n_envs == 0) return MSIK_OK;
  P.qpos = qpos; P.row_index = row_index;
  reset_masked_kernel<<<(cfg->n_envs + TPB - 1) / TPB, TPB, 0, (cudaStream_t)stream>>>(
      P, reset_mask, ee_pose, start_qpos, target_qpos, target_pose);
  g_launches.fetch_add(1);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "reset_masked_kernel launch");
  return MSIK_OK;
}

// ------------------------------------------------------------------ host-buffer pipeline
int msik_pipeline_create(int32_t chunk_envs, int32_t qpos_stride, int32_t action_stride, msik_pipeline_t* out) {
  if (!out || chunk_envs < 1 || qpos_stride < 1 || action_stride < 1) return fail(MSIK_EINVAL, "bad pipeline argument");
  msik_pipeline* p = new (std::nothrow) msik_pipeline;
  if (!p) return fail(MSIK_ENOMEM, "out of host memory");
  memset(p, 0, sizeof *p);
  p->chunk = chunk_envs; p->qs = qpos_stride; p->as = action_stride; p->nslots = kPipeSlots;
  cudaError_t e = cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->s_k, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking);
  // MSIK_PIPE_TRACE=1: timed events, and msik_delta_step_host prints when each stage of each chunk finished
  const unsigned evf = getenv("MSIK_PIPE_TRACE") ? cudaEventDefault : cudaEventDisableTiming;
  p->trace = evf == cudaEventDefault;
  if (p->trace && e == cudaSuccess) e = cudaEventCreate(&p->ev_t0);
  for (int i = 0; i < p->nslots && e == cudaSuccess; ++i) {
    e = cudaEventCreateWithFlags(&p->ev_in[i], evf);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_k[i], evf);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_out[i], evf);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_q[i], sizeof(float) * (size_t)chunk_envs * qpos_stride);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_a[i], sizeof(float) * (size_t)chunk_envs * action_stride);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_t[i], sizeof(float) * (size_t)chunk_envs * MSIK_MAX_JOINTS);
  }
  if (e != cudaSuccess) {
    msik_pipeline_destroy(p);
    return cuda_fail(e, "msik_pipeline_create");
  }
  *out = p;
  return MSIK_OK;
}

int msik_pipeline_destroy(msik_pipeline_t p) {
  if (!p) return MSIK_OK;
  for (int i = 0; i < kPipeSlots; ++i) {
    if (p->d_q[i]) cudaFree(p->d_q[i]);
    if (p->d_a[i]) cudaFree(p->d_a[i]);
    if (p->d_t[i]) cudaFree(p->d_t[i]);
    if (p->ev_in[i]) cudaEventDestroy(p->ev_in[i]);
    if (p->ev_k[i]) cudaEventDestroy(p->ev_k[i]);
    if (p->ev_out[i]) cudaEventDestroy(p->ev_out[i]);
  }
  if (p->ev_t0) cudaEventDestroy(p->ev_t0);
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_k) cudaStreamDestroy(p->s_k);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  delete p;
  return MSIK_OK;
}

int msik_delta_step_host(msik_pipeline_t p, msik_chain_t chain, const msik_step_cfg* cfg,
                         const float* h_qpos, const float* h_action, float* h_target_out) {
  if (!p || !chain || !cfg || !h_qpos || !h_action || !h_target_out) return fail(MSIK_EINVAL, "NULL argument");
  if (cfg->qpos_stride != p->qs || cfg->action_stride != p->as)
    return fail(MSIK_EINVAL, "pipeline was created for strides (%d,%d), cfg has (%d,%d)", p->qs, p->as,
                cfg->qpos_stride, cfg->action_stride);
  const int c = __builtin_popcount(cfg->ctrl_mask);
  int rc = MSIK_OK;
  cudaError_t e = cudaSuccess;
  int k = 0;
  if (p->trace) cudaEventRecord(p->ev_t0, p->s_in);
  for (int e0 = 0; e0 < cfg->n_envs && rc == MSIK_OK; e0 += p->chunk, ++k) {
    const int n = cfg->n_envs - e0 < p->chunk ? cfg->n_envs - e0 : p->chunk;
    const int slot = k % p->nslots;
    const bool reused = k >= p->nslots;
    // host -> device: the slot's inputs are free once its previous kernel has run
    if (reused) e = cudaStreamWaitEvent(p->s_in, p->ev_k[slot], 0);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(p->d_q[slot], h_qpos + (size_t)e0 * p->qs, sizeof(float) * (size_t)n * p->qs, cudaMemcpyHostToDevice, p->s_in);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(p->d_a[slot], h_action + (size_t)e0 * p->as, sizeof(float) * (size_t)n * p->as, cudaMemcpyHostToDevice, p->s_in);
    if (e == cudaSuccess) e = cudaEventRecord(p->ev_in[slot], p->s_in);
    // kernel: needs the inputs, and the slot's previous targets to have left the device
    if (e == cudaSuccess) e = cudaStreamWaitEvent(p->s_k, p->ev_in[slot], 0);
    if (e == cudaSuccess && reused) e = cudaStreamWaitEvent(p->s_k, p->ev_out[slot], 0);
    if (e != cudaSuccess) { rc = cuda_fail(e, "H2D copy"); break; }
    msik_step_cfg sub = *cfg;
    sub.n_envs = n;
    sub.flags |= MSIK_F_NO_DRIVE;
    rc = msik_delta_step(chain, &sub, p->d_q[slot], nullptr, p->d_a[slot], nullptr, p->d_t[slot], nullptr, p->s_k);
    if (rc) break;
    e = cudaEventRecord(p->ev_k[slot], p->s_k);
    // device -> host
    if (e == cudaSuccess) e = cudaStreamWaitEvent(p->s_out, p->ev_k[slot], 0);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(h_target_out + (size_t)e0 * c, p->d_t[slot], sizeof(float) * (size_t)n * c, cudaMemcpyDeviceToHost, p->s_out);
    if (e == cudaSuccess) e = cudaEventRecord(p->ev_out[slot], p->s_out);
    if (e != cudaSuccess) { rc = cuda_fail(e, "D2H copy"); break; }
  }
  for (cudaStream_t st : {p->s_in, p->s_k, p->s_out}) {
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess && rc == MSIK_OK) rc = cuda_fail(e, "pipeline sync");
  }
  if (p->trace && rc == MSIK_OK && k <= p->nslots) {  // every slot used once: its events still belong to this call
    for (int i = 0; i < k; ++i) {
      float a = 0, b = 0, d = 0;
      cudaEventElapsedTime(&a, p->ev_t0, p->ev_in[i]);
      cudaEventElapsedTime(&b, p->ev_t0, p->ev_k[i]);
      cudaEventElapsedTime(&d, p->ev_t0, p->ev_out[i]);
      fprintf(stderr, "msik pipe chunk %2d: inputs on device %.3f ms, kernel done %.3f, targets on host %.3f\n", i, a, b, d);
    }
  }
  return rc;
}

}  // extern "C"

// ------------------------------------------------------------------ no-allocation host-buffer step
// msik_delta_step_host needs a pipeline whose device staging slots come from cudaMalloc.  Callers that may not allocate
// (CUDA graph capture around them, tight serving loops, harnesses that forbid it) get the same step with the staging
// in a file-scope device array instead: every GPU that loads this library holds MSIK_NOALLOC_MAX_BYTES until exit.
// In HOST code `g_noalloc_scratch` is only a stand-in for the device symbol, never a device address: its address is
// resolved once per device with cudaGetSymbolAddress (that device current) and that pointer is what the kernel gets.
static __device__ __align__(256) unsigned char g_noalloc_scratch[MSIK_NOALLOC_MAX_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[kMaxDevices] = {};

int noalloc_scratch(unsigned char** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) return fail(MSIK_EINVAL, "device index %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // the current device's copy of the array
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
    g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return MSIK_OK;
}

size_t align256(size_t n) { return (n + 255) & ~size_t(255); }
}  // namespace

extern "C" {

int64_t msik_noalloc_bytes_needed(const msik_step_cfg* cfg) {
  if (!cfg || cfg->n_envs < 0 || cfg->qpos_stride < 1 || cfg->action_stride < 1) return -1;
  const size_t n = (size_t)cfg->n_envs;
  const size_t c = (size_t)__builtin_popcount(cfg->ctrl_mask);
  return (int64_t)(align256(n * cfg->qpos_stride * 4) + align256(n * cfg->action_stride * 4) + align256(n * c * 4));
}

int msik_delta_step_host_noalloc(msik_chain_t chain, const msik_step_cfg* cfg, const float* h_qpos, const float* h_action,
                                 float* h_target_out, void* stream) {
  if (!chain || !cfg || !h_qpos || !h_action || !h_target_out) return fail(MSIK_EINVAL, "NULL argument");
  const int64_t need = msik_noalloc_bytes_needed(cfg);
  if (need < 0) return fail(MSIK_EINVAL, "bad step configuration (n_envs / strides)");
  if (need > (int64_t)MSIK_NOALLOC_MAX_BYTES)
    return fail(MSIK_EINVAL,
                "msik_delta_step_host_noalloc: %d envs need %lld bytes of staging, the static scratch holds %lld "
                "(MSIK_NOALLOC_MAX_BYTES); split the batch or use msik_delta_step_host",
                cfg->n_envs, (long long)need, (long long)MSIK_NOALLOC_MAX_BYTES);
  if (cfg->n_envs == 0) return MSIK_OK;
  unsigned char* base = nullptr;
  int rc = noalloc_scratch(&base);
  if (rc) return rc;
  const size_t n = (size_t)cfg->n_envs;
  const size_t c = (size_t)__builtin_popcount(cfg->ctrl_mask);
  const size_t qb = n * cfg->qpos_stride * 4, ab = n * cfg->action_stride * 4, tb = n * c * 4;
  // every byte the kernel reads is overwritten here and every byte copied back is written by the kernel:
  // nothing in the shared array has to start at zero, whatever the previous call left in it
  float* d_q = reinterpret_cast<float*>(base);
  float* d_a = reinterpret_cast<float*>(base + align256(qb));
  float* d_t = reinterpret_cast<float*>(base + align256(qb) + align256(ab));
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemcpyAsync(d_q, h_qpos, qb, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_a, h_action, ab, cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess) return cuda_fail(e, "H2D copy");
  msik_step_cfg sub = *cfg;
  sub.target_stride = sub.qpos_stride;
  sub.flags |= MSIK_F_NO_DRIVE;
  rc = msik_delta_step(chain, &sub, d_q, nullptr, d_a, nullptr, d_t, nullptr, stream);
  if (rc) return rc;
  e = cudaMemcpyAsync(h_target_out, d_t, tb, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return cuda_fail(e, "D2H copy / sync");
  return MSIK_OK;
}

}  // extern "C"
