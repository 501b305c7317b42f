/*
 * msik.h -- C ABI of the B200-native batched end-effector controller hot path
 * (forward kinematics -> geometric Jacobian -> differential IK -> drive targets).
 *
 * The reference (ManiSkill, mani_skill 3.0.0b20) has no native boundary for this
 * path: it is eager PyTorch + pytorch_kinematics on the GPU and SAPIEN/Pinocchio
 * on the CPU.  The entry points below are what a binding for that path has to
 * call; each cites the reference function it replaces (paths relative to the
 * reference root).  INTEGRATION.md shows the ctypes stub a maintainer adds.
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the name starts with h_ (host);
 *   - all arrays are fp32 / int32, row-major, caller-owned, never retained;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*),
 *     re-entrant, and touches no global state; chain handles are immutable;
 *   - return value 0 = success, negative = error (MSIK_E*); the message of the
 *     last error on the calling thread is available from msik_last_error();
 *   - no C++ exception crosses the boundary.
 */
#ifndef MSIK_H_
#define MSIK_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSIK_VERSION 111 /* 0.1.1: msik_iter_cfg gained damping / guard_pivot / mapping; msik_step_* */

#define MSIK_MAX_JOINTS 16 /* movable joints along root -> end link */
#define MSIK_MAX_AUX 16    /* joint-position targets of sibling controllers fused into a launch */

/* error codes */
#define MSIK_OK 0
#define MSIK_EINVAL (-22)  /* bad argument / unsupported configuration      */
#define MSIK_ECUDA (-5)    /* a CUDA runtime call or kernel launch failed   */
#define MSIK_ENOMEM (-12)
#define MSIK_ENODEV (-19)  /* no sm_100 device                               */

/* joint types of the flattened chain */
#define MSIK_JOINT_REVOLUTE 1 /* URDF revolute / continuous */
#define MSIK_JOINT_PRISMATIC 2

/* msik_step_cfg.flags */
#define MSIK_F_NORMALIZE 0x1    /* clip + scale the action (config.normalize_action)              */
#define MSIK_F_POS_ONLY 0x2     /* PDEEPosController: 3-dof action, J[:, 0:3]                     */
#define MSIK_F_USE_DELTA 0x4    /* config.use_delta                                               */
#define MSIK_F_USE_TARGET 0x8   /* config.use_target: target pose state is read and written       */
#define MSIK_F_BODY_TRANS 0x10  /* frame contains "body_translation" (else root_translation)       */
#define MSIK_F_BODY_ROT 0x20    /* frame contains "body_aligned_body_rotation" (else root_aligned)*/
#define MSIK_F_NO_DRIVE 0x40    /* do not write the drive-target buffer (config.interpolate)      */
#define MSIK_F_TARGET_GIVEN 0x80 /* msik_target_ik: prev_pose IS the target pose (compute_ik called
                                    directly, kinematics.py:124-155); action is ignored            */

/*
 * Flattened kinematic chain:  T_ee(q) = G_0 Z_0(q_0) ... G_{n-1} Z_{n-1}(q_{n-1}) G_tool,
 * every movable joint acting about / along its local z axis, G_i = (M_i, t_i)
 * constant.  Replaces the chain object built by
 * pk.build_serial_chain_from_urdf(...) at
 * mani_skill/agents/controllers/utils/kinematics.py:106-109.
 * pre[i] = { M_i row-major (9), t_i (3) }, pre[n_joints] = tool transform.
 */
typedef struct msik_chain_desc {
  int32_t n_joints;
  int32_t joint_type[MSIK_MAX_JOINTS];
  float pre[MSIK_MAX_JOINTS + 1][12];
  float lower[MSIK_MAX_JOINTS];
  float upper[MSIK_MAX_JOINTS];
} msik_chain_desc;

typedef struct msik_chain* msik_chain_t;

/*
 * Per-controller constants: index bookkeeping of Kinematics.__init__
 * (kinematics.py:54-71,119-122) and the action pre-processing constants of
 * PDEEPoseController (pd_ee_pose.py:178-210, gym_utils.py:102-105).
 */
typedef struct msik_step_cfg {
  int32_t n_envs;
  int32_t qpos_stride;   /* floats per row of cuda_articulation_qpos (max dof of the scene)   */
  int32_t target_stride; /* floats per row of cuda_articulation_target_qpos                   */
  int32_t action_stride; /* floats per row of the (flat) action tensor                        */
  int32_t action_offset; /* first column of this controller's slice (action_mapping[uid][0])  */
  int32_t flags;         /* MSIK_F_*                                                          */
  uint32_t ctrl_mask;    /* bit i: chain joint i is driven by this controller (qmask)         */
  int32_t anc_col[MSIK_MAX_JOINTS]; /* qpos / target column of chain joint i
                                       (active_ancestor_joint_idxs, kinematics.py:65-67)     */
  float pos_low[3];      /* action_space_low[:3]  (pos_lower broadcast)                       */
  float pos_high[3];     /* action_space_high[:3]                                             */
  float rot_scale;       /* config.rot_lower: rot_action * rot_lower, pd_ee_pose.py:209       */
} msik_step_cfg;

/* parameters of the iterative solver (pk.PseudoInverseIK, kinematics.py:110-117) */
typedef struct msik_iter_cfg {
  int32_t max_iterations; /* 200 in the reference                                            */
  int32_t patience;       /* no-improvement patience (recalled library default 2); <0 = off  */
  float lr;               /* recalled library default 0.2                                     */
  float pos_tol;          /* recalled library default 1e-3                                    */
  float rot_tol;          /* recalled library default 1e-2                                    */
  float rank_rtol;        /* relative pivot cut-off of the min-norm solve; <=0 -> default     */
  float damping;          /* lambda^2 of dq = J^T (J J^T + lambda^2 I)^-1 dx, pk's `regularlization`
                             (recalled library default 1e-9).  0: the undamped limit everywhere
                             (min-norm solve with the rank cut-off above).  > 0: envs whose smallest
                             Householder pivot stays above guard_pivot take the undamped step (the
                             damping is below fp32 resolution there), an env that comes closer to a
                             singularity continues with the exact damped update (QR of [J^T; lambda I]);
                             iters_out is negative for such envs                                   */
  float guard_pivot;      /* <= 0: derived from damping (2.5e7 * damping).  Huge (e.g. 1e30): every
                             iteration of every env takes the exact damped update                  */
  int32_t mapping;        /* MSIK_IK_AUTO: by env count; MSIK_IK_THREAD_PER_ENV: one thread per env,
                             everything in registers; MSIK_IK_LANE_PER_JOINT: a group of 8 lanes per
                             env, shuffle prefix products, Cholesky of J J^T + lambda^2 I with one
                             refinement step (chains of <= 8 joints; wins on small batches, where a
                             step is the serial latency of the iterations).  The lane-per-joint
                             kernel applies the damped update in every iteration (guard_pivot unused,
                             iters_out always positive)                                            */
} msik_iter_cfg;
#define MSIK_IK_AUTO 0
#define MSIK_IK_THREAD_PER_ENV 1
#define MSIK_IK_LANE_PER_JOINT 2

int msik_version(void);
const char* msik_last_error(void);

/* number of SMs / name of the device a stream would launch on; fails with MSIK_ENODEV
 * when no CUDA device of compute capability 10.x is present (there is no CPU path). */
int msik_device_check(int* sm_count, int* cc_major, int* cc_minor);

/* Replaces Kinematics._setup_gpu chain construction (kinematics.py:91-122). */
int msik_chain_create(const msik_chain_desc* desc, msik_chain_t* out);
int msik_chain_destroy(msik_chain_t chain);

/*
 * Register a chain specialisation compiled at run time (maniskill_b200/jit.py builds a small shared
 * object from csrc/ + generated straight-line FK / Jacobian code for one URDF chain; its
 * `msik_plugin_ops()` returns the table to pass here).  Later msik_chain_create calls whose
 * description matches use its kernels instead of the table-driven ones.  The table must stay valid
 * for the life of the process.  No reference counterpart (the reference has no compiled code).
 */
int msik_register_chain_ops(const void* ops);
/* name of the compiled specialisation the table matched ("panda_tcp", "fetch_gripper",
 * ...) or "generic"; written NUL-terminated into buf. */
int msik_chain_kernel_name(msik_chain_t chain, char* buf, int cap);

/*
 * Path A: one fused launch for PDEEPosController.set_action with
 * use_delta && !use_target (pd_ee_pose.py:91-115):
 *   _preprocess_action / _clip_and_scale_action (base_controller.py:120-133,
 *   pd_ee_pose.py:197-210), q0 = qpos[:, anc] (kinematics.py:143),
 *   J = pk_chain.jacobian(q0) (kinematics.py:157), q0 + pinv(J) @ action
 *   (kinematics.py:167-172), set_drive_targets (pd_joint_pos.py:68-71 ->
 *   utils/structs/articulation.py:731-738).
 *
 * qpos        (M, qpos_stride)   cuda_articulation_qpos
 * row_index   (n_envs) int32 rows of this articulation (Articulation._data_index,
 *             articulation.py:257-263) or NULL for rows 0..n_envs-1
 * action      (n_envs, action_stride)
 * drive_target(M, target_stride) cuda_articulation_target_qpos; only controlled
 *             columns of the indexed rows are modified (NULL with MSIK_F_NO_DRIVE)
 * target_qpos_out optional dense (n_envs, c) copy (controller._target_qpos)
 * start_qpos_out  optional dense (n_envs, c) copy of qpos[:, controlled]
 *             (controller._start_qpos)
 */
int msik_delta_step(msik_chain_t chain, const msik_step_cfg* cfg, const float* qpos,
                    const int32_t* row_index, const float* action, float* drive_target,
                    float* target_qpos_out, float* start_qpos_out, void* stream);

/*
 * Sibling joint-position controllers of the same CombinedController
 * (base_controller.py:287-300) fused into the path-A launch:
 * PDJointPosController.set_action / PDJointPosMimicController without use_target /
 * interpolate (pd_joint_pos.py:73-90):
 *   v = action[:, action_col[k]]
 *   if MSIK_AUX_NORMALIZE: v = mid[k] + half[k] * clip(v, -1, 1)   (gym_utils.py:102-105,
 *                              mid = 0.5(high+low), half = 0.5(high-low), rounded like torch)
 *   if MSIK_AUX_DELTA:     v = qpos[row, col[k]] + v                (use_delta, pd_joint_pos.py:79-84)
 *   drive_target[row, col[k]] = v
 * When the arm's controlled columns plus these columns cover every column of a target
 * row, whole rows are written and the old targets are never read.
 */
#define MSIK_AUX_NORMALIZE 0x1
#define MSIK_AUX_DELTA 0x2
typedef struct msik_aux_cfg {
  int32_t n_cols;
  int32_t col[MSIK_MAX_AUX];        /* qpos / target column                                  */
  int32_t action_col[MSIK_MAX_AUX]; /* column of the flat action (mimic joints share one)    */
  int32_t flags[MSIK_MAX_AUX];      /* MSIK_AUX_*                                            */
  float mid[MSIK_MAX_AUX];
  float half[MSIK_MAX_AUX];
} msik_aux_cfg;

/* msik_delta_step plus the fused sibling columns (aux may be NULL = msik_delta_step). */
int msik_delta_step_fused(msik_chain_t chain, const msik_step_cfg* cfg, const msik_aux_cfg* aux,
                          const float* qpos, const int32_t* row_index, const float* action,
                          float* drive_target, float* target_qpos_out, float* start_qpos_out,
                          void* stream);

/*
 * Path B: fused launch for set_action with use_target or absolute targets:
 * compute_target_pose (pd_ee_pose.py:76-89,212-238) + Kinematics.compute_ik with
 * use_delta_ik_solver=False (kinematics.py:144-155; pk.PseudoInverseIK.solve).
 *
 * prev_pose   (n_envs, 7) [p, q wxyz]: controller._target_pose when use_target,
 *             ee_pose_at_base otherwise; NULL when !use_delta
 * target_pose_out (n_envs, 7) new controller._target_pose (may alias prev_pose)
 * iters_out   optional (n_envs) int32 iterations used (negated when the env finished on the damped step)
 * err_out     optional (n_envs, 2) final position / rotation error norms
 */
int msik_target_ik(msik_chain_t chain, const msik_step_cfg* cfg, const msik_iter_cfg* it,
                   const float* qpos, const int32_t* row_index, const float* action,
                   const float* prev_pose, float* target_pose_out, float* drive_target,
                   float* target_qpos_out, float* start_qpos_out, int32_t* iters_out,
                   float* err_out, void* stream);

/*
 * Forward kinematics + Jacobian (pk chain.forward_kinematics / chain.jacobian as used
 * at kinematics.py:157).  pose_out (n_envs,7) [p, q wxyz] in the chain root frame;
 * jac_out optional (n_envs, 6, n_joints), rows [linear; angular], base frame.
 */
int msik_fk(msik_chain_t chain, const msik_step_cfg* cfg, const float* qpos,
            const int32_t* row_index, float* pose_out, float* jac_out, void* stream);

/*
 * World pose of the end link from joint positions alone: root_pose (n_envs,7) * FK(qpos)
 * (Pose.__mul__, utils/structs/pose.py:187-201).  Replaces reading `agent.tcp.pose` back from
 * PhysX rigid-body data (agents/robots/panda/panda.py:232-234, utils/structs/link.py:225-237)
 * for proprioception / reward terms, and `pin_model.compute_forward_kinematics` + `get_link_pose`
 * in the trajectory conversion (trajectory/utils/actions/conversion.py:136-143).
 */
int msik_fk_world(msik_chain_t chain, const msik_step_cfg* cfg, const float* qpos,
                  const int32_t* row_index, const float* root_pose, float* pose_out, void* stream);

/*
 * PDEEPosController.ee_pose_at_base (pd_ee_pose.py:61-64) = articulation.pose.inv() *
 * ee_link.pose from PhysX rigid body data (utils/structs/link.py:225-237,
 * pose.py:187-220).  body_data (Nbody, body_stride>=7) [p, q wxyz, ...];
 * root_index / ee_index (n_envs) int32 rows.  pose_out (n_envs, 7).
 */
int msik_ee_pose_at_base(const float* body_data, int32_t body_stride, const int32_t* root_index,
                         const int32_t* ee_index, int32_t n_envs, float* pose_out, void* stream);

/*
 * compute_target_pose alone (pd_ee_pose.py:76-89,212-238) incl. the action
 * pre-processing; used to materialise controller._target_pose on path A.
 */
int msik_target_pose(const msik_step_cfg* cfg, const float* action, const float* prev_pose,
                     float* target_pose_out, void* stream);

/*
 * Trajectory action conversion, the batchable arithmetic of
 * mani_skill/trajectory/utils/actions/conversion.py:52-81,144-183 (pd_joint_pos demonstrations ->
 * pd_ee_delta_pose / pd_ee_delta_pos actions), one env per thread instead of the reference's
 * per-step numpy / sapien.Pose code ("TODO: parallelize / use GPU", conversion.py:68):
 *   delta_position = target.p - ee.p                                 (conversion.py:156)
 *   delta_q        = ee.q (x) target.q^-1                            (conversion.py:163, sapien Pose.__mul__)
 *   a = [delta_position, compact_axis_angle(delta_q)]                (conversion.py:52-57,76-80;
 *        transforms3d quat2axangle, angle wrapped to (-pi, pi])
 *   a = (a - 0.5 (high + low)) / (0.5 (high - low))                  (gym_utils.inv_scale_action :89-91)
 *   position part clipped to [-1, 1], rotation part scaled to unit norm if longer; exact_out[e] = 0
 *   when either happened (the reference then retries the step, conversion.py:168-178)
 * cfg: pos_low / pos_high and rot_low / rot_high = action-space bounds (flags: MSIK_F_POS_ONLY for the
 * 3-dof controller).  target_pose / ee_pose (n_envs, 7) [p, q wxyz] in the same frame;
 * action_out (n_envs, 6 | 3); exact_out optional (n_envs) uint8.
 */
int msik_pose_to_delta_action(const float pos_low[3], const float pos_high[3], float rot_low, float rot_high,
                              int32_t pos_only, const float* target_pose, const float* ee_pose, int32_t n_envs,
                              float* action_out, uint8_t* exact_out, void* stream);

/*
 * Velocity controllers that sit in the same CombinedController as the EE controller:
 *   ego = MSIK_VEL_JOINT    PDJointVelController.set_action        (pd_joint_vel.py:38-42)
 *   ego = MSIK_VEL_BASE     PDBaseVelController.set_action         (pd_base_vel.py:19-33)
 *   ego = MSIK_VEL_FORWARD  PDBaseForwardVelController.set_action  (pd_base_vel.py:50-72)
 * plus BaseController._preprocess_action (base_controller.py:120-133) and the scatter of
 * Articulation.set_joint_drive_velocity_targets (utils/structs/articulation.py:743-761) into
 * cuda_articulation_target_qvel.  One launch; only the n_cols listed columns of our rows are written.
 *
 * action (n_envs, action_stride): the flat action; this controller's slice starts at action_offset and has
 * n_cols entries (MSIK_VEL_FORWARD: n_cols - 1 entries -- forward and rotation for 3 joints).
 * MSIK_VEL_BASE / FORWARD: (vx, vy) = Rz(qpos[row, ori_col]) (a0, a1 | 0); columns 2.. pass through.
 * flags & MSIK_AUX_NORMALIZE: v = mid[k] + half[k] * clip(a_k, -1, 1) per ACTION entry k, rounded like the
 * torch expression (separate multiply and add).
 */
#define MSIK_VEL_JOINT 0
#define MSIK_VEL_BASE 1
#define MSIK_VEL_FORWARD 2
typedef struct msik_vel_cfg {
  int32_t n_envs;
  int32_t n_cols;                /* controlled joints (<= MSIK_MAX_AUX)                    */
  int32_t col[MSIK_MAX_AUX];     /* qvel-target column of each controlled joint            */
  int32_t ego;                   /* MSIK_VEL_*                                             */
  int32_t ori_col;               /* qpos column of the orientation joint (ego != JOINT)    */
  int32_t flags;                 /* MSIK_AUX_NORMALIZE                                     */
  int32_t qpos_stride, action_stride, action_offset, qvel_stride;
  float mid[MSIK_MAX_AUX];       /* per action entry: 0.5 (high + low)                     */
  float half[MSIK_MAX_AUX];      /* per action entry: 0.5 (high - low)                     */
} msik_vel_cfg;
int msik_vel_targets(const msik_vel_cfg* cfg, const float* qpos, const int32_t* row_index, const float* action,
                     float* target_qvel, void* stream);

/*
 * msik_delta_step_fused plus the velocity controller of the same CombinedController (vel may be NULL): the stock Fetch
 * dict (arm + gripper + body + base, fetch.py:273-278) in ONE launch.  vel->n_envs, qpos_stride and action_stride must
 * equal cfg's; vel->action_offset is a column of the same flat action; row_index is shared.  Where the persistent
 * kernel does not apply (gathered rows, unaligned bases) the velocity targets are written by a second small launch
 * inside this call.
 */
int msik_delta_step_fused_vel(msik_chain_t chain, const msik_step_cfg* cfg, const msik_aux_cfg* aux, const msik_vel_cfg* vel,
                              const float* qpos, const int32_t* row_index, const float* action, float* drive_target,
                              float* target_qvel, float* target_qpos_out, float* start_qpos_out, void* stream);

/*
 * Two EE controllers of ONE CombinedController in one launch (base_controller.py:287-300 runs them one after the other; the
 * fork's Realman robot drives both arms from one flat action, agents/robots/realman/mobile_realman.py:436-443):
 * msik_delta_step_fused_vel for `chain` / `cfg` plus the second controller `second` -- same n_envs / strides / buffers, its
 * own chain, action slice and columns, also use_delta && !use_target, same task dimension (both pos-only or both pose).
 * Where the library has a fused kernel for the pair (the two Realman arm chains, stock stepping configuration, rows
 * 0..n_envs-1, 16-byte aligned buffers) both solves share ONE pass over the rows; anywhere else the call issues the second
 * controller's launch right after the first one's -- the results are the same either way.  second == NULL: the plain step.
 */
typedef struct msik_dual_cfg {
  msik_chain_t chain;
  msik_step_cfg cfg;
} msik_dual_cfg;
int msik_delta_step_dual(msik_chain_t chain, const msik_step_cfg* cfg, const msik_dual_cfg* second, const msik_aux_cfg* aux,
                         const msik_vel_cfg* vel, const float* qpos, const int32_t* row_index, const float* action,
                         float* drive_target, float* target_qvel, void* stream);

/*
 * Prepared step: msik_delta_step_fused_vel with everything but the action pointer and the stream fixed at creation
 * (validation, kernel-variant choice and the kernel parameter block are done once).  At RL-typical env counts
 * (4096) the host issue path is the step time, so the per-step call is three words.  The handle keeps the pointers
 * it was given (PhysX buffers live as long as the scene); not thread-safe per handle (one controller, one thread --
 * the Controller API's contract).  No reference counterpart.
 */
typedef struct msik_step* msik_step_t;
int msik_step_create(msik_chain_t chain, const msik_step_cfg* cfg, const msik_aux_cfg* aux, const msik_vel_cfg* vel,
                     const float* qpos, const int32_t* row_index, float* drive_target, float* target_qvel,
                     msik_step_t* out);
int msik_step_create_dual(msik_chain_t chain, const msik_step_cfg* cfg, const msik_dual_cfg* second, const msik_aux_cfg* aux,
                          const msik_vel_cfg* vel, const float* qpos, const int32_t* row_index, float* drive_target,
                          float* target_qvel, msik_step_t* out); /* msik_delta_step_dual, prepared */
int msik_step_launch(msik_step_t step, const float* action, void* stream);
int msik_step_destroy(msik_step_t step);

/*
 * PDJointPosController.before_simulation_step with interpolate
 * (pd_joint_pos.py:92-98): drive_target[row, col_j] = start[:, j] + step_size[:, j] * step.
 * start / step_size dense (n_envs, c).
 */
int msik_interp_targets(const msik_step_cfg* cfg, const float* start, const float* step_size,
                        float step, const int32_t* row_index, float* drive_target, void* stream);

/*
 * Masked controller reset (pd_joint_pos.py:51-66, pd_ee_pose.py:66-74):
 * where reset_mask[e] != 0: start_qpos[e] = target_qpos[e] = qpos[row(e), controlled];
 * target_pose[e] = ee_pose[e] (if both non-NULL).  reset_mask (n_envs) uint8.
 */
int msik_reset_masked(const msik_step_cfg* cfg, const uint8_t* reset_mask, const float* qpos,
                      const int32_t* row_index, const float* ee_pose, float* start_qpos,
                      float* target_qpos, float* target_pose, void* stream);

/*
 * End-to-end variant of msik_delta_step for HOST buffers (pinned recommended):
 * chunks the batch, and overlaps host->device copies, the kernel and the
 * device->host copy of the joint targets on internal streams (one per engine,
 * 8 device slots; MSIK_PIPE_TRACE=1 prints the per-chunk timeline).  h_target_out is
 * dense (n_envs, c).  Synchronous: returns when h_target_out is complete.
 */
typedef struct msik_pipeline* msik_pipeline_t;
int msik_pipeline_create(int32_t chunk_envs, int32_t qpos_stride, int32_t action_stride,
                         msik_pipeline_t* out);
int msik_pipeline_destroy(msik_pipeline_t p);
int msik_delta_step_host(msik_pipeline_t p, msik_chain_t chain, const msik_step_cfg* cfg,
                         const float* h_qpos, const float* h_action, float* h_target_out);

/*
 * msik_delta_step_host without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arithmetic as msik_delta_step_host (one launch of the path-A kernel, MSIK_F_NO_DRIVE), but the
 * device staging of qpos / action / joint targets lives in a file-scope `static __device__` array of
 * MSIK_NOALLOC_MAX_BYTES inside the library -- every GPU that loads libmsik.so holds that much until the process
 * exits -- instead of a pipeline's cudaMalloc'ed slots.  The call issues no cudaMalloc / cudaFree / stream or event
 * creation: two host->device copies, the kernel and one device->host copy on `stream`, then cudaStreamSynchronize.
 *
 * Limit: align256(n_envs * qpos_stride * 4) + align256(n_envs * action_stride * 4) + align256(n_envs * c * 4)
 * <= MSIK_NOALLOC_MAX_BYTES (c = controlled joints; Panda rows: 92 B per env -> 91 180 envs); larger batches are
 * rejected with MSIK_EINVAL and a message naming both figures (msik_noalloc_bytes_needed returns the left-hand side).
 * ALL calls on a device share the one array and it keeps the previous call's contents: the entry point overwrites
 * everything the kernel reads and copies back only what the kernel wrote, so nothing needs clearing, but two calls
 * that overlap on the GPU (different streams / threads, same device) RACE on it -- serialise them.
 * h_qpos (n_envs, qpos_stride), h_action (n_envs, action_stride), h_target_out dense (n_envs, c): host buffers
 * (pinned recommended; cfg->target_stride is ignored).  Synchronous: returns when h_target_out is complete.
 */
#define MSIK_NOALLOC_MAX_BYTES (8u << 20)
int64_t msik_noalloc_bytes_needed(const msik_step_cfg* cfg);
int msik_delta_step_host_noalloc(msik_chain_t chain, const msik_step_cfg* cfg, const float* h_qpos,
                                 const float* h_action, float* h_target_out, void* stream);

/* number of kernel launches issued through this library by the calling process */
int64_t msik_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* MSIK_H_ */
