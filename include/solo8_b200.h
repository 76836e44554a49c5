/* solo8_b200.h -- C-ABI of the B200-native batched Solo 8 step path.
 *
 * Drop-in boundary.  The reference has no FFI: its
 * boundary is the set of pybullet BulletClient calls made by
 *   gym_solo/envs/solo8v2vanilla.py:72-102   (step)
 *   gym_solo/envs/solo8v2vanilla.py:104-143  (reset, 500 settle steps)
 *   gym_solo/core/obs.py:130-159, 260-282, 347-363      (observations)
 *   gym_solo/core/rewards.py:104-118, 221-234, 237-373  (rewards)
 *   gym_solo/core/termination.py:38-56, 59-98           (terminations)
 * Each entry point below names the reference call(s) it replaces.  Plain C
 * types only: pointers, sizes, PODs.  Every `*_dev` pointer is CUDA device
 * memory owned by the caller (PyTorch tensors in the Python host layer); the
 * handle owns only the constant model / config / program tables.
 *
 * All functions return 0 on success, a negative SOLO_E* code otherwise;
 * solo_last_error() gives the message.  Calls on one handle are stream
 * ordered and not thread-safe (one handle per process per GPU).
 */
#ifndef SOLO8_B200_H_
#define SOLO8_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SOLO_MAX_LEGS 4
#define SOLO_MAX_CHAIN 3
#define SOLO_MAX_DOF 12
#define SOLO_MAX_BODIES 13
#define SOLO_MAX_API 16
#define SOLO_MAX_CAND 128
#define SOLO_MAX_OBS_TERMS 8
#define SOLO_MAX_REWARD_OPS 32
#define SOLO_MAX_TERM_TERMS 8
#define SOLO_REWARD_STACK 8

#define SOLO_OK 0
#define SOLO_EINVAL (-1)
#define SOLO_ECUDA (-2)
#define SOLO_ENOGPU (-3)

/* Compiled robot model (written by gym_solo_b200/assets/compile_model.py from
 * gym_solo/assets/solo8v2/solo.urdf:8-583).  Body 0 is the floating base in
 * its inertial frame; body 1+d is the child of dof d.  Dofs are chain
 * contiguous: leg l owns dofs [sum(chain_len[:l]), +chain_len[l]). */
typedef struct SoloModel {
  int32_t n_legs, ndof, n_api, n_cand;
  int32_t chain_len[SOLO_MAX_LEGS];
  int32_t api_dof[SOLO_MAX_API];          /* pybullet joint index -> dof, -1 = fixed */
  double api_lower[SOLO_MAX_API], api_upper[SOLO_MAX_API];
  double body_mass[SOLO_MAX_BODIES];
  double body_com[SOLO_MAX_BODIES][3];      /* in body frame */
  double body_inertia[SOLO_MAX_BODIES][9];  /* about COM, body-frame axes */
  double joint_origin[SOLO_MAX_DOF][3];     /* joint frame origin in parent body frame */
  double joint_rot[SOLO_MAX_DOF][9];        /* parent<-child rotation at q = 0 */
  double joint_axis[SOLO_MAX_DOF][3];       /* unit axis in child frame */
  int32_t cand_body[SOLO_MAX_CAND];         /* contact support points, grouped by body */
  double cand_pos[SOLO_MAX_CAND][3];        /* in body frame */
  double cand_mu[SOLO_MAX_CAND];            /* link lateral friction (changeDynamics) */
  double collision_margin;                  /* 0.001: btConvexHullShape margin */
} SoloModel;

/* Physics configuration = Solo8BaseConfig (gym_solo/core/configs.py:8-38) plus
 * the Bullet world defaults pybullet sets (recollection). */
typedef struct SoloConfig {
  double dt;                      /* configs.py:10 (1e-3); BASELINE configs use 1/240 */
  double gravity[3];              /* configs.py:17 */
  double motor_torque_limit;      /* configs.py:12; max impulse = limit*dt */
  double linear_damping;          /* configs.py:21 */
  double angular_damping;         /* configs.py:22 */
  double plane_friction;          /* pybullet_data plane.urdf: 1.0 */
  double motor_kp, motor_kd;      /* pybullet POSITION_CONTROL defaults 0.1 / 1.0 */
  double contact_erp;             /* m_erp2 = 0.08 */
  double contact_slop;            /* m_linearSlop = 1e-5 */
  double contact_breaking_threshold; /* 0.02 */
  double max_coordinate_velocity; /* btMultiBody m_maxCoordinateVelocity = 100 */
  double residual_threshold;      /* m_leastSquaresResidualThreshold = 1e-7 */
  double joint_limit_erp;         /* btMultiBodyJointLimitConstraint rows (URDF limits, solo.urdf:47):
                                     positional correction, m_erp = 0.2 */
  double joint_limit_max_impulse; /* btMultiBodyConstraint m_maxAppliedImpulse = 100 */
  int32_t solver_iterations;      /* 50 */
  int32_t num_substeps;           /* physics steps of dt per action (reference: 1) */
  int32_t control_mode;           /* 0 = POSITION_CONTROL, 1 = torque */
  int32_t max_contacts;           /* cap on simultaneous contact points per env, in Bullet's order
                                     (by body, then by support point); <= SOLO_MAX_CONTACTS */
} SoloConfig;

#define SOLO_MAX_CONTACTS 56
#define SOLO_CONTROL_POSITION 0
#define SOLO_CONTROL_TORQUE 1

/* State layout, structure-of-arrays: state[f * n_envs + env], float32.
 *   f 0..2 base COM position, 3..6 base quaternion (x,y,z,w), 7..9 base linear
 *   velocity (world), 10..12 base angular velocity (world), then q[ndof],
 *   qdot[ndof].  == getBasePositionAndOrientation / getBaseVelocity /
 *   getJointState of the reference (obs.py:268-273, 354). */
#define SOLO_STATE_BASE 13
#define SOLO_STATE_DIM(ndof) (SOLO_STATE_BASE + 2 * (ndof))

/* ---- fused epilogue program ------------------------------------------------ */

/* Observation terms (obs.py:193-363), concatenated in registration order. */
#define SOLO_OBS_TORSO_IMU 1
#define SOLO_OBS_MOTOR_ENCODER 2
typedef struct SoloObsTerm {
  int32_t kind;
  int32_t degrees;        /* obs.py:275-279, 357-358 */
  int32_t clip;           /* MotorEncoder: max_rotation given (obs.py:360-361) */
  int32_t normalize;      /* ObservationFactory(normalize=True) (obs.py:148-152) */
  float clip_lo[SOLO_MAX_API], clip_hi[SOLO_MAX_API];   /* np.clip bounds */
  float norm_lo[SOLO_MAX_API], norm_hi[SOLO_MAX_API];   /* float32 Box low/high */
} SoloObsTerm;

/* Reward program: postfix (RPN) over a small stack.
 * RewardFactory.get_reward = sum_i w_i r_i (rewards.py:104-118);
 * AdditiveReward (rewards.py:121-160); MultiplicitiveReward (rewards.py:163-199). */
#define SOLO_REW_UPRIGHT 1      /* rewards.py:221-234 */
#define SOLO_REW_FLAT_TORSO 2   /* rewards.py:237-269: p0 hard, p1 soft */
#define SOLO_REW_SMALL_CONTROL 3/* rewards.py:272-301: p1 margin */
#define SOLO_REW_HSPEED 4       /* rewards.py:304-338: p0 lo, p1 hi, p2 margin */
#define SOLO_REW_HEIGHT 5       /* rewards.py:341-373: p0 lo, p1 hi, p2 margin */
#define SOLO_REW_CONST 6        /* push p0 */
#define SOLO_REW_ADD 16         /* pop n, push sum_k c_k * x_k with c in coef[] */
#define SOLO_REW_MUL 17         /* pop n, push p0 * prod x_k */
typedef struct SoloRewardOp {
  int32_t op;
  int32_t n;              /* arity for ADD / MUL */
  float p[4];             /* leaf parameters: lower, upper, margin, scale */
  float coef[SOLO_REWARD_STACK];
} SoloRewardOp;

/* Termination terms (termination.py:59-98), OR-combined with short circuit. */
#define SOLO_TERM_PERPETUAL 1
#define SOLO_TERM_TIME 2        /* max_step_delta; one int32 counter per env */
typedef struct SoloTermTerm {
  int32_t kind;
  int32_t max_step_delta;
} SoloTermTerm;

typedef struct SoloProgram {
  int32_t n_obs, n_rew, n_term;
  int32_t auto_reset;     /* 1: envs that terminate are restored from the snapshot in-kernel */
  SoloObsTerm obs[SOLO_MAX_OBS_TERMS];
  SoloRewardOp rew[SOLO_MAX_REWARD_OPS];
  SoloTermTerm term[SOLO_MAX_TERM_TERMS];
} SoloProgram;

/* Camera and shading of solo_render.  view / proj are the 16-float column-major
 * matrices pybullet's computeViewMatrixFromYawPitchRoll / computeProjectionMatrixFOV
 * return (gym_solo/envs/solo8_base_env.py:129-135); proj must be a centred
 * perspective frustum.  Colours are linear 0..1. */
typedef struct SoloCamera {
  float view[16];
  float proj[16];
  float light_direction[3];   /* from the scene towards the light, any length */
  float ambient, diffuse;     /* pixel = rgb * (ambient + diffuse * max(0, n.l)) */
  float background[3];
  float plane_rgb_a[3], plane_rgb_b[3];   /* checkerboard of the ground plane z = 0 */
  float checker_size;         /* metres per cell */
  int32_t width, height;
} SoloCamera;

typedef struct SoloHandle_ *SoloHandle;

/* BulletClient(DIRECT) + setGravity + setPhysicsEngineParameter + loadURDF +
 * changeDynamics  (solo8_base_env.py:34-48, solo8v2vanilla.py:145-172). */
int solo_create(const SoloModel *model, const SoloConfig *config, int32_t n_envs,
                int32_t device, SoloHandle *out);
/* client.disconnect (solo8_base_env.py:150-152). */
int solo_destroy(SoloHandle h);
/* obs_factory.register_observation / reward_factory.register_reward /
 * termination_factory.register_termination, compiled to tables
 * (obs.py:109-128, rewards.py:87-102, termination.py:29-36). */
int solo_set_program(SoloHandle h, const SoloProgram *program);
/* len(observation_space.low) for the current program (obs.py:161-190). */
int solo_obs_dim(SoloHandle h);
/* Bytes of solver scratch the caller must bind (256-byte aligned): the
 * per-environment records setup leaves for the solver warp, mass-matrix
 * factors, solver output and the support-point table of one step, and the
 * persistent solver's two work counters.  solo_bind fills the table and zeroes
 * the counters; between calls the scratch belongs to the library (the counters
 * must still be zero at the next step) and to ONE stream at a time. */
int64_t solo_scratch_bytes(SoloHandle h);

/* Bind caller-owned device buffers.  state_dev: float[SOLO_STATE_DIM][n_envs];
 * counters_dev: int32[SOLO_MAX_TERM_TERMS][n_envs]; snapshot_dev:
 * float[SOLO_STATE_DIM] post-reset state (solo8v2vanilla.py:124-136 computed
 * once); stats_dev: float[5][n_envs] (may be NULL): running return and length
 * of the episode in progress, then totals over FINISHED episodes -- sum of
 * returns, number of episodes, sum of lengths; scratch_dev: solo_scratch_bytes() bytes (may be NULL
 * when that is 0). */
int solo_bind(SoloHandle h, float *state_dev, int32_t *counters_dev,
              float *snapshot_dev, float *stats_dev, void *scratch_dev);

/* Physics only: setJointMotorControlArray + stepSimulation
 * (solo8v2vanilla.py:87-91), `n_steps` times with the same command.
 * actions_dev: float[n_envs][n_api], or float[n_api] broadcast when
 * `broadcast` != 0 (the 500 settle steps of reset, solo8v2vanilla.py:130-136). */
int solo_simulate(SoloHandle h, const float *actions_dev, int32_t broadcast,
                  int32_t n_steps, void *stream);

/* One env.step for every env (solo8v2vanilla.py:72-102): optional action
 * scaling `a * action_scale` (normalize_actions, :84-85; pass 1.0f otherwise),
 * motor command, num_substeps physics steps, observations, reward,
 * termination, optional in-kernel auto reset.  Outputs: obs_dev
 * float[n_envs][obs_dim], reward_dev float[n_envs], done_dev uint8[n_envs];
 * terminal_obs_dev (may be NULL) receives the pre-reset observation when
 * auto_reset is on. */
int solo_step(SoloHandle h, const float *actions_dev, float action_scale,
              float *obs_dev, float *reward_dev, uint8_t *done_dev,
              float *terminal_obs_dev, void *stream);

/* obs_factory.get_obs() alone (reset's return value, solo8v2vanilla.py:141-143). */
int solo_observe(SoloHandle h, float *obs_dev, void *stream);
/* reward_factory.get_reward() alone on the current state. */
int solo_reward(SoloHandle h, float *reward_dev, void *stream);

/* env.reset (solo8v2vanilla.py:104-143) by snapshot restore.  mask_dev: uint8
 * per env (NULL = all).  Also termination_factory.reset() (:137). */
int solo_reset(SoloHandle h, const uint8_t *mask_dev, void *stream);

/* Same as solo_step but with HOST buffers (the reference's calling convention:
 * numpy in, numpy out) and a stream synchronise before it returns.
 *  - Page-locked buffers (cudaHostAlloc / cudaHostRegister / torch pin_memory();
 *    all four): the kernels read the actions from and write obs / reward / done
 *    to the caller's memory directly, so the bytes cross the bus inside the step,
 *    overlapped with the kernels.  The device aliases of the three result
 *    buffers are cached while the same three pointers are passed: free or
 *    unpin them only after passing different ones (or destroying the handle).
 *  - Pageable buffers: H2D of actions, the step, D2H of obs/reward/done through
 *    the staging buffers bound with solo_bind_staging. */
int solo_bind_staging(SoloHandle h, float *actions_dev, float *obs_dev,
                      float *reward_dev, uint8_t *done_dev);
int solo_step_host(SoloHandle h, const float *actions_host, float action_scale,
                   float *obs_host, float *reward_host, uint8_t *done_host,
                   void *stream);

/* Execution options (A/B measurements, debugging); the first two default to 1.
 *  SOLO_OPT_GRAPH      solo_step / solo_step_host replay one captured CUDA graph per
 *                      set of output buffers instead of 3 x num_substeps launches
 *  SOLO_OPT_ZERO_COPY  solo_step_host on page-locked buffers without staging copies
 *  SOLO_OPT_SOLVER_OCCUPANCY  which build of the solver kernel runs: 0 = one block per
 *                      environment (latency: the launch is its slowest environment), 2 =
 *                      persistent: one resident wave of warps that draw environments from a
 *                      counter and prefetch the next record (many waves), -1 = by batch size
 *                      (default: 2 from 28672 environments); 1 is accepted and means 0 (it used
 *                      to select a 128-register build that no longer differs).  Same results
 *                      bit for bit.
 *  SOLO_OPT_FINISH_WARMUP  1 = finish_kernel, launched as a programmatic dependent of the solver,
 *                      runs one dry pass (no side effects) while the solver drains so that its
 *                      code and factors are warm for the pass that counts; 0 = off; -1 = by batch
 *                      size (default: on below 10240 environments) */
#define SOLO_OPT_GRAPH 1
#define SOLO_OPT_ZERO_COPY 2
#define SOLO_OPT_SOLVER_OCCUPANCY 3
#define SOLO_OPT_FINISH_WARMUP 4
int solo_set_option(SoloHandle h, int32_t option, int32_t value);

/* Layout of the snapshot buffer bound with solo_bind: per_env = 0 (default) one
 * reset state float[state_dim] shared by every environment -- the reference's
 * reset() is deterministic (solo8v2vanilla.py:104-143); per_env = 1 one reset state
 * per environment, float[state_dim][n_envs], for start poses that differ per
 * environment.  solo_reset, the in-kernel auto reset and the NaN
 * guard all restore from it. */
int solo_snapshot_layout(SoloHandle h, int32_t per_env);
/* Diagnostics: per-step solver statistics of the last solo_step/solo_simulate
 * (sum over envs of contact points, PGS iterations run, NaN resets), written to
 * a caller device buffer int64[4] when bound (NULL to disable). */
int solo_bind_diag(SoloHandle h, unsigned long long *diag_dev);
/* Per-environment lateral friction: float[n_envs] on the device,
 * entry e = friction coefficient of every non-base link of environment e times
 * the plane's (what changeDynamics(lateralFriction=...) sets for joints 0..n-1,
 * solo8v2vanilla.py:157-163, combined as Bullet does with the plane's); the base
 * keeps its model value.  NULL (default): the model's values for every env. */
int solo_bind_link_friction(SoloHandle h, const float *link_friction_dev);
/* Per-environment payload (domain randomisation): scale_dev[n_envs]
 * (device, caller-owned, must stay alive) multiplies the BASE body's mass and
 * inertia tensor in environment e; NULL restores the model's values.  The
 * reference has one robot (changeDynamics(mass=...) would be its call). */
int solo_bind_base_mass_scale(SoloHandle h, const float *scale_dev);
/* Parity harness: per-step contact normal force per support point
 * (float[n_cand][n_envs], == getContactPoints()[9]) and applied motor torque per
 * dof (float[ndof][n_envs], == getJointState()[3]) of the last physics step.
 * NULL disables; binding either selects the instrumented kernel variant. */
int solo_bind_debug(SoloHandle h, float *normal_force_dev, float *motor_torque_dev);
/* Threads per block of the one-thread-per-environment kernels (setup, finish). */
int solo_block_size(SoloHandle h);
/* Per-environment solver statistics of the last physics step, float[4][n_envs]
 * inside the bound scratch: PGS iterations, contact points, contact points that
 * carried an impulse, SM cycles spent in the solve. */
int solo_solver_stats(SoloHandle h, const float **stats_dev);
/* render() -- gym_solo/envs/solo8_base_env.py:116-142 (getCameraImage).  The
 * mesh is a triangle soup in BODY coordinates (body 0 = base inertial frame, body
 * d + 1 = link of dof d), device pointers owned by the caller: tri_body
 * int32[n_tri], tri_verts float[n_tri][9], tri_rgba float[n_tri][4].
 * solo_render rasterises environments [env_begin, env_begin + env_count) at their
 * current state into rgba_dev, uint8[env_count][height][width][4] (device), with two
 * launches on `stream`; scratch_dev holds solo_render_scratch_bytes(h, env_count)
 * bytes, 256-byte aligned.  depth_dev (float[env_count][height][width], OpenGL depth in
 * [0, 1], 1 = nothing hit) and seg_dev (int32, same shape: plane 0, robot 1, nothing -1)
 * are getCameraImage's other two images; either may be NULL.  No graphics API, no host
 * round trip. */
int solo_bind_mesh(SoloHandle h, const int32_t *tri_body_dev, const float *tri_verts_dev,
                   const float *tri_rgba_dev, int32_t n_tri);
int64_t solo_render_scratch_bytes(SoloHandle h, int32_t env_count);
int solo_render(SoloHandle h, const SoloCamera *camera, int32_t env_begin, int32_t env_count,
                uint8_t *rgba_dev, float *depth_dev, int32_t *seg_dev, void *scratch_dev, void *stream);
/* solo_render without caller scratch and without any device allocation.  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same images as
 * solo_render (the same two kernels); the projected-triangle records and bounding boxes live in a
 * file-scope `static __device__` array of SOLO_RENDER_STATIC_BYTES that every GPU loading the
 * library holds until exit.  A call whose solo_render_scratch_bytes(h, env_count) exceeds that
 * (59 environments of the 1100-triangle solo8v2 mesh) fails with SOLO_EINVAL and says so in
 * solo_last_error.  Calls no cudaMalloc, cudaFree or other allocating routine; the first call on a device
 * resolves the array's address (cudaGetSymbolAddress) -- make it before a graph capture.  Every
 * call on a device shares the array, which keeps the last call's contents (what must start clean
 * is reset on `stream` inside the call): calls that overlap on the GPU -- different streams,
 * different handles -- race; order them. */
#define SOLO_RENDER_STATIC_BYTES (4 * 1024 * 1024)
int solo_render_static(SoloHandle h, const SoloCamera *camera, int32_t env_begin, int32_t env_count,
                       uint8_t *rgba_dev, float *depth_dev, int32_t *seg_dev, void *stream);
/* Live per-kernel timing (bench.py roofline): while enabled every physics step
 * records CUDA events on the launching stream around its three kernels;
 * solo_phase_times synchronises on them and returns the number of physics
 * steps timed since enabling, with the summed setup / solve / finish
 * milliseconds in ms3[0..2]. */
int solo_phase_timing(SoloHandle h, int32_t enable);
int64_t solo_phase_times(SoloHandle h, double *ms3);
/* How many kernels this handle has launched so far (bench `gpu_launches`). */
int64_t solo_launch_count(SoloHandle h);
const char *solo_last_error(SoloHandle h);
const char *solo_version(void);

#ifdef __cplusplus
}
#endif
#endif  /* SOLO8_B200_H_ */
