// solo_kernels.cuh -- sm_100a kernels of the Solo 8 env-step path.
//
// One env.step = three launches on the caller's stream (solo_phases.h):
//
//   setup_legpar_kernel (solo_setup_legpar.cuh)  one lane per (environment, leg):
//                  the part of a step whose size does not depend on the pose.
//                  Reads the action tile and the SoA state, runs kinematics, bias
//                  forces, the mass-matrix factorisation and the motor rows (legs in
//                  parallel, their sums by xor-shuffle, the 6 x 6 base part on all
//                  four lanes); writes the factors [k][env] for finish_kernel and one
//                  contiguous KinLayout record per environment for the solver warp.
//                  (setup_step of solo_setup.h is the one-thread-per-environment
//                  statement of the same phase, which tests/hostsim runs on the CPU;
//                  as a kernel it exists only in -DSOLO_EXPERIMENTS builds.)
//   solve_warp_kernel (solo_solve_warp.cuh)  one warp per environment: contact
//                  detection and contact rows one contact per lane, then the
//                  data-dependent part: projected Gauss-Seidel, a serial chain
//                  of 8 + 3 n_contacts row updates x 2..50 iterations.
//   finish_kernel  one thread per environment: dv = T^T yh, integration, NaN
//                  guard, observations, reward, termination, optional snapshot
//                  auto-reset; writes the SoA state and the [env][k] outputs
//                  (transposed through shared memory into whole lines).
//
// Constant tables (model, config, epilogue program) travel as ONE
// __grid_constant__ kernel parameter: warp-uniform reads are LDC / direct FFMA
// operands, there is no module-global state, and handles with different robots
// coexist in one process.  Intermediates (rows, factors, yh) live in the
// caller-bound scratch and stay L2-resident for the batch sizes of the bench.
#ifndef SOLO_KERNELS_CUH_
#define SOLO_KERNELS_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include "solo_phases.h"

namespace solo {

struct KParams {
  ModelT<float> M;
  ConfigT<float> C;
  ProgramC P;
};
static_assert(sizeof(KParams) < 16384, "constant tables must fit the kernel parameter bank");

struct Buffers {
  float *state;              // [state_dim][n]
  int32_t *counters;         // [kMaxTermTerms][n]
  const float *snapshot;     // [state_dim], or [state_dim][n] when snapshot_per_env (solo_snapshot_layout)
  float *stats;              // [5][n] or null: running return, running length, sum of finished returns,
                             // finished episodes, sum of finished lengths
  unsigned long long *diag;  // [4] or null: contacts, PGS iterations, NaN restores, physics steps
  float *dbg_normal_force;   // [n_cand][n] or null
  float *dbg_motor_torque;   // [ndof][n] or null
  const float *base_mass_scale; // [n] or null: per-env factor on the base body's mass and inertia (solo_bind_base_mass_scale)
  const float *link_friction; // [n] or null: per-env friction of every non-base link x plane friction (solo_bind_link_friction)
  // solver scratch (caller-bound, solo_scratch_bytes)
  const float *cand_tab;     // [kMaxCand][kCandStride] support point xyz, friction, leg, depth
  float *yh;                 // [6 + kLegs * CH + 4][n]: solver output, then per-env solver statistics
  float *fac;                // [FactorLayout<CH>::kSize][n]
  float *kin;                // [n][KinLayout<CH>::kSize]: one contiguous record per env
  int32_t *work;             // [2] zero between launches: the persistent solver's ticket counter and its check-out count
  int32_t n;
  int32_t obs_dim;
  int32_t snapshot_per_env;  // 0: one reset state for every env; 1: one per env (randomised start poses)
};

template <int CH>
__device__ __forceinline__ bool env_finite(const EnvState<float, CH> &s) {
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 3; i++) acc += (s.pos[i] + s.vlin[i] + s.vang[i]) * 0.f;
#pragma unroll
  for (int i = 0; i < 4; i++) acc += s.quat[i] * 0.f;
#pragma unroll
  for (int l = 0; l < kLegs; l++)
#pragma unroll
    for (int j = 0; j < CH; j++) acc += (s.q[l][j] + s.qd[l][j]) * 0.f;
  return acc == 0.f;   // NaN/Inf propagate, finite -> 0
}

template <int CH>
__device__ __forceinline__ void load_snapshot(const ModelT<float> &M, const Buffers &B, const int64_t env,
                                              EnvState<float, CH> &s) {
  if (B.snapshot_per_env) load_state<float, CH, float>(M, B.snapshot + env, B.n, s);
  else load_state<float, CH, float>(M, B.snapshot, 1, s);
}

// warp-level reduction, one atomic per warp and counter
__device__ __forceinline__ void diag_add(unsigned long long *diag, int contacts, int iters, int nan, int steps) {
  const unsigned mask = __activemask();
  for (int off = 16; off > 0; off >>= 1) {
    contacts += __shfl_down_sync(mask, contacts, off);
    iters += __shfl_down_sync(mask, iters, off);
    nan += __shfl_down_sync(mask, nan, off);
    steps += __shfl_down_sync(mask, steps, off);
  }
  if ((threadIdx.x & 31) == 0) {
    if (contacts) atomicAdd(diag + 0, (unsigned long long)contacts);
    if (iters) atomicAdd(diag + 1, (unsigned long long)iters);
    if (nan) atomicAdd(diag + 2, (unsigned long long)nan);
    if (steps) atomicAdd(diag + 3, (unsigned long long)steps);
  }
}

// Block-cooperative transpose of a [rows = blockDim.x envs][k] tile between the
// reference's row-major layout in global memory and one value per thread.
// smem holds blockDim.x * k floats; rows are padded to an odd stride to keep the
// per-thread row accesses conflict free.
__device__ __forceinline__ int pad_stride(int k) { return k | 1; }
// i / k for the flat index of a [rows <= 128][k <= 128] tile without an integer division:
// (i + 1/2) / k is at least 1 / (2 k) away from every integer, far more than its float32 error
__device__ __forceinline__ int tile_row(int i, float inv_k) { return (int)(((float)i + 0.5f) * inv_k); }

__device__ __forceinline__ void tile_gather(const float *__restrict__ g, int64_t env0, int n_valid, int k,
                                            float *sm) {
  const int ps = pad_stride(k);
  const int total = n_valid * k;
  const float *src = g + env0 * k;
  const float inv_k = 1.f / (float)k;
  // loads in batches of kU per thread so that their latencies overlap
  constexpr int kU = 6;
  for (int i = threadIdx.x; i < total; i += kU * blockDim.x) {
    float v[kU];
#pragma unroll
    for (int u = 0; u < kU; u++) {
      const int ii = i + u * blockDim.x;
      v[u] = ii < total ? __ldg(src + ii) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < kU; u++) {
      const int ii = i + u * blockDim.x;
      if (ii < total) {
        const int r = tile_row(ii, inv_k), c = ii - r * k;
        sm[r * ps + c] = v[u];
      }
    }
  }
  __syncthreads();
}
// The same gather without waiting for it: every element goes global -> shared memory as a 4-byte
// cp.async; tile_gather_wait() before the first read.  The setup kernel issues it first thing and
// needs the actions only for the motor rows at its very end, so the latency of the action bytes
// (over PCIe when the caller's page-locked buffer is read in place) hides behind the dynamics.
__device__ __forceinline__ void tile_gather_async(const float *__restrict__ g, int64_t env0, int n_valid, int k,
                                                  float *sm) {
  const int ps = pad_stride(k);
  const int total = n_valid * k;
  const float *src = g + env0 * k;
  const float inv_k = 1.f / (float)k;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int r = tile_row(i, inv_k), c = i - r * k;
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(sm + r * ps + c);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src + i) : "memory");
  }
}
__device__ __forceinline__ void tile_gather_wait() {
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncthreads();
}

__device__ __forceinline__ void tile_scatter(float *__restrict__ g, int64_t env0, int n_valid, int k,
                                             const float *sm, const bool commit = true) {
  const int ps = pad_stride(k);
  const int total = n_valid * k;
  float *dst = g + env0 * k;
  const float inv_k = 1.f / (float)k;
  __syncthreads();
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int r = tile_row(i, inv_k), c = i - r * k;
    const float v = sm[r * ps + c];
    if (commit) dst[i] = v;
  }
}

#ifdef SOLO_EXPERIMENTS   // one thread per environment: A/B builds only, not in the product library
// ---------------------------------------------------------------------------------
// phase 1: setJointMotorControlArray + the pre-solver half of stepSimulation
// (solo8v2vanilla.py:87-91)
// ---------------------------------------------------------------------------------
template <int CH>
__global__ void __launch_bounds__(128)
setup_kernel(const __grid_constant__ KParams K, const Buffers B, const float *__restrict__ actions,
             const int broadcast, const float action_scale) {
  extern __shared__ float smem[];
  typedef RowLayout<CH> RL;
  typedef KinLayout<CH> KL;
  const ModelT<float> &M = K.M;
  const int64_t env0 = (int64_t)blockIdx.x * blockDim.x;
  const int64_t env = env0 + threadIdx.x;
  const int n_valid = (int)min((int64_t)blockDim.x, (int64_t)B.n - env0);
  const bool live = env < B.n;
  const int n_api = M.n_api;

  // the state loads go out first: they fly while the action tile is gathered
  EnvState<float, CH> s;
  if (live) load_state<float, CH, float>(M, B.state + env, B.n, s);
  float cmd[kLegs][CH];
  if (broadcast) {
    load_command<float, CH, float>(M, actions, action_scale, cmd);
  } else {
    // actions: [env][n_api] -> per-thread command
    tile_gather(actions, env0, n_valid, n_api, smem);
    if (live) load_command<float, CH, float>(M, smem + threadIdx.x * pad_stride(n_api), action_scale, cmd);
  }
  if (!live) return;
  float *kin = B.kin + env * KL::kSize;
  setup_step<float, CH, false>(M, K.C, s, cmd, nullptr, kin + KL::kMotor, RL::kRow, B.fac + env, B.n, kin,
                               B.base_mass_scale ? B.base_mass_scale[env] : 1.f);
}
#endif  // SOLO_EXPERIMENTS

// ---------------------------------------------------------------------------------
// phase 3: velocity update + integration, then (epilogue != 0) the rest of
// env.step (solo8v2vanilla.py:96-102): observations, reward, termination,
// optional in-kernel auto reset.
// ---------------------------------------------------------------------------------
// warm != 0 (the kernel was launched as a programmatic dependent of the solver): the whole body
// runs once BEFORE cudaGridDependencySynchronize() with every side effect switched off, on
// whatever yh holds at that moment.  At small batches one warp per SM walks ~6.6 k instructions
// exactly once and the kernel was bound by instruction fetch (no_instruction + long_scoreboard);
// the dry pass pulls the code and the factor lines in while the solver drains, and the pass that
// counts finds them warm.
template <int CH>
__global__ void __launch_bounds__(128)
finish_kernel(const __grid_constant__ KParams K, const Buffers B, const int epilogue, const int warm,
              float *__restrict__ obs_out, float *__restrict__ reward_out, uint8_t *__restrict__ done_out,
              float *__restrict__ terminal_obs_out) {
  extern __shared__ float smem[];
  const ModelT<float> &M = K.M;
  const ProgramC &P = K.P;
  const int64_t env0 = (int64_t)blockIdx.x * blockDim.x;
  const int64_t env = env0 + threadIdx.x;
  const int n_valid = (int)min((int64_t)blockDim.x, (int64_t)B.n - env0);
  const bool live = env < B.n;
  const int od = B.obs_dim;
  const int ps = pad_stride(od);

#pragma unroll 1
  for (int pass = warm ? 0 : 1; pass < 2; pass++) {
    const bool commit = pass == 1;
    // programmatic dependent of solve_warp_kernel: nothing the solver writes is trusted before this
    if (commit) cudaGridDependencySynchronize();
    EnvState<float, CH> s;
    float ob[kMaxObsDim];
    float rew = 0.f;
    bool done = false;
    int nan = 0;
    if (live) {
      load_state<float, CH, float>(M, B.state + env, B.n, s);
      float yb[6], yl[kLegs][CH];
      const float *y = B.yh + env;
#pragma unroll
      for (int a = 0; a < 6; a++) yb[a] = __ldcg(y + (int64_t)a * B.n);
#pragma unroll
      for (int l = 0; l < kLegs; l++)
#pragma unroll
        for (int k = 0; k < CH; k++) yl[l][k] = __ldcg(y + (int64_t)(6 + l * CH + k) * B.n);
      finish_step<float, CH>(M, K.C, s, B.fac + env, B.n, yb, yl);
      // A non-finite state ends the episode: the environment restarts from its reset snapshot, the
      // step reports done = 1 (epilogue steps), the termination counters and the running episode
      // statistics start over, and diag[2] counts it (pybullet would hand the NaN to the caller).
      if (!env_finite<CH>(s)) {
        load_snapshot<CH>(M, B, env, s);
        nan = 1;
        if (commit) {
          for (int t = 0; t < kMaxTermTerms; t++) B.counters[(int64_t)t * B.n + env] = 0;
          if (B.stats && !epilogue) { B.stats[env] = 0.f; B.stats[(int64_t)B.n + env] = 0.f; }
        }
      }
    }
    if (B.diag && commit) {
      // fold this step's per-env solver statistics into the running totals
      const float *st = B.yh + (int64_t)(6 + kLegs * CH) * B.n + env;
      diag_add(B.diag, live ? (int)__ldcg(st + (int64_t)B.n) : 0, live ? (int)__ldcg(st) : 0, nan, live ? 1 : 0);
    }
    if (!epilogue) {
      if (live && commit) store_state<float, CH, float>(M, s, B.state + env, B.n);
      continue;
    }

    if (live) {
      float e[3];
      euler_from_quat(s.quat, e);
      compute_obs<float, CH>(M, P, s, e, ob);
      rew = P.n_rew > 0 ? compute_reward<float, CH>(M, P, s, e) : 0.f;
      done = eval_terminations(P, B.counters + env, B.n, commit);
      if (nan) {
        done = true;   // restarted from the snapshot above: the new episode has not ticked yet
        if (commit)
          for (int t = 0; t < P.n_term; t++) B.counters[(int64_t)t * B.n + env] = 0;
      }
      if (B.stats) {
        float ret = B.stats[env] + rew;
        float len = B.stats[(int64_t)B.n + env] + 1.f;
        if (commit) {
          if (done) {
            // totals over FINISHED episodes (episode_statistics: mean return / mean length)
            B.stats[2 * (int64_t)B.n + env] += ret;
            B.stats[3 * (int64_t)B.n + env] += 1.f;
            B.stats[4 * (int64_t)B.n + env] += len;
            if (P.auto_reset || nan) { ret = 0.f; len = 0.f; }
          }
          B.stats[env] = ret;
          B.stats[(int64_t)B.n + env] = len;
        }
      }
    }

    if (P.auto_reset) {
      // terminal observation first (gym.vector convention), then the first
      // observation of the new episode from the snapshot
      if (terminal_obs_out) {
        if (live)
          for (int i = 0; i < od; i++) smem[threadIdx.x * ps + i] = ob[i];
        tile_scatter(terminal_obs_out, env0, n_valid, od, smem, commit);
        __syncthreads();
      }
      if (live && done) {
        load_snapshot<CH>(M, B, env, s);
        if (commit)
          for (int t = 0; t < P.n_term; t++) B.counters[(int64_t)t * B.n + env] = 0;
        float e[3];
        euler_from_quat(s.quat, e);
        compute_obs<float, CH>(M, P, s, e, ob);
      }
    }
    if (live) {
      if (commit) {
        store_state<float, CH, float>(M, s, B.state + env, B.n);
        reward_out[env] = rew;
        done_out[env] = done ? 1 : 0;
      }
      for (int i = 0; i < od; i++) smem[threadIdx.x * ps + i] = ob[i];
    }
    tile_scatter(obs_out, env0, n_valid, od, smem, commit);
    __syncthreads();   // the tile is reused by the next pass
  }
}

// obs_factory.get_obs() / reward_factory.get_reward() on the current state
// (solo8v2vanilla.py:141-143; rewards.py:104-118)
template <int CH>
__global__ void __launch_bounds__(128)
observe_kernel(const __grid_constant__ KParams K, const Buffers B, float *__restrict__ obs_out,
               float *__restrict__ reward_out) {
  extern __shared__ float smem[];
  const ModelT<float> &M = K.M;
  const ProgramC &P = K.P;
  const int64_t env0 = (int64_t)blockIdx.x * blockDim.x;
  const int64_t env = env0 + threadIdx.x;
  const int n_valid = (int)min((int64_t)blockDim.x, (int64_t)B.n - env0);
  const bool live = env < B.n;
  const int od = B.obs_dim, ps = pad_stride(od);
  if (live) {
    EnvState<float, CH> s;
    load_state<float, CH, float>(M, B.state + env, B.n, s);
    float e[3];
    euler_from_quat(s.quat, e);
    if (obs_out) {
      float ob[kMaxObsDim];
      compute_obs<float, CH>(M, P, s, e, ob);
      for (int i = 0; i < od; i++) smem[threadIdx.x * ps + i] = ob[i];
    }
    if (reward_out) reward_out[env] = P.n_rew > 0 ? compute_reward<float, CH>(M, P, s, e) : 0.f;
  }
  if (obs_out) tile_scatter(obs_out, env0, n_valid, od, smem);
}

// env.reset by snapshot restore + termination_factory.reset()
// (solo8v2vanilla.py:104-143)
__global__ void reset_kernel(const Buffers B, const int state_dim, const uint8_t *__restrict__ mask) {
  const int64_t env = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= B.n) return;
  if (mask && !mask[env]) return;
  for (int f = 0; f < state_dim; f++)
    B.state[(int64_t)f * B.n + env] = B.snapshot_per_env ? B.snapshot[(int64_t)f * B.n + env] : B.snapshot[f];
  for (int t = 0; t < kMaxTermTerms; t++) B.counters[(int64_t)t * B.n + env] = 0;
  if (B.stats) { B.stats[env] = 0.f; B.stats[(int64_t)B.n + env] = 0.f; }
}

}  // namespace solo
#endif  // SOLO_KERNELS_CUH_
