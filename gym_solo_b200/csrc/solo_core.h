// solo_core.h -- per-environment arithmetic of the Solo 8 env step, written once
// as __host__ __device__ templates.  The CUDA kernels (solo_kernels.cu) run it
// in float32 with one thread per environment; tests/hostsim compiles the very
// same code for the host CPU so that the kernel's arithmetic can be checked
// against the oracle without a GPU (test infrastructure only -- the product
// path is the CUDA library and has no CPU fallback).
//
// Replaces, per environment (reference file:line):
//   setJointMotorControlArray + stepSimulation   gym_solo/envs/solo8v2vanilla.py:87-91
//   ObservationFactory.get_obs / TorsoIMU / MotorEncoder   gym_solo/core/obs.py:130-159, 260-282, 347-363
//   RewardFactory.get_reward and the reward classes        gym_solo/core/rewards.py:104-118, 121-431
//   TerminationFactory.is_terminated                       gym_solo/core/termination.py:38-98
//
// Formulation (NOT the oracle's: the two are meant to check each other).
// Generalised velocity v = (v_com0, omega0, qdot), world frame.  The robot is a
// star: a floating base plus n_legs serial chains of <= CH revolute joints, so
//   M = [ Mbb  B ; B^T  D ],  D = blockdiag(D_leg).
// Per leg: D = C C^T, E = C^-1, Bh = E B^T; base Schur complement
// S = Mbb - sum Bh^T Bh = Cs Cs^T, Ls = Cs^-1.  With
//   T = [ Ls  -Ls Bh^T E ... ; 0  E ]   we have   M^-1 = T^T T,
// so every constraint row J is stored whitened, Jh = T J^T (6 base entries + CH
// entries of its own leg), the PGS state is yh = sum Jh * delta, the row
// velocity is J dv = Jh . yh and the update is yh += Jh * delta: 8 FMAs each
// instead of Bullet's 8 + 14, with the same iterates in exact arithmetic.
// Unconstrained accelerations use the same factors: a = T^T (T f), f = tau - C
// with C from a Newton-Euler pass at zero generalised acceleration.
#ifndef SOLO_CORE_H_
#define SOLO_CORE_H_

#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define SOLO_HD __host__ __device__ __forceinline__
#else
#define SOLO_HD inline
#endif

namespace solo {

constexpr int kLegs = 4;
constexpr int kMaxDof = 12;
constexpr int kMaxBodies = 13;
constexpr int kMaxApi = 16;
constexpr int kMaxCand = 128;
// simultaneous contact points per environment (SoloConfig.max_contacts <= this): the solver
// warp's shared-memory array of active contacts is sized by it, and 56 entries (16 KB per block
// with the record) let 13 one-warp blocks share an SM
constexpr int kMaxContacts = 56;
constexpr int kStateBase = 13;
constexpr int kMaxObsTerms = 8;
constexpr int kMaxRewardOps = 32;
constexpr int kMaxTermTerms = 8;
constexpr int kRewardStack = 8;
constexpr int kMaxObsDim = kMaxObsTerms * kMaxApi;

// ---- constant tables (device: __constant__; host: plain struct) --------------
template <typename T>
struct ModelT {
  int32_t n_legs, ndof, n_api, n_cand;
  int32_t chain_len[kLegs];
  int32_t dof0[kLegs];               // first dof of each leg
  int32_t api_dof[kMaxApi];          // pybullet joint index -> dof or -1
  int32_t dof_api[kMaxDof];          // dof -> pybullet joint index
  int32_t cand_begin[kMaxBodies + 1];  // candidates are grouped by body
  T api_lower[kMaxApi], api_upper[kMaxApi];
  T mass[kMaxBodies];
  T com[kMaxBodies][3];
  T inertia[kMaxBodies][6];          // xx xy xz yy yz zz, body frame, about COM
  T jorigin[kMaxDof][3];
  T jrot[kMaxDof][9];
  T jaxis[kMaxDof][3];
  T cand_pos[kMaxCand][3];
  T cand_mu[kMaxCand];               // link friction * plane friction
  T margin;
};

template <typename T>
struct ConfigT {
  T dt, inv_dt;
  T gravity[3];
  T torque_limit, lin_damp, ang_damp;
  T kp, kd, erp, slop, breaking, max_vel, res_thr;
  T lim_erp, lim_max;                // joint-limit rows: positional correction, impulse cap
  int32_t iters, substeps, control_mode, max_contacts;
};

struct ObsTermC {
  int32_t kind, degrees, clip, normalize;
  float clip_lo[kMaxApi], clip_hi[kMaxApi], norm_lo[kMaxApi], norm_hi[kMaxApi];
};
struct RewardOpC {
  int32_t op, n;
  float p[4];
  float coef[kRewardStack];
};
struct TermTermC {
  int32_t kind, max_step_delta;
};
struct ProgramC {
  int32_t n_obs, n_rew, n_term, auto_reset;
  ObsTermC obs[kMaxObsTerms];
  RewardOpC rew[kMaxRewardOps];
  TermTermC term[kMaxTermTerms];
};

enum { kObsTorsoImu = 1, kObsMotorEncoder = 2 };
enum { kRewUpright = 1, kRewFlatTorso = 2, kRewSmallControl = 3, kRewHSpeed = 4,
       kRewHeight = 5, kRewConst = 6, kRewAdd = 16, kRewMul = 17 };
enum { kTermPerpetual = 1, kTermTime = 2 };

// ---- scalar helpers -----------------------------------------------------------
SOLO_HD float m_sqrt(float x) { return sqrtf(x); }
SOLO_HD double m_sqrt(double x) { return sqrt(x); }
SOLO_HD float m_rsqrt(float x) {
#ifdef __CUDA_ARCH__
  return rsqrtf(x);
#else
  return 1.0f / sqrtf(x);
#endif
}
SOLO_HD double m_rsqrt(double x) { return 1.0 / sqrt(x); }
SOLO_HD float m_abs(float x) { return fabsf(x); }
SOLO_HD double m_abs(double x) { return fabs(x); }
SOLO_HD float m_atan2(float y, float x) { return atan2f(y, x); }
SOLO_HD double m_atan2(double y, double x) { return atan2(y, x); }
SOLO_HD float m_asin(float x) { return asinf(x); }
SOLO_HD double m_asin(double x) { return asin(x); }
SOLO_HD float m_exp(float x) { return expf(x); }
SOLO_HD double m_exp(double x) { return exp(x); }
SOLO_HD float m_log(float x) { return logf(x); }
SOLO_HD double m_log(double x) { return log(x); }
SOLO_HD void m_sincos(float x, float *s, float *c) {
#ifdef __CUDA_ARCH__
  sincosf(x, s, c);
#else
  *s = sinf(x); *c = cosf(x);
#endif
}
SOLO_HD void m_sincos(double x, double *s, double *c) { *s = sin(x); *c = cos(x); }
template <typename T> SOLO_HD T m_min(T a, T b) { return a < b ? a : b; }
template <typename T> SOLO_HD T m_max(T a, T b) { return a > b ? a : b; }
template <typename T> SOLO_HD T m_clamp(T x, T lo, T hi) { return x > hi ? hi : (x < lo ? lo : x); }

template <typename T>
struct V3 {
  T x, y, z;
};
template <typename T> SOLO_HD V3<T> mk(T x, T y, T z) { V3<T> r; r.x = x; r.y = y; r.z = z; return r; }
template <typename T> SOLO_HD V3<T> operator+(V3<T> a, V3<T> b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> SOLO_HD V3<T> operator-(V3<T> a, V3<T> b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> SOLO_HD V3<T> operator*(V3<T> a, T s) { return mk(a.x * s, a.y * s, a.z * s); }
template <typename T> SOLO_HD V3<T> cross(V3<T> a, V3<T> b) {
  return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
template <typename T> SOLO_HD T dot(V3<T> a, V3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> SOLO_HD V3<T> ld3(const T *p) { return mk(p[0], p[1], p[2]); }

// 3x3 row-major
template <typename T>
struct M3 {
  T m[9];
};
template <typename T> SOLO_HD V3<T> mul(const M3<T> &A, V3<T> v) {
  return mk(A.m[0] * v.x + A.m[1] * v.y + A.m[2] * v.z, A.m[3] * v.x + A.m[4] * v.y + A.m[5] * v.z,
            A.m[6] * v.x + A.m[7] * v.y + A.m[8] * v.z);
}
template <typename T> SOLO_HD M3<T> mul(const M3<T> &A, const M3<T> &B) {
  M3<T> C;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++)
      C.m[3 * i + j] = A.m[3 * i] * B.m[j] + A.m[3 * i + 1] * B.m[3 + j] + A.m[3 * i + 2] * B.m[6 + j];
  return C;
}
// symmetric 3x3: xx xy xz yy yz zz
template <typename T>
struct S3 {
  T xx, xy, xz, yy, yz, zz;
};
template <typename T> SOLO_HD V3<T> mul(const S3<T> &A, V3<T> v) {
  return mk(A.xx * v.x + A.xy * v.y + A.xz * v.z, A.xy * v.x + A.yy * v.y + A.yz * v.z,
            A.xz * v.x + A.yz * v.y + A.zz * v.z);
}
// R I R^T for symmetric I
template <typename T> SOLO_HD S3<T> rotate_inertia(const M3<T> &R, const T *I6) {
  S3<T> I; I.xx = I6[0]; I.xy = I6[1]; I.xz = I6[2]; I.yy = I6[3]; I.yz = I6[4]; I.zz = I6[5];
  // columns of R^T are rows of R: t_i = I * row_i(R)
  V3<T> r0 = mk(R.m[0], R.m[1], R.m[2]), r1 = mk(R.m[3], R.m[4], R.m[5]), r2 = mk(R.m[6], R.m[7], R.m[8]);
  V3<T> t0 = mul(I, r0), t1 = mul(I, r1), t2 = mul(I, r2);
  S3<T> o;
  o.xx = dot(r0, t0); o.xy = dot(r0, t1); o.xz = dot(r0, t2);
  o.yy = dot(r1, t1); o.yz = dot(r1, t2); o.zz = dot(r2, t2);
  return o;
}
// Composite (mass m, first moment h, inertia I) about a reference point -> about the point
// that sees every mass at r + d (new reference = old reference - d).
template <typename T> SOLO_HD void shift_composite(const T m, V3<T> &h, S3<T> &I, const V3<T> d) {
  const T s = 2 * dot(h, d) + m * dot(d, d);
  I.xx += s - 2 * h.x * d.x - m * d.x * d.x;
  I.yy += s - 2 * h.y * d.y - m * d.y * d.y;
  I.zz += s - 2 * h.z * d.z - m * d.z * d.z;
  I.xy -= h.x * d.y + d.x * h.y + m * d.x * d.y;
  I.xz -= h.x * d.z + d.x * h.z + m * d.x * d.z;
  I.yz -= h.y * d.z + d.y * h.z + m * d.y * d.z;
  h = h + d * m;
}
template <typename T> SOLO_HD M3<T> quat_to_mat(const T *q) {
  T x = q[0], y = q[1], z = q[2], w = q[3];
  T s = T(2) / (x * x + y * y + z * z + w * w);
  M3<T> R;
  R.m[0] = 1 - s * (y * y + z * z); R.m[1] = s * (x * y - w * z); R.m[2] = s * (x * z + w * y);
  R.m[3] = s * (x * y + w * z); R.m[4] = 1 - s * (x * x + z * z); R.m[5] = s * (y * z - w * x);
  R.m[6] = s * (x * z - w * y); R.m[7] = s * (y * z + w * x); R.m[8] = 1 - s * (x * x + y * y);
  return R;
}
template <typename T> SOLO_HD M3<T> axis_angle(V3<T> a, T th) {
  T s, c;
  m_sincos(th, &s, &c);
  T t = 1 - c;
  M3<T> R;
  R.m[0] = t * a.x * a.x + c;       R.m[1] = t * a.x * a.y - s * a.z; R.m[2] = t * a.x * a.z + s * a.y;
  R.m[3] = t * a.x * a.y + s * a.z; R.m[4] = t * a.y * a.y + c;       R.m[5] = t * a.y * a.z - s * a.x;
  R.m[6] = t * a.x * a.z - s * a.y; R.m[7] = t * a.y * a.z + s * a.x; R.m[8] = t * a.z * a.z + c;
  return R;
}

// Joint limit of one dof (URDF <limit lower upper>, solo.urdf:47; Bullet's
// btMultiBodyJointLimitConstraint): its row exists only while the limit is reached or
// violated.  Returns 0 (no row), +1 (lower limit, *pen = q - lower <= 0) or -1 (upper limit,
// *pen = upper - q <= 0).  lower > upper means "no limit" (continuous joint).
template <typename T> SOLO_HD int joint_limit_state(const T q, const T lower, const T upper, T *pen) {
  *pen = 0;
  if (!(lower <= upper)) return 0;
  const T lo = q - lower, hi = upper - q;
  if (lo <= 0) { *pen = lo; return 1; }
  if (hi <= 0) { *pen = hi; return -1; }
  return 0;
}

// ---- per-environment state (registers) ---------------------------------------
template <typename T, int CH>
struct EnvState {
  T pos[3], quat[4], vlin[3], vang[3];
  T q[kLegs][CH], qd[kLegs][CH];
};

struct StepDiag {
  int32_t n_contacts, iterations;
};

// Optional per-env debug outputs (parity harness): normal force per candidate
// and applied motor torque per dof.  Pointers may be null.
template <typename T>
struct DebugOut {
  T *normal_force;  // [n_cand] stride `stride`
  T *motor_torque;  // [ndof]   stride `stride`
  int64_t stride;
};

// Cholesky factor inverse of a symmetric positive definite N x N matrix held as
// a full array (lower part read).  Returns L = C^-1 (lower triangular) in A.
template <typename T, int N> SOLO_HD void chol_inverse_lower(T (&A)[N][N]) {
  T rd[N];
#pragma unroll
  for (int j = 0; j < N; j++) {
    T d = A[j][j];
#pragma unroll
    for (int k = 0; k < j; k++) d -= A[j][k] * A[j][k];
    d = m_sqrt(m_max(d, T(1e-30)));
    rd[j] = T(1) / d;
    A[j][j] = d;
#pragma unroll
    for (int i = j + 1; i < N; i++) {
      T s = A[i][j];
#pragma unroll
      for (int k = 0; k < j; k++) s -= A[i][k] * A[j][k];
      A[i][j] = s * rd[j];
    }
  }
  // invert the lower-triangular factor in place, column by column
#pragma unroll
  for (int j = 0; j < N; j++) {
    T col[N];
    col[j] = rd[j];
#pragma unroll
    for (int i = j + 1; i < N; i++) {
      T s = 0;
#pragma unroll
      for (int k = j; k < i; k++) s -= A[i][k] * col[k];
      col[i] = s * rd[i];
    }
#pragma unroll
    for (int i = j; i < N; i++) A[i][j] = col[i];
  }
}

// ---- pybullet getters restated -------------------------------------------------
// getEulerFromQuaternion (obs.py:271, rewards.py:233,265); quaternion NOT normalised
// first (golden: test_obs_observations.py:67-88).
template <typename T> SOLO_HD void euler_from_quat(const T *q, T *e) {
  const T x = q[0], y = q[1], z = q[2], w = q[3];
  const T sarg = T(-2) * (x * z - w * y);
  const T sqx = x * x, sqy = y * y, sqz = z * z, sqw = w * w;
  if (sarg <= T(-0.99999)) {
    e[0] = 0; e[1] = T(-0.5 * 3.14159265358979323846); e[2] = 2 * m_atan2(x, -y);
  } else if (sarg >= T(0.99999)) {
    e[0] = 0; e[1] = T(0.5 * 3.14159265358979323846); e[2] = 2 * m_atan2(-x, y);
  } else {
    e[0] = m_atan2(2 * (y * z + w * x), sqw - sqx - sqy + sqz);
    e[1] = m_asin(sarg);
    e[2] = m_atan2(2 * (x * y + w * z), sqw + sqx - sqy - sqz);
  }
}

// rewards.py:381-431 (tolerance -> gaussian); margin_value 0.1 in every call site.
template <typename T> SOLO_HD T gaussian_tol(T x, T lower, T upper, T margin) {
  const bool within = (lower <= x) && (x <= upper);
  if (within) return T(1);
  if (margin == 0) return T(0);
  const T d = ((x < lower) ? lower - x : x - upper) / margin;
  // exp(-0.5 (d * sqrt(-2 ln 0.1))^2) = exp(ln(0.1) d^2)
  return m_exp(T(-2.302585092994046) * d * d);
}

// Observations of the current state -> out[obs_dim] (stride 1).  Returns dim.
template <typename T, int CH>
SOLO_HD int compute_obs(const ModelT<T> &M, const ProgramC &P, const EnvState<T, CH> &s, const T *euler, T *out) {
  const T rad2deg = T(57.29577951308232);
  int o = 0;
  for (int t = 0; t < P.n_obs; t++) {
    const ObsTermC &ot = P.obs[t];
    if (ot.kind == kObsTorsoImu) {
      T v[9] = {euler[0], euler[1], euler[2], s.vlin[0], s.vlin[1], s.vlin[2], s.vang[0], s.vang[1], s.vang[2]};
      if (ot.degrees) {
#pragma unroll
        for (int i = 0; i < 3; i++) { v[i] *= rad2deg; v[6 + i] *= rad2deg; }
      }
#pragma unroll
      for (int i = 0; i < 9; i++) {
        T x = m_clamp(v[i], T(ot.clip_lo[i]), T(ot.clip_hi[i]));
        if (ot.normalize) x = (2 * (x - T(ot.norm_lo[i]))) / (T(ot.norm_hi[i]) - T(ot.norm_lo[i])) - 1;
        out[o + i] = x;
      }
      o += 9;
    } else if (ot.kind == kObsMotorEncoder) {
      for (int a = 0; a < M.n_api; a++) {
        T x = 0;   // fixed joints read 0.0
        if (ot.normalize) x = (2 * (x - T(ot.norm_lo[a]))) / (T(ot.norm_hi[a]) - T(ot.norm_lo[a])) - 1;
        if (M.api_dof[a] < 0) out[o + a] = x;
      }
#pragma unroll
      for (int l = 0; l < kLegs; l++)
#pragma unroll
        for (int j = 0; j < CH; j++)
          if ((l < M.n_legs) && (j < M.chain_len[l])) {
            const int a = M.dof_api[M.dof0[l] + j];
            T x = s.q[l][j];
            if (ot.degrees) x *= rad2deg;
            if (ot.clip) x = m_clamp(x, T(ot.clip_lo[a]), T(ot.clip_hi[a]));
            if (ot.normalize) x = (2 * (x - T(ot.norm_lo[a]))) / (T(ot.norm_hi[a]) - T(ot.norm_lo[a])) - 1;
            out[o + a] = x;
          }
      o += M.n_api;
    }
  }
  return o;
}

// Reward program (postfix) on the current state.
template <typename T, int CH>
SOLO_HD T compute_reward(const ModelT<T> &M, const ProgramC &P, const EnvState<T, CH> &s, const T *euler) {
  T st[kRewardStack];
  int sp = 0;
  for (int k = 0; k < P.n_rew; k++) {
    const RewardOpC &op = P.rew[k];
    T v = 0;
    switch (op.op) {
      case kRewUpright:   // rewards.py:221-234: (-pi/2) pitch / (-pi/2)^2
        v = T(-1.5707963267948966) * euler[1] / T(2.4674011002723395);
        break;
      case kRewFlatTorso: {   // rewards.py:254-269
        const T x = m_sqrt(euler[0] * euler[0] + euler[1] * euler[1]);
        v = gaussian_tol(x, T(-op.p[0]), T(op.p[0]), T(op.p[1]));
      } break;
      case kRewSmallControl: {   // rewards.py:289-301: mean |qdot| over ALL api joints
        T a = 0;
#pragma unroll
        for (int l = 0; l < kLegs; l++)
#pragma unroll
          for (int j = 0; j < CH; j++)
            if ((l < M.n_legs) && (j < M.chain_len[l])) a += m_abs(s.qd[l][j]);
        v = gaussian_tol(a / T(M.n_api), T(0), T(0), T(op.p[1]));
      } break;
      case kRewHSpeed: {   // rewards.py:326-338
        const T x = m_sqrt(s.vlin[0] * s.vlin[0] + s.vlin[1] * s.vlin[1]);
        v = gaussian_tol(x, T(op.p[0]), T(op.p[1]), T(op.p[2]));
      } break;
      case kRewHeight:   // rewards.py:362-373
        v = gaussian_tol(s.pos[2], T(op.p[0]), T(op.p[1]), T(op.p[2]));
        break;
      case kRewConst:
        v = T(op.p[0]);
        break;
      case kRewAdd: {   // rewards.py:121-160 / 104-118: sum_k c_k x_k, first pushed first
        T acc = 0;
        for (int i = 0; i < op.n; i++) acc += T(op.coef[i]) * st[sp - op.n + i];
        sp -= op.n;
        v = acc;
      } break;
      case kRewMul: {   // rewards.py:163-199: c * prod x_k
        T acc = T(op.p[0]);
        for (int i = 0; i < op.n; i++) acc *= st[sp - op.n + i];
        sp -= op.n;
        v = acc;
      } break;
      default: break;
    }
    if (sp < kRewardStack) st[sp++] = v;
  }
  return sp > 0 ? st[sp - 1] : T(0);
}

// TerminationFactory.is_terminated (termination.py:38-50): OR with short
// circuit -- once a term fires, later TimeBasedTermination counters are not
// ticked.  counters: int32 [term][stride].
// commit = false evaluates without ticking (finish_kernel's warm-up pass).
SOLO_HD bool eval_terminations(const ProgramC &P, int32_t *counters, int64_t stride, const bool commit = true) {
  for (int t = 0; t < P.n_term; t++) {
    const TermTermC &tt = P.term[t];
    if (tt.kind == kTermTime) {   // termination.py:72-83
      const int32_t c = counters[t * stride] + 1;
      if (commit) counters[t * stride] = c;
      if (c > tt.max_step_delta) return true;
    }
    // PerpetualTermination (termination.py:86-98): never fires
  }
  return false;
}

// state <-> flat API layout helpers (pos 3, quat 4, vlin 3, vang 3, q[ndof], qd[ndof])
template <typename T, int CH, typename U>
SOLO_HD void load_state(const ModelT<T> &M, const U *src, int64_t stride, EnvState<T, CH> &s) {
#pragma unroll
  for (int i = 0; i < 3; i++) {
    s.pos[i] = (T)src[i * stride];
    s.vlin[i] = (T)src[(7 + i) * stride];
    s.vang[i] = (T)src[(10 + i) * stride];
  }
#pragma unroll
  for (int i = 0; i < 4; i++) s.quat[i] = (T)src[(3 + i) * stride];
#pragma unroll
  for (int l = 0; l < kLegs; l++)
#pragma unroll
    for (int j = 0; j < CH; j++) {
      const bool a = (l < M.n_legs) && (j < M.chain_len[l]);
      const int d = a ? M.dof0[l] + j : 0;
      s.q[l][j] = a ? (T)src[(kStateBase + d) * stride] : T(0);
      s.qd[l][j] = a ? (T)src[(kStateBase + M.ndof + d) * stride] : T(0);
    }
}
template <typename T, int CH, typename U>
SOLO_HD void store_state(const ModelT<T> &M, const EnvState<T, CH> &s, U *dst, int64_t stride) {
#pragma unroll
  for (int i = 0; i < 3; i++) {
    dst[i * stride] = (U)s.pos[i];
    dst[(7 + i) * stride] = (U)s.vlin[i];
    dst[(10 + i) * stride] = (U)s.vang[i];
  }
#pragma unroll
  for (int i = 0; i < 4; i++) dst[(3 + i) * stride] = (U)s.quat[i];
#pragma unroll
  for (int l = 0; l < kLegs; l++)
#pragma unroll
    for (int j = 0; j < CH; j++)
      if ((l < M.n_legs) && (j < M.chain_len[l])) {
        const int d = M.dof0[l] + j;
        dst[(kStateBase + d) * stride] = (U)s.q[l][j];
        dst[(kStateBase + M.ndof + d) * stride] = (U)s.qd[l][j];
      }
}
// action (API-wide, fixed-joint slots ignored; solo8v2vanilla.py:87-90) -> per-dof command
template <typename T, int CH, typename U>
SOLO_HD void load_command(const ModelT<T> &M, const U *act, T scale, T (&cmd)[kLegs][CH]) {
#pragma unroll
  for (int l = 0; l < kLegs; l++)
#pragma unroll
    for (int j = 0; j < CH; j++) {
      const bool a = (l < M.n_legs) && (j < M.chain_len[l]);
      cmd[l][j] = a ? (T)act[M.dof_api[M.dof0[l] + j]] * scale : T(0);
    }
}

}  // namespace solo
#endif  // SOLO_CORE_H_
