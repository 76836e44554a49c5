// solo_phases.h -- the physics step of one environment, in three phases that the
// CUDA path runs as three kernels (solo_kernels.cuh) and tests/hostsim runs back
// to back on the host CPU (test infrastructure only):
//
//   setup_step   state + command -> whitened constraint rows, mass-matrix factors
//                (uniform work per environment: one thread per environment)
//   solve_rows   projected Gauss-Seidel over the rows -> yh   (data-dependent work:
//                2..50 iterations x 8 + 3 n_contacts rows; a serial chain per env)
//   finish_step  dv = T^T yh, integrate, (then the obs/reward/termination epilogue)
//
// Replaces setJointMotorControlArray + stepSimulation
// (gym_solo/envs/solo8v2vanilla.py:87-91).
//
// Formulation (NOT the oracle's: the two are meant to check each other).
// Generalised velocity v = (v_com0, omega0, qdot), world frame.  The robot is a
// star: a floating base plus n_legs serial chains of <= CH revolute joints, so
//   M = [ Mbb  B ; B^T  D ],  D = blockdiag(D_leg).
// Per leg: D = C C^T, E = C^-1, Bh = E B^T; base Schur complement
// S = Mbb - sum Bh^T Bh = Cs Cs^T, Ls = Cs^-1.  With
//   T = [ Ls  -Ls Bh^T E ... ; 0  E ]   we have   M^-1 = T^T T,
// so every constraint row J is stored whitened AND normalised,
//   u = T J^T / |T J^T|   (6 base entries + CH entries of its own leg),
// the PGS state is yh = sum u * delta', the scaled impulse is lambda' =
// lambda |T J^T|, and one Gauss-Seidel update is
//   delta' = b' - u . yh ;  clamp lambda' + delta' ;  yh += u * delta'
// -- 8 FMAs each way instead of Bullet's 8 + 14, with the same iterates in exact
// arithmetic (Bullet: delta = (rhs - J dv) / (J M^-1 J^T)).
// Unconstrained accelerations use the same factors: a = T^T (T f), f = tau - C
// with C from a Newton-Euler pass at zero generalised acceleration.
#ifndef SOLO_PHASES_H_
#define SOLO_PHASES_H_

#include "solo_core.h"

namespace solo {

// One stored row: [0..5] u base, [6..6+CH) u leg, then
//   kB   b' = velocity error / |T J^T|
//   kS   s  = |T J^T|        (true impulse = lambda' / s; Bullet residual = delta' * s)
//   kIs  1 / s
//   kX   normal row: contact candidate index (as float); first friction row:
//        mu / s_normal (friction limit = kX * lambda'_normal in true units);
//        motor slot: b' of the joint-limit row of the same joint (it shares the
//        motor row's u up to its sign, see `lim` below)
// Slots: kLegs * CH motor rows (leg major), then per contact: normal, t1, t2.
//
// Joint-limit rows (btMultiBodyJointLimitConstraint): J = dir e_j with dir = +1 at the
// lower and -1 at the upper limit, i.e. u_limit = dir * u_motor.  Which slots have a live
// limit row is a bit mask `lim`: bit j = slot j has one, bit 16 + j = it is the upper one.
template <int CH> struct RowLayout {
  static constexpr int kJ = 6 + CH;
  static constexpr int kB = kJ, kS = kJ + 1, kIs = kJ + 2, kX = kJ + 3;
  static constexpr int kRow = (kJ + 4 + 3) & ~3;      // 12 floats (CH 2), 16 (CH 3)
  static constexpr int kMotorSlots = kLegs * CH;
  static constexpr int kSlots = kMotorSlots + 3 * kMaxContacts;
};

// Mass-matrix factors and post-drift velocities that finish_step needs.
template <int CH> struct FactorLayout {
  static constexpr int kLs = 0;                               // 21: lower triangle, row major
  static constexpr int kLegSize = CH * 6 + CH * (CH + 1) / 2;  // Bh[CH][6], E lower
  static constexpr int kLeg = 21;
  static constexpr int kVb = kLeg + kLegs * kLegSize;         // 6
  static constexpr int kVl = kVb + 6;                         // kLegs * CH
  static constexpr int kSize = kVl + kLegs * CH;
};

// What setup_step<.., CONTACTS = false> leaves per environment for the solver
// warp (one contiguous record, floats): the motor rows, the base factor Ls,
// the post-drift velocities, and per leg the factors and joint frames a contact
// row is built from.
template <int CH> struct KinLayout {
  static constexpr int kMotor = 0;                                           // kMotorSlots x kRow
  static constexpr int kLs = RowLayout<CH>::kMotorSlots * RowLayout<CH>::kRow;   // 21: lower triangle, row major
  static constexpr int kVb = kLs + 21;                                       // 6
  static constexpr int kC0 = kVb + 6;                                        // 3  base COM, world
  static constexpr int kO0 = kC0 + 3;                                        // 3  base frame origin
  static constexpr int kR0 = kO0 + 3;                                        // 9  base rotation
  static constexpr int kLeg = kR0 + 9;
  static constexpr int kE = 0;                                               // CH (CH + 1) / 2, lower, row major
  static constexpr int kBh = CH * (CH + 1) / 2;                              // CH x 6
  static constexpr int kVl = kBh + CH * 6;                                   // CH
  static constexpr int kJoint = kVl + CH;                                    // per joint: axis 3, origin 3, rotation 9
  static constexpr int kLegSize = kJoint + CH * 15;
  static constexpr int kLim = kLeg + kLegs * kLegSize;                       // 1: joint-limit mask (bits of a float)
  static constexpr int kSize = (kLim + 1 + 3) & ~3;                          // 328 (CH 2), 524 (CH 3)
};

// Per-env header: contact-group ends (base, leg 0..3), 8 bits each.
struct RowHeader {
  uint32_t ends03;   // leg_end[0..3]
  uint32_t nc;       // leg_end[4] = number of contacts
  uint32_t lim;      // live joint-limit rows (see RowLayout)
};

SOLO_HD int lowest_bit(uint32_t m) {
#ifdef __CUDA_ARCH__
  return __ffs((int)m) - 1;
#else
  return __builtin_ctz(m);
#endif
}

SOLO_HD int lowest_bit64(uint64_t m) {
#ifdef __CUDA_ARCH__
  return __ffsll((long long)m) - 1;
#else
  return __builtin_ctzll(m);
#endif
}
SOLO_HD uint64_t low_mask64(int n) { return n >= 64 ? ~0ull : ((1ull << n) - 1ull); }

// Row IO.  The device float path moves whole 16-byte words.
template <typename T, int N> SOLO_HD void store_row(T *dst, const T (&r)[N]) {
#pragma unroll
  for (int i = 0; i < N; i++) dst[i] = r[i];
}
template <typename T, int N> SOLO_HD void load_row(const T *src, T (&r)[N]) {
#pragma unroll
  for (int i = 0; i < N; i++) r[i] = src[i];
}
#ifdef __CUDA_ARCH__
template <int N> __device__ __forceinline__ void store_row(float *dst, const float (&r)[N]) {
  static_assert(N % 4 == 0, "rows are whole float4s");
#pragma unroll
  for (int i = 0; i < N; i += 4) *reinterpret_cast<float4 *>(dst + i) = make_float4(r[i], r[i + 1], r[i + 2], r[i + 3]);
}
template <int N> __device__ __forceinline__ void load_row(const float *src, float (&r)[N]) {
  static_assert(N % 4 == 0, "rows are whole float4s");
#pragma unroll
  for (int i = 0; i < N; i += 4) {
    const float4 v = *reinterpret_cast<const float4 *>(src + i);
    r[i] = v.x; r[i + 1] = v.y; r[i + 2] = v.z; r[i + 3] = v.w;
  }
}
#endif

}  // namespace solo

#include "solo_setup.h"
#include "solo_solve.h"
#include "solo_finish.h"

#endif  // SOLO_PHASES_H_
