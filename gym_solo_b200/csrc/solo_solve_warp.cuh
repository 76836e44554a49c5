// solo_solve_warp.cuh -- the constraint solver of stepSimulation with ONE WARP PER
// ENVIRONMENT.  Same iterates as solo_solve.h (the sequential statement of the
// algorithm, which tests/hostsim runs on the CPU), different schedule:
//
//   * yh, the motor impulses and every piece of control flow are warp-uniform
//     (all 32 lanes carry the same values): no divergence, no max-over-lanes wait
//     for the slowest environment of a warp, and the per-SM warp scheduler
//     interleaves many independent environments to hide each one's serial chain;
//   * the normal rows are lane-parallel: lane l keeps normal row l (and l + 32) as
//     a dense (6 + 4 CH)-wide vector in registers for the whole solve.  About 90 %
//     of the rows inside Bullet's 2 cm breaking threshold are speculative -- zero
//     impulse and b' - u.yh <= 0, a no-op -- so one iteration's normal phase is
//     "every lane screens its row against the current yh, ballot", and only the
//     rows that are live are applied, in Bullet's order, each followed by a
//     re-screen of the rows behind it (exactly sequential Gauss-Seidel: a row is a
//     no-op iff it is idle when its turn comes);
//   * the friction phase walks the bitmask of contacts that have carried a normal
//     impulse (only they can hold friction: limit = mu * lambda_n).
//
// Rows come from the per-env record list written by setup_kernel (L2/L1
// resident); the 8..12 motor rows and all impulses sit in shared memory.
#ifndef SOLO_SOLVE_WARP_CUH_
#define SOLO_SOLVE_WARP_CUH_

#include "solo_kernels.cuh"

namespace solo {

template <int CH>
struct WarpSolveSmem {
  float lam[3 * kMaxContacts];                                  // scaled contact impulses
  float mrow[RowLayout<CH>::kMotorSlots][RowLayout<CH>::kRow];  // motor rows
};

// uniform: group (0 base, 1..4 legs) of contact c from the packed group ends
__device__ __forceinline__ int contact_group(const uint32_t ends03, const int c) {
  return (c >= (int)(ends03 & 255u)) + (c >= (int)((ends03 >> 8) & 255u)) + (c >= (int)((ends03 >> 16) & 255u)) +
         (c >= (int)((ends03 >> 24) & 255u));
}

template <int CH, bool DBG>
__global__ void __launch_bounds__(32)
solve_warp_kernel(const __grid_constant__ ConfigT<float> C, const Buffers B, const uint32_t motor_mask,
                  const int ndof_total, const int n_cand) {
  typedef RowLayout<CH> RL;
  constexpr int W = 6 + kLegs * CH;   // dense width
  constexpr unsigned FULL = 0xffffffffu;
  __shared__ __align__(16) WarpSolveSmem<CH> sm;
  const int lane = threadIdx.x;
  const int64_t env = blockIdx.x;
  const long long t_begin = clock64();

  const uint32_t ends03 = B.hdr[env];
  const int nc = (int)B.hdr[(int64_t)B.n + env];
  const float *rows = B.rows + env * (RL::kSlots * RL::kRow);
  const float *crow = rows + RL::kMotorSlots * RL::kRow;

  // ---- stage: motor rows -> smem, impulses = 0, this lane's normal rows -> registers
  {
    const float4 *src = reinterpret_cast<const float4 *>(rows);
    float4 *dst = reinterpret_cast<float4 *>(&sm.mrow[0][0]);
    for (int i = lane; i < RL::kMotorSlots * RL::kRow / 4; i += 32) dst[i] = src[i];
    for (int i = lane; i < 3 * nc; i += 32) sm.lam[i] = 0.f;
  }
  float ud[2][W], ub[2], us[2];   // dense normal row of contacts lane and lane + 32, b', s
  float lamn[2] = {0.f, 0.f};     // their scaled normal impulses
  bool valid[2];
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int c = lane + 32 * r;
    valid[r] = c < nc;
#pragma unroll
    for (int a = 0; a < W; a++) ud[r][a] = 0.f;
    ub[r] = 0.f; us[r] = 0.f;
    if (valid[r]) {
      float row[RL::kRow];
      load_row(crow + (3 * c) * RL::kRow, row);
      const int g = contact_group(ends03, c);
#pragma unroll
      for (int a = 0; a < 6; a++) ud[r][a] = row[a];
#pragma unroll
      for (int l = 0; l < kLegs; l++)
#pragma unroll
        for (int k = 0; k < CH; k++) ud[r][6 + l * CH + k] = (g == l + 1) ? row[6 + k] : 0.f;
      ub[r] = row[RL::kB]; us[r] = row[RL::kS];
    }
  }
  __syncwarp();

  float yb[6], yl[kLegs][CH], mlam[kLegs * CH];
#pragma unroll
  for (int a = 0; a < 6; a++) yb[a] = 0.f;
#pragma unroll
  for (int l = 0; l < kLegs; l++)
#pragma unroll
    for (int k = 0; k < CH; k++) yl[l][k] = 0.f;
#pragma unroll
  for (int j = 0; j < kLegs * CH; j++) mlam[j] = 0.f;

  // screen this lane's rows against the current yh: b' - u.yh per row
  auto screen = [&](float (&t)[2]) {
#pragma unroll
    for (int r = 0; r < 2; r++) {
      float a0 = ub[r], a1 = 0.f, a2 = 0.f;
#pragma unroll
      for (int a = 0; a < 6; a += 2) { a0 -= ud[r][a] * yb[a]; a1 -= ud[r][a + 1] * yb[a + 1]; }
#pragma unroll
      for (int l = 0; l < kLegs; l++)
#pragma unroll
        for (int k = 0; k < CH; k++) {
          if ((l * CH + k) % 3 == 0) a2 -= ud[r][6 + l * CH + k] * yl[l][k];
          else if ((l * CH + k) % 3 == 1) a0 -= ud[r][6 + l * CH + k] * yl[l][k];
          else a1 -= ud[r][6 + l * CH + k] * yl[l][k];
        }
      t[r] = (a0 + a1) + a2;
    }
  };
  // yh += u * d for a compact row of group g (warp-uniform g)
  auto apply = [&](const float (&r)[RL::kRow], const int g, const float d) {
#pragma unroll
    for (int a = 0; a < 6; a++) yb[a] += r[a] * d;
#pragma unroll
    for (int l = 0; l < kLegs; l++)
      if (g == l + 1) {
#pragma unroll
        for (int k = 0; k < CH; k++) yl[l][k] += r[6 + k] * d;
      }
  };
  // u . yh for a compact row of group g
  auto dotrow = [&](const float (&r)[RL::kRow], const int g) {
    float a0 = r[RL::kB], a1 = 0.f;
#pragma unroll
    for (int a = 0; a < 6; a += 2) { a0 -= r[a] * yb[a]; a1 -= r[a + 1] * yb[a + 1]; }
#pragma unroll
    for (int l = 0; l < kLegs; l++)
      if (g == l + 1) {
#pragma unroll
        for (int k = 0; k < CH; k++) { if (k & 1) a1 -= r[6 + k] * yl[l][k]; else a0 -= r[6 + k] * yl[l][k]; }
      }
    return a0 + a1;
  };

  const float mlim = C.torque_limit * C.dt;
  uint64_t active = 0;   // contacts that have carried a normal impulse
  int it = 0;
  if (motor_mask != 0 || nc > 0) {
    for (; it < C.iters;) {
      float res = 0.f;
      // ---- motor rows (btMultiBodyJointMotor), direction alternates with parity ----
      if (motor_mask != 0) {
        auto motor = [&](const int l, const int i) {
          const int j = l * CH + i;
          float r[RL::kRow];
          load_row(&sm.mrow[j][0], r);
          float a0 = r[RL::kB], a1 = 0.f;
#pragma unroll
          for (int a = 0; a < 6; a += 2) { a0 -= r[a] * yb[a]; a1 -= r[a + 1] * yb[a + 1]; }
#pragma unroll
          for (int k = 0; k < CH; k++) { if (k & 1) a1 -= r[6 + k] * yl[l][k]; else a0 -= r[6 + k] * yl[l][k]; }
          const float lim = mlim * r[RL::kS];
          const float old = mlam[j];
          const float sum = m_clamp(old + (a0 + a1), -lim, lim);
          const float d = sum - old;
          mlam[j] = sum;
#pragma unroll
          for (int a = 0; a < 6; a++) yb[a] += r[a] * d;
#pragma unroll
          for (int k = 0; k < CH; k++) yl[l][k] += r[6 + k] * d;
          const float rr = d * r[RL::kS];
          res = m_max(res, rr * rr);
        };
        if (it & 1) {
#pragma unroll
          for (int jj = 0; jj < kLegs * CH; jj++) motor(jj / CH, jj % CH);
        } else {
#pragma unroll
          for (int jj = kLegs * CH - 1; jj >= 0; jj--) motor(jj / CH, jj % CH);
        }
      }

      // ---- normal rows: screen in parallel, apply the live ones in order ----
      // Lane l owns rows l and l + 32: dense u, b', s and the impulse.  A row is
      // live iff its clamped update d is non-zero; the lowest live row is applied
      // (its u and d broadcast by shuffles), then the rows behind it are re-screened.
      if (nc > 0) {
        uint64_t todo = ~0ull;   // rows not yet passed in this iteration
        for (;;) {
          float t[2], dl[2];
          screen(t);
#pragma unroll
          for (int r = 0; r < 2; r++) dl[r] = m_max(lamn[r] + t[r], 0.f) - lamn[r];
          uint64_t live = (uint64_t)__ballot_sync(FULL, valid[0] && dl[0] != 0.f);
          if (nc > 32) live |= (uint64_t)__ballot_sync(FULL, valid[1] && dl[1] != 0.f) << 32;
          live &= todo;
          if (live == 0) break;
          const int c = lowest_bit64(live);
          const int owner = c & 31;
          todo = ~low_mask64(c + 1);
          active |= 1ull << c;
          float d;
          if (c < 32) {
            d = __shfl_sync(FULL, dl[0], owner);
#pragma unroll
            for (int a = 0; a < 6; a++) yb[a] += __shfl_sync(FULL, ud[0][a], owner) * d;
#pragma unroll
            for (int l = 0; l < kLegs; l++)
#pragma unroll
              for (int k = 0; k < CH; k++) yl[l][k] += __shfl_sync(FULL, ud[0][6 + l * CH + k], owner) * d;
            if (lane == owner) {
              lamn[0] += dl[0];
              const float rr = dl[0] * us[0];
              res = m_max(res, rr * rr);
            }
          } else {
            d = __shfl_sync(FULL, dl[1], owner);
#pragma unroll
            for (int a = 0; a < 6; a++) yb[a] += __shfl_sync(FULL, ud[1][a], owner) * d;
#pragma unroll
            for (int l = 0; l < kLegs; l++)
#pragma unroll
              for (int k = 0; k < CH; k++) yl[l][k] += __shfl_sync(FULL, ud[1][6 + l * CH + k], owner) * d;
            if (lane == owner) {
              lamn[1] += dl[1];
              const float rr = dl[1] * us[1];
              res = m_max(res, rr * rr);
            }
          }
          if (c + 1 >= nc) break;
        }
      }

      // ---- friction pairs of the contacts that can hold friction ----
      for (uint64_t rem = active; rem != 0; rem &= rem - 1) {
        const int c = lowest_bit64(rem);
        const int g = contact_group(ends03, c);
        float fa[RL::kRow], fb[RL::kRow];
        load_row(crow + (3 * c + 1) * RL::kRow, fa);
        load_row(crow + (3 * c + 2) * RL::kRow, fb);
        const float ln = __shfl_sync(FULL, c < 32 ? lamn[0] : lamn[1], c & 31);
        const float la = sm.lam[3 * c + 1], lb = sm.lam[3 * c + 2];
        float sa = la + dotrow(fa, g), sb = lb + dotrow(fb, g);
        const float limit = fa[RL::kX] * ln;                       // mu * lambda_n, true units
        const float ta = sa * fa[RL::kIs], tb = sb * fb[RL::kIs];  // true units
        const float n2 = ta * ta + tb * tb;
        if (n2 > limit * limit) {
          const float sc = limit * rsqrtf(n2);
          sa *= sc; sb *= sc;
        }
        const float da = sa - la, db = sb - lb;
        __syncwarp();
        sm.lam[3 * c + 1] = sa; sm.lam[3 * c + 2] = sb;
        __syncwarp();
        apply(fa, g, da);
        apply(fb, g, db);
        const float r1 = da * fa[RL::kS], r2 = db * fb[RL::kS];
        res = m_max(res, m_max(r1 * r1, r2 * r2));
      }
      it++;
      // the owner lanes of the normal rows hold part of the residual
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) res = m_max(res, __shfl_xor_sync(FULL, res, off));
      if (res <= C.res_thr) break;
    }
  }

  // ---- outputs ----
  {
    // yh is warp-uniform: lane a writes entry a
    float v = 0.f;
#pragma unroll
    for (int a = 0; a < 6; a++)
      if (lane == a) v = yb[a];
#pragma unroll
    for (int l = 0; l < kLegs; l++)
#pragma unroll
      for (int k = 0; k < CH; k++)
        if (lane == 6 + l * CH + k) v = yl[l][k];
    if (lane < W) B.yh[(int64_t)lane * B.n + env] = v;
  }
  if (lane == 0) {
    // per-env solver statistics: iterations, contacts, contacts that carried an
    // impulse, SM cycles spent in the solve
    float *st = B.yh + (int64_t)W * B.n + env;
    st[0] = (float)it; st[(int64_t)B.n] = (float)nc; st[2 * (int64_t)B.n] = (float)__popcll(active);
    st[3 * (int64_t)B.n] = (float)(clock64() - t_begin);
  }
  if (DBG) {
    __syncwarp();
    if (B.dbg_normal_force) {
      for (int c = lane; c < n_cand; c += 32) B.dbg_normal_force[(int64_t)c * B.n + env] = 0.f;
      __syncwarp();
#pragma unroll
      for (int r = 0; r < 2; r++) {
        const int i = lane + 32 * r;
        if (i < nc) {
          const float *nr = crow + (3 * i) * RL::kRow;
          B.dbg_normal_force[(int64_t)(int)nr[RL::kX] * B.n + env] = lamn[r] * nr[RL::kIs] * C.inv_dt;
        }
      }
    }
    if (B.dbg_motor_torque && lane == 0) {
      // dof order == (leg, joint) order with inactive slots skipped
      if (motor_mask == 0)
        for (int k = 0; k < ndof_total; k++) B.dbg_motor_torque[(int64_t)k * B.n + env] = 0.f;
      int d = 0;
#pragma unroll
      for (int j = 0; j < kLegs * CH; j++) {
        if ((motor_mask >> j) & 1u) {
          if (d < ndof_total) B.dbg_motor_torque[(int64_t)d * B.n + env] = mlam[j] * sm.mrow[j][RL::kIs] * C.inv_dt;
          d++;
        }
      }
    }
  }
}

}  // namespace solo
#endif  // SOLO_SOLVE_WARP_CUH_
