// solo_solve.h -- phase 2 of the physics step (see solo_phases.h): projected
// Gauss-Seidel over the whitened, normalised rows, in the order of
// btMultiBodyConstraintSolver::solveSingleIteration (Bullet
// recollection): the non-contact rows -- joint-limit rows (created at URDF import,
// one per joint at or beyond a limit), then the motor rows, the traversal of the
// whole list alternating in direction with iteration parity -- then every contact
// normal row, then every friction pair (both rows
// evaluated on the same velocity, radial projection onto the disc of radius
// mu * lambda_n: resolveConeFrictionConstraintRows); early exit once the
// largest squared Bullet residual of an iteration is <= res_thr.
//
// The per-environment work is a serial chain (every update reads the yh the
// previous one wrote), so the code is written for latency: rows are prefetched
// one update ahead, the 6 + CH wide dot runs as two accumulator chains, and the
// only state carried between updates is yh (6 + kLegs * CH registers).
#ifndef SOLO_SOLVE_H_
#define SOLO_SOLVE_H_

namespace solo {

// rows: as written by setup_step.  lam: scaled contact impulses, entry r at
// lam[r * lam_stride], r = 3 * contact + {0 normal, 1 t1, 2 t2} (zeroed here).
// motor_mask: bit (l * CH + i) set for every live motor row (0: no motor rows).
// Outputs: yb / yl = yh, mlam = scaled motor impulses.  Returns iterations run.
template <typename T, int CH>
SOLO_HD int solve_rows(const ConfigT<T> &C, const uint32_t motor_mask, const T *rows, const int64_t slot_stride,
                       const RowHeader h, T *lam, const int64_t lam_stride, T (&yb)[6], T (&yl)[kLegs][CH],
                       T (&mlam)[kLegs * CH], int *n_active = nullptr) {
  typedef RowLayout<CH> RL;
  const int nc = (int)h.nc;
  const int ends[kLegs + 1] = {(int)(h.ends03 & 255u), (int)((h.ends03 >> 8) & 255u), (int)((h.ends03 >> 16) & 255u),
                               (int)((h.ends03 >> 24) & 255u), nc};
#pragma unroll
  for (int a = 0; a < 6; a++) yb[a] = 0;
#pragma unroll
  for (int l = 0; l < kLegs; l++)
#pragma unroll
    for (int k = 0; k < CH; k++) yl[l][k] = 0;
#pragma unroll
  for (int j = 0; j < kLegs * CH; j++) mlam[j] = 0;
  for (int r = 0; r < 3 * nc; r++) lam[(int64_t)r * lam_stride] = 0;
  const uint32_t lim = h.lim;
  T llam[kLegs * CH];      // scaled joint-limit impulses (>= 0 along the row's own direction)
#pragma unroll
  for (int j = 0; j < kLegs * CH; j++) llam[j] = 0;
  if (motor_mask == 0 && nc == 0 && lim == 0) return 0;

  const T mlim = C.torque_limit * C.dt;
  const T *crow = rows + (int64_t)RL::kMotorSlots * slot_stride;
  // the next normal row and the next friction pair, always one update ahead
  T ncur[RL::kRow], fa[RL::kRow], fb[RL::kRow];
  uint64_t active = 0;   // contacts that have carried a normal impulse
  int fpre = 0;          // contact whose friction pair sits in fa / fb
  if (nc > 0) {
    load_row(crow, ncur);
    load_row(crow + slot_stride, fa);
    load_row(crow + 2 * slot_stride, fb);
  } else {
#pragma unroll
    for (int a = 0; a < RL::kRow; a++) { ncur[a] = 0; fa[a] = 0; fb[a] = 0; }
  }

  int it = 0;
  for (; it < C.iters;) {
    T res = 0;
    // joint-limit rows: u = dir * u_motor of the slot, b' in its kX, 0 <= lambda' <= lim_max * s
    auto limit = [&](const int l, const int i) {
      const int j = l * CH + i;
      if (((lim >> j) & 1u) == 0) return;
      T r[RL::kRow];
      load_row(rows + (int64_t)j * slot_stride, r);
      const T dir = ((lim >> (16 + j)) & 1u) ? T(-1) : T(1);
      T dotv = 0;
#pragma unroll
      for (int a = 0; a < 6; a++) dotv += r[a] * yb[a];
#pragma unroll
      for (int k = 0; k < CH; k++) dotv += r[6 + k] * yl[l][k];
      const T old = llam[j];
      const T sum = m_clamp(old + (r[RL::kX] - dir * dotv), T(0), C.lim_max * r[RL::kS]);
      const T d = (sum - old) * dir;
      llam[j] = sum;
#pragma unroll
      for (int a = 0; a < 6; a++) yb[a] += r[a] * d;
#pragma unroll
      for (int k = 0; k < CH; k++) yl[l][k] += r[6 + k] * d;
      const T rr = d * r[RL::kS];
      res = m_max(res, rr * rr);
    };
    if (lim != 0 && (it & 1)) {
#pragma unroll
      for (int jj = 0; jj < kLegs * CH; jj++) limit(jj / CH, jj % CH);
    }
    if (motor_mask != 0) {
      auto motor = [&](const int l, const int i) {
        // dead slots (shorter chains, torque mode) hold all-zero rows: no test needed
        const int j = l * CH + i;
        T r[RL::kRow];
        load_row(rows + (int64_t)j * slot_stride, r);
        T a0 = r[RL::kB], a1 = 0;
#pragma unroll
        for (int a = 0; a < 6; a += 2) { a0 -= r[a] * yb[a]; a1 -= r[a + 1] * yb[a + 1]; }
#pragma unroll
        for (int k = 0; k < CH; k++) { if (k & 1) a1 -= r[6 + k] * yl[l][k]; else a0 -= r[6 + k] * yl[l][k]; }
        const T lim = mlim * r[RL::kS];
        const T old = mlam[j];
        const T sum = m_clamp(old + (a0 + a1), -lim, lim);
        const T d = sum - old;
        mlam[j] = sum;
#pragma unroll
        for (int a = 0; a < 6; a++) yb[a] += r[a] * d;
#pragma unroll
        for (int k = 0; k < CH; k++) yl[l][k] += r[6 + k] * d;
        const T rr = d * r[RL::kS];
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
    if (lim != 0 && !(it & 1)) {
#pragma unroll
      for (int jj = kLegs * CH - 1; jj >= 0; jj--) limit(jj / CH, jj % CH);
    }

    // ---- normal rows ----
    // Most contacts inside the 2 cm breaking threshold are speculative: their
    // impulse stays 0 and yh is untouched, so the next row's dot does not wait
    // for this one (the update is skipped, which also breaks the dependency chain).
    int c = 0;
#pragma unroll
    for (int g = 0; g <= kLegs; g++) {
      const int ll = g > 0 ? g - 1 : 0;
      while (c < ends[g]) {
        const int cn = (c + 1 < nc) ? c + 1 : 0;
        T nxt[RL::kRow];
        load_row(crow + (int64_t)(3 * cn) * slot_stride, nxt);
        T a0 = ncur[RL::kB], a1 = 0;
#pragma unroll
        for (int a = 0; a < 6; a += 2) { a0 -= ncur[a] * yb[a]; a1 -= ncur[a + 1] * yb[a + 1]; }
        if (g > 0) {
#pragma unroll
          for (int k = 0; k < CH; k++) { if (k & 1) a1 -= ncur[6 + k] * yl[ll][k]; else a0 -= ncur[6 + k] * yl[ll][k]; }
        }
        T *lp = lam + (int64_t)(3 * c) * lam_stride;
        const T old = *lp;
        const T sum = m_max(old + (a0 + a1), T(0));
        const T d = sum - old;
        if (d != T(0)) {
          *lp = sum;
          active |= 1ull << c;
#pragma unroll
          for (int a = 0; a < 6; a++) yb[a] += ncur[a] * d;
          if (g > 0) {
#pragma unroll
            for (int k = 0; k < CH; k++) yl[ll][k] += ncur[6 + k] * d;
          }
          const T rr = d * ncur[RL::kS];
          res = m_max(res, rr * rr);
        }
#pragma unroll
        for (int a = 0; a < RL::kRow; a++) ncur[a] = nxt[a];
        c++;
      }
    }

    // ---- friction pairs ----
    // Only contacts that have carried a normal impulse at some point can hold a
    // non-zero friction impulse (limit = mu * lambda_n); for every other contact
    // the pair projects 0 onto 0.  `active` has one bit per such contact.
    uint64_t rem = active;
#pragma unroll
    for (int g = 0; g <= kLegs; g++) {
      const int ll = g > 0 ? g - 1 : 0;
      const uint64_t gm = low_mask64(ends[g]) & ~low_mask64(g > 0 ? ends[g - 1] : 0);
      while ((rem & gm) != 0) {
        const int cc = lowest_bit64(rem);
        rem &= rem - 1;
        if (fpre != cc) {   // a contact with a lower index became active since the prefetch
          load_row(crow + (int64_t)(3 * cc + 1) * slot_stride, fa);
          load_row(crow + (int64_t)(3 * cc + 2) * slot_stride, fb);
        }
        const int cn = lowest_bit64(rem != 0 ? rem : active);
        fpre = cn;
        T na[RL::kRow], nb[RL::kRow];
        load_row(crow + (int64_t)(3 * cn + 1) * slot_stride, na);
        load_row(crow + (int64_t)(3 * cn + 2) * slot_stride, nb);
        T *lp = lam + (int64_t)(3 * cc) * lam_stride;
        const T ln = lp[0], la = lp[lam_stride], lb = lp[2 * lam_stride];
        T a0 = fa[RL::kB], a1 = 0, b0 = fb[RL::kB], b1 = 0;
#pragma unroll
        for (int a = 0; a < 6; a += 2) {
          a0 -= fa[a] * yb[a]; a1 -= fa[a + 1] * yb[a + 1];
          b0 -= fb[a] * yb[a]; b1 -= fb[a + 1] * yb[a + 1];
        }
        if (g > 0) {
#pragma unroll
          for (int k = 0; k < CH; k++) {
            if (k & 1) { a1 -= fa[6 + k] * yl[ll][k]; b1 -= fb[6 + k] * yl[ll][k]; }
            else { a0 -= fa[6 + k] * yl[ll][k]; b0 -= fb[6 + k] * yl[ll][k]; }
          }
        }
        T sa = la + (a0 + a1), sb = lb + (b0 + b1);
        const T limit = fa[RL::kX] * ln;                       // mu * lambda_n, true units
        const T ta = sa * fa[RL::kIs], tb = sb * fb[RL::kIs];  // true units
        const T n2 = ta * ta + tb * tb;
        if (n2 > limit * limit) {
          const T sc = limit * m_rsqrt(n2);
          sa *= sc; sb *= sc;
        }
        const T da = sa - la, db = sb - lb;
        lp[lam_stride] = sa; lp[2 * lam_stride] = sb;
#pragma unroll
        for (int a = 0; a < 6; a++) yb[a] += fa[a] * da + fb[a] * db;
        if (g > 0) {
#pragma unroll
          for (int k = 0; k < CH; k++) yl[ll][k] += fa[6 + k] * da + fb[6 + k] * db;
        }
        const T r1 = da * fa[RL::kS], r2 = db * fb[RL::kS];
        res = m_max(res, m_max(r1 * r1, r2 * r2));
#pragma unroll
        for (int a = 0; a < RL::kRow; a++) { fa[a] = na[a]; fb[a] = nb[a]; }
      }
    }
    it++;
    if (res <= C.res_thr) break;
  }
  if (n_active) {
    int k = 0;
    for (uint64_t m = active; m != 0; m &= m - 1) k++;
    *n_active = k;
  }
  return it;
}

}  // namespace solo
#endif  // SOLO_SOLVE_H_
