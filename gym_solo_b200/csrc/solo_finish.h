// solo_finish.h -- phase 3 of the physics step (see solo_phases.h):
// dv = T^T yh from the stored factors, velocity clamp, semi-implicit Euler.
#ifndef SOLO_FINISH_H_
#define SOLO_FINISH_H_

namespace solo {

template <typename T, int CH>
SOLO_HD void finish_step(const ModelT<T> &M, const ConfigT<T> &C, EnvState<T, CH> &s, const T *fac,
                         const int64_t fac_stride, const T (&yb)[6], const T (&yl)[kLegs][CH]) {
  typedef FactorLayout<CH> FL;
  const T dt = C.dt;
  T S[6][6];
  {
    int k = FL::kLs;
#pragma unroll
    for (int i = 0; i < 6; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) S[i][j] = fac[(k++) * fac_stride];
  }
  // every factor is fetched before the first use: the loads are independent, so
  // their latencies overlap instead of adding up phase by phase
  T Bh[kLegs][CH][6], E[kLegs][CH][CH], vb0[6], vl0[kLegs][CH];
#pragma unroll
  for (int i = 0; i < 6; i++) vb0[i] = fac[(FL::kVb + i) * fac_stride];
#pragma unroll
  for (int l = 0; l < kLegs; l++) {
    const int base = FL::kLeg + l * FL::kLegSize;
#pragma unroll
    for (int i = 0; i < CH; i++)
#pragma unroll
      for (int a = 0; a < 6; a++) Bh[l][i][a] = fac[(base + i * 6 + a) * fac_stride];
    int k = base + CH * 6;
#pragma unroll
    for (int i = 0; i < CH; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) E[l][i][j] = fac[(k++) * fac_stride];
#pragma unroll
    for (int i = 0; i < CH; i++) vl0[l][i] = fac[(FL::kVl + l * CH + i) * fac_stride];
  }
  T vb[6], dvb[6];
#pragma unroll
  for (int i = 0; i < 6; i++) {
    T t = 0;
#pragma unroll
    for (int k = i; k < 6; k++) t += S[k][i] * yb[k];
    dvb[i] = t;
    vb[i] = m_clamp(vb0[i] + t, -C.max_vel, C.max_vel);
  }
#pragma unroll
  for (int l = 0; l < kLegs; l++) {
    T t[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) {
      T u = yl[l][i];
#pragma unroll
      for (int a = 0; a < 6; a++) u -= Bh[l][i][a] * dvb[a];
      t[i] = u;
    }
#pragma unroll
    for (int i = 0; i < CH; i++) {
      T dv = 0;
#pragma unroll
      for (int k = i; k < CH; k++) dv += E[l][k][i] * t[k];
      const bool a = (l < M.n_legs) && (i < M.chain_len[l]);
      const T v = a ? m_clamp(vl0[l][i] + dv, -C.max_vel, C.max_vel) : T(0);
      s.qd[l][i] = v;
      s.q[l][i] += dt * v;
    }
  }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    s.vlin[i] = vb[i];
    s.vang[i] = vb[3 + i];
    s.pos[i] += dt * vb[i];
  }
  {
    // base orientation: exponential map of the world angular velocity
    // (btTransformUtil::integrateTransform as used by stepPositionsMultiDof)
    const T wx = vb[3], wy = vb[4], wz = vb[5];
    T ang = m_sqrt(wx * wx + wy * wy + wz * wz);
    if (ang * dt > T(0.25 * 3.14159265358979323846)) ang = T(0.5 * 0.5 * 3.14159265358979323846) / dt;
    T sc, cs, sn;
    m_sincos(T(0.5) * ang * dt, &sn, &cs);
    if (ang < T(0.001)) sc = T(0.5) * dt - dt * dt * dt * T(0.020833333333) * ang * ang;
    else sc = sn / ang;
    const T dx = wx * sc, dy = wy * sc, dz = wz * sc, dw = cs;
    const T x = s.quat[0], y = s.quat[1], z = s.quat[2], w = s.quat[3];
    T nx = dw * x + dx * w + dy * z - dz * y;
    T ny = dw * y - dx * z + dy * w + dz * x;
    T nz = dw * z + dx * y - dy * x + dz * w;
    T nw = dw * w - dx * x - dy * y - dz * z;
    const T nn = T(1) / m_sqrt(nx * nx + ny * ny + nz * nz + nw * nw);
    s.quat[0] = nx * nn; s.quat[1] = ny * nn; s.quat[2] = nz * nn; s.quat[3] = nw * nn;
  }
}

// The three phases back to back for one environment through caller scratch
// (tests/hostsim, i.e. test infrastructure; the CUDA path runs them as three
// kernels).  scratch: RowLayout<CH>::kSlots * kRow + FactorLayout<CH>::kSize +
// 3 * kMaxContacts values of T.
template <typename T, int CH> struct HostScratch {
  static constexpr int kRows = RowLayout<CH>::kSlots * RowLayout<CH>::kRow;
  static constexpr int kFac = FactorLayout<CH>::kSize;
  static constexpr int kLam = 3 * kMaxContacts;
  static constexpr int kSize = kRows + kFac + kLam + 4 * kMaxCand;
};

template <typename T, int CH>
SOLO_HD void physics_step(const ModelT<T> &M, const ConfigT<T> &C, EnvState<T, CH> &s, const T (&cmd)[kLegs][CH],
                          T *scratch, StepDiag *diag, const DebugOut<T> *dbg) {
  typedef RowLayout<CH> RL;
  typedef HostScratch<T, CH> HS;
  T *rows = scratch, *fac = scratch + HS::kRows, *lam = fac + HS::kFac, *ctab = lam + HS::kLam;
  for (int c = 0; c < M.n_cand; c++) {
    ctab[4 * c] = M.cand_pos[c][0]; ctab[4 * c + 1] = M.cand_pos[c][1]; ctab[4 * c + 2] = M.cand_pos[c][2];
    ctab[4 * c + 3] = M.cand_mu[c];
  }
  const RowHeader h = setup_step<T, CH>(M, C, s, cmd, ctab, rows, RL::kRow, fac, 1);
  uint32_t motor_mask = 0;
  if (C.control_mode == 0)
    for (int l = 0; l < M.n_legs; l++)
      for (int i = 0; i < M.chain_len[l]; i++) motor_mask |= 1u << (l * CH + i);
  T yb[6], yl[kLegs][CH], mlam[kLegs * CH];
  const int it = solve_rows<T, CH>(C, motor_mask, rows, RL::kRow, h, lam, 1, yb, yl, mlam);
  finish_step<T, CH>(M, C, s, fac, 1, yb, yl);
  if (diag) { diag->n_contacts = (int)h.nc; diag->iterations = it; }
  if (dbg) {
    if (dbg->normal_force) {
      for (int c = 0; c < M.n_cand; c++) dbg->normal_force[c * dbg->stride] = 0;
      for (int i = 0; i < (int)h.nc; i++) {
        const T *nr = rows + (int64_t)(RL::kMotorSlots + 3 * i) * RL::kRow;
        dbg->normal_force[(int)nr[RL::kX] * dbg->stride] = lam[3 * i] * nr[RL::kIs] * C.inv_dt;
      }
    }
    if (dbg->motor_torque) {
      for (int l = 0; l < M.n_legs; l++)
        for (int i = 0; i < M.chain_len[l]; i++) {
          const T *mr = rows + (int64_t)(l * CH + i) * RL::kRow;
          dbg->motor_torque[(M.dof0[l] + i) * dbg->stride] = mlam[l * CH + i] * mr[RL::kIs] * C.inv_dt;
        }
    }
  }
}

}  // namespace solo
#endif  // SOLO_FINISH_H_
