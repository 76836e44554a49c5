// solo_setup.h -- phase 1 of the physics step (see solo_phases.h): kinematics,
// Newton-Euler bias forces, composite inertias, the star-topology block
// factorisation of the mass matrix, the unconstrained velocity update, and the
// whitened + normalised constraint rows (motor rows, then contact rows grouped
// base / leg 0 / ... / leg 3).
#ifndef SOLO_SETUP_H_
#define SOLO_SETUP_H_

// The per-leg loops stay rolled on the device: the setup kernel runs ONE warp per SM
// sub-partition at the bench size, i.e. straight-line code is fetched cold exactly
// once per warp, and the unrolled program (8.5 k SASS instructions, 137 KB) was
// bound by instruction-cache misses (ncu: stall_no_instruction 5.7 per issue).
#ifdef __CUDA_ARCH__
#define SOLO_LEG_LOOP _Pragma("unroll 1")
#else
#define SOLO_LEG_LOOP
#endif

namespace solo {

// cand_tab: [n_cand][4] = body-frame support point xyz + combined friction (a
// lane-indexed copy of M.cand_pos / M.cand_mu: shared memory on the device).
// rows: row r lives at rows[r * slot_stride + 0..kRow); fac: fac[k * fac_stride].
// Returns the header (contact counts per group).
//
// CONTACTS = false (the CUDA path): stop after the motor rows and leave, in
// `kin` (KinLayout<CH>, one contiguous record per environment, motor rows
// included: pass rows = kin + KinLayout::kMotor), what the solver warp needs to
// build the contact rows itself, one contact per lane (solo_contact.cuh).
template <typename T, int CH, bool CONTACTS = true>
SOLO_HD RowHeader setup_step(const ModelT<T> &M, const ConfigT<T> &C, const EnvState<T, CH> &s,
                             const T (&cmd)[kLegs][CH], const T *cand_tab, T *rows, int64_t slot_stride, T *fac,
                             int64_t fac_stride, T *kin = nullptr, T base_scale = T(1)) {
  typedef RowLayout<CH> RL;
  typedef FactorLayout<CH> FL;
  const T dt = C.dt;

  // ---------------- base kinematics ----------------
  const M3<T> R0 = quat_to_mat(s.quat);
  const V3<T> o0 = ld3(s.pos);
  const V3<T> c0 = o0 + mul(R0, ld3(M.com[0]));
  const V3<T> w0 = ld3(s.vang), vc0 = ld3(s.vlin);
  const V3<T> grav = ld3(C.gravity);

  // totals about c0: bias force/moment and composite inertia (mass, m*r, I_O)
  V3<T> Ftot, Ntot;
  // base_scale: per-environment factor on the base body's mass and inertia (solo_bind_base_mass_scale)
  const T mass0 = M.mass[0] * base_scale;
  T mtot = mass0;
  V3<T> htot = mk(T(0), T(0), T(0));
  S3<T> Itot = rotate_inertia(R0, M.inertia[0]);
  Itot.xx *= base_scale; Itot.xy *= base_scale; Itot.xz *= base_scale;
  Itot.yy *= base_scale; Itot.yz *= base_scale; Itot.zz *= base_scale;
  {
    T vn = m_sqrt(dot(vc0, vc0)), wn = m_sqrt(dot(w0, w0));
    V3<T> Iw = mul(Itot, w0);
    V3<T> fext = grav * mass0 - vc0 * (mass0 * (C.lin_damp + C.lin_damp * vn));
    V3<T> next = Iw * (-(C.ang_damp + C.ang_damp * wn));
    Ftot = mk(T(0), T(0), T(0)) - fext;
    Ntot = cross(w0, Iw) - next;
  }

  // Schur complement accumulator (full 6x6, lower part used) and reduced force
  T S[6][6];
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j < 6; j++) S[i][j] = 0;
  T gf[6] = {0, 0, 0, 0, 0, 0};  // sum_l Bh_l^T fh_l

  // per-leg data kept for the row phase
  V3<T> jo[kLegs][CH], ja[kLegs][CH];   // joint origin / axis, world
  M3<T> jR[kLegs][CH];                  // body rotation
  T E[kLegs][CH][CH];                   // C^-1 of the leg block (lower)
  T Bh[kLegs][CH][6];                   // E B^T
  T fh[kLegs][CH];                      // E f_leg

SOLO_LEG_LOOP
  for (int l = 0; l < kLegs; l++) {
    V3<T> bc[CH], F[CH], N[CH];
    S3<T> bI[CH];
    T bm[CH];
    bool act[CH];
    {
      M3<T> Rp = R0;
      V3<T> op = o0, cp = c0, wp = w0, vcp = vc0;
      V3<T> alp = mk(T(0), T(0), T(0)), accp = alp;
#pragma unroll
      for (int j = 0; j < CH; j++) {
        act[j] = (l < M.n_legs) && (j < M.chain_len[l]);
        const int d = act[j] ? M.dof0[l] + j : 0;
        const int b = d + 1;
        if (act[j]) {
          const V3<T> ax = ld3(M.jaxis[d]);
          M3<T> Jr;
#pragma unroll
          for (int k = 0; k < 9; k++) Jr.m[k] = M.jrot[d][k];
          const M3<T> Rj = mul(mul(Rp, Jr), axis_angle(ax, s.q[l][j]));
          const V3<T> oj = op + mul(Rp, ld3(M.jorigin[d]));
          const V3<T> aj = mul(Rj, ax);
          const V3<T> cj = oj + mul(Rj, ld3(M.com[b]));
          const S3<T> Iw = rotate_inertia(Rj, M.inertia[b]);
          const T qd = s.qd[l][j];
          const V3<T> wj = wp + aj * qd;
          const V3<T> rp = oj - cp, rc = cj - oj;
          const V3<T> vo = vcp + cross(wp, rp);
          const V3<T> vcj = vo + cross(wj, rc);
          const V3<T> alj = alp + cross(wp, aj) * qd;
          const V3<T> acco = accp + cross(alp, rp) + cross(wp, cross(wp, rp));
          const V3<T> accc = acco + cross(alj, rc) + cross(wj, cross(wj, rc));
          const T mass = M.mass[b];
          const T vn = m_sqrt(dot(vcj, vcj)), wn = m_sqrt(dot(wj, wj));
          const V3<T> Iww = mul(Iw, wj);
          const V3<T> fext = grav * mass - vcj * (mass * (C.lin_damp + C.lin_damp * vn));
          const V3<T> next = Iww * (-(C.ang_damp + C.ang_damp * wn));
          F[j] = accc * mass - fext;
          N[j] = mul(Iw, alj) + cross(wj, Iww) - next;
          jo[l][j] = oj; ja[l][j] = aj; jR[l][j] = Rj;
          bc[j] = cj; bI[j] = Iw; bm[j] = mass;
          Rp = Rj; op = oj; cp = cj; wp = wj; vcp = vcj; alp = alj; accp = accc;
        } else {
          jo[l][j] = o0; ja[l][j] = mk(T(0), T(0), T(0)); jR[l][j] = R0;
          bc[j] = c0; bm[j] = 0;
          F[j] = mk(T(0), T(0), T(0)); N[j] = F[j];
          bI[j].xx = bI[j].xy = bI[j].xz = bI[j].yy = bI[j].yz = bI[j].zz = 0;
        }
      }
    }
    // tip -> root: bias generalised forces and composite inertia.  Everything a joint sees is
    // accumulated ABOUT THAT JOINT'S OWN ORIGIN (shifted link by link), not about the base COM:
    // the leg block D of a few-gram distal link is then a sum of small terms instead of the
    // difference of two large ones (MMR v4: joint-rate error 5e-3 -> the common tolerance).
    V3<T> Fs = mk(T(0), T(0), T(0)), Ns = Fs, ch = Fs;   // Ns, ch, cI: about jo[l][j]
    T cm = 0;
    S3<T> cI; cI.xx = cI.xy = cI.xz = cI.yy = cI.yz = cI.zz = 0;
    V3<T> Bp[CH], BL[CH];
    T D[CH][CH];
    T fl[CH];
#pragma unroll
    for (int j = CH - 1; j >= 0; j--) {
      if (j < CH - 1) {
        const V3<T> d = jo[l][j + 1] - jo[l][j];
        Ns = Ns + cross(d, Fs);
        shift_composite(cm, ch, cI, d);
      }
      const V3<T> r = bc[j] - jo[l][j];
      Fs = Fs + F[j];
      Ns = Ns + cross(r, F[j]) + N[j];
      const T Cj = dot(ja[l][j], Ns);
      T tau = 0;
      if (C.control_mode == 1) tau = m_clamp(cmd[l][j], -C.torque_limit, C.torque_limit);
      fl[j] = act[j] ? tau - Cj : T(0);
      const T mb = bm[j];
      cm += mb;
      ch = ch + r * mb;
      const T rr = dot(r, r);
      cI.xx += bI[j].xx + mb * (rr - r.x * r.x); cI.xy += bI[j].xy - mb * r.x * r.y;
      cI.xz += bI[j].xz - mb * r.x * r.z;        cI.yy += bI[j].yy + mb * (rr - r.y * r.y);
      cI.yz += bI[j].yz - mb * r.y * r.z;        cI.zz += bI[j].zz + mb * (rr - r.z * r.z);
      // momentum of the subtree under a unit rate of joint j: linear p, angular n about jo[l][j]
      const V3<T> p = cross(ja[l][j], ch);
      const V3<T> n = mul(cI, ja[l][j]);
      Bp[j] = p; BL[j] = n + cross(jo[l][j] - c0, p);   // base coupling: angular part about c0
#pragma unroll
      for (int jj = 0; jj <= j; jj++) {
        T v = dot(ja[l][jj], n + cross(jo[l][j] - jo[l][jj], p));
        D[j][jj] = v; D[jj][j] = v;
      }
      if (!act[j]) D[j][j] = 1;   // dummy unit dof, zero coupling
    }
    {
      // leg totals about the base COM
      const V3<T> d = jo[l][0] - c0;
      Ns = Ns + cross(d, Fs);
      shift_composite(cm, ch, cI, d);
    }
    Ftot = Ftot + Fs; Ntot = Ntot + Ns;
    mtot += cm; htot = htot + ch;
    Itot.xx += cI.xx; Itot.xy += cI.xy; Itot.xz += cI.xz; Itot.yy += cI.yy; Itot.yz += cI.yz; Itot.zz += cI.zz;
    chol_inverse_lower<T, CH>(D);   // D <- C^-1 (lower)
#pragma unroll
    for (int i = 0; i < CH; i++) {
#pragma unroll
      for (int k = 0; k < CH; k++) E[l][i][k] = (k <= i) ? D[i][k] : T(0);
      T b6[6] = {0, 0, 0, 0, 0, 0};
      T f = 0;
#pragma unroll
      for (int k = 0; k <= i; k++) {
        const T e = D[i][k];
        b6[0] += e * Bp[k].x; b6[1] += e * Bp[k].y; b6[2] += e * Bp[k].z;
        b6[3] += e * BL[k].x; b6[4] += e * BL[k].y; b6[5] += e * BL[k].z;
        f += e * fl[k];
      }
      fh[l][i] = f;
#pragma unroll
      for (int a = 0; a < 6; a++) {
        Bh[l][i][a] = b6[a];
        gf[a] += b6[a] * f;
#pragma unroll
        for (int b = 0; b <= a; b++) S[a][b] -= b6[a] * b6[b];
      }
    }
  }

  // ---------------- base Schur complement ----------------
  S[0][0] += mtot; S[1][1] += mtot; S[2][2] += mtot;
  // lin-ang block: p = m v + w x h  ->  row x: (0, hz, -hy) etc.; lower part = its transpose
  S[3][1] += -htot.z; S[3][2] += htot.y;
  S[4][0] += htot.z;  S[4][2] += -htot.x;
  S[5][0] += -htot.y; S[5][1] += htot.x;
  S[3][3] += Itot.xx; S[4][3] += Itot.xy; S[5][3] += Itot.xz;
  S[4][4] += Itot.yy; S[5][4] += Itot.yz; S[5][5] += Itot.zz;
  chol_inverse_lower<T, 6>(S);   // S <- Ls (lower)

  // ---------------- unconstrained acceleration, v += a dt ----------------
  T vb[6] = {s.vlin[0], s.vlin[1], s.vlin[2], s.vang[0], s.vang[1], s.vang[2]};
  T vl[kLegs][CH];
  {
    T g[6] = {-Ftot.x - gf[0], -Ftot.y - gf[1], -Ftot.z - gf[2], -Ntot.x - gf[3], -Ntot.y - gf[4], -Ntot.z - gf[5]};
    T fb[6], ab[6];
#pragma unroll
    for (int i = 0; i < 6; i++) {
      T t = 0;
#pragma unroll
      for (int k = 0; k <= i; k++) t += S[i][k] * g[k];
      fb[i] = t;
    }
#pragma unroll
    for (int i = 0; i < 6; i++) {
      T t = 0;
#pragma unroll
      for (int k = i; k < 6; k++) t += S[k][i] * fb[k];
      ab[i] = t;
      vb[i] = m_clamp(vb[i] + t * dt, -C.max_vel, C.max_vel);
    }
SOLO_LEG_LOOP
    for (int l = 0; l < kLegs; l++) {
      T t[CH];
#pragma unroll
      for (int i = 0; i < CH; i++) {
        T u = fh[l][i];
#pragma unroll
        for (int a = 0; a < 6; a++) u -= Bh[l][i][a] * ab[a];
        t[i] = u;
      }
#pragma unroll
      for (int i = 0; i < CH; i++) {
        T acc = 0;
#pragma unroll
        for (int k = i; k < CH; k++) acc += E[l][k][i] * t[k];
        vl[l][i] = m_clamp(s.qd[l][i] + acc * dt, -C.max_vel, C.max_vel);
      }
    }
  }

  // ---------------- factors for finish_step ----------------
  {
    int k = FL::kLs;
#pragma unroll
    for (int i = 0; i < 6; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) fac[(k++) * fac_stride] = S[i][j];
SOLO_LEG_LOOP
    for (int l = 0; l < kLegs; l++) {
#pragma unroll
      for (int i = 0; i < CH; i++)
#pragma unroll
        for (int a = 0; a < 6; a++) fac[(k++) * fac_stride] = Bh[l][i][a];
#pragma unroll
      for (int i = 0; i < CH; i++)
#pragma unroll
        for (int j = 0; j <= i; j++) fac[(k++) * fac_stride] = E[l][i][j];
    }
#pragma unroll
    for (int i = 0; i < 6; i++) fac[(k++) * fac_stride] = vb[i];
SOLO_LEG_LOOP
    for (int l = 0; l < kLegs; l++)
#pragma unroll
      for (int i = 0; i < CH; i++) fac[(k++) * fac_stride] = vl[l][i];
  }

  // whiten g (base part, pre-Ls) + leg part already in o[6..], normalise, store
  auto finish_row = [&](T (&o)[RL::kRow], const T (&g)[6], T den_leg, T verr, T extra, T *dst,
                        const bool scale_extra = false) {
    T den = den_leg;
#pragma unroll
    for (int a = 0; a < 6; a++) {
      T t = 0;
#pragma unroll
      for (int kk = 0; kk <= a; kk++) t += S[a][kk] * g[kk];
      o[a] = t;
      den += t * t;
    }
    const T sn = m_sqrt(den);
    const T is = T(1) / sn;
#pragma unroll
    for (int a = 0; a < RL::kJ; a++) o[a] *= is;
    o[RL::kB] = verr * is; o[RL::kS] = sn; o[RL::kIs] = is; o[RL::kX] = scale_extra ? extra * is : extra;
#pragma unroll
    for (int a = RL::kX + 1; a < RL::kRow; a++) o[a] = 0;
    store_row(dst, o);
    return is;
  };

  // ---------------- motor rows (POSITION_CONTROL) and joint-limit rows ----------------
  // A joint at or beyond a URDF limit gets Bullet's unilateral limit row; it shares the motor
  // row's whitened direction (J = +-e_j), so the slot is built whenever either row is live.
  const bool motors = (C.control_mode == 0);
  uint32_t lim_bits = 0;
SOLO_LEG_LOOP
  for (int l = 0; l < kLegs; l++)
#pragma unroll
    for (int i = 0; i < CH; i++) {
      T o[RL::kRow];
      T *dst = rows + (int64_t)(l * CH + i) * slot_stride;
      const bool on = (l < M.n_legs) && (i < M.chain_len[l]);
      T pen = 0;
      int ldir = 0;
      if (on) {
        const int api = M.dof_api[M.dof0[l] + i];
        ldir = joint_limit_state(s.q[l][i], M.api_lower[api], M.api_upper[api], &pen);
      }
      if (ldir != 0) lim_bits |= (1u << (l * CH + i)) | (ldir < 0 ? (1u << (16 + l * CH + i)) : 0u);
      const bool act = on && (motors || ldir != 0);
      if (act) {
        T g[6] = {0, 0, 0, 0, 0, 0};
        T den = 0;
#pragma unroll
        for (int k = 0; k < CH; k++) {
          const T e = (k >= i) ? E[l][k][i] : T(0);
          o[6 + k] = e;
          den += e * e;
#pragma unroll
          for (int a = 0; a < 6; a++) g[a] -= Bh[l][k][a] * e;
        }
        // btMultiBodyJointMotor: velocity error kp (q* - q)/dt - kd v
        const T verr = motors ? C.kp * (cmd[l][i] - s.q[l][i]) * C.inv_dt - C.kd * vl[l][i] : T(0);
        // limit row (along dir e_j): positional correction -pen erp / dt, velocity error -dir v
        const T lerr = ldir != 0 ? -pen * C.lim_erp * C.inv_dt - (T)ldir * vl[l][i] : T(0);
        finish_row(o, g, den, verr, lerr, dst, true);
      } else {
#pragma unroll
        for (int a = 0; a < RL::kRow; a++) o[a] = 0;
        store_row(dst, o);
      }
    }

  if (!CONTACTS) {
    typedef KinLayout<CH> KL;
    int k = KL::kLs;
#pragma unroll
    for (int i = 0; i < 6; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) kin[k++] = S[i][j];
#pragma unroll
    for (int i = 0; i < 6; i++) kin[KL::kVb + i] = vb[i];
    kin[KL::kC0] = c0.x; kin[KL::kC0 + 1] = c0.y; kin[KL::kC0 + 2] = c0.z;
    kin[KL::kO0] = o0.x; kin[KL::kO0 + 1] = o0.y; kin[KL::kO0 + 2] = o0.z;
#pragma unroll
    for (int i = 0; i < 9; i++) kin[KL::kR0 + i] = R0.m[i];
SOLO_LEG_LOOP
    for (int l = 0; l < kLegs; l++) {
      T *kl = kin + KL::kLeg + l * KL::kLegSize;
      int e = KL::kE;
#pragma unroll
      for (int i = 0; i < CH; i++)
#pragma unroll
        for (int j = 0; j <= i; j++) kl[e++] = E[l][i][j];
#pragma unroll
      for (int i = 0; i < CH; i++)
#pragma unroll
        for (int a = 0; a < 6; a++) kl[KL::kBh + i * 6 + a] = Bh[l][i][a];
#pragma unroll
      for (int i = 0; i < CH; i++) kl[KL::kVl + i] = vl[l][i];
#pragma unroll
      for (int j = 0; j < CH; j++) {
        T *kj = kl + KL::kJoint + j * 15;
        kj[0] = ja[l][j].x; kj[1] = ja[l][j].y; kj[2] = ja[l][j].z;
        kj[3] = jo[l][j].x; kj[4] = jo[l][j].y; kj[5] = jo[l][j].z;
#pragma unroll
        for (int i = 0; i < 9; i++) kj[6 + i] = jR[l][j].m[i];
      }
    }
    union { uint32_t u; float f; } lim_word;   // the mask travels as the bits of one float of the record
    lim_word.u = lim_bits;
    kin[KL::kLim] = (T)lim_word.f;
    RowHeader h0;
    h0.ends03 = 0; h0.nc = 0; h0.lim = lim_bits;
    return h0;
  }

  // ---------------- contact rows ----------------
  int nc = 0;
  uint32_t ends03 = 0;
  auto add_contacts = [&](const int body, const int l, const int depth, const V3<T> ob, const M3<T> &Rb) {
    const int cb = M.cand_begin[body], ce = M.cand_begin[body + 1];
    // pass 1 (same candidate in every lane): which support points are inside the
    // breaking threshold
    uint32_t mask = 0;
    for (int c = cb; c < ce; c++) {
      const T z = ob.z + Rb.m[6] * M.cand_pos[c][0] + Rb.m[7] * M.cand_pos[c][1] + Rb.m[8] * M.cand_pos[c][2];
      if (z - M.margin < C.breaking) mask |= 1u << (c - cb);
    }
    // pass 2: lanes walk their own hit lists in step
    while (mask != 0 && nc < C.max_contacts) {
      const int c = cb + lowest_bit(mask);
      mask &= mask - 1;
      const T *ct = cand_tab + 4 * c;
      const V3<T> p = ob + mul(Rb, mk(ct[0], ct[1], ct[2]));
      const T dist = p.z - M.margin;
      const V3<T> pc = mk(p.x, p.y, dist);
      const V3<T> r = pc - c0;
      V3<T> vp = mk(vb[0], vb[1], vb[2]) + cross(mk(vb[3], vb[4], vb[5]), r);
      V3<T> w[CH];
      const int ll = l < 0 ? 0 : l;
#pragma unroll
      for (int j = 0; j < CH; j++) {
        if (l >= 0 && j <= depth) {
          w[j] = cross(ja[ll][j], pc - jo[ll][j]);
          vp = vp + w[j] * vl[ll][j];
        } else {
          w[j] = mk(T(0), T(0), T(0));
        }
      }
      T *dst = rows + (int64_t)(RL::kMotorSlots + 3 * nc) * slot_stride;
      T is_n = 0;
      // directions: n = +z, t1 = -y, t2 = +x  (btPlaneSpace1 of (0,0,1))
#pragma unroll
      for (int k = 0; k < 3; k++) {
        T o[RL::kRow];
        T g[6], jl[CH];
        if (k == 0) {
          g[0] = 0; g[1] = 0; g[2] = 1; g[3] = r.y; g[4] = -r.x; g[5] = 0;
#pragma unroll
          for (int j = 0; j < CH; j++) jl[j] = w[j].z;
        } else if (k == 1) {
          g[0] = 0; g[1] = -1; g[2] = 0; g[3] = r.z; g[4] = 0; g[5] = -r.x;
#pragma unroll
          for (int j = 0; j < CH; j++) jl[j] = -w[j].y;
        } else {
          g[0] = 1; g[1] = 0; g[2] = 0; g[3] = 0; g[4] = r.z; g[5] = -r.y;
#pragma unroll
          for (int j = 0; j < CH; j++) jl[j] = w[j].x;
        }
        T den = 0;
        if (l >= 0) {
#pragma unroll
          for (int i = 0; i < CH; i++) {
            T t = 0;
#pragma unroll
            for (int kk = 0; kk <= i; kk++) t += E[ll][i][kk] * jl[kk];
            o[6 + i] = t;
            den += t * t;
#pragma unroll
            for (int a = 0; a < 6; a++) g[a] -= Bh[ll][i][a] * t;
          }
        } else {
#pragma unroll
          for (int i = 0; i < CH; i++) o[6 + i] = 0;
        }
        T verr, extra;
        if (k == 0) {
          const T pen = dist + C.slop;
          verr = -vp.z;   // restitution 0
          if (pen > 0) verr -= pen * C.inv_dt; else verr += -pen * C.erp * C.inv_dt;
          extra = (T)c;
        } else {
          verr = (k == 1) ? vp.y : -vp.x;
          extra = (k == 1) ? ct[3] * is_n : T(0);
        }
        const T is = finish_row(o, g, den, verr, extra, dst + (int64_t)k * slot_stride);
        if (k == 0) is_n = is;
      }
      nc++;
    }
  };
  add_contacts(0, -1, 0, o0, R0);
  ends03 = (uint32_t)nc;
#pragma unroll
  for (int l = 0; l < kLegs; l++) {
#pragma unroll
    for (int j = 0; j < CH; j++) {
      if ((l < M.n_legs) && (j < M.chain_len[l])) add_contacts(M.dof0[l] + j + 1, l, j, jo[l][j], jR[l][j]);
    }
    if (l < 3) ends03 |= (uint32_t)nc << (8 * (l + 1));
  }
  RowHeader h;
  h.ends03 = ends03;
  h.nc = (uint32_t)nc;
  h.lim = lim_bits;
  return h;
}

}  // namespace solo
#endif  // SOLO_SETUP_H_
