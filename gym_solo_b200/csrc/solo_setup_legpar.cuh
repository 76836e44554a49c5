// solo_setup_legpar.cuh -- setup_step<float, CH, false> (solo_setup.h) with the four
// legs of an environment on four adjacent lanes.
//
// setup_kernel runs one thread per environment: at the bench size that is one warp
// per SM walking ~7.7 k dependent instructions, with its per-leg arrays in local
// memory (the leg loops are rolled for the instruction cache) -- a pure latency
// chain.  The robot is a star, so the per-leg part (chain kinematics, bias forces,
// the leg block of the mass matrix, E, Bh) is independent work: here lane (4 e + l)
// does leg l of environment e with everything in registers, the legs' contributions
// to the base totals and to the Schur complement are summed with two xor-shuffles
// (every lane of the group ends with bit-identical sums), the 6 x 6 base part runs
// redundantly on the four lanes, and each lane writes its own leg's factors, motor
// rows and joint frames.  Same outputs (FactorLayout, KinLayout) as setup_step;
// the sums over legs are associated differently, i.e. float32 rounding only.
// Used for batches where the setup phase is latency bound; the one-thread-per-env
// kernel stays for large batches (4 x the lanes, ~1.8 x the instructions).
#ifndef SOLO_SETUP_LEGPAR_CUH_
#define SOLO_SETUP_LEGPAR_CUH_

#include "solo_kernels.cuh"

namespace solo {

__device__ __forceinline__ float grp4_sum(float x) {
  x += __shfl_xor_sync(0xffffffffu, x, 1);
  x += __shfl_xor_sync(0xffffffffu, x, 2);
  return x;
}
__device__ __forceinline__ V3<float> grp4_sum(V3<float> v) {
  return mk(grp4_sum(v.x), grp4_sum(v.y), grp4_sum(v.z));
}

template <int CH>
__global__ void __launch_bounds__(256)
setup_legpar_kernel(const __grid_constant__ KParams K, const Buffers B, const float *__restrict__ actions,
                    const int broadcast, const float action_scale) {
  extern __shared__ float smem[];
  typedef RowLayout<CH> RL;
  typedef KinLayout<CH> KL;
  typedef FactorLayout<CH> FL;
  typedef float T;
  const ModelT<float> &M = K.M;
  const ConfigT<float> &C = K.C;
  const int epb = blockDim.x >> 2;                     // environments per block
  const int slot = threadIdx.x >> 2, l = threadIdx.x & 3;
  const int64_t env0 = (int64_t)blockIdx.x * epb;
  const int n_valid = (int)min((int64_t)epb, (int64_t)B.n - env0);
  const bool live = slot < n_valid;
  // a dead group shadows the last environment so that its shuffles stay defined; it writes nothing
  const int64_t env = live ? env0 + slot : (int64_t)B.n - 1;
  const int n_api = M.n_api;
  const float dt = C.dt;

  // ---- state: base + own leg (the loads fly while the action tile is gathered) ----
  const float *src = B.state + env;
  const int64_t st = B.n;
  float pos[3], quat[4], vlin[3], vang[3], q[CH], qd[CH];
  bool act[CH];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    pos[i] = src[i * st]; vlin[i] = src[(7 + i) * st]; vang[i] = src[(10 + i) * st];
  }
#pragma unroll
  for (int i = 0; i < 4; i++) quat[i] = src[(3 + i) * st];
#pragma unroll
  for (int j = 0; j < CH; j++) {
    act[j] = (l < M.n_legs) && (j < M.chain_len[l]);
    const int d = act[j] ? M.dof0[l] + j : 0;
    q[j] = act[j] ? src[(kStateBase + d) * st] : 0.f;
    qd[j] = act[j] ? src[(kStateBase + M.ndof + d) * st] : 0.f;
  }
  // ---- actions: [env][n_api] tile -> shared memory, asynchronously.  Torque control needs them in
  // the bias forces below; position control only in the motor rows at the very end ----
  float cmd[CH];
  const bool cmd_late = !broadcast && C.control_mode == 0;
  auto read_cmd = [&]() {
    const float *a = broadcast ? actions : smem + (live ? slot : 0) * pad_stride(n_api);
#pragma unroll
    for (int j = 0; j < CH; j++) cmd[j] = act[j] ? a[M.dof_api[M.dof0[l] + j]] * action_scale : 0.f;
  };
  if (!broadcast) tile_gather_async(actions, env0, n_valid, n_api, smem);
  if (!cmd_late) {
    if (!broadcast) tile_gather_wait();
    read_cmd();
  } else {
#pragma unroll
    for (int j = 0; j < CH; j++) cmd[j] = 0.f;
  }
  const float base_scale = B.base_mass_scale ? B.base_mass_scale[env] : 1.f;

  // ---- base kinematics (all four lanes) ----
  const M3<T> R0 = quat_to_mat(quat);
  const V3<T> o0 = ld3(pos);
  const V3<T> c0 = o0 + mul(R0, ld3(M.com[0]));
  const V3<T> w0 = ld3(vang), vc0 = ld3(vlin);
  const V3<T> grav = ld3(C.gravity);

  // ---- own leg: root -> tip kinematics and bias forces ----
  V3<T> jo[CH], ja[CH], bc[CH], F[CH], N[CH];
  M3<T> jR[CH];
  S3<T> bI[CH];
  T bm[CH];
  {
    M3<T> Rp = R0;
    V3<T> op = o0, cp = c0, wp = w0, vcp = vc0;
    V3<T> alp = mk(0.f, 0.f, 0.f), accp = alp;
#pragma unroll
    for (int j = 0; j < CH; j++) {
      const int d = act[j] ? M.dof0[l] + j : 0;
      const int b = d + 1;
      if (act[j]) {
        const V3<T> ax = ld3(M.jaxis[d]);
        M3<T> Jr;
#pragma unroll
        for (int k = 0; k < 9; k++) Jr.m[k] = M.jrot[d][k];
        const M3<T> Rj = mul(mul(Rp, Jr), axis_angle(ax, q[j]));
        const V3<T> oj = op + mul(Rp, ld3(M.jorigin[d]));
        const V3<T> aj = mul(Rj, ax);
        const V3<T> cj = oj + mul(Rj, ld3(M.com[b]));
        const S3<T> Iw = rotate_inertia(Rj, M.inertia[b]);
        const T qdj = qd[j];
        const V3<T> wj = wp + aj * qdj;
        const V3<T> rp = oj - cp, rc = cj - oj;
        const V3<T> vo = vcp + cross(wp, rp);
        const V3<T> vcj = vo + cross(wj, rc);
        const V3<T> alj = alp + cross(wp, aj) * qdj;
        const V3<T> acco = accp + cross(alp, rp) + cross(wp, cross(wp, rp));
        const V3<T> accc = acco + cross(alj, rc) + cross(wj, cross(wj, rc));
        const T mass = M.mass[b];
        const T vn = m_sqrt(dot(vcj, vcj)), wn = m_sqrt(dot(wj, wj));
        const V3<T> Iww = mul(Iw, wj);
        const V3<T> fext = grav * mass - vcj * (mass * (C.lin_damp + C.lin_damp * vn));
        const V3<T> next = Iww * (-(C.ang_damp + C.ang_damp * wn));
        F[j] = accc * mass - fext;
        N[j] = mul(Iw, alj) + cross(wj, Iww) - next;
        jo[j] = oj; ja[j] = aj; jR[j] = Rj;
        bc[j] = cj; bI[j] = Iw; bm[j] = mass;
        Rp = Rj; op = oj; cp = cj; wp = wj; vcp = vcj; alp = alj; accp = accc;
      } else {
        jo[j] = o0; ja[j] = mk(0.f, 0.f, 0.f); jR[j] = R0;
        bc[j] = c0; bm[j] = 0.f;
        F[j] = mk(0.f, 0.f, 0.f); N[j] = F[j];
        bI[j].xx = bI[j].xy = bI[j].xz = bI[j].yy = bI[j].yz = bI[j].zz = 0.f;
      }
    }
  }
  // ---- own leg: tip -> root, generalised bias forces, composite inertia, leg block ----
  // accumulated about each joint's own origin, shifted link by link (see setup_step): no
  // cancellation in the leg block of a few-gram distal link
  V3<T> Fs = mk(0.f, 0.f, 0.f), Ns = Fs, ch = Fs;
  T cm = 0.f;
  S3<T> cI; cI.xx = cI.xy = cI.xz = cI.yy = cI.yz = cI.zz = 0.f;
  V3<T> Bp[CH], BL[CH];
  T D[CH][CH], fl[CH];
#pragma unroll
  for (int j = CH - 1; j >= 0; j--) {
    if (j < CH - 1) {
      const V3<T> d = jo[j + 1] - jo[j];
      Ns = Ns + cross(d, Fs);
      shift_composite(cm, ch, cI, d);
    }
    const V3<T> r = bc[j] - jo[j];
    Fs = Fs + F[j];
    Ns = Ns + cross(r, F[j]) + N[j];
    const T Cj = dot(ja[j], Ns);
    T tau = 0.f;
    if (C.control_mode == 1) tau = m_clamp(cmd[j], -C.torque_limit, C.torque_limit);
    fl[j] = act[j] ? tau - Cj : 0.f;
    const T mb = bm[j];
    cm += mb;
    ch = ch + r * mb;
    const T rr = dot(r, r);
    cI.xx += bI[j].xx + mb * (rr - r.x * r.x); cI.xy += bI[j].xy - mb * r.x * r.y;
    cI.xz += bI[j].xz - mb * r.x * r.z;        cI.yy += bI[j].yy + mb * (rr - r.y * r.y);
    cI.yz += bI[j].yz - mb * r.y * r.z;        cI.zz += bI[j].zz + mb * (rr - r.z * r.z);
    const V3<T> p = cross(ja[j], ch);
    const V3<T> n = mul(cI, ja[j]);
    Bp[j] = p; BL[j] = n + cross(jo[j] - c0, p);
#pragma unroll
    for (int jj = 0; jj <= j; jj++) {
      const T v = dot(ja[jj], n + cross(jo[j] - jo[jj], p));
      D[j][jj] = v; D[jj][j] = v;
    }
    if (!act[j]) D[j][j] = 1.f;   // dummy unit dof, zero coupling
  }
  {
    // leg totals about the base COM
    const V3<T> d = jo[0] - c0;
    Ns = Ns + cross(d, Fs);
    shift_composite(cm, ch, cI, d);
  }
  chol_inverse_lower<T, CH>(D);   // D <- C^-1 (lower) = E
  T Bh[CH][6], fh[CH];
  T gfp[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  T Sp[21];                        // this leg's -Bh^T Bh, lower triangle row major
#pragma unroll
  for (int i = 0; i < 21; i++) Sp[i] = 0.f;
#pragma unroll
  for (int i = 0; i < CH; i++) {
    T b6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    T f = 0.f;
#pragma unroll
    for (int k = 0; k <= i; k++) {
      const T e = D[i][k];
      b6[0] += e * Bp[k].x; b6[1] += e * Bp[k].y; b6[2] += e * Bp[k].z;
      b6[3] += e * BL[k].x; b6[4] += e * BL[k].y; b6[5] += e * BL[k].z;
      f += e * fl[k];
    }
    fh[i] = f;
    int idx = 0;
#pragma unroll
    for (int a = 0; a < 6; a++) {
      Bh[i][a] = b6[a];
      gfp[a] += b6[a] * f;
#pragma unroll
      for (int b = 0; b <= a; b++) Sp[idx++] -= b6[a] * b6[b];
    }
  }

  // ---- sums over the four legs (every lane of the group gets the same bits) ----
  Fs = grp4_sum(Fs); Ns = grp4_sum(Ns); ch = grp4_sum(ch);
  cm = grp4_sum(cm);
  cI.xx = grp4_sum(cI.xx); cI.xy = grp4_sum(cI.xy); cI.xz = grp4_sum(cI.xz);
  cI.yy = grp4_sum(cI.yy); cI.yz = grp4_sum(cI.yz); cI.zz = grp4_sum(cI.zz);
  T gf[6];
#pragma unroll
  for (int a = 0; a < 6; a++) gf[a] = grp4_sum(gfp[a]);
  T S[6][6];
  {
    int idx = 0;
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
      for (int b = 0; b < 6; b++) {
        if (b <= a) S[a][b] = grp4_sum(Sp[idx++]);
        else S[a][b] = 0.f;
      }
  }

  // ---- base totals and Schur complement (all four lanes) ----
  const T mass0 = M.mass[0] * base_scale;
  S3<T> Itot = rotate_inertia(R0, M.inertia[0]);
  Itot.xx *= base_scale; Itot.xy *= base_scale; Itot.xz *= base_scale;
  Itot.yy *= base_scale; Itot.yz *= base_scale; Itot.zz *= base_scale;
  V3<T> Ftot, Ntot;
  {
    const T vn = m_sqrt(dot(vc0, vc0)), wn = m_sqrt(dot(w0, w0));
    const V3<T> Iw = mul(Itot, w0);
    const V3<T> fext = grav * mass0 - vc0 * (mass0 * (C.lin_damp + C.lin_damp * vn));
    const V3<T> next = Iw * (-(C.ang_damp + C.ang_damp * wn));
    Ftot = mk(0.f, 0.f, 0.f) - fext + Fs;
    Ntot = cross(w0, Iw) - next + Ns;
  }
  const T mtot = mass0 + cm;
  const V3<T> htot = ch;
  Itot.xx += cI.xx; Itot.xy += cI.xy; Itot.xz += cI.xz; Itot.yy += cI.yy; Itot.yz += cI.yz; Itot.zz += cI.zz;
  S[0][0] += mtot; S[1][1] += mtot; S[2][2] += mtot;
  S[3][1] += -htot.z; S[3][2] += htot.y;
  S[4][0] += htot.z;  S[4][2] += -htot.x;
  S[5][0] += -htot.y; S[5][1] += htot.x;
  S[3][3] += Itot.xx; S[4][3] += Itot.xy; S[5][3] += Itot.xz;
  S[4][4] += Itot.yy; S[5][4] += Itot.yz; S[5][5] += Itot.zz;
  chol_inverse_lower<T, 6>(S);   // S <- Ls (lower)

  // ---- unconstrained acceleration, v += a dt ----
  T vb[6] = {vlin[0], vlin[1], vlin[2], vang[0], vang[1], vang[2]};
  T vl[CH];
  {
    const T g[6] = {-Ftot.x - gf[0], -Ftot.y - gf[1], -Ftot.z - gf[2], -Ntot.x - gf[3], -Ntot.y - gf[4], -Ntot.z - gf[5]};
    T fb[6], ab[6];
#pragma unroll
    for (int i = 0; i < 6; i++) {
      T t = 0.f;
#pragma unroll
      for (int k = 0; k <= i; k++) t += S[i][k] * g[k];
      fb[i] = t;
    }
#pragma unroll
    for (int i = 0; i < 6; i++) {
      T t = 0.f;
#pragma unroll
      for (int k = i; k < 6; k++) t += S[k][i] * fb[k];
      ab[i] = t;
      vb[i] = m_clamp(vb[i] + t * dt, -C.max_vel, C.max_vel);
    }
    T t[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) {
      T u = fh[i];
#pragma unroll
      for (int a = 0; a < 6; a++) u -= Bh[i][a] * ab[a];
      t[i] = u;
    }
#pragma unroll
    for (int i = 0; i < CH; i++) {
      T acc = 0.f;
#pragma unroll
      for (int k = i; k < CH; k++) acc += D[k][i] * t[k];
      vl[i] = m_clamp(qd[i] + acc * dt, -C.max_vel, C.max_vel);
    }
  }
  // ---- joint limits of the own leg; the environment's mask is the OR over the four lanes ----
  T lpen[CH];
  int ldir[CH];
  uint32_t lim_bits = 0;
#pragma unroll
  for (int i = 0; i < CH; i++) {
    lpen[i] = 0.f;
    ldir[i] = 0;
    if (act[i]) {
      const int api = M.dof_api[M.dof0[l] + i];
      ldir[i] = joint_limit_state(q[i], M.api_lower[api], M.api_upper[api], &lpen[i]);
    }
    if (ldir[i] != 0) lim_bits |= (1u << (l * CH + i)) | (ldir[i] < 0 ? (1u << (16 + l * CH + i)) : 0u);
  }
  lim_bits |= __shfl_xor_sync(0xffffffffu, lim_bits, 1);
  lim_bits |= __shfl_xor_sync(0xffffffffu, lim_bits, 2);

  // The records of the block's environments are contiguous in memory ([env][KinLayout::kSize]):
  // they are assembled in shared memory and leave as ONE bulk copy (cp.async.bulk, the 1-D form of
  // TMA) issued by one thread, instead of 328 scattered 4-byte stores per environment whose queue
  // (lg_throttle) the four lanes would wait on.
  float *rec = smem + ((epb * pad_stride(n_api) + 3) & ~3);
  if (cmd_late) {
    tile_gather_wait();
    read_cmd();
  }
  if (live) {

  // ---- factors for finish_step (FactorLayout) ----
  {
    float *fac = B.fac + env;
    const int64_t fs = B.n;
    if (l == 0) {
      int k = FL::kLs;
#pragma unroll
      for (int i = 0; i < 6; i++)
#pragma unroll
        for (int j = 0; j <= i; j++) fac[(k++) * fs] = S[i][j];
#pragma unroll
      for (int i = 0; i < 6; i++) fac[(FL::kVb + i) * fs] = vb[i];
    }
    int k = FL::kLeg + l * FL::kLegSize;
#pragma unroll
    for (int i = 0; i < CH; i++)
#pragma unroll
      for (int a = 0; a < 6; a++) fac[(k++) * fs] = Bh[i][a];
#pragma unroll
    for (int i = 0; i < CH; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) fac[(k++) * fs] = D[i][j];
#pragma unroll
    for (int i = 0; i < CH; i++) fac[(FL::kVl + l * CH + i) * fs] = vl[i];
  }

  // ---- motor rows of the own leg (POSITION_CONTROL) and joint-limit rows ----
  // a joint at or beyond a URDF limit gets Bullet's unilateral limit row; it shares the motor
  // row's whitened direction (J = +-e_j), so the slot is built whenever either row is live
  float *kin = rec + slot * KL::kSize;
  const bool motors = (C.control_mode == 0);
#pragma unroll
  for (int i = 0; i < CH; i++) {
    T o[RL::kRow];
    float *dst = kin + KL::kMotor + (l * CH + i) * RL::kRow;
    if (act[i] && (motors || ldir[i] != 0)) {
      T g[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      T den = 0.f;
#pragma unroll
      for (int k = 0; k < CH; k++) {
        const T e = (k >= i) ? D[k][i] : 0.f;
        o[6 + k] = e;
        den += e * e;
#pragma unroll
        for (int a = 0; a < 6; a++) g[a] -= Bh[k][a] * e;
      }
      // btMultiBodyJointMotor: velocity error kp (q* - q)/dt - kd v
      const T verr = motors ? C.kp * (cmd[i] - q[i]) * C.inv_dt - C.kd * vl[i] : 0.f;
      // limit row (along dir e_j): positional correction -pen erp / dt, velocity error -dir v
      const T lerr = ldir[i] != 0 ? -lpen[i] * C.lim_erp * C.inv_dt - (T)ldir[i] * vl[i] : 0.f;
#pragma unroll
      for (int a = 0; a < 6; a++) {
        T t = 0.f;
#pragma unroll
        for (int kk = 0; kk <= a; kk++) t += S[a][kk] * g[kk];
        o[a] = t;
        den += t * t;
      }
      const T sn = m_sqrt(den);
      const T is = 1.f / sn;
#pragma unroll
      for (int a = 0; a < RL::kJ; a++) o[a] *= is;
      o[RL::kB] = verr * is; o[RL::kS] = sn; o[RL::kIs] = is; o[RL::kX] = lerr * is;
#pragma unroll
      for (int a = RL::kX + 1; a < RL::kRow; a++) o[a] = 0.f;
    } else {
#pragma unroll
      for (int a = 0; a < RL::kRow; a++) o[a] = 0.f;
    }
    store_row(dst, o);
  }

  // ---- the record the solver warp builds contact rows from (KinLayout) ----
  if (l == 0) {
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
    kin[KL::kLim] = __uint_as_float(lim_bits);
#pragma unroll
    for (int i = KL::kLim + 1; i < KL::kSize; i++) kin[i] = 0.f;   // padding of the record
  }
  {
    float *kl = kin + KL::kLeg + l * KL::kLegSize;
    int e = KL::kE;
#pragma unroll
    for (int i = 0; i < CH; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) kl[e++] = D[i][j];
#pragma unroll
    for (int i = 0; i < CH; i++)
#pragma unroll
      for (int a = 0; a < 6; a++) kl[KL::kBh + i * 6 + a] = Bh[i][a];
#pragma unroll
    for (int i = 0; i < CH; i++) kl[KL::kVl + i] = vl[i];
#pragma unroll
    for (int j = 0; j < CH; j++) {
      float *kj = kl + KL::kJoint + j * 15;
      kj[0] = ja[j].x; kj[1] = ja[j].y; kj[2] = ja[j].z;
      kj[3] = jo[j].x; kj[4] = jo[j].y; kj[5] = jo[j].z;
#pragma unroll
      for (int i = 0; i < 9; i++) kj[6 + i] = jR[j].m[i];
    }
  }
  }   // live
  __syncthreads();
  if (threadIdx.x == 0 && n_valid > 0) {
    const uint32_t src = (uint32_t)__cvta_generic_to_shared(rec);
    float *dst = B.kin + env0 * KL::kSize;
    const uint32_t bytes = (uint32_t)n_valid * KL::kSize * sizeof(float);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the generic-proxy writes above, before the async copy reads them
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // complete (and visible) before the block exits
  }
}

}  // namespace solo
#endif  // SOLO_SETUP_LEGPAR_CUH_
