// solo_contact.cuh -- contact rows built by the solver warp, one contact per lane
// (device only).  Same arithmetic as the contact section of setup_step
// (solo_setup.h: add_contacts / finish_row, which tests/hostsim runs on the CPU),
// fed from the per-environment KinLayout record that setup_kernel leaves: the
// serial part of a physics step (kinematics, factorisation) stays one thread per
// environment, the part whose size depends on the pose (13 contacts on average,
// up to 64) is spread over the lanes of the environment's warp and never goes
// through memory.
#ifndef SOLO_CONTACT_CUH_
#define SOLO_CONTACT_CUH_

#include "solo_phases.h"

namespace solo {

// device support-point table: [kMaxCand][8] = body-frame xyz, combined friction,
// leg (-1: base) and depth in the chain (as floats), 2 spare
constexpr int kCandStride = 8;

struct Cand {
  float x, y, z, mu;
  int leg, depth;
};
__device__ __forceinline__ Cand load_cand(const float *tab, const int c) {
  const float4 a = __ldg(reinterpret_cast<const float4 *>(tab + kCandStride * c));
  const float4 b = __ldg(reinterpret_cast<const float4 *>(tab + kCandStride * c + 4));
  Cand k;
  k.x = a.x; k.y = a.y; k.z = a.z; k.mu = a.w;
  k.leg = (int)b.x; k.depth = (int)b.y;
  return k;
}

// frame (origin, rotation) of the body the support point belongs to
template <int CH>
__device__ __forceinline__ const float *cand_frame(const float *kin, const Cand &k) {
  typedef KinLayout<CH> KL;
  // base: o0 (3) + R0 (9) are contiguous at kO0; joints: origin at +3, rotation at +6
  return k.leg < 0 ? kin + KL::kO0 : kin + KL::kLeg + k.leg * KL::kLegSize + KL::kJoint + k.depth * 15 + 3;
}

// signed distance of the support point to the ground plane (margin removed)
template <int CH>
__device__ __forceinline__ float cand_distance(const float *kin, const Cand &k, const float margin) {
  const float *f = cand_frame<CH>(kin, k);
  return (f[2] + f[3 + 6] * k.x + f[3 + 7] * k.y + f[3 + 8] * k.z) - margin;
}

// The contact's rows in RowLayout<CH> form: NROWS = 1 (normal only) or 3 (normal,
// t1, t2).  cand_index goes to kX of the normal row.
template <int CH, int NROWS>
__device__ __forceinline__ void build_contact_rows(const float *kin, const ConfigT<float> &C, const float margin,
                                                   const Cand &k, const int cand_index,
                                                   float (&out)[NROWS][RowLayout<CH>::kRow]) {
  typedef RowLayout<CH> RL;
  typedef KinLayout<CH> KL;
  const float *f = cand_frame<CH>(kin, k);
  const float *Rb = f + 3;
  const float px = f[0] + Rb[0] * k.x + Rb[1] * k.y + Rb[2] * k.z;
  const float py = f[1] + Rb[3] * k.x + Rb[4] * k.y + Rb[5] * k.z;
  const float pz = f[2] + Rb[6] * k.x + Rb[7] * k.y + Rb[8] * k.z;
  const float dist = pz - margin;
  const V3<float> pc = mk(px, py, dist);
  const V3<float> r = pc - mk(kin[KL::kC0], kin[KL::kC0 + 1], kin[KL::kC0 + 2]);
  const float *vb = kin + KL::kVb;
  V3<float> vp = mk(vb[0], vb[1], vb[2]) + cross(mk(vb[3], vb[4], vb[5]), r);
  const bool on_leg = k.leg >= 0;
  const float *kl = kin + KL::kLeg + (on_leg ? k.leg : 0) * KL::kLegSize;
  V3<float> w[CH];
#pragma unroll
  for (int j = 0; j < CH; j++) {
    if (on_leg && j <= k.depth) {
      const float *kj = kl + KL::kJoint + j * 15;
      w[j] = cross(mk(kj[0], kj[1], kj[2]), pc - mk(kj[3], kj[4], kj[5]));
      vp = vp + w[j] * kl[KL::kVl + j];
    } else {
      w[j] = mk(0.f, 0.f, 0.f);
    }
  }
  float S[21];
#pragma unroll
  for (int i = 0; i < 21; i++) S[i] = kin[KL::kLs + i];
  float is_n = 0.f;
  // directions: n = +z, t1 = -y, t2 = +x  (btPlaneSpace1 of (0,0,1))
#pragma unroll
  for (int d = 0; d < NROWS; d++) {
    float (&o)[RL::kRow] = out[d];
    float g[6], jl[CH];
    if (d == 0) {
      g[0] = 0; g[1] = 0; g[2] = 1; g[3] = r.y; g[4] = -r.x; g[5] = 0;
#pragma unroll
      for (int j = 0; j < CH; j++) jl[j] = w[j].z;
    } else if (d == 1) {
      g[0] = 0; g[1] = -1; g[2] = 0; g[3] = r.z; g[4] = 0; g[5] = -r.x;
#pragma unroll
      for (int j = 0; j < CH; j++) jl[j] = -w[j].y;
    } else {
      g[0] = 1; g[1] = 0; g[2] = 0; g[3] = 0; g[4] = r.z; g[5] = -r.y;
#pragma unroll
      for (int j = 0; j < CH; j++) jl[j] = w[j].x;
    }
    float den = 0.f;
    {
      int e = KL::kE;
#pragma unroll
      for (int i = 0; i < CH; i++) {
        float t = 0.f;
#pragma unroll
        for (int kk = 0; kk <= i; kk++) t += kl[e++] * jl[kk];
        t = on_leg ? t : 0.f;
        o[6 + i] = t;
        den += t * t;
#pragma unroll
        for (int a = 0; a < 6; a++) g[a] -= kl[KL::kBh + i * 6 + a] * t;
      }
    }
    float verr, extra;
    if (d == 0) {
      const float pen = dist + C.slop;
      verr = -vp.z;   // restitution 0
      if (pen > 0) verr -= pen * C.inv_dt; else verr += -pen * C.erp * C.inv_dt;
      extra = (float)cand_index;
    } else {
      verr = (d == 1) ? vp.y : -vp.x;
      extra = (d == 1) ? k.mu * is_n : 0.f;
    }
    {
      int e = 0;
#pragma unroll
      for (int a = 0; a < 6; a++) {
        float t = 0.f;
#pragma unroll
        for (int kk = 0; kk <= a; kk++) t += S[e++] * g[kk];
        o[a] = t;
        den += t * t;
      }
    }
    const float sn = sqrtf(den);
    const float is = 1.f / sn;
#pragma unroll
    for (int a = 0; a < RL::kJ; a++) o[a] *= is;
    o[RL::kB] = verr * is; o[RL::kS] = sn; o[RL::kIs] = is; o[RL::kX] = extra;
#pragma unroll
    for (int a = RL::kX + 1; a < RL::kRow; a++) o[a] = 0.f;
    if (d == 0) is_n = is;
  }
}

// One row of the contact, direction chosen at run time (0 normal, 1 t1, 2 t2): the same arithmetic as
// row `dir` of build_contact_rows<CH, 3>, so that three lanes of a warp can build the three rows of a
// promoted contact side by side from one instruction stream.  kX of a friction row is left to the
// caller (it needs 1 / s of the normal row).
template <int CH>
__device__ __forceinline__ void build_contact_row_dir(const float *kin, const ConfigT<float> &C, const float margin,
                                                      const Cand &k, const int cand_index, const int dir,
                                                      float (&o)[RowLayout<CH>::kRow]) {
  typedef RowLayout<CH> RL;
  typedef KinLayout<CH> KL;
  const float *f = cand_frame<CH>(kin, k);
  const float *Rb = f + 3;
  const float px = f[0] + Rb[0] * k.x + Rb[1] * k.y + Rb[2] * k.z;
  const float py = f[1] + Rb[3] * k.x + Rb[4] * k.y + Rb[5] * k.z;
  const float pz = f[2] + Rb[6] * k.x + Rb[7] * k.y + Rb[8] * k.z;
  const float dist = pz - margin;
  const V3<float> pc = mk(px, py, dist);
  const V3<float> r = pc - mk(kin[KL::kC0], kin[KL::kC0 + 1], kin[KL::kC0 + 2]);
  const float *vb = kin + KL::kVb;
  V3<float> vp = mk(vb[0], vb[1], vb[2]) + cross(mk(vb[3], vb[4], vb[5]), r);
  const bool on_leg = k.leg >= 0;
  const float *kl = kin + KL::kLeg + (on_leg ? k.leg : 0) * KL::kLegSize;
  V3<float> w[CH];
#pragma unroll
  for (int j = 0; j < CH; j++) {
    if (on_leg && j <= k.depth) {
      const float *kj = kl + KL::kJoint + j * 15;
      w[j] = cross(mk(kj[0], kj[1], kj[2]), pc - mk(kj[3], kj[4], kj[5]));
      vp = vp + w[j] * kl[KL::kVl + j];
    } else {
      w[j] = mk(0.f, 0.f, 0.f);
    }
  }
  // directions: n = +z, t1 = -y, t2 = +x  (btPlaneSpace1 of (0,0,1))
  const bool d0 = dir == 0, d1 = dir == 1;
  float g[6], jl[CH];
  g[0] = d0 ? 0.f : d1 ? 0.f : 1.f;
  g[1] = d0 ? 0.f : d1 ? -1.f : 0.f;
  g[2] = d0 ? 1.f : 0.f;
  g[3] = d0 ? r.y : d1 ? r.z : 0.f;
  g[4] = d0 ? -r.x : d1 ? 0.f : r.z;
  g[5] = d0 ? 0.f : d1 ? -r.x : -r.y;
#pragma unroll
  for (int j = 0; j < CH; j++) jl[j] = d0 ? w[j].z : d1 ? -w[j].y : w[j].x;
  float den = 0.f;
  {
    int e = KL::kE;
#pragma unroll
    for (int i = 0; i < CH; i++) {
      float t = 0.f;
#pragma unroll
      for (int kk = 0; kk <= i; kk++) t += kl[e++] * jl[kk];
      t = on_leg ? t : 0.f;
      o[6 + i] = t;
      den += t * t;
#pragma unroll
      for (int a = 0; a < 6; a++) g[a] -= kl[KL::kBh + i * 6 + a] * t;
    }
  }
  float verr;
  {
    const float pen = dist + C.slop;
    float vn = -vp.z;   // restitution 0
    if (pen > 0) vn -= pen * C.inv_dt; else vn += -pen * C.erp * C.inv_dt;
    verr = d0 ? vn : d1 ? vp.y : -vp.x;
  }
  {
    int e = 0;
#pragma unroll
    for (int a = 0; a < 6; a++) {
      float t = 0.f;
#pragma unroll
      for (int kk = 0; kk <= a; kk++) t += kin[KL::kLs + e++] * g[kk];
      o[a] = t;
      den += t * t;
    }
  }
  const float sn = sqrtf(den);
  const float is = 1.f / sn;
#pragma unroll
  for (int a = 0; a < RL::kJ; a++) o[a] *= is;
  o[RL::kB] = verr * is; o[RL::kS] = sn; o[RL::kIs] = is; o[RL::kX] = d0 ? (float)cand_index : 0.f;
#pragma unroll
  for (int a = RL::kX + 1; a < RL::kRow; a++) o[a] = 0.f;
}

}  // namespace solo
#endif  // SOLO_CONTACT_CUH_
