// solo_solve_warp.cuh -- the constraint solver of stepSimulation with ONE WARP PER
// ENVIRONMENT.  Same iterates as solo_solve.h (the sequential statement of the
// algorithm, which tests/hostsim runs on the CPU), scheduled for LATENCY: the
// launch lasts as long as its slowest environment (50 iterations x ~10 contacts
// that carry an impulse), so what counts is the length of the dependent chain of
// one Gauss-Seidel update.
//
//   * the environment's row list is copied once from the L2-resident record list
//     into shared memory (coalesced float4 copy); every later row access is a
//     broadcast LDS whose address does not depend on yh, i.e. off the chain;
//   * yh, all impulses and every piece of control flow are warp-uniform; impulses
//     live in shared memory, written and read back by every lane (no __syncwarp);
//   * lane l keeps the normal rows of contacts l and l + 32 as dense vectors in
//     registers.  ~90 % of the rows inside Bullet's 2 cm breaking threshold are
//     speculative (zero impulse, b' - u.yh <= 0: a no-op), so the normal phase
//     walks the bitmask of contacts that HAVE carried an impulse with uniform
//     updates (dot -> clamp -> axpy, ~50 cycles each) while the lanes check in
//     parallel that no speculative row between two of them has come alive under
//     the current yh; one that has is promoted on the spot, at its place in
//     Bullet's order -- exactly sequential Gauss-Seidel;
//   * the friction phase walks the same bitmask (only those contacts can hold
//     friction: limit = mu * lambda_n), cone projection without a branch.
#ifndef SOLO_SOLVE_WARP_CUH_
#define SOLO_SOLVE_WARP_CUH_

#include "solo_kernels.cuh"

namespace solo {

template <int CH>
struct WarpSolveSmem {
  float row[RowLayout<CH>::kSlots][RowLayout<CH>::kRow];   // the env's row list (motor slots, then contacts)
  float4 lam[kMaxContacts];                                // scaled impulses: normal, t1, t2, -
};

// uniform: group (0 base, 1..4 legs) of contact c from the packed group ends
__device__ __forceinline__ int contact_group(const uint32_t ends03, const int c) {
  return (c >= (int)(ends03 & 255u)) + (c >= (int)((ends03 >> 8) & 255u)) + (c >= (int)((ends03 >> 16) & 255u)) +
         (c >= (int)((ends03 >> 24) & 255u));
}

// contact bitmasks: 32 bits when the env has <= 32 contacts (99.8 % of env-steps), else 64
template <int NR> struct CMask;
template <> struct CMask<1> {
  typedef uint32_t T;
  static __device__ __forceinline__ int first(T m) { return __ffs((int)m) - 1; }
  static __device__ __forceinline__ T bit(int c) { return 1u << c; }
  static __device__ __forceinline__ T below(int c) { return c >= 32 ? ~0u : ((1u << c) - 1u); }
  static __device__ __forceinline__ int count(T m) { return __popc(m); }
};
template <> struct CMask<2> {
  typedef uint64_t T;
  static __device__ __forceinline__ int first(T m) {
    const uint32_t lo = (uint32_t)m;
    return lo ? __ffs((int)lo) - 1 : 31 + __ffs((int)(uint32_t)(m >> 32));
  }
  static __device__ __forceinline__ T bit(int c) { return 1ull << c; }
  static __device__ __forceinline__ T below(int c) { return c >= 64 ? ~0ull : ((1ull << c) - 1ull); }
  static __device__ __forceinline__ int count(T m) { return __popcll(m); }
};

__device__ __forceinline__ float fast_rsqrt(const float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// A stored row as fetched from shared memory (broadcast LDS): the record, the
// contact's impulses and its group.  Nothing here depends on yh, so a fetch can
// be issued one update ahead.
template <int CH>
struct CompactRow {
  float r[RowLayout<CH>::kRow];
  float4 lam;
  int g;
  __device__ __forceinline__ void fetch(const float *crow, const float4 *lams, const int c, const int k,
                                        const uint32_t ends03) {
    load_row(crow + (3 * c + k) * RowLayout<CH>::kRow, r);
    lam = lams[c];
    g = contact_group(ends03, c);
  }
};

template <int G> struct IntC { static constexpr int value = G; };

// (b + extra) - u . y for a compact row of group G (0 base, 1..4 leg G - 1): the
// leg entries of y it touches are static
template <int CH, int G>
__device__ __forceinline__ float row_gap(const float (&r)[RowLayout<CH>::kRow], const float (&y)[6 + kLegs * CH],
                                         const float extra) {
  float a0 = r[RowLayout<CH>::kB] + extra, a1 = 0.f;
#pragma unroll
  for (int a = 0; a < 6; a += 2) { a0 = fmaf(-r[a], y[a], a0); a1 = fmaf(-r[a + 1], y[a + 1], a1); }
  if (G > 0) {
#pragma unroll
    for (int k = 0; k < CH; k++) {
      if (k & 1) a1 = fmaf(-r[6 + k], y[6 + (G > 0 ? G - 1 : 0) * CH + k], a1);
      else a0 = fmaf(-r[6 + k], y[6 + (G > 0 ? G - 1 : 0) * CH + k], a0);
    }
  }
  return a0 + a1;
}
// y += u * d for a compact row of group G
template <int CH, int G>
__device__ __forceinline__ void row_axpy(const float (&r)[RowLayout<CH>::kRow], const float d,
                                         float (&y)[6 + kLegs * CH]) {
#pragma unroll
  for (int a = 0; a < 6; a++) y[a] = fmaf(r[a], d, y[a]);
  if (G > 0) {
#pragma unroll
    for (int k = 0; k < CH; k++) y[6 + (G > 0 ? G - 1 : 0) * CH + k] = fmaf(r[6 + k], d, y[6 + (G > 0 ? G - 1 : 0) * CH + k]);
  }
}

// A compact row (6 base + CH own-leg entries) expanded to the dense width with
// selects on the (uniform) group: the selects depend on the row and the group
// only, so they run ahead of the yh chain.
template <int CH>
struct DenseRow {
  static constexpr int W = 6 + kLegs * CH;
  float u[W];
  float b, s, is, x;
  __device__ __forceinline__ void load(const float *p, const int g) {
    typedef RowLayout<CH> RL;
    float r[RL::kRow];
    load_row(p, r);
#pragma unroll
    for (int a = 0; a < 6; a++) u[a] = r[a];
#pragma unroll
    for (int l = 0; l < kLegs; l++)
#pragma unroll
      for (int k = 0; k < CH; k++) u[6 + l * CH + k] = (g == l + 1) ? r[6 + k] : 0.f;
    b = r[RL::kB]; s = r[RL::kS]; is = r[RL::kIs]; x = r[RL::kX];
  }
  __device__ __forceinline__ void expand(const CompactRow<CH> &c) {
    typedef RowLayout<CH> RL;
#pragma unroll
    for (int a = 0; a < 6; a++) u[a] = c.r[a];
#pragma unroll
    for (int l = 0; l < kLegs; l++)
#pragma unroll
      for (int k = 0; k < CH; k++) u[6 + l * CH + k] = (c.g == l + 1) ? c.r[6 + k] : 0.f;
    b = c.r[RL::kB]; s = c.r[RL::kS]; is = c.r[RL::kIs]; x = c.r[RL::kX];
  }
  // (b + extra) - u . y in four accumulator chains
  __device__ __forceinline__ float gap(const float (&y)[W], const float extra = 0.f) const {
    float a0 = b + extra, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int a = 0; a < W; a++) {
      if ((a & 3) == 0) a0 = fmaf(-u[a], y[a], a0);
      else if ((a & 3) == 1) a1 = fmaf(-u[a], y[a], a1);
      else if ((a & 3) == 2) a2 = fmaf(-u[a], y[a], a2);
      else a3 = fmaf(-u[a], y[a], a3);
    }
    return (a0 + a1) + (a2 + a3);
  }
};

// The solve proper.  NR = dense normal rows per lane (1: <= 32 contacts, 2: <= 64).
// Returns iterations; y = yh, mlam = scaled motor impulses, *n_active = contacts
// that carried an impulse.
template <int CH, int NR>
__device__ __forceinline__ int warp_solve(const ConfigT<float> &C, WarpSolveSmem<CH> &sm, const uint32_t motor_mask,
                                          const uint32_t ends03, const int nc, const int lane,
                                          float (&y)[6 + kLegs * CH], float (&mlam)[kLegs * CH], int *n_active) {
  typedef RowLayout<CH> RL;
  typedef CMask<NR> MK;
  typedef typename MK::T mask_t;
  constexpr int W = 6 + kLegs * CH;
  constexpr unsigned FULL = 0xffffffffu;
  const float *crow = &sm.row[RL::kMotorSlots][0];

  // this lane's normal rows, dense, for the whole solve
  float ud[NR][W], ub[NR];
  bool valid[NR];
#pragma unroll
  for (int r = 0; r < NR; r++) {
    const int c = lane + 32 * r;
    valid[r] = c < nc;
    DenseRow<CH> d;
    d.load(crow + (3 * (valid[r] ? c : 0)) * RL::kRow, contact_group(ends03, c));
#pragma unroll
    for (int a = 0; a < W; a++) ud[r][a] = valid[r] ? d.u[a] : 0.f;
    ub[r] = valid[r] ? d.b : -1.f;   // an absent row never comes alive
  }
#pragma unroll
  for (int a = 0; a < W; a++) y[a] = 0.f;
#pragma unroll
  for (int j = 0; j < kLegs * CH; j++) mlam[j] = 0.f;

  const float mlim = C.torque_limit * C.dt;
  mask_t act = 0;   // contacts that have carried a normal impulse
  int it = 0;
  for (; it < C.iters;) {
    float res = 0.f;
    // ---- motor rows (btMultiBodyJointMotor), direction alternates with parity ----
    if (motor_mask != 0) {
      auto motor = [&](const int l, const int i) {
        // dead slots (shorter chains) hold all-zero rows: no test needed
        const int j = l * CH + i;
        float r[RL::kRow];
        load_row(&sm.row[j][0], r);
        float a0 = r[RL::kB], a1 = 0.f;
#pragma unroll
        for (int a = 0; a < 6; a += 2) { a0 = fmaf(-r[a], y[a], a0); a1 = fmaf(-r[a + 1], y[a + 1], a1); }
#pragma unroll
        for (int k = 0; k < CH; k++) {
          if (k & 1) a1 = fmaf(-r[6 + k], y[6 + l * CH + k], a1);
          else a0 = fmaf(-r[6 + k], y[6 + l * CH + k], a0);
        }
        const float lim = mlim * r[RL::kS];
        const float old = mlam[j];
        const float sum = fminf(fmaxf(old + (a0 + a1), -lim), lim);
        const float d = sum - old;
        mlam[j] = sum;
#pragma unroll
        for (int a = 0; a < 6; a++) y[a] = fmaf(r[a], d, y[a]);
#pragma unroll
        for (int k = 0; k < CH; k++) y[6 + l * CH + k] = fmaf(r[6 + k], d, y[6 + l * CH + k]);
        const float rr = d * r[RL::kS];
        res = fmaxf(res, rr * rr);
      };
      if (it & 1) {
#pragma unroll
        for (int jj = 0; jj < kLegs * CH; jj++) motor(jj / CH, jj % CH);
      } else {
#pragma unroll
        for (int jj = kLegs * CH - 1; jj >= 0; jj--) motor(jj / CH, jj % CH);
      }
    }

    if (nc > 0) {
      // ---- normal rows ----
      // Walk `act` in contact order with uniform updates.  Before the update of
      // the next active row every lane tests its own speculative rows that sit
      // between the last processed row and that one against the current yh; the
      // first that has come alive (b' - u.yh > 0 with a zero impulse) is promoted:
      // it is applied now, at its place in the order, and joins `act`.
      // Contacts are sorted by group (base, leg 0..3), so the walk is unrolled over
      // the group: inside a group the leg a compact row touches is static.
      const mask_t all = MK::below(nc);
      int lo = 0;
      auto normal_group = [&](auto gtag) {
        constexpr int G = decltype(gtag)::value;
        const int end = G < kLegs ? (int)((ends03 >> (8 * G)) & 255u) : nc;
        mask_t rem = act & MK::below(end) & ~MK::below(lo);
        for (;;) {
          const int nxt = rem != 0 ? MK::first(rem) : end;
          if (G < kLegs && nxt >= end) break;   // the pending range is screened by the next group
          const mask_t range = MK::below(nxt) & ~MK::below(lo) & ~act & all;   // speculative rows in [lo, nxt)
          if (range != 0) {
            mask_t viol = 0;
            float t[NR];
#pragma unroll
            for (int q = 0; q < NR; q++) {
              float a0 = ub[q], a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
              for (int a = 0; a < W; a++) {
                if ((a & 3) == 0) a0 = fmaf(-ud[q][a], y[a], a0);
                else if ((a & 3) == 1) a1 = fmaf(-ud[q][a], y[a], a1);
                else if ((a & 3) == 2) a2 = fmaf(-ud[q][a], y[a], a2);
                else a3 = fmaf(-ud[q][a], y[a], a3);
              }
              t[q] = (a0 + a1) + (a2 + a3);
              viol |= (mask_t)__ballot_sync(FULL, t[q] > 0.f) << (32 * q);
            }
            viol &= range;
            if (viol != 0) {
              // rare (at most once per contact and solve): dense update, any group
              const int v = MK::first(viol);
              float dv = __shfl_sync(FULL, t[0], v & 31);
              if (NR > 1) {
                const float d1 = __shfl_sync(FULL, t[NR - 1], v & 31);
                dv = v < 32 ? dv : d1;
              }
              DenseRow<CH> p;
              p.load(crow + (3 * v) * RL::kRow, contact_group(ends03, v));
#pragma unroll
              for (int a = 0; a < W; a++) y[a] = fmaf(p.u[a], dv, y[a]);
              sm.lam[v].x = dv;
              act |= MK::bit(v);
              const float rr = dv * p.s;
              res = fmaxf(res, rr * rr);
              lo = v + 1;
              continue;
            }
          }
          if (nxt >= end) break;   // G == kLegs: everything up to nc has been screened
          float r[RL::kRow];
          load_row(crow + (3 * nxt) * RL::kRow, r);
          const float lam_n = sm.lam[nxt].x;
          const float sum = fmaxf(lam_n + row_gap<CH, G>(r, y, 0.f), 0.f);
          const float d = sum - lam_n;
          sm.lam[nxt].x = sum;
          row_axpy<CH, G>(r, d, y);
          const float rr = d * r[RL::kS];
          res = fmaxf(res, rr * rr);
          rem &= rem - 1;
          lo = nxt + 1;
        }
      };
      normal_group(IntC<0>());
      normal_group(IntC<1>());
      normal_group(IntC<2>());
      normal_group(IntC<3>());
      normal_group(IntC<4>());

      // ---- friction pairs of the contacts that can hold friction ----
      // both rows on the same velocity, radial projection onto the disc of radius
      // mu * lambda_n (resolveConeFrictionConstraintRows)
      auto friction_group = [&](auto gtag) {
        constexpr int G = decltype(gtag)::value;
        const int begin = G > 0 ? (int)((ends03 >> (8 * (G - 1))) & 255u) : 0;
        const int end = G < kLegs ? (int)((ends03 >> (8 * G)) & 255u) : nc;
        for (mask_t fr = act & MK::below(end) & ~MK::below(begin); fr != 0; fr &= fr - 1) {
          const int c = MK::first(fr);
          float fa[RL::kRow], fb[RL::kRow];
          load_row(crow + (3 * c + 1) * RL::kRow, fa);
          load_row(crow + (3 * c + 2) * RL::kRow, fb);
          const float4 l4 = sm.lam[c];
          float sa = row_gap<CH, G>(fa, y, l4.y), sb = row_gap<CH, G>(fb, y, l4.z);
          const float limit = fa[RL::kX] * l4.x;                   // mu * lambda_n, true units
          const float ta = sa * fa[RL::kIs], tb = sb * fb[RL::kIs];  // true units
          const float n2 = fmaf(ta, ta, fmaf(tb, tb, 1e-30f));
          const float sc = n2 > limit * limit ? limit * fast_rsqrt(n2) : 1.f;
          sa *= sc; sb *= sc;
          const float da = sa - l4.y, db = sb - l4.z;
          sm.lam[c].y = sa; sm.lam[c].z = sb;
          row_axpy<CH, G>(fa, da, y);
          row_axpy<CH, G>(fb, db, y);
          const float r1 = da * fa[RL::kS], r2 = db * fb[RL::kS];
          res = fmaxf(res, fmaxf(r1 * r1, r2 * r2));
        }
      };
      if (act != 0) {
        friction_group(IntC<0>());
        friction_group(IntC<1>());
        friction_group(IntC<2>());
        friction_group(IntC<3>());
        friction_group(IntC<4>());
      }
    }
    it++;
    if (res <= C.res_thr) break;
  }
  *n_active = MK::count(act);
  return it;
}

template <int CH, bool DBG>
__global__ void __launch_bounds__(32)
solve_warp_kernel(const __grid_constant__ ConfigT<float> C, const Buffers B, const uint32_t motor_mask,
                  const int ndof_total, const int n_cand) {
  typedef RowLayout<CH> RL;
  constexpr int W = 6 + kLegs * CH;   // dense width
  __shared__ __align__(16) WarpSolveSmem<CH> sm;
  const int lane = threadIdx.x;
  const int64_t env = blockIdx.x;
  const long long t_begin = clock64();

  const uint32_t ends03 = B.hdr[env];
  const int nc = (int)B.hdr[(int64_t)B.n + env];

  // ---- stage: the env's row list -> smem, impulses = 0 ----
  {
    const float4 *src = reinterpret_cast<const float4 *>(B.rows + env * (RL::kSlots * RL::kRow));
    float4 *dst = reinterpret_cast<float4 *>(&sm.row[0][0]);
    const int n4 = (RL::kMotorSlots + 3 * nc) * (RL::kRow / 4);
    for (int i = lane; i < n4; i += 32) dst[i] = __ldcg(src + i);
    for (int i = lane; i < nc; i += 32) sm.lam[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncwarp();

  float y[W], mlam[kLegs * CH];
  int it = 0, n_active = 0;
  if (motor_mask != 0 || nc > 0) {
    if (nc <= 32) it = warp_solve<CH, 1>(C, sm, motor_mask, ends03, nc, lane, y, mlam, &n_active);
    else it = warp_solve<CH, 2>(C, sm, motor_mask, ends03, nc, lane, y, mlam, &n_active);
  } else {
#pragma unroll
    for (int a = 0; a < W; a++) y[a] = 0.f;
#pragma unroll
    for (int j = 0; j < kLegs * CH; j++) mlam[j] = 0.f;
  }

  // ---- outputs ----
  {
    // yh is warp-uniform: lane a writes entry a
    float v = 0.f;
#pragma unroll
    for (int a = 0; a < W; a++)
      if (lane == a) v = y[a];
    if (lane < W) B.yh[(int64_t)lane * B.n + env] = v;
  }
  if (lane == 0) {
    // per-env solver statistics: iterations, contacts, contacts that carried an
    // impulse, SM cycles spent in the solve
    float *st = B.yh + (int64_t)W * B.n + env;
    st[0] = (float)it; st[(int64_t)B.n] = (float)nc; st[2 * (int64_t)B.n] = (float)n_active;
    st[3 * (int64_t)B.n] = (float)(clock64() - t_begin);
  }
  if (DBG) {
    __syncwarp();
    if (B.dbg_normal_force) {
      for (int c = lane; c < n_cand; c += 32) B.dbg_normal_force[(int64_t)c * B.n + env] = 0.f;
      __syncwarp();
      for (int i = lane; i < nc; i += 32) {
        const float *nr = &sm.row[RL::kMotorSlots + 3 * i][0];
        B.dbg_normal_force[(int64_t)(int)nr[RL::kX] * B.n + env] = sm.lam[i].x * nr[RL::kIs] * C.inv_dt;
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
          if (d < ndof_total) B.dbg_motor_torque[(int64_t)d * B.n + env] = mlam[j] * sm.row[j][RL::kIs] * C.inv_dt;
          d++;
        }
      }
    }
  }
}

}  // namespace solo
#endif  // SOLO_SOLVE_WARP_CUH_
