// solo_solve_warp.cuh -- contact detection, contact rows and the constraint solver
// of stepSimulation with ONE WARP PER ENVIRONMENT.  Same iterates as solo_solve.h
// (the sequential statement of the algorithm, which tests/hostsim runs on the
// CPU), scheduled for LATENCY: a launch lasts as long as its slowest environment
// (50 iterations x ~10 contacts that carry an impulse), and a warp running alone
// issues one instruction every ~2 cycles, so what counts is the instruction count
// and the dependent chain of one Gauss-Seidel update.
//
//   * prologue: the environment's KinLayout record (written by setup_kernel) is
//     copied to shared memory; the lanes test the model's support points against
//     the ground plane, compact the hits by ballot (Bullet's order: by body, then
//     by support point) and lane l builds the normal row of contact l (and l + 32)
//     as a dense vector in registers -- no row ever goes through memory;
//   * yh, all impulses and every piece of control flow are warp-uniform;
//   * ~90 % of the rows inside Bullet's 2 cm breaking threshold are speculative
//     (zero impulse, b' - u.yh <= 0: a no-op), so only the contacts that HAVE
//     carried an impulse are updated, uniformly, while the lanes check in parallel
//     that no speculative row between two of them has come alive under the
//     current yh; one that has is promoted on the spot, at its place in Bullet's
//     order -- exactly sequential Gauss-Seidel;
//   * the promoted ("active") contacts live in shared memory as an array of dense
//     records sorted by contact index: normal row, the friction pair, the three
//     impulses, and the mask of speculative rows that precede the contact.  The
//     walk over it is a counted loop: addresses are k * stride, nothing on the
//     chain but the arithmetic; the normal rows go through two register buffers,
//     entry k + 1 loaded while entry k is on the chain.  A promotion (at most once
//     per contact and solve) shifts the tail of the array; the owner lane builds
//     the friction pair then;
//   * friction: both rows on the same velocity, cone projection without a branch.
#ifndef SOLO_SOLVE_WARP_CUH_
#define SOLO_SOLVE_WARP_CUH_

#include "solo_contact.cuh"
#include "solo_kernels.cuh"

// -DSOLO_CHECKS (tools/sanitizer_run.py; compute-sanitizer is closed on this pool): every index into
// the warp's shared arrays and the candidate table is checked; a violation prints and traps.
#ifdef SOLO_CHECKS
#include <stdio.h>
#define SOLO_CHECK(cond)                                                                     \
  do {                                                                                       \
    if (!(cond)) {                                                                           \
      printf("SOLO_CHECK failed: %s (%s:%d) block %d lane %d\n", #cond, __FILE__, __LINE__, \
             (int)blockIdx.x, (int)threadIdx.x);                                             \
      __trap();                                                                              \
    }                                                                                        \
  } while (0)
#else
#define SOLO_CHECK(cond) do { } while (0)
#endif

namespace solo {

// one active contact in shared memory (floats)
template <int CH>
struct ActiveLayout {
  static constexpr int W = 6 + kLegs * CH;           // dense width
  static constexpr int NW = (W + 2 + 3) & ~3;        // normal: u[W], b, s
  static constexpr int FW = (W + 4 + 3) & ~3;        // friction: u[W], b, s, 1/s, mu/s_n
  static constexpr int kFA = NW, kFB = NW + FW;
  static constexpr int kMeta = NW + 2 * FW;          // lam_n, lam_a, lam_b, idx | range lo, range hi, cand, -
  static constexpr int kSize = kMeta + 8;            // 64 floats (CH 2), 76 (CH 3)
};

// TAB: the block keeps its own copy of the support-point table (persistent build: filled once per
// warp, read by every environment it walks over; the L1 is all but gone to the shared-memory
// carve-out, so the global table costs an L2 round trip twice per environment)
// STASH: every listed contact's support point is kept beside its index, so that a promotion
// (on the dependent chain) does not go back to the global table.
template <int CH, bool TAB = false, bool STASH = false>
struct WarpSolveSmem {
  float kin[KinLayout<CH>::kSize];                   // setup_kernel's record (motor rows first)
  float ent[kMaxContacts][ActiveLayout<CH>::kSize];  // active contacts, sorted by contact index
  int32_t list[kMaxContacts];                        // contact -> support point
  float llam[kLegs * CH];                            // scaled joint-limit impulses (rows are rare)
  float4 ctab[TAB ? kMaxCand : 1];                   // support point xyz, friction
  uint8_t cbody[TAB ? kMaxCand : 4];                 // (leg + 1) << 4 | depth in the chain
  float4 pstash[STASH ? kMaxContacts : 1];           // contact -> support point xyz, friction
  uint8_t pbody[STASH ? kMaxContacts : 4];
#if defined(SOLO_SECTION_PROBE) || defined(SOLO_PROLOGUE_PROBE)
  long long probe[4];                                // probe build: cycles in normal rows / friction pairs, screens
#endif
};

__device__ __forceinline__ Cand unpack_cand(const float4 a, const int b) {
  Cand k;
  k.x = a.x; k.y = a.y; k.z = a.z; k.mu = a.w;
  k.leg = (b >> 4) - 1; k.depth = b & 15;
  return k;
}
__device__ __forceinline__ uint8_t pack_body(const Cand &k) { return (uint8_t)(((k.leg + 1) << 4) | (k.depth & 15)); }

// support point of listed contact v, from the nearest place that has it
template <int CH, bool TAB, bool STASH>
__device__ __forceinline__ Cand cand_of_contact(const WarpSolveSmem<CH, TAB, STASH> &sm, const float *tab, const int v) {
  SOLO_CHECK(v >= 0 && v < kMaxContacts && sm.list[v] >= 0 && sm.list[v] < kMaxCand);
  if (STASH) return unpack_cand(sm.pstash[v], sm.pbody[v]);
  if (TAB) return unpack_cand(sm.ctab[sm.list[v]], sm.cbody[sm.list[v]]);
  return load_cand(tab, sm.list[v]);
}

template <int CH, bool TAB, bool STASH>
__device__ __forceinline__ Cand cand_at(const WarpSolveSmem<CH, TAB, STASH> &sm, const float *tab, const int c) {
  SOLO_CHECK(c >= 0 && c < kMaxCand);
  if (!TAB) return load_cand(tab, c);
  return unpack_cand(sm.ctab[c], sm.cbody[c]);
}

// contact bitmasks: 32 bits when the env has <= 32 contacts (99.8 % of env-steps), else 64
template <int NR> struct CMask;
template <> struct CMask<1> {
  typedef uint32_t T;
  static __device__ __forceinline__ int first(T m) { return __ffs((int)m) - 1; }
  static __device__ __forceinline__ int last(T m) { return 31 - __clz((int)m); }
  static __device__ __forceinline__ T bit(int c) { return 1u << c; }
  static __device__ __forceinline__ T below(int c) { return c >= 32 ? ~0u : ((1u << c) - 1u); }
  static __device__ __forceinline__ T make(uint32_t lo, uint32_t) { return lo; }
};
template <> struct CMask<2> {
  typedef uint64_t T;
  static __device__ __forceinline__ int first(T m) {
    const uint32_t lo = (uint32_t)m;
    return lo ? __ffs((int)lo) - 1 : 31 + __ffs((int)(uint32_t)(m >> 32));
  }
  static __device__ __forceinline__ int last(T m) { return 63 - __clzll((long long)m); }
  static __device__ __forceinline__ T bit(int c) { return 1ull << c; }
  static __device__ __forceinline__ T below(int c) { return c >= 64 ? ~0ull : ((1ull << c) - 1ull); }
  static __device__ __forceinline__ T make(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; }
};

__device__ __forceinline__ float fast_rsqrt(const float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// compact row of a contact on leg `leg` (-1: base) -> dense u[W] followed by n_tail scalars
template <int CH>
__device__ __forceinline__ void store_dense(float *dst, const float (&row)[RowLayout<CH>::kRow], const int leg,
                                            const int n_tail, const int width) {
  typedef RowLayout<CH> RL;
  constexpr int W = 6 + kLegs * CH;
#pragma unroll
  for (int a = 0; a < 6; a++) dst[a] = row[a];
#pragma unroll
  for (int l = 0; l < kLegs; l++)
#pragma unroll
    for (int k = 0; k < CH; k++) dst[6 + l * CH + k] = (leg == l) ? row[6 + k] : 0.f;
#pragma unroll
  for (int t = 0; t < 4; t++)
    if (W + t < width) dst[W + t] = t < n_tail ? row[RL::kJ + t] : 0.f;
  for (int a = W + 4; a < width; a++) dst[a] = 0.f;
}

// b - u . y over a dense record (u at r[0..W), b at r[W]) in four accumulator chains
template <int W, int N>
__device__ __forceinline__ float dense_gap(const float (&r)[N], const float (&y)[W], const float extra) {
  float a0 = r[W] + extra, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
  for (int a = 0; a < W; a++) {
    if ((a & 3) == 0) a0 = fmaf(-r[a], y[a], a0);
    else if ((a & 3) == 1) a1 = fmaf(-r[a], y[a], a1);
    else if ((a & 3) == 2) a2 = fmaf(-r[a], y[a], a2);
    else a3 = fmaf(-r[a], y[a], a3);
  }
  return (a0 + a1) + (a2 + a3);
}
template <int W, int N>
__device__ __forceinline__ void dense_axpy(float (&y)[W], const float (&r)[N], const float d) {
#pragma unroll
  for (int a = 0; a < W; a++) y[a] = fmaf(r[a], d, y[a]);
}

// The three rows of a promoted contact, expanded into entry `e` by lanes 0, 1, 2 side by side (one
// instruction stream, the direction is the lane: a third of the warp instructions of one lane building
// all three; the promotion sits on the warp's dependent chain).  Called by the whole warp.
template <int CH>
__device__ __forceinline__ void build_entry_inline(float *e, const float *kin, const ConfigT<float> &C, const float margin,
                                         Cand k, const int cand, const int contact, const float lam_n,
                                         const float link_mu, const int lane) {
  typedef RowLayout<CH> RL;
  typedef ActiveLayout<CH> AL;
  if (link_mu >= 0.f && k.leg >= 0) k.mu = link_mu;   // per-env friction of the non-base links
  const int dir = lane < 2 ? lane : 2;
  float row[RL::kRow];
  build_contact_row_dir<CH>(kin, C, margin, k, cand, dir, row);
  const float is_n = __shfl_sync(0xffffffffu, row[RL::kIs], 0);
  if (dir == 1) row[RL::kX] = k.mu * is_n;
  if (lane < 3) {
    if (dir == 0) store_dense<CH>(e, row, k.leg, 2, AL::NW);
    else store_dense<CH>(e + (dir == 1 ? AL::kFA : AL::kFB), row, k.leg, 4, AL::FW);
  }
  if (lane == 0) {
    e[AL::kMeta] = lam_n; e[AL::kMeta + 1] = 0.f; e[AL::kMeta + 2] = 0.f; e[AL::kMeta + 3] = __int_as_float(contact);
    e[AL::kMeta + 4] = 0.f; e[AL::kMeta + 5] = 0.f; e[AL::kMeta + 6] = __int_as_float(cand); e[AL::kMeta + 7] = 0.f;
  }
}

// Out of line for the one-block-per-environment build (a shorter hot loop is worth more there than the
// call: 4096 environments 0.1321 against 0.1332 ms); the persistent CH = 2 build inlines it (65536: 0.643 -> 0.629).
template <int CH>
__device__ __noinline__ void build_entry(float *e, const float *kin, const ConfigT<float> &C, const float margin,
                                         const Cand k, const int cand, const int contact, const float lam_n,
                                         const float link_mu, const int lane) {
  build_entry_inline<CH>(e, kin, C, margin, k, cand, contact, lam_n, link_mu, lane);
}

// The solve proper.  NR = dense normal rows per lane (1: <= 32 contacts, 2: <= 64).
// Returns iterations; y = yh, mlam = scaled motor impulses, *n_active = entries of sm.ent.
template <int CH, int NR, bool TAB, bool STASH>
__device__ __forceinline__ int warp_solve(const ConfigT<float> &C, WarpSolveSmem<CH, TAB, STASH> &sm, const float margin,
                                          const float *cand_tab, const float link_mu, const uint32_t motor_mask,
                                          const uint32_t lim, const int nc, const int lane, float (&y)[6 + kLegs * CH],
                                          float (&mlam)[kLegs * CH], int *n_active) {
  typedef RowLayout<CH> RL;
  typedef KinLayout<CH> KL;
  typedef ActiveLayout<CH> AL;
  typedef CMask<NR> MK;
  typedef typename MK::T mask_t;
  constexpr int W = AL::W;
  constexpr unsigned FULL = 0xffffffffu;

  // this lane's normal rows, dense, for the whole solve
  float ud[NR][W], ub[NR];
#pragma unroll
  for (int r = 0; r < NR; r++) {
    const int c = lane + 32 * r;
    const bool valid = c < nc;
    const int cand = valid ? sm.list[c] : 0;
    SOLO_CHECK(cand >= 0 && cand < kMaxCand && nc <= kMaxContacts && nc <= 32 * NR);
    const Cand k = cand_at(sm, cand_tab, cand);
    if (STASH && valid) { sm.pstash[c] = make_float4(k.x, k.y, k.z, k.mu); sm.pbody[c] = pack_body(k); }
    float row[1][RL::kRow];
    build_contact_rows<CH, 1>(sm.kin, C, margin, k, cand, row);
#pragma unroll
    for (int a = 0; a < 6; a++) ud[r][a] = valid ? row[0][a] : 0.f;
#pragma unroll
    for (int l = 0; l < kLegs; l++)
#pragma unroll
      for (int j = 0; j < CH; j++) ud[r][6 + l * CH + j] = (valid && k.leg == l) ? row[0][6 + j] : 0.f;
    ub[r] = valid ? row[0][RL::kB] : -1.f;   // an absent row never comes alive
  }
#pragma unroll
  for (int a = 0; a < W; a++) y[a] = 0.f;
#pragma unroll
  for (int j = 0; j < kLegs * CH; j++) mlam[j] = 0.f;

  const float mlim = C.torque_limit * C.dt;
  const mask_t all = MK::below(nc);
  mask_t act = 0;          // contacts that have carried a normal impulse (= the entries of sm.ent)
  mask_t tail_range = all; // speculative rows behind the last active contact
  // Screening budget.  Rows are unit vectors, so an update of size d moves any gap
  // b' - u.yh by at most |d|.  After an exact screen the smallest distance of a
  // speculative row from coming alive is known; until the |d| of the updates since
  // then add up to it, no speculative row can be alive and the screens are skipped
  // (exactly: a skipped screen would have found nothing).
  float budget = 0.f;
  int n_act = 0;
  int it = 0;
#ifdef SOLO_SECTION_PROBE
  long long pr_n = 0, pr_f = 0, pr_s = 0, pr_t = 0, pr_m = 0, pr_nu = 0, pr_fu = 0;
#define PROBE_T() do { pr_t = clock64(); } while (0)
#define PROBE_ADD(x) do { const long long _n = clock64(); x += _n - pr_t; pr_t = _n; } while (0)
#else
#define PROBE_T() do { } while (0)
#define PROBE_ADD(x) do { } while (0)
#endif
#ifdef SOLO_IDLE_PROBE
  int probe_visits = 0, probe_idle = 0;   // probe build: how many promoted-contact visits found the contact idle
#endif
  // Joint-limit rows (btMultiBodyJointLimitConstraint; a joint at or beyond a URDF limit).  They
  // sit in front of the motor rows in Bullet's non-contact list, whose traversal alternates
  // with iteration parity.  u = dir * u_motor of the slot (expanded to the dense width with
  // selects: yh stays in registers), b' in the slot's kX, 0 <= lambda' <= lim_max * s.  Rare,
  // so one rolled loop: nothing of it is on the path of an environment without limit rows.
  auto limit_rows = [&](const bool forward, float &res) {
#pragma unroll 1
    for (int jj = 0; jj < kLegs * CH; jj++) {
      const int j = forward ? jj : kLegs * CH - 1 - jj;
      if (((lim >> j) & 1u) == 0) continue;
      const int l = j / CH;
      float r[RL::kRow];
      load_row(sm.kin + KL::kMotor + j * RL::kRow, r);
      const float dir = ((lim >> (16 + j)) & 1u) ? -1.f : 1.f;
      float dotv = 0.f;
#pragma unroll
      for (int a = 0; a < 6; a++) dotv = fmaf(r[a], y[a], dotv);
#pragma unroll
      for (int l2 = 0; l2 < kLegs; l2++)
#pragma unroll
        for (int k = 0; k < CH; k++) dotv = fmaf(l2 == l ? r[6 + k] : 0.f, y[6 + l2 * CH + k], dotv);
      const float old = sm.llam[j];
      const float sum = fminf(fmaxf(old + (r[RL::kX] - dir * dotv), 0.f), C.lim_max * r[RL::kS]);
      const float d = (sum - old) * dir;
      __syncwarp();
      if (lane == 0) sm.llam[j] = sum;
      __syncwarp();
#pragma unroll
      for (int a = 0; a < 6; a++) y[a] = fmaf(r[a], d, y[a]);
#pragma unroll
      for (int l2 = 0; l2 < kLegs; l2++)
#pragma unroll
        for (int k = 0; k < CH; k++) y[6 + l2 * CH + k] = fmaf(l2 == l ? r[6 + k] : 0.f, d, y[6 + l2 * CH + k]);
      const float rr = d * r[RL::kS];
      res = fmaxf(res, rr * rr);
      budget -= fabsf(d);
    }
  };
#ifdef SOLO_PROLOGUE_PROBE
  if (lane == 0) sm.probe[2] = clock64();   // rows built: the iterations start
#endif
  for (; it < C.iters;) {
    float res = 0.f;
    PROBE_T();
    if (lim != 0 && (it & 1)) limit_rows(true, res);
    // ---- motor rows (btMultiBodyJointMotor), direction alternates with parity ----
    if (motor_mask != 0) {
      auto motor = [&](const int l, const int i) {
        // dead slots (shorter chains) hold all-zero rows: no test needed
        const int j = l * CH + i;
        float r[RL::kRow];
        load_row(sm.kin + KL::kMotor + j * RL::kRow, r);
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
        budget -= fabsf(d);
      };
      if (it & 1) {
#pragma unroll
        for (int jj = 0; jj < kLegs * CH; jj++) motor(jj / CH, jj % CH);
      } else {
#pragma unroll
        for (int jj = kLegs * CH - 1; jj >= 0; jj--) motor(jj / CH, jj % CH);
      }
    }
    if (lim != 0 && !(it & 1)) limit_rows(false, res);
    PROBE_ADD(pr_m);

    if (nc > 0) {
      // ---- normal rows ----
      // Entry k of sm.ent is the k-th active contact; its `range` is the set of speculative
      // rows between the previous active contact and itself: they are screened against the
      // current yh before entry k is updated, and the first that has come alive is promoted
      // into slot k and applied.  The walk is software-pipelined over two register buffers:
      // entry k + 1 is loaded while entry k is on the dependent chain, and the entry's own gap
      // is formed before the (rare) screen is decided -- a screen without a promotion leaves
      // yh alone, one with a promotion reloads.
      auto load_n = [&](const int k, float (&n)[AL::NW], float &lam, mask_t &range) {
        const float *e = &sm.ent[k < kMaxContacts ? k : kMaxContacts - 1][0];
        load_row(e, n);
        lam = e[AL::kMeta];
        range = MK::make(__float_as_uint(e[AL::kMeta + 4]), NR > 1 ? __float_as_uint(e[AL::kMeta + 5]) : 0u);
      };
      float t[NR];
      PROBE_T();
      auto screen = [&](const mask_t range) -> mask_t {
#ifdef SOLO_SECTION_PROBE
        pr_s++;
#endif
        mask_t viol = 0;
        float slack = 3.0e38f;
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
          // distance from coming alive, speculative rows only (absent rows sit at 1)
          if (((act >> (lane + 32 * q)) & 1) == 0) slack = fminf(slack, fmaxf(-t[q], 0.f));
          viol |= (mask_t)__ballot_sync(FULL, t[q] > 0.f) << (32 * q);
        }
        budget = 0.999f * __uint_as_float(__reduce_min_sync(FULL, __float_as_uint(slack)));
        return viol & range;
      };
      auto update_n = [&](const int k, const float (&n)[AL::NW], const float lam, const float gap) {
        SOLO_CHECK(k >= 0 && k < n_act);
#ifdef SOLO_SECTION_PROBE
        pr_nu++;
#endif
        const float sum = fmaxf(gap, 0.f);
        const float d = sum - lam;
        sm.ent[k][AL::kMeta] = sum;
        dense_axpy<W>(y, n, d);
        const float rr = d * n[W + 1];
        res = fmaxf(res, rr * rr);
        budget -= fabsf(d);
      };
      for (int k = 0;;) {
        mask_t viol = 0;
        if (k < n_act) {
          float nA[AL::NW], nB[AL::NW], lamA, lamB;
          mask_t rA, rB;
          load_n(k, nA, lamA, rA);
          for (;;) {
            // the entry's own gap in front of the decision (which hangs off the previous update's
            // |d| through the budget): a screen without a promotion leaves yh alone
            float g = dense_gap<W>(nA, y, lamA);
            if (rA != 0 && budget <= 0.f) {
              viol = screen(rA);
              if (viol != 0) break;
            }
            load_n(k + 1, nB, lamB, rB);   // in the update's block: the loads fill its dependency stalls
            update_n(k, nA, lamA, g);
            k++;
            if (k >= n_act) break;
            g = dense_gap<W>(nB, y, lamB);
            if (rB != 0 && budget <= 0.f) {
              viol = screen(rB);
              if (viol != 0) break;
            }
            load_n(k + 1, nA, lamA, rA);
            update_n(k, nB, lamB, g);
            k++;
            if (k >= n_act) break;
          }
        }
        if (viol == 0) {
          // behind the last active contact
          if (tail_range != 0 && budget <= 0.f) viol = screen(tail_range);
          if (viol == 0) break;
        }
        // ---- promotion of contact v into slot k (rare) ----
        const int v = MK::first(viol);
        SOLO_CHECK(v >= 0 && v < nc && ((act >> v) & 1) == 0 && n_act < nc && n_act < kMaxContacts && k <= n_act);
        float dv = __shfl_sync(FULL, t[0], v & 31);
        if (NR > 1) {
          const float d1 = __shfl_sync(FULL, t[NR - 1], v & 31);
          dv = v < 32 ? dv : d1;
        }
        // shift entries [k, n_act) up by one; lane j moves the same words of every entry
        for (int s = n_act - 1; s >= k; s--)
          for (int j = lane; j < AL::kSize; j += 32) sm.ent[s + 1][j] = sm.ent[s][j];
        __syncwarp();
        // lanes 0-2 build the contact's three rows (the friction pair for the first time)
        if (TAB && CH == 2) build_entry_inline<CH>(&sm.ent[k][0], sm.kin, C, margin, cand_of_contact(sm, cand_tab, v), sm.list[v], v, dv, link_mu, lane);
        else build_entry<CH>(&sm.ent[k][0], sm.kin, C, margin, cand_of_contact(sm, cand_tab, v), sm.list[v], v, dv, link_mu, lane);
        act |= MK::bit(v);
        n_act++;
        __syncwarp();
        // ranges of every entry and of the tail
        for (int s = lane; s < n_act; s += 32) {
          const int idx = __float_as_int(sm.ent[s][AL::kMeta + 3]);
          const int prev = s > 0 ? __float_as_int(sm.ent[s - 1][AL::kMeta + 3]) : -1;
          const mask_t m = MK::below(idx) & ~MK::below(prev + 1);
          sm.ent[s][AL::kMeta + 4] = __uint_as_float((uint32_t)m);
          sm.ent[s][AL::kMeta + 5] = __uint_as_float((uint32_t)((uint64_t)m >> 32));
        }
        tail_range = all & ~MK::below(MK::last(act) + 1);
        __syncwarp();
        float pr[AL::NW];
        load_row(&sm.ent[k][0], pr);
#pragma unroll
        for (int a = 0; a < W; a++) y[a] = fmaf(pr[a], dv, y[a]);
        const float rr = dv * pr[W + 1];
        res = fmaxf(res, rr * rr);
        budget = 0.f;   // a row elsewhere may be alive too: screen again
        k++;
      }

      PROBE_ADD(pr_n);
      // ---- friction pairs of the contacts that can hold friction ----
      // both rows on the same velocity, radial projection onto the disc of radius
      // mu * lambda_n (resolveConeFrictionConstraintRows)
      for (int k = 0; k < n_act; k++) {
        float *e = &sm.ent[k][0];
        float fa[AL::FW], fb[AL::FW];
        load_row(e + AL::kFA, fa);
        load_row(e + AL::kFB, fb);
        const float4 l4 = *reinterpret_cast<const float4 *>(e + AL::kMeta);
        // no normal impulse and no friction impulse: the disc has radius 0, the update is exactly a no-op
#ifdef SOLO_IDLE_PROBE
        probe_visits++;
        if (l4.x == 0.f && l4.y == 0.f && l4.z == 0.f) probe_idle++;
#endif
        if (l4.x == 0.f && l4.y == 0.f && l4.z == 0.f) continue;
#ifdef SOLO_SECTION_PROBE
        pr_fu++;
#endif
        float sa = dense_gap<W>(fa, y, l4.y), sb = dense_gap<W>(fb, y, l4.z);
        const float limit = fa[W + 3] * l4.x;                  // mu * lambda_n, true units
        const float ta = sa * fa[W + 2], tb = sb * fb[W + 2];  // true units
        const float n2 = fmaf(ta, ta, fmaf(tb, tb, 1e-30f));
        const float sc = n2 > limit * limit ? limit * fast_rsqrt(n2) : 1.f;
        sa *= sc; sb *= sc;
        const float da = sa - l4.y, db = sb - l4.z;
        e[AL::kMeta + 1] = sa; e[AL::kMeta + 2] = sb;
        dense_axpy<W>(y, fa, da);
        dense_axpy<W>(y, fb, db);
        const float r1 = da * fa[W + 1], r2 = db * fb[W + 1];
        res = fmaxf(res, fmaxf(r1 * r1, r2 * r2));
        budget -= fabsf(da) + fabsf(db);
      }
      PROBE_ADD(pr_f);
    }
    it++;
    if (res <= C.res_thr) break;
  }
  *n_active = n_act;
#ifdef SOLO_SECTION_PROBE
  if (lane == 0) { sm.probe[0] = (pr_n / 100) * 1000 + min(999ll, pr_nu); sm.probe[1] = (pr_f / 100) * 1000 + min(999ll, pr_fu); sm.probe[2] = pr_s; sm.probe[3] = pr_m; }
#endif
#ifdef SOLO_IDLE_PROBE
  *n_active = n_act + 1000 * (probe_visits > 0 ? (100 * probe_idle) / probe_visits : 0);
#endif
  return it;
}

// MINB: minimum resident blocks per SM the register allocation is held to (0: none; the CH = 3
// instantiation is held to 10 = 168 registers).  With the three rows of a promoted contact built by
// three lanes side by side the CH = 2 solve takes 127 registers, so shared memory (16.9 KB a block)
// is what limits residency: 13 one-warp blocks per SM.
// PERSIST: the grid is ONE resident wave of warps that walk over the environments (below): for
// batches of many waves, where the launch is throughput bound.
template <int CH, bool DBG, int MINB, bool PERSIST>
__global__ void __launch_bounds__(32, MINB)
solve_warp_kernel(const __grid_constant__ ConfigT<float> C, const Buffers B, const uint32_t motor_mask,
                  const int ndof_total, const int n_cand, const float margin) {
  typedef RowLayout<CH> RL;
  typedef KinLayout<CH> KL;
  typedef ActiveLayout<CH> AL;
  constexpr int W = 6 + kLegs * CH;   // dense width
  constexpr unsigned FULL = 0xffffffffu;
  constexpr bool STASH = !PERSIST;
  __shared__ __align__(16) WarpSolveSmem<CH, PERSIST, STASH> sm;
  const int lane = threadIdx.x;
  // PERSIST: the grid is one wave of warps, each of which walks over environments: its own block
  // index, that plus the grid size, then whatever the launch's counter hands out (B.work[0]; drawn
  // one environment ahead, so the atomic's latency is off the path).  While an environment is
  // solved the NEXT one's record is copied into the unused tail of the active-contact array with
  // cp.async: a throughput-bound launch then pays neither a block launch nor an exposed memory
  // latency per environment.  The order changes nothing in the results: environments are
  // independent.
  constexpr int kStageEntries = (KL::kSize + AL::kSize - 1) / AL::kSize;
  constexpr int kStageAt = kMaxContacts - kStageEntries;   // first entry of the staging area
  static_assert(KL::kSize % 4 == 0 && (AL::kSize * 4) % 16 == 0, "records are copied 16 bytes at a time");
  int64_t env = blockIdx.x;
  int64_t nxt = PERSIST ? (int64_t)blockIdx.x + gridDim.x : (int64_t)B.n;
  bool staged = false;
#ifdef SOLO_TIMELINE
  unsigned tl_begin;
  asm volatile("mov.u32 %0, %%globaltimer_lo;" : "=r"(tl_begin));
#endif

  if (PERSIST) {
    // the warp's copy of the support-point table (not written by setup_legpar_kernel: in front of the wait on it)
    for (int c = lane; c < kMaxCand; c += 32) {
      const Cand k = load_cand(B.cand_tab, c);
      sm.ctab[c] = make_float4(k.x, k.y, k.z, k.mu);
      sm.cbody[c] = pack_body(k);
    }
    __syncwarp();
  }
  // programmatic dependent of setup_legpar_kernel: wait for its records, then let the finish
  // kernel's blocks be scheduled behind the last solver block
  cudaGridDependencySynchronize();
  cudaTriggerProgrammaticLaunchCompletion();

  // start the copy of the next environment's record into the staging area (if there is one and
  // the area is free)
  auto prefetch_next = [&](const bool stage_free) {
    staged = stage_free && nxt < B.n;
    SOLO_CHECK(nxt >= 0 && kStageAt >= 0 && (kMaxContacts - kStageAt) * AL::kSize >= KL::kSize);
    if (staged) {
      const float4 *src = reinterpret_cast<const float4 *>(B.kin + nxt * KL::kSize);
      const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&sm.ent[kStageAt][0]);
      for (int i = lane; i < KL::kSize / 4; i += 32)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * i), "l"(src + i) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
  };
  for (; env < B.n;) {
    const long long t_begin = clock64();

    // ---- the environment's record -> smem ----
#ifdef SOLO_PROLOGUE_PROBE
    const bool was_staged = PERSIST && staged;
#endif
    if (PERSIST && staged) {
      asm volatile("cp.async.wait_all;" ::: "memory");
      __syncwarp();
      const float4 *src = reinterpret_cast<const float4 *>(&sm.ent[kStageAt][0]);
      float4 *dst = reinterpret_cast<float4 *>(sm.kin);
      for (int i = lane; i < KL::kSize / 4; i += 32) dst[i] = src[i];
    } else {
      const float4 *src = reinterpret_cast<const float4 *>(B.kin + env * KL::kSize);
      float4 *dst = reinterpret_cast<float4 *>(sm.kin);
      for (int i = lane; i < KL::kSize / 4; i += 32) dst[i] = __ldcg(src + i);
    }
    __syncwarp();
#ifdef SOLO_PROLOGUE_PROBE
    if (lane == 0) { sm.probe[0] = clock64(); sm.probe[2] = 0; }   // record in shared memory
#endif
    // the ticket for the environment after next; read at the end of this one.  Drawn BEHIND the wait
    // on the staged record: cp.async.wait_all waits on the scoreboard the atomic's return sits on too
    int ticket = 0;
    if (PERSIST && lane == 0) ticket = atomicAdd(B.work, 1);
    // ---- contact detection: support points inside the breaking threshold, in
    // Bullet's order (by body, then by support point), capped at max_contacts ----
    int nc = 0;
    {
      constexpr int kRounds = kMaxCand / 32;
      uint32_t hits[kRounds];
#pragma unroll
      for (int i = 0; i < kRounds; i++) {
        const int c = lane + 32 * i;
        bool hit = false;
        if (c < n_cand) hit = cand_distance<CH>(sm.kin, cand_at(sm, B.cand_tab, c), margin) < C.breaking;
        hits[i] = __ballot_sync(FULL, hit);
      }
      int before = 0;
#pragma unroll
      for (int i = 0; i < kRounds; i++) {
        const int rank = before + __popc(hits[i] & ((1u << lane) - 1u));
        SOLO_CHECK(C.max_contacts <= kMaxContacts);
        if (((hits[i] >> lane) & 1u) && rank < C.max_contacts) sm.list[rank] = lane + 32 * i;
        before += __popc(hits[i]);
      }
      nc = min(before, C.max_contacts);
    }
    __syncwarp();
#ifdef SOLO_PROLOGUE_PROBE
    if (lane == 0) sm.probe[1] = clock64();   // contacts listed
#endif
    // the solve touches entries [0, nc) only: the staging area is free unless the environment has
    // more contacts than that (rare; the next record is then loaded the plain way)
    if (PERSIST) prefetch_next(nc <= kStageAt);

    float y[W], mlam[kLegs * CH];
    int it = 0, n_active = 0;
    const uint32_t lim = __float_as_uint(sm.kin[KL::kLim]);   // live joint-limit rows (setup)
    if (lane < kLegs * CH) sm.llam[lane] = 0.f;
    __syncwarp();
    if (motor_mask != 0 || nc > 0 || lim != 0) {
      const float link_mu = B.link_friction ? B.link_friction[env] : -1.f;   // < 0: the model's value
      if (nc <= 32) it = warp_solve<CH, 1, PERSIST, STASH>(C, sm, margin, B.cand_tab, link_mu, motor_mask, lim, nc, lane, y, mlam, &n_active);
      else it = warp_solve<CH, 2, PERSIST, STASH>(C, sm, margin, B.cand_tab, link_mu, motor_mask, lim, nc, lane, y, mlam, &n_active);
    } else {
#pragma unroll
      for (int a = 0; a < W; a++) y[a] = 0.f;
#pragma unroll
      for (int j = 0; j < kLegs * CH; j++) mlam[j] = 0.f;
    }

#ifdef SOLO_PROLOGUE_PROBE
    const long long t_solved = clock64();
#endif
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
#ifdef SOLO_PROLOGUE_PROBE
      // probe build (tools/solver_prologue.py): cycles until the record is in shared memory, from
      // there until the contacts are listed, from there until the iterations start (0: no solve);
      // row 3: the iterations themselves
      st[0] = (float)(sm.probe[0] - t_begin) * (was_staged ? 1.f : -1.f);   // < 0: the record came the plain way
      st[(int64_t)B.n] = (float)(sm.probe[1] - sm.probe[0]);
      st[2 * (int64_t)B.n] = sm.probe[2] ? (float)(sm.probe[2] - sm.probe[1]) : 0.f;
      st[3 * (int64_t)B.n] = sm.probe[2] ? (float)(t_solved - sm.probe[2]) : 0.f;
#endif
#ifdef SOLO_SECTION_PROBE
      // probe build (tools/solver_sections.py): the contact-count rows carry cycles in the normal rows
      // (low 3 decimal digits: active contacts) and in the friction pairs (low 3 digits: screens / 10)
      if (it > 0) {
        st[0] = (float)(it + 100 * (sm.probe[3] / 100));
        st[(int64_t)B.n] = (float)sm.probe[0];
        st[2 * (int64_t)B.n] = (float)sm.probe[1];
      }
#endif
#ifdef SOLO_TIMELINE
      {
        // probe build (tools/solver_timeline.py): the contact-count rows carry the block's start
        // and end time in ns (low 32 bits of %globaltimer) instead
        unsigned tl_end;
        asm volatile("mov.u32 %0, %%globaltimer_lo;" : "=r"(tl_end));
        st[(int64_t)B.n] = __uint_as_float(tl_begin);
        st[2 * (int64_t)B.n] = __uint_as_float(tl_end);
      }
#endif
    }
    if (DBG) {
      __syncwarp();
      if (B.dbg_normal_force) {
        for (int c = lane; c < n_cand; c += 32) B.dbg_normal_force[(int64_t)c * B.n + env] = 0.f;
        __syncwarp();
        for (int k = lane; k < n_active; k += 32) {
          const float *e = &sm.ent[k][0];
          const int cand = __float_as_int(e[AL::kMeta + 6]);
          B.dbg_normal_force[(int64_t)cand * B.n + env] = e[AL::kMeta] / e[W + 1] * C.inv_dt;
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
            if (d < ndof_total)
              B.dbg_motor_torque[(int64_t)d * B.n + env] = mlam[j] * sm.kin[KL::kMotor + j * RL::kRow + RL::kIs] * C.inv_dt;
            d++;
          }
        }
      }
    }
    if (!PERSIST) break;
    // ---- next environment ----
    __syncwarp();
    env = nxt;
    nxt = 2 * (int64_t)gridDim.x + __shfl_sync(FULL, ticket, 0);
  }
  if (PERSIST && lane == 0) {
    // the last warp of the launch leaves the counters at zero for the next one: every warp's
    // tickets are drawn before it checks out (work[1] wraps to 0 at the grid size by itself)
    __threadfence();
    if (atomicInc(reinterpret_cast<unsigned *>(B.work) + 1, gridDim.x - 1) == gridDim.x - 1) {
      B.work[0] = 0;
    }
  }
}

}  // namespace solo
#endif  // SOLO_SOLVE_WARP_CUH_
