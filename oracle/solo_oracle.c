/* solo_oracle.c -- CPU restatement (double precision) of the physics step that
 * gym_solo delegates to pybullet.  TEST INFRASTRUCTURE ONLY: imported by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs as the checker and the timed CPU baseline; never by the
 * product path.
 *
 * PARITY UNPINNED for the dynamics: the arithmetic of
 *   client.setJointMotorControlArray + client.stepSimulation
 *   (gym_solo/envs/solo8v2vanilla.py:87-91)
 * lives in the third-party `pybullet` wheel (unpinned, setup.py:7), which is
 * neither under /root/reference nor installable here or on the GPU box
 * (profiles/r2_box_probe.txt).  The reference's own tests pin no post-step numbers
 * (SURVEY.md section 8c).  What the reference does hold pins this file on the CPU
 * (tests/test_joint_info_golden.py: the 12 golden getJointInfo tuples for the model;
 * tests/test_oracle_reference_properties.py: the examples' standing height 0.33698, the
 * rest-state / response / reset properties of gym_solo/envs/test_solo8v2vanilla.py).
 * This file restates Bullet's published
 * btMultiBodyDynamicsWorld step from recollection:
 *   B2 step order: contacts at current pose -> forces + ABA -> v += a dt ->
 *      rows -> PGS -> v += dv -> integrate (semi-implicit, exponential map)
 *   B3 POSITION_CONTROL = btMultiBodyJointMotor row, kp .1, kd 1, +-force*dt
 *   B4 hull(+1 mm margin) vs ground, normal row (erp2 .08, slop 1e-5,
 *      speculative rows inside the 2 cm breaking threshold) + two friction
 *      rows (btPlaneSpace1 directions) with the implicit friction cone
 *   B6 iteration order: non-contact rows (joint-limit rows, then motor rows;
 *      direction of traversal alternates with iteration parity), normal rows,
 *      friction pairs; early exit on residual^2 <= 1e-7
 *   joint limits (URDF <limit lower upper>, solo.urdf:47) =
 *      btMultiBodyJointLimitConstraint, created at import before the motors: one
 *      unilateral row J = +-e_j per joint at or beyond a limit, 0 <= lambda <= 100,
 *      rhs = -pen * erp / dt - rel_vel with erp = m_erp = 0.2
 * Deliberately written with DIFFERENT algorithms from the CUDA kernel so the
 * two check each other: Featherstone ABA in world Plucker coordinates, a dense
 * Jacobian-built mass matrix + Cholesky for M^-1 J^T, and dense (6+ndof)-wide
 * PGS rows -- the kernel uses CRBA + Newton-Euler bias + a base Schur
 * complement and 8-sparse rows in reduced coordinates.
 *
 * Documented deviations from Bullet: contact points are a
 * fixed support-point set per link instead of GJK + persistent manifold;
 * fixed-joint children are merged into their parent body.
 */
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "../include/solo8_b200.h"

#define NMAX (6 + SOLO_MAX_DOF)
#define MAXROWS (SOLO_MAX_DOF + 3 * SOLO_MAX_CAND)

typedef struct OracleDiag {
  int32_t n_contacts;
  int32_t iterations;
  double normal_force[SOLO_MAX_CAND]; /* lambda_n / dt per support point */
  double motor_torque[SOLO_MAX_DOF];  /* applied motor impulse / dt */
} OracleDiag;

/* ---- small linear algebra ---------------------------------------------------*/
static void cross(const double *a, const double *b, double *o) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2],
         z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
static double dot3(const double *a, const double *b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
static void matvec3(const double *M, const double *v, double *o) {
  double x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2];
  double y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2];
  double z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
static void matmul3(const double *A, const double *B, double *C) {
  double T[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      T[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  memcpy(C, T, sizeof T);
}
static void matmul3_bt(const double *A, const double *B, double *C) { /* A B^T */
  double T[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      T[3 * i + j] = A[3 * i] * B[3 * j] + A[3 * i + 1] * B[3 * j + 1] + A[3 * i + 2] * B[3 * j + 2];
  memcpy(C, T, sizeof T);
}
static void quat_to_mat(const double *q, double *R) { /* q = x,y,z,w */
  double x = q[0], y = q[1], z = q[2], w = q[3];
  double n = x * x + y * y + z * z + w * w;
  double s = 2.0 / n;
  R[0] = 1 - s * (y * y + z * z); R[1] = s * (x * y - w * z); R[2] = s * (x * z + w * y);
  R[3] = s * (x * y + w * z); R[4] = 1 - s * (x * x + z * z); R[5] = s * (y * z - w * x);
  R[6] = s * (x * z - w * y); R[7] = s * (y * z + w * x); R[8] = 1 - s * (x * x + y * y);
}
static void axis_angle_mat(const double *a, double th, double *R) {
  double c = cos(th), s = sin(th), t = 1 - c;
  R[0] = t * a[0] * a[0] + c;        R[1] = t * a[0] * a[1] - s * a[2]; R[2] = t * a[0] * a[2] + s * a[1];
  R[3] = t * a[0] * a[1] + s * a[2]; R[4] = t * a[1] * a[1] + c;        R[5] = t * a[1] * a[2] - s * a[0];
  R[6] = t * a[0] * a[2] - s * a[1]; R[7] = t * a[1] * a[2] + s * a[0]; R[8] = t * a[2] * a[2] + c;
}

/* spatial vectors: (angular[3], linear[3]), world Plucker coordinates about the
 * world origin */
static void crm(const double *v, const double *s, double *o) { /* v x s */
  double a[3], b[3], c[3];
  cross(v, s, a); cross(v, s + 3, b); cross(v + 3, s, c);
  o[0] = a[0]; o[1] = a[1]; o[2] = a[2];
  o[3] = b[0] + c[0]; o[4] = b[1] + c[1]; o[5] = b[2] + c[2];
}
static void crf(const double *v, const double *f, double *o) { /* v x* f */
  double a[3], b[3], c[3];
  cross(v, f, a); cross(v + 3, f + 3, b); cross(v, f + 3, c);
  o[0] = a[0] + b[0]; o[1] = a[1] + b[1]; o[2] = a[2] + b[2];
  o[3] = c[0]; o[4] = c[1]; o[5] = c[2];
}
static void mv6(const double *M, const double *v, double *o) {
  double t[6];
  for (int i = 0; i < 6; i++) {
    double s = 0;
    for (int j = 0; j < 6; j++) s += M[6 * i + j] * v[j];
    t[i] = s;
  }
  memcpy(o, t, sizeof t);
}
static double dot6(const double *a, const double *b) {
  double s = 0;
  for (int i = 0; i < 6; i++) s += a[i] * b[i];
  return s;
}
static void spatial_inertia(double m, const double *c, const double *Ic, double *I) {
  /* rows/cols: (ang, lin) */
  double cx[9] = {0, -c[2], c[1], c[2], 0, -c[0], -c[1], c[0], 0};
  double cc = dot3(c, c);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      I[6 * i + j] = Ic[3 * i + j] + m * ((i == j ? cc : 0) - c[i] * c[j]);
      I[6 * i + 3 + j] = m * cx[3 * i + j];
      I[6 * (3 + i) + j] = -m * cx[3 * i + j];
      I[6 * (3 + i) + 3 + j] = (i == j) ? m : 0;
    }
}
static int solve_dense(int n, double *A, double *b) { /* Gaussian elimination, partial pivoting */
  for (int k = 0; k < n; k++) {
    int p = k;
    for (int i = k + 1; i < n; i++)
      if (fabs(A[i * n + k]) > fabs(A[p * n + k])) p = i;
    if (A[p * n + k] == 0) return -1;
    if (p != k) {
      for (int j = 0; j < n; j++) { double t = A[k * n + j]; A[k * n + j] = A[p * n + j]; A[p * n + j] = t; }
      double t = b[k]; b[k] = b[p]; b[p] = t;
    }
    for (int i = k + 1; i < n; i++) {
      double f = A[i * n + k] / A[k * n + k];
      for (int j = k; j < n; j++) A[i * n + j] -= f * A[k * n + j];
      b[i] -= f * b[k];
    }
  }
  for (int i = n - 1; i >= 0; i--) {
    double s = b[i];
    for (int j = i + 1; j < n; j++) s -= A[i * n + j] * b[j];
    b[i] = s / A[i * n + i];
  }
  return 0;
}
static int cholesky(int n, double *A) { /* in place, lower */
  for (int j = 0; j < n; j++) {
    double d = A[j * n + j];
    for (int k = 0; k < j; k++) d -= A[j * n + k] * A[j * n + k];
    if (d <= 0) return -1;
    d = sqrt(d);
    A[j * n + j] = d;
    for (int i = j + 1; i < n; i++) {
      double s = A[i * n + j];
      for (int k = 0; k < j; k++) s -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = s / d;
    }
  }
  return 0;
}
static void chol_solve(int n, const double *L, double *b) {
  for (int i = 0; i < n; i++) {
    double s = b[i];
    for (int k = 0; k < i; k++) s -= L[i * n + k] * b[k];
    b[i] = s / L[i * n + i];
  }
  for (int i = n - 1; i >= 0; i--) {
    double s = b[i];
    for (int k = i + 1; k < n; k++) s -= L[k * n + i] * b[k];
    b[i] = s / L[i * n + i];
  }
}

/* ---- kinematics --------------------------------------------------------------*/
typedef struct Kin {
  int nb, ndof;
  int parent[SOLO_MAX_BODIES];
  double R[SOLO_MAX_BODIES][9];  /* world <- body */
  double o[SOLO_MAX_BODIES][3];  /* body frame origin (= joint origin), world */
  double c[SOLO_MAX_BODIES][3];  /* COM, world */
  double Iw[SOLO_MAX_BODIES][9]; /* inertia about COM, world axes */
  double a[SOLO_MAX_DOF][3];     /* joint axis, world */
} Kin;

static void forward_kinematics(const SoloModel *m, const double *st, Kin *k) {
  const double *pos = st, *quat = st + 3, *q = st + SOLO_STATE_BASE;
  k->ndof = m->ndof; k->nb = m->ndof + 1;
  quat_to_mat(quat, k->R[0]);
  memcpy(k->o[0], pos, 3 * sizeof(double));
  k->parent[0] = -1;
  int d = 0;
  for (int l = 0; l < m->n_legs; l++)
    for (int j = 0; j < m->chain_len[l]; j++, d++) {
      int b = d + 1, p = (j == 0) ? 0 : d;
      k->parent[b] = p;
      double t[3], Rj[9], Rq[9];
      matvec3(k->R[p], m->joint_origin[d], t);
      for (int i = 0; i < 3; i++) k->o[b][i] = k->o[p][i] + t[i];
      matmul3(k->R[p], m->joint_rot[d], Rj);
      axis_angle_mat(m->joint_axis[d], q[d], Rq);
      matmul3(Rj, Rq, k->R[b]);
      matvec3(k->R[b], m->joint_axis[d], k->a[d]);
    }
  for (int b = 0; b < k->nb; b++) {
    double t[3], T[9];
    matvec3(k->R[b], m->body_com[b], t);
    for (int i = 0; i < 3; i++) k->c[b][i] = k->o[b][i] + t[i];
    matmul3(k->R[b], m->body_inertia[b], T);
    matmul3_bt(T, k->R[b], k->Iw[b]);
  }
}

/* body velocities (classical): omega_b and v of the body COM */
static void body_velocities(const Kin *k, const double *vgen, double w[][3], double vc[][3]) {
  /* vgen = (v_com0[3], omega0[3], qd[ndof]) */
  for (int i = 0; i < 3; i++) { w[0][i] = vgen[3 + i]; vc[0][i] = vgen[i]; }
  for (int b = 1; b < k->nb; b++) {
    int p = k->parent[b], d = b - 1;
    double r[3], t[3], vo[3];
    /* velocity of joint origin as a point of the parent */
    for (int i = 0; i < 3; i++) r[i] = k->o[b][i] - k->c[p][i];
    cross(w[p], r, t);
    for (int i = 0; i < 3; i++) vo[i] = vc[p][i] + t[i];
    for (int i = 0; i < 3; i++) w[b][i] = w[p][i] + k->a[d][i] * vgen[6 + d];
    for (int i = 0; i < 3; i++) r[i] = k->c[b][i] - k->o[b][i];
    cross(w[b], r, t);
    for (int i = 0; i < 3; i++) vc[b][i] = vo[i] + t[i];
  }
}

/* ---- forward dynamics: Featherstone ABA, world Plucker coordinates ---------- */
static int aba(const SoloModel *m, const SoloConfig *cfg, const Kin *k, const double *vgen,
               const double *tau, double *acc /* (vdot_com0, omegadot0, qdd) */) {
  int nb = k->nb;
  double w[SOLO_MAX_BODIES][3], vc[SOLO_MAX_BODIES][3];
  body_velocities(k, vgen, w, vc);
  double v[SOLO_MAX_BODIES][6], S[SOLO_MAX_BODIES][6], cJ[SOLO_MAX_BODIES][6];
  double IA[SOLO_MAX_BODIES][36], pA[SOLO_MAX_BODIES][6];
  double U[SOLO_MAX_BODIES][6], D[SOLO_MAX_BODIES], u[SOLO_MAX_BODIES];
  for (int b = 0; b < nb; b++) {
    /* spatial velocity: (omega, v of the body point at the world origin) */
    double t[3];
    cross(w[b], k->c[b], t);
    for (int i = 0; i < 3; i++) { v[b][i] = w[b][i]; v[b][3 + i] = vc[b][i] - t[i]; }
    if (b > 0) {
      int d = b - 1;
      double oxa[3];
      cross(k->o[b], k->a[d], oxa);
      for (int i = 0; i < 3; i++) { S[b][i] = k->a[d][i]; S[b][3 + i] = oxa[i]; }
      double sq[6];
      for (int i = 0; i < 6; i++) sq[i] = S[b][i] * vgen[6 + d];
      crm(v[b], sq, cJ[b]);
    }
    spatial_inertia(m->body_mass[b], k->c[b], k->Iw[b], IA[b]);
    /* external force at the COM: gravity + Bullet's velocity damping
     * f = -m v (k + k|v|), n = -(I w)(k + k|w|)  (App. B2) */
    double mass = m->body_mass[b], f[3], n[3], Iw_[3];
    double vn = sqrt(dot3(vc[b], vc[b])), wn = sqrt(dot3(w[b], w[b]));
    matvec3(k->Iw[b], w[b], Iw_);
    for (int i = 0; i < 3; i++) {
      f[i] = mass * cfg->gravity[i] - mass * vc[b][i] * (cfg->linear_damping + cfg->linear_damping * vn);
      n[i] = -Iw_[i] * (cfg->angular_damping + cfg->angular_damping * wn);
    }
    double fext[6], cxf[3];
    cross(k->c[b], f, cxf);
    for (int i = 0; i < 3; i++) { fext[i] = n[i] + cxf[i]; fext[3 + i] = f[i]; }
    double Iv[6], vIv[6];
    mv6(IA[b], v[b], Iv);
    crf(v[b], Iv, vIv);
    for (int i = 0; i < 6; i++) pA[b][i] = vIv[i] - fext[i];
  }
  for (int b = nb - 1; b >= 1; b--) {
    int p = k->parent[b];
    mv6(IA[b], S[b], U[b]);
    D[b] = dot6(S[b], U[b]);
    u[b] = tau[b - 1] - dot6(S[b], pA[b]);
    double Ia[36], pa[6], Iac[6];
    for (int i = 0; i < 6; i++)
      for (int j = 0; j < 6; j++) Ia[6 * i + j] = IA[b][6 * i + j] - U[b][i] * U[b][j] / D[b];
    mv6(Ia, cJ[b], Iac);
    for (int i = 0; i < 6; i++) pa[i] = pA[b][i] + Iac[i] + U[b][i] * u[b] / D[b];
    for (int i = 0; i < 36; i++) IA[p][i] += Ia[i];
    for (int i = 0; i < 6; i++) pA[p][i] += pa[i];
  }
  double a[SOLO_MAX_BODIES][6];
  {
    double A[36], rhs[6];
    memcpy(A, IA[0], sizeof A);
    for (int i = 0; i < 6; i++) rhs[i] = -pA[0][i];
    if (solve_dense(6, A, rhs)) return -1;
    memcpy(a[0], rhs, sizeof rhs);
  }
  for (int b = 1; b < nb; b++) {
    int p = k->parent[b];
    double ap[6];
    for (int i = 0; i < 6; i++) ap[i] = a[p][i] + cJ[b][i];
    double qdd = (u[b] - dot6(U[b], ap)) / D[b];
    acc[6 + b - 1] = qdd;
    for (int i = 0; i < 6; i++) a[b][i] = ap[i] + S[b][i] * qdd;
  }
  /* spatial -> classical acceleration of the base COM */
  double t1[3], t2[3];
  cross(a[0], k->c[0], t1);
  cross(w[0], vc[0], t2);
  for (int i = 0; i < 3; i++) {
    acc[i] = a[0][3 + i] + t1[i] + t2[i];
    acc[3 + i] = a[0][i];
  }
  return 0;
}

/* ---- dense mass matrix from body Jacobians ---------------------------------- */
static int on_path(const SoloModel *m, int body, int dof) {
  /* is dof an ancestor joint of body (same chain, not deeper)? */
  if (body == 0) return 0;
  int d0 = 0;
  for (int l = 0; l < m->n_legs; l++) {
    int d1 = d0 + m->chain_len[l];
    if (body - 1 >= d0 && body - 1 < d1) return dof >= d0 && dof <= body - 1;
    d0 = d1;
  }
  return 0;
}
/* Jacobian row of the world velocity of point p (fixed in `body`) along dir */
static void point_jacobian(const SoloModel *m, const Kin *k, int body, const double *p,
                           const double *dir, double *J) {
  int n = 6 + k->ndof;
  for (int i = 0; i < n; i++) J[i] = 0;
  double r[3], t[3];
  for (int i = 0; i < 3; i++) { J[i] = dir[i]; r[i] = p[i] - k->c[0][i]; }
  cross(r, dir, t); /* dir . (w x r) = w . (r x dir) */
  for (int i = 0; i < 3; i++) J[3 + i] = t[i];
  for (int d = 0; d < k->ndof; d++)
    if (on_path(m, body, d)) {
      for (int i = 0; i < 3; i++) r[i] = p[i] - k->o[d + 1][i];
      cross(k->a[d], r, t);
      J[6 + d] = dot3(dir, t);
    }
}
static void mass_matrix(const SoloModel *m, const Kin *k, double *M) {
  int n = 6 + k->ndof;
  memset(M, 0, sizeof(double) * n * n);
  static const double E[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int b = 0; b < k->nb; b++) {
    double Jv[3][NMAX], Jw[3][NMAX];
    for (int i = 0; i < 3; i++) {
      point_jacobian(m, k, b, k->c[b], E[i], Jv[i]);
      for (int j = 0; j < n; j++) Jw[i][j] = 0;
      Jw[i][3 + i] = 1;
      for (int d = 0; d < k->ndof; d++)
        if (on_path(m, b, d)) Jw[i][6 + d] = k->a[d][i];
    }
    for (int r = 0; r < n; r++)
      for (int c = 0; c < n; c++) {
        double s = 0;
        for (int i = 0; i < 3; i++) {
          s += m->body_mass[b] * Jv[i][r] * Jv[i][c];
          for (int j = 0; j < 3; j++) s += Jw[i][r] * k->Iw[b][3 * i + j] * Jw[j][c];
        }
        M[r * n + c] += s;
      }
  }
}

/* ---- constraint rows (Bullet btMultiBodySolverConstraint, dense) ------------- */
typedef struct Row {
  double J[NMAX], u[NMAX]; /* jacobian, M^-1 J^T */
  double rhs, dinv, lo, hi, lambda;
} Row;

typedef struct Contact {
  int cand, body;
  double p[3]; /* world point on the inflated hull */
  double dist, mu;
} Contact;

static void fill_row(int n, const double *L, Row *r) {
  memcpy(r->u, r->J, sizeof(double) * n);
  chol_solve(n, L, r->u);
  double den = 0;
  for (int i = 0; i < n; i++) den += r->J[i] * r->u[i];
  r->dinv = 1.0 / den; /* m_jacDiagABInv, relaxation (sor) = 1, cfm = 0 */
  r->lambda = 0;       /* no warm starting for btMultiBody rows */
}
static double resolve_row(int n, Row *c, double *dv) { /* resolveSingleConstraintRowGeneric */
  double jdv = 0;
  for (int i = 0; i < n; i++) jdv += c->J[i] * dv[i];
  double delta = c->rhs - jdv * c->dinv;
  double sum = c->lambda + delta;
  if (sum < c->lo) { delta = c->lo - c->lambda; c->lambda = c->lo; }
  else if (sum > c->hi) { delta = c->hi - c->lambda; c->lambda = c->hi; }
  else c->lambda = sum;
  for (int i = 0; i < n; i++) dv[i] += c->u[i] * delta;
  return delta / c->dinv;
}
static double resolve_cone(int n, Row *a, Row *b, double limit, double *dv, double *res_b) {
  /* resolveConeFrictionConstraintRows: both rows evaluated on the same dv, the
   * pair projected radially onto the disc of radius mu*lambda_n, then applied */
  double ja = 0, jb = 0;
  for (int i = 0; i < n; i++) { ja += a->J[i] * dv[i]; jb += b->J[i] * dv[i]; }
  double sa = a->lambda + (a->rhs - ja * a->dinv);
  double sb = b->lambda + (b->rhs - jb * b->dinv);
  double nrm2 = sa * sa + sb * sb;
  if (nrm2 > limit * limit) {
    double s = limit / sqrt(nrm2);
    sa *= s; sb *= s;
  }
  double da = sa - a->lambda, db = sb - b->lambda;
  a->lambda = sa; b->lambda = sb;
  for (int i = 0; i < n; i++) dv[i] += a->u[i] * da + b->u[i] * db;
  *res_b = db / b->dinv;
  return da / a->dinv;
}

static void clamp_vel(int n, double *v, double lim) {
  for (int i = 0; i < n; i++) v[i] = v[i] > lim ? lim : (v[i] < -lim ? -lim : v[i]);
}

/* One Bullet step of cfg->dt.  st: double[SOLO_STATE_DIM(ndof)], in place.
 * cmd: per-dof command (position target or torque). */
static int physics_step(const SoloModel *m, const SoloConfig *cfg, double *st, const double *cmd,
                        OracleDiag *diag) {
  int nd = m->ndof, n = 6 + nd;
  double dt = cfg->dt;
  double *pos = st, *quat = st + 3, *q = st + SOLO_STATE_BASE, *qd = q + nd;
  Kin k;
  forward_kinematics(m, st, &k);

  /* 1. contact points at the current pose (collision detection runs first) */
  Contact *ct = (Contact *)malloc(sizeof(Contact) * SOLO_MAX_CAND);
  int nc = 0;
  for (int c = 0; c < m->n_cand && nc < cfg->max_contacts; c++) {
    int b = m->cand_body[c];
    double t[3];
    matvec3(k.R[b], m->cand_pos[c], t);
    double z = k.o[b][2] + t[2];
    double dist = z - m->collision_margin;
    if (dist < cfg->contact_breaking_threshold) {
      ct[nc].cand = c; ct[nc].body = b;
      ct[nc].p[0] = k.o[b][0] + t[0]; ct[nc].p[1] = k.o[b][1] + t[1]; ct[nc].p[2] = dist;
      ct[nc].dist = dist;
      ct[nc].mu = m->cand_mu[c] * cfg->plane_friction;
      nc++;
    }
  }

  /* 2. forward dynamics, v += a dt */
  double v[NMAX], acc[NMAX], tau[SOLO_MAX_DOF] = {0};
  for (int i = 0; i < 3; i++) { v[i] = st[7 + i]; v[3 + i] = st[10 + i]; }
  for (int d = 0; d < nd; d++) {
    v[6 + d] = qd[d];
    tau[d] = 0;
    if (cfg->control_mode == SOLO_CONTROL_TORQUE) {
      double t = cmd[d], lim = cfg->motor_torque_limit;
      tau[d] = t > lim ? lim : (t < -lim ? -lim : t);
    }
  }
  if (aba(m, cfg, &k, v, tau, acc)) { free(ct); return -1; }
  for (int i = 0; i < n; i++) v[i] += acc[i] * dt;
  clamp_vel(n, v, cfg->max_coordinate_velocity);

  /* 3. rows */
  double *M = (double *)malloc(sizeof(double) * n * n);
  mass_matrix(m, &k, M);
  if (cholesky(n, M)) { free(M); free(ct); return -2; }
  int n_motor = (cfg->control_mode == SOLO_CONTROL_POSITION) ? nd : 0;
  /* joint-limit rows first (the constraints exist since URDF import; a row is emitted only
   * while q is at or beyond the limit: createConstraintRows skips penetration > 0) */
  Row *limr = (Row *)calloc((size_t)SOLO_MAX_DOF + 1, sizeof(Row));
  int n_lim = 0;
  {
    int dof_api[SOLO_MAX_DOF];
    for (int j = 0; j < m->n_api; j++)
      if (m->api_dof[j] >= 0) dof_api[m->api_dof[j]] = j;
    for (int d = 0; d < nd; d++) {
      double lower = m->api_lower[dof_api[d]], upper = m->api_upper[dof_api[d]];
      if (!(lower <= upper)) continue;
      double dir = 0, pen = 0;
      if (q[d] - lower <= 0) { dir = 1; pen = q[d] - lower; }
      else if (upper - q[d] <= 0) { dir = -1; pen = upper - q[d]; }
      if (dir == 0) continue;
      Row *r = &limr[n_lim++];
      r->J[6 + d] = dir;
      fill_row(n, M, r);
      double rel = dir * v[6 + d];
      r->rhs = (-pen * cfg->joint_limit_erp / dt - rel) * r->dinv;
      r->lo = 0; r->hi = cfg->joint_limit_max_impulse;
    }
  }
  Row *motor = (Row *)calloc((size_t)SOLO_MAX_DOF + 1, sizeof(Row));
  Row *nrm = (Row *)calloc((size_t)SOLO_MAX_CAND + 1, sizeof(Row));
  Row *fr = (Row *)calloc(2 * (size_t)SOLO_MAX_CAND + 1, sizeof(Row));
  for (int d = 0; d < n_motor; d++) {
    Row *r = &motor[d];
    r->J[6 + d] = 1;
    fill_row(n, M, r);
    /* btMultiBodyJointMotor: rhs = kp*(q*-q)/dt + qd + kd*(0 - qd); the row's
     * velocity error is rhs - J.v */
    double target = cfg->motor_kp * (cmd[d] - q[d]) / dt + v[6 + d] + cfg->motor_kd * (0 - v[6 + d]);
    r->rhs = (target - v[6 + d]) * r->dinv;
    r->hi = cfg->motor_torque_limit * dt; r->lo = -r->hi;
  }
  static const double NZ[3] = {0, 0, 1}, T1[3] = {0, -1, 0}, T2[3] = {1, 0, 0}; /* btPlaneSpace1(0,0,1) */
  for (int i = 0; i < nc; i++) {
    Row *r = &nrm[i];
    point_jacobian(m, &k, ct[i].body, ct[i].p, NZ, r->J);
    fill_row(n, M, r);
    double rel = 0;
    for (int j = 0; j < n; j++) rel += r->J[j] * v[j];
    double pen = ct[i].dist + cfg->contact_slop;
    double poserr = 0, velerr = -rel; /* restitution 0 */
    if (pen > 0) velerr -= pen / dt; else poserr = -pen * cfg->contact_erp / dt;
    r->rhs = (poserr + velerr) * r->dinv;
    r->lo = 0; r->hi = 1e10;
    for (int t = 0; t < 2; t++) {
      Row *f = &fr[2 * i + t];
      point_jacobian(m, &k, ct[i].body, ct[i].p, t ? T2 : T1, f->J);
      fill_row(n, M, f);
      double relf = 0;
      for (int j = 0; j < n; j++) relf += f->J[j] * v[j];
      f->rhs = -relf * f->dinv;
    }
  }

  /* 4. PGS (btMultiBodyConstraintSolver::solveSingleIteration order) */
  double dv[NMAX];
  for (int i = 0; i < n; i++) dv[i] = 0;
  int it;
  for (it = 0; it < cfg->solver_iterations; it++) {
    double res = 0, r, r2;
    for (int j = 0; j < n_lim + n_motor; j++) {
      int idx = (it & 1) ? j : n_lim + n_motor - 1 - j;
      r = resolve_row(n, idx < n_lim ? &limr[idx] : &motor[idx - n_lim], dv); if (r * r > res) res = r * r;
    }
    for (int j = 0; j < nc; j++) { r = resolve_row(n, &nrm[j], dv); if (r * r > res) res = r * r; }
    for (int j = 0; j < nc; j++) {
      r = resolve_cone(n, &fr[2 * j], &fr[2 * j + 1], ct[j].mu * nrm[j].lambda, dv, &r2);
      if (r * r > res) res = r * r;
      if (r2 * r2 > res) res = r2 * r2;
    }
    if (res <= cfg->residual_threshold) { it++; break; }
  }
  for (int i = 0; i < n; i++) v[i] += dv[i];
  clamp_vel(n, v, cfg->max_coordinate_velocity);

  if (diag) {
    diag->n_contacts = nc; diag->iterations = it;
    for (int c = 0; c < SOLO_MAX_CAND; c++) diag->normal_force[c] = 0;
    for (int i = 0; i < nc; i++) diag->normal_force[ct[i].cand] = nrm[i].lambda / dt;
    for (int d = 0; d < SOLO_MAX_DOF; d++) diag->motor_torque[d] = d < n_motor ? motor[d].lambda / dt : 0;
  }
  free(M); free(limr); free(motor); free(nrm); free(fr); free(ct);

  /* 5. integrate positions with the new velocities (stepPositionsMultiDof) */
  for (int i = 0; i < 3; i++) pos[i] += dt * v[i];
  {
    const double *w = v + 3;
    double ang = sqrt(dot3(w, w));
    if (ang * dt > 0.25 * M_PI) ang = 0.5 * (0.5 * M_PI) / dt; /* ANGULAR_MOTION_THRESHOLD */
    double ax[3], sc;
    if (ang < 0.001) sc = 0.5 * dt - dt * dt * dt * 0.020833333333 * ang * ang;
    else sc = sin(0.5 * ang * dt) / ang;
    for (int i = 0; i < 3; i++) ax[i] = w[i] * sc;
    double dq[4] = {ax[0], ax[1], ax[2], cos(0.5 * ang * dt)};
    /* quat <- dq (x) quat   (world-frame angular velocity) */
    double x = quat[0], y = quat[1], z = quat[2], ww = quat[3];
    double nq[4] = {dq[3] * x + dq[0] * ww + dq[1] * z - dq[2] * y,
                    dq[3] * y - dq[0] * z + dq[1] * ww + dq[2] * x,
                    dq[3] * z + dq[0] * y - dq[1] * x + dq[2] * ww,
                    dq[3] * ww - dq[0] * x - dq[1] * y - dq[2] * z};
    double nn = 1.0 / sqrt(nq[0] * nq[0] + nq[1] * nq[1] + nq[2] * nq[2] + nq[3] * nq[3]);
    for (int i = 0; i < 4; i++) quat[i] = nq[i] * nn;
  }
  for (int i = 0; i < 3; i++) { st[7 + i] = v[i]; st[10 + i] = v[3 + i]; }
  for (int d = 0; d < nd; d++) { qd[d] = v[6 + d]; q[d] += dt * qd[d]; }
  return 0;
}

/* ---- exported API -------------------------------------------------------------*/
/* One env.step worth of physics: action is API-wide (n_api), fixed-joint slots
 * ignored (solo8v2vanilla.py:87-91); cfg->num_substeps Bullet steps. */
int oracle_step(const SoloModel *m, const SoloConfig *cfg, double *state, const double *action_api,
                OracleDiag *diag) {
  double cmd[SOLO_MAX_DOF] = {0};
  for (int j = 0; j < m->n_api; j++)
    if (m->api_dof[j] >= 0) cmd[m->api_dof[j]] = action_api[j];
  int ns = cfg->num_substeps > 0 ? cfg->num_substeps : 1;
  for (int s = 0; s < ns; s++) {
    int rc = physics_step(m, cfg, state, cmd, diag);
    if (rc) return rc;
  }
  return 0;
}

/* Generalised accelerations (vdot_com, omegadot, qdd) at a state; for tests. */
int oracle_forward_dynamics(const SoloModel *m, const SoloConfig *cfg, const double *state,
                            const double *tau, double *acc) {
  Kin k;
  forward_kinematics(m, state, &k);
  double v[NMAX];
  for (int i = 0; i < 3; i++) { v[i] = state[7 + i]; v[3 + i] = state[10 + i]; }
  for (int d = 0; d < m->ndof; d++) v[6 + d] = state[SOLO_STATE_BASE + m->ndof + d];
  return aba(m, cfg, &k, v, tau, acc);
}
/* Dense mass matrix in (v_com, omega, qd) coordinates; for tests. */
int oracle_mass_matrix(const SoloModel *m, const double *state, double *M) {
  Kin k;
  forward_kinematics(m, state, &k);
  mass_matrix(m, &k, M);
  return 6 + m->ndof;
}
/* World positions of the contact support points; for tests. */
int oracle_support_points(const SoloModel *m, const double *state, double *out /* [n_cand][3] */) {
  Kin k;
  forward_kinematics(m, state, &k);
  for (int c = 0; c < m->n_cand; c++) {
    int b = m->cand_body[c];
    double t[3];
    matvec3(k.R[b], m->cand_pos[c], t);
    for (int i = 0; i < 3; i++) out[3 * c + i] = k.o[b][i] + t[i];
  }
  return m->n_cand;
}

typedef struct Job {
  const SoloModel *m; const SoloConfig *cfg;
  double *states; const double *actions; int sdim, lo, hi, n_steps, rc;
} Job;
static void *worker(void *arg) {
  Job *j = (Job *)arg;
  for (int s = 0; s < j->n_steps; s++)
    for (int e = j->lo; e < j->hi; e++) {
      int rc = oracle_step(j->m, j->cfg, j->states + (size_t)e * j->sdim,
                           j->actions + ((size_t)s * 0 + e) * j->m->n_api, NULL);
      if (rc) j->rc = rc;
    }
  return NULL;
}
/* Batch of independent envs over `n_threads` host threads (the reference's
 * "N multiprocessing environments").  states: [n][sdim], actions: [n][n_api]. */
int oracle_step_batch(const SoloModel *m, const SoloConfig *cfg, double *states, const double *actions,
                      int n, int n_steps, int n_threads) {
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 256) n_threads = 256;
  pthread_t th[256];
  Job jobs[256];
  int sdim = SOLO_STATE_DIM(m->ndof);
  for (int t = 0; t < n_threads; t++) {
    jobs[t] = (Job){m, cfg, states, actions, sdim, (int)((long)n * t / n_threads),
                    (int)((long)n * (t + 1) / n_threads), n_steps, 0};
    pthread_create(&th[t], NULL, worker, &jobs[t]);
  }
  int rc = 0;
  for (int t = 0; t < n_threads; t++) { pthread_join(th[t], NULL); if (jobs[t].rc) rc = jobs[t].rc; }
  return rc;
}
int oracle_sizeof_model(void) { return (int)sizeof(SoloModel); }
int oracle_sizeof_config(void) { return (int)sizeof(SoloConfig); }
int oracle_sizeof_diag(void) { return (int)sizeof(OracleDiag); }
