// hostsim.cpp -- TEST INFRASTRUCTURE ONLY.  Compiles the per-environment core of
// the CUDA kernels (gym_solo_b200/csrc/solo_core.h) for the host CPU so that the
// `-m "not gpu"` tests can check the kernel's arithmetic (float32 and, for
// diagnosis, float64) against the oracle without a GPU.  Nothing in the product
// path (gym_solo_b200/) loads this library.
#include <string>
#include <vector>

#include "../../gym_solo_b200/csrc/solo_convert.h"

using namespace solo;

template <typename T, int CH>
static int run_step(const SoloModel *m, const SoloConfig *c, double *state, const double *action, int32_t *diag2,
                    double *normal_force, double *motor_torque) {
  ModelT<T> M;
  ConfigT<T> C;
  if (!convert_model<T>(*m, c->plane_friction, &M).empty()) return -1;
  if (!convert_config<T>(*c, &C).empty()) return -1;
  EnvState<T, CH> s;
  load_state<T, CH, double>(M, state, 1, s);
  T cmd[kLegs][CH];
  load_command<T, CH, double>(M, action, T(1), cmd);
  std::vector<T> scratch((size_t)HostScratch<T, CH>::kSize);
  std::vector<T> nf(kMaxCand), mt(kMaxDof);
  DebugOut<T> dbg = {nf.data(), mt.data(), 1};
  StepDiag d = {0, 0};
  for (int k = 0; k < C.substeps; k++) physics_step<T, CH>(M, C, s, cmd, scratch.data(), &d, &dbg);
  store_state<T, CH, double>(M, s, state, 1);
  if (diag2) { diag2[0] = d.n_contacts; diag2[1] = d.iterations; }
  if (normal_force) for (int i = 0; i < M.n_cand; i++) normal_force[i] = (double)nf[i];
  if (motor_torque) for (int i = 0; i < M.ndof; i++) motor_torque[i] = (double)mt[i];
  return 0;
}

template <typename T, int CH>
static int run_epilogue(const SoloModel *m, const SoloProgram *p, const double *state, double *obs, double *reward) {
  ModelT<T> M;
  if (!convert_model<T>(*m, 1.0, &M).empty()) return -1;
  int dim = 0;
  if (!check_program(*p, M.n_api, &dim).empty()) return -2;
  ProgramC P;
  memcpy(&P, p, sizeof P);
  EnvState<T, CH> s;
  load_state<T, CH, double>(M, state, 1, s);
  T e[3];
  euler_from_quat(s.quat, e);
  std::vector<T> o(kMaxObsDim);
  int n = compute_obs<T, CH>(M, P, s, e, o.data());
  for (int i = 0; i < n; i++) obs[i] = (double)o[i];
  if (reward) *reward = P.n_rew > 0 ? (double)compute_reward<T, CH>(M, P, s, e) : 0.0;
  return n;
}

// Diagnosis only: the three phases in different precisions (TS setup, TV solve, finish in TS),
// to see which phase's float32 rounding an error comes from.
template <typename TS, typename TV, int CH>
static int run_mixed(const SoloModel *m, const SoloConfig *c, double *state, const double *action) {
  typedef RowLayout<CH> RL;
  ModelT<TS> M; ConfigT<TS> C; ConfigT<TV> CV;
  if (!convert_model<TS>(*m, c->plane_friction, &M).empty()) return -1;
  if (!convert_config<TS>(*c, &C).empty() || !convert_config<TV>(*c, &CV).empty()) return -1;
  EnvState<TS, CH> s;
  load_state<TS, CH, double>(M, state, 1, s);
  TS cmd[kLegs][CH];
  load_command<TS, CH, double>(M, action, TS(1), cmd);
  std::vector<TS> rows((size_t)RL::kSlots * RL::kRow), fac(FactorLayout<CH>::kSize), ctab(4 * kMaxCand);
  for (int i = 0; i < M.n_cand; i++) { for (int k = 0; k < 3; k++) ctab[4 * i + k] = M.cand_pos[i][k]; ctab[4 * i + 3] = M.cand_mu[i]; }
  const RowHeader h = setup_step<TS, CH>(M, C, s, cmd, ctab.data(), rows.data(), RL::kRow, fac.data(), 1);
  std::vector<TV> rv(rows.size()), lam(3 * kMaxContacts);
  for (size_t i = 0; i < rows.size(); i++) rv[i] = (TV)rows[i];
  uint32_t motor_mask = 0;
  if (C.control_mode == 0)
    for (int l = 0; l < M.n_legs; l++) for (int i = 0; i < M.chain_len[l]; i++) motor_mask |= 1u << (l * CH + i);
  TV yb[6], yl[kLegs][CH], mlam[kLegs * CH];
  solve_rows<TV, CH>(CV, motor_mask, rv.data(), RL::kRow, h, lam.data(), 1, yb, yl, mlam);
  TS yb2[6], yl2[kLegs][CH];
  for (int a = 0; a < 6; a++) yb2[a] = (TS)yb[a];
  for (int l = 0; l < kLegs; l++) for (int k = 0; k < CH; k++) yl2[l][k] = (TS)yl[l][k];
  finish_step<TS, CH>(M, C, s, fac.data(), 1, yb2, yl2);
  store_state<TS, CH, double>(M, s, state, 1);
  return 0;
}

// splice: everything in double except one piece taken from the float32 setup
// piece 0: motor rows, 1: contact rows, 2: factors (Ls, Bh, E), 3: post-drift velocities vb / vl
template <int CH>
static int run_splice(const SoloModel *m, const SoloConfig *c, int piece, double *state, const double *action) {
  typedef RowLayout<CH> RL;
  typedef FactorLayout<CH> FL;
  ModelT<double> M; ConfigT<double> C; ModelT<float> Mf; ConfigT<float> Cf;
  convert_model<double>(*m, c->plane_friction, &M); convert_config<double>(*c, &C);
  convert_model<float>(*m, c->plane_friction, &Mf); convert_config<float>(*c, &Cf);
  EnvState<double, CH> s; EnvState<float, CH> sf;
  load_state<double, CH, double>(M, state, 1, s); load_state<float, CH, double>(Mf, state, 1, sf);
  double cmd[kLegs][CH]; float cmdf[kLegs][CH];
  load_command<double, CH, double>(M, action, 1.0, cmd); load_command<float, CH, double>(Mf, action, 1.f, cmdf);
  std::vector<double> rows((size_t)RL::kSlots * RL::kRow), fac(FL::kSize), ctab(4 * kMaxCand), lam(3 * kMaxContacts);
  std::vector<float> rowsf(rows.size()), facf(FL::kSize), ctabf(4 * kMaxCand);
  for (int i = 0; i < M.n_cand; i++) { for (int k = 0; k < 3; k++) { ctab[4 * i + k] = M.cand_pos[i][k]; ctabf[4 * i + k] = Mf.cand_pos[i][k]; } ctab[4 * i + 3] = M.cand_mu[i]; ctabf[4 * i + 3] = Mf.cand_mu[i]; }
  const RowHeader h = setup_step<double, CH>(M, C, s, cmd, ctab.data(), rows.data(), RL::kRow, fac.data(), 1);
  const RowHeader hf = setup_step<float, CH>(Mf, Cf, sf, cmdf, ctabf.data(), rowsf.data(), RL::kRow, facf.data(), 1);
  if (hf.nc != h.nc) return -5;
  if (piece == 0) for (int i = 0; i < RL::kMotorSlots * RL::kRow; i++) rows[i] = rowsf[i];
  if (piece == 1) for (size_t i = (size_t)RL::kMotorSlots * RL::kRow; i < rows.size(); i++) rows[i] = rowsf[i];
  if (piece == 2) for (int i = 0; i < FL::kVb; i++) fac[i] = facf[i];
  if (piece == 3) for (int i = FL::kVb; i < FL::kSize; i++) fac[i] = facf[i];
  uint32_t motor_mask = 0;
  if (C.control_mode == 0)
    for (int l = 0; l < M.n_legs; l++) for (int i = 0; i < M.chain_len[l]; i++) motor_mask |= 1u << (l * CH + i);
  double yb[6], yl[kLegs][CH], mlam[kLegs * CH];
  solve_rows<double, CH>(C, motor_mask, rows.data(), RL::kRow, h, lam.data(), 1, yb, yl, mlam);
  finish_step<double, CH>(M, C, s, fac.data(), 1, yb, yl);
  store_state<double, CH, double>(M, s, state, 1);
  return 0;
}

extern "C" {
int hostsim_step_splice(const SoloModel *m, const SoloConfig *c, int piece, double *state, const double *action) {
  return max_chain(*m) <= 2 ? run_splice<2>(m, c, piece, state, action) : run_splice<3>(m, c, piece, state, action);
}
// mode 0: setup double, solve float; mode 1: setup float, solve double
int hostsim_step_mixed(const SoloModel *m, const SoloConfig *c, int mode, double *state, const double *action) {
  const int ch = max_chain(*m);
  if (ch <= 2) return mode == 0 ? run_mixed<double, float, 2>(m, c, state, action) : run_mixed<float, double, 2>(m, c, state, action);
  return mode == 0 ? run_mixed<double, float, 3>(m, c, state, action) : run_mixed<float, double, 3>(m, c, state, action);
}
int hostsim_step(const SoloModel *m, const SoloConfig *c, int use_double, double *state, const double *action,
                 int32_t *diag2, double *normal_force, double *motor_torque) {
  const int ch = max_chain(*m);
  if (ch <= 2) {
    return use_double ? run_step<double, 2>(m, c, state, action, diag2, normal_force, motor_torque)
                      : run_step<float, 2>(m, c, state, action, diag2, normal_force, motor_torque);
  }
  return use_double ? run_step<double, 3>(m, c, state, action, diag2, normal_force, motor_torque)
                    : run_step<float, 3>(m, c, state, action, diag2, normal_force, motor_torque);
}
int hostsim_epilogue(const SoloModel *m, const SoloProgram *p, int use_double, const double *state, double *obs,
                     double *reward) {
  const int ch = max_chain(*m);
  if (ch <= 2) return use_double ? run_epilogue<double, 2>(m, p, state, obs, reward) : run_epilogue<float, 2>(m, p, state, obs, reward);
  return use_double ? run_epilogue<double, 3>(m, p, state, obs, reward) : run_epilogue<float, 3>(m, p, state, obs, reward);
}
int hostsim_sizeof_model(void) { return (int)sizeof(SoloModel); }
int hostsim_sizeof_program(void) { return (int)sizeof(SoloProgram); }
}
