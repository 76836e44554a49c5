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

extern "C" {
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
