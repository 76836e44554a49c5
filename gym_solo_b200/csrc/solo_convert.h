// solo_convert.h -- host-side conversion of the C-ABI PODs (include/solo8_b200.h,
// double precision, as written by the model compiler) into the constant tables
// the per-environment core consumes.  Host only.
#ifndef SOLO_CONVERT_H_
#define SOLO_CONVERT_H_

#include <string.h>

#include <string>

#include "../../include/solo8_b200.h"
#include "solo_phases.h"

namespace solo {

static_assert(kMaxDof == SOLO_MAX_DOF && kMaxBodies == SOLO_MAX_BODIES && kMaxApi == SOLO_MAX_API &&
                  kMaxCand == SOLO_MAX_CAND && kLegs == SOLO_MAX_LEGS && kStateBase == SOLO_STATE_BASE &&
                  kMaxObsTerms == SOLO_MAX_OBS_TERMS && kMaxRewardOps == SOLO_MAX_REWARD_OPS &&
                  kMaxTermTerms == SOLO_MAX_TERM_TERMS && kRewardStack == SOLO_REWARD_STACK,
              "core limits must match the C-ABI header");
static_assert(sizeof(ObsTermC) == sizeof(SoloObsTerm) && sizeof(RewardOpC) == sizeof(SoloRewardOp) &&
                  sizeof(TermTermC) == sizeof(SoloTermTerm) && sizeof(ProgramC) == sizeof(SoloProgram),
              "program tables are passed through bit for bit");

// Returns "" on success, an error message otherwise.
template <typename T>
inline std::string convert_model(const SoloModel &m, double plane_friction, ModelT<T> *out) {
  ModelT<T> &o = *out;
  memset(&o, 0, sizeof o);
  if (m.n_legs < 1 || m.n_legs > kLegs) return "n_legs out of range";
  if (m.ndof < 1 || m.ndof > kMaxDof) return "ndof out of range";
  if (m.n_api < m.ndof || m.n_api > kMaxApi) return "n_api out of range";
  if (m.n_cand < 0 || m.n_cand > kMaxCand) return "n_cand out of range";
  o.n_legs = m.n_legs; o.ndof = m.ndof; o.n_api = m.n_api; o.n_cand = m.n_cand;
  int d = 0;
  for (int l = 0; l < kLegs; l++) {
    o.chain_len[l] = l < m.n_legs ? m.chain_len[l] : 0;
    if (o.chain_len[l] < 0 || o.chain_len[l] > SOLO_MAX_CHAIN) return "chain_len out of range";
    o.dof0[l] = d;
    d += o.chain_len[l];
  }
  if (d != m.ndof) return "sum(chain_len) != ndof";
  for (int i = 0; i < kMaxDof; i++) o.dof_api[i] = 0;
  int seen = 0;
  for (int a = 0; a < kMaxApi; a++) {
    o.api_dof[a] = a < m.n_api ? m.api_dof[a] : -1;
    if (o.api_dof[a] >= m.ndof) return "api_dof out of range";
    if (o.api_dof[a] >= 0) { o.dof_api[o.api_dof[a]] = a; seen++; }
    o.api_lower[a] = (T)m.api_lower[a];
    o.api_upper[a] = (T)m.api_upper[a];
  }
  if (seen != m.ndof) return "api_dof does not cover every dof";
  for (int b = 0; b <= m.ndof; b++) {
    o.mass[b] = (T)m.body_mass[b];
    for (int i = 0; i < 3; i++) o.com[b][i] = (T)m.body_com[b][i];
    const double *I = m.body_inertia[b];
    o.inertia[b][0] = (T)I[0]; o.inertia[b][1] = (T)(0.5 * (I[1] + I[3])); o.inertia[b][2] = (T)(0.5 * (I[2] + I[6]));
    o.inertia[b][3] = (T)I[4]; o.inertia[b][4] = (T)(0.5 * (I[5] + I[7])); o.inertia[b][5] = (T)I[8];
  }
  for (int j = 0; j < m.ndof; j++) {
    for (int i = 0; i < 3; i++) { o.jorigin[j][i] = (T)m.joint_origin[j][i]; o.jaxis[j][i] = (T)m.joint_axis[j][i]; }
    for (int i = 0; i < 9; i++) o.jrot[j][i] = (T)m.joint_rot[j][i];
  }
  // candidates must be grouped by body in ascending order
  int c = 0;
  for (int b = 0; b <= m.ndof; b++) {
    o.cand_begin[b] = c;
    while (c < m.n_cand && m.cand_body[c] == b) c++;
  }
  if (c != m.n_cand) return "contact candidates are not grouped by ascending body";
  for (int b = m.ndof + 1; b <= kMaxBodies; b++) o.cand_begin[b] = c;
  for (int i = 0; i < m.n_cand; i++) {
    for (int k = 0; k < 3; k++) o.cand_pos[i][k] = (T)m.cand_pos[i][k];
    o.cand_mu[i] = (T)(m.cand_mu[i] * plane_friction);
  }
  o.margin = (T)m.collision_margin;
  return "";
}

template <typename T>
inline std::string convert_config(const SoloConfig &c, ConfigT<T> *out) {
  ConfigT<T> &o = *out;
  if (!(c.dt > 0)) return "dt must be positive";
  if (c.solver_iterations < 0) return "solver_iterations must be >= 0";
  if (c.max_contacts < 0 || c.max_contacts > kMaxContacts) return "max_contacts out of range (0..56)";
  if (c.control_mode != SOLO_CONTROL_POSITION && c.control_mode != SOLO_CONTROL_TORQUE) return "bad control_mode";
  o.dt = (T)c.dt; o.inv_dt = (T)(1.0 / c.dt);
  for (int i = 0; i < 3; i++) o.gravity[i] = (T)c.gravity[i];
  o.torque_limit = (T)c.motor_torque_limit;
  o.lin_damp = (T)c.linear_damping; o.ang_damp = (T)c.angular_damping;
  o.kp = (T)c.motor_kp; o.kd = (T)c.motor_kd;
  o.erp = (T)c.contact_erp; o.slop = (T)c.contact_slop; o.breaking = (T)c.contact_breaking_threshold;
  o.max_vel = (T)c.max_coordinate_velocity; o.res_thr = (T)c.residual_threshold;
  if (!(c.joint_limit_erp >= 0) || !(c.joint_limit_max_impulse >= 0)) return "bad joint limit parameters";
  o.lim_erp = (T)c.joint_limit_erp; o.lim_max = (T)c.joint_limit_max_impulse;
  o.iters = c.solver_iterations; o.substeps = c.num_substeps > 0 ? c.num_substeps : 1;
  o.control_mode = c.control_mode; o.max_contacts = c.max_contacts;

  return "";
}

inline int max_chain(const SoloModel &m) {
  int mc = 0;
  for (int l = 0; l < m.n_legs && l < kLegs; l++) mc = m.chain_len[l] > mc ? m.chain_len[l] : mc;
  return mc;
}

inline std::string check_program(const SoloProgram &p, int n_api, int *obs_dim) {
  if (p.n_obs < 0 || p.n_obs > kMaxObsTerms) return "n_obs out of range";
  if (p.n_rew < 0 || p.n_rew > kMaxRewardOps) return "n_rew out of range";
  if (p.n_term < 0 || p.n_term > kMaxTermTerms) return "n_term out of range";
  int dim = 0;
  for (int i = 0; i < p.n_obs; i++) {
    if (p.obs[i].kind == SOLO_OBS_TORSO_IMU) dim += 9;
    else if (p.obs[i].kind == SOLO_OBS_MOTOR_ENCODER) dim += n_api;
    else return "unknown observation kind";
  }
  int sp = 0;
  for (int i = 0; i < p.n_rew; i++) {
    const SoloRewardOp &op = p.rew[i];
    switch (op.op) {
      case SOLO_REW_UPRIGHT: case SOLO_REW_FLAT_TORSO: case SOLO_REW_SMALL_CONTROL: case SOLO_REW_HSPEED:
      case SOLO_REW_HEIGHT: case SOLO_REW_CONST:
        sp++;
        break;
      case SOLO_REW_ADD: case SOLO_REW_MUL:
        if (op.n < 1 || op.n > kRewardStack || op.n > sp) return "reward program: stack underflow";
        sp -= op.n - 1;
        break;
      default: return "unknown reward op";
    }
    if (sp > kRewardStack) return "reward program: stack overflow";
  }
  if (p.n_rew > 0 && sp != 1) return "reward program must leave exactly one value";
  for (int i = 0; i < p.n_term; i++)
    if (p.term[i].kind != SOLO_TERM_PERPETUAL && p.term[i].kind != SOLO_TERM_TIME) return "unknown termination kind";
  *obs_dim = dim;
  return "";
}

}  // namespace solo
#endif  // SOLO_CONVERT_H_
