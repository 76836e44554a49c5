"""Solo8VanillaEnv -- batched drop-in for gym_solo/envs/solo8v2vanilla.py:18-172.

`Solo8VanillaEnv(num_envs=N)` keeps the reference's constructor flags, the
12-wide action layout (fixed ankle slots ignored), action normalisation, the
factories, `step -> (obs, reward, done, {'labels': ...})` and `reset -> obs`,
with a leading num_envs axis on every value.  One `step()` is one C-ABI call:
three kernels per physics substep (setup -> solve -> finish; motor rows, physics,
observations, reward and terminations all inside them), replayed as ONE captured
CUDA graph when everything registered is a built-in; user-defined plug-ins fall
back to the physics launches followed by their own torch code.
"""
from dataclasses import dataclass
from typing import Any, Dict, Tuple
import time

import numpy as np
import torch

from gym_solo_b200 import _abi
from gym_solo_b200 import client as bclient
from gym_solo_b200 import solo_types
from gym_solo_b200 import spaces
from gym_solo_b200.core import _epoch
from gym_solo_b200.core.configs import Solo8BaseConfig
from gym_solo_b200.envs.solo8_base_env import Solo8BaseEnv


@dataclass
class Solo8VanillaConfig(Solo8BaseConfig):
  urdf_path: str = 'assets/solo8v2/solo.urdf'
  starting_joint_pos = {
    'FL_HFE': np.pi / 2,
    'FL_KFE': np.pi,
    'FL_ANKLE': 0,
    'FR_HFE': np.pi / 2,
    'FR_KFE': np.pi,
    'FR_ANKLE': 0,
    'HL_HFE': -np.pi / 2,
    'HL_KFE': -np.pi,
    'HL_ANKLE': 0,
    'HR_HFE': -np.pi / 2,
    'HR_KFE': -np.pi,
    'HR_ANKLE': 0
  }


@dataclass
class Solo8MMRConfig(Solo8VanillaConfig):
  """examples/solo8_mmr/interactive_pos_control.py:15-18: the same env class
  with the MMR v4 URDF."""
  urdf_path: str = 'assets/solo8_URDF_v4/solo8_URDF_v4.urdf'


_CONTROL = {'position': _abi.CONTROL_POSITION, 'torque': _abi.CONTROL_TORQUE}


class Solo8VanillaEnv(Solo8BaseEnv):
  """An unmodified solo8 gym environment, num_envs at a time."""

  def __init__(self, use_gui: bool = False, realtime: bool = False,
               config=None, normalize_actions: bool = False,
               normalize_observations: bool = False, num_envs: int = 1,
               device='cuda:0', auto_reset: bool = False, output_ring: int = 2, **kwargs):
    """`output_ring` (batched-only): step() hands out its result tensors from a ring of that
    many pre-allocated sets (obs, reward, done, terminal observation), so a result stays valid
    until `output_ring` further steps have been taken and the step replays a captured graph;
    0 allocates fresh tensors on every step."""
    self._ring_n = max(0, int(output_ring))
    self._ring = None
    self._ring_at = 0
    self._realtime = realtime
    self._normalize = normalize_actions
    self._auto_reset = bool(auto_reset)
    self._action_space = None
    self._plan_sig = None
    self._plan = None
    self._host = None
    super().__init__(config or Solo8VanillaConfig(), use_gui,
                     normalize_observations=normalize_observations,
                     num_envs=num_envs, device=device)

  # ---- construction ---------------------------------------------------------------
  def _create_client(self):
    cfg = self.config
    if cfg.control_mode not in _CONTROL:
      raise ValueError('control_mode must be "position" or "torque"')
    self._blob = _abi.load_model_blob(cfg.urdf)
    # changeDynamics on joints 0..n-1 (solo8v2vanilla.py:157-163): every link but
    # the base gets config.lateral_friction; the base keeps its URDF value
    # (a [num_envs] array of lateral frictions = one value per environment)
    per_env_friction = np.ndim(cfg.lateral_friction) == 1
    self._model = _abi.make_model(self._blob, link_friction=float(np.mean(cfg.lateral_friction)))
    self._sim_config = _abi.make_config(
        dt=cfg.dt, gravity=tuple(cfg.gravity), motor_torque_limit=cfg.motor_torque_limit,
        linear_damping=cfg.linear_damping, angular_damping=cfg.angular_damping,
        num_substeps=cfg.num_substeps, control_mode=_CONTROL[cfg.control_mode],
        max_contacts=cfg.max_contacts, solver_iterations=cfg.solver_iterations)
    client = bclient.BatchedClient(self._blob, self._model, self._sim_config, self.num_envs,
                                   self.device)
    if per_env_friction:
      client.set_link_friction(cfg.lateral_friction)
    # config.base_mass_scale ([num_envs], optional; not a reference field): per-env payload
    if getattr(cfg, 'base_mass_scale', None) is not None:
      client.set_base_mass_scale(cfg.base_mass_scale)
    return client

  def load_bodies(self):
    """solo8v2vanilla.py:145-172."""
    self.robot = bclient.ROBOT_ID
    self._joint_cnt = self.client.getNumJoints(self.robot)
    self._zero_gains = np.zeros(self._joint_cnt)
    self.joint_ordering = [self.client.getJointInfo(self.robot, j)[1].decode('UTF-8')
                           for j in range(self._joint_cnt)]
    bound = (self.config.motor_torque_limit if self.config.control_mode == 'torque'
             else self.config.max_motor_rotation)
    self._action_space = spaces.Box(-bound, bound, shape=(self._joint_cnt,))

  @property
  def action_space(self) -> spaces.Space:
    """solo8v2vanilla.py:51-70."""
    if not self._action_space:
      raise ValueError('No valid action space')

    if self._normalize:
      return spaces.Box(low=-1, high=1, shape=self._action_space.shape)
    else:
      return self._action_space

  # ---- fused program --------------------------------------------------------------
  def _build_plan(self):
    """Compile the registered observations / rewards / terminations into the
    kernel's descriptor tables; None when any of them is a user subclass (or a
    factory is empty: the Python path then raises the reference's ValueError)."""
    # one integer moves whenever a registration, an add_term or an attribute of a built-in
    # plug-in changes (core/_epoch.py): the descriptor tables are recompiled then
    sig = (_epoch.value, self._auto_reset)
    if sig == self._plan_sig:
      return self._plan
    self._plan_sig, self._plan = sig, None
    o = self.obs_factory.fused_terms()
    r = self.reward_factory.fused_ops()
    t = self.termination_factory.fused_terms()
    if o is None or r is None or t is None:
      return None
    prog = _abi.SoloProgram()
    prog.n_obs, prog.n_rew, prog.n_term = len(o), len(r), len(t)
    prog.auto_reset = 1 if self._auto_reset else 0
    for i, x in enumerate(o):
      prog.obs[i] = x
    for i, x in enumerate(r):
      prog.rew[i] = x
    for i, x in enumerate(t):
      prog.term[i] = x
    dim = sum(9 if x.kind == _abi.OBS_TORSO_IMU else self._joint_cnt for x in o)
    self._plan = dict(program=prog, obs_dim=dim, labels=self.obs_factory.labels(), token=object())
    return self._plan

  def _scale(self):
    # np.array(action) * self._action_space.high: a float32 bound
    return float(self._action_space.high[0]) if self._normalize else 1.0

  # ---- step -----------------------------------------------------------------------
  def step(self, action) -> Tuple[solo_types.obs, Any, Any, Dict[Any, Any]]:
    """solo8v2vanilla.py:72-102.  action: [num_envs, 12] tensor / array ([12]
    accepted when num_envs == 1).  Returns obs [num_envs, D], reward
    [num_envs], done [num_envs] bool, {'labels': [...]}."""
    c = self.client
    plan = self._build_plan()
    if plan is not None:
      c.set_program(plan['program'], token=plan['token'])
      obs_values, reward, done, term_obs = self._outputs(plan)
      c.step(action, self._scale(), obs_values, reward, done, term_obs)
      if self._realtime:
        time.sleep(self.config.dt)
      info = {'labels': plan['labels']}
      if self._auto_reset:
        info['terminal_observation'] = term_obs
      return obs_values, reward, done.view(torch.bool), info

    # user-defined plug-ins: physics launch, then the factories in Python
    a = c._actions(action)
    if self._normalize:
      a = a * self._scale()
    c.simulate(a, n_steps=max(1, int(self.config.num_substeps)))
    if self._realtime:
      time.sleep(self.config.dt)

    obs_values, obs_labels = self.obs_factory.get_obs()
    reward = self.reward_factory.get_reward()
    if not torch.is_tensor(reward):
      # a user reward that returns one number (the reference's signature): the same for every env
      reward = torch.full((c.n,), float(reward), dtype=torch.float32, device=c.device)
    done = self.termination_factory.is_terminated()
    if not torch.is_tensor(done):
      done = torch.full((c.n,), bool(done), dtype=torch.bool, device=c.device)
    info = {'labels': obs_labels}
    if self._auto_reset and bool(done.any()):
      info['terminal_observation'] = obs_values
      c.reset(done)
      self.termination_factory.reset(mask=done)
      fresh, _ = self.obs_factory.get_obs()
      obs_values = torch.where(done[:, None], fresh, obs_values)
    return obs_values, reward, done, info

  def _outputs(self, plan):
    c = self.client

    def fresh():
      dev = c.device
      o = torch.empty((c.n, plan['obs_dim']), dtype=torch.float32, device=dev)
      return (o, torch.empty((c.n,), dtype=torch.float32, device=dev),
              torch.empty((c.n,), dtype=torch.uint8, device=dev),
              torch.empty_like(o) if self._auto_reset else None)
    if self._ring_n == 0:
      return fresh()
    key = (plan['obs_dim'], self._auto_reset)
    if self._ring is None or self._ring[0] != key:
      self._ring = (key, [fresh() for _ in range(self._ring_n)])
      self._ring_at = 0
    self._ring_at = (self._ring_at + 1) % self._ring_n
    return self._ring[1][self._ring_at]

  def step_host(self, action: np.ndarray):
    """The reference's calling convention -- numpy in, numpy out -- through the
    C-ABI's host entry (solo_step_host), synchronous.  The kernels read the actions from and
    write observations / reward / done to page-locked host memory directly (no staging copies:
    the bytes cross the bus inside the step).  Needs every registered plug-in to be a
    built-in.  `action`: a numpy array (copied into the env's pinned buffer first) or a pinned
    float32 CPU tensor (read in place).  The returned arrays alias the env's pinned result
    block and are overwritten by the next call."""
    plan = self._build_plan()
    if plan is None:
      raise ValueError('step_host needs built-in observations, rewards and terminations')
    c = self.client
    if c._program_token is not plan['token']:
      c.set_program(plan['program'], token=plan['token'])
    h = self._host
    if h is None or h['token'] is not plan['token']:
      c.bind_staging(plan['obs_dim'])
      # one pinned block laid out as the device staging (obs | reward | done)
      d, n = plan['obs_dim'], c.n
      out = torch.empty((n * (d * 4 + 4 + 1),), dtype=torch.uint8).pin_memory()
      h = self._host = dict(token=plan['token'], out=out,
                            act=torch.empty((n, c.n_api), dtype=torch.float32).pin_memory(),
                            obs=out[:n * d * 4].view(torch.float32).view(n, d),
                            rew=out[n * d * 4:n * d * 4 + n * 4].view(torch.float32),
                            done=out[n * d * 4 + n * 4:], pinned=set())
      # numpy views of the pinned block, made once (the returned arrays alias it: they are
      # overwritten by the next step_host, like the staging buffers of any vector env)
      h['np'] = (h['act'].numpy(), h['obs'].numpy(), h['rew'].numpy(), h['done'].numpy())
      h['ret'] = (h['np'][1], h['np'][2], h['np'][3].view(np.bool_), {'labels': plan['labels']})
      h['act_ptr'] = h['act'].data_ptr()
      h['call'] = c.host_stepper(h['obs'].data_ptr(), h['rew'].data_ptr(), h['done'].data_ptr())
    ptr = None
    if type(action) is torch.Tensor:
      p = action.data_ptr()
      if p in h['pinned']:
        ptr = p
      elif (action.device.type == 'cpu' and action.dtype == torch.float32 and action.is_contiguous()
            and action.numel() == c.n * c.n_api and action.is_pinned()):
        # actions already in page-locked host memory (torch.empty(...).pin_memory()): read in place
        # (the C side checks the pointer again on every call and stages it if it is not page-locked)
        h['pinned'].add(p)
        ptr = p
    if ptr is None:
      h['np'][0][...] = np.asarray(action, dtype=np.float32).reshape(c.n, c.n_api)
      ptr = h['act_ptr']
    h['call'](ptr, self._scale())
    return h['ret']

  # ---- reset ----------------------------------------------------------------------
  def reset(self, init_call: bool = False, mask=None):
    """solo8v2vanilla.py:104-143.  The world rebuild + 500 settle steps are
    deterministic, so they run once on the GPU (BatchedClient.settle_snapshot)
    and every reset restores that snapshot.  `mask` ([num_envs] bool, batched
    only) resets a subset."""
    positions = [self.config.starting_joint_pos.get(j, 0) for j in self.joint_ordering]
    self.client.settle_snapshot(positions, self.config.robot_start_pos,
                                self.config.robot_start_orientation_euler, n_steps=500)
    self.client.reset(mask)
    self.termination_factory.reset(mask=mask)

    if init_call:
      return np.empty(shape=(0,)), []
    else:
      obs_values, _ = self.obs_factory.get_obs()
      return obs_values

  # ---- checkpoint / statistics ------------------------------------------------------
  def get_state(self):
    """[num_envs, 13 + 2 ndof]: base pos 3, quat 4 (x,y,z,w), lin vel 3, ang vel
    3, q, qdot -- everything a checkpoint needs."""
    return self.client.get_state()

  def set_state(self, state):
    self.client.set_state(state)

  def episode_statistics(self, all_gather: bool = False):
    """Totals over the FINISHED episodes of this rank's shard -- sum of returns, sum of
    lengths, number of episodes -- and the NaN restarts (sharding.summarize turns them into
    episode-weighted means); with all_gather=True (torch.distributed initialised) the four
    numbers of every rank, [world, 4] -- the only collective of the framework.  The kernel
    accumulates the totals when an episode ends (fused built-in programs)."""
    st = self.client.stats
    mine = torch.stack([st[2].sum(), st[4].sum(), st[3].sum(),
                        self.client.diag[2].to(torch.float32)])
    if not all_gather:
      return mine
    from gym_solo_b200 import sharding
    return sharding.gather_statistics(mine)
