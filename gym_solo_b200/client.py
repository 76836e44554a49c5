"""BatchedClient -- what `pybullet_utils.bullet_client.BulletClient` is to the
reference env, for `num_envs` environments resident on one B200.

It owns the C-ABI handle (include/solo8_b200.h) and the torch tensors the
handle is bound to, and exposes the subset of the BulletClient surface that
gym_solo's env and plug-ins call, batched:

  setJointMotorControlArray + stepSimulation   -> simulate(actions)
  getBasePositionAndOrientation(robot)         -> (pos [N,3], quat [N,4])
  getEulerFromQuaternion(quat)                 -> [N,3]
  getBaseVelocity(robot)                       -> (lin [N,3], ang [N,3])
  getJointState(robot, j)                      -> (q [N], qdot [N], zeros, zeros)
  getNumJoints(robot) / getJointInfo(robot, j)

so that user-written Observation / Reward subclasses keep working with tensors
in place of floats.  torch is used for device memory and streams only; all
arithmetic of the step path happens in the CUDA library.
"""
import ctypes as C
import math

import numpy as np
import torch

from gym_solo_b200 import _abi
from gym_solo_b200 import _lib
from gym_solo_b200 import camera

ROBOT_ID = 1   # plane is body 0, robot body 1 (solo8_base_env.py:47, solo8v2vanilla.py:151)

_SNAPSHOT_CACHE = {}
_RAW_STREAM = getattr(torch._C, '_cuda_getCurrentRawStream', None)


def _ptr(t):
  return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _struct_key(s):
  return bytes(memoryview(s))


class BatchedClient:
  def __init__(self, model_blob, model, config, num_envs, device):
    self.lib = _lib.load()
    self.device = torch.device(device)
    if self.device.type != 'cuda':
      raise ValueError('gym_solo_b200 runs on CUDA devices only (got %r)' % (device,))
    if not torch.cuda.is_available():
      raise _lib.SoloError('no CUDA device: the Solo 8 step path has no CPU fallback')
    self.index = self.device.index if self.device.index is not None else torch.cuda.current_device()
    self.device = torch.device('cuda', self.index)
    self.blob, self.model, self.config = model_blob, model, config
    self.n = int(num_envs)
    self.ndof, self.n_api = model.ndof, model.n_api
    self.sdim = _abi.state_dim(model.ndof)
    self.api_dof = [model.api_dof[i] for i in range(model.n_api)]
    self.handle = C.c_void_p()
    _lib.check(self.lib.solo_create(C.byref(model), C.byref(config), self.n, self.index,
                                    C.byref(self.handle)))
    dev = self.device
    self.state = torch.zeros((self.sdim, self.n), dtype=torch.float32, device=dev)
    self.counters = torch.zeros((_abi.MAX_TERM_TERMS, self.n), dtype=torch.int32, device=dev)
    self.snapshot = torch.zeros((self.sdim,), dtype=torch.float32, device=dev)
    self.stats = torch.zeros((5, self.n), dtype=torch.float32, device=dev)
    self.diag = torch.zeros((4,), dtype=torch.int64, device=dev)
    # solver scratch: per-env records for the solver warp, mass-matrix factors, yh, support points
    nbytes = int(self.lib.solo_scratch_bytes(self.handle))
    if nbytes < 0:
      raise _lib.SoloError('solo_scratch_bytes failed')
    self.scratch = torch.empty((max(nbytes, 256),), dtype=torch.uint8, device=dev)
    _lib.check(self.lib.solo_bind(self.handle, _ptr(self.state), _ptr(self.counters),
                                  _ptr(self.snapshot), _ptr(self.stats), _ptr(self.scratch)), self.handle)
    _lib.check(self.lib.solo_bind_diag(self.handle, _ptr(self.diag)), self.handle)
    self._program_token = None
    self._dbg = None
    self._staging = None
    self._mesh = None
    self._render_scratch = None
    self._link_friction = None
    self._base_mass_scale = None
    self._joint_info = None
    self._randomisation_version = 0       # bumped by set_link_friction / set_base_mass_scale
    self._per_env_snapshot_key = None

  # ---- render (solo8_base_env.py:116-142) ---------------------------------------
  def computeProjectionMatrixFOV(self, fov, aspect, nearVal, farVal):
    return camera.compute_projection_matrix_fov(fov, aspect, nearVal, farVal)

  def computeViewMatrixFromYawPitchRoll(self, cameraTargetPosition, distance, yaw, pitch, roll,
                                        upAxisIndex):
    return camera.compute_view_matrix_from_yaw_pitch_roll(cameraTargetPosition, distance, yaw, pitch,
                                                          roll, upAxisIndex)

  def bind_mesh(self, visual=None):
    """Upload the render mesh (default: the one compiled next to the model blob)."""
    v = visual if visual is not None else _abi.load_visual(self.blob)
    dev = self.device
    body = torch.tensor(v['tri_body'], dtype=torch.int32, device=dev).reshape(-1)
    verts = torch.tensor(v['tri_verts'], dtype=torch.float32, device=dev).reshape(-1, 9).contiguous()
    rgba = torch.tensor(v['tri_rgba'], dtype=torch.float32, device=dev).reshape(-1, 4).contiguous()
    if not (body.numel() == verts.shape[0] == rgba.shape[0]):
      raise ValueError('render mesh tables disagree in length')
    if body.numel() and (int(body.min()) < 0 or int(body.max()) > self.ndof):
      raise ValueError('render mesh refers to a body the model does not have')
    _lib.check(self.lib.solo_bind_mesh(self.handle, _ptr(body), _ptr(verts), _ptr(rgba), int(body.numel())),
               self.handle)
    self._mesh = (body, verts, rgba)
    self._render_scratch = None

  def getCameraImage(self, width, height, viewMatrix, projectionMatrix, env_begin=0, env_count=1,
                     out=None, depth=True, segmentation=True, **shading):
    """pybullet.getCameraImage for environments [env_begin, env_begin + env_count): returns
    pybullet's five values (width, height, rgba, depth, segmentation) as device tensors --
    rgba uint8 [env_count, height, width, 4]; depth float32 [env_count, height, width], the
    OpenGL depth buffer in [0, 1]; segmentation int32 (plane 0, robot 1, nothing -1).
    depth=False / segmentation=False skip that image (None in its place)."""
    if self._mesh is None:
      self.bind_mesh()
    cam, _ = _abi.make_camera(viewMatrix, projectionMatrix, width, height, **shading)
    cnt_all, hh, ww = int(env_count), int(height), int(width)
    need = int(self.lib.solo_render_scratch_bytes(self.handle, cnt_all))
    if need < 0:
      raise ValueError('env_count must be >= 1')
    if self._render_scratch is None or self._render_scratch.numel() < need:
      self._render_scratch = torch.empty((need,), dtype=torch.uint8, device=self.device)
    if out is None:
      out = torch.empty((cnt_all, hh, ww, 4), dtype=torch.uint8, device=self.device)
    elif (out.dtype != torch.uint8 or not out.is_contiguous() or out.device != self.device
          or out.numel() != cnt_all * hh * ww * 4):
      raise ValueError('out must be a contiguous uint8 [env_count, height, width, 4] tensor on the device')
    dep = torch.empty((cnt_all, hh, ww), dtype=torch.float32, device=self.device) if depth else None
    seg = torch.empty((cnt_all, hh, ww), dtype=torch.int32, device=self.device) if segmentation else None
    # one call takes at most 65535 environments (grid z): larger ranges go in chunks
    pix = hh * ww
    for b in range(0, cnt_all, 65535):
      cnt = min(65535, cnt_all - b)
      _lib.check(self.lib.solo_render(
          self.handle, C.byref(cam), int(env_begin) + b, cnt, C.c_void_p(out.data_ptr() + b * pix * 4),
          C.c_void_p(dep.data_ptr() + b * pix * 4) if dep is not None else C.c_void_p(0),
          C.c_void_p(seg.data_ptr() + b * pix * 4) if seg is not None else C.c_void_p(0),
          _ptr(self._render_scratch), self._stream()), self.handle)
    return ww, hh, out, dep, seg

  # ---- lifetime ---------------------------------------------------------------
  def disconnect(self):
    if self.handle:
      self.lib.solo_destroy(self.handle)
      self.handle = C.c_void_p()

  def __del__(self):
    try:
      self.disconnect()
    except Exception:
      pass

  def _stream(self):
    # torch's current stream on this device; the raw getter skips building a Stream object
    raw = _RAW_STREAM
    if raw is not None:
      return C.c_void_p(raw(self.index))
    return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

  # ---- program ----------------------------------------------------------------
  def set_program(self, program, token=None):
    """Install the fused epilogue tables; `token` lets callers skip re-installing
    the program that is already active."""
    if token is not None and token == self._program_token:
      return
    _lib.check(self.lib.solo_set_program(self.handle, C.byref(program)), self.handle)
    self._program_token = token

  def obs_dim(self):
    return self.lib.solo_obs_dim(self.handle)

  # ---- state ------------------------------------------------------------------
  def initial_state(self, pos, euler):
    """State right after loadURDF (solo8v2vanilla.py:151-155)."""
    s = np.zeros(self.sdim, dtype=np.float64)
    s[0:3] = pos
    s[3:7] = quaternion_from_euler(euler)
    return s

  def set_state(self, state):
    """state: [sdim] (broadcast) or [N, sdim]; numpy or tensor."""
    t = torch.as_tensor(np.asarray(state) if not torch.is_tensor(state) else state)
    t = t.to(device=self.device, dtype=torch.float32)
    if t.dim() == 1:
      self.state.copy_(t[:, None].expand(self.sdim, self.n))
    else:
      self.state.copy_(t.t())

  def get_state(self):
    """[N, sdim] float32 copy (pos 3, quat 4, lin 3, ang 3, q, qdot)."""
    return self.state.t().contiguous()

  def set_snapshot(self, state):
    """state: [sdim] -- one reset state for every env (the reference's reset is
    deterministic) -- or [N, sdim], one per env (solo_snapshot_layout)."""
    t = torch.as_tensor(np.asarray(state, dtype=np.float32) if not torch.is_tensor(state) else state)
    t = t.to(device=self.device, dtype=torch.float32)
    per_env = t.dim() == 2
    self._per_env_snapshot_key = None
    shape = (self.sdim, self.n) if per_env else (self.sdim,)
    if tuple(self.snapshot.shape) != shape:
      self.snapshot = torch.zeros(shape, dtype=torch.float32, device=self.device)
      _lib.check(self.lib.solo_bind(self.handle, _ptr(self.state), _ptr(self.counters),
                                    _ptr(self.snapshot), _ptr(self.stats), _ptr(self.scratch)), self.handle)
    _lib.check(self.lib.solo_snapshot_layout(self.handle, 1 if per_env else 0), self.handle)
    self.snapshot.copy_(t.t() if per_env else t)

  def settle_snapshot(self, start_targets_api, pos, euler, n_steps=500):
    """reset()'s settle (solo8v2vanilla.py:124-136) computed ONCE per (model,
    config, start pose): a one-environment handle runs the 500
    POSITION_CONTROL steps on the GPU; every later reset restores the result."""
    cfg = _abi.SoloConfig.from_buffer_copy(self.config)
    cfg.control_mode = _abi.CONTROL_POSITION
    cfg.num_substeps = 1
    if (np.ndim(pos) == 2 or np.ndim(euler) == 2 or self._link_friction is not None
        or self._base_mass_scale is not None):
      # start poses, friction or payload that differ per environment: the
      # settle runs for the whole batch, one snapshot per env
      pos_n = np.broadcast_to(np.asarray(pos, dtype=np.float64), (self.n, 3))
      eul_n = np.broadcast_to(np.asarray(euler, dtype=np.float64), (self.n, 3))
      # settled once per (config, targets, poses, friction / payload version): later resets --
      # masked ones included -- only run reset_kernel on the snapshot already bound
      key = (_struct_key(cfg), tuple(float(x) for x in start_targets_api), pos_n.tobytes(), eul_n.tobytes(),
             n_steps, self._randomisation_version)
      if self._per_env_snapshot_key == key:
        return None
      full = BatchedClient(self.blob, self.model, cfg, self.n, self.device)
      if self._link_friction is not None:
        _lib.check(self.lib.solo_bind_link_friction(full.handle, _ptr(self._link_friction)), full.handle)
      if self._base_mass_scale is not None:
        _lib.check(self.lib.solo_bind_base_mass_scale(full.handle, _ptr(self._base_mass_scale)), full.handle)
      full.set_state(np.stack([full.initial_state(pos_n[i], eul_n[i]) for i in range(self.n)]))
      a = torch.tensor(np.asarray(start_targets_api, dtype=np.float32), device=self.device)
      full.simulate(a, n_steps=n_steps, broadcast=True)
      snap = full.get_state()
      full.disconnect()
      self.set_snapshot(snap)
      self._per_env_snapshot_key = key
      return snap.double().cpu().numpy()
    key = (_struct_key(self.model), _struct_key(cfg), tuple(float(x) for x in start_targets_api),
           tuple(pos), tuple(euler), n_steps, self.index)
    snap = _SNAPSHOT_CACHE.get(key)
    if snap is None:
      one = BatchedClient(self.blob, self.model, cfg, 1, self.device)
      one.set_state(one.initial_state(pos, euler))
      a = torch.tensor(np.asarray(start_targets_api, dtype=np.float32), device=self.device)
      one.simulate(a, n_steps=n_steps, broadcast=True)
      snap = one.get_state()[0].double().cpu().numpy()
      one.disconnect()
      _SNAPSHOT_CACHE[key] = snap
    self.set_snapshot(snap)
    return snap

  def reset(self, mask=None):
    m = None
    if mask is not None:
      m = torch.as_tensor(mask).to(device=self.device, dtype=torch.uint8).contiguous()
    _lib.check(self.lib.solo_reset(self.handle, _ptr(m), self._stream()), self.handle)

  # ---- stepping ---------------------------------------------------------------
  def _actions(self, actions, broadcast=False):
    a = actions if torch.is_tensor(actions) else torch.as_tensor(np.asarray(actions, dtype=np.float32))
    a = a.to(device=self.device, dtype=torch.float32)
    if broadcast:
      if a.shape != (self.n_api,):
        raise ValueError('broadcast action must have shape (%d,)' % self.n_api)
    else:
      if a.dim() == 1 and self.n == 1:
        a = a[None]
      if a.shape != (self.n, self.n_api):
        raise ValueError('actions must have shape (%d, %d), got %s' % (self.n, self.n_api, tuple(a.shape)))
    return a.contiguous()

  def simulate(self, actions, n_steps=1, broadcast=False):
    """setJointMotorControlArray + stepSimulation, n_steps times."""
    a = self._actions(actions, broadcast)
    _lib.check(self.lib.solo_simulate(self.handle, _ptr(a), 1 if broadcast else 0, int(n_steps),
                                      self._stream()), self.handle)

  def step(self, actions, scale, obs, reward, done, terminal_obs=None):
    a = self._actions(actions)
    _lib.check(self.lib.solo_step(self.handle, _ptr(a), float(scale), _ptr(obs), _ptr(reward), _ptr(done),
                                  _ptr(terminal_obs), self._stream()), self.handle)

  def observe(self, obs):
    _lib.check(self.lib.solo_observe(self.handle, _ptr(obs), self._stream()), self.handle)

  def reward(self, out):
    _lib.check(self.lib.solo_reward(self.handle, _ptr(out), self._stream()), self.handle)

  def bind_staging(self, obs_dim):
    dev = self.device
    # observations | reward | done in ONE allocation: solo_step_host then brings the
    # three results back with a single device->host copy when the host side is laid
    # out the same way (Solo8VanillaEnv.step_host does that)
    n, nb = self.n, self.n * (obs_dim * 4 + 4 + 1)
    self._staging_out = torch.empty((nb,), dtype=torch.uint8, device=dev)
    self._staging = (torch.empty((n, self.n_api), dtype=torch.float32, device=dev),
                     self._staging_out[:n * obs_dim * 4].view(torch.float32).view(n, obs_dim),
                     self._staging_out[n * obs_dim * 4:n * obs_dim * 4 + n * 4].view(torch.float32),
                     self._staging_out[n * obs_dim * 4 + n * 4:])
    _lib.check(self.lib.solo_bind_staging(self.handle, *[_ptr(t) for t in self._staging]), self.handle)

  def step_host(self, actions_np, scale, obs_np, reward_np, done_np, actions_ptr=None):
    """Host buffers in, host buffers out (the reference's calling convention)."""
    key = (id(actions_np), id(obs_np), id(reward_np), id(done_np))
    cached = getattr(self, '_host_ptrs', None)
    if cached is None or cached[0] != key:
      cached = (key, tuple(C.c_void_p(x.ctypes.data) for x in (actions_np, obs_np, reward_np, done_np)),
                (actions_np, obs_np, reward_np, done_np))      # keeps the arrays (and their ids) alive
      self._host_ptrs = cached
    pa, po, pr, pd = cached[1]
    if actions_ptr is not None:
      pa = C.c_void_p(actions_ptr)     # the caller's own pinned buffer: no staging copy
    _lib.check(self.lib.solo_step_host(self.handle, pa, float(scale), po, pr, pd, self._stream()), self.handle)

  def host_stepper(self, obs_ptr, reward_ptr, done_ptr):
    """A closure over solo_step_host with everything but the action pointer and the scale bound:
    the per-step host cost of Solo8VanillaEnv.step_host is one ctypes call."""
    fn, handle, check, stream = self.lib.solo_step_host, self.handle, _lib.check, self._stream

    def call(actions_ptr, scale):
      rc = fn(handle, actions_ptr, scale, obs_ptr, reward_ptr, done_ptr, stream())
      if rc:
        check(rc, handle)
    return call

  def set_link_friction(self, friction):
    """Per-env lateral friction of the non-base links ([N], None restores the model's
    value for every env): changeDynamics(lateralFriction=...) per environment."""
    self._randomisation_version += 1
    if friction is None:
      self._link_friction = None
      _lib.check(self.lib.solo_bind_link_friction(self.handle, None), self.handle)
      return
    t = torch.as_tensor(np.asarray(friction, dtype=np.float32) if not torch.is_tensor(friction) else friction)
    t = t.to(device=self.device, dtype=torch.float32).reshape(self.n) * float(self.config.plane_friction)
    self._link_friction = t.contiguous()
    _lib.check(self.lib.solo_bind_link_friction(self.handle, _ptr(self._link_friction)), self.handle)

  def set_base_mass_scale(self, scale):
    """Per-env factor on the base body's mass and inertia ([N] > 0; None restores the
    model's base for every env)."""
    self._randomisation_version += 1
    if scale is None:
      self._base_mass_scale = None
      _lib.check(self.lib.solo_bind_base_mass_scale(self.handle, None), self.handle)
      return
    t = torch.as_tensor(np.asarray(scale, dtype=np.float32) if not torch.is_tensor(scale) else scale)
    t = t.to(device=self.device, dtype=torch.float32).reshape(self.n).contiguous()
    if not bool((t > 0).all()):
      raise ValueError('base mass scales must be positive')
    self._base_mass_scale = t
    _lib.check(self.lib.solo_bind_base_mass_scale(self.handle, _ptr(t)), self.handle)

  def enable_debug(self, on=True):
    """Record per-step contact normal forces [n_cand, N] and applied motor
    torques [ndof, N] (getContactPoints()[9] / getJointState()[3])."""
    if on:
      self._dbg = (torch.zeros((self.model.n_cand, self.n), dtype=torch.float32, device=self.device),
                   torch.zeros((self.ndof, self.n), dtype=torch.float32, device=self.device))
      _lib.check(self.lib.solo_bind_debug(self.handle, _ptr(self._dbg[0]), _ptr(self._dbg[1])), self.handle)
    else:
      self._dbg = None
      _lib.check(self.lib.solo_bind_debug(self.handle, None, None), self.handle)
    return self._dbg

  def set_option(self, option, value):
    """solo_set_option: _abi.OPT_GRAPH / _abi.OPT_ZERO_COPY (both on by default)."""
    _lib.check(self.lib.solo_set_option(self.handle, int(option), int(value)), self.handle)

  def launch_count(self):
    return int(self.lib.solo_launch_count(self.handle))

  def block_size(self):
    return int(self.lib.solo_block_size(self.handle))

  def phase_timing(self, on=True):
    _lib.check(self.lib.solo_phase_timing(self.handle, 1 if on else 0), self.handle)

  def phase_times(self):
    """(physics steps timed, [setup_ms, solve_ms, finish_ms] summed) since phase_timing(True)."""
    ms = (C.c_double * 3)()
    n = int(self.lib.solo_phase_times(self.handle, ms))
    if n < 0:
      raise _lib.SoloError('solo_phase_times failed')
    return n, [float(ms[0]), float(ms[1]), float(ms[2])]

  def solver_stats(self):
    """[4, N] float32 copy: PGS iterations, contacts, contacts that carried an
    impulse, SM cycles of the last physics step's solve, per environment."""
    ptr = C.c_void_p()
    _lib.check(self.lib.solo_solver_stats(self.handle, C.byref(ptr)), self.handle)
    off = (ptr.value - self.scratch.data_ptr())
    return self.scratch[off:off + 16 * self.n].view(torch.float32).view(4, self.n).clone()

  # ---- BulletClient getters, batched -------------------------------------------
  def getBasePositionAndOrientation(self, body=ROBOT_ID):
    return self.state[0:3].t(), self.state[3:7].t()

  def getBaseVelocity(self, body=ROBOT_ID):
    return self.state[7:10].t(), self.state[10:13].t()

  def getEulerFromQuaternion(self, quat):
    return euler_from_quaternion(quat)

  def getQuaternionFromEuler(self, euler):
    return tuple(quaternion_from_euler(euler))

  def getNumJoints(self, body=ROBOT_ID):
    return self.n_api

  def getJointInfo(self, body, j):
    """pybullet's 17-tuple, derived from the compiled model (assets/joint_info.py; the
    golden tuples of gym_solo/core/test_obs_observations.py:123-162 are reproduced to 1e-9)."""
    if self._joint_info is None:
      from gym_solo_b200.assets.joint_info import joint_info
      self._joint_info = joint_info(self.blob)
    return self._joint_info[j]

  def getJointState(self, body, j):
    d = self.api_dof[j]
    z = torch.zeros((self.n,), dtype=torch.float32, device=self.device)
    if d < 0:
      return z, z, (0.,) * 6, z
    return self.state[_abi.STATE_BASE + d], self.state[_abi.STATE_BASE + self.ndof + d], (0.,) * 6, z


def quaternion_from_euler(e):
  """pybullet.getQuaternionFromEuler (solo8v2vanilla.py:153): (x, y, z, w)."""
  r, p, y = (0.5 * float(v) for v in e)
  cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
  return np.array([sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy,
                   cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy])


def euler_from_quaternion(quat):
  """pybullet.getEulerFromQuaternion on a [N,4] tensor (for user plug-ins; the
  fused path computes this in the kernel: solo_core.h euler_from_quat)."""
  q = torch.as_tensor(quat)
  x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
  sarg = -2.0 * (x * z - w * y)
  roll = torch.atan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z)
  pitch = torch.asin(torch.clamp(sarg, -1.0, 1.0))
  yaw = torch.atan2(2 * (x * y + w * z), w * w + x * x - y * y - z * z)
  hi, lo = sarg >= 0.99999, sarg <= -0.99999
  zero = torch.zeros_like(roll)
  roll = torch.where(hi | lo, zero, roll)
  pitch = torch.where(hi, zero + 0.5 * math.pi, torch.where(lo, zero - 0.5 * math.pi, pitch))
  yaw = torch.where(hi, 2 * torch.atan2(-x, y), torch.where(lo, 2 * torch.atan2(x, -y), yaw))
  return torch.stack([roll, pitch, yaw], dim=-1)
