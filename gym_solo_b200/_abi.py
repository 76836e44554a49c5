"""ctypes mirror of include/solo8_b200.h (PODs only) + model blob loader.

Layout must match the header field for field; tests/test_abi.py checks the
struct sizes against the compiled libraries.
"""
import ctypes as C
import json
import os

MAX_LEGS = 4
MAX_CHAIN = 3
MAX_DOF = 12
MAX_BODIES = 13
MAX_API = 16
MAX_CAND = 128
MAX_CONTACTS = 56
RENDER_STATIC_BYTES = 4 * 1024 * 1024   # SOLO_RENDER_STATIC_BYTES (include/solo8_b200.h)
MAX_OBS_TERMS = 8
MAX_REWARD_OPS = 32
MAX_TERM_TERMS = 8
REWARD_STACK = 8
STATE_BASE = 13

CONTROL_POSITION = 0
CONTROL_TORQUE = 1

OBS_TORSO_IMU = 1
OBS_MOTOR_ENCODER = 2

REW_UPRIGHT = 1
REW_FLAT_TORSO = 2
REW_SMALL_CONTROL = 3
REW_HSPEED = 4
REW_HEIGHT = 5
REW_CONST = 6
REW_ADD = 16
REW_MUL = 17

OPT_GRAPH = 1
OPT_ZERO_COPY = 2
OPT_SOLVER_OCCUPANCY = 3
OPT_FINISH_WARMUP = 4

TERM_PERPETUAL = 1
TERM_TIME = 2


def state_dim(ndof):
  return STATE_BASE + 2 * ndof


class SoloModel(C.Structure):
  _fields_ = [
      ('n_legs', C.c_int32), ('ndof', C.c_int32), ('n_api', C.c_int32),
      ('n_cand', C.c_int32),
      ('chain_len', C.c_int32 * MAX_LEGS),
      ('api_dof', C.c_int32 * MAX_API),
      ('api_lower', C.c_double * MAX_API), ('api_upper', C.c_double * MAX_API),
      ('body_mass', C.c_double * MAX_BODIES),
      ('body_com', (C.c_double * 3) * MAX_BODIES),
      ('body_inertia', (C.c_double * 9) * MAX_BODIES),
      ('joint_origin', (C.c_double * 3) * MAX_DOF),
      ('joint_rot', (C.c_double * 9) * MAX_DOF),
      ('joint_axis', (C.c_double * 3) * MAX_DOF),
      ('cand_body', C.c_int32 * MAX_CAND),
      ('cand_pos', (C.c_double * 3) * MAX_CAND),
      ('cand_mu', C.c_double * MAX_CAND),
      ('collision_margin', C.c_double),
  ]


class SoloConfig(C.Structure):
  _fields_ = [
      ('dt', C.c_double), ('gravity', C.c_double * 3),
      ('motor_torque_limit', C.c_double),
      ('linear_damping', C.c_double), ('angular_damping', C.c_double),
      ('plane_friction', C.c_double),
      ('motor_kp', C.c_double), ('motor_kd', C.c_double),
      ('contact_erp', C.c_double), ('contact_slop', C.c_double),
      ('contact_breaking_threshold', C.c_double),
      ('max_coordinate_velocity', C.c_double),
      ('residual_threshold', C.c_double),
      ('joint_limit_erp', C.c_double), ('joint_limit_max_impulse', C.c_double),
      ('solver_iterations', C.c_int32), ('num_substeps', C.c_int32),
      ('control_mode', C.c_int32), ('max_contacts', C.c_int32),
  ]


class SoloObsTerm(C.Structure):
  _fields_ = [
      ('kind', C.c_int32), ('degrees', C.c_int32), ('clip', C.c_int32),
      ('normalize', C.c_int32),
      ('clip_lo', C.c_float * MAX_API), ('clip_hi', C.c_float * MAX_API),
      ('norm_lo', C.c_float * MAX_API), ('norm_hi', C.c_float * MAX_API),
  ]


class SoloRewardOp(C.Structure):
  _fields_ = [
      ('op', C.c_int32), ('n', C.c_int32),
      ('p', C.c_float * 4), ('coef', C.c_float * REWARD_STACK),
  ]


class SoloTermTerm(C.Structure):
  _fields_ = [('kind', C.c_int32), ('max_step_delta', C.c_int32)]


class SoloProgram(C.Structure):
  _fields_ = [
      ('n_obs', C.c_int32), ('n_rew', C.c_int32), ('n_term', C.c_int32),
      ('auto_reset', C.c_int32),
      ('obs', SoloObsTerm * MAX_OBS_TERMS),
      ('rew', SoloRewardOp * MAX_REWARD_OPS),
      ('term', SoloTermTerm * MAX_TERM_TERMS),
  ]


class SoloCamera(C.Structure):
  _fields_ = [
      ('view', C.c_float * 16), ('proj', C.c_float * 16),
      ('light_direction', C.c_float * 3), ('ambient', C.c_float), ('diffuse', C.c_float),
      ('background', C.c_float * 3), ('plane_rgb_a', C.c_float * 3), ('plane_rgb_b', C.c_float * 3),
      ('checker_size', C.c_float), ('width', C.c_int32), ('height', C.c_int32),
  ]


# render() shading defaults (ours: the reference leaves them to pybullet's renderer, which is
# absent here -- DESIGN.md section 9): light towards (-5, -40, 200)/|.|, ambient 0.6, diffuse 0.35
RENDER_DEFAULTS = dict(light_direction=(-5., -40., 200.), ambient=0.6, diffuse=0.35,
                       background=(1., 1., 1.), plane_rgb_a=(1., 1., 1.), plane_rgb_b=(0.55, 0.65, 0.85),
                       checker_size=0.5)


def make_camera(view, proj, width, height, **shading):
  opts = dict(RENDER_DEFAULTS)
  for k, v in shading.items():
    if k not in opts:
      raise ValueError('unknown render option %r' % k)
    opts[k] = v
  if len(view) != 16 or len(proj) != 16:
    raise ValueError('view and projection matrices are 16 floats, column-major')
  cam = SoloCamera()
  for i in range(16):
    cam.view[i], cam.proj[i] = float(view[i]), float(proj[i])
  for k in ('light_direction', 'background', 'plane_rgb_a', 'plane_rgb_b'):
    for i in range(3):
      getattr(cam, k)[i] = float(opts[k][i])
  cam.ambient, cam.diffuse = float(opts['ambient']), float(opts['diffuse'])
  cam.checker_size = float(opts['checker_size'])
  cam.width, cam.height = int(width), int(height)
  return cam, opts


def load_visual(blob):
  """The render() mesh compiled next to the model blob (compile_model.py `visual`):
  dict(n_tri, tri_body, tri_verts [n_tri][9] body coordinates, tri_rgba)."""
  path = os.path.join(blob.get('_dir', MODELS_DIR), blob['name'] + '_visual.json')
  if not os.path.exists(path):
    raise ValueError('no compiled render mesh for model %r' % blob['name'])
  with open(path) as f:
    return json.load(f)


MODELS_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'assets')


def load_model_blob(name_or_path):
  """Read a compiled model JSON (gym_solo_b200/assets/compile_model.py)."""
  path = name_or_path
  if not os.path.exists(path):
    path = os.path.join(MODELS_DIR, name_or_path + '.json')
  with open(path) as f:
    blob = json.load(f)
  blob['_dir'] = os.path.dirname(os.path.abspath(path))   # where its render mesh lives
  return blob


def make_model(blob, link_friction=None):
  """Blob -> SoloModel.  `link_friction` applies the env's changeDynamics call
  (gym_solo/envs/solo8v2vanilla.py:157-163): joints 0..n-1, i.e. every link
  except the base, get that lateral friction; the base keeps its URDF value."""
  m = SoloModel()
  m.n_legs, m.ndof, m.n_api, m.n_cand = (blob['n_legs'], blob['ndof'],
                                         blob['n_api'], blob['n_cand'])
  if m.n_cand > MAX_CAND or m.ndof > MAX_DOF or m.n_api > MAX_API:
    raise ValueError('model exceeds compiled limits')
  for i, v in enumerate(blob['chain_len']):
    m.chain_len[i] = v
  for i in range(m.n_api):
    m.api_dof[i] = blob['api_dof'][i]
    m.api_lower[i] = blob['api_lower'][i]
    m.api_upper[i] = blob['api_upper'][i]
  for b in range(m.ndof + 1):
    m.body_mass[b] = blob['body_mass'][b]
    for i in range(3):
      m.body_com[b][i] = blob['body_com'][b][i]
      for j in range(3):
        m.body_inertia[b][3 * i + j] = blob['body_inertia'][b][i][j]
  for d in range(m.ndof):
    for i in range(3):
      m.joint_origin[d][i] = blob['joint_origin'][d][i]
      m.joint_axis[d][i] = blob['joint_axis'][d][i]
      for j in range(3):
        m.joint_rot[d][3 * i + j] = blob['joint_rot'][d][i][j]
  for c in range(m.n_cand):
    body = blob['cand_body'][c]
    m.cand_body[c] = body
    for i in range(3):
      m.cand_pos[c][i] = blob['cand_pos'][c][i]
    mu = blob['cand_mu'][c]
    if link_friction is not None and body != 0:
      mu = link_friction
    m.cand_mu[c] = mu
  m.collision_margin = blob['collision_margin']
  return m


def make_config(dt=1e-3, gravity=(0., 0., -9.81), motor_torque_limit=2.,
                linear_damping=.04, angular_damping=.04, num_substeps=1,
                control_mode=CONTROL_POSITION, max_contacts=56,
                solver_iterations=50, **kw):
  """Solo8BaseConfig physics fields (gym_solo/core/configs.py:8-24) + the
  Bullet world defaults pybullet sets."""
  c = SoloConfig()
  c.dt = dt
  for i in range(3):
    c.gravity[i] = gravity[i]
  c.motor_torque_limit = motor_torque_limit
  c.linear_damping = linear_damping
  c.angular_damping = angular_damping
  c.plane_friction = kw.get('plane_friction', 1.0)
  c.motor_kp = kw.get('motor_kp', 0.1)
  c.motor_kd = kw.get('motor_kd', 1.0)
  c.contact_erp = kw.get('contact_erp', 0.08)
  c.contact_slop = kw.get('contact_slop', 1e-5)
  c.contact_breaking_threshold = kw.get('contact_breaking_threshold', 0.02)
  c.max_coordinate_velocity = kw.get('max_coordinate_velocity', 100.0)
  c.residual_threshold = kw.get('residual_threshold', 1e-7)
  c.joint_limit_erp = kw.get('joint_limit_erp', 0.2)
  c.joint_limit_max_impulse = kw.get('joint_limit_max_impulse', 100.0)
  c.solver_iterations = solver_iterations
  c.num_substeps = num_substeps
  c.control_mode = control_mode
  c.max_contacts = max_contacts
  return c
