"""Solo8BaseConfig (gym_solo/core/configs.py:8-38) plus the batched-only knobs.

Field names, defaults and the dataclass-field / class-attribute split are the
reference's (tests mutate `max_motor_rotation` on the instance,
test_solo8v2vanilla.py:109-110).  New fields are appended at the end.
"""
from dataclasses import dataclass
from typing import Tuple
import os

import numpy as np


@dataclass
class Solo8BaseConfig:
  dt: float = 1e-3
  # Max torque supplied by the motors
  motor_torque_limit: float = 2

  robot_start_pos: Tuple[float] = (0., 0., 0.5)
  robot_start_orientation_euler: Tuple[float] = (0., 0., 0.)
  gravity: Tuple[float] = (0., 0., -9.81)

  max_motor_rotation = 2 * np.pi

  linear_damping: float = .04
  angular_damping: float = .04
  # accepted and without effect, as in the reference: Bullet combines restitution as the product
  # of the two bodies' values and the plane's (plane.urdf) is 0
  restitution: float = 0.
  lateral_friction: float = 0.5

  render_width: int = 369
  render_height: int = 369
  render_fov: int = 80
  render_aspect: float = render_width / render_height
  render_pos = [0, 0, .2]
  render_cam_distance = 1
  render_yaw = 0.
  render_pitch = -20.
  render_roll = 0.

  # ---- batched-only additions ------------------------------
  # physics steps of `dt` per env.step (reference: numSubSteps=1, one
  # stepSimulation per step; solo8_base_env.py:40-41)
  num_substeps: int = 1
  # 'position' = p.POSITION_CONTROL (solo8v2vanilla.py:88); 'torque' is new
  control_mode: str = 'position'
  # Bullet world defaults the reference inherits from pybullet
  solver_iterations: int = 50
  max_contacts: int = 56

  @property
  def urdf(self):
    """Reference: pkg_resources path of the URDF.  Here: the compiled model blob
    (gym_solo_b200/assets/compile_model.py output) for that URDF."""
    from gym_solo_b200.assets import blob_for_urdf
    return blob_for_urdf(self.urdf_path)
