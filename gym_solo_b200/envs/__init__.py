"""Custom solo environments (gym_solo/envs/__init__.py)."""
from gym_solo_b200.envs.solo8_base_env import Solo8BaseEnv
from gym_solo_b200.envs.solo8v2vanilla import Solo8VanillaEnv, Solo8VanillaConfig, Solo8MMRConfig
