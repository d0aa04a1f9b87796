"""Drop-in replacement for examples/locomotion/go2_env_stair_lidar_2.py (its own header names it go2_env_stair5_info.py; imported by
go2_stair_train_lidar_2.py:45): `from go2_env_stair_lidar_2 import Go2Env`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from go2_b200.envs import CurriculumManager, StairInfoConfig  # noqa: E402,F401
from go2_b200.envs import Go2EnvStairInfo as Go2Env  # noqa: E402,F401
from go2_b200.terrain import build_stair_terrain_variable_depth as build_stair_terrain  # noqa: E402,F401
