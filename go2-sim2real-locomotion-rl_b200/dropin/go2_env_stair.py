"""Drop-in replacement for examples/locomotion/final/go2_env_stair.py: `from go2_env_stair import Go2Env`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from go2_b200.envs import CurriculumManager  # noqa: E402,F401
from go2_b200.envs import Go2EnvStair as Go2Env  # noqa: E402,F401
from go2_b200.terrain import build_stair_terrain  # noqa: E402,F401
