"""Drop-in replacement for examples/locomotion/go2_env_terrain.py (height-map walk env with the ground height in the privileged obs,
ground-relative foot clearance and the stumble reward; TerrainManager in custom heightmap mode): `from go2_env_terrain import Go2Env`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from go2_b200.envs import CurriculumManager  # noqa: E402,F401
from go2_b200.envs import Go2EnvTerrain as Go2Env  # noqa: E402,F401
