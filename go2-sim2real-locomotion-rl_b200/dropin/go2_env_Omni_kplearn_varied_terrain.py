"""Drop-in replacement for examples/locomotion/go2_env_Omni_kplearn_varied_terrain.py (the walk env on a TerrainManager terrain, custom
heightmap mode): `from go2_env_Omni_kplearn_varied_terrain import Go2Env`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from go2_b200.envs import CurriculumManager  # noqa: E402,F401
from go2_b200.envs import Go2EnvWalkTerrain as Go2Env  # noqa: E402,F401
