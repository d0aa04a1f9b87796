"""Drop-in replacement for examples/locomotion/go2_env_stair_lidar.py (imported as `go2_env_stair4_lidar` by
go2_stair_train_lidar.py:46): `from go2_env_stair4_lidar import Go2Env`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from go2_b200.envs import CurriculumManager, LidarScanConfig  # noqa: E402,F401
from go2_b200.envs import Go2EnvStairLidar as Go2Env  # noqa: E402,F401
from go2_b200.terrain import build_stair_terrain  # noqa: E402,F401
