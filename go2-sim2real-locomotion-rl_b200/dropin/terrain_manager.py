"""Drop-in replacement for examples/locomotion/terrain_manager.py: `from terrain_manager import TerrainManager` (custom heightmap mode)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from go2_b200.envs import TerrainManager  # noqa: E402,F401
