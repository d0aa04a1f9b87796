"""Drop-in replacement for examples/locomotion/go2_env_stair4.py - the same env as final/go2_env_stair.py (stair6: file-identical; stair4:
docstrings only; stair3: without the orientation_roll_only reward term, which a cfg simply does not name)."""
from go2_env_stair import CurriculumManager, Go2Env, build_stair_terrain  # noqa: F401
