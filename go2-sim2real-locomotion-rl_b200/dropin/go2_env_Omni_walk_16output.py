"""Drop-in replacement for examples/locomotion/go2_env_Omni_walk_16output.py (file-identical to final/go2_env_walk.py)."""
from go2_env_walk import CurriculumManager, Go2Env  # noqa: F401
