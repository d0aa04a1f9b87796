"""Drop-in replacement for examples/locomotion/go2_env.py (Genesis' stock Go2 example env; file-identical to
final/go2_env_base.py): `from go2_env import Go2Env` as go2_train.py / go2_eval.py / go2_backflip.py do."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from go2_b200.envs import Go2EnvBaseFlavour as Go2Env  # noqa: E402,F401
