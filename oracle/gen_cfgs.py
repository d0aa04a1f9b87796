"""TEST / PACKAGING INFRASTRUCTURE - regenerates go2_b200/cfgs/*.json, the shipped configuration tables, by executing ONLY the
`get_cfgs()` function of the reference's train scripts (their other imports need rsl-rl / genesis).

    python oracle/gen_cfgs.py            # needs /root/reference; the JSON files are committed
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from gen_golden import FINAL, REF, ROOT, load_funcs  # noqa: E402

OUT = os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs")
SCRIPTS = {
    "walk": os.path.join(FINAL, "go2_train_walk.py"), "stair": os.path.join(FINAL, "go2_train_stair.py"),
    "crouch": os.path.join(FINAL, "go2_train_crouch.py"), "jump": os.path.join(FINAL, "go2_train_jump.py"),
    "stair_lidar": os.path.join(REF, "examples/locomotion/go2_stair_train_lidar.py"),
    "stair_info": os.path.join(REF, "examples/locomotion/go2_stair_train_lidar_2.py"),
    "go2": os.path.join(REF, "examples/locomotion/go2_train.py"),  # Genesis' stock example (go2_env.py == final/go2_env_base.py)
}

if __name__ == "__main__":
    for name in sys.argv[1:] or list(SCRIPTS):
        cfgs = load_funcs(SCRIPTS[name], ["get_cfgs"]).get_cfgs()
        with open(os.path.join(OUT, name + ".json"), "w") as f:
            json.dump(list(cfgs), f)
        print(name, "num_obs", cfgs[1]["num_obs"], "num_privileged_obs", cfgs[1].get("num_privileged_obs"))
