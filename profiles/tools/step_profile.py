"""Minimal driver for ncu: N ticks of the fused step on the headline workload (walk, 65,536 envs, random actions).
usage: python profiles/tools/step_profile.py [envs] [workload] [ticks]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200")]
import torch  # noqa: E402

from go2_b200.envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvWalk  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
wl = sys.argv[2] if len(sys.argv) > 2 else "walk"
ticks = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs", wl + ".json")))
cls = {"walk": Go2EnvWalk, "stair": Go2EnvStair}.get(wl, Go2EnvBaseFlavour)
env = cls(B, *cfgs, device="cuda:0", seed=1, stats_sync_interval=24, lib_path=os.environ.get("GO2_B200_LIB") or None)
env.reset()
g = torch.Generator(device="cuda").manual_seed(1234)
acts = [0.5 * torch.randn(B, env.num_actions, device="cuda", generator=g) for _ in range(8)]
for i in range(ticks):
    env.step(acts[i % 8])
torch.cuda.synchronize()
print("done", ticks, "ticks", env.core.launch_shape())
