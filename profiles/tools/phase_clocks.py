"""Per-phase share of the step kernel's time from a -DG2_PHASE_CLOCKS build (SM clock ticks of thread 0 of every block).
usage: GO2_B200_LIB=build/var_clk.so python profiles/tools/phase_clocks.py [envs] [workload]"""
import ctypes, json, os, sys
import torch
sys.path.insert(0, "go2-sim2real-locomotion-rl_b200")
from go2_b200.envs import Go2EnvStair, Go2EnvWalk
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
wl = sys.argv[2] if len(sys.argv) > 2 else "walk"
cfgs = json.load(open(f"go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/{wl}.json"))
env = (Go2EnvStair if wl == "stair" else Go2EnvWalk)(B, *cfgs, device="cuda:0", seed=1, lib_path=os.environ["GO2_B200_LIB"], stats_sync_interval=24)
env.reset()
g = torch.Generator(device="cuda").manual_seed(1234)
acts = [0.5 * torch.randn(B, env.num_actions, device="cuda", generator=g) for _ in range(8)]
lib = ctypes.CDLL(os.environ["GO2_B200_LIB"])
out = (ctypes.c_ulonglong * 8)()
for i in range(30):
    env.step(acts[i % 8])
lib.go2_phase_clocks(out)
for i in range(20):
    env.step(acts[i % 8])
lib.go2_phase_clocks(out)
v = list(out)[:7]
names = ["load + FK + env_pre", "smooth dynamics (CRBA, LtDL, RNE) x2", "collision detection x2", "constraint assembly x2", "Newton solve x2",
         "integrate + FK x2", "env_post + store"]
tot = sum(v)
for n, x in zip(names, v):
    print(f"{n:40s} {100 * x / tot:5.1f} %")
