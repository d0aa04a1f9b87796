"""Shares of the collision phase's parts from a -DG2_PHASE_CLOCKS build (thread 0 of every block): usage like phase_clocks.py."""
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
ph, sub = (ctypes.c_ulonglong * 8)(), (ctypes.c_ulonglong * 8)()
for i in range(60):
    env.step(acts[i % 8])
lib.go2_phase_clocks(ph); lib.go2_sub_clocks(sub)
for i in range(20):
    env.step(acts[i % 8])
lib.go2_phase_clocks(ph); lib.go2_sub_clocks(sub)
tot, s = sum(list(ph)[:7]), list(sub)
print(f"collision phase = {100 * ph[2] / tot:.1f} % of the tick; of the tick:")
names = {0: "geom poses + boxes", 1: "link-pair box tests", 2: "pair sweep (collect)", 3: "narrow phase (all)", 5: "  terrain: max-level cull", 6: "  terrain: geom extents (6 supports)", 7: "  terrain: strip walk + sphere cull"}
for k, n in names.items():
    print(f"{n:40s} {100 * s[k] / tot:5.1f} %")
print(f"{'  terrain / pair MPR (rest of narrow)':40s} {100 * (s[3] - s[5] - s[6] - s[7]) / tot:5.1f} %")
