"""Step time as a function of simulated time since reset(): the env population leaves the all-standing start state (random actions
-> falls, resets, more / fewer contacts), so a benchmark window right after reset() is not the steady state.
usage: python profiles/tools/tick_time_drift.py [ticks] [workload] [envs]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200")]
import torch  # noqa: E402

from go2_b200.envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvWalk  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1500
wl = sys.argv[2] if len(sys.argv) > 2 else "walk"
B = int(sys.argv[3]) if len(sys.argv) > 3 else 65536
cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs", wl + ".json")))
cls = {"walk": Go2EnvWalk, "stair": Go2EnvStair}.get(wl, Go2EnvBaseFlavour)
env = cls(B, *cfgs, device="cuda:0", seed=1, stats_sync_interval=24)
env.reset()
g = torch.Generator(device="cuda").manual_seed(1234)
acts = [0.5 * torch.randn(B, env.num_actions, device="cuda", generator=g) for _ in range(8)]
W = 50
rows = []
for w in range(T // W):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    nres = 0
    for i in range(W):
        _, _, d, _ = env.step(acts[i % 8])
    e1.record()
    torch.cuda.synchronize()
    rows.append({"ticks": (w + 1) * W, "ms_per_tick": round(e0.elapsed_time(e1) / W, 4), "mean_contacts": round(float(env.core.buf["ncontact"].float().mean()), 3),
                 "mean_newton": round(float(env.core.buf["nnewton"].float().mean()), 3), "resets_last_tick": int(d.sum()),
                 "mean_ep_len": round(float(env.episode_length_buf.float().mean()), 1), "level": round(float(env.curriculum.level), 3)})
    print(json.dumps(rows[-1]), flush=True)
