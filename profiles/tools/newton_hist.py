"""Distribution of Newton iterations / contacts per env (last substep of a tick) and of their per-warp and per-block maxima - the
phase-locked block runs the Newton loop max-over-block times.  usage: python profiles/tools/newton_hist.py [envs] [workload]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200")]
import torch  # noqa: E402

from go2_b200.envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvWalk  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
wl = sys.argv[2] if len(sys.argv) > 2 else "walk"
cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs", wl + ".json")))
cls = {"walk": Go2EnvWalk, "stair": Go2EnvStair}.get(wl, Go2EnvBaseFlavour)
env = cls(B, *cfgs, device="cuda:0", seed=1, stats_sync_interval=24)
env.reset()
g = torch.Generator(device="cuda").manual_seed(1234)
tpb = env.core.launch_shape()["threads"]
out = {"launch": env.core.launch_shape()}
hist = torch.zeros(64, device="cuda")
wmax, bmax, cw, cb = [], [], [], []
for i in range(60):
    env.step(0.5 * torch.randn(B, env.num_actions, device="cuda", generator=g))
    if i >= 20:
        nn = env.core.buf["nnewton"].view(-1).float()
        nc = env.core.buf["ncontact"].view(-1).float()
        hist += torch.bincount(nn.long().clamp(max=63), minlength=64).float()
        pad = (-B) % tpb
        nb = torch.nn.functional.pad(nn, (0, pad)).view(-1, tpb)
        wmax.append(float(nn.view(-1, 32).max(1).values.mean())); bmax.append(float(nb.max(1).values.mean()))
        cw.append(float(nc.view(-1, 32).max(1).values.mean())); cb.append(float(torch.nn.functional.pad(nc, (0, pad)).view(-1, tpb).max(1).values.mean()))
h = (hist / hist.sum()).tolist()
out.update(mean_iters=sum(i * p for i, p in enumerate(h)), hist={i: round(p, 5) for i, p in enumerate(h) if p > 0},
           mean_warp_max_iters=sum(wmax) / len(wmax), mean_block_max_iters=sum(bmax) / len(bmax),
           mean_contacts=float(env.core.buf["ncontact"].float().mean()), mean_warp_max_contacts=sum(cw) / len(cw), mean_block_max_contacts=sum(cb) / len(cb))
print(json.dumps(out))
