import json, os, sys, ctypes, numpy as np, torch
sys.path.insert(0, "tests"); sys.path.insert(0, "go2-sim2real-locomotion-rl_b200")
from util import make_env
lib = os.path.abspath("build/libgo2_hostemu_stats.so")
cfgs = json.load(open("go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json"))
B = int(os.environ.get("SB", 448))
env = make_env("walk", cfgs, B, "hostemu", lib, seed=1)
env.reset()
g = torch.Generator().manual_seed(0)
nn_hist = np.zeros(64, int); nc_hist = np.zeros(40, int); mx = []
for t in range(int(os.environ.get("ST", 120))):
    env.step(0.5 * torch.randn(B, 16, generator=g))
    if t >= int(os.environ.get("S0", 40)):
        nn = env.core.buf["nnewton"].numpy().ravel(); nc = env.core.buf["ncontact"].numpy().ravel()
        nn_hist += np.bincount(nn, minlength=64)[:64]; nc_hist += np.bincount(nc, minlength=40)[:40]
        mx.append((nn.max(), nn.reshape(-1, 32).max(1).mean(), nn.mean()))
print("newton hist", nn_hist[:30]); print("contacts hist", nc_hist[:20])
mx = np.array(mx); print("block max, warp-max mean, mean:", mx.mean(0))
L = ctypes.CDLL(lib); out = (ctypes.c_long * 210)(); L.go2_ls_stats(out); o = np.array(out[:])
print("ls calls", o[0]); print("A", o[1:65]); print("B", o[65:129]); print("end", o[129:133]); print("nfl", o[133:150]); print("refac over/fail/total", o[197:200]); print("pair kinds ground/base-leg/same-leg/two-legs", o[200:204])
