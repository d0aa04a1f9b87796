import json, os, sys, numpy as np, torch
sys.path.insert(0, "tests"); sys.path.insert(0, "go2-sim2real-locomotion-rl_b200")
from util import make_env
cfgs = json.load(open("go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json"))
B = 256
def run(lib, T=60):
    env = make_env("walk", cfgs, B, "hostemu", os.path.abspath(lib), seed=1); env.reset()
    g = torch.Generator().manual_seed(0); out = []; nn = []
    for t in range(T):
        o, r, d, _ = env.step(0.5 * torch.randn(B, 16, generator=g))
        out.append((o.clone(), r.clone(), d.clone())); nn.append(env.core.buf["nnewton"].numpy().mean())
    return out, np.array(nn)
a, na = run("tests/host_emul/libgo2_hostemu.so"); b, nb = run("build/libgo2_hostemu_dense.so")
for t in (0, 1, 2, 5, 10, 20, 40, 59):
    print(t, "max|dobs|", float((a[t][0] - b[t][0]).abs().max()), "max|drew|", float((a[t][1] - b[t][1]).abs().max()), "reset mismatches", int((a[t][2] != b[t][2]).sum()))
print("mean newton arrow", na.mean(), "dense", nb.mean())
