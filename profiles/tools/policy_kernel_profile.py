import sys, os, torch
sys.path.insert(0, "go2-sim2real-locomotion-rl_b200")
from go2_b200.rollout import FusedMLP
torch.manual_seed(0)
dims = [49, 512, 256, 128, 16]
L = []
for i in range(4):
    L.append(torch.nn.Linear(dims[i], dims[i + 1]))
    if i < 3: L.append(torch.nn.ELU())
net = torch.nn.Sequential(*L).cuda()
f = FusedMLP(net, "cuda:0")
x = torch.randn(65536, 49, device="cuda")
out = torch.empty(65536, 16, device="cuda")
for _ in range(5): f(x, out)
torch.cuda.synchronize()
