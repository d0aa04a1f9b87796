"""e2e overhead at steady state, legs interleaved on ONE env: device-resident env.step vs go2_step_host with pinned buffers (zero-copy
result streaming) vs go2_step_host with pageable result buffers (staged D2H copies).  usage: python profiles/tools/e2e_legs.py"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200")]
import torch  # noqa: E402

from go2_b200.envs import Go2EnvWalk  # noqa: E402

B = 65536
cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
env = Go2EnvWalk(B, *cfgs, device="cuda:0", seed=1, stats_sync_interval=24, fast_math=True)
env.reset()
g = torch.Generator(device="cuda").manual_seed(1)
env.episode_length_buf = torch.randint(0, int(env.max_episode_length), (B,), device="cuda", generator=g, dtype=torch.int32)
d_acts = [0.5 * torch.randn(B, env.num_actions, device="cuda", generator=g) for _ in range(8)]
h_acts = [a.cpu().pin_memory() for a in d_acts]
pin = lambda *s, dt=torch.float32: torch.empty(*s, dtype=dt).pin_memory()  # noqa: E731
h_obs, h_rew, h_done = pin(B, env.num_obs), pin(B), pin(B, dt=torch.uint8)
p_obs, p_rew, p_done = torch.empty(B, env.num_obs).numpy(), torch.empty(B).numpy(), torch.empty(B, dtype=torch.uint8).numpy()
for i in range(300):
    env.step(d_acts[i % 8])
torch.cuda.synchronize()
K = 48
out = {"device": [], "host_pinned": [], "host_pageable": [], "h2d_only_ms": None}
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        env.step(d_acts[i % 8])
    e1.record(); torch.cuda.synchronize()
    out["device"].append(round(e0.elapsed_time(e1) / K, 4))
    t0 = time.perf_counter()
    for i in range(K):
        env.step_host(h_acts[i % 8], h_obs, h_rew, h_done)
    out["host_pinned"].append(round(1e3 * (time.perf_counter() - t0) / K, 4))
    t0 = time.perf_counter()
    for i in range(K):
        env.step_host(h_acts[i % 8], p_obs, p_rew, p_done)
    out["host_pageable"].append(round(1e3 * (time.perf_counter() - t0) / K, 4))
dst = torch.empty(B, env.num_actions, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(50):
    dst.copy_(h_acts[i % 8], non_blocking=True); torch.cuda.synchronize()
out["h2d_only_ms"] = round(1e3 * (time.perf_counter() - t0) / 50, 4)
print(json.dumps(out))
