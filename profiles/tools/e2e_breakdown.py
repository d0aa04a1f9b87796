"""Where the host-buffer tick (go2_step_host, bench `e2e`) spends its time beyond the step kernel: host-side Python before / after
the native call vs the native call itself (H2D actions + step + post kernels + stream sync).  usage: python profiles/tools/e2e_breakdown.py"""
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
env = Go2EnvWalk(B, *cfgs, device="cuda:0", seed=1, stats_sync_interval=24)
env.reset()
g = torch.Generator().manual_seed(1)
acts = [(0.5 * torch.randn(B, env.num_actions, generator=g)).pin_memory() for _ in range(4)]
h_obs = torch.empty(B, env.num_obs).pin_memory(); h_rew = torch.empty(B).pin_memory(); h_done = torch.empty(B, dtype=torch.uint8).pin_memory()
orig_make, orig_c = env._make_params, env.core.step_host
T = {"make_params": 0.0, "native_call": 0.0}


def make(*a, **k):
    t0 = time.perf_counter(); r = orig_make(*a, **k); T["make_params"] += time.perf_counter() - t0; return r


def native(*a, **k):
    t0 = time.perf_counter(); r = orig_c(*a, **k); T["native_call"] += time.perf_counter() - t0; return r


env._make_params, env.core.step_host = make, native
for i in range(20):
    env.step_host(acts[i % 4], h_obs, h_rew, h_done)
for k in T:
    T[k] = 0.0
N = 100
t0 = time.perf_counter()
for i in range(N):
    env.step_host(acts[i % 4], h_obs, h_rew, h_done)
tot = time.perf_counter() - t0
# the same ticks device-resident (async; events)
d_acts = [a.cuda() for a in acts]
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(N):
    env.step(d_acts[i % 4])
e1.record(); torch.cuda.synchronize()
print(json.dumps({"ticks": N, "e2e_ms_per_tick": 1e3 * tot / N, "make_params_ms": 1e3 * T["make_params"] / N, "native_call_ms": 1e3 * T["native_call"] / N,
                  "other_python_ms": 1e3 * (tot - T["make_params"] - T["native_call"]) / N, "device_resident_ms_per_tick": e0.elapsed_time(e1) / N}))
