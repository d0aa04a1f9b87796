"""Fused tcgen05 actor / critic kernels against a plain PyTorch fp32 reference of the same ops.

Tolerance: the kernel multiplies bf16 operands (8-bit mantissa) with fp32 accumulation, the reference is fp32 throughout;
for the 49-512-256-128-16 ELU actor with unit-scale inputs the outputs agree within 3e-2 + 3e-2 |x| (stated here, checked
below); the sampled noise, log-prob and GAE are fp32 and agree within 1e-4.
"""
import json
import math
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def mlp(dims, seed):
    torch.manual_seed(seed)
    layers = []
    for i in range(len(dims) - 1):
        layers.append(torch.nn.Linear(dims[i], dims[i + 1]))
        if i + 2 < len(dims):
            layers.append(torch.nn.ELU())
    return torch.nn.Sequential(*layers).cuda().eval()


@pytest.mark.parametrize("dims,B", [([49, 512, 256, 128, 16], 300), ([104, 512, 256, 128, 1], 128), ([45, 512, 256, 128, 12], 1000),
                                     ([49, 400, 200, 100, 16], 200), ([182, 300, 272, 16], 130),
                                    ([182, 512, 256, 128, 1], 4096), ([49, 128, 16], 77)])
def test_fused_mlp_matches_torch_fp32(dims, B):
    from go2_b200.rollout import FusedMLP

    net = mlp(dims, 3)
    fused = FusedMLP(net, "cuda:0")
    x = torch.randn(B, dims[0], device="cuda")
    with torch.no_grad():
        ref = net(x)
    out = fused(x)
    torch.cuda.synchronize()
    err = (out - ref).abs()
    assert torch.isfinite(out).all()
    assert (err <= 3e-2 + 3e-2 * ref.abs()).all(), float(err.max())
    # and against the same arithmetic in torch (bf16 operands, fp32 accumulate): tight
    with torch.no_grad():
        h = x
        lin = [m for m in net if isinstance(m, torch.nn.Linear)]
        for i, m in enumerate(lin):
            h = (h.bfloat16().float() @ m.weight.bfloat16().float().T) + m.bias
            if i + 1 < len(lin):
                h = torch.nn.functional.elu(h)
    assert (out - h).abs().max() < 2e-3, float((out - h).abs().max())


def test_policy_act_sampling_and_logprob():
    from go2_b200.rollout import FusedActorCritic

    actor, critic = mlp([49, 512, 256, 128, 16], 1), mlp([104, 512, 256, 128, 1], 2)
    std = torch.linspace(0.3, 1.2, 16)
    ac = FusedActorCritic(actor, critic, std, "cuda:0", seed=7)
    B = 8192
    obs = torch.randn(B, 49, device="cuda")
    a, mu, sigma, logp = ac.act(obs, step=5)
    a2, mu2, _, _ = ac.act(obs, step=5)
    a3, _, _, _ = ac.act(obs, step=6)
    torch.cuda.synchronize()
    assert torch.equal(a, a2) and torch.equal(mu, mu2) and not torch.equal(a, a3)  # counter-based: same key, same draw
    assert torch.equal(mu, ac.act_inference(obs))
    assert torch.allclose(sigma, std.cuda().expand(B, 16))
    z = (a - mu) / sigma
    assert abs(float(z.mean())) < 0.02 and abs(float(z.std()) - 1.0) < 0.02
    ref_lp = torch.distributions.Normal(mu, sigma).log_prob(a).sum(-1, keepdim=True)
    assert torch.allclose(logp, ref_lp, atol=2e-3, rtol=1e-4)
    v = ac.evaluate(torch.randn(B, 104, device="cuda"))
    assert v.shape == (B, 1) and torch.isfinite(v).all()


def test_gae_matches_rsl_rl_recurrence():
    from go2_b200.rollout import gae

    T, B = 24, 1000
    g = torch.Generator(device="cuda").manual_seed(0)
    rew = torch.randn(T, B, 1, device="cuda", generator=g)
    val = torch.randn(T, B, 1, device="cuda", generator=g)
    last = torch.randn(B, 1, device="cuda", generator=g)
    done = (torch.rand(T, B, 1, device="cuda", generator=g) < 0.1).to(torch.uint8)
    ret, adv = gae(rew, done, val, last, 0.99, 0.95)
    # rsl_rl/storage/rollout_storage.py compute_returns
    ref = torch.zeros_like(rew)
    a = 0
    for t in reversed(range(T)):
        nv = last if t == T - 1 else val[t + 1]
        nt = 1.0 - done[t].float()
        delta = rew[t] + nt * 0.99 * nv - val[t]
        a = delta + nt * 0.99 * 0.95 * a
        ref[t] = a + val[t]
    assert torch.allclose(ret, ref, atol=1e-4, rtol=1e-5)
    assert torch.allclose(adv, ref - val, atol=1e-4, rtol=1e-5)


def test_rollout_collector_runs_on_the_env():
    from go2_b200.envs import Go2EnvWalk
    from go2_b200.rollout import FusedActorCritic, RolloutCollector

    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    env = Go2EnvWalk(256, *cfgs, device="cuda:0", seed=1)
    env.reset()
    ac = FusedActorCritic(mlp([49, 512, 256, 128, 16], 1), mlp([104, 512, 256, 128, 1], 2), 0.5 * torch.ones(16), "cuda:0", seed=3)
    col = RolloutCollector(env, ac, num_steps=24)
    col.collect()
    torch.cuda.synchronize()
    assert col.returns.shape == (24, 256, 1) and torch.isfinite(col.returns).all() and torch.isfinite(col.advantages).all()
    assert abs(float(col.advantages.mean())) < 1e-3
    assert torch.isfinite(col.actions).all() and float(col.actions.abs().max()) < 10.0
    assert (col.observations[1] != col.observations[0]).any()


def test_policy_noise_is_shard_invariant():
    """Two ranks holding envs [0, B/2) and [B/2, B) draw the same exploration noise as one rank holding all B envs
    (the generator is keyed by the GLOBAL env id, like the env kernel's)."""
    from go2_b200.rollout import FusedActorCritic

    actor, critic = mlp([49, 512, 256, 128, 16], 1), mlp([104, 512, 256, 128, 1], 2)
    std = 0.7 * torch.ones(16)
    B = 512
    obs = torch.randn(B, 49, device="cuda")
    whole = FusedActorCritic(actor, critic, std, "cuda:0", seed=11).act(obs, step=3)[0]
    lo = FusedActorCritic(actor, critic, std, "cuda:0", seed=11, env_offset=0).act(obs[: B // 2].contiguous(), step=3)[0]
    hi = FusedActorCritic(actor, critic, std, "cuda:0", seed=11, env_offset=B // 2).act(obs[B // 2:].contiguous(), step=3)[0]
    torch.cuda.synchronize()
    assert torch.equal(whole, torch.cat([lo, hi]))


def test_actor_and_critic_in_one_launch_match_the_separate_calls():
    from go2_b200.rollout import FusedActorCritic

    ac = FusedActorCritic(mlp([49, 512, 256, 128, 16], 1), mlp([104, 512, 256, 128, 1], 2), 0.5 * torch.ones(16), "cuda:0", seed=5)
    B = 1000
    obs, cobs = torch.randn(B, 49, device="cuda"), torch.randn(B, 104, device="cuda")
    a, mu, sg, lp = ac.act(obs, step=9)
    v = ac.evaluate(cobs)
    a2, mu2, sg2, lp2, v2 = (torch.empty_like(x) for x in (a, mu, sg, lp, v))
    ac.act_and_evaluate(obs, cobs, 9, a2, mu2, sg2, lp2, v2)
    torch.cuda.synchronize()
    for x, y in ((a, a2), (mu, mu2), (sg, sg2), (lp, lp2), (v, v2)):
        assert torch.equal(x, y)
