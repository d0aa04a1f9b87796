"""Rollout-side networks and storage around the env step.

What rsl-rl 2.2.4's `OnPolicyRunner.learn` does per control tick while collecting (`PPO.act` -> `env.step` ->
`PPO.process_env_step`, `rsl_rl/algorithms/ppo.py`, `rsl_rl/storage/rollout_storage.py`), on the fused tcgen05 MLP kernel of
`csrc/go2_policy.cu` (C ABI: `include/go2_policy.h`): one launch for the actor (mean, Gaussian sample, log-prob written
straight into the storage rows of the tick), one for the critic, then the env's single step launch.

There is no fallback: the classes raise if the CUDA library is missing or the tensors are not on a CUDA device.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import abi


_LIB = None


def _lib():
    global _LIB
    if _LIB is None:  # one dlopen per process
        _LIB = abi.Go2Lib(abi.CUDA_LIB_PATH).lib
        _LIB.go2_mlp_create.restype = C.c_int
    return _LIB


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _linear_layers(module_or_list):
    if isinstance(module_or_list, torch.nn.Module):
        return [(m.weight.detach(), m.bias.detach()) for m in module_or_list.modules() if isinstance(m, torch.nn.Linear)]
    return [(w.detach(), b.detach()) for w, b in module_or_list]


class FusedMLP:
    """ELU MLP (<= 4 Linear layers, hidden <= 512, output <= 16) evaluated by one tcgen05 kernel launch."""

    def __init__(self, layers, device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("go2_b200.rollout: the fused MLP runs on a CUDA device only (no CPU fallback)")
        layers = _linear_layers(layers)
        self.dims = [int(layers[0][0].shape[1])] + [int(w.shape[0]) for w, _ in layers]
        flat = np.concatenate([np.concatenate([w.float().cpu().numpy().ravel(), b.float().cpu().numpy().ravel()]) for w, b in layers]).astype(np.float32)
        self._lib = _lib()
        self._h = C.c_void_p()
        dims = (C.c_int32 * len(self.dims))(*self.dims)
        rc = self._lib.go2_mlp_create(dims, len(layers), flat.ctypes.data_as(C.c_void_p), self.device.index if self.device.index is not None else torch.cuda.current_device(), C.byref(self._h))
        if rc != 0:
            raise RuntimeError(f"go2_mlp_create failed rc={rc} for dims {self.dims}")

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.go2_mlp_destroy(self._h)
        except Exception:
            pass

    def _check(self, x):
        if x.device != self.device or x.dtype != torch.float32 or not x.is_contiguous() or x.shape[-1] != self.dims[0]:
            raise ValueError(f"expected a contiguous f32 [B, {self.dims[0]}] tensor on {self.device}")

    def __call__(self, x, out=None):
        self._check(x)
        B = x.shape[0]
        if out is None:
            out = torch.empty((B, self.dims[-1]), device=self.device, dtype=torch.float32)
        rc = self._lib.go2_mlp_forward(self._h, _ptr(x), B, _ptr(out), _stream(self.device))
        if rc != 0:
            raise RuntimeError(f"go2_mlp_forward failed rc={rc}")
        return out


class FusedActorCritic:
    """`ActorCritic` of rsl-rl 2.2.4 (`actor.*`, `critic.*`, `std` state-dict keys) on the fused kernels."""

    def __init__(self, actor_layers, critic_layers, std, device, seed=0, env_offset=0):
        self.actor = FusedMLP(actor_layers, device)
        self.critic = FusedMLP(critic_layers, device)
        self.device = self.actor.device
        self.std = std.detach().to(self.device, torch.float32).contiguous()
        self.num_actions = self.actor.dims[-1]
        self.seed, self.env_offset = int(seed), int(env_offset)

    @classmethod
    def from_state_dict(cls, sd, device, **kw):
        def layers(prefix):
            idx = sorted({int(k.split(".")[1]) for k in sd if k.startswith(prefix + ".") and k.endswith(".weight")})
            return [(sd[f"{prefix}.{i}.weight"], sd[f"{prefix}.{i}.bias"]) for i in idx]

        return cls(layers("actor"), layers("critic"), sd["std"], device, **kw)

    def act_inference(self, obs, out=None):
        return self.actor(obs, out)

    def evaluate(self, critic_obs, out=None):
        return self.critic(critic_obs, out)

    def act(self, obs, step, actions=None, mu=None, sigma=None, logp=None):
        """actions ~ N(actor(obs), std); every output may be a row of the rollout storage."""
        self.actor._check(obs)
        B, A = obs.shape[0], self.num_actions
        mk = lambda t, shape: t if t is not None else torch.empty(shape, device=self.device, dtype=torch.float32)  # noqa: E731
        actions, mu, sigma, logp = mk(actions, (B, A)), mk(mu, (B, A)), mk(sigma, (B, A)), mk(logp, (B, 1))
        rc = self.actor._lib.go2_policy_act(self.actor._h, _ptr(obs), B, _ptr(self.std), C.c_uint64(self.seed), C.c_uint32(int(step)),
                                            C.c_uint32(self.env_offset), _ptr(actions), _ptr(mu), _ptr(sigma), _ptr(logp), _stream(self.device))
        if rc != 0:
            raise RuntimeError(f"go2_policy_act failed rc={rc}")
        return actions, mu, sigma, logp


    def act_and_evaluate(self, obs, critic_obs, step, actions, mu, sigma, logp, values):
        """`act` + `evaluate` in one launch (actor tiles and critic tiles share the grid)."""
        self.actor._check(obs)
        self.critic._check(critic_obs)
        rc = self.actor._lib.go2_policy_act_critic(self.actor._h, self.critic._h, _ptr(obs), _ptr(critic_obs), obs.shape[0], _ptr(self.std),
                                                   C.c_uint64(self.seed), C.c_uint32(int(step)), C.c_uint32(self.env_offset), _ptr(actions),
                                                   _ptr(mu), _ptr(sigma), _ptr(logp), _ptr(values), _stream(self.device))
        if rc != 0:
            raise RuntimeError(f"go2_policy_act_critic failed rc={rc}")
        return actions, values


def gae(rewards, dones, values, last_values, gamma, lam):
    """[T, B] returns and (un-normalised) advantages; `RolloutStorage.compute_returns` without the normalisation."""
    T, B = rewards.shape[:2]
    ret, adv = torch.empty_like(rewards), torch.empty_like(rewards)
    rc = _lib().go2_gae(_ptr(rewards), _ptr(dones), _ptr(values), _ptr(last_values), T, B, C.c_float(gamma), C.c_float(lam), _ptr(ret), _ptr(adv),
                        _stream(rewards.device))
    if rc != 0:
        raise RuntimeError(f"go2_gae failed rc={rc}")
    return ret, adv


class RolloutCollector:
    """Collects `num_steps` transitions per env into tensors laid out like rsl-rl's `RolloutStorage` ([T, B, ...])."""

    def __init__(self, env, ac: FusedActorCritic, num_steps=24, gamma=0.99, lam=0.95):
        self.env, self.ac, self.T, self.gamma, self.lam = env, ac, int(num_steps), float(gamma), float(lam)
        B, dev, A = env.num_envs, ac.device, ac.num_actions
        f = lambda *s: torch.zeros(s, device=dev, dtype=torch.float32)  # noqa: E731
        self.observations = f(self.T, B, env.num_obs)
        self.privileged_observations = f(self.T, B, ac.critic.dims[0])
        self.actions, self.mu, self.sigma = f(self.T, B, A), f(self.T, B, A), f(self.T, B, A)
        self.rewards, self.values, self.actions_log_prob = f(self.T, B, 1), f(self.T, B, 1), f(self.T, B, 1)
        self.dones = torch.zeros((self.T, B, 1), device=dev, dtype=torch.uint8)
        self.last_values = f(B, 1)
        self.returns = self.advantages = None
        self.tick = 0

    def collect(self, obs=None, critic_obs=None):
        env, ac = self.env, self.ac
        if obs is None:
            obs = env.get_observations()[0]
            critic_obs = env.get_privileged_observations()
            critic_obs = obs if critic_obs is None else critic_obs
        for t in range(self.T):
            self.observations[t].copy_(obs)
            self.privileged_observations[t].copy_(critic_obs)
            ac.act_and_evaluate(obs, critic_obs, self.tick, self.actions[t], self.mu[t], self.sigma[t], self.actions_log_prob[t], self.values[t])
            obs, rew, done, extras = env.step(self.actions[t])
            critic_obs = env.get_privileged_observations()
            critic_obs = obs if critic_obs is None else critic_obs  # base envs: the critic sees the actor's observations
            # PPO.process_env_step: bootstrap on time-outs
            torch.addcmul(rew.unsqueeze(1), self.values[t], extras["time_outs"].unsqueeze(1).to(torch.float32), value=self.gamma, out=self.rewards[t])
            self.dones[t].copy_(done.unsqueeze(1))
            self.tick += 1
        ac.evaluate(critic_obs, self.last_values)
        self.returns, adv = gae(self.rewards, self.dones, self.values, self.last_values, self.gamma, self.lam)
        self.advantages = (adv - adv.mean()) / (adv.std() + 1e-8)
        return obs, critic_obs
