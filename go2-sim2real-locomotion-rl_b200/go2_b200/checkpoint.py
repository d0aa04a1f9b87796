"""On-disk formats either side of the hot path.

* `cfgs.pkl` - what the reference's train scripts dump and the eval scripts load
  (`go2_train_walk.py:462-465`, `go2_eval_walk.py:500`): a pickled 5-list
  `[env_cfg, obs_cfg, reward_cfg, command_cfg, train_cfg]` of plain dicts.  Loaded with a restricted unpickler
  (plain containers / numbers / strings only - a cfg file must not be able to execute code).
* rsl-rl 2.2.4 `model_<iter>.pt` (`OnPolicyRunner.save`): a torch file with `model_state_dict`,
  `optimizer_state_dict`, `iter`, `infos`.  `load_actor` rebuilds the actor MLP from the `actor.<2k>.weight/bias` keys.
* env state: an additive `env_state_<iter>.pt` with the device buffers of the native handle plus the host-side
  curriculum / global-DR state, so that a run can resume bit-for-bit (the reference restarts its envs from scratch).
"""
from __future__ import annotations

import io
import pickle

import numpy as np
import torch

_SAFE_BUILTINS = {"dict", "list", "tuple", "set", "frozenset", "str", "int", "float", "bool", "complex", "bytes", "NoneType"}


class _RestrictedUnpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module == "builtins" and name in _SAFE_BUILTINS:
            return getattr(__import__("builtins"), name)
        if module == "collections" and name == "OrderedDict":
            import collections

            return collections.OrderedDict
        raise pickle.UnpicklingError(f"cfgs.pkl may only contain plain containers, not {module}.{name}")


def load_cfgs(path):
    """-> [env_cfg, obs_cfg, reward_cfg, command_cfg, train_cfg] (train_cfg may be missing in older files)."""
    with open(path, "rb") as f:
        obj = _RestrictedUnpickler(io.BytesIO(f.read())).load()
    if not isinstance(obj, (list, tuple)) or len(obj) < 4 or not all(isinstance(c, dict) for c in obj):
        raise ValueError(f"{path}: expected a list of 4-5 cfg dicts")
    return list(obj)


def save_cfgs(path, env_cfg, obs_cfg, reward_cfg, command_cfg, train_cfg=None):
    """Same layout as the reference's `pickle.dump([env_cfg, obs_cfg, reward_cfg, command_cfg, train_cfg], f)`."""
    with open(path, "wb") as f:
        pickle.dump([env_cfg, obs_cfg, reward_cfg, command_cfg, train_cfg or {}], f)


# ---------------------------------------------------------------------------------------------------
# rsl-rl 2.2.4 checkpoints
# ---------------------------------------------------------------------------------------------------
def _mlp_from_state(sd, prefix, activation=torch.nn.ELU):
    idx = sorted({int(k.split(".")[1]) for k in sd if k.startswith(prefix + ".") and k.endswith(".weight")})
    if not idx:
        raise KeyError(f"no '{prefix}.<i>.weight' entries in the checkpoint")
    layers = []
    for n, i in enumerate(idx):
        w, b = sd[f"{prefix}.{i}.weight"], sd[f"{prefix}.{i}.bias"]
        lin = torch.nn.Linear(w.shape[1], w.shape[0])
        with torch.no_grad():
            lin.weight.copy_(w)
            lin.bias.copy_(b)
        layers.append(lin)
        if n + 1 < len(idx):
            layers.append(activation())
    return torch.nn.Sequential(*layers)


def load_runner_checkpoint(path, map_location="cpu"):
    """rsl-rl `OnPolicyRunner.load` layout -> dict(model_state_dict, optimizer_state_dict, iter, infos)."""
    ck = torch.load(path, map_location=map_location, weights_only=True)
    for k in ("model_state_dict",):
        if k not in ck:
            raise KeyError(f"{path}: not an rsl-rl checkpoint (missing {k})")
    return ck


def load_actor(path, device="cpu"):
    """Inference policy of a `model_<iter>.pt`: obs -> action mean (what `runner.get_inference_policy()` evaluates)."""
    ck = load_runner_checkpoint(path)
    return _mlp_from_state(ck["model_state_dict"], "actor").to(device).eval()


def save_runner_checkpoint(path, actor, critic, std, optimizer_state=None, it=0, infos=None):
    """Write a file `OnPolicyRunner.load` accepts: ActorCritic state dict keys `actor.*`, `critic.*`, `std`."""
    sd = {"std": std.detach().cpu()}
    for prefix, net in (("actor", actor), ("critic", critic)):
        for k, v in net.state_dict().items():
            sd[f"{prefix}.{k}"] = v.detach().cpu()
    torch.save({"model_state_dict": sd, "optimizer_state_dict": optimizer_state or {}, "iter": int(it), "infos": infos}, path)


# ---------------------------------------------------------------------------------------------------
# env state (additive)
# ---------------------------------------------------------------------------------------------------
_HOST_ATTRS = ["_step_count", "_action_history_write_idx", "_push_counter", "_curr_ep_total", "_curr_timeout_total", "_curr_tracking_sum",
               "_curr_tracking_n", "_lock_terrain_rows", "_ticks_since_stats"]


def env_state_dict(env):
    """Everything needed to continue `env.step` bit-for-bit: all device buffers (they include the global physics DR in
    force and the statistics accumulator) + the double-buffer parity + host-side scalars."""
    env.flush_stats()
    st = {"buffers": {k: v.detach().cpu().clone() for k, v in env.core.buf.items()},
          "host": {a: getattr(env, a) for a in _HOST_ATTRS if hasattr(env, a)},
          "parity": env.core.tick_parity(),
          "rng": env._np_rng.bit_generator.state}
    if hasattr(env, "curriculum"):
        c = env.curriculum
        st["curriculum"] = {"level": c.level, "ema": dict(c.ema), "_ready_streak": c._ready_streak, "_hard_streak": c._hard_streak,
                            "_cooldown": c._cooldown}
    return st


def load_env_state_dict(env, st):
    for k, v in st["buffers"].items():
        env.core.buf[k].copy_(v.to(env.device))
    if st.get("parity", 0) != env.core.tick_parity():  # rows of the double-buffered tables swap roles with the parity
        for k in ("globals", "stats"):
            env.core.buf[k][:2] = env.core.buf[k][:2].flip(0)
    for a, v in st["host"].items():
        setattr(env, a, v)
    env._np_rng.bit_generator.state = st["rng"]
    if "curriculum" in st and hasattr(env, "curriculum"):
        env.curriculum.__dict__.update(st["curriculum"])
        env._apply_curriculum_level(force=True)
