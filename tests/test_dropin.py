"""The drop-in boundary: the modules a train / eval script of the reference imports, used the way those
scripts use them (go2_train_walk.py:17-20,382,467-476; go2_train_crouch.py:144-162; go2_eval_walk.py:399-490,610-617,649,
746-790), plus a stand-in for rsl-rl 2.2.4's OnPolicyRunner collection loop that - like PPO.act / process_env_step - keeps
the REFERENCE to the observation tensor across env.step() and only copies it into the storage afterwards."""
import importlib
import json
import os
import sys

import numpy as np
import pytest
import torch

from util import ROOT, np_

DROPIN = os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200", "dropin")
CFG_DIR = os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs")
BACKENDS = ["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)]


def _cfgs(name):
    with open(os.path.join(CFG_DIR, name + ".json")) as f:
        return json.load(f)


@pytest.fixture
def dropin(backend, hostemu_lib, monkeypatch):
    """sys.path as INTEGRATION.md prescribes; on the CPU backend the env's native core is the host emulation of the kernel."""
    monkeypatch.syspath_prepend(DROPIN)
    for m in ("genesis", "genesis.utils", "genesis.utils.geom", "go2_env_walk", "go2_env_stair", "go2_env_base", "go2_env", "go2_env_stair6", "go2_env_stair4", "go2_env_stair3", "go2_env_Omni_walk_16output"):
        sys.modules.pop(m, None)
    gs = importlib.import_module("genesis")
    assert os.path.dirname(gs.__file__).startswith(DROPIN)
    import go2_b200.envs as envs

    monkeypatch.setattr(envs, "DEFAULT_FAST_MATH", [False])  # gs.init() switches the default to the fast-math build: undone after the test
    if backend == "hostemu":
        from hostemu_core import HostEmuCore

        monkeypatch.setattr(envs, "_DEFAULT_CORE", [HostEmuCore])
        monkeypatch.setattr(gs, "device", torch.device("cpu"))
    else:
        gs.init(backend=gs.gpu, precision="32", performance_mode=True)
    yield gs
    for m in ("genesis", "genesis.utils", "genesis.utils.geom", "go2_env_walk", "go2_env_stair", "go2_env_base", "go2_env", "go2_env_stair6", "go2_env_stair4", "go2_env_stair3", "go2_env_Omni_walk_16output"):
        sys.modules.pop(m, None)


def _collect(env, gs, n_iters=2, steps=24, n_act=None):
    """OnPolicyRunner.learn(init_at_random_ep_len=True), collection part."""
    torch.manual_seed(0)
    obs, extras = env.get_observations()
    critic_obs = extras["observations"].get("critic", obs)
    env.episode_length_buf = torch.randint_like(env.episode_length_buf, high=int(env.max_episode_length))
    B, A = env.num_envs, env.num_actions
    storage = torch.zeros(steps, B, env.num_obs, device=gs.device)
    n_done = 0
    for it in range(n_iters):
        for i in range(steps):
            actions = 0.3 * torch.randn(B, A, device=gs.device)
            transition_obs = obs  # PPO.act: self.transition.observations = obs  (by reference)
            snapshot = obs.clone()
            obs, rewards, dones, infos = env.step(actions)
            storage[i].copy_(transition_obs)  # storage.add_transitions, after the step
            assert torch.equal(storage[i], snapshot), "env.step() overwrote the observation tensor rsl-rl still holds"
            assert obs.shape == (B, env.num_obs) and obs.dtype == torch.float32 and obs.device.type == gs.device.type
            assert rewards.shape == (B,) and rewards.dtype == torch.float32
            assert dones.shape == (B,) and dones.dtype == torch.bool
            assert infos["time_outs"].shape == (B,)
            critic_obs = infos["observations"]["critic"]
            n_done += int(dones.sum())
            assert torch.isfinite(obs).all() and torch.isfinite(rewards).all()
    return n_done, critic_obs


@pytest.mark.parametrize("backend", BACKENDS)
def test_walk_module_as_the_train_script_uses_it(backend, dropin):
    gs = dropin
    from go2_env_walk import CurriculumManager, Go2Env

    env_cfg, obs_cfg, reward_cfg, command_cfg = _cfgs("walk")
    scale_before = reward_cfg["reward_scales"]["tracking_lin_vel"]
    env = Go2Env(num_envs=64, env_cfg=env_cfg, obs_cfg=obs_cfg, reward_cfg=reward_cfg, command_cfg=command_cfg)
    assert reward_cfg["reward_scales"]["tracking_lin_vel"] == pytest.approx(scale_before * 0.02)  # the ctor scales the cfg dict in place
    assert (env.num_envs, env.num_obs, env.num_privileged_obs, env.num_actions) == (64, 49, 104, 16)
    assert isinstance(env.max_episode_length, int) and env.device == gs.device
    assert isinstance(env.curriculum, CurriculumManager) and CurriculumManager(env_cfg["curriculum"]).level == env_cfg["curriculum"]["level_init"]
    obs, none = env.reset()
    assert none is None
    n_done, critic = _collect(env, gs)
    assert critic.shape == (64, 104)
    assert n_done > 0  # random episode lengths -> time-outs inside 48 ticks
    assert env.get_privileged_observations() is not None
    assert "episode" in env.extras and all(k.startswith("rew_") for k in env.extras["episode"])


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("module,cfg,n_obs,n_act,n_priv", [("go2_env_stair", "stair", 49, 16, 182), ("go2_env_stair6", "stair", 49, 16, 182), ("go2_env_Omni_walk_16output", "walk", 49, 16, 104), ("go2_env_stair4_lidar", "stair_lidar", 64, 16, 165),
                                                           ("go2_env_stair_lidar", "stair_lidar", 64, 16, 165), ("go2_env_stair_lidar_2", "stair_info", 53, 16, 113), ("go2_env_stair5_info", "stair_info", 53, 16, 113),
                                                           ("go2_env_base", "crouch", 45, 12, None), ("go2_env", "go2", 45, 12, None),
                                                           ("go2_env_base", "jump", 45, 12, None)])
def test_stair_and_base_modules(backend, dropin, module, cfg, n_obs, n_act, n_priv):
    gs = dropin
    Go2Env = importlib.import_module(module).Go2Env
    env_cfg, obs_cfg, reward_cfg, command_cfg = _cfgs(cfg)
    env = Go2Env(num_envs=32, env_cfg=env_cfg, obs_cfg=obs_cfg, reward_cfg=reward_cfg, command_cfg=command_cfg, show_viewer=False)
    assert (env.num_obs, env.num_actions) == (n_obs, n_act)
    env.reset()
    n_done, critic = _collect(env, gs, n_iters=1)
    assert critic.shape == (32, n_priv or n_obs)
    assert (env.get_privileged_observations() is None) == (n_priv is None)


@pytest.mark.parametrize("backend", BACKENDS)
def test_eval_script_surface(backend, dropin):
    """The calls go2_eval_walk.py / go2_eval_stairs.py make on the env between steps."""
    gs = dropin
    from go2_env_walk import Go2Env

    env_cfg, obs_cfg, reward_cfg, command_cfg = _cfgs("walk")
    env = Go2Env(num_envs=1, env_cfg=env_cfg, obs_cfg=obs_cfg, reward_cfg=reward_cfg, command_cfg=command_cfg, show_viewer=False)
    env.ground.set_friction(0.7)
    env.robot.set_friction(0.9)
    env._motor_strength[:] = 0.95
    env._kp_factors[:] = 1.0
    env._kd_factors[:] = 1.0
    base_link_idx = env.robot.links[1].idx
    obs, _ = env.reset()
    zero = torch.zeros((1, env.num_actions), device=gs.device)
    for _ in range(5):
        obs, *_ = env.step(zero)
    obs, _ = env.get_observations()
    # respawn_on_tile (flat ground branch)
    env.robot.set_dofs_position(position=env.default_dof_pos.unsqueeze(0), dofs_idx_local=env.motors_dof_idx, zero_velocity=True, envs_idx=[0])
    env.robot.set_pos(env.base_init_pos.unsqueeze(0), zero_velocity=True, envs_idx=[0])
    env.robot.set_quat(env.base_init_quat.unsqueeze(0), zero_velocity=True, envs_idx=[0])
    env.robot.zero_all_dofs_velocity([0])
    env.last_actions[0] = 0.0
    env.last_dof_vel[0] = 0.0
    env._applied_actions[0] = 0.0
    env._action_history[0] = 0.0
    env.episode_length_buf[0] = 0
    assert int(env.episode_length_buf[0]) == 0
    assert torch.allclose(env.robot.get_pos()[0], env.base_init_pos) and float(env.core.buf["qvel"].abs().max()) == 0.0
    env.commands[:] = torch.tensor([[0.5, 0.0, 0.0]], device=gs.device)
    env.robot.set_mass_shift([2.0], [0])  # dynamic payload
    assert env.global_dr_values()["mass_shift"][0] == 2.0
    x0 = float(env.base_pos[0, 0])
    for step in range(12):
        if step < 6:  # a 6-tick push of 80 N along +y on the link the scripts push (robot.links[1])
            env.scene.sim.rigid_solver.apply_links_external_force(force=torch.tensor([[0.0, 80.0, 0.0]], device=gs.device), links_idx=[base_link_idx], envs_idx=[0])
        obs, _, resets, _ = env.step(zero)
    assert float(env.base_lin_vel[0, 1]) > 0.05 or float(env.base_pos[0, 1]) > 0.005, "the external push had no effect"
    assert float(env.commands[0, 0]) == 0.5
    assert env.pls_kp_default == env_cfg["pls_kp_default"] and env.num_pos_actions == 12 and env.dt == 0.02
    assert float(env.core.buf["ext_force"][0, 0]) == -1.0  # cleared after the tick
    assert abs(float(env.base_pos[0, 0]) - x0) < 0.5
    with pytest.raises(NotImplementedError):
        env.scene.step()
