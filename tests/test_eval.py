"""Headless eval loop on the host-emulated kernel: zero-action policy stands, metrics are sane."""
import json
import os

import torch

from go2_b200 import eval as g2eval

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_headless_eval_zero_policy_stands(hostemu_lib):
    from util import make_env

    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    env = make_env("walk", cfgs, 8, "hostemu", hostemu_lib, seed=11)
    zeros = torch.zeros(8, env.num_actions)
    res = g2eval.evaluate(env, lambda obs: zeros, 40, command_schedule=[(0, [0.0, 0.0, 0.0]), (20, [0.3, 0.0, 0.0])])
    assert res["envs"] == 8 and res["steps"] == 40
    assert res["falls"] == 0 and 0.2 < res["mean_base_height"] < 0.45
    assert torch.allclose(env.commands[:, 0], torch.full((8,), 0.3))
    assert res["mean_lin_vel_error"] < 0.5 and abs(res["mean_forward_distance"]) < 0.2


def test_eval_scenario_pushes_payload_friction_and_respawn(hostemu_lib):
    """EvalScenario = the disturbance knobs of go2_eval_walk.py:31-112, batched: each one must reach the physics."""
    from go2_b200 import abi
    from util import make_env

    cfgs = g2eval.eval_cfgs(json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json"))))
    assert "push_force_range" not in cfgs[0] and cfgs[0]["curriculum"]["enabled"] is False and cfgs[2]["reward_scales"] == {}
    assert cfgs[1]["num_privileged_obs"] is None and cfgs[0]["termination_if_roll_greater_than"] == 1e9
    B = 6
    zeros = torch.zeros(B, 16)  # walk cfg: 12 position + 4 stiffness actions

    def run(scn, steps=30, seed=5):
        env = make_env("walk", cfgs, B, "hostemu", hostemu_lib, seed=seed)
        res = g2eval.evaluate(env, lambda obs: zeros, steps, [(0, [0.0, 0.0, 0.0])], scenario=scn)
        return env, res

    env0, r0 = run(None)
    # friction + payload + delay line: the globals carry the friction, the payload lowers the standing height
    scn = g2eval.EvalScenario(ground_friction=0.4, robot_friction=0.6, payload_mass=6.0, action_delay_steps=2, force_link=0)
    env1, r1 = run(scn)
    gl = env1.global_dr_values()
    assert abs(gl["friction_ground"] - 0.4) < 1e-6 and abs(gl["friction_robot"] - 0.6) < 1e-6
    assert r1["mean_base_height"] < r0["mean_base_height"] - 1e-3 and r1["falls"] == 0 and r1["respawns"] == 0
    assert len(scn.delay_line) == 3
    # pushes: 100 N on the base for 3 ticks every 10 ticks moves the base sideways; different directions per env
    scn = g2eval.EvalScenario(push=dict(force_range=(100.0, 100.0), duration_s=(0.06, 0.06), interval_s=0.2), force_link=0, seed=3)
    env2, r2 = run(scn)
    d = (env2.base_pos[:, :2] - env0.base_pos[:, :2]).norm(dim=1)
    assert float(d.min()) > 5e-3 and float((env2.base_pos[:, :2] - env2.base_pos[0, :2]).norm(dim=1).max()) > 1e-3
    # a dynamic payload changes the mass shift of the base at the given step (go2_eval_walk.py:697-707)
    scn = g2eval.EvalScenario(dynamic_payload=(10, 3.0), force_link=0)
    env3, r3 = run(scn, steps=20)
    assert abs(float(env3.core.buf["globals"][env3.core.tick_parity(), abi.G_MASS_SHIFT]) - 3.0) < 1e-6
    assert abs(float(scn.const_force[0, 2]) + 3.0 * 9.81) < 1e-4
    # obs noise / gravity offset act on the policy input only, action noise reaches the env
    seen = []

    def spy(obs):
        seen.append(obs.clone())
        return zeros

    scn = g2eval.EvalScenario(obs_noise_level=1.0, gravity_obs_offset=[0.5, 0.0, 0.0], action_noise_std=0.05)
    env4 = make_env("walk", cfgs, B, "hostemu", hostemu_lib, seed=5)
    g2eval.evaluate(env4, spy, 5, [(0, [0.0, 0.0, 0.0])], scenario=scn)
    assert float((seen[-1][:, 3] - env4.obs_buf[:, 3]).mean()) > 0.2 and float(seen[-1][:, 21:33].std()) > 0.1
    assert float(seen[-1][:, 6:9].abs().max()) == 0.0  # commands carry no noise (go2_eval_walk.py:64)
    assert float(env4.actions.abs().max()) > 0.0
    # respawn of fallen envs: drop two envs below the fall height, they come back at the spawn pose
    scn = g2eval.EvalScenario(respawn_on_fall=True, fall_height=0.15)
    env5 = make_env("walk", cfgs, B, "hostemu", hostemu_lib, seed=5)
    env5.reset()
    scn.setup(env5)
    env5.core.buf["qpos"][torch.tensor([1, 4]), 2] = 0.05
    env5.base_pos[torch.tensor([1, 4]), 2] = 0.05
    assert scn.after_step(env5) == 2 and scn.respawns == 2
    assert torch.allclose(env5.core.buf["qpos"][torch.tensor([1, 4]), 2], env5.base_init_pos[2].expand(2))


def test_step_rejects_a_short_action_tensor(hostemu_lib):
    """The kernel reads num_actions floats per env; a [B, 12] tensor on the 16-action walk env must not be read past its end."""
    import pytest
    from util import make_env

    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    env = make_env("walk", cfgs, 4, "hostemu", hostemu_lib, seed=1)
    env.reset()
    with pytest.raises(ValueError, match="actions must be"):
        env.step(torch.zeros(4, 12))


def test_respawn_on_a_terrain_row_of_the_stair_env(hostemu_lib):
    """respawn() = respawn_at_start of go2_eval_stairs.py:314-361: row centre + base height, default pose, cleared action buffers."""
    from util import make_env

    cfgs = g2eval.eval_cfgs(json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/stair.json"))))
    env = make_env("stair", cfgs, 4, "hostemu", hostemu_lib, seed=2)
    env._lock_terrain_rows = True
    env.reset()
    a = 0.3 * torch.ones(4, env.num_actions)
    for _ in range(5):
        env.step(a)
    assert float(env.last_actions.abs().max()) > 0
    g2eval.respawn(env, [1, 3], terrain_row=5)
    cx, cy, cz = env._terrain_row_centers[5]
    q = env.core.buf["qpos"]
    for e in (1, 3):
        assert int(env._env_terrain_row[e]) == 5
        assert torch.allclose(q[e, :3], torch.tensor([cx, cy, cz + float(env.base_init_pos[2])]), atol=1e-6)
        assert torch.allclose(q[e, 3:7], env.base_init_quat) and float(env.core.buf["qvel"][e].abs().max()) == 0.0
        assert float(env.last_actions[e].abs().max()) == 0.0 and float(env._action_history[e].abs().max()) == 0.0
        assert int(env.episode_length_buf[e]) == 0 and abs(float(env._last_base_pos_x[e]) - cx) < 1e-6
    assert float(env.last_actions[0].abs().max()) > 0  # the other envs are untouched
    obs, rew, done, _ = env.step(torch.zeros(4, env.num_actions))
    assert torch.isfinite(obs).all() and not bool(done[1]) and not bool(done[3])
