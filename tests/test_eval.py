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
