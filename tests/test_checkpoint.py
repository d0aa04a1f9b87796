"""On-disk formats around the hot path: cfgs.pkl, rsl-rl model_<iter>.pt, env-state resume."""
import json
import os
import pickle

import numpy as np
import pytest
import torch

from go2_b200 import checkpoint as ck

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LOGS = "/root/reference/logs"


def test_cfgs_pkl_roundtrip_and_restriction(tmp_path):
    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    p = tmp_path / "cfgs.pkl"
    ck.save_cfgs(p, *cfgs, train_cfg={"algorithm": {"gamma": 0.99}})
    back = ck.load_cfgs(p)
    assert back[:4] == cfgs and back[4]["algorithm"]["gamma"] == 0.99
    evil = tmp_path / "evil.pkl"
    evil.write_bytes(pickle.dumps([os.system, {}, {}, {}]))
    with pytest.raises(pickle.UnpicklingError):
        ck.load_cfgs(evil)


@pytest.mark.skipif(not os.path.isdir(REF_LOGS), reason="reference checkout not present (GPU box)")
@pytest.mark.parametrize("run", ["go2-walk", "go2-stairs", "go2-crouch", "go2-jump"])
def test_reference_cfgs_pkl_load(run):
    cfgs = ck.load_cfgs(os.path.join(REF_LOGS, run, "cfgs.pkl"))
    assert cfgs[0]["num_actions"] in (12, 16) and "num_obs" in cfgs[1] and "reward_scales" in cfgs[2]


def test_runner_checkpoint_roundtrip(tmp_path):
    torch.manual_seed(0)
    mk = lambda i, o: torch.nn.Sequential(torch.nn.Linear(i, 512), torch.nn.ELU(), torch.nn.Linear(512, 256), torch.nn.ELU(),
                                          torch.nn.Linear(256, 128), torch.nn.ELU(), torch.nn.Linear(128, o))
    actor, critic = mk(49, 16), mk(104, 1)
    p = tmp_path / "model_100.pt"
    ck.save_runner_checkpoint(p, actor, critic, torch.ones(16), it=100)
    d = ck.load_runner_checkpoint(p)
    assert d["iter"] == 100 and "actor.0.weight" in d["model_state_dict"] and "std" in d["model_state_dict"]
    x = torch.randn(7, 49)
    assert torch.equal(ck.load_actor(p)(x), actor(x))


def test_env_state_resume_is_bit_exact(hostemu_lib):
    from util import make_env

    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/walk.json")))
    env = make_env("walk", cfgs, 16, "hostemu", hostemu_lib, seed=3)
    env.reset()
    g = torch.Generator().manual_seed(5)
    acts = [0.5 * torch.randn(16, 16, generator=g) for _ in range(12)]
    for a in acts[:6]:
        env.step(a)
    st = ck.env_state_dict(env)
    first = [tuple(t.clone() for t in env.step(a)[:3]) for a in acts[6:]]
    ck.load_env_state_dict(env, st)
    again = [tuple(t.clone() for t in env.step(a)[:3]) for a in acts[6:]]
    for x, y in zip(first, again):
        for u, v in zip(x, y):
            assert torch.equal(u, v)
