"""Host-side state machine and facade pieces of the boundary checked against the reference / the oracle:
  * CurriculumManager: step for step against the reference's own class (go2_env_walk.py:42-151), imported from /root/reference
    when the checkout is present (build container), for a long pseudo-random metric stream;
  * apply_links_external_force: one tick with an external force against the oracle's kernel_apply_links_external_force
    restatement (abd/misc.py:655-715);
  * robot facade: set_pos / set_quat / set_dofs_position / zero_all_dofs_velocity + get_links_pos / get_links_vel
    (go2_links_kernel) against the oracle's setters / getters."""
import ast
import json
import os
import types

import numpy as np
import pytest
import torch

from util import ROOT, close, load_golden, make_env, np_

REF_ENV = "/root/reference/examples/locomotion/final/go2_env_walk.py"
BACKENDS = ["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)]
MDI = [7, 11, 15, 6, 10, 14, 9, 13, 17, 8, 12, 16]


def _reference_curriculum_class():
    tree = ast.parse(open(REF_ENV).read())
    keep = [n for n in tree.body if (isinstance(n, ast.ClassDef) and n.name == "CurriculumManager")
            or (isinstance(n, ast.FunctionDef) and n.name in ("_clamp01", "_lerp", "_lerp_range"))]
    ns = {"torch": torch}
    exec(compile(ast.Module(body=keep, type_ignores=[]), REF_ENV, "exec"), ns)
    return ns["CurriculumManager"]


@pytest.mark.skipif(not os.path.exists(REF_ENV), reason="reference checkout not present (GPU box)")
@pytest.mark.parametrize("cfg_name", ["walk", "stair"])
def test_curriculum_manager_matches_reference_class(cfg_name):
    from go2_b200.envs import CurriculumManager

    cfg = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs", cfg_name + ".json")))[0]["curriculum"]
    Ref = _reference_curriculum_class()
    for variant in (dict(cfg), dict(cfg, level_init=0.5, step_up=0.05, step_down=0.07, cooldown_updates=2, ready_streak=2), {}, {"enabled": True}):
        ref, mine = Ref(dict(variant)), CurriculumManager(dict(variant))
        assert mine.enabled == ref.enabled and mine.level == ref.level
        rng = np.random.default_rng(3)
        changes = 0
        for i in range(4000):
            phase = (i // 400) % 3  # good stretch, bad stretch, mixed
            tr = float(np.clip(rng.normal([0.9, 0.3, 0.7][phase], 0.1), 0, 1))
            tk = float(np.clip(rng.normal([0.9, 0.4, 0.7][phase], 0.1), 0, 2))
            a, b = ref.update(tr, tk, 1.0 - tr), mine.update(tr, tk, 1.0 - tr)
            assert bool(a) == bool(b), i
            assert mine.level == ref.level, i
            changes += int(bool(a))
            sr, sm = ref.state_dict(), mine.state_dict()
            assert sr.keys() == sm.keys()
            for k in sr:
                assert sr[k] == sm[k] or (sr[k] is not None and sr[k] == pytest.approx(sm[k], rel=0, abs=0)), (i, k)
        if variant.get("enabled"):
            assert changes > 5
        # sample_level: same support and mixture weights (the reference draws from torch's global stream)
        lv = np.array([mine.sample_level() for _ in range(4000)])
        torch.manual_seed(0)
        lr = np.array([ref.sample_level() for _ in range(4000)])
        # same mixture: P(current level) and the uniform part's range [mix_level_low, min(level, mix_level_high)]
        assert abs((lv == mine.level).mean() - (lr == ref.level).mean()) < 0.05
        uv, ur = lv[lv != mine.level], lr[lr != ref.level]
        if len(uv) > 50 and len(ur) > 50:
            assert abs(uv.mean() - ur.mean()) < 0.05 and abs(uv.max() - ur.max()) < 0.02 and abs(uv.min() - ur.min()) < 0.02


@pytest.mark.parametrize("backend", BACKENDS)
def test_external_link_force_matches_oracle(backend, hostemu_lib):
    from oracle.rigid import OracleRigid

    _, cfgs = load_golden("walk_det")
    B = 6
    env = make_env("walk", cfgs, B, backend, hostemu_lib)
    env.reset()
    q0 = np_(env.core.buf["qpos"]).astype(np.float64)
    o = OracleRigid(B, kind="f32")
    o.set_qpos(q0)
    o.set_qvel(np.zeros((B, 18)))
    force = np.array([[40.0, 0, 0], [0, -60.0, 0], [0, 0, 30.0], [25.0, 25.0, 0], [0, 0, 0], [-80.0, 10.0, 5.0]])
    link_global = env.robot.links[1].idx
    for tick in range(3):
        env.scene.sim.rigid_solver.apply_links_external_force(force=torch.as_tensor(force, dtype=torch.float32, device=env.device),
                                                              links_idx=[link_global], envs_idx=torch.arange(B, device=env.device))
        env.step(torch.zeros(B, 16, device=env.device))
        # oracle: PLS torque law of the env at zero action (go2_env_walk.py:969-979,1007-1019), force held over both substeps
        kp = 40.0
        dq = np.array(list(json.loads(json.dumps(cfgs[0]))["default_joint_angles"][n] for n in cfgs[0]["joint_names"]))
        tq = kp * (dq - o.get_qpos()[:, [d + 1 for d in MDI]]) - 0.2 * np.sqrt(kp) * o.get_qvel()[:, MDI]
        lim = np.array(cfgs[0].get("torque_limits", [23.7] * 12))
        o.control_dofs_force(np.clip(tq, -lim, lim), MDI)
        o.apply_link_external_force(force, link_global)
        o.step(2)
        o.lib.orc_clear_external_force(o.h) if hasattr(o.lib, "orc_clear_external_force") else None
        assert close(np_(env.core.buf["qvel"]), o.get_qvel(), 2e-4, 2e-3).all(), tick
        assert close(np_(env.core.buf["qpos"]), o.get_qpos(), 1e-4, 1e-3).all(), tick
    assert abs(np_(env.core.buf["qvel"])[0, 0]) > 1e-3  # the force did something


@pytest.mark.parametrize("backend", BACKENDS)
def test_robot_facade_setters_and_link_getters_match_oracle(backend, hostemu_lib):
    from oracle.rigid import OracleRigid

    _, cfgs = load_golden("walk_det")
    B = 5
    env = make_env("walk", cfgs, B, backend, hostemu_lib)
    env.reset()
    rng = np.random.default_rng(1)
    o = OracleRigid(B, kind="f32")
    pos = np.array([0, 0, 0.5]) + 0.2 * rng.standard_normal((B, 3))
    quat = rng.standard_normal((B, 4))
    quat /= np.linalg.norm(quat, axis=1, keepdims=True)
    dofs = 0.4 * rng.standard_normal((B, 12)) + np.array([0, 0.8, -1.5] * 4)
    idx = [0, 2, 3]
    t = lambda a: torch.as_tensor(a, dtype=torch.float32, device=env.device)  # noqa: E731
    env.robot.set_dofs_position(position=t(dofs[idx]), dofs_idx_local=MDI, zero_velocity=True, envs_idx=idx)
    env.robot.set_pos(t(pos[idx]), zero_velocity=False, envs_idx=idx)
    env.robot.set_quat(t(quat[idx]), zero_velocity=False, envs_idx=idx)
    o.set_qpos(np_(env.core.buf["qpos"]).astype(np.float64))  # envs 1, 4 keep the reset state
    ref = OracleRigid(B, kind="f32")
    ref.set_qpos(np_(env.core.buf["qpos"]).astype(np.float64) * 0 + o.get_qpos())
    o2 = OracleRigid(B, kind="f32")
    q_reset = np.tile(o.get_qpos()[1], (B, 1))
    o2.set_qpos(q_reset)
    o2.set_dofs_position(dofs[idx], MDI, envs=idx, zero_velocity=True)
    o2.set_base_pos(pos[idx], envs=idx, zero_velocity=False)
    o2.set_base_quat(quat[idx], envs=idx, zero_velocity=False)
    assert np.allclose(np_(env.core.buf["qpos"]), o2.get_qpos(), atol=1e-6)
    assert np.allclose(np_(env.robot.get_pos()), o2.get_links_pos()[:, 1], atol=1e-6)
    assert np.allclose(np_(env.robot.get_dofs_position(MDI)), o2.get_qpos()[:, [d + 1 for d in MDI]], atol=1e-6)
    # velocities + link getters (go2_links_kernel): link-origin positions and velocities of all 13 robot links
    qvel = 0.5 * rng.standard_normal((B, 18))
    env.core.buf["qvel"].copy_(t(qvel))
    o2.set_qvel(qvel)
    assert close(np_(env.robot.get_links_pos()), o2.get_links_pos()[:, 1:], 1e-5, 1e-5).all()
    assert close(np_(env.robot.get_links_vel()), o2.get_links_vel()[:, 1:], 1e-5, 1e-4).all()
    assert close(np_(env.robot.get_vel()), o2.get_links_vel()[:, 1], 1e-5, 1e-4).all()
    env.robot.zero_all_dofs_velocity([1])
    assert float(env.core.buf["qvel"][1].abs().max()) == 0.0 and float(env.core.buf["qvel"][0].abs().max()) > 0.0
    assert np_(env.core.buf["fk_fresh"]).reshape(-1).tolist() == [1, 1, 1, 1, 1]
