"""Parity with domain randomisation ON: golden vectors produced by the UNMODIFIED reference env classes with their
shipped configurations (oracle/gen_golden_dr.py: walk 32 envs x 1000 ticks = BASELINE.json configs[0]; walk at
curriculum level 1; stair; base env with every reward term), every random draw of the reference recorded and
replayed here.  Per tick the env is teacher-forced from the fixture's pre-state and must reproduce

  * bit-exactly: reset flags, time-outs, episode lengths, the commands (hence the command-resample mask), foot-contact
    flags, every per-episode DR buffer a reset rewrites (kp / kd / motor-strength factors, gravity offset, delay,
    terrain row), push state, and the global DR values (friction with its reset-count throttle, mass / COM / leg shifts);
  * within |d| <= 1e-4 + 1e-3 |x|: observations incl. noise, privileged observations, the reward and EACH
    reward term, episode sums, qpos, qvel.

Env-ticks outside the float band are not budgeted: they are counted, the worst one is printed, and the count must
stay within MAX_OUTSIDE (measured, per fixture; 0 where the measurement is 0).
"""
import numpy as np
import pytest
import torch

from util import Derived, close, force_state_dr, load_golden, make_env, np_, pre

ATOL, RTOL = 1e-4, 1e-3
BACKENDS = ["hostemu", pytest.param("cuda", marks=pytest.mark.gpu)]
# the library build with approximate division / sqrt (the reference's own fast_math arithmetic class, what `bench.py` and the drop-in
# `genesis.init()` select): same code, so the same fixtures - but an operation that is 1-2 ulp off can flip a borderline decision (a
# foot-contact flag at the 3 N threshold, a contact mode), after which that env-tick is off by far more than the band.  Measured on
# a B200: 1 such env-tick in walk_c1's 32,000, 2 in walk_hard's 4,800 (profiles/r02_fast_math_parity.txt).  Budget: 1 in 2,000.
FAST_EXTRA = lambda n: max(2, n // 2000)  # noqa: E731
# fixture, env kind, env-ticks allowed outside the float band (contact-mode flips at the solver tolerance)
# walk_terrain_c1 (walk env on a custom height map, smoothed +-8 cm bumps): 3.7 %, contacts on prism edges again
# stair_lidar_c1 (go2_stair_train_lidar cfg: 20 cm risers, 25 cm treads): 7.5 % for the same reason; its lidar outputs (actor scan in the
# obs, critic scan in the privileged obs, the kept scan) differ in 1 of 4,800 env-ticks
# stair_info_c1 (go2_stair_train_lidar_2 cfg, spawn 0.25 m before the first 20 cm riser so that the edge detector fires from tick 0: the
# robots spend the fixture ON the riser edges): 9.4 %, all of them qvel / qpos (the stair-info outputs - noisy copy in the obs, clean
# copy in the privileged obs, the kept copy - are inside the band on all 4,800 env-ticks)
# stair: 4.9 % of the env-ticks (hostemu) sit on rounded step-edge contacts whose MPR output is discontinuous at rounding level
# (test_oracle_contact_output_is_discontinuous_at_rounding_level); every bit-exact output still matches on all of them
CASES = [("walk_c1", "walk", 0), ("walk_hard", "walk", 0), ("base_allrew", "base", 0), ("stair_c1", "stair", 400), ("stair_lidar_c1", "stair_lidar", 420), ("stair_info_c1", "stair_info", 480), ("walk_terrain_c1", "walk_terrain", 240), ("walk_grid_c1", "walk_terrain", 240), ("terrain_env_c1", "terrain_env", 240)]
EXACT_AFTER = ["kp_factors", "kd_factors", "motor_strength", "grav_off", "delay", "commands", "ep_len"]


def _run(name, kind, backend, hostemu_lib, stride=1):
    g, cfgs = load_golden(name)
    T, B = g["actions"].shape[:2]
    env = make_env(kind, cfgs, B, backend, hostemu_lib, replay=True, export_reward_terms=True)
    env.reset()
    der = Derived(g, kind, cfgs[0]["clip_actions"])
    der.start(getattr(env, "_delay_buf_size", 1))
    if kind == "base":
        der.ep_sums[:] = 0.0
    bad_exact, bad_float, worst = [], [], (0.0, None)
    for t in range(T):
        if t % stride == 0:
            force_state_dr(env, g, t, kind, der)
            obs, rew, done, extras = env.step(torch.as_tensor(g["actions"][t], device=env.device))
            b = env.core.buf
            # ---- exact
            ex = {"reset": (np_(done).astype(np.uint8), g["o_reset"][t]), "timeout": (np_(extras["time_outs"]), g["o_timeout"][t])}
            for k in EXACT_AFTER:
                if k in g:
                    ex[k] = (np_(b[k]).reshape(pre(g, k, t + 1).shape), pre(g, k, t + 1))
            if kind in ("walk", "stair", "stair_lidar", "stair_info", "walk_terrain", "terrain_env"):
                ex["foot_contact"] = (np_(env._foot_contact), g["o_foot_contact"][t])
                if "push_force" in g:
                    ex["push_force"] = (np_(b["push_force"]), pre(g, "push_force", t + 1))
                    ex["push_remain"] = (np_(b["push_remain"]).reshape(-1), pre(g, "push_remain", t + 1))
                gv = env.global_dr_values()
                ex["g_friction"] = (np.float32(gv["friction_ground"]), np.float32(pre(g, "g_friction", t + 1)))
                ex["g_mass"] = (np.float32(gv["mass_shift"][0]), np.float32(pre(g, "g_mass", t + 1)))
                ex["g_com"] = (np.asarray(gv["com_shift"], np.float32), pre(g, "g_com", t + 1))
                ex["g_counter"] = (np.float32(gv["reset_counter"]), np.float32(pre(g, "g_counter", t + 1)))
            if kind in ("walk_terrain", "terrain_env"):
                ex["spawn_pos_xy"] = (np_(b["spawn_pos"])[:, :2], pre(g, "spawn_pos", t + 1)[:, :2])  # (u * 2 - 1) * half: exact in f32
            if kind in ("stair_lidar", "stair_info"):  # the scan kept for the next tick's "latency" replay (noise branches decided by replayed uniforms)
                fl_scan = (np_(b["actor_scan"]), pre(g, "actor_scan", t + 1))
            if kind in ("stair", "stair_lidar", "stair_info"):
                ex["terrain_row"] = (np_(b["terrain_row"]).reshape(-1), pre(g, "terrain_row", t + 1))
            for k, (a, r) in ex.items():
                if not np.array_equal(np.asarray(a), np.asarray(r)):
                    bad_exact.append((name, t, k, np.nonzero(np.atleast_1d(np.asarray(a) != np.asarray(r)).reshape(B, -1).any(1))[0].tolist()
                                      if np.asarray(a).shape[:1] == (B,) else "global"))
            # ---- floats
            same = np_(done).astype(np.uint8) == g["o_reset"][t]
            fl = {"obs": (np_(obs), g["o_obs"][t]), "rew": (np_(rew), g["o_rew"][t]), "rew_terms": (np_(env.reward_terms), g["o_rew_terms"][t]),
                  "qpos": (np_(b["qpos"]), pre(g, "qpos", t + 1)), "qvel": (np_(b["qvel"]), pre(g, "qvel", t + 1))}
            if kind in ("walk", "stair", "stair_lidar", "stair_info", "walk_terrain", "terrain_env"):
                fl["priv"] = (np_(extras["observations"]["critic"])[:, env.num_obs:], g["o_priv_tail"][t])
                fl["air_time"] = (np_(b["air_time"]), pre(g, "air_time", t + 1))
            if kind in ("stair_lidar", "stair_info"):
                fl["actor_scan"] = fl_scan
            out = np.zeros(B, bool)
            for k, (a, r) in fl.items():
                a2, r2 = np.asarray(a, np.float64).reshape(B, -1), np.asarray(r, np.float64).reshape(B, -1)
                ok = close(a2, r2, ATOL, RTOL).all(1)
                out |= ~ok
                err = np.abs(a2 - r2) / (ATOL + RTOL * np.abs(r2))
                if err.max() > worst[0]:
                    e = int(err.max(1).argmax())
                    worst = (float(err.max()), (name, t, e, k, float(np.abs(a2 - r2)[e].max())))
            for e in np.nonzero(out & same)[0]:
                bad_float.append((name, t, int(e)))
        der.advance(g, t)
    return bad_exact, bad_float, worst, T * B // stride


def _drop_force_derived(bad_exact, bad_float, max_outside):
    """On the terrain fixtures (the ones with a float budget: contacts on prism edges, where MPR's output is discontinuous at rounding
    level) an env-tick that is already counted outside the float band has different contact forces, so its foot-contact flag - the
    3 N threshold on those forces - may differ too.  Such a flag mismatch belongs to the same counted env-tick; anything else stays exact."""
    if max_outside == 0:
        return bad_exact
    out = {(t, e) for _, t, e in bad_float}
    return [m for m in bad_exact if not (m[2] == "foot_contact" and all((m[1], e) in out for e in m[3]))]


@pytest.mark.gpu
@pytest.mark.parametrize("name,kind,max_outside", CASES)
def test_teacher_forced_fast_math_build(name, kind, max_outside, hostemu_lib):
    bad_exact, bad_float, worst, n = _run(name, kind, "cuda_fast", hostemu_lib)
    print(f"\n[{name}/cuda_fast] {n} env-ticks, {len(bad_exact)} exact mismatches {bad_exact[:4]}, {len(bad_float)} outside the band, worst {worst}")
    assert len(bad_exact) <= FAST_EXTRA(n), bad_exact[:10]
    assert len(bad_float) <= max_outside + FAST_EXTRA(n), (len(bad_float), worst)


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("name,kind,max_outside", CASES)
def test_teacher_forced_with_replayed_draws(name, kind, max_outside, backend, hostemu_lib):
    stride = 4 if (backend == "hostemu" and name == "walk_c1") else 1  # keep the CPU suite short; the GPU run checks every tick
    bad_exact, bad_float, worst, n = _run(name, kind, backend, hostemu_lib, stride)
    print(f"\n[{name}/{backend}] {n} env-ticks, worst float error = {worst[0]:.3f} x band at {worst[1]}, outside: {bad_float[:20]}")
    bad_exact = _drop_force_derived(bad_exact, bad_float, max_outside)
    assert not bad_exact, f"bit-exact outputs differ (fixture, tick, buffer, envs): {bad_exact[:10]}"
    assert len(bad_float) <= max_outside, f"{len(bad_float)} env-ticks outside |d| <= {ATOL} + {RTOL}|x|: {bad_float[:20]}; worst {worst}"


def _free_run(name, kind, backend, hostemu_lib, horizon):
    """Start from the fixture's state at tick 0 and run `horizon` ticks on the fixture's actions with the reference's draws
    replayed; per env: first tick outside the float band (qpos, qvel, reward) and first integer mismatch (-1 = never)."""
    g, cfgs = load_golden(name)
    T, B = g["actions"].shape[:2]
    env = make_env(kind, cfgs, B, backend, hostemu_lib, replay=True)
    env.reset()
    der = Derived(g, kind, cfgs[0]["clip_actions"])
    der.start(getattr(env, "_delay_buf_size", 1))
    force_state_dr(env, g, 0, kind, der)
    first_out, first_int = np.full(B, -1), np.full(B, -1)
    for t in range(horizon):
        if t > 0:
            env.replay_draws(g["replay"][t])
            if kind != "base":
                env._cand_override = {"friction": float(pre(g, "g_friction", t + 1)), "mass": float(pre(g, "g_mass", t + 1)),
                                      "com": [float(x) for x in pre(g, "g_com", t + 1)], "leg": [float(x) for x in pre(g, "g_leg", t + 1)]}
        obs, rew, done, extras = env.step(torch.as_tensor(g["actions"][t], device=env.device))
        b = env.core.buf
        ok = close(np_(b["qpos"]), pre(g, "qpos", t + 1), ATOL, RTOL).all(1) & close(np_(b["qvel"]), pre(g, "qvel", t + 1), ATOL, RTOL).all(1)
        ok &= close(np_(rew), g["o_rew"][t], ATOL, RTOL)
        iok = (np_(done).astype(np.uint8) == g["o_reset"][t]) & (np_(b["commands"]) == pre(g, "commands", t + 1)).all(1)
        iok &= (np_(b["ep_len"]).reshape(-1) == pre(g, "ep_len", t + 1)) & (np_(extras["time_outs"]) == g["o_timeout"][t])
        first_out[(first_out < 0) & ~ok] = t
        first_int[(first_int < 0) & ~iok] = t
    return first_out, first_int


@pytest.mark.parametrize("backend", BACKENDS)
def test_free_running_200_ticks_with_replayed_draws(backend, hostemu_lib):
    """SURVEY.md 8(d) parity acceptance on BASELINE configs[0] (walk, 32 envs, full DR): free-running (no teacher forcing)
    for the first 200 ticks.  Both fp32 implementations are chaotic past contact-mode switches, so an env may leave the
    float band; it must not happen to more than a few, and the integer outputs (reset flags, time-outs, episode lengths,
    commands = resample mask) must stay bit-exact for every env that is still inside the band."""
    first_out, first_int = _free_run("walk_c1", "walk", backend, hostemu_lib, 200)
    B = len(first_out)
    inside = int((first_out < 0).sum())
    print(f"\n[walk_c1 free-running/{backend}] envs inside the float band for all 200 ticks: {inside}/{B}; "
          f"first tick outside: {sorted(first_out[first_out >= 0].tolist())}; first integer mismatch: {sorted(first_int[first_int >= 0].tolist())}")
    assert inside >= (28 if backend == "hostemu" else 20), "too many envs left the float band within 200 free-running ticks"
    # an integer output may only differ after the env's trajectory has left the band
    for e in range(B):
        if first_int[e] >= 0:
            assert first_out[e] >= 0 and first_out[e] <= first_int[e], f"env {e}: integer mismatch at tick {first_int[e]} while inside the float band"


def test_oracle_contact_output_is_discontinuous_at_rounding_level():
    """Why stair env-ticks can leave the float band (MAX_OUTSIDE of stair_c1): MPR stops when the portal is within
    CCD_TOLERANCE = 1e-6 of the support plane; on the 22 mm foot sphere resting on a step EDGE that is ~1e-2 rad of contact
    normal, and which iteration reaches the tolerance flips with the last bit of the input.  Shown on the ORACLE ALONE:
    perturbing qpos by one ulp changes a sizeable share of its own terrain contact normals by more than 1e-4."""
    from go2_b200.terrain import build_stair_terrain
    from oracle.rigid import OracleRigid

    g, cfgs = load_golden("stair_c1")
    B = g["actions"].shape[1]
    hf, info = build_stair_terrain(cfgs[0]["terrain"])

    def mk():
        o = OracleRigid(B, kind="f32")
        o.set_terrain(hf.astype(np.float64) * info["vertical_scale"], info["horizontal_scale"], np.asarray(info["terrain_origin"], dtype=np.float64))
        return o

    oa, ob = mk(), mk()
    rng = np.random.default_rng(0)
    tot = big = 0
    for t in range(14, 34):
        q = g["qpos"][t].astype(np.float32)
        q2 = np.nextafter(q, q + np.where(rng.random(q.shape) < 0.5, -1, 1).astype(np.float32)).astype(np.float32)
        for o, qq in ((oa, q), (ob, q2)):
            o.set_qpos(qq)
            o.set_qvel(g["qvel"][t])
            o.step(1)
        for e in range(B):
            (_, da), (_, db) = oa.get_contacts(e), ob.get_contacts(e)
            if len(da) == len(db):
                tot += len(da)
                big += int((np.abs(da[:, 3:6] - db[:, 3:6]).max(1) > 1e-4).sum()) if len(da) else 0
    print(f"\noracle vs oracle(+1 ulp): {big}/{tot} terrain contact normals differ by more than 1e-4")
    assert tot > 500 and big > tot // 20
