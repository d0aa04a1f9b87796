"""CPU-side checks: the C-ABI library loads and exports every symbol include/go2_b200.h declares (no compute
calls), the model table matches known Go2 facts, and the oracle satisfies physical invariants."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    from go2_b200 import abi

    hdr = open(os.path.join(ROOT, "include/go2_b200.h")).read()
    declared = set(re.findall(r"\b(go2_[a-z_]+)\s*\(", hdr))
    assert declared == set(abi._EXPORTS), declared ^ set(abi._EXPORTS)
    if not os.path.exists(abi.CUDA_LIB_PATH):
        pytest.skip("libgo2_b200.so not built (run __graft_entry__.build())")
    lib = abi.Go2Lib(abi.CUDA_LIB_PATH)  # dlopen + symbol + struct-size checks; no CUDA call is made
    assert lib.stats_len == 36
    hdr2 = open(os.path.join(ROOT, "include/go2_policy.h")).read()
    declared2 = set(re.findall(r"\b(go2_[a-z_]+)\s*\(", hdr2))
    assert declared2 == set(abi._POLICY_EXPORTS), declared2 ^ set(abi._POLICY_EXPORTS)
    for name in abi._POLICY_EXPORTS:
        assert hasattr(lib.lib, name), name
    fast = abi.Go2Lib(abi.CUDA_LIB_FAST_PATH)  # the fast-math build of the same sources exports the same ABI
    assert fast.stats_len == lib.stats_len and all(hasattr(fast.lib, n) for n in abi._EXPORTS + abi._POLICY_EXPORTS)


def test_missing_library_fails_loudly(tmp_path):
    from go2_b200 import abi

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        abi.Go2Lib(str(tmp_path / "nope.so"))


def test_cpu_device_is_refused():
    from go2_b200 import abi
    from go2_b200.core import Go2Core

    with pytest.raises(RuntimeError, match="CUDA devices only"):
        Go2Core(abi.EnvConfig(), device="cpu")


def test_model_table():
    from go2_b200.model import load_model

    m = load_model()
    assert [l["name"] for l in m["links"]] == ["base", "FL_hip", "FR_hip", "RL_hip", "RR_hip", "FL_thigh", "FR_thigh", "RL_thigh",
                                               "RR_thigh", "FL_calf", "FR_calf", "RL_calf", "RR_calf"]
    assert abs(m["total_mass"] - 15.019) < 1e-9
    assert m["n_geoms"] == 28 and m["geoms"][0]["type"] == 5 and m["geoms"][0]["data"] == [200.0, 200.0, 10.0]
    by = {l["joint_name"]: l["dof_start"] for l in m["links"]}
    names = ["FR_hip_joint", "FR_thigh_joint", "FR_calf_joint", "FL_hip_joint", "FL_thigh_joint", "FL_calf_joint",
             "RR_hip_joint", "RR_thigh_joint", "RR_calf_joint", "RL_hip_joint", "RL_thigh_joint", "RL_calf_joint"]
    assert [by[n] for n in names] == [7, 11, 15, 6, 10, 14, 9, 13, 17, 8, 12, 16]
    calf = m["links"][9]
    assert calf["force_range"] == [-35.55, 35.55] and abs(calf["limit"][1] + 0.83776) < 1e-9


@pytest.mark.parametrize("kind", ["f32", "f64"])
def test_oracle_static_stance(kind):
    from oracle.rigid import OracleRigid

    o = OracleRigid(2, kind=kind)
    mdi = [7, 11, 15, 6, 10, 14, 9, 13, 17, 8, 12, 16]
    q0 = np.array([0, 0.8, -1.5, 0, 0.8, -1.5, 0, 1.0, -1.5, 0, 1.0, -1.5])
    o.set_dofs_position(np.tile(q0, (2, 1)), mdi)
    o.set_dofs_kp([60.0] * 12, mdi)
    o.set_dofs_kv([2.0] * 12, mdi)
    for _ in range(150):
        o.control_dofs_position(np.tile(q0, (2, 1)), mdi)
        o.step(2)
    fz = o.get_links_net_contact_force()[0, 1:, 2].sum()
    assert abs(fz - 15.019 * 9.81) < 1.5  # static stance: sum Fz = m g
    M = o.get_mass_matrix(0)
    assert np.allclose(M, M.T) and np.linalg.eigvalsh(M).min() > 0
    assert o.get_errno().max() == 0
    g, d = o.get_contacts(0)
    assert set(g[:, 1].tolist()) == {0} and set(g[:, 0].tolist()) == {15, 19, 23, 27}  # the four foot spheres on the ground box


def test_oracle_free_fall():
    from oracle.rigid import OracleRigid

    o = OracleRigid(1, kind="f64", base_pos=(0, 0, 3.0))
    mdi = [7, 11, 15, 6, 10, 14, 9, 13, 17, 8, 12, 16]
    q0 = np.array([0, 0.8, -1.5, 0, 0.8, -1.5, 0, 1.0, -1.5, 0, 1.0, -1.5])
    o.set_dofs_position(q0[None], mdi)  # inside the joint limits: no internal forces while falling
    for _ in range(10):
        o.step(2)
    t = 10 * 0.02
    # semi-implicit Euler free fall: z = z0 - g h^2 n(n+1)/2
    n, h = 20, 0.01
    assert abs(o.get_qpos()[0, 2] - (3.0 - 9.81 * h * h * n * (n + 1) / 2)) < 1e-9
    assert abs(o.get_qvel()[0, 2] + 9.81 * t) < 1e-9


STAIR_HF_SHA256 = "9ccdb2c405cef8d826eebb46dd0f6b022728a6bca4618b9597040ba77463bc84"  # reference build_stair_terrain output (int16, C order)


def test_stair_terrain_matches_reference_fixture():
    """build_stair_terrain restated in go2_b200/terrain.py vs the digest of the reference's own function
    (go2_env_stair.py:47-185, executed in the build container on go2_train_stair.get_cfgs()['terrain'])."""
    import hashlib
    import json

    from go2_b200.terrain import build_stair_terrain

    cfgs = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/stair.json")))
    hf, info = build_stair_terrain(cfgs[0]["terrain"])
    assert hf.shape == (704, 1560) and hf.dtype == np.int16
    assert int(hf.astype(np.int64).sum()) == 49645440 and int(hf.max()) == 180
    assert hashlib.sha256(hf.tobytes()).hexdigest() == STAIR_HF_SHA256
    assert abs(info["row_centers"][0][0] - 1.0) < 1e-9 and abs(info["terrain_origin"][1] + 39.0) < 1e-9
