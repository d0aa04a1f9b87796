"""TEST INFRASTRUCTURE ONLY - a minimal `genesis` look-alike backed by the CPU oracle.

It exists so that the UNMODIFIED reference env classes (examples/locomotion/final/go2_env_*.py) can be
executed in the build container to generate golden vectors (oracle/gen_golden.py).  It mirrors only the
attribute surface those files touch; it is never imported by the product.
"""
import os
import sys
import types

import numpy as np
import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
from oracle.rigid import OracleRigid  # noqa: E402

from .utils import geom as _geom  # noqa: F401,E402

device = torch.device("cpu")
tc_float = torch.float32
tc_int = torch.int32
gpu = "gpu"
cpu = "cpu"
EPS = 1.1920929e-07
ORACLE_KIND = os.environ.get("GO2_ORACLE_KIND", "f32")


def init(*args, **kwargs):
    return None


class _Logger:
    def info(self, *a, **k):
        pass

    warning = debug = error = info


logger = _Logger()
constraint_solver = types.SimpleNamespace(Newton="Newton", CG="CG")


class _Opt:
    def __init__(self, **kw):
        self.__dict__.update(kw)


options = types.SimpleNamespace(SimOptions=_Opt, ViewerOptions=_Opt, VisOptions=_Opt, RigidOptions=_Opt)


class _URDF(_Opt):
    pass


class _Terrain(_Opt):
    pass


morphs = types.SimpleNamespace(URDF=_URDF, Terrain=_Terrain)


def _np(x):
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy().astype(np.float64)
    return np.asarray(x, dtype=np.float64)


def _idx(x):
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy().astype(np.int32)
    return np.asarray(x, dtype=np.int32)


def _t(a, dtype=torch.float32):
    return torch.as_tensor(np.asarray(a), dtype=dtype)


class _Link:
    def __init__(self, name, idx):
        self.name, self.idx = name, idx


class _Joint:
    def __init__(self, name, dof_start):
        self.name, self.dof_start = name, dof_start


class _Ground:
    def __init__(self, scene):
        self._scene = scene
        self.links = [_Link("planeLink", 0)]

    def set_friction(self, mu):
        self._scene._rigid.set_friction(0, mu)


class _Robot:
    """Entity-local indices: link i <-> global link i + 1 (the ground entity owns global link 0)."""

    def __init__(self, scene, model):
        self._scene = scene
        self.links = [_Link(e["name"], i + 1) for i, e in enumerate(model["links"])]
        self._joints = {e["joint_name"]: _Joint(e["joint_name"], e["dof_start"]) for e in model["links"]}

    @property
    def _r(self):
        return self._scene._rigid

    def get_joint(self, name):
        return self._joints[name]

    def set_dofs_kp(self, kp, dofs):
        self._r.set_dofs_kp(_np(kp), dofs)

    def set_dofs_kv(self, kv, dofs):
        self._r.set_dofs_kv(_np(kv), dofs)

    def control_dofs_force(self, force, dofs):
        self._r.control_dofs_force(_np(force), dofs)

    def control_dofs_position(self, pos, dofs):
        self._r.control_dofs_position(_np(pos), dofs)

    def set_dofs_position(self, position, dofs_idx_local=None, zero_velocity=True, envs_idx=None):
        self._r.set_dofs_position(_np(position), dofs_idx_local, _idx(envs_idx), zero_velocity)

    def set_pos(self, pos, zero_velocity=True, envs_idx=None):
        self._r.set_base_pos(_np(pos), _idx(envs_idx), zero_velocity)

    def set_quat(self, quat, zero_velocity=True, envs_idx=None):
        self._r.set_base_quat(_np(quat), _idx(envs_idx), zero_velocity)

    def zero_all_dofs_velocity(self, envs_idx=None):
        self._r.zero_all_dofs_velocity(_idx(envs_idx))

    def get_pos(self):
        return _t(self._r.get_links_pos()[:, 1])

    def get_quat(self):
        return _t(self._r.get_links_quat()[:, 1])

    def get_vel(self):
        return _t(self._r.get_links_vel()[:, 1])

    def get_ang(self):
        return _t(self._r.get_links_ang()[:, 1])

    def get_dofs_position(self, dofs=None):
        p = self._r.get_dofs_position()
        return _t(p if dofs is None else p[:, list(dofs)])

    def get_dofs_velocity(self, dofs=None):
        v = self._r.get_qvel()
        return _t(v if dofs is None else v[:, list(dofs)])

    def get_dofs_control_force(self, dofs=None):
        f = self._r.get_dofs_control_force()
        return _t(f if dofs is None else f[:, list(dofs)])

    def get_links_net_contact_force(self):
        return _t(self._r.get_links_net_contact_force()[:, 1:])

    def get_links_vel(self):
        return _t(self._r.get_links_vel()[:, 1:])

    def get_links_pos(self):
        return _t(self._r.get_links_pos()[:, 1:])

    def set_friction(self, mu):
        self._r.set_friction(1, mu)

    def set_mass_shift(self, shifts, links_idx_local):
        for s, l in zip(shifts, links_idx_local):
            self._r.set_mass_shift(int(l) + 1, float(s))

    def set_COM_shift(self, shifts, links_idx_local):
        for s, l in zip(shifts, links_idx_local):
            self._r.set_com_shift(int(l) + 1, _np(s))


class _RigidSolverView:
    def __init__(self, scene):
        self._scene = scene

    def apply_links_external_force(self, force, links_idx, envs_idx=None):
        f = _np(force)
        if f.ndim == 3:
            f = f[:, 0]
        self._scene._rigid.apply_link_external_force(f, int(links_idx[0]))


class Scene:
    def __init__(self, sim_options=None, rigid_options=None, **kw):
        self._substeps = getattr(sim_options, "substeps", 1)
        self._entities = []
        self._rigid = None
        self._model = None
        self._robot_morph = None
        self._terrain_morph = None
        self.sim = types.SimpleNamespace(rigid_solver=_RigidSolverView(self))

    def add_entity(self, morph, **kw):
        from go2_b200.model import load_model

        if isinstance(morph, _Terrain):  # heightfield terrain in place of the plane (go2_env_stair.py:423-433)
            self._terrain_morph = morph
            ent = _Ground(self)
        elif "plane" in morph.file:
            ent = _Ground(self)
        else:
            self._model = load_model()
            self._robot_morph = morph
            ent = _Robot(self, self._model)
        self._entities.append(ent)
        return ent

    def build(self, n_envs=1, **kw):
        m = self._robot_morph
        self._rigid = OracleRigid(n_envs, self._model, kind=ORACLE_KIND, base_pos=tuple(np.asarray(m.pos, dtype=float)),
                                  base_quat=tuple(np.asarray(m.quat, dtype=float)))
        t = self._terrain_morph
        if t is not None:
            if getattr(t, "height_field", None) is None:  # sub-terrain grid: the restatement of genesis/utils/terrain.py parse_terrain
                from go2_b200.terrain import build_subterrain_grid  # (pinned against the reference generators by oracle/ref_terrain_tiles.py)

                t.generated_field = build_subterrain_grid(t.subterrain_types, t.subterrain_size, t.horizontal_scale, t.vertical_scale,
                                                          getattr(t, "randomize", False), getattr(t, "subterrain_parameters", None))
                hf = np.asarray(t.generated_field, dtype=np.float64) * float(t.vertical_scale)
            else:
                hf = np.asarray(t.height_field, dtype=np.float64) * float(t.vertical_scale)
            self._rigid.set_terrain(hf, float(t.horizontal_scale), np.asarray(t.pos, dtype=np.float64))

    def step(self):
        self._rigid.step(self._substeps)
        err = self._rigid.get_errno()
        if err.any():
            raise RuntimeError(f"oracle errno {np.bitwise_or.reduce(err)}")
