"""TEST INFRASTRUCTURE ONLY - ctypes binding of the CPU oracle (oracle/go2_oracle.cpp).

`OracleRigid` exposes the Genesis accessor surface the Go2 envs use,
batched over B envs, numpy in / numpy out.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.join(os.path.dirname(_HERE), "go2-sim2real-locomotion-rl_b200")
if _PKG not in sys.path:
    sys.path.insert(0, _PKG)
from go2_b200.model import flatten_model, load_model  # noqa: E402

NL, NV, NQ = 14, 18, 19
_LIBS = {"f32": "libgo2oracle_f32.so", "f64": "libgo2oracle_f64.so", "f32_fast": "libgo2oracle_f32_fast.so",
         "f32_fast_nc": "libgo2oracle_f32_fast_nc.so"}


def build(force=False):
    out = os.path.join(_HERE, "_build")
    if force or not all(os.path.exists(os.path.join(out, n)) for n in _LIBS.values()):
        subprocess.check_call(["make", "-C", _HERE, "-j4"], stdout=subprocess.DEVNULL)
    return out


def _load(kind):
    lib = C.CDLL(os.path.join(build(), _LIBS[kind]))
    lib.orc_create.restype = C.c_void_p
    return lib


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleRigid:
    def __init__(self, num_envs, model=None, kind="f32", base_pos=(0.0, 0.0, 0.42), base_quat=(1.0, 0, 0, 0), n_threads=1):
        self.model = model or load_model()
        self.B = int(num_envs)
        self.lib = _load(kind)
        iv, fv = flatten_model(self.model)
        qpos0 = np.zeros(NQ)
        qpos0[:3] = base_pos
        qpos0[3:7] = base_quat
        self.h = C.c_void_p(self.lib.orc_create(_p(iv), _p(fv), self.B, _p(qpos0), int(n_threads)))
        self.npairs = len(self.model["pairs"])

    def __del__(self):
        try:
            self.lib.orc_destroy(self.h)
        except Exception:
            pass

    # ------------------------------------------------------------------ stepping / control
    def step(self, substeps=2):
        self.lib.orc_step(self.h, int(substeps))

    def set_terrain(self, hf_m, h_scale, origin):
        """Heightfield terrain (metres, [rows][cols]) instead of the ground plane."""
        hf = _d(hf_m)
        self.lib.orc_set_terrain(self.h, _p(hf), int(hf.shape[0]), int(hf.shape[1]), C.c_double(float(h_scale)), _p(_d(origin)))

    def control_dofs_force(self, force, dofs):
        f, d = _d(force), _i(dofs)
        self.lib.orc_control_dofs_force(self.h, _p(d), len(d), _p(f))

    def control_dofs_position(self, pos, dofs):
        f, d = _d(pos), _i(dofs)
        self.lib.orc_control_dofs_position(self.h, _p(d), len(d), _p(f))

    def set_dofs_kp(self, kp, dofs):
        f, d = _d(kp), _i(dofs)
        self.lib.orc_set_dofs_kp(self.h, _p(d), len(d), _p(f))

    def set_dofs_kv(self, kv, dofs):
        f, d = _d(kv), _i(dofs)
        self.lib.orc_set_dofs_kv(self.h, _p(d), len(d), _p(f))

    def apply_link_external_force(self, force, link):
        f = _d(force)
        assert f.shape == (self.B, 3)
        self.lib.orc_apply_link_external_force(self.h, int(link), _p(f))

    def set_friction(self, entity, mu):
        self.lib.orc_set_friction(self.h, int(entity), C.c_double(float(mu)))

    def set_mass_shift(self, link, shift):
        self.lib.orc_set_mass_shift(self.h, int(link), C.c_double(float(shift)))

    def set_com_shift(self, link, xyz):
        f = _d(xyz)
        self.lib.orc_set_com_shift(self.h, int(link), _p(f))

    # ------------------------------------------------------------------ state setters
    def _envs(self, envs):
        return _i(np.arange(self.B) if envs is None else envs)

    def set_qpos(self, qpos, envs=None):
        e, q = self._envs(envs), _d(qpos)
        self.lib.orc_set_qpos(self.h, _p(e), len(e), _p(q))

    def set_qvel(self, qvel, envs=None):
        e, q = self._envs(envs), _d(qvel)
        self.lib.orc_set_qvel(self.h, _p(e), len(e), _p(q))

    def set_dofs_position(self, pos, dofs, envs=None, zero_velocity=True):
        e, q, d = self._envs(envs), _d(pos), _i(dofs)
        self.lib.orc_set_dofs_position(self.h, _p(e), len(e), _p(d), len(d), _p(q), int(zero_velocity))

    def set_base_pos(self, pos, envs=None, zero_velocity=True):
        e, q = self._envs(envs), _d(pos)
        self.lib.orc_set_base_pos(self.h, _p(e), len(e), _p(q), int(zero_velocity))

    def set_base_quat(self, quat, envs=None, zero_velocity=True):
        e, q = self._envs(envs), _d(quat)
        self.lib.orc_set_base_quat(self.h, _p(e), len(e), _p(q), int(zero_velocity))

    def zero_all_dofs_velocity(self, envs=None):
        e = self._envs(envs)
        self.lib.orc_zero_all_dofs_velocity(self.h, _p(e), len(e))

    # ------------------------------------------------------------------ getters
    def _get(self, fn, shape):
        out = np.zeros(shape, dtype=np.float64)
        fn(self.h, _p(out))
        return out

    def get_qpos(self):
        return self._get(self.lib.orc_get_qpos, (self.B, NQ))

    def get_qvel(self):
        return self._get(self.lib.orc_get_qvel, (self.B, NV))

    def get_dofs_position(self):
        return self._get(self.lib.orc_get_dofs_position, (self.B, NV))

    def get_links_pos(self):
        return self._get(self.lib.orc_get_links_pos, (self.B, NL, 3))

    def get_links_quat(self):
        return self._get(self.lib.orc_get_links_quat, (self.B, NL, 4))

    def get_links_vel(self):
        return self._get(self.lib.orc_get_links_vel, (self.B, NL, 3))

    def get_links_ang(self):
        return self._get(self.lib.orc_get_links_ang, (self.B, NL, 3))

    def get_links_net_contact_force(self):
        return self._get(self.lib.orc_get_links_net_contact_force, (self.B, NL, 3))

    def get_dofs_control_force(self):
        return self._get(self.lib.orc_get_dofs_control_force, (self.B, NV))

    def get_errno(self):
        out = np.zeros(self.B, dtype=np.int32)
        self.lib.orc_get_errno(self.h, _p(out))
        return out

    def get_contacts(self, b):
        g = np.zeros((150, 2), dtype=np.int32)
        d = np.zeros((150, 8), dtype=np.float64)
        n = self.lib.orc_get_contacts(self.h, int(b), _p(g), _p(d))
        return g[:n], d[:n]

    def get_gjk_stats(self):
        """[sphere-pair, other-pair first, other-pair perturbed] detections where the reference would prefer GJK, [all MPR collisions]."""
        out = np.zeros(4, dtype=np.int64)
        self.lib.orc_get_gjk_stats(self.h, _p(out))
        return out

    def get_solver_stats(self):
        out = np.zeros((self.B, 3), dtype=np.int32)
        self.lib.orc_get_solver_stats(self.h, _p(out))
        return out

    def get_mass_matrix(self, b):
        out = np.zeros((NV, NV))
        self.lib.orc_get_mass_matrix(self.h, int(b), _p(out))
        return out

    def get_acc(self):
        a, b = np.zeros((self.B, NV)), np.zeros((self.B, NV))
        self.lib.orc_get_acc(self.h, _p(a), _p(b))
        return a, b

    def get_warm_state(self):
        q = np.zeros((self.B, NV))
        w = np.zeros(self.B, dtype=np.int32)
        c = np.zeros((self.B, self.npairs, 3))
        self.lib.orc_get_warm_state(self.h, _p(q), _p(w), _p(c))
        return q, w, c
