"""The reference answers support queries of mesh-backed geoms (the Go2's 17 cylinders) from a 180 x 180 direction grid
(collider/support_field.py:22-92 build, :139-182 `_func_support_mesh`: the 4 grid cells around the query direction, best dot
product wins).  Kernel and oracle use the closed form for the 32-gon prism instead (csrc/go2_collide.cuh: support(), nearest
ring vertex to the direction's azimuth, top or bottom ring by the sign of d.z).  This test restates the reference's grid and
lookup in numpy for every cylinder of the model and shows that the two return the SAME vertex for a million random
directions - so the replacement is not a deviation for these shapes."""
import json
import math
import os

import numpy as np

from util import ROOT

MODEL = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/go2_model.json")))
RES = 180


def _grid():  # support_field.py:22-35
    theta = np.arange(RES) / RES * 2 * math.pi - math.pi
    phi = np.arange(RES) / RES * math.pi
    T, P = np.meshgrid(theta, phi, indexing="ij")
    return np.stack((np.sin(P) * np.cos(T), np.sin(P) * np.sin(T), np.cos(P)), -1).reshape(-1, 3)


def _prism(r, half_h):  # trimesh.creation.cylinder(sections=32): two 32-gon rings starting at azimuth 0 + the two cap centres
    a = np.arange(32) / 32 * 2 * math.pi
    ring = np.stack((r * np.cos(a), r * np.sin(a)), 1)
    return np.concatenate([np.c_[ring, np.full(32, -half_h)], np.c_[ring, np.full(32, half_h)], [[0, 0, -half_h], [0, 0, half_h]]])


def _lookup(verts, cell_vid, d):  # support_field.py:139-182
    th, ph = np.arctan2(d[:, 1], d[:, 0]), np.arccos(np.clip(d[:, 2], -1, 1))
    ii, jj = (th + math.pi) / math.pi / 2 * RES, ph / math.pi * RES
    best, out = np.full(len(d), -1e20), np.zeros((len(d), 3))
    for i4 in range(4):
        i = (np.ceil(ii) if i4 % 2 else np.floor(ii)).astype(int) % RES
        if i4 // 2 > 0:
            j = np.clip(np.ceil(jj), 0, RES - 1).astype(int)
            j[j == RES - 1] = RES - 2
        else:
            j = np.clip(np.floor(jj), 0, RES - 1).astype(int)
            j[j == 0] = 1
        p = verts[cell_vid[i * RES + j]]
        dot = (p * d).sum(1)
        m = dot > best
        best[m], out[m] = dot[m], p[m]
    return out


def _closed_form(r, half_h, d):  # go2_collide.cuh: support(), cylinder branch
    step = 2 * math.pi / 32
    k = np.floor(np.arctan2(d[:, 1], d[:, 0]) / step + 0.5)
    return np.stack((r * np.cos(k * step), r * np.sin(k * step), np.where(d[:, 2] < 0, -half_h, half_h)), 1)


def test_support_field_lookup_equals_the_closed_form_for_every_cylinder_of_the_model():
    V = _grid()
    rng = np.random.default_rng(0)
    d = rng.standard_normal((1_000_000, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    shapes = sorted({(g["data"][0], g["data"][1]) for g in MODEL["geoms"] if g["type"] == 3})
    assert len(shapes) >= 4
    for r, h in shapes:
        verts = _prism(r, 0.5 * h)
        cell_vid = np.argmax(V @ verts.T, 1)  # support_field.py:52-63
        a, b = _lookup(verts, cell_vid, d), _closed_form(r, 0.5 * h, d)
        assert int((np.abs(a - b).max(1) > 1e-9).sum()) == 0, (r, h)
