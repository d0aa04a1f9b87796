"""Go2 model compiler: URDF -> constant tables for the step kernel (runs once on the host).

Re-derives, from `go2.urdf` + `plane.urdf`, exactly what Genesis' URDF front end produces for the
two entities the Go2 envs build (reference: genesis/utils/urdf.py:52-90 BFS link order, :396-537
merge_fixed_links/merge_inertia, :115-330 per-joint defaults; options/morphs.py:995-1000
default_armature=0.1; rigid_solver.py:188-215 sol_params sanitisation; collider.py:220-329
collision-pair table; rigid_solver.py:529-682 invweight / meaninertia).

Only primitives are handled (box / cylinder / sphere) - that is all the Go2 and plane URDFs use, so
no trimesh/urdfpy is needed.  The output is a plain dict (see `compile_model`) that is serialised to
`go2_model.json` next to this file; the CUDA library and the oracle both consume that table.

Geometry/link numbering (global, as in the reference scene: ground entity first):
  link 0 = ground (fixed), links 1..13 = base, FL_hip, FR_hip, RL_hip, RR_hip, FL_thigh, ... RR_calf
  geom 0 = ground box, geoms 1..27 = robot geoms in (link, URDF-merge) order.
"""
from __future__ import annotations

import json
import math
import os
import xml.etree.ElementTree as ET

import numpy as np

GEOM_SPHERE, GEOM_CYLINDER, GEOM_BOX = 1, 3, 5  # genesis/constants.py:18-27
EPS32 = float(np.finfo(np.float32).eps)
CYL_SECTIONS = 32  # trimesh.creation.cylinder default (sections=32 -> 32-gon rings)
_HERE = os.path.dirname(os.path.abspath(__file__))
MODEL_JSON = os.path.join(_HERE, "go2_model.json")


# ----------------------------------------------------------------------------- small math
def rpy_to_R(rpy):
    r, p, y = rpy
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def R_to_quat(R):
    """Rotation matrix -> (w, x, y, z), w >= 0."""
    t = np.trace(R)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = math.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = [(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s]
    elif R[1, 1] > R[2, 2]:
        s = math.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = [(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s]
    else:
        s = math.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = [(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s]
    q = np.array(q)
    if q[0] < 0:
        q = -q
    return q / np.linalg.norm(q)


def quat_to_R(q):
    w, x, y, z = q
    return np.array(
        [
            [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
            [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
            [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)],
        ]
    )


def _origin_T(elem):
    T = np.eye(4)
    if elem is not None:
        xyz = [float(v) for v in elem.get("xyz", "0 0 0").split()]
        rpy = [float(v) for v in elem.get("rpy", "0 0 0").split()]
        T[:3, :3] = rpy_to_R(rpy)
        T[:3, 3] = xyz
    return T


def _translate_inertia(I, m, d):
    d = np.asarray(d)
    return I + m * (d.dot(d) * np.eye(3) - np.outer(d, d))


# ----------------------------------------------------------------------------- URDF front end
class _Link:
    def __init__(self, name):
        self.name = name
        self.inertial = None  # dict(T, mass, I)
        self.collisions = []  # list of dict(T, type, data)


class _Joint:
    def __init__(self):
        self.name = self.type = self.parent = self.child = None
        self.T = np.eye(4)
        self.axis = np.array([1.0, 0, 0])
        self.lower = self.upper = self.effort = None
        self.damping = self.friction = 0.0


def _parse_urdf(path):
    root = ET.parse(path).getroot()
    links, joints = [], []
    for le in root.findall("link"):
        lk = _Link(le.get("name"))
        ine = le.find("inertial")
        if ine is not None:
            I = ine.find("inertia")
            Im = np.array(
                [
                    [float(I.get("ixx")), float(I.get("ixy")), float(I.get("ixz"))],
                    [float(I.get("ixy")), float(I.get("iyy")), float(I.get("iyz"))],
                    [float(I.get("ixz")), float(I.get("iyz")), float(I.get("izz"))],
                ]
            )
            lk.inertial = dict(T=_origin_T(ine.find("origin")), mass=float(ine.find("mass").get("value")), I=Im)
        for ce in le.findall("collision"):
            g = list(ce.find("geometry"))[0]
            if g.tag == "box":
                gtype, data = GEOM_BOX, [float(v) for v in g.get("size").split()]
            elif g.tag == "cylinder":
                gtype, data = GEOM_CYLINDER, [float(g.get("radius")), float(g.get("length"))]
            elif g.tag == "sphere":
                gtype, data = GEOM_SPHERE, [float(g.get("radius"))]
            else:
                raise ValueError(f"unsupported collision primitive <{g.tag}> in {path}")
            lk.collisions.append(dict(T=_origin_T(ce.find("origin")), type=gtype, data=data))
        links.append(lk)
    for je in root.findall("joint"):
        j = _Joint()
        j.name, j.type = je.get("name"), je.get("type")
        j.parent, j.child = je.find("parent").get("link"), je.find("child").get("link")
        j.T = _origin_T(je.find("origin"))
        ax = je.find("axis")
        if ax is not None:
            j.axis = np.array([float(v) for v in ax.get("xyz").split()])
        lim = je.find("limit")
        if lim is not None:
            j.lower = float(lim.get("lower")) if lim.get("lower") is not None else None
            j.upper = float(lim.get("upper")) if lim.get("upper") is not None else None
            j.effort = float(lim.get("effort")) if lim.get("effort") is not None else None
        dyn = je.find("dynamics")
        if dyn is not None:
            j.damping = float(dyn.get("damping", 0.0))
            j.friction = float(dyn.get("friction", 0.0))
        joints.append(j)
    return links, joints


def _merge_fixed(links, joints):
    """Same traversal as genesis/utils/urdf.py:396-455: repeatedly take the FIRST fixed joint in file
    order, fold the child (inertia, collisions, outgoing joint origins) into the ultimate parent."""
    links, joints = list(links), list(joints)
    merged_into = {}
    while True:
        found = False
        for jt in joints:
            if jt.type != "fixed":
                continue
            pname, cname = jt.parent, jt.child
            while pname in merged_into:
                pname = merged_into[pname]
            while cname in merged_into:
                cname = merged_into[cname]
            by_name = {lk.name: lk for lk in links}
            if pname not in by_name or cname not in by_name:
                continue
            parent, child = by_name[pname], by_name[cname]
            merged_into[jt.child] = pname
            for k in list(merged_into):
                if merged_into[k] == cname:
                    merged_into[k] = pname
            # update_subtree (urdf.py:551-568): push the joint transform into the child's payload
            if child.inertial is not None:
                child.inertial["T"] = jt.T @ child.inertial["T"]
            for c in child.collisions:
                c["T"] = jt.T @ c["T"]
            for j2 in joints:
                if j2.parent == cname:
                    j2.T = jt.T @ j2.T
            # merge_inertia (urdf.py:505-541)
            if child.inertial is not None:
                if parent.inertial is None:
                    parent.inertial = child.inertial
                else:
                    m1, m2 = parent.inertial["mass"], child.inertial["mass"]
                    c1, c2 = parent.inertial["T"][:3, 3].copy(), child.inertial["T"][:3, 3].copy()
                    R1, R2 = parent.inertial["T"][:3, :3], child.inertial["T"][:3, :3]
                    m = m1 + m2
                    com = (m1 * c1 + m2 * c2) / m if m > 0 else c1
                    I1 = _translate_inertia(R1 @ parent.inertial["I"] @ R1.T, m1, com - c1)
                    I2 = _translate_inertia(R2 @ child.inertial["I"] @ R2.T, m2, com - c2)
                    T = np.eye(4)
                    T[:3, 3] = com
                    parent.inertial = dict(T=T, mass=m, I=I1 + I2)
            parent.collisions.extend(child.collisions)
            links.remove(child)
            joints.remove(jt)
            found = True
            break
        if not found:
            break
    for jt in joints:
        jt.parent = merged_into.get(jt.parent, jt.parent)
        jt.child = merged_into.get(jt.child, jt.child)
    return links, joints


def _bfs_order(links, joints):
    """Breadth-first link order, children in original index order (urdf.py:52-90)."""
    idx = {lk.name: i for i, lk in enumerate(links)}
    parent = [-1] * len(links)
    for jt in joints:
        parent[idx[jt.child]] = idx[jt.parent]
    children = {i: [] for i in range(len(links))}
    for c, p in enumerate(parent):
        if p >= 0:
            children[p].append(c)
    level = [i for i, p in enumerate(parent) if p < 0]
    order = []
    while level:
        nxt = []
        for i in level:
            order.append(i)
            nxt += children[i]
        level = nxt
    return order, parent


# ----------------------------------------------------------------------------- tables
def _local_aabb(gtype, data):
    """Initial (geom-frame) AABB of the collision mesh Genesis builds for the primitive
    (box: exact; cylinder: 32-gon prism whose rings contain the +-x/+-y points; sphere: icosphere
    subdivisions=2, which contains the six axis points) -> all three are the analytic extents."""
    if gtype == GEOM_BOX:
        h = 0.5 * np.array(data[:3])
    elif gtype == GEOM_CYLINDER:
        h = np.array([data[0], data[0], 0.5 * data[1]])
    else:
        h = np.array([data[0]] * 3)
    return -h, h


def _geom_verts(gtype, data, n_sphere=3):
    """Vertex cloud of the primitive (only used by the qpos0 overlap filter)."""
    if gtype == GEOM_BOX:
        h = 0.5 * np.array(data[:3])
        return np.array([[sx * h[0], sy * h[1], sz * h[2]] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)])
    if gtype == GEOM_CYLINDER:
        r, hl = data[0], 0.5 * data[1]
        th = 2 * math.pi * np.arange(CYL_SECTIONS) / CYL_SECTIONS
        ring = np.stack([r * np.cos(th), r * np.sin(th)], 1)
        return np.concatenate([np.c_[ring, np.full(len(th), -hl)], np.c_[ring, np.full(len(th), hl)]])
    # sphere: lat/long cloud (conservative enough for the filter)
    pts = []
    for i in range(0, 13):
        for j in range(24):
            ph, th = math.pi * i / 12, 2 * math.pi * j / 24
            pts.append([math.sin(ph) * math.cos(th), math.sin(ph) * math.sin(th), math.cos(ph)])
    return data[0] * np.array(pts)


def _hulls_intersect(A, B):
    """Convex-hull intersection test by LP feasibility (scipy)."""
    from scipy.optimize import linprog

    na, nb = len(A), len(B)
    Aeq = np.zeros((5, na + nb))
    Aeq[:3, :na] = A.T
    Aeq[:3, na:] = -B.T
    Aeq[3, :na] = 1
    Aeq[4, na:] = 1
    beq = np.array([0, 0, 0, 1, 1.0])
    res = linprog(np.zeros(na + nb), A_eq=Aeq, b_eq=beq, bounds=(0, None), method="highs")
    return res.status == 0


def compile_model(go2_urdf, plane_urdf, substep_dt=0.01, constraint_timeconst=0.01, default_armature=0.1):
    links, joints = _parse_urdf(go2_urdf)
    links, joints = _merge_fixed(links, joints)
    order, parent_old = _bfs_order(links, joints)
    new_idx = {old: new for new, old in enumerate(order)}
    joint_of = {jt.child: jt for jt in joints}
    nl = len(order)
    assert nl == 13, nl

    timeconst = max(constraint_timeconst, 2.0 * substep_dt)  # rigid_solver.py:188-215,260-261
    sol_params = [timeconst, 1.0, 0.9, 0.95, 0.001, 0.5, 2.0]

    L = []  # robot links, local index 0..12
    geoms = []
    dof = 6
    for new, old in enumerate(order):
        lk = links[old]
        ent = dict(name=lk.name, parent=(new_idx[parent_old[old]] if parent_old[old] >= 0 else -1))
        if ent["parent"] < 0:
            ent.update(pos=[0.0, 0.0, 0.0], quat=[1.0, 0, 0, 0], joint_name="base_free", dof_start=0, n_dofs=6,
                       q_start=0, axis=[0.0, 0, 1], limit=[-1e30, 1e30], force_range=[-1e30, 1e30], armature=0.0)
        else:
            jt = joint_of[lk.name]
            assert jt.type == "revolute", jt.type
            ent.update(pos=jt.T[:3, 3].tolist(), quat=R_to_quat(jt.T[:3, :3]).tolist(), joint_name=jt.name,
                       dof_start=dof, n_dofs=1, q_start=dof + 1, axis=jt.axis.tolist(),
                       limit=[jt.lower, jt.upper], force_range=[-jt.effort, jt.effort], armature=default_armature,
                       damping=jt.damping, frictionloss=jt.friction)
            assert jt.damping == 0.0 and jt.friction == 0.0  # Go2 URDF has no <dynamics>
            dof += 1
        ine = lk.inertial
        ent.update(inertial_pos=ine["T"][:3, 3].tolist(), inertial_quat=R_to_quat(ine["T"][:3, :3]).tolist(),
                   inertial_i=ine["I"].reshape(-1).tolist(), mass=max(ine["mass"], EPS32))
        L.append(ent)
    assert dof == 18

    # geoms: ground first (plane.urdf: one box on a fixed link), then robot geoms in link order
    gl, _ = _parse_urdf(plane_urdf)
    gc = gl[0].collisions[0]
    lo, hi = _local_aabb(gc["type"], gc["data"])
    geoms.append(dict(link=0, type=gc["type"], data=list(gc["data"]) + [0.0] * (3 - len(gc["data"])),
                      pos=gc["T"][:3, 3].tolist(), quat=R_to_quat(gc["T"][:3, :3]).tolist(),
                      aabb_min=lo.tolist(), aabb_max=hi.tolist()))
    for new, old in enumerate(order):
        for c in links[old].collisions:
            lo, hi = _local_aabb(c["type"], c["data"])
            geoms.append(dict(link=new + 1, type=c["type"], data=list(c["data"]) + [0.0] * (3 - len(c["data"])),
                              pos=c["T"][:3, 3].tolist(), quat=R_to_quat(c["T"][:3, :3]).tolist(),
                              aabb_min=lo.tolist(), aabb_max=hi.tolist()))
    ng = len(geoms)

    # ---- neutral-pose kinematics (float64) for invweight / meaninertia / qpos0 overlap filter
    Rw = [None] * nl
    pw = [None] * nl
    for i, ent in enumerate(L):
        if ent["parent"] < 0:
            Rw[i], pw[i] = np.eye(3), np.zeros(3)
        else:
            p = ent["parent"]
            Rw[i] = Rw[p] @ quat_to_R(ent["quat"])
            pw[i] = pw[p] + Rw[p] @ np.array(ent["pos"])
    nv = 18
    M = np.zeros((nv, nv))
    Jp_com, Jr = [], []
    for i, ent in enumerate(L):
        com = pw[i] + Rw[i] @ np.array(ent["inertial_pos"])
        Ri = Rw[i] @ quat_to_R(ent["inertial_quat"])
        Iw = Ri @ np.array(ent["inertial_i"]).reshape(3, 3) @ Ri.T
        jp, jr = np.zeros((3, nv)), np.zeros((3, nv))
        k = i
        while k >= 0:
            e = L[k]
            if e["parent"] < 0:
                jp[:, 0:3] = np.eye(3)
                for a in range(3):
                    ax = Rw[k][:, a]
                    jr[:, 3 + a] = ax
                    jp[:, 3 + a] = np.cross(ax, com - pw[k])
            else:
                ax = Rw[k] @ np.array(e["axis"])
                d = e["dof_start"]
                jr[:, d] = ax
                jp[:, d] = np.cross(ax, com - pw[k])
            k = e["parent"]
        M += ent["mass"] * jp.T @ jp + jr.T @ Iw @ jr
        Jp_com.append(jp)
        Jr.append(jr)
    for e in L:
        if e["parent"] >= 0:
            M[e["dof_start"], e["dof_start"]] += e["armature"]
    Minv = np.linalg.inv(M)
    links_invweight = [[0.0, 0.0]]  # ground link
    for i in range(nl):
        J = np.concatenate([Jp_com[i], Jr[i]], 0)
        A = np.diag(J @ Minv @ J.T)
        links_invweight.append([float(A[:3].mean()), float(A[3:].mean())])
    dmi = np.diag(Minv)
    dofs_invweight = [float(dmi[:3].mean())] * 3 + [float(dmi[3:6].mean())] * 3 + [float(v) for v in dmi[6:]]
    meaninertia = float(np.trace(M) / nv)

    # ---- collision-pair table (collider.py:220-329)
    link_parent_global = [-1] + [(-1 if e["parent"] < 0 else e["parent"] + 1) for e in L]
    gverts = []
    for g in geoms:
        V = _geom_verts(g["type"], g["data"])
        V = V.mean(0, keepdims=True) + (1.0 - 1e-3) * (V - V.mean(0, keepdims=True))
        Rg = quat_to_R(g["quat"])
        if g["link"] == 0:
            gverts.append(V @ Rg.T + np.array(g["pos"]))
        else:
            i = g["link"] - 1
            gverts.append((V @ Rg.T + np.array(g["pos"])) @ Rw[i].T + pw[i])
    pair_idx = -np.ones((ng, ng), dtype=int)
    pairs, filtered = [], []
    for a in range(ng):
        for b in range(a + 1, ng):
            la, lb = geoms[a]["link"], geoms[b]["link"]
            if la == lb:
                continue
            if la != 0 and lb != 0:  # self-collision of the robot
                lo_, hi_ = (la, lb) if la < lb else (lb, la)
                if link_parent_global[hi_] == lo_:  # adjacent links
                    continue
                if _hulls_intersect(gverts[a], gverts[b]):  # active at qpos0 -> filtered
                    filtered.append((a, b))
                    continue
            pair_idx[a, b] = len(pairs)
            pairs.append([a, b])

    model = dict(
        n_links=nl, n_dofs=nv, n_q=19, n_geoms=ng, substep_dt=substep_dt, gravity=[0.0, 0.0, -9.81],
        sol_params=sol_params, links=L, geoms=geoms, pairs=pairs, filtered_pairs=filtered,
        links_invweight=links_invweight, dofs_invweight=dofs_invweight, meaninertia=meaninertia,
        total_mass=float(sum(e["mass"] for e in L)),
    )
    return model


def default_urdf_paths(reference_root="/root/reference"):
    a = os.path.join(reference_root, "genesis/assets/urdf")
    return os.path.join(a, "go2/urdf/go2.urdf"), os.path.join(a, "plane/plane.urdf")


def load_model(path=MODEL_JSON):
    with open(path) as f:
        return json.load(f)


def motors_dof_idx(model, joint_names):
    by = {e["joint_name"]: e["dof_start"] for e in model["links"]}
    return [by[n] for n in joint_names]


if __name__ == "__main__":
    import sys

    go2, plane = default_urdf_paths(*(sys.argv[1:2]))
    m = compile_model(go2, plane)
    with open(MODEL_JSON, "w") as f:
        json.dump(m, f, indent=1)
    print("links:", [e["name"] for e in m["links"]])
    print("n_geoms", m["n_geoms"], "n_pairs", len(m["pairs"]), "filtered", m["filtered_pairs"])
    print("total mass", m["total_mass"], "meaninertia", m["meaninertia"])
    print("dofs_invweight", np.round(m["dofs_invweight"], 4))


def flatten_model(model):
    """Pack the model dict into (int32[], float64[]) - the flat layout both native libraries parse.

    ints  : n_links, n_geoms, n_pairs, per robot link (parent_local, dof_start, n_dofs, q_start),
            per geom (link_global, type), per pair (geom_a, geom_b)
    floats: per robot link: pos3 quat4 inertial_pos3 inertial_quat4 inertial_i9 mass axis3 limit2
            force_range2 armature (=32); per geom: pos3 quat4 data3 aabb_min3 aabb_max3 (=16);
            sol_params7; links_invweight[14] (translational); dofs_invweight[18]; meaninertia; gravity3; dt
    """
    iv = [model["n_links"], model["n_geoms"], len(model["pairs"])]
    fv = []
    for e in model["links"]:
        iv += [e["parent"], e["dof_start"], e["n_dofs"], e["q_start"]]
        fv += list(e["pos"]) + list(e["quat"]) + list(e["inertial_pos"]) + list(e["inertial_quat"])
        fv += list(e["inertial_i"]) + [e["mass"]] + list(e["axis"]) + list(e["limit"]) + list(e["force_range"])
        fv += [e["armature"]]
    for g in model["geoms"]:
        iv += [g["link"], g["type"]]
        fv += list(g["pos"]) + list(g["quat"]) + list(g["data"]) + list(g["aabb_min"]) + list(g["aabb_max"])
    for a, b in model["pairs"]:
        iv += [a, b]
    fv += list(model["sol_params"])
    fv += [w[0] for w in model["links_invweight"]]
    fv += list(model["dofs_invweight"])
    fv += [model["meaninertia"]] + list(model["gravity"]) + [model["substep_dt"]]
    return np.asarray(iv, dtype=np.int32), np.asarray(fv, dtype=np.float64)
