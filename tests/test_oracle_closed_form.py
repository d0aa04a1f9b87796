"""Closed-form / independent checks of the CPU oracle (L1 of SURVEY.md 8): the reference ships no golden vectors for its
rigid solver, so the restatement is pinned against things that can be derived on paper or computed along a second route:

  * MPR sphere-vs-box-face contacts against the analytic answer - which is also what the reference's GJK/EPA fallback
    (narrowphase.py:724-736, not restated) converges to for the only pairs on which it fires in the walk / crouch / jump envs;
  * the soft-contact law: resting normal force of every contact = 4 K imp(pen) pen / diag from MuJoCo's published solref /
    solimp formulas (geom.py:405-427, solver.py:498-595);
  * Coulomb's slip threshold of the friction pyramid and the closed-form minimiser of a single joint-limit row;
  * free flight: the centre of mass accelerates with exactly g and the angular momentum about it is conserved;
  * `invweight` / `meaninertia` of the model table (shared by product and oracle) recomputed from the oracle's mass matrix and
    a velocity-kinematics Jacobian, following rigid_solver.py:529-682.
"""
import json
import os

import numpy as np
import pytest

from oracle.rigid import OracleRigid
from util import ROOT, load_golden

MODEL = json.load(open(os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200/go2_b200/go2_model.json")))
MDI = [7, 11, 15, 6, 10, 14, 9, 13, 17, 8, 12, 16]
Q0 = np.array([0, 0.8, -1.5, 0, 0.8, -1.5, 0, 1.0, -1.5, 0, 1.0, -1.5])


def _quat_rot(q, v):
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    return R @ v


def test_mpr_sphere_vs_ground_box_matches_the_analytic_contact():
    g, _ = load_golden("walk_c1")
    B = g["actions"].shape[1]
    o = OracleRigid(B, kind="f32")
    geoms = MODEL["geoms"]
    n = worst = 0
    for t in range(0, 1000, 25):
        o.set_qpos(g["qpos"][t])
        o.set_qvel(g["qvel"][t])
        o.step(1)
        gp = None
        for e in range(B):
            gg, d = o.get_contacts(e)
            for (ga, gb), c in zip(gg, d):
                a, b = (ga, gb) if geoms[ga]["type"] == 1 else (gb, ga)
                if geoms[a]["type"] != 1 or b != 0:
                    continue
                r = geoms[a]["data"][0]
                # analytic: sphere of radius r whose centre is (cx, cy, cz) over the box face z = 0:
                # normal = +z (from the box towards the sphere), pen = r - cz, position = midway between the two surface points
                pos, nrm, pen = c[0:3], c[3:6], c[6]
                assert abs(abs(nrm[2]) - 1.0) < 1e-5 and abs(nrm[0]) < 2e-3 and abs(nrm[1]) < 2e-3
                # the two witness points are (cx, cy, cz - r) on the sphere and (cx, cy, 0) on the face: pos.z = (cz - r) / 2, pen = r - cz
                worst = max(worst, abs(pos[2] + pen / 2))
                assert abs(pos[2] + pen / 2) < 2e-6, (pos, pen)
                assert 0 <= pen < r
                n += 1
    assert n > 500
    print(f"\n{n} foot-ground MPR contacts match the analytic sphere / plane contact (max |pos.z + pen/2| = {worst:.2e})")


def _sol(neg_pen, vel, pos):
    timeconst, dampratio, dmin, dmax, width, mid, power = MODEL["sol_params"]
    x = abs(neg_pen) / width
    y = x ** power / mid ** (power - 1) if x < mid else 1 - (1 - x) ** power / (1 - mid) ** (power - 1)
    imp = min(max(dmin + y * (dmax - dmin), dmin), dmax) if x <= 1 else dmax
    b, k = 2 / (dmax * timeconst), 1 / (dmax ** 2 * timeconst ** 2 * dampratio ** 2)
    return imp, -b * vel - k * imp * pos


def test_resting_contact_force_follows_the_soft_constraint_law():
    """Robot settled on the ground under PD control: for every foot contact the solver's normal force must equal
    4 * D * aref with D = 1 / diag, diag = invweight (1 + mu^2) * 2 mu^2 (1 - imp) / imp (4 pyramid rows, all active, no
    tangential load left once the robot is at rest), up to the residual acceleration / velocity terms."""
    o = OracleRigid(1, kind="f64")
    o.set_dofs_position(Q0[None], MDI)
    o.set_dofs_kp(np.full(12, 60.0), MDI)
    o.set_dofs_kv(np.full(12, 5.0), MDI)
    for _ in range(3000):
        o.control_dofs_position(Q0[None], MDI)
        o.step(1)
    assert np.abs(o.get_qvel()).max() < 2e-3, np.abs(o.get_qvel()).max()
    gg, d = o.get_contacts(0)
    cf = o.get_links_net_contact_force()[0]
    inv_link = MODEL["links_invweight"]
    feet = 0
    for (ga, gb), c in zip(gg, d):
        a, b = (ga, gb) if MODEL["geoms"][ga]["type"] == 1 else (gb, ga)
        if MODEL["geoms"][a]["type"] != 1 or b != 0:
            continue
        link = MODEL["geoms"][a]["link"]
        pen, mu = c[6], c[7]
        imp, aref = _sol(-pen, 0.0, -pen)
        invw = inv_link[link][0] + inv_link[0][0]
        diag = invw * (1 + mu * mu) * 2 * mu * mu * (1 - imp) / imp
        n_expected = 4 * aref / diag
        assert cf[link][2] == pytest.approx(n_expected, rel=2e-2), (link, cf[link], n_expected)
        feet += 1
    assert feet == 4
    assert cf[1:, 2].sum() == pytest.approx(MODEL["total_mass"] * 9.81, rel=1e-3)  # robot links (link 0 = ground takes the reaction)


def test_free_flight_momentum():
    """No contacts, zero torques, 0.5 s of flight: the linear momentum of the links changes by M g t (semi-implicit Euler in
    generalised coordinates is exact for the free joint's translation; the configuration-dependent coupling and the rotor
    armature leave an O(h) residual, bounded here at 1 %), and the float and double builds agree to 1e-4."""
    masses = np.array([0.0] + [l["mass"] for l in MODEL["links"]])
    ipos = [np.zeros(3)] + [np.array(l["inertial_pos"]) for l in MODEL["links"]]
    res = {}
    for kind in ("f64", "f32"):
        rng = np.random.default_rng(0)
        o = OracleRigid(1, kind=kind)
        q = np.zeros((1, 19))
        q[0, :3] = [0, 0, 5.0]
        q[0, 3] = 1.0
        q[0, 7:] = Q0[np.argsort(np.argsort(MDI))] * 0 + np.array([0, 0, 0, 0, 0.8, 0.8, 1.0, 1.0, -1.5, -1.5, -1.5, -1.5]) + 0.1 * rng.standard_normal(12)
        o.set_qpos(q)
        o.set_qvel(0.5 * rng.standard_normal((1, 18)))

        def momentum():
            quat, vel, ang = o.get_links_quat()[0], o.get_links_vel()[0], o.get_links_ang()[0]
            p = np.zeros(3)
            for l in range(1, 14):
                p += masses[l] * (vel[l] + np.cross(ang[l], _quat_rot(quat[l], ipos[l])))
            return p

        o.control_dofs_force(np.zeros((1, 12)), MDI)
        p0 = momentum()
        n = 50
        for _ in range(n):
            o.step(1)
        assert o.get_solver_stats()[0, 0] == 0  # no contacts
        p1 = momentum()
        expect = p0 + masses.sum() * np.array(MODEL["gravity"]) * n * MODEL["substep_dt"]
        assert np.abs(p1 - expect).max() < 1e-2 * np.abs(expect).max(), (p1, expect)
        res[kind] = p1
    assert np.abs(res["f32"] - res["f64"]).max() < 1e-4 * np.abs(res["f64"]).max()


def test_invweight_and_meaninertia_recomputed_independently():
    """rigid_solver.py:529-682: M at qpos0, dof invweight = diag(M^-1) (free joint: means over the 3 translational / 3
    rotational dofs), link invweight = mean diag(J M^-1 J^T) of the link-COM translational Jacobian, meaninertia = mean
    diag(M).  Here M comes from the f64 oracle's CRBA and J from its velocity kinematics (unit dof velocities)."""
    o = OracleRigid(1, kind="f64")
    q = np.zeros((1, 19))
    q[0, :3] = [0, 0, 0.42]
    q[0, 3] = 1.0
    o.set_qpos(q)
    o.set_qvel(np.zeros((1, 18)))
    o.control_dofs_force(np.zeros((1, 12)), MDI)
    o.step(1)  # refreshes the mass matrix at (nearly) qpos0: one 10 ms free-fall substep moves the base by 0.5 mm only
    o.set_qpos(q)
    M = o.get_mass_matrix(0)
    M = np.tril(M) + np.tril(M, -1).T
    Minv = np.linalg.inv(M)
    dinv = np.diag(Minv).copy()
    dinv[:3], dinv[3:6] = dinv[:3].mean(), dinv[3:6].mean()
    assert np.allclose(dinv, MODEL["dofs_invweight"], rtol=2e-3), (dinv, MODEL["dofs_invweight"])
    assert np.trace(M) / 18 == pytest.approx(MODEL["meaninertia"], rel=2e-3)
    ipos = [np.zeros(3)] + [np.array(l["inertial_pos"]) for l in MODEL["links"]]
    J = np.zeros((14, 3, 18))
    for d in range(18):
        v = np.zeros((1, 18))
        v[0, d] = 1.0
        o.set_qvel(v)
        quat, vel, ang = o.get_links_quat()[0], o.get_links_vel()[0], o.get_links_ang()[0]
        for l in range(1, 14):
            J[l, :, d] = vel[l] + np.cross(ang[l], _quat_rot(quat[l], ipos[l]))
    for l in range(1, 14):
        w = np.diag(J[l] @ Minv @ J[l].T).mean()
        assert w == pytest.approx(MODEL["links_invweight"][l][0], rel=5e-3), (l, w, MODEL["links_invweight"][l])


def _geom_vertices(g):
    """Local vertices of a box (full sizes) or of the 32-gon prism the reference meshes a cylinder into (radius, height)."""
    if g["type"] == 5:
        sx, sy, sz = (0.5 * v for v in g["data"])
        return np.array([[a * sx, b * sy, c * sz] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)])
    r, h = g["data"][0], g["data"][1]
    a = np.arange(32) / 32 * 2 * np.pi
    ring = np.stack([r * np.cos(a), r * np.sin(a)], 1)
    return np.concatenate([np.c_[ring, np.full(32, -h / 2)], np.c_[ring, np.full(32, h / 2)]])


def test_mpr_box_and_cylinder_vs_ground_match_the_deepest_point_answer():
    """The pairs on which the reference's GJK/EPA fallback can fire with a non-sphere geom (base env: calf / thigh cylinders and
    calf boxes on the ground in a crouch; counted by orc_get_gjk_stats).  For a convex body against the ground's top face both
    MPR and EPA must return the minimum-translation answer: normal = +z, depth = how far the body's lowest vertex is under z = 0,
    witness midway.  The oracle's MPR (first detection of the pair) is checked against exactly that."""
    B = 16
    o = OracleRigid(B, kind="f32")
    geoms = MODEL["geoms"]
    n = 0
    worst_pen = worst_n = 0.0
    rng = np.random.default_rng(5)
    for t in range(24):
        # robots lying low in random attitudes with random leg poses: base box, hip / thigh / calf cylinders and calf boxes on the ground
        qpos = np.zeros((B, 19))
        qpos[:, 0:2] = rng.uniform(-1, 1, (B, 2))
        qpos[:, 2] = rng.uniform(0.05, 0.16, B)
        rpy = rng.uniform(-1.2, 1.2, (B, 3)) * np.array([1.0, 0.6, 1.0])
        cr, sr, cp, sp, cy, sy = (f(rpy[:, i] / 2) for i in range(3) for f in (np.cos, np.sin))
        qpos[:, 3:7] = np.stack([cr * cp * cy + sr * sp * sy, sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy], 1)
        qpos[:, 7:19] = rng.uniform(-0.8, 0.8, (B, 12)) + np.tile([0, 0, 0, 0, 1.0, 1.0, 1.0, 1.0, -1.8, -1.8, -1.8, -1.8], (B, 1))
        o.set_qpos(qpos)
        o.set_qvel(np.zeros((B, 18)))
        lp, lq = o.get_links_pos(), o.get_links_quat()  # poses the detection of the next substep runs on
        o.step(1)
        for e in range(B):
            gg, d = o.get_contacts(e)
            seen = set()
            for (ga, gb), c in zip(gg, d):
                a = ga if gb == 0 else (gb if ga == 0 else None)
                if a is None or geoms[a]["type"] == 1 or a in seen:
                    continue
                seen.add(a)  # first contact of the pair = the unperturbed detection
                G = geoms[a]
                V = _geom_vertices(G)
                local = np.array([_quat_rot(G["quat"], v) for v in V]) + np.array(G["pos"])
                world = np.array([_quat_rot(lq[e, G["link"]], v) for v in local]) + lp[e, G["link"]]  # link 0 = the ground
                depth = -world[:, 2].min()
                pos, nrm, pen = c[0:3], c[3:6], c[6]
                assert depth > 0
                worst_pen, worst_n = max(worst_pen, abs(pen - depth)), max(worst_n, 1.0 - abs(nrm[2]))
                if depth > 0.03:
                    continue  # deep overlaps are not what a running env sees (MPR's portal may settle on another face)
                assert abs(pen - depth) < 2e-5, (t, e, a, pen, depth)
                assert 1.0 - abs(nrm[2]) < 1e-5 and abs(pos[2] + depth / 2) < 2e-5, (t, e, a, nrm, pos, depth)
                n += 1
    assert n > 50
    print(f"\n{n} box / cylinder vs ground MPR contacts equal the deepest-point answer (max |d pen| = {worst_pen:.2e}, max 1 - |n.z| = {worst_n:.2e})")


def test_friction_pyramid_slip_threshold():
    """Standing robot (PD hold) on mu = 0.3 pushed horizontally at the base along a pyramid axis (world x): below mu m g it stays put
    (the soft constraint only creeps), above it slides with a = (F - mu m g) / m.  Exercises contact-row assembly, the pyramid's
    friction limit and the Newton solve's active-set logic against Coulomb's law."""
    m, mu, g = MODEL["total_mass"], 0.3, 9.81
    out = {}
    for frac in (0.5, 1.6):
        o = OracleRigid(1, kind="f64")
        o.set_friction(0, mu)
        o.set_friction(1, mu)
        o.set_dofs_position(Q0[None], MDI)
        o.set_dofs_kp(np.full(12, 60.0), MDI)
        o.set_dofs_kv(np.full(12, 5.0), MDI)
        for _ in range(1500):  # settle
            o.control_dofs_position(Q0[None], MDI)
            o.step(1)
        F = frac * mu * m * g
        vx = []
        for k in range(60):  # 0.6 s of pushing (the force is cleared by every step call, like Scene.step does)
            o.control_dofs_position(Q0[None], MDI)
            o.apply_link_external_force(np.array([[F, 0.0, 0.0]]), 1)
            o.step(1)
            vx.append(o.get_qvel()[0, 0])
        out[frac] = (F, np.array(vx))
    F, vx = out[0.5]
    assert abs(vx[-1]) < 0.01 and np.abs(vx).max() < 0.15, vx[-5:]  # sticks: the legs deflect under the PD hold (transient sway), then rest
    F, vx = out[1.6]
    a = (vx[-1] - vx[19]) / (40 * 0.01)  # mean acceleration over the last 0.4 s
    assert a == pytest.approx((F - mu * m * g) / m, rel=0.15), (a, (F - mu * m * g) / m)


def test_single_joint_limit_row_has_the_closed_form_solution():
    """Robot in free flight with ONE joint pushed 0.05 rad past its upper limit and nothing else active: the constrained
    acceleration minimises 1/2 (a - a0)' M (a - a0) + 1/2 D (J a - aref)^2, i.e. a = a0 + M^-1 J' f with
    f = -D (J a0 - aref) / (1 + D J M^-1 J') - checked against the oracle's limit row (solver.py:1088-1143: jac = -1 at the upper
    limit, diag = invweight_dof (1 - imp) / imp) and its Newton solve."""
    link = next(l for l in MODEL["links"] if l["name"] == "FL_calf")
    j, over = link["dof_start"], 0.05
    o = OracleRigid(1, kind="f64", base_pos=(0.0, 0.0, 1.0))
    o.set_dofs_position(Q0[None], MDI)
    q = o.get_qpos()
    q[0, link["q_start"]] = link["limit"][1] + over
    o.set_qpos(q)
    o.set_qvel(np.zeros((1, 18)))
    o.control_dofs_force(np.zeros((1, 12)), MDI)
    o.step(1)
    assert len(o.get_contacts(0)[0]) == 0
    a0, a = o.get_acc()
    M = o.get_mass_matrix(0)
    imp, aref = _sol(-over, 0.0, -over)
    D = 1.0 / (MODEL["dofs_invweight"][j] * (1 - imp) / imp)
    J = np.zeros(18)
    J[j] = -1.0
    Minv_Jt = np.linalg.solve(M, J)
    f = -D * (J @ a0[0] - aref) / (1.0 + D * (J @ Minv_Jt))
    assert f > 0
    want = a0[0] + Minv_Jt * f
    np.testing.assert_allclose(a[0], want, rtol=1e-6, atol=1e-6 * np.abs(want).max())
    assert a[0, j] < a0[0, j] - 1.0  # the limit pushes the joint back (negative direction at the upper limit)
