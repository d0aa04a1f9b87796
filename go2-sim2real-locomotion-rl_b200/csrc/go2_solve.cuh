// go2_b200 - per-environment constraint assembly and primal Newton solve.
//
// Soft-constraint model of the reference (MuJoCo style): 4 pyramid rows per contact, 1 row per violated
// joint limit, convex cost 0.5 (Ma - f).(a - a0) + sum_active 0.5 D (J a - aref)^2 minimised by Newton
// steps with an exact line search and rank-1 Cholesky up/down-dates of H = M + J^T D J when the active
// set changes (constraint/solver.py:498-595, 1088-1143, 1286-1343, 1467-1494, 1632-1675, 1747-1765,
// 1888-2417, 2428-2966, 2974-3037).  All working data is thread-private; matrices are packed lower
// triangles.
#pragma once
#include "go2_collide.cuh"

namespace go2 {

G2_HDN void assemble_constraints(const Go2Model& m, Phys& p) {
    int r = 0;
    for (int ic = 0; ic < p.n_contacts; ++ic) {
        const Contact& c = p.con[ic];
        v3 d1, d2;
        orthobasis(c.normal, d1, d2);
        float invw = (c.la >= 0 ? m.invweight_link[c.la] : 0.0f) + (c.lb >= 0 ? m.invweight_link[c.lb] : 0.0f);
        v3 tp = c.pos - p.root_com;
        for (int k = 0; k < 4; ++k) {
            v3 d = (k < 2 ? d1 : d2) * (float)(2 * (k % 2) - 1);
            v3 n = d * c.mu - c.normal;
            float* Jr = p.J[r];
            for (int i = 0; i < NV; ++i) Jr[i] = 0.0f;
            float jqv = 0.0f;
            for (int side = 0; side < 2; ++side) {
                float sgn = side == 0 ? -1.0f : 1.0f;
                int l = side == 0 ? c.la : c.lb;
                while (l >= 0) {
                    int d0 = l == 0 ? 0 : 5 + l, nd = l == 0 ? 6 : 1;
                    for (int i = d0 + nd - 1; i >= d0; --i) {
                        v3 vel = p.cdof_v[i] - cross(tp, p.cdof_a[i]);
                        float j = sgn * dot(vel, n);
                        jqv += j * p.qvel[i];
                        Jr[i] += j;
                    }
                    l = l == 0 ? -1 : m.parent[l];
                }
            }
            float imp, ar;
            imp_aref(m.sol_params, -c.pen, jqv, -c.pen, imp, ar);
            float diag = invw + c.mu * c.mu * invw;
            diag *= 2.0f * c.mu * c.mu * (1.0f - imp) / imp;
            diag = fmaxf(diag, EPSF);
            p.aref[r] = ar;
            p.Dd[r] = 1.0f / diag;
            ++r;
        }
    }
    for (int d = 6; d < NV; ++d) {
        float q = p.qpos[d + 1];
        float dmin = q - m.limit_lo[d], dmax = m.limit_hi[d] - q;
        float pd = fminf(dmin, dmax);
        if (pd < 0.0f) {
            float j = dmin < dmax ? 1.0f : -1.0f;
            float imp, ar;
            imp_aref(m.sol_params, pd, j * p.qvel[d], pd, imp, ar);
            float diag = fmaxf(m.invweight_dof[d] * (1.0f - imp) / imp, EPSF);
            float* Jr = p.J[r];
            for (int i = 0; i < NV; ++i) Jr[i] = 0.0f;
            Jr[d] = j;
            p.aref[r] = ar;
            p.Dd[r] = 1.0f / diag;
            ++r;
        }
    }
    p.n_rows = r;
}

G2_HD float Mat(const Phys& p, int i, int j) { return i >= j ? p.M[tri(i, j)] : p.M[tri(j, i)]; }

// active set, constraint forces, J^T f, cost
G2_HDN void update_constraint(Phys& p) {
    p.prev_cost = p.cost;
    for (int d = 0; d < NV; ++d) p.qfc[d] = 0.0f;
    float cost = 0.0f, gauss = 0.0f;
    float ccost = 0.0f;
    for (int c = 0; c < p.n_rows; ++c) {
        p.pact[c] = p.act[c];
        float ja = p.Jaref[c];
        bool a = ja < 0.0f;
        p.act[c] = a;
        float f = a ? -ja * p.Dd[c] : 0.0f;
        p.efc[c] = f;
        if (a) {
            const float* Jr = p.J[c];
            for (int d = 0; d < NV; ++d) p.qfc[d] += Jr[d] * f;
            ccost += 0.5f * ja * ja * p.Dd[c];
        }
    }
    for (int d = 0; d < NV; ++d) {
        float v = 0.5f * (p.Ma[d] - p.force[d]) * (p.qacc[d] - p.acc_smooth[d]);
        gauss += v;
    }
    cost = gauss + ccost;
    p.gauss = gauss;
    p.cost = cost;
}

// H = M + J^T D J over active rows, then dense Cholesky H = L L^T (L overwrites H)
G2_HDN void hessian_factor(Phys& p) {
    for (int k = 0; k < NTRI; ++k) p.H[k] = p.M[k];
    for (int c = 0; c < p.n_rows; ++c) {
        if (!p.act[c]) continue;
        const float* Jr = p.J[c];
        float D = p.Dd[c];
        for (int i = 0; i < NV; ++i) {
            float ji = Jr[i];
            if (fabsf(ji) > EPSF) {
                float s = ji * D;
                for (int j = 0; j <= i; ++j) p.H[tri(i, j)] += Jr[j] * s;
            }
        }
    }
    for (int i = 0; i < NV; ++i) {
        float t = p.H[tri(i, i)];
        for (int j = 0; j < i; ++j) t -= p.H[tri(i, j)] * p.H[tri(i, j)];
        float lii = sqrtf(fmaxf(t, EPSF));
        p.H[tri(i, i)] = lii;
        float inv = 1.0f / lii;
        for (int j = i + 1; j < NV; ++j) {
            float dt = 0.0f;
            for (int k = 0; k < i; ++k) dt += p.H[tri(j, k)] * p.H[tri(i, k)];
            p.H[tri(j, i)] = (p.H[tri(j, i)] - dt) * inv;
        }
    }
}
// rank-1 up/down-dates for rows whose active flag flipped; returns false if a downdate degenerated
G2_HDN bool hessian_update(Phys& p) {
    bool ok = true;
    float vec[NV];
    for (int c = 0; c < p.n_rows; ++c) {
        if (p.act[c] == p.pact[c]) continue;
        float sgn = p.act[c] ? 1.0f : -1.0f;
        float sd = sqrtf(p.Dd[c]);
        for (int d = 0; d < NV; ++d) vec[d] = p.J[c][d] * sd;
        for (int k = 0; k < NV; ++k) {
            if (fabsf(vec[k]) > EPSF) {
                float Lkk = p.H[tri(k, k)];
                float t = Lkk * Lkk + sgn * vec[k] * vec[k];
                if (t < EPSF) { ok = false; break; }
                float r = sqrtf(t), cc = r / Lkk, ci = 1.0f / cc, s = vec[k] / Lkk;
                p.H[tri(k, k)] = r;
                for (int i = k + 1; i < NV; ++i) {
                    float h = (p.H[tri(i, k)] + s * vec[i] * sgn) * ci;
                    p.H[tri(i, k)] = h;
                    vec[i] = vec[i] * cc - s * h;
                }
            }
        }
    }
    return ok;
}
G2_HDN void update_gradient(Phys& p) {
    for (int d = 0; d < NV; ++d) p.grad[d] = p.Ma[d] - p.force[d] - p.qfc[d];
    for (int i = 0; i < NV; ++i) {
        float c = p.grad[i];
        for (int j = 0; j < i; ++j) c -= p.H[tri(i, j)] * p.Mgrad[j];
        p.Mgrad[i] = c / p.H[tri(i, i)];
    }
    for (int i = NV - 1; i >= 0; --i) {
        float c = p.Mgrad[i];
        for (int j = i + 1; j < NV; ++j) c -= p.H[tri(j, i)] * p.Mgrad[j];
        p.Mgrad[i] = c / p.H[tri(i, i)];
    }
}

struct LsPt {
    float alpha, cost, d0, d1;
};
G2_HD LsPt ls_eval(Phys& p, float alpha) {
    float q0 = p.qg0, q1 = p.qg1, q2 = p.qg2;
    for (int c = 0; c < p.n_rows; ++c) {
        float ja = p.Jaref[c], jv = p.jv[c], D = p.Dd[c];
        if (ja + alpha * jv < 0.0f) {
            q0 += D * (0.5f * ja * ja);
            q1 += D * (jv * ja);
            q2 += D * (0.5f * jv * jv);
        }
    }
    LsPt r;
    r.alpha = alpha;
    r.cost = alpha * alpha * q2 + alpha * q1 + q0;
    r.d0 = 2.0f * alpha * q2 + q1;
    r.d1 = 2.0f * q2;
    if (r.d1 <= 0.0f) r.d1 = EPSF;
    p.ls_it += 1;
    return r;
}
G2_HD void ls_eval3(Phys& p, const float* a, LsPt* out) {
    float t[3][3];
    for (int k = 0; k < 3; ++k) { t[k][0] = p.qg0; t[k][1] = p.qg1; t[k][2] = p.qg2; }
    for (int c = 0; c < p.n_rows; ++c) {
        float ja = p.Jaref[c], jv = p.jv[c], D = p.Dd[c];
        float f0 = D * (0.5f * ja * ja), f1 = D * (jv * ja), f2 = D * (0.5f * jv * jv);
        for (int k = 0; k < 3; ++k) {
            if (ja + a[k] * jv < 0.0f) { t[k][0] += f0; t[k][1] += f1; t[k][2] += f2; }
        }
    }
    for (int k = 0; k < 3; ++k) {
        out[k].alpha = a[k];
        out[k].cost = a[k] * a[k] * t[k][2] + a[k] * t[k][1] + t[k][0];
        out[k].d0 = 2.0f * a[k] * t[k][2] + t[k][1];
        out[k].d1 = 2.0f * t[k][2];
        if (out[k].d1 <= 0.0f) out[k].d1 = EPSF;
    }
    p.ls_it += 3;
}
G2_HD int bracket_update(LsPt& pt, const LsPt* cand, float& next_alpha) {
    int flag = 0;
    for (int i = 0; i < 3; ++i) {
        if (pt.d0 < 0.0f && cand[i].d0 < 0.0f && pt.d0 < cand[i].d0) { pt = cand[i]; flag = 1; }
        else if (pt.d0 > 0.0f && cand[i].d0 > 0.0f && pt.d0 > cand[i].d0) { pt = cand[i]; flag = 2; }
    }
    next_alpha = pt.alpha;
    if (flag > 0) next_alpha = pt.alpha - pt.d0 / pt.d1;
    return flag;
}

// exact line search along p.search; fills p.mv / p.jv; returns the step length
G2_HDN float linesearch(const Go2Model& m, Phys& p) {
    float sn = 0.0f;
    for (int d = 0; d < NV; ++d) sn += p.search[d] * p.search[d];
    sn = sqrtf(sn);
    float gtol = SOLVER_TOL * LS_TOL * sn * (m.meaninertia * (float)NV);
    p.ls_it = 0;
    if (sn < EPSF) return 0.0f;
    for (int i = 0; i < NV; ++i) {
        float s = 0.0f;
        for (int j = 0; j < NV; ++j) s += Mat(p, i, j) * p.search[j];
        p.mv[i] = s;
    }
    for (int c = 0; c < p.n_rows; ++c) {
        const float* Jr = p.J[c];
        float s = 0.0f;
        for (int d = 0; d < NV; ++d) s += Jr[d] * p.search[d];
        p.jv[c] = s;
    }
    float q1 = 0.0f, q2 = 0.0f;
    for (int d = 0; d < NV; ++d) {
        q1 += p.search[d] * p.Ma[d] - p.search[d] * p.force[d];
        q2 += 0.5f * p.search[d] * p.mv[d];
    }
    p.qg0 = p.gauss; p.qg1 = q1; p.qg2 = q2;
    LsPt p0 = ls_eval(p, 0.0f);
    LsPt p1 = ls_eval(p, p0.alpha - p0.d0 / p0.d1);
    if (p0.cost < p1.cost) p1 = p0;
    if (fabsf(p1.d0) < gtol) return p1.alpha;
    float dir = p1.d0 < 0.0f ? 1.0f : -1.0f;
    bool p2upd = false;
    LsPt p2 = p1;
    while (p1.d0 * dir <= -gtol && p.ls_it < LS_ITERS) {
        p2 = p1;
        p2upd = true;
        p1 = ls_eval(p, p1.alpha - p1.d0 / p1.d1);
        if (fabsf(p1.d0) < gtol) return p1.alpha;
    }
    if (p.ls_it >= LS_ITERS) return p1.alpha;
    if (!p2upd) return p1.alpha;
    float a[3] = {p1.alpha - p1.d0 / p1.d1, p1.alpha, (p1.alpha + p2.alpha) * 0.5f};
    while (p.ls_it < LS_ITERS) {
        LsPt cand[3];
        ls_eval3(p, a, cand);
        bool found = false;
        float ba = 0.0f, bc = 0.0f;
        for (int i = 0; i < 3; ++i) {
            if (fabsf(cand[i].d0) < gtol && (!found || cand[i].cost < bc)) { ba = cand[i].alpha; bc = cand[i].cost; found = true; }
        }
        if (found) return ba;
        float n1, n2;
        int b1 = bracket_update(p1, cand, n1);
        int b2 = bracket_update(p2, cand, n2);
        if (b1 == 0 && b2 == 0) return a[2];
        a[0] = n1; a[1] = n2; a[2] = (p1.alpha + p2.alpha) * 0.5f;
    }
    if (p1.cost <= p2.cost && p1.cost < p0.cost) return p1.alpha;
    if (p2.cost <= p1.cost && p2.cost < p0.cost) return p2.alpha;
    return 0.0f;
}

// Newton solve for the constrained acceleration; writes p.acc, warm start, per-link contact forces.
G2_HDN void solve_constraints(const Go2Model& m, Phys& p) {
    bool ws = p.n_rows > 0 && p.warm;
    for (int d = 0; d < NV; ++d) p.qacc[d] = ws ? p.qacc_ws[d] : p.acc_smooth[d];
    for (int i = 0; i < NV; ++i) {
        float s = 0.0f;
        for (int j = 0; j < NV; ++j) s += Mat(p, i, j) * p.qacc[j];
        p.Ma[i] = s;
    }
    for (int c = 0; c < p.n_rows; ++c) {
        const float* Jr = p.J[c];
        float s = -p.aref[c];
        for (int d = 0; d < NV; ++d) s += Jr[d] * p.qacc[d];
        p.Jaref[c] = s;
    }
    update_constraint(p);
    p.n_newton = 0;
    if (p.n_rows > 0) {
        hessian_factor(p);
        update_gradient(p);
        for (int d = 0; d < NV; ++d) p.search[d] = -p.Mgrad[d];
        float tol = m.meaninertia * (float)NV * SOLVER_TOL;
        for (int it = 0; it < SOLVER_ITERS; ++it) {
            p.n_newton++;
            float alpha = linesearch(m, p);
            if (fabsf(alpha) < EPSF) break;
            for (int d = 0; d < NV; ++d) {
                p.qacc[d] += p.search[d] * alpha;
                p.Ma[d] += p.mv[d] * alpha;
            }
            for (int c = 0; c < p.n_rows; ++c) p.Jaref[c] += p.jv[c] * alpha;
            update_constraint(p);
            if (!hessian_update(p)) hessian_factor(p);
            update_gradient(p);
            float gn = 0.0f;
            for (int d = 0; d < NV; ++d) gn += p.grad[d] * p.grad[d];
            gn = sqrtf(gn);
            bool improved = gn > tol && (p.prev_cost - p.cost) > tol;
            if (!improved) break;
            for (int d = 0; d < NV; ++d) p.search[d] = -p.Mgrad[d];
        }
    }
    bool bad = false;
    for (int d = 0; d < NV; ++d) {
        p.acc[d] = p.qacc[d];
        p.qacc_ws[d] = p.qacc[d];
        bad = bad || !(p.qacc[d] == p.qacc[d]);
    }
    if (bad) p.err |= ERR_NAN_FORCE;
    p.warm = 1;
    for (int l = 0; l < NLINK; ++l) p.cforce[l] = mk3(0, 0, 0);
    for (int ic = 0; ic < p.n_contacts; ++ic) {
        const Contact& c = p.con[ic];
        v3 d1, d2;
        orthobasis(c.normal, d1, d2);
        v3 f = mk3(0, 0, 0);
        for (int k = 0; k < 4; ++k) {
            v3 d = (k < 2 ? d1 : d2) * (float)(2 * (k % 2) - 1);
            f = f + (d * c.mu - c.normal) * p.efc[ic * 4 + k];
        }
        if (c.la >= 0) p.cforce[c.la] = p.cforce[c.la] - f;
        if (c.lb >= 0) p.cforce[c.lb] = p.cforce[c.lb] + f;
    }
}

// One rigid substep: smooth dynamics -> collision -> constraints -> integrate -> refresh kinematics.
G2_HDN void substep(const Go2Model& m, Phys& p, float friction, const float* mass_shift, const float* com_shift) {
    smooth_dynamics(m, p);
    detect_collisions(m, p, friction);
    assemble_constraints(m, p);
    solve_constraints(m, p);
    integrate(m, p);
    kinematics(m, mass_shift, com_shift, p);
    velocities(m, p);
}

}  // namespace go2
