// go2_b200 - per-environment constraint assembly and primal Newton solve (layout v3: one THREAD per env).
//
// Soft-constraint model of the reference (MuJoCo style): 4 pyramid rows per contact, 1 row per violated
// joint limit, convex cost 0.5 (Ma - f).(a - a0) + sum_active 0.5 D (J a - aref)^2 minimised by Newton
// steps with an exact line search and rank-1 Cholesky up/down-dates of H = M + J^T D J when the active
// set changes (constraint/solver.py:498-595, 1088-1143, 1286-1343, 1467-1494, 1632-1675, 1747-1765,
// 1888-2417, 2428-2966, 2974-3037).
//
// Written for SIMT convergence (ncu on the first thread-per-env kernel: 75 % of all issued instructions
// were solver code running with 4 of 32 lanes, because envs called the same helper from different call
// sites / loop trips):
//   * every dof loop is fully unrolled with compile-time packed-triangle indices -> vectors and the
//     Hessian being factored live in registers, no index arithmetic;
//   * the line search is a two-loop state machine with ONE evaluation site per loop;
//   * active-set flips are collected in a packed list while the constraint state is updated, and the rank-1
//     updates run "k-th flip of every lane" together instead of row by row;
//   * M keeps only its tree-structured nonzeros in products.
#pragma once
#include "go2_collide.cuh"

namespace go2 {

#if defined(__CUDACC__) && defined(G2_ROLLED)
#define G2_UNROLL _Pragma("unroll 1")
#elif defined(__CUDACC__)
#define G2_UNROLL _Pragma("unroll")
#else
#define G2_UNROLL
#endif
#if defined(__CUDACC__) && defined(G2_ROLLED) && !defined(G2_ROLLED_ALL)
#define G2_UNROLL_H _Pragma("unroll")
#else
#define G2_UNROLL_H G2_UNROLL
#endif

// M[i][j] is structurally nonzero iff one of the dofs is a base dof or both hinges belong to the same leg
G2_HD bool m_nz(int i, int j) { return i < 6 || j < 6 || (((i - 6) & 3) == ((j - 6) & 3)); }
G2_HD int trs(int i, int j) { return i >= j ? tri(i, j) : tri(j, i); }

// y = M x over the structural nonzeros (packed lower triangle)
G2_HD void mul_M(const float* M, const float* x, float* y) {
    G2_UNROLL
    for (int i = 0; i < NV; ++i) {
        float s = 0.0f;
        G2_UNROLL
        for (int j = 0; j < NV; ++j)
            if (m_nz(i, j)) s += M[trs(i, j)] * x[j];
        y[i] = s;
    }
}

G2_HDN void assemble_constraints(const Go2Model& m, Phys& p) {
    int r = 0;
    for (int ic = 0; ic < p.n_contacts; ++ic) {
        const Contact& c = p.con[ic];
        v3 d1, d2;
        orthobasis(c.normal, d1, d2);
        float invw = (c.la >= 0 ? m.invweight_link[c.la] : 0.0f) + (c.lb >= 0 ? m.invweight_link[c.lb] : 0.0f);
        v3 tp = c.pos - p.root_com;
        v3 n[4];
        float jqv[4];
        G2_UNROLL
        for (int k = 0; k < 4; ++k) {
            v3 d = (k < 2 ? d1 : d2) * (float)(2 * (k % 2) - 1);
            n[k] = d * c.mu - c.normal;
            jqv[k] = 0.0f;
            float* Jr = p.J[r + k];
            G2_UNROLL
            for (int i = 0; i < NV; ++i) Jr[i] = 0.0f;
        }
        for (int side = 0; side < 2; ++side) {
            float sgn = side == 0 ? -1.0f : 1.0f;
            int l = side == 0 ? c.la : c.lb;
            while (l >= 0) {
                int d0 = l == 0 ? 0 : 5 + l, nd = l == 0 ? 6 : 1;
                for (int i = d0 + nd - 1; i >= d0; --i) {
                    v3 vel = p.cdof_v[i] - cross(tp, p.cdof_a[i]);
                    float qv = p.qvel[i];
                    G2_UNROLL
                    for (int k = 0; k < 4; ++k) {
                        float j = sgn * dot(vel, n[k]);
                        jqv[k] += j * qv;
                        p.J[r + k][i] += j;
                    }
                }
                l = l == 0 ? -1 : m.parent[l];
            }
        }
        G2_UNROLL
        for (int k = 0; k < 4; ++k) {
            float imp, ar;
            imp_aref(m.sol_params, -c.pen, jqv[k], -c.pen, imp, ar);
            float diag = invw + c.mu * c.mu * invw;
            diag *= 2.0f * c.mu * c.mu * (1.0f - imp) / imp;
            diag = fmaxf(diag, EPSF);
            p.aref[r + k] = ar;
            p.Dd[r + k] = 1.0f / diag;
        }
        r += 4;
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
            G2_UNROLL
            for (int i = 0; i < NV; ++i) Jr[i] = 0.0f;
            Jr[d] = j;
            p.aref[r] = ar;
            p.Dd[r] = 1.0f / diag;
            ++r;
        }
    }
    p.n_rows = r;
}

// Active set, constraint forces, qfc = J^T f, cost.  Rows whose active flag changed are appended to a packed
// list (8 bits per row index, newest in the low byte); `nfl` counts all of them.
G2_HD void update_constraint(Phys& p, const float* qacc, const float* Ma, float* qfc, unsigned long long& flist, int& nfl) {
    p.prev_cost = p.cost;
    G2_UNROLL
    for (int d = 0; d < NV; ++d) qfc[d] = 0.0f;
    float ccost = 0.0f;
    flist = 0ull;
    nfl = 0;
    int n = p.n_rows;
    for (int c = 0; c < n; ++c) {
        float ja = p.Jaref[c], D = p.Dd[c];
        bool a = ja < 0.0f;
        bool old = p.act[c] != 0;
        p.act[c] = a ? 1 : 0;
        float f = a ? -ja * D : 0.0f;
        p.efc[c] = f;
        if (a) {
            const float* Jr = p.J[c];
            G2_UNROLL
            for (int d = 0; d < NV; ++d) qfc[d] += Jr[d] * f;
            ccost += 0.5f * ja * ja * D;
        }
        if (a != old) {
            flist = (flist << 8) | (unsigned long long)c;
            nfl++;
        }
    }
    float gauss = 0.0f;
    G2_UNROLL
    for (int d = 0; d < NV; ++d) gauss += 0.5f * (Ma[d] - p.force[d]) * (qacc[d] - p.acc_smooth[d]);
    p.gauss = gauss;
    p.cost = gauss + ccost;
}

// G2_H_REGS (experiment): the Cholesky factor lives in a function-local array of solve_constraints and the helpers are
// force-inlined, so that with 255 registers the compiler can keep it in registers across the Newton loop.
#ifdef G2_H_REGS
#define G2_HSOLVE G2_HD
#else
#define G2_HSOLVE G2_HDN
#endif

// H = M + J^T D J over active rows, then dense Cholesky H = L L^T; assembled and factored in registers,
// L written to p.H once.
G2_HSOLVE void hessian_factor(Phys& p, float* Hs) {
#ifdef G2_H_REGS
    float* H = Hs;
#else
    float H[NTRI];
#endif
    G2_UNROLL_H
    for (int k = 0; k < NTRI; ++k) H[k] = p.M[k];
    int n = p.n_rows;
    for (int c = 0; c < n; ++c) {
        if (!p.act[c]) continue;
        const float* Jr = p.J[c];
        float D = p.Dd[c];
        float jr[NV];
        G2_UNROLL_H
        for (int i = 0; i < NV; ++i) jr[i] = Jr[i];
        G2_UNROLL_H
        for (int i = 0; i < NV; ++i) {
            float s = jr[i] * D;
            G2_UNROLL_H
            for (int j = 0; j <= i; ++j) H[tri(i, j)] += jr[j] * s;
        }
    }
    G2_UNROLL_H
    for (int i = 0; i < NV; ++i) {
        float t = H[tri(i, i)];
        G2_UNROLL_H
        for (int j = 0; j < i; ++j) t -= H[tri(i, j)] * H[tri(i, j)];
        float lii = sqrtf(fmaxf(t, EPSF));
        H[tri(i, i)] = lii;
        float inv = 1.0f / lii;
        G2_UNROLL_H
        for (int j = i + 1; j < NV; ++j) {
            float dt = 0.0f;
            G2_UNROLL_H
            for (int k = 0; k < i; ++k) dt += H[tri(j, k)] * H[tri(i, k)];
            H[tri(j, i)] = (H[tri(j, i)] - dt) * inv;
        }
    }
#ifndef G2_H_REGS
    G2_UNROLL_H
    for (int k = 0; k < NTRI; ++k) Hs[k] = H[k];
#endif
}

// rank-1 up/down-date of the Cholesky factor for row c; returns false if a downdate degenerated
G2_HSOLVE bool hessian_rank1(Phys& p, float* Hs, int c, float sgn) {
    float vec[NV];
    float sd = sqrtf(p.Dd[c]);
    const float* Jr = p.J[c];
    G2_UNROLL
    for (int d = 0; d < NV; ++d) vec[d] = Jr[d] * sd;
    bool ok = true;
    G2_UNROLL
    for (int k = 0; k < NV; ++k) {
        if (ok && fabsf(vec[k]) > EPSF) {
            float Lkk = Hs[tri(k, k)];
            float t = Lkk * Lkk + sgn * vec[k] * vec[k];
            if (t < EPSF) {
                ok = false;
            } else {
                float r = sqrtf(t), cc = r / Lkk, ci = 1.0f / cc, s = vec[k] / Lkk;
                Hs[tri(k, k)] = r;
                G2_UNROLL
                for (int i = k + 1; i < NV; ++i) {
                    float h = (Hs[tri(i, k)] + s * vec[i] * sgn) * ci;
                    Hs[tri(i, k)] = h;
                    vec[i] = vec[i] * cc - s * h;
                }
            }
        }
    }
    return ok;
}

// grad = Ma - force - qfc;  Mgrad = H^-1 grad (two triangular solves with the Cholesky factor)
G2_HD void update_gradient(Phys& p, const float* Hs, const float* Ma, const float* qfc, float* grad, float* Mgrad) {
    G2_UNROLL
    for (int d = 0; d < NV; ++d) grad[d] = Ma[d] - p.force[d] - qfc[d];
    G2_UNROLL
    for (int i = 0; i < NV; ++i) {
        float c = grad[i];
        G2_UNROLL
        for (int j = 0; j < i; ++j) c -= Hs[tri(i, j)] * Mgrad[j];
        Mgrad[i] = c / Hs[tri(i, i)];
    }
    G2_UNROLL
    for (int i = NV - 1; i >= 0; --i) {
        float c = Mgrad[i];
        G2_UNROLL
        for (int j = i + 1; j < NV; ++j) c -= Hs[tri(j, i)] * Mgrad[j];
        Mgrad[i] = c / Hs[tri(i, i)];
    }
}

#if defined(GO2_LS_STATS) && !defined(__CUDA_ARCH__)
// host-only debugging counters of the line search (tests/host_emul with -DGO2_LS_STATS)
static long g_ls_calls = 0, g_lsA[64] = {0}, g_lsB[64] = {0}, g_lsend[4] = {0};
static long g_nfl[64] = {0}, g_refac[3] = {0};
#define G2_LS_COUNT(nA, nB, k) (g_ls_calls++, g_lsA[nA]++, g_lsB[nB]++, g_lsend[k]++)
#define G2_NT_COUNT(nfl, over, fail) (g_nfl[(nfl) < 63 ? (nfl) : 63]++, g_refac[0] += (over), g_refac[1] += (fail), g_refac[2]++)
#else
#define G2_NT_COUNT(nfl, over, fail)
#define G2_LS_COUNT(nA, nB, k)
#endif
constexpr int LS_REG = 16;  // rows of the line search kept in registers (4 contacts)
struct LsPt {
    float alpha, cost, d0, d1;
};
G2_HD LsPt ls_point(float alpha, float q0, float q1, float q2) {
    LsPt r;
    r.alpha = alpha;
    r.cost = alpha * alpha * q2 + alpha * q1 + q0;
    r.d0 = 2.0f * alpha * q2 + q1;
    r.d1 = 2.0f * q2;
    if (r.d1 <= 0.0f) r.d1 = EPSF;
    return r;
}
G2_HD int bracket_update(LsPt& pt, const LsPt* cand, float& next_alpha) {
    int flag = 0;
    G2_UNROLL
    for (int i = 0; i < 3; ++i) {
        if (pt.d0 < 0.0f && cand[i].d0 < 0.0f && pt.d0 < cand[i].d0) { pt = cand[i]; flag = 1; }
        else if (pt.d0 > 0.0f && cand[i].d0 > 0.0f && pt.d0 > cand[i].d0) { pt = cand[i]; flag = 2; }
    }
    next_alpha = pt.alpha;
    if (flag > 0) next_alpha = pt.alpha - pt.d0 / pt.d1;
    return flag;
}

// Exact line search along `search` (constraint/solver.py:2246-2417).  Fills mv = M search and p.jv = J search;
// returns the step length.  Same decisions as the reference, arranged as two loops with one evaluation site each.
G2_HD float linesearch(const Go2Model& m, Phys& p, const float* search, const float* Ma, float* mv) {
    float sn = 0.0f;
    G2_UNROLL
    for (int d = 0; d < NV; ++d) sn += search[d] * search[d];
    sn = sqrtf(sn);
    float gtol = SOLVER_TOL * LS_TOL * sn * (m.meaninertia * (float)NV);
    if (sn < EPSF) return 0.0f;
    mul_M(p.M, search, mv);
    int n = p.n_rows;
    // the first LS_REG rows are cached in registers for the evaluations (rows >= n contribute exact zeros);
    // any further rows (more than 4 contacts) are streamed from thread-local memory
    float rja[LS_REG], rjv[LS_REG], rD[LS_REG];
    G2_UNROLL
    for (int c = 0; c < LS_REG; ++c) {
        float s = 0.0f, ja = 0.0f, D = 0.0f;
        if (c < n) {
            const float* Jr = p.J[c];
            G2_UNROLL
            for (int d = 0; d < NV; ++d) s += Jr[d] * search[d];
            p.jv[c] = s;
            ja = p.Jaref[c];
            D = p.Dd[c];
        }
        rja[c] = ja; rjv[c] = s; rD[c] = D;
    }
    for (int c = LS_REG; c < n; ++c) {
        const float* Jr = p.J[c];
        float s = 0.0f;
        G2_UNROLL
        for (int d = 0; d < NV; ++d) s += Jr[d] * search[d];
        p.jv[c] = s;
    }
    float g1 = 0.0f, g2 = 0.0f;
    G2_UNROLL
    for (int d = 0; d < NV; ++d) {
        g1 += search[d] * Ma[d] - search[d] * p.force[d];
        g2 += 0.5f * search[d] * mv[d];
    }
    const float g0 = p.gauss;
    int ls_it = 0;
    // ---- loop A: p0 at alpha = 0, p1 at the Newton step, then Newton steps until the derivative changes sign
    LsPt p0, p1, p2;
    p0.alpha = p0.cost = p0.d0 = 0.0f; p0.d1 = 1.0f;
    p1 = p0; p2 = p0;
    float dir = 1.0f, alpha = 0.0f, result = 0.0f;
    bool p2upd = false, finished = false;
    int stage = 0;
    for (;;) {
        float q0 = g0, q1 = g1, q2 = g2;
        G2_UNROLL
        for (int c = 0; c < LS_REG; ++c) {
            float ja = rja[c], jv = rjv[c], D = rD[c];
            if (ja + alpha * jv < 0.0f) {
                q0 += D * (0.5f * ja * ja);
                q1 += D * (jv * ja);
                q2 += D * (0.5f * jv * jv);
            }
        }
        for (int c = LS_REG; c < n; ++c) {
            float ja = p.Jaref[c], jv = p.jv[c], D = p.Dd[c];
            if (ja + alpha * jv < 0.0f) {
                q0 += D * (0.5f * ja * ja);
                q1 += D * (jv * ja);
                q2 += D * (0.5f * jv * jv);
            }
        }
        LsPt e = ls_point(alpha, q0, q1, q2);
        ls_it += 1;
        bool more;
        if (stage == 0) {
            p0 = e;
            alpha = p0.alpha - p0.d0 / p0.d1;
            stage = 1;
            more = true;
        } else {
            if (stage == 1) {
                p1 = e;
                if (p0.cost < p1.cost) p1 = p0;
            } else {
                p1 = e;
            }
            if (fabsf(p1.d0) < gtol) {
                result = p1.alpha;
                finished = true;
                more = false;
            } else {
                if (stage == 1) {
                    dir = p1.d0 < 0.0f ? 1.0f : -1.0f;
                    p2 = p1;
                    stage = 2;
                }
                if (p1.d0 * dir <= -gtol && ls_it < LS_ITERS) {
                    p2 = p1;
                    p2upd = true;
                    alpha = p1.alpha - p1.d0 / p1.d1;
                    more = true;
                } else {
                    more = false;
                }
            }
        }
        if (!more) break;
    }
    if (!finished && (ls_it >= LS_ITERS || !p2upd)) {
        result = p1.alpha;
        finished = true;
    }
    // ---- loop B: three-point refinement of the bracket [p1, p2]
    float a[3] = {p1.alpha - p1.d0 / p1.d1, p1.alpha, (p1.alpha + p2.alpha) * 0.5f};
    int nA = ls_it, nB = 0, endk = finished ? 0 : 3;
    (void)nA; (void)nB; (void)endk;
    while (!finished && ls_it < LS_ITERS) {
        nB++;
        float t[3][3];
        G2_UNROLL
        for (int k = 0; k < 3; ++k) { t[k][0] = g0; t[k][1] = g1; t[k][2] = g2; }
        G2_UNROLL
        for (int c = 0; c < LS_REG; ++c) {
            float ja = rja[c], jv = rjv[c], D = rD[c];
            float f0 = D * (0.5f * ja * ja), f1 = D * (jv * ja), f2 = D * (0.5f * jv * jv);
            G2_UNROLL
            for (int k = 0; k < 3; ++k) {
                if (ja + a[k] * jv < 0.0f) { t[k][0] += f0; t[k][1] += f1; t[k][2] += f2; }
            }
        }
        for (int c = LS_REG; c < n; ++c) {
            float ja = p.Jaref[c], jv = p.jv[c], D = p.Dd[c];
            float f0 = D * (0.5f * ja * ja), f1 = D * (jv * ja), f2 = D * (0.5f * jv * jv);
            G2_UNROLL
            for (int k = 0; k < 3; ++k) {
                if (ja + a[k] * jv < 0.0f) { t[k][0] += f0; t[k][1] += f1; t[k][2] += f2; }
            }
        }
        LsPt cand[3];
        G2_UNROLL
        for (int k = 0; k < 3; ++k) cand[k] = ls_point(a[k], t[k][0], t[k][1], t[k][2]);
        ls_it += 3;
        bool found = false;
        float ba = 0.0f, bc = 0.0f;
        G2_UNROLL
        for (int i = 0; i < 3; ++i) {
            if (fabsf(cand[i].d0) < gtol && (!found || cand[i].cost < bc)) { ba = cand[i].alpha; bc = cand[i].cost; found = true; }
        }
        if (found) {
            result = ba;
            finished = true;
            endk = 1;
        } else {
            float n1, n2;
            int b1 = bracket_update(p1, cand, n1);
            int b2 = bracket_update(p2, cand, n2);
            if (b1 == 0 && b2 == 0) {
                result = a[2];
                finished = true;
                endk = 2;
            } else {
                a[0] = n1; a[1] = n2; a[2] = (p1.alpha + p2.alpha) * 0.5f;
            }
        }
    }
    G2_LS_COUNT(nA, nB, endk);
    if (!finished) {
        if (p1.cost <= p2.cost && p1.cost < p0.cost) result = p1.alpha;
        else if (p2.cost <= p1.cost && p2.cost < p0.cost) result = p2.alpha;
        else result = 0.0f;
    }
    return result;
}

constexpr int MAX_RANK1 = 6;  // more flips than this in one Newton iteration -> assemble + factor from scratch

// Block-wide "does any thread still iterate": keeps the warps of a block in the same Newton iteration so that
// they share instruction fetches (the solver body is ~60 KB of straight-line code).  Host: a single env.
#if defined(__CUDA_ARCH__) && defined(G2_WARP_ANY)
#define G2_BLOCK_ANY(x) (__any_sync(0xffffffffu, (x)))
#elif defined(__CUDA_ARCH__)
#define G2_BLOCK_ANY(x) (__syncthreads_or((x) ? 1 : 0) != 0)
#else
#define G2_BLOCK_ANY(x) (x)
#endif

// Newton solve for the constrained acceleration; writes p.acc, warm start, per-link contact forces.
// Called by EVERY thread of the block (`valid` = the thread owns an env) because the iteration loop contains
// block barriers.  Trip 0 of the loop is the initial evaluation (active set, Hessian factorisation, gradient);
// trips >= 1 are the Newton iterations of constraint/solver.py:2862-2938, each helper having ONE call site.
G2_HDN void solve_constraints(const Go2Model& m, Phys& p, bool valid) {
    const int n = valid ? p.n_rows : 0;
    float qacc[NV], Ma[NV], qfc[NV], grad[NV], Mgrad[NV], search[NV], mv[NV];
#ifdef G2_H_REGS
    float Hloc[NTRI];
    float* Hs = Hloc;
#else
    float* Hs = p.H;
#endif
    if (valid) {
        bool ws = n > 0 && p.warm;
        G2_UNROLL
        for (int d = 0; d < NV; ++d) qacc[d] = ws ? p.qacc_ws[d] : p.acc_smooth[d];
        if (n > 0) mul_M(p.M, qacc, Ma);
        for (int c = 0; c < n; ++c) {
            const float* Jr = p.J[c];
            float s = -p.aref[c];
            G2_UNROLL
            for (int d = 0; d < NV; ++d) s += Jr[d] * qacc[d];
            p.Jaref[c] = s;
            p.act[c] = 0;
        }
        p.cost = 0.0f;
    }
    int n_newton = 0;
    bool active = n > 0;
    const float tol = m.meaninertia * (float)NV * SOLVER_TOL;
    for (int it = 0; it <= SOLVER_ITERS; ++it) {
        if (!G2_BLOCK_ANY(active)) break;
        if (active) {
            bool first = it == 0;
            if (!first) {
                n_newton++;
                float alpha = linesearch(m, p, search, Ma, mv);
                if (fabsf(alpha) < EPSF) {
                    active = false;
                } else {
                    G2_UNROLL
                    for (int d = 0; d < NV; ++d) {
                        qacc[d] += search[d] * alpha;
                        Ma[d] += mv[d] * alpha;
                    }
                    for (int c = 0; c < n; ++c) p.Jaref[c] += p.jv[c] * alpha;
                }
            }
            if (active) {
                unsigned long long flist;
                int nfl;
                update_constraint(p, qacc, Ma, qfc, flist, nfl);
                // active-set changes: k-th flip of every lane together, oldest first (row order, like the reference)
                bool refactor = first || nfl > MAX_RANK1;
                for (int k = nfl - 1; k >= 0 && !refactor; --k) {
                    int c = (int)((flist >> (8 * k)) & 0xffull);
                    if (!hessian_rank1(p, Hs, c, p.act[c] ? 1.0f : -1.0f)) refactor = true;
                }
                if (!first) G2_NT_COUNT(nfl, nfl > MAX_RANK1 ? 1 : 0, (refactor && nfl <= MAX_RANK1) ? 1 : 0);
                if (refactor) hessian_factor(p, Hs);
                update_gradient(p, Hs, Ma, qfc, grad, Mgrad);
                if (!first) {
                    float gn = 0.0f;
                    G2_UNROLL
                    for (int d = 0; d < NV; ++d) gn += grad[d] * grad[d];
                    gn = sqrtf(gn);
                    bool improved = gn > tol && (p.prev_cost - p.cost) > tol;
                    if (!improved || it == SOLVER_ITERS) active = false;
                }
                G2_UNROLL
                for (int d = 0; d < NV; ++d) search[d] = -Mgrad[d];
            }
        }
    }
    if (!valid) return;
    p.n_newton = n_newton;
    bool bad = false;
    G2_UNROLL
    for (int d = 0; d < NV; ++d) {
        p.acc[d] = qacc[d];
        p.qacc_ws[d] = qacc[d];
        bad = bad || !(qacc[d] == qacc[d]);
    }
    if (bad) p.err |= ERR_NAN_FORCE;
    p.warm = 1;
    for (int l = 0; l < NLINK; ++l) p.cforce[l] = mk3(0, 0, 0);
    for (int ic = 0; ic < p.n_contacts; ++ic) {
        const Contact& c = p.con[ic];
        v3 d1, d2;
        orthobasis(c.normal, d1, d2);
        v3 f = mk3(0, 0, 0);
        G2_UNROLL
        for (int k = 0; k < 4; ++k) {
            v3 d = (k < 2 ? d1 : d2) * (float)(2 * (k % 2) - 1);
            f = f + (d * c.mu - c.normal) * p.efc[ic * 4 + k];
        }
        if (c.la >= 0) p.cforce[c.la] = p.cforce[c.la] - f;
        if (c.lb >= 0) p.cforce[c.lb] = p.cforce[c.lb] + f;
    }
}

}  // namespace go2
