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
//   * M keeps only its tree-structured nonzeros in products;
//   * ARROW mode (every row touches the 6 base dofs + ONE leg - true unless two different legs touch each other):
//     rows keep 9 Jacobian entries, and the Newton Hessian is factored with the legs eliminated first, which
//     makes the factor an arrowhead (four 3x3 leg blocks, four 6x3 base-leg blocks, one 6x6 base block = 117 of
//     171 entries); a rank-1 update touches 45 entries instead of 171.  Same minimiser, other rounding than
//     the reference's dense factor in dof order (constraint/solver.py:1088-1143), which the DENSE mode keeps.
#pragma once
#include "go2_collide.cuh"

namespace go2 {

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

// ---- row access, dense (J[c][18]) or arrow (Js[c][9] + leg id) ---------------------------------------
G2_HD float sel4(int k, float a, float b, float c, float d) {
    float lo = (k & 1) ? b : a, hi = (k & 1) ? d : c;
    return (k & 2) ? hi : lo;
}
// J[c] . x   (arrow: same summation order as the dense row with its zeros skipped -> bit-identical)
template <bool AR>
G2_HD float row_dot(const Phys& p, int c, const float* x, float s0) {
    float s = s0;
    if (AR) {
        const float* Jr = p.Js[c];
        int leg = p.rleg[c];
        G2_UNROLL
        for (int d = 0; d < 6; ++d) s += Jr[d] * x[d];
        s += Jr[6] * sel4(leg, x[6], x[7], x[8], x[9]);
        s += Jr[7] * sel4(leg, x[10], x[11], x[12], x[13]);
        s += Jr[8] * sel4(leg, x[14], x[15], x[16], x[17]);
    } else {
        const float* Jr = p.J[c];
        G2_UNROLL
        for (int d = 0; d < NV; ++d) s += Jr[d] * x[d];
    }
    return s;
}
// y += J[c]^T f
template <bool AR>
G2_HD void row_axpy(const Phys& p, int c, float f, float* y) {
    if (AR) {
        const float* Jr = p.Js[c];
        int leg = p.rleg[c];
        G2_UNROLL
        for (int d = 0; d < 6; ++d) y[d] += Jr[d] * f;
        G2_UNROLL
        for (int a = 0; a < 3; ++a) {
            float t = Jr[6 + a] * f;
            G2_UNROLL
            for (int l = 0; l < 4; ++l) y[6 + 4 * a + l] += (leg == l) ? t : 0.0f;
        }
    } else {
        const float* Jr = p.J[c];
        G2_UNROLL
        for (int d = 0; d < NV; ++d) y[d] += Jr[d] * f;
    }
}

template <bool AR>
G2_HDN void assemble_rows(const Go2Model& m, Phys& p) {
    constexpr int NC = AR ? 9 : NV;
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
            float* Jr = AR ? p.Js[r + k] : p.J[r + k];
            G2_UNROLL
            for (int i = 0; i < NC; ++i) Jr[i] = 0.0f;
        }
        int leg = 0;
        for (int side = 0; side < 2; ++side) {
            float sgn = side == 0 ? -1.0f : 1.0f;
            int l = side == 0 ? c.la : c.lb;
            if (l > 0) leg = (l - 1) & 3;
            while (l >= 0) {
                int d0 = l == 0 ? 0 : 5 + l, nd = l == 0 ? 6 : 1;
                for (int i = d0 + nd - 1; i >= d0; --i) {
                    v3 vel = p.cdof_v[i] - cross(tp, p.cdof_a[i]);
                    float qv = p.qvel[i];
                    int col = AR ? (i < 6 ? i : 6 + ((i - 6) >> 2)) : i;
                    G2_UNROLL
                    for (int k = 0; k < 4; ++k) {
                        float j = sgn * dot(vel, n[k]);
                        jqv[k] += j * qv;
                        if (AR) p.Js[r + k][col] += j;
                        else p.J[r + k][col] += j;
                    }
                }
                l = par_link(l);
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
            if (AR) p.rleg[r + k] = (uint8_t)leg;
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
            float* Jr = AR ? p.Js[r] : p.J[r];
            G2_UNROLL
            for (int i = 0; i < NC; ++i) Jr[i] = 0.0f;
            Jr[AR ? 6 + ((d - 6) >> 2) : d] = j;
            if (AR) p.rleg[r] = (uint8_t)((d - 6) & 3);
            p.aref[r] = ar;
            p.Dd[r] = 1.0f / diag;
            ++r;
        }
    }
    p.n_rows = r;
}

// Arrow-mode rows with the 4 pyramid rows of a contact built in registers (static columns: 6 base dofs, then the hip /
// thigh / calf of the contact's leg) and stored once.  Same operation order as assemble_rows (link -> root, side a
// then side b), so the two produce identical numbers.
G2_HDN void assemble_rows_a(const Go2Model& m, Phys& p) {
    int r = 0;
    for (int ic = 0; ic < p.n_contacts; ++ic) {
        const Contact& c = p.con[ic];
        v3 d1, d2;
        orthobasis(c.normal, d1, d2);
        float invw = (c.la >= 0 ? m.invweight_link[c.la] : 0.0f) + (c.lb >= 0 ? m.invweight_link[c.lb] : 0.0f);
        v3 tp = c.pos - p.root_com;
        v3 n[4];
        float jqv[4], jr[4][9];
        G2_UNROLL
        for (int k = 0; k < 4; ++k) {
            v3 d = (k < 2 ? d1 : d2) * (float)(2 * (k % 2) - 1);
            n[k] = d * c.mu - c.normal;
            jqv[k] = 0.0f;
            G2_UNROLL
            for (int i = 0; i < 9; ++i) jr[k][i] = 0.0f;
        }
        int leg = 0;
        G2_UNROLL
        for (int side = 0; side < 2; ++side) {
            const float sgn = side == 0 ? -1.0f : 1.0f;
            const int l = side == 0 ? c.la : c.lb;
            if (l >= 0) {
                const int depth = l == 0 ? 0 : 1 + ((l - 1) >> 2);  // hinges between the link and the base
                if (l > 0) leg = (l - 1) & 3;
                G2_UNROLL
                for (int col = 8; col >= 0; --col) {
                    if (col < 6 || col - 6 < depth) {
                        const int i = col < 6 ? col : 6 + 4 * (col - 6) + ((l - 1) & 3);
                        v3 vel = p.cdof_v[i] - cross(tp, p.cdof_a[i]);
                        float qv = p.qvel[i];
                        G2_UNROLL
                        for (int k = 0; k < 4; ++k) {
                            float j = sgn * dot(vel, n[k]);
                            jqv[k] += j * qv;
                            jr[k][col] += j;
                        }
                    }
                }
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
            p.rleg[r + k] = (uint8_t)leg;
            G2_UNROLL
            for (int i = 0; i < 9; ++i) p.Js[r + k][i] = jr[k][i];
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
            float* Jr = p.Js[r];
            G2_UNROLL
            for (int i = 0; i < 9; ++i) Jr[i] = 0.0f;
            Jr[6 + ((d - 6) >> 2)] = j;
            p.rleg[r] = (uint8_t)((d - 6) & 3);
            p.aref[r] = ar;
            p.Dd[r] = 1.0f / diag;
            ++r;
        }
    }
    p.n_rows = r;
}

// Picks the row format (arrow unless two different legs touch each other) and builds the rows.
G2_HD void assemble_constraints(const Go2Model& m, Phys& p, int force_dense) {
    int arrow = 1;
    for (int ic = 0; ic < p.n_contacts; ++ic) {
        int a = p.con[ic].la, b = p.con[ic].lb;
        if (a > 0 && b > 0 && ((a - 1) & 3) != ((b - 1) & 3)) arrow = 0;
    }
    if (force_dense) arrow = 0;
    p.arrow = arrow;
    if (arrow) assemble_rows_a(m, p);
    else assemble_rows<false>(m, p);
}

// Active set, constraint forces, qfc = J^T f, cost.  Rows whose active flag changed are appended to a packed
// list (8 bits per row index, newest in the low byte); `nfl` counts all of them.
template <bool AR>
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
            row_axpy<AR>(p, c, f, qfc);
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

#define G2_HSOLVE G2_HDN

// H = M + J^T D J over active rows, then dense Cholesky H = L L^T; assembled and factored in registers,
// L written to p.H once.
G2_HSOLVE void hessian_factor(Phys& p, float* Hs) {
    float H[NTRI];
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
    G2_UNROLL_H
    for (int k = 0; k < NTRI; ++k) Hs[k] = H[k];
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

// ---- arrow mode: H = M + J^T D J with the dofs ordered (leg 0, leg 1, leg 2, leg 3, base) ------------------
// Leg l owns dofs 6+l (hip), 10+l (thigh), 14+l (calf).  H has no leg-leg coupling between different legs
// (M: tree structure; rows: one leg each), so L = [diag(La_l), 0; Lb_l, Lc].
G2_HD void arrow_acc_row(const float* jr, float D, float* C, float* tA, float* tB) {
    G2_UNROLL
    for (int i = 0; i < 6; ++i) {
        float sb = jr[i] * D;
        G2_UNROLL
        for (int j = 0; j <= i; ++j) C[tri(i, j)] += jr[j] * sb;
        G2_UNROLL
        for (int a = 0; a < 3; ++a) tB[i * 3 + a] += jr[6 + a] * sb;
    }
    G2_UNROLL
    for (int a = 0; a < 3; ++a) {
        float sa = jr[6 + a] * D;
        G2_UNROLL
        for (int b = 0; b <= a; ++b) tA[tri(a, b)] += jr[6 + b] * sa;
    }
}

G2_HDN void hessian_factor_a(Phys& p) {
    float C[21];
    G2_UNROLL
    for (int k = 0; k < 21; ++k) C[k] = p.M[k];  // base block: dofs 0..5 = first 21 packed entries
    G2_UNROLL
    for (int l = 0; l < 4; ++l) {
        G2_UNROLL
        for (int a = 0; a < 3; ++a) {
            G2_UNROLL
            for (int b = 0; b <= a; ++b) p.Hl[l][tri(a, b)] = p.M[tri(6 + 4 * a + l, 6 + 4 * b + l)];
            G2_UNROLL
            for (int r = 0; r < 6; ++r) p.Hbl[l][r * 3 + a] = p.M[tri(6 + 4 * a + l, r)];
        }
    }
    // rows come in groups sharing a leg: 4 per contact, then single joint-limit rows
    const int n = p.n_rows, ncr = 4 * p.n_contacts;
    for (int c0 = 0; c0 < n; c0 += (c0 < ncr ? 4 : 1)) {
        float tA[6], tB[18];
        G2_UNROLL
        for (int k = 0; k < 6; ++k) tA[k] = 0.0f;
        G2_UNROLL
        for (int k = 0; k < 18; ++k) tB[k] = 0.0f;
        const int leg = p.rleg[c0];
        const bool grp = c0 < ncr;
        bool any = false;
        G2_UNROLL
        for (int k = 0; k < 4; ++k) {
            if ((k == 0 || grp) && p.act[c0 + k]) {
                float jr[9];
                G2_UNROLL
                for (int i = 0; i < 9; ++i) jr[i] = p.Js[c0 + k][i];
                arrow_acc_row(jr, p.Dd[c0 + k], C, tA, tB);
                any = true;
            }
        }
        if (any) {
            float* A = p.Hl[leg];
            float* Bm = p.Hbl[leg];
            G2_UNROLL
            for (int k = 0; k < 6; ++k) A[k] += tA[k];
            G2_UNROLL
            for (int k = 0; k < 18; ++k) Bm[k] += tB[k];
        }
    }
    G2_UNROLL
    for (int l = 0; l < 4; ++l) {
        float* A = p.Hl[l];
        float* Bm = p.Hbl[l];
        float l00 = sqrtf(fmaxf(A[0], EPSF)), i00 = 1.0f / l00;
        float l10 = A[1] * i00, l20 = A[3] * i00;
        float l11 = sqrtf(fmaxf(A[2] - l10 * l10, EPSF)), i11 = 1.0f / l11;
        float l21 = (A[4] - l20 * l10) * i11;
        float l22 = sqrtf(fmaxf(A[5] - l20 * l20 - l21 * l21, EPSF)), i22 = 1.0f / l22;
        A[0] = l00; A[1] = l10; A[2] = l11; A[3] = l20; A[4] = l21; A[5] = l22;
        float X[6][3];
        G2_UNROLL
        for (int r = 0; r < 6; ++r) {
            float x0 = Bm[r * 3 + 0] * i00;
            float x1 = (Bm[r * 3 + 1] - x0 * l10) * i11;
            float x2 = (Bm[r * 3 + 2] - x0 * l20 - x1 * l21) * i22;
            Bm[r * 3 + 0] = x0; Bm[r * 3 + 1] = x1; Bm[r * 3 + 2] = x2;
            X[r][0] = x0; X[r][1] = x1; X[r][2] = x2;
        }
        G2_UNROLL
        for (int i = 0; i < 6; ++i) {
            G2_UNROLL
            for (int j = 0; j <= i; ++j) C[tri(i, j)] -= X[i][0] * X[j][0] + X[i][1] * X[j][1] + X[i][2] * X[j][2];
        }
    }
    G2_UNROLL
    for (int i = 0; i < 6; ++i) {
        float t = C[tri(i, i)];
        G2_UNROLL
        for (int j = 0; j < i; ++j) t -= C[tri(i, j)] * C[tri(i, j)];
        float lii = sqrtf(fmaxf(t, EPSF));
        C[tri(i, i)] = lii;
        float inv = 1.0f / lii;
        G2_UNROLL
        for (int j = i + 1; j < 6; ++j) {
            float dt = 0.0f;
            G2_UNROLL
            for (int k = 0; k < i; ++k) dt += C[tri(j, k)] * C[tri(i, k)];
            C[tri(j, i)] = (C[tri(j, i)] - dt) * inv;
        }
    }
    G2_UNROLL
    for (int k = 0; k < 21; ++k) p.Hbb[k] = C[k];
}

// rank-1 up/down-date of the arrow factor for row c: the leg's 3 columns, then the 6 base columns.  The 45 entries it
// touches are staged in registers first (one burst of independent loads instead of a load per dependent step).
G2_HDN bool hessian_rank1_a(Phys& p, int c, float sgn) {
    float vl[3], vb[6], A[6], Bm[18], Cb[21];
    const float sd = sqrtf(p.Dd[c]);
    const float* Jr = p.Js[c];
    float* Ap = p.Hl[p.rleg[c]];
    float* Bp = p.Hbl[p.rleg[c]];
    G2_UNROLL
    for (int d = 0; d < 6; ++d) vb[d] = Jr[d] * sd;
    G2_UNROLL
    for (int a = 0; a < 3; ++a) vl[a] = Jr[6 + a] * sd;
    G2_UNROLL
    for (int k = 0; k < 6; ++k) A[k] = Ap[k];
    G2_UNROLL
    for (int k = 0; k < 18; ++k) Bm[k] = Bp[k];
    G2_UNROLL
    for (int k = 0; k < 21; ++k) Cb[k] = p.Hbb[k];
    bool ok = true;
    G2_UNROLL
    for (int k = 0; k < 3; ++k) {
        if (ok && fabsf(vl[k]) > EPSF) {
            float Lkk = A[tri(k, k)];
            float t = Lkk * Lkk + sgn * vl[k] * vl[k];
            if (t < EPSF) {
                ok = false;
            } else {
                float r = sqrtf(t), cc = r / Lkk, ci = 1.0f / cc, s = vl[k] / Lkk;
                A[tri(k, k)] = r;
                G2_UNROLL
                for (int i = k + 1; i < 3; ++i) {
                    float h = (A[tri(i, k)] + s * vl[i] * sgn) * ci;
                    A[tri(i, k)] = h;
                    vl[i] = vl[i] * cc - s * h;
                }
                G2_UNROLL
                for (int i = 0; i < 6; ++i) {
                    float h = (Bm[i * 3 + k] + s * vb[i] * sgn) * ci;
                    Bm[i * 3 + k] = h;
                    vb[i] = vb[i] * cc - s * h;
                }
            }
        }
    }
    G2_UNROLL
    for (int k = 0; k < 6; ++k) {
        if (ok && fabsf(vb[k]) > EPSF) {
            float Lkk = Cb[tri(k, k)];
            float t = Lkk * Lkk + sgn * vb[k] * vb[k];
            if (t < EPSF) {
                ok = false;
            } else {
                float r = sqrtf(t), cc = r / Lkk, ci = 1.0f / cc, s = vb[k] / Lkk;
                Cb[tri(k, k)] = r;
                G2_UNROLL
                for (int i = k + 1; i < 6; ++i) {
                    float h = (Cb[tri(i, k)] + s * vb[i] * sgn) * ci;
                    Cb[tri(i, k)] = h;
                    vb[i] = vb[i] * cc - s * h;
                }
            }
        }
    }
    // a degenerate downdate is followed by a refactorisation from scratch, so partial results may be stored too
    G2_UNROLL
    for (int k = 0; k < 6; ++k) Ap[k] = A[k];
    G2_UNROLL
    for (int k = 0; k < 18; ++k) Bp[k] = Bm[k];
    G2_UNROLL
    for (int k = 0; k < 21; ++k) p.Hbb[k] = Cb[k];
    return ok;
}

// grad = Ma - force - qfc;  Mgrad = H^-1 grad through the arrow factor
G2_HD void update_gradient_a(Phys& p, const float* Ma, const float* qfc, float* grad, float* Mgrad) {
    G2_UNROLL
    for (int d = 0; d < NV; ++d) grad[d] = Ma[d] - p.force[d] - qfc[d];
    float yb[6];
    G2_UNROLL
    for (int r = 0; r < 6; ++r) yb[r] = grad[r];
    G2_UNROLL
    for (int l = 0; l < 4; ++l) {
        const float* A = p.Hl[l];
        const float* Bm = p.Hbl[l];
        float y0 = grad[6 + l] / A[0];
        float y1 = (grad[10 + l] - A[1] * y0) / A[2];
        float y2 = (grad[14 + l] - A[3] * y0 - A[4] * y1) / A[5];
        Mgrad[6 + l] = y0; Mgrad[10 + l] = y1; Mgrad[14 + l] = y2;
        G2_UNROLL
        for (int r = 0; r < 6; ++r) yb[r] -= Bm[r * 3 + 0] * y0 + Bm[r * 3 + 1] * y1 + Bm[r * 3 + 2] * y2;
    }
    G2_UNROLL
    for (int i = 0; i < 6; ++i) {
        float c = yb[i];
        G2_UNROLL
        for (int j = 0; j < i; ++j) c -= p.Hbb[tri(i, j)] * yb[j];
        yb[i] = c / p.Hbb[tri(i, i)];
    }
    G2_UNROLL
    for (int i = 5; i >= 0; --i) {
        float c = yb[i];
        G2_UNROLL
        for (int j = i + 1; j < 6; ++j) c -= p.Hbb[tri(j, i)] * yb[j];
        yb[i] = c / p.Hbb[tri(i, i)];
    }
    G2_UNROLL
    for (int r = 0; r < 6; ++r) Mgrad[r] = yb[r];
    G2_UNROLL
    for (int l = 0; l < 4; ++l) {
        const float* A = p.Hl[l];
        const float* Bm = p.Hbl[l];
        float z0 = Mgrad[6 + l], z1 = Mgrad[10 + l], z2 = Mgrad[14 + l];
        G2_UNROLL
        for (int r = 0; r < 6; ++r) {
            z0 -= Bm[r * 3 + 0] * yb[r];
            z1 -= Bm[r * 3 + 1] * yb[r];
            z2 -= Bm[r * 3 + 2] * yb[r];
        }
        float x2 = z2 / A[5];
        float x1 = (z1 - A[4] * x2) / A[2];
        float x0 = (z0 - A[1] * x1 - A[3] * x2) / A[0];
        Mgrad[6 + l] = x0; Mgrad[10 + l] = x1; Mgrad[14 + l] = x2;
    }
}

// The line search stops on |derivative| < gtol with gtol at fp32 rounding level (LS_TOL * SOLVER_TOL).  That test is
// met through the exact cancellation of SEPARATELY rounded products; with fused multiply-adds the residual never
// cancels and 13 % of the searches run to the 50-evaluation cap (measured, host build with -mfma: 0 % without).  The
// evaluations therefore use explicitly unfused arithmetic on the device, like the CPU builds (-ffp-contract=off).
#if defined(__CUDA_ARCH__)
G2_HD float mul_nf(float a, float b) { return __fmul_rn(a, b); }
G2_HD float add_nf(float a, float b) { return __fadd_rn(a, b); }
#else
G2_HD float mul_nf(float a, float b) { return a * b; }
G2_HD float add_nf(float a, float b) { return a + b; }
#endif
constexpr int LS_REG = 16;  // rows of the line search kept in registers (16 = 4 contacts)
struct LsPt {
    float alpha, cost, d0, d1;
};
G2_HD LsPt ls_point(float alpha, float q0, float q1, float q2) {
    LsPt r;
    r.alpha = alpha;
    r.cost = add_nf(add_nf(mul_nf(mul_nf(alpha, alpha), q2), mul_nf(alpha, q1)), q0);
    r.d0 = add_nf(mul_nf(mul_nf(2.0f, alpha), q2), q1);
    r.d1 = mul_nf(2.0f, q2);
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
    if (flag > 0) next_alpha = add_nf(pt.alpha, -(pt.d0 / pt.d1));
    return flag;
}

// Exact line search along `search` (constraint/solver.py:2246-2417).  Fills mv = M search and p.jv = J search;
// returns the step length.  Same decisions as the reference, arranged as two loops with one evaluation site each.
template <bool AR>
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
            s = row_dot<AR>(p, c, search, 0.0f);
            p.jv[c] = s;
            ja = p.Jaref[c];
            D = p.Dd[c];
        }
        rja[c] = ja; rjv[c] = s; rD[c] = D;
    }
    for (int c = LS_REG; c < n; ++c) p.jv[c] = row_dot<AR>(p, c, search, 0.0f);
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
            if (add_nf(ja, mul_nf(alpha, jv)) < 0.0f) {
                q0 = add_nf(q0, mul_nf(D, mul_nf(mul_nf(0.5f, ja), ja)));
                q1 = add_nf(q1, mul_nf(D, mul_nf(jv, ja)));
                q2 = add_nf(q2, mul_nf(D, mul_nf(mul_nf(0.5f, jv), jv)));
            }
        }
        for (int c = LS_REG; c < n; ++c) {
            float ja = p.Jaref[c], jv = p.jv[c], D = p.Dd[c];
            if (add_nf(ja, mul_nf(alpha, jv)) < 0.0f) {
                q0 = add_nf(q0, mul_nf(D, mul_nf(mul_nf(0.5f, ja), ja)));
                q1 = add_nf(q1, mul_nf(D, mul_nf(jv, ja)));
                q2 = add_nf(q2, mul_nf(D, mul_nf(mul_nf(0.5f, jv), jv)));
            }
        }
        LsPt e = ls_point(alpha, q0, q1, q2);
        ls_it += 1;
        bool more;
        if (stage == 0) {
            p0 = e;
            alpha = add_nf(p0.alpha, -(p0.d0 / p0.d1));
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
                    alpha = add_nf(p1.alpha, -(p1.d0 / p1.d1));
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
    float a[3] = {add_nf(p1.alpha, -(p1.d0 / p1.d1)), p1.alpha, mul_nf(add_nf(p1.alpha, p2.alpha), 0.5f)};
    while (!finished && ls_it < LS_ITERS) {
        float t[3][3];
        G2_UNROLL
        for (int k = 0; k < 3; ++k) { t[k][0] = g0; t[k][1] = g1; t[k][2] = g2; }
        G2_UNROLL
        for (int c = 0; c < LS_REG; ++c) {
            float ja = rja[c], jv = rjv[c], D = rD[c];
            float f0 = mul_nf(D, mul_nf(mul_nf(0.5f, ja), ja)), f1 = mul_nf(D, mul_nf(jv, ja)), f2 = mul_nf(D, mul_nf(mul_nf(0.5f, jv), jv));
            G2_UNROLL
            for (int k = 0; k < 3; ++k) {
                if (add_nf(ja, mul_nf(a[k], jv)) < 0.0f) { t[k][0] = add_nf(t[k][0], f0); t[k][1] = add_nf(t[k][1], f1); t[k][2] = add_nf(t[k][2], f2); }
            }
        }
        for (int c = LS_REG; c < n; ++c) {
            float ja = p.Jaref[c], jv = p.jv[c], D = p.Dd[c];
            float f0 = mul_nf(D, mul_nf(mul_nf(0.5f, ja), ja)), f1 = mul_nf(D, mul_nf(jv, ja)), f2 = mul_nf(D, mul_nf(mul_nf(0.5f, jv), jv));
            G2_UNROLL
            for (int k = 0; k < 3; ++k) {
                if (add_nf(ja, mul_nf(a[k], jv)) < 0.0f) { t[k][0] = add_nf(t[k][0], f0); t[k][1] = add_nf(t[k][1], f1); t[k][2] = add_nf(t[k][2], f2); }
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
        } else {
            float n1, n2;
            int b1 = bracket_update(p1, cand, n1);
            int b2 = bracket_update(p2, cand, n2);
            if (b1 == 0 && b2 == 0) {
                result = a[2];
                finished = true;
            } else {
                a[0] = n1; a[1] = n2; a[2] = mul_nf(add_nf(p1.alpha, p2.alpha), 0.5f);
            }
        }
    }
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
#if defined(__CUDA_ARCH__) && (defined(G2_WARP_ASYNC) || defined(G2_WARP_NEWTON))  // experiments: warps run free of each other / only inside the Newton loop (profiles/r02_experiments.md)
#define G2_BLOCK_ANY(x) (__any_sync(0xffffffffu, (x)) != 0)
#elif defined(__CUDA_ARCH__)
#define G2_BLOCK_ANY(x) (__syncthreads_or((x) ? 1 : 0) != 0)
#else
#define G2_BLOCK_ANY(x) (x)
#endif

#if defined(__CUDA_ARCH__) && defined(G2_FINE_SYNC_ON)
#define G2_FINE_SYNC() __syncthreads()
#else
#define G2_FINE_SYNC()
#endif

// Newton solve for the constrained acceleration; writes p.acc, warm start, per-link contact forces.
// Called by EVERY thread of the block (`valid` = the thread owns an env) because the iteration loop contains
// block barriers.  Trip 0 of the loop is the initial evaluation (active set, Hessian factorisation, gradient);
// trips >= 1 are the Newton iterations of constraint/solver.py:2862-2938, each helper having ONE call site.
template <bool AR>
G2_HDN void solve_constraints_t(const Go2Model& m, Phys& p, bool valid) {
    const int n = valid ? p.n_rows : 0;
    float qacc[NV], Ma[NV], qfc[NV], grad[NV], Mgrad[NV], search[NV], mv[NV];
    float* Hs = p.H;
    if (valid) {
        bool ws = n > 0 && p.warm;
        G2_UNROLL
        for (int d = 0; d < NV; ++d) qacc[d] = ws ? p.qacc_ws[d] : p.acc_smooth[d];
        if (n > 0) mul_M(p.M, qacc, Ma);
        for (int c = 0; c < n; ++c) {
            p.Jaref[c] = row_dot<AR>(p, c, qacc, -p.aref[c]);
            p.act[c] = 0;
        }
        p.cost = 0.0f;
    }
    int n_newton = 0;
    bool active = n > 0;
    const float tol = m.meaninertia * (float)NV * SOLVER_TOL;
    for (int it = 0; it <= SOLVER_ITERS; ++it) {
        if (!G2_BLOCK_ANY(active)) break;
        const bool first = it == 0;
        if (active && !first) {
            n_newton++;
            float alpha = linesearch<AR>(m, p, search, Ma, mv);
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
        G2_FINE_SYNC();
        if (active) {
            unsigned long long flist;
            int nfl;
            update_constraint<AR>(p, qacc, Ma, qfc, flist, nfl);
            // active-set changes: k-th flip of every lane together, oldest first (row order, like the reference)
            bool refactor = first || nfl > MAX_RANK1;
            for (int k = nfl - 1; k >= 0 && !refactor; --k) {
                int c = (int)((flist >> (8 * k)) & 0xffull);
                float sg = p.act[c] ? 1.0f : -1.0f;
                if (!(AR ? hessian_rank1_a(p, c, sg) : hessian_rank1(p, Hs, c, sg))) refactor = true;
            }
            if (refactor) {
                if (AR) hessian_factor_a(p);
                else hessian_factor(p, Hs);
            }
        }
        G2_FINE_SYNC();
        if (active) {
            if (AR) update_gradient_a(p, Ma, qfc, grad, Mgrad);
            else update_gradient(p, Hs, Ma, qfc, grad, Mgrad);
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

// Every thread of the block calls this (see solve_constraints_t); envs in dense mode are rare, so the second loop
// normally ends at its first block vote.
G2_HD void solve_constraints(const Go2Model& m, Phys& p, bool valid) {
    solve_constraints_t<true>(m, p, valid && p.arrow != 0);
    solve_constraints_t<false>(m, p, valid && p.arrow == 0);
}

}  // namespace go2
