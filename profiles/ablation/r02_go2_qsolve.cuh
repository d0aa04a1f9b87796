// go2_b200 - constraint solve with FOUR LANES PER ENV (one lane per leg), registers + warp shuffles only.
//
// Same convex problem and the same Newton iteration as go2_solve.cuh (constraint/solver.py:2739-2966: warm start, active set,
// H = M + J^T D J, exact line search, termination on gradient norm / cost improvement), laid out along the structure the arrow
// mode already exploits: every constraint row touches the 6 base dofs plus ONE leg, and M couples a leg only with the base.
// Lane l of a quad owns leg l (dofs 6+l, 10+l, 14+l), the rows of that leg and its blocks of the factor
//     L = [diag(La_l) 0; Lb_l Lc]   (La_l 3x3, Lb_l 6x3 per lane; Lc 6x6 replicated in the 4 lanes);
// base quantities are replicated and every sum over legs is a 2-step xor butterfly (bit-identical in the 4 lanes).
// A thread-per-env solve keeps ~600 words of working set per env in thread-local memory (the 35x DRAM traffic of the fused
// kernel); here a lane holds ~150 words, i.e. the solve runs out of registers, and the dependent chain per warp is the chain of
// ONE leg + the base instead of four legs + the base.
// Differences to go2_solve.cuh (same minimiser, other rounding): H is refactored whenever the active set changed (no rank-1
// updates: with the legs in parallel a refactorisation is ~600 instructions), sums over legs are butterfly sums.
//
// The code is written once for device and host: QL = lanes emulated by one caller (1 on the device - each thread is a lane -,
// 4 on the host), every per-lane quantity is an array [QL], QFOR(q) runs over the caller's lanes.
#pragma once
#include "go2_solve.cuh"

namespace go2 {

constexpr int QMAXR = 24;      // rows per lane (6 contacts of one leg); more -> the env takes the thread-per-env solver
constexpr int QMAXC = 6;       // contacts per lane

#ifdef __CUDA_ARCH__
constexpr int QL = 1;
#define QFOR(q) for (int q = 0; q < 1; ++q)
G2_HD int q_leg(int) { return (int)(threadIdx.x & 3u); }
// butterfly over the 4 lanes of the quad; the mask names only the quad, so quads of a warp may sit in different iterations
__device__ __forceinline__ float q_sum(const float* x) {
    const unsigned mask = 0xFu << (threadIdx.x & 28u);
    float v = x[0];
    v += __shfl_xor_sync(mask, v, 1);
    v += __shfl_xor_sync(mask, v, 2);
    return v;
}
__device__ __forceinline__ int q_sumi(const int* x) {
    const unsigned mask = 0xFu << (threadIdx.x & 28u);
    int v = x[0];
    v += __shfl_xor_sync(mask, v, 1);
    v += __shfl_xor_sync(mask, v, 2);
    return v;
}
#else
constexpr int QL = 4;
#define QFOR(q) for (int q = 0; q < 4; ++q)
inline int q_leg(int q) { return q; }
inline float q_sum(const float* x) { return (x[0] + x[1]) + (x[2] + x[3]); }  // what every lane of the butterfly ends up with
inline int q_sumi(const int* x) { return x[0] + x[1] + x[2] + x[3]; }
#endif

// all-lanes sum of a per-lane value, result replicated: out[q] = sum for every lane the caller emulates
#define Q_ALLSUM(dst, src)                 \
    do {                                   \
        float s_ = q_sum(src);             \
        QFOR(q_) (dst)[q_] = s_;           \
    } while (0)

// One lane's share of the problem (filled by q_pack from the thread-per-env working set, or read from the HBM handoff).
struct QLane {
    // M: leg block (packed lower 3x3: hip, thigh, calf), base x leg coupling [r*3 + a], base block (replicated, packed 6x6)
    float Mll[6], Mbl[18], Mbb[21];
    float force_l[3], force_b[6], accs_l[3], accs_b[6], ws_l[3], ws_b[6];
    int n_rows, n_con;
    float J[QMAXR][9], D[QMAXR], aref[QMAXR];
    // contacts owned by the lane (rows 4k .. 4k+3): for the per-link contact forces
    v3 c_n[QMAXC];
    float c_mu[QMAXC];
    int c_la[QMAXC], c_lb[QMAXC];
};
struct QOut {
    float qacc_l[3], qacc_b[6];
    v3 cf_leg[3], cf_base;  // net contact force on the lane's hip / thigh / calf; on the base (replicated sum)
    int n_newton, bad;
};


// Split one env's arrow-mode problem (go2_solve.cuh: assemble_rows_a) over the four lanes.  Returns false if a lane would get
// more rows / contacts than it holds (the env then takes the thread-per-env solver).  Contacts that touch no leg (trunk vs
// ground) have zero leg columns and are dealt round-robin.
G2_HD bool q_pack(const Phys& p, QLane* L) {
    for (int l = 0; l < 4; ++l) {
        QLane& q = L[l];
        for (int a = 0; a < 3; ++a) {
            for (int b = 0; b <= a; ++b) q.Mll[tri(a, b)] = p.M[tri(6 + 4 * a + l, 6 + 4 * b + l)];
            for (int r = 0; r < 6; ++r) q.Mbl[r * 3 + a] = p.M[tri(6 + 4 * a + l, r)];
            q.force_l[a] = p.force[6 + 4 * a + l];
            q.accs_l[a] = p.acc_smooth[6 + 4 * a + l];
            q.ws_l[a] = p.qacc_ws[6 + 4 * a + l];
        }
        for (int k = 0; k < 21; ++k) q.Mbb[k] = p.M[k];
        for (int r = 0; r < 6; ++r) { q.force_b[r] = p.force[r]; q.accs_b[r] = p.acc_smooth[r]; q.ws_b[r] = p.qacc_ws[r]; }
        q.n_rows = 0;
        q.n_con = 0;
    }
    const int ncr = 4 * p.n_contacts;
    for (int ic = 0; ic < p.n_contacts; ++ic) {
        const Contact& c = p.con[ic];
        const bool leg_contact = c.la > 0 || c.lb > 0;
        QLane& q = L[leg_contact ? p.rleg[4 * ic] : (ic & 3)];
        if (q.n_con >= QMAXC || q.n_rows + 4 > QMAXR) return false;
        q.c_n[q.n_con] = c.normal; q.c_mu[q.n_con] = c.mu; q.c_la[q.n_con] = c.la; q.c_lb[q.n_con] = c.lb;
        q.n_con++;
        for (int k = 0; k < 4; ++k) {
            const int r = 4 * ic + k, o = q.n_rows++;
            for (int i = 0; i < 9; ++i) q.J[o][i] = p.Js[r][i];
            q.D[o] = p.Dd[r];
            q.aref[o] = p.aref[r];
        }
    }
    for (int r = ncr; r < p.n_rows; ++r) {
        QLane& q = L[p.rleg[r]];
        if (q.n_rows >= QMAXR) return false;
        const int o = q.n_rows++;
        for (int i = 0; i < 9; ++i) q.J[o][i] = p.Js[r][i];
        q.D[o] = p.Dd[r];
        q.aref[o] = p.aref[r];
    }
    return true;
}

G2_HD int q_popc(unsigned x) {
#ifdef __CUDA_ARCH__
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}

// ---- per-lane linear algebra on the (leg 3, base 6) split ------------------------------------------------------------------
// y = M x : leg part from the lane's blocks, base part = Mbb x_b (replicated) + butterfly sum of Mbl x_l
G2_HD void q_mul_M(const QLane* const* L, const float (*xl)[3], const float (*xb)[6], float (*yl)[3], float (*yb)[6]) {
    float part[6][QL];
    QFOR(q) {
        const QLane& a = *L[q];
        G2_UNROLL
        for (int i = 0; i < 3; ++i) {
            float s = 0.0f;
            G2_UNROLL
            for (int j = 0; j < 3; ++j) s += a.Mll[trs(i, j)] * xl[q][j];
            G2_UNROLL
            for (int r = 0; r < 6; ++r) s += a.Mbl[r * 3 + i] * xb[q][r];
            yl[q][i] = s;
        }
        G2_UNROLL
        for (int r = 0; r < 6; ++r) part[r][q] = a.Mbl[r * 3 + 0] * xl[q][0] + a.Mbl[r * 3 + 1] * xl[q][1] + a.Mbl[r * 3 + 2] * xl[q][2];
    }
    G2_UNROLL
    for (int r = 0; r < 6; ++r) {
        float s = q_sum(part[r]);
        QFOR(q) {
            float b = 0.0f;
            G2_UNROLL
            for (int j = 0; j < 6; ++j) b += L[q]->Mbb[trs(r, j)] * xb[q][j];
            yb[q][r] = b + s;
        }
    }
}
G2_HD float q_row_dot(const float* Jr, const float* xl, const float* xb, float s0) {
    float s = s0;
    G2_UNROLL
    for (int d = 0; d < 6; ++d) s += Jr[d] * xb[d];
    G2_UNROLL
    for (int a = 0; a < 3; ++a) s += Jr[6 + a] * xl[a];
    return s;
}

// per-lane solver state
struct QWork {
    float qacc_l[3], qacc_b[6], Ma_l[3], Ma_b[6], qfc_l[3], qfc_b[6], grad_l[3], grad_b[6], Mg_l[3], Mg_b[6], se_l[3], se_b[6], mv_l[3], mv_b[6];
    float A[6], Bm[18], C[21];           // factor blocks: La (3x3 packed), Lb (6x3), Lc (6x6 packed, replicated)
    float Jaref[QMAXR], jv[QMAXR], efc[QMAXR];
    unsigned act;                        // active-set bits of the lane's rows
};

// active set, constraint forces, qfc = J^T f, cost; returns (replicated) whether any row of the env flipped
G2_HD bool q_update_constraint(const QLane* const* L, QWork* w, float& cost, float& prev_cost, float& gauss) {
    float pb[6][QL], pc[QL], pg[QL];
    int flips[QL];
    QFOR(q) {
        const QLane& a = *L[q];
        QWork& s = w[q];
        G2_UNROLL
        for (int i = 0; i < 3; ++i) s.qfc_l[i] = 0.0f;
        float qb[6] = {0, 0, 0, 0, 0, 0};
        float cc = 0.0f;
        unsigned act = 0u;
        for (int c = 0; c < a.n_rows; ++c) {
            const float ja = s.Jaref[c], D = a.D[c];
            const bool on = ja < 0.0f;
            const float f = on ? -ja * D : 0.0f;
            s.efc[c] = f;
            if (on) {
                act |= 1u << c;
                G2_UNROLL
                for (int d = 0; d < 6; ++d) qb[d] += a.J[c][d] * f;
                G2_UNROLL
                for (int i = 0; i < 3; ++i) s.qfc_l[i] += a.J[c][6 + i] * f;
                cc += 0.5f * ja * ja * D;
            }
        }
        flips[q] = q_popc(act ^ s.act);
        s.act = act;
        G2_UNROLL
        for (int d = 0; d < 6; ++d) pb[d][q] = qb[d];
        pc[q] = cc;
        float g = 0.0f;
        G2_UNROLL
        for (int i = 0; i < 3; ++i) g += 0.5f * (s.Ma_l[i] - a.force_l[i]) * (s.qacc_l[i] - a.accs_l[i]);
        pg[q] = g;
    }
    const int nfl = q_sumi(flips);
    const float ccost = q_sum(pc), gl = q_sum(pg);
    float gb = 0.0f;
    G2_UNROLL
    for (int d = 0; d < 6; ++d) {
        float sd = q_sum(pb[d]);
        QFOR(q) w[q].qfc_b[d] = sd;
    }
    {
        const QLane& a = *L[0];
        const QWork& s = w[0];
        G2_UNROLL
        for (int d = 0; d < 6; ++d) gb += 0.5f * (s.Ma_b[d] - a.force_b[d]) * (s.qacc_b[d] - a.accs_b[d]);
    }
    prev_cost = cost;
    gauss = gb + gl;
    cost = gauss + ccost;
    return nfl > 0;
}

// H = M + J^T D J over the active rows, factored with the legs eliminated first (go2_solve.cuh: hessian_factor_a):
// per lane A = La La^T, X = B La^-T; base block C = Mbb + sum_lanes (sum_rows D jb jb^T - X X^T) = Lc Lc^T (replicated).
G2_HD void q_factor(const QLane* const* L, QWork* w) {
    float pc[21][QL];
    QFOR(q) {
        const QLane& a = *L[q];
        QWork& s = w[q];
        float A[6], Bm[18], Cp[21];
        G2_UNROLL
        for (int k = 0; k < 6; ++k) A[k] = a.Mll[k];
        G2_UNROLL
        for (int k = 0; k < 18; ++k) Bm[k] = a.Mbl[k];
        G2_UNROLL
        for (int k = 0; k < 21; ++k) Cp[k] = 0.0f;
        for (int c = 0; c < a.n_rows; ++c) {
            if (!((s.act >> c) & 1u)) continue;
            float jr[9];
            G2_UNROLL
            for (int i = 0; i < 9; ++i) jr[i] = a.J[c][i];
            arrow_acc_row(jr, a.D[c], Cp, A, Bm);
        }
        const float l00 = sqrtf(fmaxf(A[0], EPSF)), i00 = 1.0f / l00;
        const float l10 = A[1] * i00, l20 = A[3] * i00;
        const float l11 = sqrtf(fmaxf(A[2] - l10 * l10, EPSF)), i11 = 1.0f / l11;
        const float l21 = (A[4] - l20 * l10) * i11;
        const float l22 = sqrtf(fmaxf(A[5] - l20 * l20 - l21 * l21, EPSF)), i22 = 1.0f / l22;
        s.A[0] = l00; s.A[1] = l10; s.A[2] = l11; s.A[3] = l20; s.A[4] = l21; s.A[5] = l22;
        float X[6][3];
        G2_UNROLL
        for (int r = 0; r < 6; ++r) {
            const float x0 = Bm[r * 3 + 0] * i00;
            const float x1 = (Bm[r * 3 + 1] - x0 * l10) * i11;
            const float x2 = (Bm[r * 3 + 2] - x0 * l20 - x1 * l21) * i22;
            s.Bm[r * 3 + 0] = x0; s.Bm[r * 3 + 1] = x1; s.Bm[r * 3 + 2] = x2;
            X[r][0] = x0; X[r][1] = x1; X[r][2] = x2;
        }
        G2_UNROLL
        for (int i = 0; i < 6; ++i) {
            G2_UNROLL
            for (int j = 0; j <= i; ++j) pc[tri(i, j)][q] = Cp[tri(i, j)] - (X[i][0] * X[j][0] + X[i][1] * X[j][1] + X[i][2] * X[j][2]);
        }
    }
    float C[21];
    G2_UNROLL
    for (int k = 0; k < 21; ++k) C[k] = L[0]->Mbb[k] + q_sum(pc[k]);
    G2_UNROLL
    for (int i = 0; i < 6; ++i) {
        float t = C[tri(i, i)];
        G2_UNROLL
        for (int j = 0; j < i; ++j) t -= C[tri(i, j)] * C[tri(i, j)];
        const float lii = sqrtf(fmaxf(t, EPSF));
        C[tri(i, i)] = lii;
        const float inv = 1.0f / lii;
        G2_UNROLL
        for (int j = i + 1; j < 6; ++j) {
            float dt = 0.0f;
            G2_UNROLL
            for (int k = 0; k < i; ++k) dt += C[tri(j, k)] * C[tri(i, k)];
            C[tri(j, i)] = (C[tri(j, i)] - dt) * inv;
        }
    }
    QFOR(q) {
        G2_UNROLL
        for (int k = 0; k < 21; ++k) w[q].C[k] = C[k];
    }
}

// grad = Ma - force - qfc;  Mgrad = H^-1 grad through the factor (go2_solve.cuh: update_gradient_a)
G2_HD void q_gradient(const QLane* const* L, QWork* w) {
    float py[6][QL];
    QFOR(q) {
        const QLane& a = *L[q];
        QWork& s = w[q];
        G2_UNROLL
        for (int i = 0; i < 3; ++i) s.grad_l[i] = s.Ma_l[i] - a.force_l[i] - s.qfc_l[i];
        G2_UNROLL
        for (int d = 0; d < 6; ++d) s.grad_b[d] = s.Ma_b[d] - a.force_b[d] - s.qfc_b[d];
        const float y0 = s.grad_l[0] / s.A[0];
        const float y1 = (s.grad_l[1] - s.A[1] * y0) / s.A[2];
        const float y2 = (s.grad_l[2] - s.A[3] * y0 - s.A[4] * y1) / s.A[5];
        s.Mg_l[0] = y0; s.Mg_l[1] = y1; s.Mg_l[2] = y2;
        G2_UNROLL
        for (int r = 0; r < 6; ++r) py[r][q] = s.Bm[r * 3 + 0] * y0 + s.Bm[r * 3 + 1] * y1 + s.Bm[r * 3 + 2] * y2;
    }
    float yb[6];
    G2_UNROLL
    for (int r = 0; r < 6; ++r) yb[r] = w[0].grad_b[r] - q_sum(py[r]);
    const float* C = w[0].C;
    G2_UNROLL
    for (int i = 0; i < 6; ++i) {
        float c = yb[i];
        G2_UNROLL
        for (int j = 0; j < i; ++j) c -= C[tri(i, j)] * yb[j];
        yb[i] = c / C[tri(i, i)];
    }
    G2_UNROLL
    for (int i = 5; i >= 0; --i) {
        float c = yb[i];
        G2_UNROLL
        for (int j = i + 1; j < 6; ++j) c -= C[tri(j, i)] * yb[j];
        yb[i] = c / C[tri(i, i)];
    }
    QFOR(q) {
        QWork& s = w[q];
        G2_UNROLL
        for (int r = 0; r < 6; ++r) s.Mg_b[r] = yb[r];
        float z0 = s.Mg_l[0], z1 = s.Mg_l[1], z2 = s.Mg_l[2];
        G2_UNROLL
        for (int r = 0; r < 6; ++r) {
            z0 -= s.Bm[r * 3 + 0] * yb[r];
            z1 -= s.Bm[r * 3 + 1] * yb[r];
            z2 -= s.Bm[r * 3 + 2] * yb[r];
        }
        const float x2 = z2 / s.A[5];
        const float x1 = (z1 - s.A[4] * x2) / s.A[2];
        const float x0 = (z0 - s.A[1] * x1 - s.A[3] * x2) / s.A[0];
        s.Mg_l[0] = x0; s.Mg_l[1] = x1; s.Mg_l[2] = x2;
    }
}

// one evaluation of the line-search polynomial at up to 3 step lengths: t[k] = (cost, slope, curvature) pieces summed over
// the lane's rows that are active at alpha_k (unfused arithmetic, see go2_solve.cuh), butterfly-summed and added to (g0, g1, g2)
template <int NA>
G2_HD void q_ls_eval(const QLane* const* L, const QWork* w, const float* alpha, float g0, float g1, float g2, LsPt* out) {
    float part[NA][3][QL];
    QFOR(q) {
        const QLane& a = *L[q];
        const QWork& s = w[q];
        float t[NA][3];
        G2_UNROLL
        for (int k = 0; k < NA; ++k) t[k][0] = t[k][1] = t[k][2] = 0.0f;
        for (int c = 0; c < a.n_rows; ++c) {
            const float ja = s.Jaref[c], jv = s.jv[c], D = a.D[c];
            const float f0 = mul_nf(D, mul_nf(mul_nf(0.5f, ja), ja)), f1 = mul_nf(D, mul_nf(jv, ja)), f2 = mul_nf(D, mul_nf(mul_nf(0.5f, jv), jv));
            G2_UNROLL
            for (int k = 0; k < NA; ++k) {
                if (add_nf(ja, mul_nf(alpha[k], jv)) < 0.0f) { t[k][0] = add_nf(t[k][0], f0); t[k][1] = add_nf(t[k][1], f1); t[k][2] = add_nf(t[k][2], f2); }
            }
        }
        G2_UNROLL
        for (int k = 0; k < NA; ++k) { part[k][0][q] = t[k][0]; part[k][1][q] = t[k][1]; part[k][2][q] = t[k][2]; }
    }
    G2_UNROLL
    for (int k = 0; k < NA; ++k)
        out[k] = ls_point(alpha[k], add_nf(g0, q_sum(part[k][0])), add_nf(g1, q_sum(part[k][1])), add_nf(g2, q_sum(part[k][2])));
}

// Exact line search along `search` (same decisions as go2_solve.cuh: linesearch / constraint/solver.py:2246-2417); fills
// mv = M search and jv = J search; returns the (replicated) step length.
G2_HD float q_linesearch(const Go2Model& m, const QLane* const* L, QWork* w, float gauss) {
    float pl[QL], p1[QL], p2[QL];
    float xl[QL][3], xb[QL][6], yl[QL][3], yb[QL][6];
    QFOR(q) {
        float s = 0.0f;
        G2_UNROLL
        for (int i = 0; i < 3; ++i) { s += w[q].se_l[i] * w[q].se_l[i]; xl[q][i] = w[q].se_l[i]; }
        G2_UNROLL
        for (int d = 0; d < 6; ++d) xb[q][d] = w[q].se_b[d];
        pl[q] = s;
    }
    float sn = q_sum(pl);
    G2_UNROLL
    for (int d = 0; d < 6; ++d) sn += w[0].se_b[d] * w[0].se_b[d];
    sn = sqrtf(sn);
    const float gtol = SOLVER_TOL * LS_TOL * sn * (m.meaninertia * (float)NV);
    if (sn < EPSF) return 0.0f;
    q_mul_M(L, xl, xb, yl, yb);
    QFOR(q) {
        const QLane& a = *L[q];
        QWork& s = w[q];
        G2_UNROLL
        for (int i = 0; i < 3; ++i) s.mv_l[i] = yl[q][i];
        G2_UNROLL
        for (int d = 0; d < 6; ++d) s.mv_b[d] = yb[q][d];
        for (int c = 0; c < a.n_rows; ++c) s.jv[c] = q_row_dot(a.J[c], s.se_l, s.se_b, 0.0f);
        float a1 = 0.0f, a2 = 0.0f;
        G2_UNROLL
        for (int i = 0; i < 3; ++i) {
            a1 += s.se_l[i] * s.Ma_l[i] - s.se_l[i] * a.force_l[i];
            a2 += 0.5f * s.se_l[i] * s.mv_l[i];
        }
        p1[q] = a1; p2[q] = a2;
    }
    float g1 = q_sum(p1), g2 = q_sum(p2);
    {
        const QLane& a = *L[0];
        const QWork& s = w[0];
        G2_UNROLL
        for (int d = 0; d < 6; ++d) {
            g1 += s.se_b[d] * s.Ma_b[d] - s.se_b[d] * a.force_b[d];
            g2 += 0.5f * s.se_b[d] * s.mv_b[d];
        }
    }
    const float g0 = gauss;
    int ls_it = 0;
    LsPt p0, pa, pb;
    p0.alpha = p0.cost = p0.d0 = 0.0f; p0.d1 = 1.0f;
    pa = p0; pb = p0;
    float dir = 1.0f, alpha = 0.0f, result = 0.0f;
    bool pbupd = false, finished = false;
    int stage = 0;
    for (;;) {
        LsPt e;
        q_ls_eval<1>(L, w, &alpha, g0, g1, g2, &e);
        ls_it += 1;
        bool more;
        if (stage == 0) {
            p0 = e;
            alpha = add_nf(p0.alpha, -(p0.d0 / p0.d1));
            stage = 1;
            more = true;
        } else {
            if (stage == 1) {
                pa = e;
                if (p0.cost < pa.cost) pa = p0;
            } else {
                pa = e;
            }
            if (fabsf(pa.d0) < gtol) {
                result = pa.alpha;
                finished = true;
                more = false;
            } else {
                if (stage == 1) {
                    dir = pa.d0 < 0.0f ? 1.0f : -1.0f;
                    pb = pa;
                    stage = 2;
                }
                if (pa.d0 * dir <= -gtol && ls_it < LS_ITERS) {
                    pb = pa;
                    pbupd = true;
                    alpha = add_nf(pa.alpha, -(pa.d0 / pa.d1));
                    more = true;
                } else {
                    more = false;
                }
            }
        }
        if (!more) break;
    }
    if (!finished && (ls_it >= LS_ITERS || !pbupd)) {
        result = pa.alpha;
        finished = true;
    }
    float al[3] = {add_nf(pa.alpha, -(pa.d0 / pa.d1)), pa.alpha, mul_nf(add_nf(pa.alpha, pb.alpha), 0.5f)};
    while (!finished && ls_it < LS_ITERS) {
        LsPt cand[3];
        q_ls_eval<3>(L, w, al, g0, g1, g2, cand);
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
            int b1 = bracket_update(pa, cand, n1);
            int b2 = bracket_update(pb, cand, n2);
            if (b1 == 0 && b2 == 0) {
                result = al[2];
                finished = true;
            } else {
                al[0] = n1; al[1] = n2; al[2] = mul_nf(add_nf(pa.alpha, pb.alpha), 0.5f);
            }
        }
    }
    if (!finished) {
        if (pa.cost <= pb.cost && pa.cost < p0.cost) result = pa.alpha;
        else if (pb.cost <= pa.cost && pb.cost < p0.cost) result = pb.alpha;
        else result = 0.0f;
    }
    return result;
}

// Newton solve of one env by its quad.  `warm`: the env's warm-start flag (replicated).  L[q] / out[q]: the caller's lanes.
G2_HD void q_solve(const Go2Model& m, const QLane* const* L, int warm, QOut* out) {
    QWork w[QL];
    int nr[QL];
    QFOR(q) nr[q] = L[q]->n_rows;
    const int n = q_sumi(nr);
    const bool ws = n > 0 && warm;
    float xl[QL][3], xb[QL][6], yl[QL][3], yb[QL][6];
    QFOR(q) {
        const QLane& a = *L[q];
        QWork& s = w[q];
        G2_UNROLL
        for (int i = 0; i < 3; ++i) { s.qacc_l[i] = ws ? a.ws_l[i] : a.accs_l[i]; xl[q][i] = s.qacc_l[i]; }
        G2_UNROLL
        for (int d = 0; d < 6; ++d) { s.qacc_b[d] = ws ? a.ws_b[d] : a.accs_b[d]; xb[q][d] = s.qacc_b[d]; }
        s.act = 0u;
    }
    int n_newton = 0;
    if (n > 0) {
        q_mul_M(L, xl, xb, yl, yb);
        QFOR(q) {
            const QLane& a = *L[q];
            QWork& s = w[q];
            G2_UNROLL
            for (int i = 0; i < 3; ++i) s.Ma_l[i] = yl[q][i];
            G2_UNROLL
            for (int d = 0; d < 6; ++d) s.Ma_b[d] = yb[q][d];
            for (int c = 0; c < a.n_rows; ++c) s.Jaref[c] = q_row_dot(a.J[c], s.qacc_l, s.qacc_b, -a.aref[c]);
        }
        const float tol = m.meaninertia * (float)NV * SOLVER_TOL;
        float cost = 0.0f, prev_cost = 0.0f, gauss = 0.0f;
        bool active = true;
        for (int it = 0; it <= SOLVER_ITERS && active; ++it) {
            const bool first = it == 0;
            if (!first) {
                n_newton++;
                const float alpha = q_linesearch(m, L, w, gauss);
                if (fabsf(alpha) < EPSF) break;
                QFOR(q) {
                    const QLane& a = *L[q];
                    QWork& s = w[q];
                    G2_UNROLL
                    for (int i = 0; i < 3; ++i) { s.qacc_l[i] += s.se_l[i] * alpha; s.Ma_l[i] += s.mv_l[i] * alpha; }
                    G2_UNROLL
                    for (int d = 0; d < 6; ++d) { s.qacc_b[d] += s.se_b[d] * alpha; s.Ma_b[d] += s.mv_b[d] * alpha; }
                    for (int c = 0; c < a.n_rows; ++c) s.Jaref[c] += s.jv[c] * alpha;
                }
            }
            const bool flipped = q_update_constraint(L, w, cost, prev_cost, gauss);
            if (first || flipped) q_factor(L, w);
            q_gradient(L, w);
            if (!first) {
                float pg[QL];
                QFOR(q) pg[q] = w[q].grad_l[0] * w[q].grad_l[0] + w[q].grad_l[1] * w[q].grad_l[1] + w[q].grad_l[2] * w[q].grad_l[2];
                float gn = q_sum(pg);
                G2_UNROLL
                for (int d = 0; d < 6; ++d) gn += w[0].grad_b[d] * w[0].grad_b[d];
                gn = sqrtf(gn);
                const bool improved = gn > tol && (prev_cost - cost) > tol;
                if (!improved || it == SOLVER_ITERS) active = false;
            }
            QFOR(q) {
                G2_UNROLL
                for (int i = 0; i < 3; ++i) w[q].se_l[i] = -w[q].Mg_l[i];
                G2_UNROLL
                for (int d = 0; d < 6; ++d) w[q].se_b[d] = -w[q].Mg_b[d];
            }
        }
    }
    // results: acceleration, NaN flag, per-link contact forces (solver.py:2974-3013)
    float pfx[QL], pfy[QL], pfz[QL];
    int bad[QL];
    QFOR(q) {
        const QLane& a = *L[q];
        const QWork& s = w[q];
        QOut& o = out[q];
        bool b = false;
        G2_UNROLL
        for (int i = 0; i < 3; ++i) { o.qacc_l[i] = s.qacc_l[i]; b = b || !(s.qacc_l[i] == s.qacc_l[i]); }
        G2_UNROLL
        for (int d = 0; d < 6; ++d) { o.qacc_b[d] = s.qacc_b[d]; b = b || !(s.qacc_b[d] == s.qacc_b[d]); }
        bad[q] = b ? 1 : 0;
        o.n_newton = n_newton;
        v3 fb = mk3(0, 0, 0);
        G2_UNROLL
        for (int i = 0; i < 3; ++i) o.cf_leg[i] = mk3(0, 0, 0);
        const int leg = q_leg(q);
        for (int k = 0; k < a.n_con; ++k) {
            v3 d1, d2;
            orthobasis(a.c_n[k], d1, d2);
            v3 f = mk3(0, 0, 0);
            G2_UNROLL
            for (int j = 0; j < 4; ++j) {
                v3 d = (j < 2 ? d1 : d2) * (float)(2 * (j % 2) - 1);
                f = f + (d * a.c_mu[k] - a.c_n[k]) * s.efc[4 * k + j];
            }
            const int la = a.c_la[k], lb = a.c_lb[k];
            if (la == 0) fb = fb - f;
            else if (la > 0 && ((la - 1) & 3) == leg) o.cf_leg[(la - 1) >> 2] = o.cf_leg[(la - 1) >> 2] - f;
            if (lb == 0) fb = fb + f;
            else if (lb > 0 && ((lb - 1) & 3) == leg) o.cf_leg[(lb - 1) >> 2] = o.cf_leg[(lb - 1) >> 2] + f;
        }
        pfx[q] = fb.x; pfy[q] = fb.y; pfz[q] = fb.z;
    }
    const v3 cfb = mk3(q_sum(pfx), q_sum(pfy), q_sum(pfz));
    const int anybad = q_sumi(bad);
    QFOR(q) {
        out[q].cf_base = cfb;
        out[q].bad = anybad;
    }
}

#ifndef __CUDA_ARCH__
// Host build (tests/host_emul): the quad solver on one env of the thread-per-env working set - how the CPU suite exercises
// q_pack / q_solve inside the fused tick.  Returns false if the env does not fit the quad layout.
inline bool q_solve_host(const Go2Model& m, Phys& p) {
    if (!p.arrow) return false;
    QLane L[4];
    if (!q_pack(p, L)) return false;
    const QLane* Lp[4] = {&L[0], &L[1], &L[2], &L[3]};
    QOut out[4];
    q_solve(m, Lp, p.warm, out);
    for (int d = 0; d < 6; ++d) p.acc[d] = p.qacc_ws[d] = out[0].qacc_b[d];
    for (int l = 0; l < NLINK; ++l) p.cforce[l] = mk3(0, 0, 0);
    p.cforce[0] = out[0].cf_base;
    for (int l = 0; l < 4; ++l)
        for (int a = 0; a < 3; ++a) {
            p.acc[6 + 4 * a + l] = p.qacc_ws[6 + 4 * a + l] = out[l].qacc_l[a];
            p.cforce[1 + 4 * a + l] = out[l].cf_leg[a];
        }
    p.n_newton = out[0].n_newton;
    if (out[0].bad) p.err |= ERR_NAN_FORCE;
    p.warm = 1;
    return true;
}
#endif

}  // namespace go2
