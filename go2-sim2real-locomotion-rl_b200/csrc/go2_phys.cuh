// go2_b200 - per-environment rigid-body dynamics of the Go2 (kinematics, CRBA, LtDL, RNE, integration).
//
// One environment = one `Phys` working set owned by one thread (registers + thread-local memory); nothing
// here touches global memory.  Follows the reference's MuJoCo-style formulation (spatial quantities in
// the root-COM frame) so results agree with the oracle to fp32 rounding:
//   genesis/engine/solvers/rigid/abd/forward_kinematics.py:224-618,871-994
//   genesis/engine/solvers/rigid/abd/forward_dynamics.py:292-687,819-900,962-1699
// but is restructured for the fixed Go2 topology: robot-local link ids, packed lower-triangular
// matrices, no intermediate arrays that a later stage does not read.
#pragma once
#include "go2_math.cuh"

namespace go2 {

struct Contact {
    v3 pos, normal;
    float pen, mu;
    int la, lb;  // robot-local link ids (-1 = ground)
};

struct Phys {
    // generalized state
    float qpos[NQ], qvel[NV];
    // solver warm start + compact contact-normal cache (pairs in contact at the previous detection)
    float qacc_ws[NV];
    int warm;
    int n_cache;
    int cache_pair[MAX_CPAIRS];
    v3 cache_n[MAX_CPAIRS];
    // actuation (one of: direct torque / PD target evaluated each substep)
    float ctrl_force[NV];   // used when !pd_mode
    float ctrl_pos[NV];
    int pd_mode;
    float pd_kp, pd_kv;
    v3 ext_f, ext_t;        // -cfrc_applied of the pushed link (force, torque about root COM)
    int ext_link;
    v3 ext2_f, ext2_t;      // second slot: apply_links_external_force from the host (B_EXT_FORCE)
    int ext2_link;
    // kinematics
    v3 lpos[NLINK];
    q4 lquat[NLINK];
    v3 root_com;
    v3 cdof_a[NV], cdof_v[NV];
    v3 cd_a[NLINK], cd_v[NLINK];
    // dynamics
    float M[NTRI];
    float force[NV], acc_smooth[NV], acc[NV];
    // contacts
    int n_contacts, n_rows;
    Contact con[MAX_CONTACTS];
    float cost, prev_cost, gauss;
    // Phase-private scratch sharing one address range (thread-local memory is cached per touched word, so
    // overlapping the phases keeps the per-thread cache footprint at max(phase) instead of sum(phase)):
    //   kin -> smooth : inertias about the root COM, cdof time derivatives (written by kinematics()/velocities()
    //                   at the end of a substep, read by the next smooth_dynamics()), LtDL factor of M
    //   collide       : geom AABBs
    //   solve         : constraint rows, then Cholesky factor of the Newton Hessian + Jacobian (dense or arrow form)
    union {
        struct {
            float cin_I[NLINK][6];  // symmetric inertia about root COM: xx,xy,xz,yy,yz,zz
            v3 cin_h[NLINK];        // mass * (link COM - root COM)
            float cin_m[NLINK];
            v3 cdofd_a[NV], cdofd_v[NV];
            float L[NTRI], Dinv[NV];
        };
        struct {
            v3 g_c[NGEOM], g_h[NGEOM];  // world AABB centre / half extents
            v3 lk_lo[NLINK + 1], lk_hi[NLINK + 1];  // union of the geom boxes of the ground (0) and of each robot link (l + 1)
        };
        struct {
            uint8_t act[MAX_ROWS];
            float Dd[MAX_ROWS], aref[MAX_ROWS], Jaref[MAX_ROWS], jv[MAX_ROWS], efc[MAX_ROWS];
            union {
                struct {  // dense mode (some contact couples two different legs)
                    float H[NTRI];
                    float J[MAX_ROWS][NV];
                };
                struct {  // arrow mode: every row touches the base dofs + ONE leg; dofs ordered legs first, base last
                    float Hbb[21];      // Cholesky factor, base-base block (packed lower triangle)
                    float Hl[4][6];     // per leg: 3x3 diagonal block (packed lower triangle; hip, thigh, calf)
                    float Hbl[4][18];   // per leg: base rows x leg columns [r*3 + a]
                    uint8_t rleg[MAX_ROWS];
                    float Js[MAX_ROWS][9];  // row Jacobian: 6 base dofs, then hip/thigh/calf of leg rleg
                };
            };
        };
    };
    int arrow;  // 1 = arrow mode (see above)
    v3 cforce[NLINK];  // net contact force per link
    int err;
    int n_newton;
};

// ---------------------------------------------------------------------------------------------------
// Kinematics: link poses, root COM, composite-frame inertias, motion subspaces (cdof).
// ---------------------------------------------------------------------------------------------------
// `gl` = the global physics DR slot in force (B_GLOBALS): link mass shifts and the base COM shift.
G2_HDN void kinematics(const Go2Model& m, const float* __restrict__ gl, Phys& p) {
    const float* mass_shift = gl + G_MASS_SHIFT;
    const float* com_shift = gl + G_COM_SHIFT;
    // base (free joint): pose straight from qpos, quaternion renormalised
    p.lpos[0] = mk3(p.qpos[0], p.qpos[1], p.qpos[2]);
    p.lquat[0] = qnormalize(mk4(p.qpos[3], p.qpos[4], p.qpos[5], p.qpos[6]));
    v3 anchor[NLINK], axis_w[NLINK];
    anchor[0] = p.lpos[0];
    G2_UNROLL
    for (int l = 1; l < NLINK; ++l) {
        const int par = par_link(l);
        v3 pos = p.lpos[par] + qrot(ld3(m.l_pos[l]), p.lquat[par]);
        q4 quat = qcompose(ld4(m.l_quat[l]), p.lquat[par]);
        v3 ax = ld3(m.axis[l]);
        anchor[l] = pos;
        axis_w[l] = qrot(ax, quat);
        q4 qloc = rotvec_quat(ax * p.qpos[l + 6]);  // hinge l uses qpos index 6 + l (qpos0 = 0)
        p.lquat[l] = qcompose(qloc, quat);
        p.lpos[l] = pos;
    }
    // root COM and inertias about it
    v3 comw[NLINK];
    v3 acc = mk3(0, 0, 0);
    float msum = 0.0f;
    G2_UNROLL
    for (int l = 0; l < NLINK; ++l) {
        float mass = m.mass[l] + mass_shift[l];
        v3 ip = ld3(m.inertial_pos[l]);
        if (l == 0) ip = ip + mk3(com_shift[0], com_shift[1], com_shift[2]);
        comw[l] = p.lpos[l] + qrot(ip, p.lquat[l]);
        msum += mass;
        acc = acc + mass * comw[l];
        p.cin_m[l] = mass;
    }
    p.root_com = acc * (1.0f / msum);
    G2_UNROLL
    for (int l = 0; l < NLINK; ++l) {
        v3 t = comw[l] - p.root_com;
        float mass = p.cin_m[l];
        m3 R = qmat(qcompose(ld4(m.inertial_quat[l]), p.lquat[l]));
        const float* I = m.inertial_i[l];
        // R I R^T
        float RI[3][3];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) RI[i][j] = R.a[i][0] * I[0 * 3 + j] + R.a[i][1] * I[1 * 3 + j] + R.a[i][2] * I[2 * 3 + j];
        float W[3][3];
        for (int i = 0; i < 3; ++i)
            for (int j = i; j < 3; ++j) W[i][j] = RI[i][0] * R.a[j][0] + RI[i][1] * R.a[j][1] + RI[i][2] * R.a[j][2];
        p.cin_I[l][0] = W[0][0] + mass * (t.y * t.y + t.z * t.z);
        p.cin_I[l][1] = W[0][1] - mass * t.x * t.y;
        p.cin_I[l][2] = W[0][2] - mass * t.x * t.z;
        p.cin_I[l][3] = W[1][1] + mass * (t.x * t.x + t.z * t.z);
        p.cin_I[l][4] = W[1][2] - mass * t.y * t.z;
        p.cin_I[l][5] = W[2][2] + mass * (t.x * t.x + t.y * t.y);
        p.cin_h[l] = t * mass;
    }
    // motion subspaces
    {
        v3 off = p.root_com - anchor[0];
        m3 R = qmat(p.lquat[0]);
        for (int i = 0; i < 3; ++i) {
            p.cdof_a[i] = mk3(0, 0, 0);
            p.cdof_v[i] = mk3(i == 0 ? 1.0f : 0.0f, i == 1 ? 1.0f : 0.0f, i == 2 ? 1.0f : 0.0f);
            v3 c = mcol(R, i);
            p.cdof_a[3 + i] = c;
            p.cdof_v[3 + i] = cross(c, off);
        }
    }
    G2_UNROLL
    for (int l = 1; l < NLINK; ++l) {
        v3 off = p.root_com - anchor[l];
        p.cdof_a[5 + l] = axis_w[l];
        p.cdof_v[5 + l] = cross(axis_w[l], off);
    }
}

// Link spatial velocities (COM frame) and cdof time derivatives.
G2_HDN void velocities(const Go2Model& m, Phys& p) {
    v3 cv = mk3(0, 0, 0), ca = mk3(0, 0, 0);
    for (int i = 0; i < 3; ++i) {
        cv = cv + p.cdof_v[i] * p.qvel[i];
        ca = ca + p.cdof_a[i] * p.qvel[i];
        p.cdofd_a[i] = mk3(0, 0, 0);
        p.cdofd_v[i] = mk3(0, 0, 0);
    }
    for (int i = 3; i < 6; ++i) {
        p.cdofd_a[i] = cross(ca, p.cdof_a[i]);
        p.cdofd_v[i] = cross(ca, p.cdof_v[i]) + cross(cv, p.cdof_a[i]);
    }
    for (int i = 3; i < 6; ++i) {
        cv = cv + p.cdof_v[i] * p.qvel[i];
        ca = ca + p.cdof_a[i] * p.qvel[i];
    }
    p.cd_v[0] = cv;
    p.cd_a[0] = ca;
    G2_UNROLL
    for (int l = 1; l < NLINK; ++l) {
        const int par = par_link(l), d = 5 + l;
        v3 pv = p.cd_v[par], pa = p.cd_a[par];
        p.cdofd_a[d] = cross(pa, p.cdof_a[d]);
        p.cdofd_v[d] = cross(pa, p.cdof_v[d]) + cross(pv, p.cdof_a[d]);
        p.cd_v[l] = pv + p.cdof_v[d] * p.qvel[d];
        p.cd_a[l] = pa + p.cdof_a[d] * p.qvel[d];
    }
}

G2_HD v3 symmul(const float* I, v3 a) {
    return mk3(I[0] * a.x + I[1] * a.y + I[2] * a.z, I[1] * a.x + I[3] * a.y + I[4] * a.z, I[2] * a.x + I[4] * a.y + I[5] * a.z);
}
// spatial inertia times motion (geom.py inertial_mul): returns (ang, vel) force components
G2_HD void inertia_mul(const float* I, v3 h, float mass, v3 vel, v3 ang, v3& fa, v3& fv) {
    fa = symmul(I, ang) + cross(h, vel);
    fv = mass * vel - cross(h, ang);
}

// Mass matrix by composite rigid bodies; only the tree-structured nonzeros are computed
// (base block, each leg's 3x3 block and its coupling with the base).  Packed lower triangle.
G2_HDN void mass_matrix(const Go2Model& m, Phys& p, float diag_extra_hinge) {
    (void)m;
    float cI[NLINK][6];
    v3 ch[NLINK];
    float cm[NLINK];
    G2_UNROLL
    for (int l = 0; l < NLINK; ++l) {
        G2_UNROLL
        for (int k = 0; k < 6; ++k) cI[l][k] = p.cin_I[l][k];
        ch[l] = p.cin_h[l];
        cm[l] = p.cin_m[l];
    }
    G2_UNROLL
    for (int l = NLINK - 1; l >= 1; --l) {
        const int par = par_link(l);
        G2_UNROLL
        for (int k = 0; k < 6; ++k) cI[par][k] += cI[l][k];
        ch[par] = ch[par] + ch[l];
        cm[par] += cm[l];
    }
    // row dof i (link li): M[i][j] = f_i . cdof_j for the dofs j <= i on li's chain (= structural nonzeros)
    G2_UNROLL
    for (int i = 0; i < NV; ++i) {
        const int li = i < 6 ? 0 : i - 5;
        v3 fa, fv;
        inertia_mul(cI[li], ch[li], cm[li], p.cdof_v[i], p.cdof_a[i], fa, fv);
        G2_UNROLL
        for (int j = 0; j <= i; ++j) {
            if (m_nz(i, j)) p.M[tri(i, j)] = dot(fa, p.cdof_a[j]) + dot(fv, p.cdof_v[j]);
            else p.M[tri(i, j)] = 0.0f;
        }
    }
    G2_UNROLL
    for (int d = 6; d < NV; ++d) p.M[tri(d, d)] += m.armature[d] + diag_extra_hinge;
}

// In-place LtDL of M (MuJoCo order: last dof first) over the structural nonzeros (the factor of a tree has no fill).
G2_HDN void factor_mass(Phys& p) {
    G2_UNROLL
    for (int k = 0; k < NTRI; ++k) p.L[k] = p.M[k];
    G2_UNROLL
    for (int i = NV - 1; i >= 0; --i) {
        float dinv = 1.0f / p.L[tri(i, i)];
        p.Dinv[i] = dinv;
        G2_UNROLL
        for (int j = i - 1; j >= 0; --j) {
            if (m_nz(i, j)) {
                float a = p.L[tri(i, j)] * dinv;
                G2_UNROLL
                for (int k = 0; k <= j; ++k)
                    if (m_nz(i, k)) p.L[tri(j, k)] -= a * p.L[tri(i, k)];
                p.L[tri(i, j)] = a;
            }
        }
        p.L[tri(i, i)] = 1.0f;
    }
}
G2_HDN void solve_mass(const Phys& p, const float* vec, float* out) {
    G2_UNROLL
    for (int i = NV - 1; i >= 0; --i) {
        float c = vec[i];
        G2_UNROLL
        for (int j = i + 1; j < NV; ++j)
            if (m_nz(i, j)) c -= p.L[tri(j, i)] * out[j];
        out[i] = c;
    }
    G2_UNROLL
    for (int i = 0; i < NV; ++i) out[i] *= p.Dinv[i];
    G2_UNROLL
    for (int i = 0; i < NV; ++i) {
        float c = out[i];
        G2_UNROLL
        for (int j = 0; j < i; ++j)
            if (m_nz(i, j)) c -= p.L[tri(i, j)] * out[j];
        out[i] = c;
    }
}

G2_HD float applied_force(const Go2Model& m, const Phys& p, int d) {
    float f;
    if (p.pd_mode) f = p.pd_kp * (p.ctrl_pos[d] - p.qpos[d + 1]) + p.pd_kv * (0.0f - p.qvel[d]);
    else f = p.ctrl_force[d];
    return clampf(f, -m.frange[d], m.frange[d]);
}

// Smooth dynamics: M, LtDL, bias forces by recursive Newton-Euler, unconstrained acceleration.
G2_HDN void smooth_dynamics(const Go2Model& m, Phys& p) {
    mass_matrix(m, p, p.pd_mode ? p.pd_kv * m.dt : 0.0f);  // approximate_implicitfast: + kv*dt for PD dofs
    factor_mass(p);
    // link accelerations without qacc (gravity enters as base acceleration -g)
    v3 dda[NLINK], ddv[NLINK];
    {
        v3 a = mk3(0, 0, 0), v = mk3(-m.gravity[0], -m.gravity[1], -m.gravity[2]);
        for (int d = 0; d < 6; ++d) {
            v = v + p.cdofd_v[d] * p.qvel[d];
            a = a + p.cdofd_a[d] * p.qvel[d];
        }
        dda[0] = a;
        ddv[0] = v;
    }
    G2_UNROLL
    for (int l = 1; l < NLINK; ++l) {
        const int par = par_link(l), d = 5 + l;
        ddv[l] = ddv[par] + p.cdofd_v[d] * p.qvel[d];
        dda[l] = dda[par] + p.cdofd_a[d] * p.qvel[d];
    }
    // per-link inertial + gyroscopic force, external push, then child -> parent accumulation
    v3 fa[NLINK], fv[NLINK];
    G2_UNROLL
    for (int l = 0; l < NLINK; ++l) {
        v3 a1, v1, a2, v2;
        inertia_mul(p.cin_I[l], p.cin_h[l], p.cin_m[l], ddv[l], dda[l], a1, v1);
        inertia_mul(p.cin_I[l], p.cin_h[l], p.cin_m[l], p.cd_v[l], p.cd_a[l], a2, v2);
        fv[l] = v1 + cross(p.cd_a[l], v2);
        fa[l] = a1 + cross(p.cd_a[l], a2) + cross(p.cd_v[l], v2);
        if (l == p.ext_link) {
            fv[l] = fv[l] - p.ext_f;
            fa[l] = fa[l] - p.ext_t;
        }
        if (l == p.ext2_link) {
            fv[l] = fv[l] - p.ext2_f;
            fa[l] = fa[l] - p.ext2_t;
        }
    }
    G2_UNROLL
    for (int l = NLINK - 1; l >= 1; --l) {
        const int par = par_link(l);
        fv[par] = fv[par] + fv[l];
        fa[par] = fa[par] + fa[l];
    }
    G2_UNROLL
    for (int d = 0; d < NV; ++d) {
        const int l = d < 6 ? 0 : d - 5;
        float bias = dot(p.cdof_a[d], fa[l]) + dot(p.cdof_v[d], fv[l]);
        float ap = d < 6 ? 0.0f : applied_force(m, p, d);
        p.force[d] = ap - bias;
    }
    solve_mass(p, p.force, p.acc_smooth);
    for (int d = 0; d < NV; ++d) p.acc[d] = p.acc_smooth[d];
}

// Semi-implicit Euler with NaN guard (keeps the old state and raises an errno bit).
G2_HDN void integrate(const Go2Model& m, Phys& p) {
    float vn[NV], qn[NQ];
    bool ok = true;
    for (int d = 0; d < NV; ++d) {
        vn[d] = p.qvel[d] + p.acc[d] * m.dt;
        ok = ok && (vn[d] == vn[d]);
    }
    for (int j = 0; j < 3; ++j) qn[j] = p.qpos[j] + vn[j] * m.dt;
    q4 r0 = mk4(p.qpos[3], p.qpos[4], p.qpos[5], p.qpos[6]);
    q4 dq = rotvec_quat(mk3(vn[3], vn[4], vn[5]) * m.dt);
    q4 r1 = qnormalize(qmul(r0, dq));  // body-frame angular velocity: q <- q * dq
    qn[3] = r1.w; qn[4] = r1.x; qn[5] = r1.y; qn[6] = r1.z;
    for (int d = 6; d < NV; ++d) qn[d + 1] = p.qpos[d + 1] + vn[d] * m.dt;
    for (int q = 0; q < NQ; ++q) ok = ok && (qn[q] == qn[q]);
    if (ok) {
        for (int d = 0; d < NV; ++d) p.qvel[d] = vn[d];
        for (int q = 0; q < NQ; ++q) p.qpos[q] = qn[q];
    } else {
        p.err |= ERR_NAN_ACC;
    }
}

}  // namespace go2
