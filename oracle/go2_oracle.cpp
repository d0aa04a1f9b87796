// TEST INFRASTRUCTURE ONLY - see go2_oracle.hpp.  CPU restatement of the reference rigid step.
#include "go2_oracle.hpp"
#include <cstdio>
#include <limits>
#include <thread>

namespace orc {

// =====================================================================================================
// abd/forward_kinematics.py:462-618  func_forward_kinematics_entity
// =====================================================================================================
static void forward_kinematics(const Model& m, Env& e) {
    e.pos[0] = V3(0, 0, 0);  // ground: fixed root link, pose never updated (:614-618)
    e.quat[0] = Q4(1, 0, 0, 0);
    for (int l = 1; l < NL; ++l) {
        V3 pos = m.l_pos[l];
        Q4 quat = m.l_quat[l];
        int p = m.parent[l];
        if (p != -1) {
            pos = e.pos[p] + transform_by_quat(m.l_pos[l], e.quat[p]);
            quat = transform_quat_by_quat(m.l_quat[l], e.quat[p]);
        }
        int qs = m.q_start[l], ds = m.dof_start[l];
        if (m.n_dofs[l] == 6) {  // FREE joint
            e.xanchor[l] = V3(e.qpos[qs], e.qpos[qs + 1], e.qpos[qs + 2]);
            e.xaxis[l] = V3(0, 0, 1);
            V3 pos_(e.qpos[qs], e.qpos[qs + 1], e.qpos[qs + 2]);
            real qw = e.qpos[qs + 3], qx = e.qpos[qs + 4], qy = e.qpos[qs + 5], qz = e.qpos[qs + 6];
            real n = std::sqrt(qw * qw + qx * qx + qy * qy + qz * qz);
            Q4 quat_(qw / n, qx / n, qy / n, qz / n);
            pos = pos_;
            quat = quat_;
            V3 xyz = quat_to_xyz(quat);
            for (int j = 0; j < 3; ++j) {
                e.dpos[ds + j] = pos[j];
                e.dpos[ds + 3 + j] = xyz[j];
            }
        } else {  // REVOLUTE (joint pos = 0 in link frame)
            V3 axis = m.axis[l];
            e.xanchor[l] = transform_by_quat(V3(0, 0, 0), quat) + pos;
            e.xaxis[l] = transform_by_quat(axis, quat);
            e.dpos[ds] = e.qpos[qs] - real(0);  // qpos0 = 0 for every Go2 hinge
            Q4 qloc = rotvec_to_quat(axis * e.dpos[ds]);
            quat = transform_quat_by_quat(qloc, quat);
            pos = e.xanchor[l] - transform_by_quat(V3(0, 0, 0), quat);
        }
        e.pos[l] = pos;
        e.quat[l] = quat;
    }
}

// utils/geom.py:321-336 qd_transform_inertia_by_trans_quat
static void transform_inertia(const M3& I, real mass, V3 t, Q4 q, M3& out_i, V3& out_pos) {
    real x = t.x, y = t.y, z = t.z;
    real xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    real hh[3][3] = {{yy + zz, -xy, -xz}, {-xy, xx + zz, -yz}, {-xz, -yz, xx + yy}};
    M3 R = quat_to_R(q);
    M3 RI;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            real s = 0;
            for (int k = 0; k < 3; ++k) s += R.m[i][k] * I.m[k][j];
            RI.m[i][j] = s;
        }
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            real s = 0;
            for (int k = 0; k < 3; ++k) s += RI.m[i][k] * R.m[j][k];
            out_i.m[i][j] = s + hh[i][j] * mass;
        }
    out_pos = t * mass;
}

// abd/forward_kinematics.py:224-459  func_COM_links_entity (robot entity; the ground has no dofs)
static void com_links(const Model& m, Env& e) {
    V3 com_bw(0, 0, 0);
    real mass_sum = 0;
    V3 i_pos_bw[NL];
    for (int l = 1; l < NL; ++l) {
        real mass = m.inertial_mass[l] + e.mass_shift[l];
        i_pos_bw[l] = e.pos[l] + transform_by_quat(m.inertial_pos[l] + e.i_pos_shift[l], e.quat[l]);
        e.i_quat[l] = transform_quat_by_quat(m.inertial_quat[l], e.quat[l]);
        mass_sum = mass_sum + mass;
        com_bw = com_bw + mass * i_pos_bw[l];
    }
    e.root_COM = com_bw / mass_sum;
    for (int l = 1; l < NL; ++l) {
        e.i_pos[l] = i_pos_bw[l] - e.root_COM;
        real i_mass = m.inertial_mass[l] + e.mass_shift[l];
        transform_inertia(m.inertial_i[l], i_mass, e.i_pos[l], e.i_quat[l], e.cinr_inertial[l], e.cinr_pos[l]);
        e.cinr_mass[l] = i_mass;
    }
    for (int l = 1; l < NL; ++l) {
        V3 offset_pos = e.root_COM - e.xanchor[l];
        int ds = m.dof_start[l];
        if (m.n_dofs[l] == 1) {
            e.cdof_ang[ds] = e.xaxis[l];
            e.cdof_vel[ds] = cross(e.xaxis[l], offset_pos);
        } else {
            for (int i = 0; i < 3; ++i) {
                e.cdof_ang[ds + i] = V3(0, 0, 0);
                e.cdof_vel[ds + i] = V3(0, 0, 0);
                e.cdof_vel[ds + i][i] = 1;
            }
            M3 R = quat_to_R(e.quat[l]);
            for (int i = 0; i < 3; ++i) {
                V3 col(R.m[0][i], R.m[1][i], R.m[2][i]);  // xmat_T[i, :]
                e.cdof_ang[ds + i + 3] = col;
                e.cdof_vel[ds + i + 3] = cross(col, offset_pos);
            }
        }
    }
}

// abd/forward_kinematics.py:708-744 func_update_geoms_entity + :1171-1193 kernel_update_geom_aabbs
static void update_geoms(const Model& m, Env& e) {
    for (int g = 0; g < m.ng; ++g) {
        int l = m.g_link[g];
        e.gpos[g] = e.pos[l] + transform_by_quat(m.g_pos[g], e.quat[l]);
        e.gquat[g] = transform_quat_by_quat(m.g_quat[g], e.quat[l]);
    }
}
static void update_geom_aabbs(const Model& m, Env& e) {
    const real inf = std::numeric_limits<real>::infinity();
    for (int g = 0; g < m.ng; ++g) {
        V3 lo(inf, inf, inf), hi(-inf, -inf, -inf);
        for (int c = 0; c < 8; ++c) {
            V3 corner((c & 4) ? m.g_aabb_max[g].x : m.g_aabb_min[g].x, (c & 2) ? m.g_aabb_max[g].y : m.g_aabb_min[g].y,
                      (c & 1) ? m.g_aabb_max[g].z : m.g_aabb_min[g].z);
            V3 w = transform_by_trans_quat(corner, e.gpos[g], e.gquat[g]);
            for (int k = 0; k < 3; ++k) {
                lo[k] = std::min(lo[k], w[k]);
                hi[k] = std::max(hi[k], w[k]);
            }
        }
        e.aabb_min[g] = lo;
        e.aabb_max[g] = hi;
    }
}

// abd/forward_kinematics.py:871-994 func_forward_velocity_entity
static void forward_velocity(const Model& m, Env& e) {
    for (int l = 1; l < NL; ++l) {
        V3 cvel_vel(0, 0, 0), cvel_ang(0, 0, 0);
        int p = m.parent[l];
        if (p != -1) {
            cvel_vel = e.cd_vel[p];
            cvel_ang = e.cd_ang[p];
        }
        int ds = m.dof_start[l];
        if (m.n_dofs[l] == 6) {
            for (int i = 0; i < 3; ++i) {
                cvel_vel = cvel_vel + e.cdof_vel[ds + i] * e.vel[ds + i];
                cvel_ang = cvel_ang + e.cdof_ang[ds + i] * e.vel[ds + i];
            }
            for (int i = 0; i < 3; ++i) {
                e.cdofd_ang[ds + i] = V3(0, 0, 0);
                e.cdofd_vel[ds + i] = V3(0, 0, 0);
                motion_cross_motion(cvel_ang, cvel_vel, e.cdof_ang[ds + i + 3], e.cdof_vel[ds + i + 3],
                                    e.cdofd_ang[ds + i + 3], e.cdofd_vel[ds + i + 3]);
            }
            for (int i = 0; i < 3; ++i) {
                cvel_vel = cvel_vel + e.cdof_vel[ds + i + 3] * e.vel[ds + i + 3];
                cvel_ang = cvel_ang + e.cdof_ang[ds + i + 3] * e.vel[ds + i + 3];
            }
        } else {
            motion_cross_motion(cvel_ang, cvel_vel, e.cdof_ang[ds], e.cdof_vel[ds], e.cdofd_ang[ds], e.cdofd_vel[ds]);
            cvel_vel = cvel_vel + e.cdof_vel[ds] * e.vel[ds];
            cvel_ang = cvel_ang + e.cdof_ang[ds] * e.vel[ds];
        }
        e.cd_vel[l] = cvel_vel;
        e.cd_ang[l] = cvel_ang;
    }
}

// rigid_solver.py:3155-3180  func_update_cartesian_space + func_forward_velocity
static void update_cartesian_space(const Model& m, Env& e) {
    forward_kinematics(m, e);
    com_links(m, e);
    update_geoms(m, e);
    forward_velocity(m, e);
}

// =====================================================================================================
// abd/forward_dynamics.py:292-541  func_compute_mass_matrix  (implicit_damping = approximate_implicitfast)
// =====================================================================================================
static void compute_mass_matrix(const Model& m, Env& e, bool implicit_damping) {
    M3 crb_i[NL];
    V3 crb_pos[NL];
    real crb_mass[NL];
    for (int l = 1; l < NL; ++l) {
        crb_i[l] = e.cinr_inertial[l];
        crb_pos[l] = e.cinr_pos[l];
        crb_mass[l] = e.cinr_mass[l];
    }
    for (int l = NL - 1; l >= 1; --l) {
        int p = m.parent[l];
        if (p != -1) {
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) crb_i[p].m[i][j] += crb_i[l].m[i][j];
            crb_mass[p] += crb_mass[l];
            crb_pos[p] = crb_pos[p] + crb_pos[l];
        }
    }
    V3 f_ang[NV], f_vel[NV];
    for (int l = 1; l < NL; ++l)
        for (int d = m.dof_start[l]; d < m.dof_start[l] + m.n_dofs[l]; ++d)
            inertial_mul(crb_pos[l], crb_i[l], crb_mass[l], e.cdof_vel[d], e.cdof_ang[d], f_ang[d], f_vel[d]);
    for (int i = 0; i < NV; ++i)
        for (int j = 0; j < NV; ++j)
            e.M[i][j] = (dot(f_ang[i], e.cdof_ang[j]) + dot(f_vel[i], e.cdof_vel[j])) * real(m.mass_parent_mask[i][j]);
    for (int i = 0; i < NV; ++i)
        for (int j = i + 1; j < NV; ++j) e.M[i][j] = e.M[j][i];
    for (int d = 0; d < NV; ++d) e.M[d][d] += m.armature[d];
    if (implicit_damping) {
        for (int d = 0; d < NV; ++d) {
            e.M[d][d] += real(0) * m.dt;  // damping = 0 (no <dynamics> in the URDF)
            if (e.ctrl_mode[d] == CTRL_POSITION || e.ctrl_mode[d] == CTRL_VELOCITY) e.M[d][d] += m.kv[d] * m.dt;
        }
    }
}

// abd/forward_dynamics.py:545-604  func_factor_mass (non-tiled branch), LtDL in reverse order
static void factor_mass(Env& e) {
    for (int i = 0; i < NV; ++i)
        for (int j = 0; j <= i; ++j) e.ML[i][j] = e.M[i][j];
    for (int i_ = 0; i_ < NV; ++i_) {
        int i = NV - i_ - 1;
        real D_inv = real(1.0) / e.ML[i][i];
        e.MDinv[i] = D_inv;
        for (int j_ = 0; j_ < i; ++j_) {
            int j = i - j_ - 1;
            real a = e.ML[i][j] * D_inv;
            for (int k = 0; k <= j; ++k) e.ML[j][k] -= a * e.ML[i][k];
            e.ML[i][j] = a;
        }
        e.ML[i][i] = 1;
    }
}
// abd/forward_dynamics.py:819-900 func_solve_mass_entity
static void solve_mass(const Env& e, const real* vec, real* out) {
    for (int i_ = 0; i_ < NV; ++i_) {
        int i = NV - i_ - 1;
        real c = vec[i];
        for (int j = i + 1; j < NV; ++j) c = c - e.ML[j][i] * out[j];
        out[i] = c;
    }
    for (int i = 0; i < NV; ++i) out[i] = out[i] * e.MDinv[i];
    for (int i = 0; i < NV; ++i) {
        real c = out[i];
        for (int j = 0; j < i; ++j) c = c - e.ML[i][j] * out[j];
        out[i] = c;
    }
}

// abd/forward_dynamics.py:962-1174 func_torque_and_passive_force (damping = stiffness = 0 for Go2)
static real control_force_of(const Model& m, const Env& e, int d) {
    real force = 0;
    if (e.ctrl_mode[d] == CTRL_FORCE)
        force = e.ctrl_force[d];
    else if (e.ctrl_mode[d] == CTRL_VELOCITY)
        force = m.kv[d] * (e.ctrl_vel[d] - e.vel[d]);
    else if (e.ctrl_mode[d] == CTRL_POSITION && !(d >= 3 && d < 6))
        force = m.kp[d] * (e.ctrl_pos[d] - e.dpos[d]) + m.kv[d] * (e.ctrl_vel[d] - e.vel[d]);
    return std::min(std::max(force, m.force_range[d][0]), m.force_range[d][1]);
}
static void torque_and_passive_force(const Model& m, Env& e) {
    for (int d = 0; d < NV; ++d) e.qf_applied[d] = control_force_of(m, e, d);
}

// abd/forward_dynamics.py:1178-1277 func_update_acc(update_cacc=False)
static void update_acc(const Model& m, Env& e) {
    for (int l = 1; l < NL; ++l) {
        int p = m.parent[l];
        if (p == -1) {
            e.cdd_vel[l] = -m.gravity;
            e.cdd_ang[l] = V3(0, 0, 0);
        } else {
            e.cdd_vel[l] = e.cdd_vel[p];
            e.cdd_ang[l] = e.cdd_ang[p];
        }
        for (int d = m.dof_start[l]; d < m.dof_start[l] + m.n_dofs[l]; ++d) {
            e.cdd_vel[l] = e.cdd_vel[l] + e.cdofd_vel[d] * e.vel[d];
            e.cdd_ang[l] = e.cdd_ang[l] + e.cdofd_ang[d] * e.vel[d];
        }
    }
}
// abd/forward_dynamics.py:1281-1392 func_update_force
static void update_force(const Model& m, Env& e) {
    for (int l = 1; l < NL; ++l) {
        V3 f1a, f1v, f2a, f2v, f3a, f3v;
        inertial_mul(e.cinr_pos[l], e.cinr_inertial[l], e.cinr_mass[l], e.cdd_vel[l], e.cdd_ang[l], f1a, f1v);
        inertial_mul(e.cinr_pos[l], e.cinr_inertial[l], e.cinr_mass[l], e.cd_vel[l], e.cd_ang[l], f2a, f2v);
        motion_cross_force(e.cd_ang[l], e.cd_vel[l], f2a, f2v, f3a, f3v);
        e.cfrc_vel[l] = f1v + f3v + e.cfrc_applied_vel[l];
        e.cfrc_ang[l] = f1a + f3a + e.cfrc_applied_ang[l];
    }
    for (int l = NL - 1; l >= 1; --l) {
        int p = m.parent[l];
        if (p != -1) {
            e.cfrc_vel[p] = e.cfrc_vel[p] + e.cfrc_vel[l];
            e.cfrc_ang[p] = e.cfrc_ang[p] + e.cfrc_ang[l];
        }
    }
}
// abd/forward_dynamics.py:1420-1478 func_bias_force
static void bias_force(const Model& m, Env& e) {
    for (int l = 1; l < NL; ++l)
        for (int d = m.dof_start[l]; d < m.dof_start[l] + m.n_dofs[l]; ++d) {
            e.qf_bias[d] = dot(e.cdof_ang[d], e.cfrc_ang[l]) + dot(e.cdof_vel[d], e.cfrc_vel[l]);
            e.force[d] = real(0) - e.qf_bias[d] + e.qf_applied[d];
            e.qf_smooth[d] = e.force[d];
        }
}
// abd/forward_dynamics.py:147-221 func_forward_dynamics
static void forward_dynamics(const Model& m, Env& e) {
    compute_mass_matrix(m, e, true);
    factor_mass(e);
    torque_and_passive_force(m, e);
    update_acc(m, e);
    update_force(m, e);
    bias_force(m, e);
    solve_mass(e, e.force, e.acc_smooth);  // :1499-1555 func_compute_qacc
    for (int d = 0; d < NV; ++d) e.acc[d] = e.acc_smooth[d];
}

// abd/forward_dynamics.py:1559-1699 func_integrate + abd/diff.py:25-55 func_copy_next_to_curr
static void integrate(const Model& m, Env& e) {
    real vel_next[NV], qpos_next[NQ];
    for (int d = 0; d < NV; ++d) vel_next[d] = e.vel[d] + e.acc[d] * m.dt;
    for (int l = 1; l < NL; ++l) {
        int ds = m.dof_start[l], qs = m.q_start[l];
        if (m.n_dofs[l] == 6) {
            for (int j = 0; j < 3; ++j) qpos_next[qs + j] = e.qpos[qs + j] + vel_next[ds + j] * m.dt;
            Q4 rot0(e.qpos[qs + 3], e.qpos[qs + 4], e.qpos[qs + 5], e.qpos[qs + 6]);
            V3 ang = V3(vel_next[ds + 3], vel_next[ds + 4], vel_next[ds + 5]) * m.dt;
            Q4 qrot = rotvec_to_quat(ang);
            Q4 rot = transform_quat_by_quat(qrot, rot0);
            qpos_next[qs + 3] = rot.w;
            qpos_next[qs + 4] = rot.x;
            qpos_next[qs + 5] = rot.y;
            qpos_next[qs + 6] = rot.z;
        } else {
            qpos_next[qs] = e.qpos[qs] + vel_next[ds] * m.dt;
        }
    }
    bool valid = true;
    for (int d = 0; d < NV; ++d) valid &= !std::isnan(vel_next[d]);
    for (int q = 0; q < NQ; ++q) valid &= !std::isnan(qpos_next[q]);
    if (valid) {
        for (int d = 0; d < NV; ++d) e.vel[d] = vel_next[d];
        for (int q = 0; q < NQ; ++q) e.qpos[q] = qpos_next[q];
    } else {
        e.err |= ERR_NAN_ACC;
    }
}

}  // namespace orc
#include "go2_oracle_collide.inc"
#include "go2_oracle_solve.inc"
#include "go2_oracle_api.inc"
