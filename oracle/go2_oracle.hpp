// TEST INFRASTRUCTURE ONLY - CPU oracle for the Go2 rigid step.  Not part of the product path.
//
// A plain C++ restatement (one env at a time, array-of-structs, scalar loops) of the rigid-body step
// that the reference executes through its quadrants DSL kernels.  Every function cites the reference
// file:line (relative to /root/reference/genesis/engine/solvers/rigid/) it follows, in the reference's
// own operation order.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference leg may load the library built from this file.
//
// PARITY UNPINNED: the reference ships no golden vectors for this path (it pins itself against a live
// MuJoCo that is not installed here) and cannot be imported in the build container (quadrants JIT is
// absent).  Documented deviations from the reference:
//   * GJK/EPA fallback (collider/narrowphase.py:727-845) is not restated: MPR's answer is kept.
//   * cylinder support uses the exact max over the 32-gon prism vertices instead of the 180x180
//     direction-grid lookup (collider/support_field.py:139-182); identical whenever the grid resolves
//     the best vertex.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>

namespace orc {

#ifdef ORC_DOUBLE
typedef double real;
static const real EPS = 1e-15;  // genesis/__init__.py:229  EPS = max(1e-15, finfo.eps)
#else
typedef float real;
static const real EPS = 1.1920929e-07f;
#endif

static const int NL = 14;   // links incl. ground (global index; ground = 0)
static const int NV = 18;   // dofs
static const int NQ = 19;
static const int NG_MAX = 32;
static const int MAX_PAIRS = 512;
static const int MAX_BROAD = 240;     // max_collision_pairs(30) * multiplier_collision_broad_phase(8)
static const int MAX_CONTACTS = 150;  // max_collision_pairs(30) * n_contacts_per_pair(5)  collider.py:364-372
static const int MAX_CON = 4 * MAX_CONTACTS + 12 + 18 + 48;  // constraint/solver.py:43-48

struct V3 {
    real x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(real a, real b, real c) : x(a), y(b), z(c) {}
    real& operator[](int i) { return (&x)[i]; }
    real operator[](int i) const { return (&x)[i]; }
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, real s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(real s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator/(V3 a, real s) { return V3(a.x / s, a.y / s, a.z / s); }
inline real dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline real norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalized(V3 a) { return a / norm(a); }

struct Q4 {
    real w, x, y, z;
    Q4() : w(1), x(0), y(0), z(0) {}
    Q4(real a, real b, real c, real d) : w(a), x(b), y(c), z(d) {}
};
struct M3 {
    real m[3][3];
};

// utils/geom.py:230-243 qd_quat_mul / qd_transform_quat_by_quat(v,u) = normalize(u*v)
inline Q4 quat_mul(Q4 u, Q4 v) {
    return Q4(u.w * v.w - u.x * v.x - u.y * v.y - u.z * v.z, u.w * v.x + u.x * v.w + u.y * v.z - u.z * v.y,
              u.w * v.y - u.x * v.z + u.y * v.w + u.z * v.x, u.w * v.z + u.x * v.y - u.y * v.x + u.z * v.w);
}
inline Q4 transform_quat_by_quat(Q4 v, Q4 u) {
    Q4 r = quat_mul(u, v);
    real n = std::sqrt(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z);
    return Q4(r.w / n, r.x / n, r.y / n, r.z / n);
}
inline Q4 inv_quat(Q4 q) { return Q4(q.w, -q.x, -q.y, -q.z); }
// utils/geom.py:246-262 qd_transform_by_quat
inline V3 transform_by_quat(V3 v, Q4 q) {
    real xx = q.x * q.x, xy = q.x * q.y, xz = q.x * q.z, wx = q.x * q.w;
    real yy = q.y * q.y, yz = q.y * q.z, wy = q.y * q.w, zz = q.z * q.z, wz = q.z * q.w, ww = q.w * q.w;
    V3 r(v.x * (xx + ww - yy - zz) + v.y * (2 * xy - 2 * wz) + v.z * (2 * xz + 2 * wy),
         v.x * (2 * wz + 2 * xy) + v.y * (ww - xx + yy - zz) + v.z * (-2 * wx + 2 * yz),
         v.x * (-2 * wy + 2 * xz) + v.y * (2 * wx + 2 * yz) + v.z * (ww - xx - yy + zz));
    return r / (ww + xx + yy + zz);
}
inline V3 inv_transform_by_quat(V3 v, Q4 q) { return transform_by_quat(v, inv_quat(q)); }
inline V3 transform_by_trans_quat(V3 p, V3 t, Q4 q) { return transform_by_quat(p, q) + t; }
// utils/geom.py:108-127 qd_rotvec_to_quat
inline Q4 rotvec_to_quat(V3 r) {
    Q4 q(0, 0, 0, 0);
    real thetasq = dot(r, r);
    if (thetasq > EPS * EPS) {
        real theta = std::sqrt(thetasq), h = real(0.5) * theta;
        real c = std::cos(h), s = std::sin(h);
        V3 xyz = (s / theta) * r;
        q = Q4(c, xyz.x, xyz.y, xyz.z);
        real k = real(0.5) * (real(3.0) - (q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z));
        q = Q4(q.w * k, q.x * k, q.y * k, q.z * k);
    } else {
        q = Q4(1, 0, 0, 0);
    }
    return q;
}
// utils/geom.py:130-151 qd_quat_to_R
inline M3 quat_to_R(Q4 q) {
    M3 R = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};
    real d = q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z;
    if (d > EPS) {
        real s = real(2.0) / d;
        real xs = q.x * s, ys = q.y * s, zs = q.z * s;
        real wx = q.w * xs, wy = q.w * ys, wz = q.w * zs, xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
        real yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
        M3 r = {{{1 - (yy + zz), xy - wz, xz + wy}, {xy + wz, 1 - (xx + zz), yz - wx}, {xz - wy, yz + wx, 1 - (xx + yy)}}};
        R = r;
    }
    return R;
}
// utils/geom.py:154-183 qd_quat_to_xyz
inline V3 quat_to_xyz(Q4 q) {
    real roll = 0, pitch = 0, yaw = 0;
    real n2 = q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z;
    if (n2 > EPS) {
        real s = real(2.0) / n2;
        real xs = q.x * s, ys = q.y * s, zs = q.z * s;
        real wx = q.w * xs, wy = q.w * ys, wz = q.w * zs, xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
        real yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
        real sinycosp = wz - xy, cosycosp = 1 - (yy + zz);
        real cosp = std::sqrt(cosycosp * cosycosp + sinycosp * sinycosp);
        pitch = std::atan2(xz + wy, cosp);
        if (cosp > EPS) {
            roll = std::atan2(wx - yz, 1 - (xx + yy));
            yaw = std::atan2(sinycosp, cosycosp);
        } else {
            yaw = std::atan2(wz + xy, 1 - (xx + zz));
        }
    }
    return V3(roll, pitch, yaw);
}
inline V3 mat_vec(const M3& A, V3 v) {
    return V3(A.m[0][0] * v.x + A.m[0][1] * v.y + A.m[0][2] * v.z, A.m[1][0] * v.x + A.m[1][1] * v.y + A.m[1][2] * v.z,
              A.m[2][0] * v.x + A.m[2][1] * v.y + A.m[2][2] * v.z);
}
// utils/geom.py:376-388 inertial_mul / motion_cross_force / motion_cross_motion
inline void inertial_mul(V3 pos, const M3& I, real mass, V3 vel, V3 ang, V3& o_ang, V3& o_vel) {
    o_ang = mat_vec(I, ang) + cross(pos, vel);
    o_vel = mass * vel - cross(pos, ang);
}
inline void motion_cross_force(V3 m_ang, V3 m_vel, V3 f_ang, V3 f_vel, V3& o_ang, V3& o_vel) {
    o_vel = cross(m_ang, f_vel);
    o_ang = cross(m_ang, f_ang) + cross(m_vel, f_vel);
}
inline void motion_cross_motion(V3 s_ang, V3 s_vel, V3 m_ang, V3 m_vel, V3& o_ang, V3& o_vel) {
    o_vel = cross(s_ang, m_vel) + cross(s_vel, m_ang);
    o_ang = cross(s_ang, m_ang);
}
// utils/geom.py:391-405 qd_orthogonals
inline void orthogonals(V3 a, V3& b, V3& c) {
    if (std::fabs(a.y) < real(0.5)) {
        b = V3(-a.x * a.y, 1 - a.y * a.y, -a.z * a.y);
    } else {
        b = V3(-a.x * a.z, -a.y * a.z, 1 - a.z * a.z);
    }
    b = normalized(b);
    c = cross(a, b);
}
// utils/geom.py:408-424 imp_aref
inline void imp_aref(const real* p, real neg_pen, real vel, real pos, real& imp, real& aref) {
    real timeconst = p[0], dampratio = p[1], dmin = p[2], dmax = p[3], width = p[4], mid = p[5], power = p[6];
    real imp_x = std::fabs(neg_pen) / width;
    real imp_a = (real(1.0) / std::pow(mid, power - 1)) * std::pow(imp_x, power);
    real imp_b = 1 - (real(1.0) / std::pow(1 - mid, power - 1)) * std::pow(1 - imp_x, power);
    real imp_y = imp_x < mid ? imp_a : imp_b;
    imp = dmin + imp_y * (dmax - dmin);
    imp = std::min(std::max(imp, dmin), dmax);
    if (imp_x > real(1.0)) imp = dmax;
    real b = real(2.0) / (dmax * timeconst);
    real k = real(1.0) / (dmax * dmax * timeconst * timeconst * dampratio * dampratio);
    aref = -b * vel - k * imp * pos;
}

enum { GEOM_SPHERE = 1, GEOM_CYLINDER = 3, GEOM_BOX = 5, GEOM_TERRAIN = 7 };
enum { CTRL_FORCE = 0, CTRL_VELOCITY = 1, CTRL_POSITION = 2 };
enum { ERR_BROAD_OVERFLOW = 1, ERR_CONTACT_OVERFLOW = 2, ERR_NAN_FORCE = 4, ERR_NAN_ACC = 8 };

struct Model {
    int ng, npairs;
    // links (global idx; 0 = ground)
    int parent[NL];       // global parent idx, -1 for roots
    int dof_start[NL], n_dofs[NL], q_start[NL];
    V3 l_pos[NL];
    Q4 l_quat[NL];
    V3 inertial_pos[NL];
    Q4 inertial_quat[NL];
    M3 inertial_i[NL];
    real inertial_mass[NL];
    V3 axis[NL];
    real limit[NV][2], force_range[NV][2], armature[NV], invweight_dof[NV];
    real invweight_link[NL];
    // geoms
    int g_link[NG_MAX], g_type[NG_MAX];
    V3 g_pos[NG_MAX], g_data[NG_MAX], g_aabb_min[NG_MAX], g_aabb_max[NG_MAX];
    Q4 g_quat[NG_MAX];
    real g_friction[NG_MAX];
    int pair_idx[NG_MAX][NG_MAX];
    real sol_params[7];
    real meaninertia, dt;
    V3 gravity;
    real kp[NV], kv[NV];
    int mass_parent_mask[NV][NV];
    // heightfield terrain replacing the ground box (collider/collider.py:374-394 _init_terrain_state)
    bool has_terrain;
    const real* terrain_hf;      // [rc0][rc1] metres
    int terrain_rc[2];
    real terrain_scale;          // horizontal scale
    real terrain_xyz_maxmin[6];  // x max, y max, z max, x min, y min, z min - 1
};

struct Contact {
    int geom_a, geom_b, link_a, link_b;
    V3 pos, normal;
    real penetration, friction;
};

struct Env {
    // persistent state
    real qpos[NQ], vel[NV], qacc_ws[NV];
    bool is_warmstart, first_time;
    V3 cache_normal[MAX_PAIRS];
    int sort_ig[2 * NG_MAX];
    bool sort_ismax[2 * NG_MAX];
    real sort_val[2 * NG_MAX];
    int ctrl_mode[NV];
    real ctrl_force[NV], ctrl_pos[NV], ctrl_vel[NV];
    V3 cfrc_applied_ang[NL], cfrc_applied_vel[NL];
    real mass_shift[NL];
    V3 i_pos_shift[NL];
    int err;
    // derived (refreshed by forward kinematics / velocity)
    V3 pos[NL], root_COM, i_pos[NL], cinr_pos[NL], xanchor[NL], xaxis[NL];
    Q4 quat[NL], i_quat[NL];
    M3 cinr_inertial[NL];
    real cinr_mass[NL];
    real dpos[NV];
    V3 cdof_ang[NV], cdof_vel[NV], cdofd_ang[NV], cdofd_vel[NV];
    V3 cd_ang[NL], cd_vel[NL], cdd_ang[NL], cdd_vel[NL], cfrc_ang[NL], cfrc_vel[NL];
    V3 gpos[NG_MAX], aabb_min[NG_MAX], aabb_max[NG_MAX];
    Q4 gquat[NG_MAX];
    V3 prism[6];          // collider_state.prism (terrain narrow phase)
    real xyz_max_min[6];  // collider_state.xyz_max_min
    // dynamics
    real M[NV][NV], ML[NV][NV], MDinv[NV];
    real qf_applied[NV], qf_bias[NV], force[NV], qf_smooth[NV], acc_smooth[NV], acc[NV];
    // collision
    int n_broad, broad[MAX_BROAD][2], n_contacts;
    Contact contacts[MAX_CONTACTS];
    V3 contact_force[NL];
    // constraints
    int n_con;
    std::vector<real> jac;  // [MAX_CON][NV]
    real diag[MAX_CON], aref[MAX_CON], efc_D[MAX_CON], efc_force[MAX_CON], Jaref[MAX_CON], jv[MAX_CON];
    bool active[MAX_CON], prev_active[MAX_CON];
    real qacc[NV], Ma[NV], grad[NV], Mgrad[NV], search[NV], mv[NV], qfrc_constraint[NV], H[NV][NV];
    real cost, prev_cost, gauss, quad_gauss[3];
    int ls_it, n_newton_iters;
    bool improved;
    // diagnostics: detections for which the reference would prefer GJK over MPR's answer (narrowphase.py:724-736), see
    // convex_convex_contact: [sphere pair, other pair first detection, other pair perturbed detection, all MPR collisions]
    long gjk_wanted[4] = {0, 0, 0, 0};
};

}  // namespace orc
