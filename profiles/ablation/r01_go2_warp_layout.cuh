// go2_b200 - layout v2: ONE WARP PER ENVIRONMENT, working set in shared memory.
//
// Chosen from the v1 ncu evidence (profiles/r01_v1_*): one-thread-per-env spilled 30 KB of thread-local
// Jacobian / Hessian data per env through L1/L2 to HBM (18-27x the algorithmic bytes), ran 12 warps/SM and lost
// half of the lanes to divergence between envs.  Here the 32 lanes of a warp cooperate on one env: lanes map
// to legs (4), links (13), dofs (18), geoms (28), candidate pairs and constraint rows; everything between the
// state load and the state store lives in `EnvS` (shared memory, ~12 KB per env).
//
// The code is written in "lane-loop" form so that the very same source runs on the host for the CPU test
// suite (tests/host_emul): statements inside LANES(l){...} are executed by lane l; statements outside are
// warp-uniform (every lane executes them redundantly on the device, once on the host); WSYNC() separates a
// phase that writes EnvS from a phase that reads it.
#pragma once
#include "go2_collide.cuh"

namespace go2 {

#ifdef __CUDA_ARCH__
#define LANES(l) for (int l = (int)(threadIdx.x & 31u), _g2_once = 1; _g2_once; _g2_once = 0)
#define WSYNC() __syncwarp()
#else
#define LANES(l) for (int l = 0; l < 32; ++l)
#define WSYNC()
#endif

constexpr int W_MAXC = 16;                     // contacts per env kept by the warp layout
constexpr int W_MAXR = 4 * W_MAXC + NHINGE;    // 76 constraint rows
constexpr int W_MAXCAND = 64;                  // broad-phase candidates per env
constexpr int W_MAXTASK = 64;                  // perturbed re-detections (4 per colliding non-sphere pair)

struct WContact {
    v3 pos, normal;
    float pen, mu;
    int la, lb;
};

struct EnvS {
    // generalized state (+ warm start, compact contact-normal cache)
    float qpos[NQ], qvel[NV], qacc_ws[NV];
    int warm, n_cache;
    int cache_pair[MAX_CPAIRS];
    v3 cache_n[MAX_CPAIRS];
    // actuation
    float ctrl_force[NV], ctrl_pos[NV];
    int pd_mode, ext_link;
    float pd_kp, pd_kv;
    v3 ext_f, ext_t;
    // kinematics
    v3 lpos[NLINK];
    q4 lquat[NLINK];
    v3 root_com;
    float cin_I[NLINK][6];
    v3 cin_h[NLINK];
    float cin_m[NLINK];
    v3 cdof_a[NV], cdof_v[NV], cdofd_a[NV], cdofd_v[NV];
    v3 cd_a[NLINK], cd_v[NLINK];
    // scratch per link (composite inertias, then RNE forces)
    float sc_I[NLINK][6];
    v3 sc_h[NLINK], sc_a[NLINK], sc_v[NLINK];
    float sc_m[NLINK];
    float red[32], red2[32], red3[32];
    float schur[4][21];
    // dynamics
    float M[NTRI], L[NTRI], Dinv[NV];
    float force[NV], qf_smooth[NV], acc_smooth[NV], acc[NV];
    // collision
    v3 g_c[NGEOM], g_h[NGEOM];
    int n_cand, n_task, n_colpair;
    short cand_pair[W_MAXCAND];
    unsigned char flags[32];
    // per candidate: first detection result
    unsigned char c_col[W_MAXCAND], c_ncon[W_MAXCAND];
    v3 c_n[W_MAXCAND], c_pos[W_MAXCAND];
    float c_pen[W_MAXCAND];
    // perturbed detection results: task t = (candidate slot, k)
    short t_cand[W_MAXTASK];
    unsigned char t_k[W_MAXTASK], t_col[W_MAXTASK];
    v3 t_n[W_MAXTASK], t_pos[W_MAXTASK];
    float t_pen[W_MAXTASK];
    int n_contacts, n_rows;
    WContact con[W_MAXC];
    // constraints + solver
    float J[W_MAXR][NV];
    float Dd[W_MAXR], aref[W_MAXR], Jaref[W_MAXR], jv[W_MAXR], efc[W_MAXR];
    unsigned char act[W_MAXR], pact[W_MAXR];
    float H[NTRI];
    float qacc[NV], Ma[NV], grad[NV], Mgrad[NV], search[NV], mv[NV], qfc[NV], vec[NV];
    float cost, prev_cost, gauss, qg0, qg1, qg2;
    int ls_it, n_newton, err;
    v3 cforce[NLINK];
};

// ---- warp helpers (uniform code) ------------------------------------------------------------------
// sum of red[0..31]; every lane gets the result
G2_HD float wsum32(const float* red) {
#ifdef __CUDA_ARCH__
    float v = red[threadIdx.x & 31u];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
#else
    float s = 0.0f;
    for (int i = 0; i < 32; ++i) s += red[i];
    return s;
#endif
}
// bit mask of flags[0..31] != 0
G2_HD unsigned wmask32(const unsigned char* flags) {
#ifdef __CUDA_ARCH__
    return __ballot_sync(0xffffffffu, flags[threadIdx.x & 31u] != 0);
#else
    unsigned m = 0;
    for (int i = 0; i < 32; ++i)
        if (flags[i]) m |= 1u << i;
    return m;
#endif
}
G2_HD int lowbit(unsigned m) {
#ifdef __CUDA_ARCH__
    return __ffs((int)m) - 1;
#else
    return __builtin_ctz(m);
#endif
}

G2_HD int leg_link(int leg, int level) { return 1 + leg + 4 * level; }  // hip / thigh / calf link of a leg (BFS order)

// ---------------------------------------------------------------------------------------------------
// Kinematics: lanes 0-3 walk one leg each; lanes 0-12 own one link; lanes 0-17 own one dof.
// ---------------------------------------------------------------------------------------------------
G2_HD void w_kinematics(const Go2Model& m, const float* mass_shift, const float* com_shift, EnvS& s) {
    LANES(l) {
        if (l == 0) {
            s.lpos[0] = mk3(s.qpos[0], s.qpos[1], s.qpos[2]);
            s.lquat[0] = qnormalize(mk4(s.qpos[3], s.qpos[4], s.qpos[5], s.qpos[6]));
        }
    }
    WSYNC();
    LANES(l) {
        if (l < 4) {
            v3 ppos = s.lpos[0];
            q4 pq = s.lquat[0];
            for (int lev = 0; lev < 3; ++lev) {
                int k = leg_link(l, lev);
                v3 pos = ppos + qrot(ld3(m.l_pos[k]), pq);
                q4 quat = qcompose(ld4(m.l_quat[k]), pq);
                q4 qloc = rotvec_quat(ld3(m.axis[k]) * s.qpos[k + 6]);
                pq = qcompose(qloc, quat);
                ppos = pos;
                s.lpos[k] = pos;
                s.lquat[k] = pq;
            }
        }
    }
    WSYNC();
    // mass-weighted link COMs -> root COM
    LANES(l) {
        if (l < NLINK) {
            float mass = m.mass[l] + mass_shift[l];
            v3 ip = ld3(m.inertial_pos[l]);
            if (l == 0) ip = ip + mk3(com_shift[0], com_shift[1], com_shift[2]);
            v3 c = s.lpos[l] + qrot(ip, s.lquat[l]);
            s.sc_a[l] = c;
            s.cin_m[l] = mass;
        }
    }
    WSYNC();
    {
        v3 acc = mk3(0, 0, 0);
        float ms = 0.0f;
        for (int k = 0; k < NLINK; ++k) {
            ms += s.cin_m[k];
            acc = acc + s.cin_m[k] * s.sc_a[k];
        }
        v3 rc = acc * (1.0f / ms);
        LANES(l) { if (l == 0) s.root_com = rc; }
        WSYNC();
    }
    LANES(l) {
        if (l < NLINK) {
            v3 t = s.sc_a[l] - s.root_com;
            float mass = s.cin_m[l];
            m3 R = qmat(qcompose(ld4(m.inertial_quat[l]), s.lquat[l]));
            const float* I = m.inertial_i[l];
            float RI[3][3];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) RI[i][j] = R.a[i][0] * I[j] + R.a[i][1] * I[3 + j] + R.a[i][2] * I[6 + j];
            float W[3][3];
            for (int i = 0; i < 3; ++i)
                for (int j = i; j < 3; ++j) W[i][j] = RI[i][0] * R.a[j][0] + RI[i][1] * R.a[j][1] + RI[i][2] * R.a[j][2];
            s.cin_I[l][0] = W[0][0] + mass * (t.y * t.y + t.z * t.z);
            s.cin_I[l][1] = W[0][1] - mass * t.x * t.y;
            s.cin_I[l][2] = W[0][2] - mass * t.x * t.z;
            s.cin_I[l][3] = W[1][1] + mass * (t.x * t.x + t.z * t.z);
            s.cin_I[l][4] = W[1][2] - mass * t.y * t.z;
            s.cin_I[l][5] = W[2][2] + mass * (t.x * t.x + t.y * t.y);
            s.cin_h[l] = t * mass;
        }
        if (l < NV) {
            if (l < 3) {
                s.cdof_a[l] = mk3(0, 0, 0);
                s.cdof_v[l] = mk3(l == 0 ? 1.0f : 0.0f, l == 1 ? 1.0f : 0.0f, l == 2 ? 1.0f : 0.0f);
            } else if (l < 6) {
                v3 c = mcol(qmat(s.lquat[0]), l - 3);
                s.cdof_a[l] = c;
                s.cdof_v[l] = cross(c, s.root_com - s.lpos[0]);
            } else {
                int k = l - 5;
                v3 ax = qrot(ld3(m.axis[k]), s.lquat[k]);  // a hinge rotation leaves its own axis invariant
                s.cdof_a[l] = ax;
                s.cdof_v[l] = cross(ax, s.root_com - s.lpos[k]);
            }
        }
    }
    WSYNC();
}

G2_HD void w_velocities(const Go2Model& m, EnvS& s) {
    LANES(l) {
        if (l == 0) {
            v3 cv = mk3(0, 0, 0), ca = mk3(0, 0, 0);
            for (int i = 0; i < 3; ++i) {
                cv = cv + s.cdof_v[i] * s.qvel[i];
                s.cdofd_a[i] = mk3(0, 0, 0);
                s.cdofd_v[i] = mk3(0, 0, 0);
            }
            for (int i = 3; i < 6; ++i) {
                s.cdofd_a[i] = mk3(0, 0, 0);  // base angular velocity is still zero at this point of the recursion
                s.cdofd_v[i] = cross(cv, s.cdof_a[i]);
            }
            for (int i = 3; i < 6; ++i) {
                cv = cv + s.cdof_v[i] * s.qvel[i];
                ca = ca + s.cdof_a[i] * s.qvel[i];
            }
            s.cd_v[0] = cv;
            s.cd_a[0] = ca;
        }
    }
    WSYNC();
    LANES(l) {
        if (l < 4) {
            v3 pv = s.cd_v[0], pa = s.cd_a[0];
            for (int lev = 0; lev < 3; ++lev) {
                int k = leg_link(l, lev), d = 5 + k;
                s.cdofd_a[d] = cross(pa, s.cdof_a[d]);
                s.cdofd_v[d] = cross(pa, s.cdof_v[d]) + cross(pv, s.cdof_a[d]);
                pv = pv + s.cdof_v[d] * s.qvel[d];
                pa = pa + s.cdof_a[d] * s.qvel[d];
                s.cd_v[k] = pv;
                s.cd_a[k] = pa;
            }
        }
    }
    WSYNC();
}


// ---------------------------------------------------------------------------------------------------
// Smooth dynamics.  Mass matrix rows on lanes 0-17 (only the tree-structured nonzeros), LtDL per leg on
// lanes 0-3 with the four Schur complements of the base block summed afterwards, RNE per link / per leg.
// ---------------------------------------------------------------------------------------------------
G2_HD void w_zero_M(EnvS& s) {
    LANES(l) {
        for (int k = l; k < NTRI; k += 32) s.M[k] = 0.0f;
    }
    WSYNC();
}

G2_HD void w_mass_matrix(const Go2Model& m, EnvS& s, float diag_extra_hinge) {
    // composite inertias: each leg bottom-up, then the base
    LANES(l) {
        if (l < 4) {
            float I[6];
            v3 h = mk3(0, 0, 0);
            float mm = 0.0f;
            for (int k = 0; k < 6; ++k) I[k] = 0.0f;
            for (int lev = 2; lev >= 0; --lev) {
                int k = leg_link(l, lev);
                for (int q = 0; q < 6; ++q) I[q] += s.cin_I[k][q];
                h = h + s.cin_h[k];
                mm += s.cin_m[k];
                for (int q = 0; q < 6; ++q) s.sc_I[k][q] = I[q];
                s.sc_h[k] = h;
                s.sc_m[k] = mm;
            }
        }
    }
    WSYNC();
    LANES(l) {
        if (l == 0) {
            float I[6];
            for (int q = 0; q < 6; ++q) I[q] = s.cin_I[0][q];
            v3 h = s.cin_h[0];
            float mm = s.cin_m[0];
            for (int leg = 0; leg < 4; ++leg) {
                int k = leg_link(leg, 0);
                for (int q = 0; q < 6; ++q) I[q] += s.sc_I[k][q];
                h = h + s.sc_h[k];
                mm += s.sc_m[k];
            }
            for (int q = 0; q < 6; ++q) s.sc_I[0][q] = I[q];
            s.sc_h[0] = h;
            s.sc_m[0] = mm;
        }
        for (int k = l; k < NTRI; k += 32) s.M[k] = 0.0f;
    }
    WSYNC();
    LANES(l) {
        if (l < NV) {
            int i = l, li = i < 6 ? 0 : i - 5;
            v3 fa, fv;
            inertia_mul(s.sc_I[li], s.sc_h[li], s.sc_m[li], s.cdof_v[i], s.cdof_a[i], fa, fv);
            if (i < 6) {
                for (int j = 0; j <= i; ++j) s.M[tri(i, j)] = dot(fa, s.cdof_a[j]) + dot(fv, s.cdof_v[j]);
            } else {
                int k = li;
                while (k > 0) {
                    int j = 5 + k;
                    s.M[tri(i, j)] = dot(fa, s.cdof_a[j]) + dot(fv, s.cdof_v[j]);
                    k = m.parent[k];
                }
                for (int j = 0; j < 6; ++j) s.M[tri(i, j)] = dot(fa, s.cdof_a[j]) + dot(fv, s.cdof_v[j]);
                s.M[tri(i, i)] += m.armature[i] + diag_extra_hinge;
            }
        }
    }
    WSYNC();
}

// LtDL (last dof first).  A leg's three dofs only touch that leg's rows and the base block, so the legs are
// eliminated in parallel, each accumulating its Schur complement of the 6x6 base block separately.
G2_HD void w_factor_mass(EnvS& s) {
    LANES(l) {
        for (int k = l; k < NTRI; k += 32) s.L[k] = s.M[k];
    }
    WSYNC();
    LANES(l) {
        if (l < 4) {
            int dl[3] = {5 + leg_link(l, 0), 5 + leg_link(l, 1), 5 + leg_link(l, 2)};  // hip < thigh < calf dof ids
            float sch[21];
            for (int k = 0; k < 21; ++k) sch[k] = 0.0f;
            for (int a = 2; a >= 0; --a) {
                int i = dl[a];
                float dinv = 1.0f / s.L[tri(i, i)];
                s.Dinv[i] = dinv;
                // ancestors of i in descending order: leg dofs above it, then base dofs 5..0
                for (int b = a - 1; b >= -6; --b) {
                    int j = b >= 0 ? dl[b] : 6 + b;  // b=-1 -> 5 ... b=-6 -> 0
                    float aij = s.L[tri(i, j)] * dinv;
                    // L[j][k] -= aij * L[i][k] for k <= j among the ancestors of i
                    if (j >= 6) {
                        for (int c = b; c >= 0; --c) s.L[tri(j, dl[c])] -= aij * s.L[tri(i, dl[c])];
                        for (int k = 0; k < 6; ++k) s.L[tri(j, k)] -= aij * s.L[tri(i, k)];
                    } else {
                        for (int k = 0; k <= j; ++k) sch[tri(j, k)] -= aij * s.L[tri(i, k)];
                    }
                    s.L[tri(i, j)] = aij;
                }
                s.L[tri(i, i)] = 1.0f;
            }
            for (int k = 0; k < 21; ++k) s.schur[l][k] = sch[k];
        }
    }
    WSYNC();
    LANES(l) {
        if (l < 21) s.L[l] = s.L[l] + ((s.schur[3][l] + s.schur[2][l]) + (s.schur[1][l] + s.schur[0][l]));
    }
    WSYNC();
    LANES(l) {
        if (l == 0) {
            for (int i = 5; i >= 0; --i) {
                float dinv = 1.0f / s.L[tri(i, i)];
                s.Dinv[i] = dinv;
                for (int j = i - 1; j >= 0; --j) {
                    float a = s.L[tri(i, j)] * dinv;
                    for (int k = 0; k <= j; ++k) s.L[tri(j, k)] -= a * s.L[tri(i, k)];
                    s.L[tri(i, j)] = a;
                }
                s.L[tri(i, i)] = 1.0f;
            }
        }
    }
    WSYNC();
}

// out = (L^T D L)^-1 vec, with the same leg-parallel structure
G2_HD void w_solve_mass(EnvS& s, const float* vec, float* out) {
    LANES(l) {
        if (l < 4) {
            int dl[3] = {5 + leg_link(l, 0), 5 + leg_link(l, 1), 5 + leg_link(l, 2)};
            for (int a = 2; a >= 0; --a) {
                float c = vec[dl[a]];
                for (int b = a + 1; b < 3; ++b) c -= s.L[tri(dl[b], dl[a])] * out[dl[b]];
                out[dl[a]] = c;
            }
        }
    }
    WSYNC();
    LANES(l) {
        if (l < 6) {
            float c = vec[l];
            for (int j = NV - 1; j >= 6; --j) c -= s.L[tri(j, l)] * out[j];
            s.vec[l] = c;
        }
    }
    WSYNC();
    LANES(l) {
        if (l == 0) {
            for (int i = 5; i >= 0; --i) {
                float c = s.vec[i];
                for (int j = i + 1; j < 6; ++j) c -= s.L[tri(j, i)] * out[j];
                out[i] = c;
            }
            for (int i = 0; i < 6; ++i) out[i] *= s.Dinv[i];
            for (int i = 0; i < 6; ++i) {
                float c = out[i];
                for (int j = 0; j < i; ++j) c -= s.L[tri(i, j)] * out[j];
                out[i] = c;
            }
        }
    }
    WSYNC();
    LANES(l) {
        if (l < 4) {
            int dl[3] = {5 + leg_link(l, 0), 5 + leg_link(l, 1), 5 + leg_link(l, 2)};
            for (int a = 0; a < 3; ++a) {
                int i = dl[a];
                float c = out[i] * s.Dinv[i];
                for (int j = 0; j < 6; ++j) c -= s.L[tri(i, j)] * out[j];
                for (int b = 0; b < a; ++b) c -= s.L[tri(i, dl[b])] * out[dl[b]];
                out[i] = c;
            }
        }
    }
    WSYNC();
}

G2_HD float w_applied_force(const Go2Model& m, const EnvS& s, int d) {
    float f;
    if (s.pd_mode) f = s.pd_kp * (s.ctrl_pos[d] - s.qpos[d + 1]) + s.pd_kv * (0.0f - s.qvel[d]);
    else f = s.ctrl_force[d];
    return clampf(f, -m.frange[d], m.frange[d]);
}

G2_HD void w_smooth_dynamics(const Go2Model& m, EnvS& s) {
    w_mass_matrix(m, s, s.pd_mode ? s.pd_kv * m.dt : 0.0f);
    w_factor_mass(s);
    // link accelerations without qacc: base on lane 0, then one leg per lane
    LANES(l) {
        if (l == 0) {
            v3 a = mk3(0, 0, 0), v = mk3(-m.gravity[0], -m.gravity[1], -m.gravity[2]);
            for (int d = 0; d < 6; ++d) {
                v = v + s.cdofd_v[d] * s.qvel[d];
                a = a + s.cdofd_a[d] * s.qvel[d];
            }
            s.sc_a[0] = a;
            s.sc_v[0] = v;
        }
    }
    WSYNC();
    LANES(l) {
        if (l < 4) {
            v3 a = s.sc_a[0], v = s.sc_v[0];
            for (int lev = 0; lev < 3; ++lev) {
                int k = leg_link(l, lev), d = 5 + k;
                v = v + s.cdofd_v[d] * s.qvel[d];
                a = a + s.cdofd_a[d] * s.qvel[d];
                s.sc_a[k] = a;
                s.sc_v[k] = v;
            }
        }
    }
    WSYNC();
    LANES(l) {
        if (l < NLINK) {
            v3 a1, v1, a2, v2;
            inertia_mul(s.cin_I[l], s.cin_h[l], s.cin_m[l], s.sc_v[l], s.sc_a[l], a1, v1);
            inertia_mul(s.cin_I[l], s.cin_h[l], s.cin_m[l], s.cd_v[l], s.cd_a[l], a2, v2);
            v3 fv = v1 + cross(s.cd_a[l], v2);
            v3 fa = a1 + cross(s.cd_a[l], a2) + cross(s.cd_v[l], v2);
            if (l == s.ext_link) {
                fv = fv - s.ext_f;
                fa = fa - s.ext_t;
            }
            s.sc_v[l] = fv;  // now: per-link force (vel part) / torque (ang part)
            s.sc_a[l] = fa;
        }
    }
    WSYNC();
    LANES(l) {
        if (l < 4) {
            v3 fv = mk3(0, 0, 0), fa = mk3(0, 0, 0);
            for (int lev = 2; lev >= 0; --lev) {
                int k = leg_link(l, lev);
                fv = fv + s.sc_v[k];
                fa = fa + s.sc_a[k];
                s.sc_v[k] = fv;
                s.sc_a[k] = fa;
            }
        }
    }
    WSYNC();
    LANES(l) {
        if (l == 0) {
            v3 fv = s.sc_v[0], fa = s.sc_a[0];
            for (int leg = 3; leg >= 0; --leg) {
                fv = fv + s.sc_v[leg_link(leg, 0)];
                fa = fa + s.sc_a[leg_link(leg, 0)];
            }
            s.sc_v[0] = fv;
            s.sc_a[0] = fa;
        }
    }
    WSYNC();
    LANES(l) {
        if (l < NV) {
            int k = l < 6 ? 0 : l - 5;
            float bias = dot(s.cdof_a[l], s.sc_a[k]) + dot(s.cdof_v[l], s.sc_v[k]);
            float ap = l < 6 ? 0.0f : w_applied_force(m, s, l);
            s.force[l] = ap - bias;
            s.qf_smooth[l] = ap - bias;
        }
    }
    WSYNC();
    w_solve_mass(s, s.force, s.acc_smooth);
    LANES(l) {
        if (l < NV) s.acc[l] = s.acc_smooth[l];
    }
    WSYNC();
}

G2_HD void w_integrate(const Go2Model& m, EnvS& s) {
    LANES(l) {
        if (l < NV) s.vec[l] = s.qvel[l] + s.acc[l] * m.dt;
    }
    WSYNC();
    bool ok = true;
    for (int d = 0; d < NV; ++d) ok = ok && (s.vec[d] == s.vec[d]);
    q4 r1;
    {
        q4 r0 = mk4(s.qpos[3], s.qpos[4], s.qpos[5], s.qpos[6]);
        q4 dq = rotvec_quat(mk3(s.vec[3], s.vec[4], s.vec[5]) * m.dt);
        r1 = qnormalize(qmul(r0, dq));
        ok = ok && (r1.w == r1.w) && (r1.x == r1.x) && (r1.y == r1.y) && (r1.z == r1.z);
        for (int j = 0; j < 3; ++j) ok = ok && ((s.qpos[j] + s.vec[j] * m.dt) == (s.qpos[j] + s.vec[j] * m.dt));
        for (int d = 6; d < NV; ++d) ok = ok && ((s.qpos[d + 1] + s.vec[d] * m.dt) == (s.qpos[d + 1] + s.vec[d] * m.dt));
    }
    WSYNC();
    LANES(l) {
        if (ok) {
            if (l < NV) s.qvel[l] = s.vec[l];
            if (l < 3) s.qpos[l] = s.qpos[l] + s.vec[l] * m.dt;
            if (l >= 6 && l < NV) s.qpos[l + 1] = s.qpos[l + 1] + s.vec[l] * m.dt;
            if (l == 3) { s.qpos[3] = r1.w; s.qpos[4] = r1.x; s.qpos[5] = r1.y; s.qpos[6] = r1.z; }
        } else if (l == 0) {
            s.err |= ERR_NAN_ACC;
        }
    }
    WSYNC();
}


// ---------------------------------------------------------------------------------------------------
// Collision detection: lane per geom (AABBs), lane per static pair (AABB test, ballot-compacted in pair
// order), lane per candidate pair (first MPR detection), lane per (colliding non-sphere pair, k) for the four
// perturbed re-detections, then a short serial pass that de-duplicates and orders the contacts.
// ---------------------------------------------------------------------------------------------------
G2_HD void w_pair_ctx(const Go2Model& m, const EnvS& s, int ip, int& ga, int& gb, GeomPose& A, GeomPose& B, PairCtx& c) {
    int a = m.pair_a[ip], b = m.pair_b[ip];
    ga = a; gb = b;
    if (m.g_type[ga] > m.g_type[gb]) { ga = b; gb = a; }
    A = geom_pose(m, s, ga);
    B = geom_pose(m, s, gb);
    c.ga = ga; c.gb = gb; c.pa = A.pos; c.pb = B.pos; c.Ra = qmat(A.quat); c.Rb = qmat(B.quat); c.prism = nullptr;
}

G2_HD void w_detect_collisions(const Go2Model& m, EnvS& s, float friction) {
    LANES(l) {
        if (l < NGEOM) {
            GeomPose gp = geom_pose(m, s, l);
            m3 R = qmat(gp.quat);
            v3 h = ld3(m.g_half[l]);
            s.g_c[l] = gp.pos;
            s.g_h[l] = mk3(fabsf(R.a[0][0]) * h.x + fabsf(R.a[0][1]) * h.y + fabsf(R.a[0][2]) * h.z,
                           fabsf(R.a[1][0]) * h.x + fabsf(R.a[1][1]) * h.y + fabsf(R.a[1][2]) * h.z,
                           fabsf(R.a[2][0]) * h.x + fabsf(R.a[2][1]) * h.y + fabsf(R.a[2][2]) * h.z);
        }
    }
    WSYNC();
    int n_cand = 0;
    for (int base = 0; base < m.npairs; base += 32) {
        LANES(l) {
            int ip = base + l;
            unsigned char f = 0;
            if (ip < m.npairs) {
                int a = m.pair_a[ip], b = m.pair_b[ip];
                v3 dc = s.g_c[a] - s.g_c[b], hs = s.g_h[a] + s.g_h[b];
                f = (fabsf(dc.x) < hs.x && fabsf(dc.y) < hs.y && fabsf(dc.z) < hs.z) ? 1 : 0;
            }
            s.flags[l] = f;
        }
        WSYNC();
        unsigned mask = wmask32(s.flags);
        while (mask) {
            int i = lowbit(mask);
            mask &= mask - 1;
            if (n_cand < W_MAXCAND) {
                LANES(l) { if (l == 0) s.cand_pair[n_cand] = (short)(base + i); }
                n_cand++;
            } else {
                LANES(l) { if (l == 0) s.err |= ERR_CPAIR_OVERFLOW; }
            }
        }
        WSYNC();
    }
    // first detection, one candidate per lane
    for (int base = 0; base < n_cand; base += 32) {
        LANES(l) {
            int ic = base + l;
            if (ic < n_cand) {
                int ip = s.cand_pair[ic], ga, gb;
                GeomPose A, B;
                PairCtx c;
                w_pair_ctx(m, s, ip, ga, gb, A, B, c);
                v3 nws = mk3(0, 0, 0);
                bool have = false;
                for (int k = 0; k < s.n_cache; ++k)
                    if (s.cache_pair[k] == ip) { nws = s.cache_n[k]; have = true; }
                v3 ca, cb, n0 = mk3(0, 0, 0), pos0 = mk3(0, 0, 0);
                float pen0 = 0.0f;
                seed_centres(m, c, nws, have, ca, cb);
                bool col = mpr(m, c, ca, cb, n0, pen0, pos0);
                if (!col && have) {
                    seed_centres(m, c, nws, false, ca, cb);
                    col = mpr(m, c, ca, cb, n0, pen0, pos0);
                }
                s.c_col[ic] = col ? 1 : 0;
                s.c_n[ic] = n0; s.c_pos[ic] = pos0; s.c_pen[ic] = pen0;
            }
        }
    }
    WSYNC();
    // perturbed re-detections for colliding pairs without a sphere: 4 tasks per pair
    int n_task = 0;
    for (int ic = 0; ic < n_cand; ++ic) {
        if (s.c_col[ic]) {
            int ip = s.cand_pair[ic];
            bool multi = m.g_type[m.pair_a[ip]] != GEOM_SPHERE && m.g_type[m.pair_b[ip]] != GEOM_SPHERE;
            if (multi && n_task + 4 <= W_MAXTASK) {
                LANES(l) { if (l < 4) { s.t_cand[n_task + l] = (short)ic; s.t_k[n_task + l] = (unsigned char)(l + 1); } }
                n_task += 4;
            }
        }
    }
    WSYNC();
    for (int base = 0; base < n_task; base += 32) {
        LANES(l) {
            int it = base + l;
            if (it < n_task) {
                int ic = s.t_cand[it], k = s.t_k[it], ip = s.cand_pair[ic], ga, gb;
                GeomPose A, B;
                PairCtx c;
                w_pair_ctx(m, s, ip, ga, gb, A, B, c);
                v3 n0 = s.c_n[ic], pos0 = s.c_pos[ic];
                // perturbation axes from the principal axes of the smaller geom's link inertia frame
                float va = m.g_half[ga][0] * m.g_half[ga][1] * m.g_half[ga][2], vb = m.g_half[gb][0] * m.g_half[gb][1] * m.g_half[gb][2];
                int lk = m.g_link[va < vb ? ga : gb];
                m3 R;
                if (lk < 0) { q4 id = mk4(1, 0, 0, 0); R = qmat(id); }
                else R = qmat(qcompose(ld4(m.inertial_quat[lk]), s.lquat[lk]));
                int best = 0;
                float amax = 0.0f;
                for (int i = 0; i < 3; ++i) {
                    float a = fabsf(dot(mcol(R, i), n0));
                    if (a > amax) { amax = a; best = i; }
                }
                v3 ax0 = mcol(R, (best + 1) % 3);
                ax0 = normalize3(ax0 - dot(n0, ax0) * n0);
                v3 ax1 = cross(n0, ax0);
                v3 axis = (float)(2 * (k % 2) - 1) * ax0 + (float)(1 - 2 * ((k / 2) % 2)) * ax1;
                q4 qr = rotvec_quat(axis * MC_PERTURB), qri = qconj(qr);
                PairCtx cp = c;
                v3 rel = pos0 - A.pos;
                cp.pa = A.pos - (qrot(rel, qr) - rel);
                cp.Ra = qmat(qcompose(A.quat, qr));
                rel = pos0 - B.pos;
                cp.pb = B.pos - (qrot(rel, qri) - rel);
                cp.Rb = qmat(qcompose(B.quat, qri));
                v3 ca, cb, n = mk3(0, 0, 0), pos = mk3(0, 0, 0);
                float pen = 0.0f;
                seed_centres(m, cp, n0, true, ca, cb);
                bool col = mpr(m, cp, ca, cb, n, pen, pos);
                if (col) {
                    v3 pA = qrot((pos - 0.5f * pen * n) - pos0, qri) + pos0;
                    v3 pB = qrot((pos + 0.5f * pen * n) - pos0, qr) + pos0;
                    pos = 0.5f * (pA + pB);
                    v3 tw = cross(n, n0);
                    tw = mk3(clampf(tw.x, -MC_PERTURB, MC_PERTURB), clampf(tw.y, -MC_PERTURB, MC_PERTURB), clampf(tw.z, -MC_PERTURB, MC_PERTURB));
                    n = n + cross(tw, n);
                    pen = dot(n, pB - pA);
                }
                s.t_col[it] = col ? 1 : 0;
                s.t_n[it] = n; s.t_pos[it] = pos; s.t_pen[it] = pen;
            }
        }
    }
    WSYNC();
    // serial pass (lane 0): contact list in candidate order, duplicate filtering, new normal cache
    LANES(l) {
        if (l == 0) {
            float mu = fmaxf(friction, 1e-2f);
            int nc = 0, ncache = 0, it = 0;
            for (int ic = 0; ic < n_cand; ++ic) {
                if (!s.c_col[ic]) continue;
                int ip = s.cand_pair[ic], ga = m.pair_a[ip], gb = m.pair_b[ip];
                if (m.g_type[ga] > m.g_type[gb]) { int t = ga; ga = gb; gb = t; }
                if (ncache < MAX_CPAIRS) { s.cache_pair[ncache] = ip; s.cache_n[ncache] = s.c_n[ic]; ncache++; }
                else s.err |= ERR_CPAIR_OVERFLOW;
                int first = nc;
                if (nc < W_MAXC) {
                    WContact& c = s.con[nc++];
                    c.pos = s.c_pos[ic]; c.normal = s.c_n[ic]; c.pen = s.c_pen[ic]; c.mu = mu; c.la = m.g_link[ga]; c.lb = m.g_link[gb];
                } else s.err |= ERR_CONTACT_OVERFLOW;
                if (it < n_task && s.t_cand[it] == ic) {
                    float tol = 0.5f * MC_TOL * fminf(aabb_diag(m, ga), aabb_diag(m, gb));
                    for (int k = 0; k < 4; ++k, ++it) {
                        if (!s.t_col[it]) continue;
                        bool rep = false;
                        for (int q = first; q < nc && !rep; ++q)
                            if (norm3(s.t_pos[it] - s.con[q].pos) < tol) rep = true;
                        if (!rep && s.t_pen[it] > -tol) {
                            if (nc < W_MAXC) {
                                WContact& c = s.con[nc++];
                                c.pos = s.t_pos[it]; c.normal = s.t_n[it]; c.pen = fmaxf(s.t_pen[it], 0.0f); c.mu = mu;
                                c.la = m.g_link[ga]; c.lb = m.g_link[gb];
                            } else s.err |= ERR_CONTACT_OVERFLOW;
                        }
                    }
                }
            }
            s.n_contacts = nc;
            s.n_cache = ncache;
        }
    }
    WSYNC();
}


// ---------------------------------------------------------------------------------------------------
// Constraint rows: lane per row.
// ---------------------------------------------------------------------------------------------------
G2_HD void w_assemble_constraints(const Go2Model& m, EnvS& s) {
    int nc = s.n_contacts;
    for (int base = 0; base < 4 * nc; base += 32) {
        LANES(l) {
            int r = base + l;
            if (r < 4 * nc) {
                const WContact& c = s.con[r >> 2];
                int k = r & 3;
                v3 d1, d2;
                orthobasis(c.normal, d1, d2);
                float invw = (c.la >= 0 ? m.invweight_link[c.la] : 0.0f) + (c.lb >= 0 ? m.invweight_link[c.lb] : 0.0f);
                v3 tp = c.pos - s.root_com;
                v3 d = (k < 2 ? d1 : d2) * (float)(2 * (k % 2) - 1);
                v3 n = d * c.mu - c.normal;
                float* Jr = s.J[r];
                for (int i = 0; i < NV; ++i) Jr[i] = 0.0f;
                float jqv = 0.0f;
                for (int side = 0; side < 2; ++side) {
                    float sgn = side == 0 ? -1.0f : 1.0f;
                    int lk = side == 0 ? c.la : c.lb;
                    while (lk >= 0) {
                        int d0 = lk == 0 ? 0 : 5 + lk, nd = lk == 0 ? 6 : 1;
                        for (int i = d0 + nd - 1; i >= d0; --i) {
                            v3 vel = s.cdof_v[i] - cross(tp, s.cdof_a[i]);
                            float j = sgn * dot(vel, n);
                            jqv += j * s.qvel[i];
                            Jr[i] += j;
                        }
                        lk = lk == 0 ? -1 : m.parent[lk];
                    }
                }
                float imp, ar;
                imp_aref(m.sol_params, -c.pen, jqv, -c.pen, imp, ar);
                float diag = invw + c.mu * c.mu * invw;
                diag *= 2.0f * c.mu * c.mu * (1.0f - imp) / imp;
                diag = fmaxf(diag, EPSF);
                s.aref[r] = ar;
                s.Dd[r] = 1.0f / diag;
            }
        }
    }
    // joint limits: lanes 6..17 test their hinge, violated ones are appended in dof order
    LANES(l) {
        unsigned char f = 0;
        if (l >= 6 && l < NV) {
            float q = s.qpos[l + 1];
            f = fminf(q - m.limit_lo[l], m.limit_hi[l] - q) < 0.0f ? 1 : 0;
        }
        s.flags[l] = f;
    }
    WSYNC();
    unsigned mask = wmask32(s.flags);
    int nlim = 0;
    {
        unsigned mm = mask;
        while (mm) { mm &= mm - 1; nlim++; }
    }
    LANES(l) {
        if (mask & (1u << l)) {
            int below = 0;
            for (int b = 0; b < l; ++b) below += (mask >> b) & 1u;
            int r = 4 * nc + below;
            float q = s.qpos[l + 1];
            float dmin = q - m.limit_lo[l], dmax = m.limit_hi[l] - q;
            float pd = fminf(dmin, dmax);
            float j = dmin < dmax ? 1.0f : -1.0f;
            float imp, ar;
            imp_aref(m.sol_params, pd, j * s.qvel[l], pd, imp, ar);
            float diag = fmaxf(m.invweight_dof[l] * (1.0f - imp) / imp, EPSF);
            float* Jr = s.J[r];
            for (int i = 0; i < NV; ++i) Jr[i] = 0.0f;
            Jr[l] = j;
            s.aref[r] = ar;
            s.Dd[r] = 1.0f / diag;
        }
        if (l == 0) s.n_rows = 4 * nc + nlim;
    }
    WSYNC();
}

// ---------------------------------------------------------------------------------------------------
// Newton solver.  Rows r = lane, lane+32, lane+64; dofs and Cholesky columns on lanes 0..17;
// the 171 Hessian entries are spread over all lanes.
// ---------------------------------------------------------------------------------------------------
G2_HD float w_Mat(const EnvS& s, int i, int j) { return i >= j ? s.M[tri(i, j)] : s.M[tri(j, i)]; }

// M is tree-sparse: row i only couples to the base, its own leg
G2_HD float w_Mrow_dot(const EnvS& s, int i, const float* x) {
    float a = 0.0f;
    if (i < 6) {
        for (int j = 0; j < NV; ++j) a += w_Mat(s, i, j) * x[j];
    } else {
        for (int j = 0; j < 6; ++j) a += s.M[tri(i, j)] * x[j];
        int leg = (i - 6) & 3;
        for (int lev = 0; lev < 3; ++lev) {
            int j = 6 + leg + 4 * lev;
            a += w_Mat(s, i, j) * x[j];
        }
    }
    return a;
}

G2_HD void w_update_constraint(EnvS& s) {
    int n = s.n_rows;
    LANES(l) {
        float cc = 0.0f;
        for (int r = l; r < n; r += 32) {
            s.pact[r] = s.act[r];
            float ja = s.Jaref[r];
            bool a = ja < 0.0f;
            s.act[r] = a ? 1 : 0;
            s.efc[r] = a ? -ja * s.Dd[r] : 0.0f;
            if (a) cc += 0.5f * ja * ja * s.Dd[r];
        }
        float g = 0.0f;
        if (l < NV) g = 0.5f * (s.Ma[l] - s.force[l]) * (s.qacc[l] - s.acc_smooth[l]);
        s.red[l] = cc;
        s.red2[l] = g;
    }
    WSYNC();
    float ccost = wsum32(s.red), gauss = wsum32(s.red2);
    LANES(l) {
        if (l < NV) {
            float q = 0.0f;
            for (int r = 0; r < n; ++r) q += s.J[r][l] * s.efc[r];
            s.qfc[l] = q;
        }
        if (l == 0) {
            s.prev_cost = s.cost;
            s.gauss = gauss;
            s.cost = gauss + ccost;
        }
    }
    WSYNC();
}

G2_HD void w_tri_ij(int e, int& i, int& j) {
    i = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
    if (tri(i + 1, 0) <= e) i++;
    if (tri(i, 0) > e) i--;
    j = e - tri(i, 0);
}

G2_HD void w_hessian_factor(EnvS& s) {
    int n = s.n_rows;
    LANES(l) {
        for (int e = l; e < NTRI; e += 32) {
            int i, j;
            w_tri_ij(e, i, j);
            float h = s.M[e];
            for (int r = 0; r < n; ++r)
                if (s.act[r]) h += s.J[r][i] * s.J[r][j] * s.Dd[r];
            s.H[e] = h;
        }
    }
    WSYNC();
    for (int i = 0; i < NV; ++i) {
        LANES(l) {
            if (l >= i && l < NV) {
                float dii = s.H[tri(i, i)], c = s.H[tri(l, i)];
                for (int k = 0; k < i; ++k) {
                    float lik = s.H[tri(i, k)];
                    dii -= lik * lik;
                    c -= s.H[tri(l, k)] * lik;
                }
                float lii = sqrtf(fmaxf(dii, EPSF));
                s.vec[l] = l == i ? lii : c / lii;
            }
        }
        WSYNC();
        LANES(l) {
            if (l >= i && l < NV) s.H[tri(l, i)] = s.vec[l];
        }
        WSYNC();
    }
}

G2_HD bool w_hessian_update(EnvS& s) {
    int n = s.n_rows;
    bool ok = true;
    for (int c = 0; c < n; ++c) {
        if (s.act[c] == s.pact[c]) continue;
        float sgn = s.act[c] ? 1.0f : -1.0f;
        float sd = sqrtf(s.Dd[c]);
        LANES(l) { if (l < NV) s.vec[l] = s.J[c][l] * sd; }
        WSYNC();
        for (int k = 0; k < NV && ok; ++k) {
            float vk = s.vec[k];
            if (fabsf(vk) > EPSF) {
                float Lkk = s.H[tri(k, k)];
                float t = Lkk * Lkk + sgn * vk * vk;
                if (t < EPSF) { ok = false; break; }
                float r = sqrtf(t), cc = r / Lkk, ci = 1.0f / cc, sv = vk / Lkk;
                WSYNC();
                LANES(l) {
                    if (l == k) s.H[tri(k, k)] = r;
                    if (l > k && l < NV) {
                        float h = (s.H[tri(l, k)] + sv * s.vec[l] * sgn) * ci;
                        s.H[tri(l, k)] = h;
                        s.vec[l] = s.vec[l] * cc - sv * h;
                    }
                }
                WSYNC();
            }
        }
        if (!ok) break;
    }
    return ok;
}

G2_HD void w_update_gradient(EnvS& s) {
    LANES(l) {
        if (l < NV) {
            float g = s.Ma[l] - s.force[l] - s.qfc[l];
            s.grad[l] = g;
            s.Mgrad[l] = g;
        }
    }
    WSYNC();
    for (int i = 0; i < NV; ++i) {  // L y = g (column oriented)
        float xi = s.Mgrad[i] / s.H[tri(i, i)];
        WSYNC();
        LANES(l) {
            if (l == i) s.Mgrad[i] = xi;
            if (l > i && l < NV) s.Mgrad[l] -= s.H[tri(l, i)] * xi;
        }
        WSYNC();
    }
    for (int i = NV - 1; i >= 0; --i) {  // L^T x = y
        float xi = s.Mgrad[i] / s.H[tri(i, i)];
        WSYNC();
        LANES(l) {
            if (l == i) s.Mgrad[i] = xi;
            if (l < i) s.Mgrad[l] -= s.H[tri(i, l)] * xi;
        }
        WSYNC();
    }
}

struct WLs {
    float alpha, cost, d0, d1;
};
G2_HD WLs w_ls_finish(float alpha, float q0, float q1, float q2) {
    WLs r;
    r.alpha = alpha;
    r.cost = alpha * alpha * q2 + alpha * q1 + q0;
    r.d0 = 2.0f * alpha * q2 + q1;
    r.d1 = 2.0f * q2;
    if (r.d1 <= 0.0f) r.d1 = EPSF;
    return r;
}
G2_HD WLs w_ls_eval(EnvS& s, float alpha, float g0, float g1, float g2, int& ls_it) {
    int n = s.n_rows;
    LANES(l) {
        float q0 = 0.0f, q1 = 0.0f, q2 = 0.0f;
        for (int r = l; r < n; r += 32) {
            float ja = s.Jaref[r], jv = s.jv[r], D = s.Dd[r];
            if (ja + alpha * jv < 0.0f) {
                q0 += D * (0.5f * ja * ja);
                q1 += D * (jv * ja);
                q2 += D * (0.5f * jv * jv);
            }
        }
        s.red[l] = q0; s.red2[l] = q1; s.red3[l] = q2;
    }
    WSYNC();
    float q0 = g0 + wsum32(s.red), q1 = g1 + wsum32(s.red2), q2 = g2 + wsum32(s.red3);
    WSYNC();
    ls_it += 1;
    return w_ls_finish(alpha, q0, q1, q2);
}
G2_HD int w_bracket_update(WLs& pt, const WLs* cand, float& next_alpha) {
    int flag = 0;
    for (int i = 0; i < 3; ++i) {
        if (pt.d0 < 0.0f && cand[i].d0 < 0.0f && pt.d0 < cand[i].d0) { pt = cand[i]; flag = 1; }
        else if (pt.d0 > 0.0f && cand[i].d0 > 0.0f && pt.d0 > cand[i].d0) { pt = cand[i]; flag = 2; }
    }
    next_alpha = pt.alpha;
    if (flag > 0) next_alpha = pt.alpha - pt.d0 / pt.d1;
    return flag;
}


// three step lengths in one pass over the rows
G2_HD void w_ls_eval3(EnvS& s, const float* a, float g0, float g1, float g2, WLs* out, int& ls_it) {
    int n = s.n_rows;
    float q[3][3];
    for (int k = 0; k < 3; ++k) {
        LANES(l) {
            float q0 = 0.0f, q1 = 0.0f, q2 = 0.0f;
            for (int r = l; r < n; r += 32) {
                float ja = s.Jaref[r], jv = s.jv[r], D = s.Dd[r];
                if (ja + a[k] * jv < 0.0f) {
                    q0 += D * (0.5f * ja * ja);
                    q1 += D * (jv * ja);
                    q2 += D * (0.5f * jv * jv);
                }
            }
            s.red[l] = q0; s.red2[l] = q1; s.red3[l] = q2;
        }
        WSYNC();
        q[k][0] = wsum32(s.red); q[k][1] = wsum32(s.red2); q[k][2] = wsum32(s.red3);
        WSYNC();
    }
    for (int k = 0; k < 3; ++k) out[k] = w_ls_finish(a[k], g0 + q[k][0], g1 + q[k][1], g2 + q[k][2]);
    ls_it += 3;
}

// exact line search along s.search; fills s.mv / s.jv; returns the step length (uniform)
G2_HD float w_linesearch(const Go2Model& m, EnvS& s) {
    int n = s.n_rows;
    LANES(l) {
        if (l < NV) s.mv[l] = w_Mrow_dot(s, l, s.search);
        for (int r = l; r < n; r += 32) {
            const float* Jr = s.J[r];
            float a = 0.0f;
            for (int d = 0; d < NV; ++d) a += Jr[d] * s.search[d];
            s.jv[r] = a;
        }
    }
    WSYNC();
    float sn = 0.0f, g1 = 0.0f, g2 = 0.0f;
    for (int d = 0; d < NV; ++d) {
        float sd = s.search[d];
        sn += sd * sd;
        g1 += sd * s.Ma[d] - sd * s.force[d];
        g2 += 0.5f * sd * s.mv[d];
    }
    sn = sqrtf(sn);
    float g0 = s.gauss;
    float gtol = SOLVER_TOL * LS_TOL * sn * (m.meaninertia * (float)NV);
    int ls_it = 0;
    if (sn < EPSF) return 0.0f;
    WLs p0 = w_ls_eval(s, 0.0f, g0, g1, g2, ls_it);
    WLs p1 = w_ls_eval(s, p0.alpha - p0.d0 / p0.d1, g0, g1, g2, ls_it);
    if (p0.cost < p1.cost) p1 = p0;
    if (fabsf(p1.d0) < gtol) return p1.alpha;
    float dir = p1.d0 < 0.0f ? 1.0f : -1.0f;
    bool p2upd = false;
    WLs p2 = p1;
    while (p1.d0 * dir <= -gtol && ls_it < LS_ITERS) {
        p2 = p1;
        p2upd = true;
        p1 = w_ls_eval(s, p1.alpha - p1.d0 / p1.d1, g0, g1, g2, ls_it);
        if (fabsf(p1.d0) < gtol) return p1.alpha;
    }
    if (ls_it >= LS_ITERS) return p1.alpha;
    if (!p2upd) return p1.alpha;
    float a[3] = {p1.alpha - p1.d0 / p1.d1, p1.alpha, (p1.alpha + p2.alpha) * 0.5f};
    while (ls_it < LS_ITERS) {
        WLs cand[3];
        w_ls_eval3(s, a, g0, g1, g2, cand, ls_it);
        bool found = false;
        float ba = 0.0f, bc = 0.0f;
        for (int i = 0; i < 3; ++i) {
            if (fabsf(cand[i].d0) < gtol && (!found || cand[i].cost < bc)) { ba = cand[i].alpha; bc = cand[i].cost; found = true; }
        }
        if (found) return ba;
        float n1, n2;
        int b1 = w_bracket_update(p1, cand, n1);
        int b2 = w_bracket_update(p2, cand, n2);
        if (b1 == 0 && b2 == 0) return a[2];
        a[0] = n1; a[1] = n2; a[2] = (p1.alpha + p2.alpha) * 0.5f;
    }
    if (p1.cost <= p2.cost && p1.cost < p0.cost) return p1.alpha;
    if (p2.cost <= p1.cost && p2.cost < p0.cost) return p2.alpha;
    return 0.0f;
}

G2_HD void w_solve_constraints(const Go2Model& m, EnvS& s) {
    int n = s.n_rows;
    bool ws = n > 0 && s.warm;
    LANES(l) {
        if (l < NV) s.qacc[l] = ws ? s.qacc_ws[l] : s.acc_smooth[l];
    }
    WSYNC();
    LANES(l) {
        if (l < NV) s.Ma[l] = w_Mrow_dot(s, l, s.qacc);
        for (int r = l; r < n; r += 32) {
            const float* Jr = s.J[r];
            float a = -s.aref[r];
            for (int d = 0; d < NV; ++d) a += Jr[d] * s.qacc[d];
            s.Jaref[r] = a;
        }
    }
    WSYNC();
    w_update_constraint(s);
    int n_newton = 0;
    if (n > 0) {
        w_hessian_factor(s);
        w_update_gradient(s);
        LANES(l) { if (l < NV) s.search[l] = -s.Mgrad[l]; }
        WSYNC();
        float tol = m.meaninertia * (float)NV * SOLVER_TOL;
        for (int it = 0; it < SOLVER_ITERS; ++it) {
            n_newton++;
            float alpha = w_linesearch(m, s);
            if (fabsf(alpha) < EPSF) break;
            WSYNC();
            LANES(l) {
                if (l < NV) {
                    s.qacc[l] += s.search[l] * alpha;
                    s.Ma[l] += s.mv[l] * alpha;
                }
                for (int r = l; r < n; r += 32) s.Jaref[r] += s.jv[r] * alpha;
            }
            WSYNC();
            w_update_constraint(s);
            if (!w_hessian_update(s)) w_hessian_factor(s);
            w_update_gradient(s);
            float gn = 0.0f;
            for (int d = 0; d < NV; ++d) gn += s.grad[d] * s.grad[d];
            gn = sqrtf(gn);
            bool improved = gn > tol && (s.prev_cost - s.cost) > tol;
            if (!improved) break;
            WSYNC();
            LANES(l) { if (l < NV) s.search[l] = -s.Mgrad[l]; }
            WSYNC();
        }
    }
    bool bad = false;
    for (int d = 0; d < NV; ++d) bad = bad || !(s.qacc[d] == s.qacc[d]);
    WSYNC();
    LANES(l) {
        if (l < NV) {
            s.acc[l] = s.qacc[l];
            s.qacc_ws[l] = s.qacc[l];
        }
        if (l < NLINK) s.cforce[l] = mk3(0, 0, 0);
        if (l == 0) {
            s.warm = 1;
            s.n_newton = n_newton;
            if (bad) s.err |= ERR_NAN_FORCE;
        }
    }
    WSYNC();
    LANES(l) {
        if (l == 0) {
            for (int ic = 0; ic < s.n_contacts; ++ic) {
                const WContact& c = s.con[ic];
                v3 d1, d2;
                orthobasis(c.normal, d1, d2);
                v3 f = mk3(0, 0, 0);
                for (int k = 0; k < 4; ++k) {
                    v3 d = (k < 2 ? d1 : d2) * (float)(2 * (k % 2) - 1);
                    f = f + (d * c.mu - c.normal) * s.efc[ic * 4 + k];
                }
                if (c.la >= 0) s.cforce[c.la] = s.cforce[c.la] - f;
                if (c.lb >= 0) s.cforce[c.lb] = s.cforce[c.lb] + f;
            }
        }
    }
    WSYNC();
}

G2_HD void w_substep(const Go2Model& m, EnvS& s, float friction, const float* mass_shift, const float* com_shift) {
    w_smooth_dynamics(m, s);
    w_detect_collisions(m, s, friction);
    w_assemble_constraints(m, s);
    w_solve_constraints(m, s);
    w_integrate(m, s);
    w_kinematics(m, mass_shift, com_shift, s);
    w_velocities(m, s);
}

}  // namespace go2
