// go2_b200 - host-side unpacking of the flat model table (go2_b200/model.py:flatten_model) into Go2Model.
#pragma once
#include "go2_types.cuh"
#include <stdlib.h>
#include <string.h>

namespace go2 {

inline bool unpack_model(Go2Model& m, const int32_t* iv, int n_iv, const double* fv, int n_fv) {
    memset(&m, 0, sizeof(m));
    int k = 0, f = 0;
    int nl = iv[k++], ng = iv[k++], np = iv[k++];
    if (nl != NLINK || ng != NGEOM || np > MAX_PAIRS_STATIC) return false;
    int dof_start[NLINK], n_dofs[NLINK];
    for (int l = 0; l < NLINK; ++l) {
        m.parent[l] = iv[k++];
        dof_start[l] = iv[k++];
        n_dofs[l] = iv[k++];
        k++;  // q_start (implied: hinge l <-> dof 5+l, qpos 6+l)
        if (l > 0 && (dof_start[l] != 5 + l || n_dofs[l] != 1)) return false;
        if (m.parent[l] != par_link(l)) return false;  // the kernels unroll the fixed Go2 tree (go2_math.cuh)
    }
    for (int g = 0; g < NGEOM; ++g) {
        int lg = iv[k++];
        m.g_link[g] = lg - 1;  // global link id (0 = ground) -> robot-local (-1 = ground)
        m.g_type[g] = iv[k++];
    }
    m.npairs = np;
    for (int p = 0; p < np; ++p) {
        m.pair_a[p] = (uint8_t)iv[k++];
        m.pair_b[p] = (uint8_t)iv[k++];
    }
    if (k != n_iv) return false;
    // link-pair table + runs for the two-level broad phase (go2_collide.cuh: detect_collisions)
    m.n_lp = 0;
    m.n_run = 0;
    for (int p = 0; p < np; ++p) {
        int la = m.g_link[m.pair_a[p]] + 1, lb = m.g_link[m.pair_b[p]] + 1;
        int id = -1;
        for (int q = 0; q < m.n_lp; ++q)
            if (m.lp_a[q] == la && m.lp_b[q] == lb) id = q;
        if (id < 0) {
            if (m.n_lp >= MAX_LINK_PAIRS) return false;
            id = m.n_lp++;
            m.lp_a[id] = (uint8_t)la;
            m.lp_b[id] = (uint8_t)lb;
        }
        if (m.n_run > 0 && m.run_lp[m.n_run - 1] == id) {
            m.run_end[m.n_run - 1] = (uint16_t)(p + 1);
        } else {
            if (m.n_run >= MAX_PAIR_RUNS) return false;
            m.run_lp[m.n_run] = (uint8_t)id;
            m.run_end[m.n_run] = (uint16_t)(p + 1);
            m.n_run++;
        }
    }
    for (int d = 0; d < NV; ++d) {
        m.limit_lo[d] = -1e30f; m.limit_hi[d] = 1e30f; m.frange[d] = 1e30f; m.armature[d] = 0.0f;
    }
    for (int l = 0; l < NLINK; ++l) {
        for (int j = 0; j < 3; ++j) m.l_pos[l][j] = (float)fv[f++];
        for (int j = 0; j < 4; ++j) m.l_quat[l][j] = (float)fv[f++];
        for (int j = 0; j < 3; ++j) m.inertial_pos[l][j] = (float)fv[f++];
        for (int j = 0; j < 4; ++j) m.inertial_quat[l][j] = (float)fv[f++];
        for (int j = 0; j < 9; ++j) m.inertial_i[l][j] = (float)fv[f++];
        m.mass[l] = (float)fv[f++];
        for (int j = 0; j < 3; ++j) m.axis[l][j] = (float)fv[f++];
        float lo = (float)fv[f++], hi = (float)fv[f++];
        f++;  // force_range lower (= -upper)
        float fr = (float)fv[f++], arm = (float)fv[f++];
        if (l > 0) {
            int d = 5 + l;
            m.limit_lo[d] = lo; m.limit_hi[d] = hi; m.frange[d] = fr; m.armature[d] = arm;
        }
    }
    for (int g = 0; g < NGEOM; ++g) {
        for (int j = 0; j < 3; ++j) m.g_pos[g][j] = (float)fv[f++];
        for (int j = 0; j < 4; ++j) m.g_quat[g][j] = (float)fv[f++];
        for (int j = 0; j < 3; ++j) m.g_size[g][j] = (float)fv[f++];
        f += 3;  // aabb_min = -aabb_max for the three primitives
        for (int j = 0; j < 3; ++j) m.g_half[g][j] = (float)fv[f++];
    }
    for (int j = 0; j < 7; ++j) m.sol_params[j] = (float)fv[f++];
    f++;  // ground link invweight
    for (int l = 0; l < NLINK; ++l) m.invweight_link[l] = (float)fv[f++];
    for (int d = 0; d < NV; ++d) m.invweight_dof[d] = (float)fv[f++];
    m.meaninertia = (float)fv[f++];
    for (int j = 0; j < 3; ++j) m.gravity[j] = (float)fv[f++];
    m.dt = (float)fv[f++];
    return f == n_fv;
}

}  // namespace go2
