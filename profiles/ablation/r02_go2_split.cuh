// go2_b200 - the control tick as a pipeline of kernels around a four-lanes-per-env constraint solve (cfg.solver_quad).
//
//   A  (thread per env)  load, FK, env_pre (action -> torques, pushes), smooth dynamics, collision, constraint assembly, q_pack
//   S  (4 lanes per env) Newton solve of substep 1 (go2_qsolve.cuh)
//   B  (thread per env)  integrate, FK, smooth dynamics, collision, constraint assembly, q_pack
//   S                    Newton solve of substep 2
//   C  (thread per env)  integrate, FK, read-back, rewards, termination, reset + DR draws, observations
//
// Same per-env arithmetic as the fused kernel (go2_env.cuh: t_env_step) outside the solve; what crosses a kernel boundary goes
// through HBM: qpos / qvel (their own buffers), the tick's actuation (B_SC_CTRL, 56 words), the solver hand-off (B_SC_LANES,
// ~150 words + 11 per row, written by q_pack) and its result (B_SC_QACC 18 words, contact forces, warm start).  Envs whose
// problem does not fit the quad layout (a contact between two different legs, more than 6 contacts on one leg) are solved
// inside A / B by the thread-per-env solver of go2_solve.cuh.
#pragma once
#include "go2_env.cuh"

namespace go2 {

#define F32(id) ((float*)bf.p[id])
#define I32(id) ((int*)bf.p[id])
#define U8(id) ((uint8_t*)bf.p[id])

G2_HD void sc_store_ctrl(const Buffers& bf, int b, const Phys& p) {
    float* s = F32(B_SC_CTRL) + (size_t)b * SC_CTRL_LEN;
    for (int d = 0; d < NV; ++d) { s[d] = p.ctrl_force[d]; s[NV + d] = p.ctrl_pos[d]; }
    float* t = s + 2 * NV;
    t[0] = (float)p.pd_mode; t[1] = p.pd_kp; t[2] = p.pd_kv;
    t[3] = (float)p.ext_link; t[4] = p.ext_f.x; t[5] = p.ext_f.y; t[6] = p.ext_f.z; t[7] = p.ext_t.x; t[8] = p.ext_t.y; t[9] = p.ext_t.z;
    t[10] = (float)p.ext2_link; t[11] = p.ext2_f.x; t[12] = p.ext2_f.y; t[13] = p.ext2_f.z; t[14] = p.ext2_t.x; t[15] = p.ext2_t.y; t[16] = p.ext2_t.z;
}
G2_HD void sc_load_ctrl(const Buffers& bf, int b, Phys& p) {
    const float* s = F32(B_SC_CTRL) + (size_t)b * SC_CTRL_LEN;
    for (int d = 0; d < NV; ++d) { p.ctrl_force[d] = s[d]; p.ctrl_pos[d] = s[NV + d]; }
    const float* t = s + 2 * NV;
    p.pd_mode = (int)t[0]; p.pd_kp = t[1]; p.pd_kv = t[2];
    p.ext_link = (int)t[3]; p.ext_f = mk3(t[4], t[5], t[6]); p.ext_t = mk3(t[7], t[8], t[9]);
    p.ext2_link = (int)t[10]; p.ext2_f = mk3(t[11], t[12], t[13]); p.ext2_t = mk3(t[14], t[15], t[16]);
}
G2_HD void sc_store_cforce(const Buffers& bf, int b, const Phys& p) {
    for (int l = 0; l < NLINK; ++l) {
        float* o = F32(B_CONTACT_FORCE) + ((size_t)b * NLINK + l) * 3;
        o[0] = p.cforce[l].x; o[1] = p.cforce[l].y; o[2] = p.cforce[l].z;
    }
}

// smooth dynamics -> collision -> assembly -> hand-off to the solve kernel (or the in-kernel solve for envs that do not fit it).
// Every thread of the block calls it (phase barriers; the fallback solver's iteration loop is block-synchronous).
G2_HD void split_front(const Go2Model& m, const EnvConfig& cfg, const Buffers& bf, int b, bool valid, Phys& p, const float* gl) {
    if (valid) smooth_dynamics(m, p);
    G2_PHASE_SYNC();
    if (valid) detect_collisions(m, p, gl[G_FRIC_GROUND], gl[G_FRIC_ROBOT]);
    G2_PHASE_SYNC();
    int state = 0;
    bool fallback = false;
    if (valid) {
        assemble_constraints(m, p, cfg.solver_dense);
        if (p.n_rows > 0) {
            if (p.arrow && q_pack(p, (QLane*)F32(B_SC_LANES) + (size_t)b * 4)) state = 1;
            else fallback = true;
        }
    }
    solve_constraints(m, p, valid && fallback);
    if (valid) {
        int* si = I32(B_SC_INT) + (size_t)b * 4;
        if (state == 0) {
            if (!fallback) {  // no constraint row: the unconstrained acceleration is the answer (go2_solve.cuh, n == 0)
                for (int d = 0; d < NV; ++d) p.acc[d] = p.qacc_ws[d] = p.acc_smooth[d];
                p.warm = 1;
                p.n_newton = 0;
                for (int l = 0; l < NLINK; ++l) p.cforce[l] = mk3(0, 0, 0);
            }
            for (int d = 0; d < NV; ++d) { F32(B_SC_QACC)[b * NV + d] = p.acc[d]; F32(B_QACC_WS)[b * NV + d] = p.qacc_ws[d]; }
            I32(B_WARM)[b] = p.warm;
            sc_store_cforce(bf, b, p);
            si[SCI_NNEWTON] = p.n_newton;
        }
        si[SCI_STATE] = state;
        si[SCI_ERR] = p.err;
        si[SCI_NCON] = p.n_contacts;
        // contact-normal cache for the next detection
        I32(B_NCACHE)[b] = p.n_cache;
        for (int k = 0; k < p.n_cache; ++k) {
            I32(B_CACHE_PAIR)[b * MAX_CPAIRS + k] = p.cache_pair[k];
            float* cn = F32(B_CACHE_N) + (size_t)b * MAX_CPAIRS * 3 + 3 * k;
            cn[0] = p.cache_n[k].x; cn[1] = p.cache_n[k].y; cn[2] = p.cache_n[k].z;
        }
    }
}

// ---- A: tick opening + front half of substep 1
G2_HD void split_a(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, const float* actions, int b, bool valid,
                   Phys& p, EnvLocal& e, EnvStepOut& out) {
    const float* gl = F32(B_GLOBALS) + sp.parity * GLOBALS_LEN;
    if (valid) {
        load_phys(bf, b, p);
        const float* gl0 = U8(B_FK_FRESH)[b] ? gl : F32(B_GLOBALS) + (sp.parity ^ 1) * GLOBALS_LEN;  // see t_env_step
        kinematics(m, gl0, p);
        velocities(m, p);
        env_pre(m, cfg, sp, bf, actions, b, p, e, out);
        sc_store_ctrl(bf, b, p);
        // what phase C reloads instead of keeping it alive across four kernels
        const int A = cfg.num_actions;
        for (int i = 0; i < A; ++i) F32(B_APPLIED)[b * A + i] = e.applied[i];
        for (int j = 0; j < NHINGE; ++j) F32(B_TARGET)[b * NHINGE + j] = e.target[j];
    }
    G2_PHASE_SYNC();
    split_front(m, cfg, bf, b, valid, p, gl);
}

// ---- B: back half of substep 1 (integrate, FK) + front half of substep 2
G2_HD void split_b(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, int b, bool valid, Phys& p) {
    const float* gl = F32(B_GLOBALS) + sp.parity * GLOBALS_LEN;
    if (valid) {
        load_phys(bf, b, p);
        sc_load_ctrl(bf, b, p);
        p.err = I32(B_SC_INT)[b * 4 + SCI_ERR];
        for (int d = 0; d < NV; ++d) p.acc[d] = F32(B_SC_QACC)[b * NV + d];
        integrate(m, p);
        kinematics(m, gl, p);
        velocities(m, p);
        for (int i = 0; i < NQ; ++i) F32(B_QPOS)[b * NQ + i] = p.qpos[i];
        for (int i = 0; i < NV; ++i) F32(B_QVEL)[b * NV + i] = p.qvel[i];
    }
    G2_PHASE_SYNC();
    split_front(m, cfg, bf, b, valid, p, gl);
}

// ---- C: back half of substep 2 + the env logic after the physics
G2_HD void split_c(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, const float* actions, int b, Phys& p,
                   EnvLocal& e, EnvStepOut& out) {
    const float* gl = F32(B_GLOBALS) + sp.parity * GLOBALS_LEN;
    const int A = cfg.num_actions;
    load_phys(bf, b, p);
    sc_load_ctrl(bf, b, p);
    const int* si = I32(B_SC_INT) + (size_t)b * 4;
    p.err = si[SCI_ERR];
    p.n_contacts = si[SCI_NCON];
    p.n_newton = si[SCI_NNEWTON];
    for (int d = 0; d < NV; ++d) p.acc[d] = F32(B_SC_QACC)[b * NV + d];
    integrate(m, p);
    kinematics(m, gl, p);
    velocities(m, p);
    for (int l = 0; l < NLINK; ++l) p.cforce[l] = ld3(F32(B_CONTACT_FORCE) + ((size_t)b * NLINK + l) * 3);
    // env-level locals as env_pre left them
    out.reset = 0; out.timeout = 0; out.tracking_per_sec = 0.0f;
    for (int i = 0; i < A; ++i) {
        e.last_act[i] = F32(B_LAST_ACTIONS)[b * A + i];
        e.act[i] = clampf(actions[b * A + i], -cfg.clip_actions, cfg.clip_actions);
        e.applied[i] = cfg.kind != ENV_BASE ? F32(B_APPLIED)[b * A + i] : 0.0f;
    }
    for (int j = 0; j < NHINGE; ++j) {
        e.dof_pos[j] = F32(B_DOF_POS)[b * NHINGE + j];
        e.dof_vel[j] = F32(B_DOF_VEL)[b * NHINGE + j];
        e.last_dof_vel[j] = F32(B_LAST_DOF_VEL)[b * NHINGE + j];
        e.target[j] = F32(B_TARGET)[b * NHINGE + j];
        e.torque[j] = 0.0f;
    }
    for (int i = 0; i < 3; ++i) e.cmd[i] = F32(B_COMMANDS)[b * 3 + i];
    e.ep_len = I32(B_EP_LEN)[b];
    e.last_x = F32(B_LAST_X)[b];
    e.terrain_row = I32(B_TERRAIN_ROW)[b];
    for (int i = 0; i < 4; ++i) { e.air[i] = 0.0f; e.fc[i] = 0; e.lfc[i] = 0; }
    if (cfg.has_foot_tracking) {
        for (int i = 0; i < 4; ++i) { e.air[i] = F32(B_AIR_TIME)[b * 4 + i]; e.fc[i] = U8(B_FOOT_CONTACT)[b * 4 + i]; }
    }
    env_post(m, cfg, sp, bf, b, p, e, out);
    store_phys(bf, b, p);
}

// ---- S: one lane (device) / one env (host) of the solve.  `b` = env; the caller's lanes are the env's QLane records.
G2_HD void split_s(const Go2Model& m, const Buffers& bf, int b) {
    int* si = I32(B_SC_INT) + (size_t)b * 4;
    if (si[SCI_STATE] != 1) return;
    const QLane* base = (const QLane*)F32(B_SC_LANES) + (size_t)b * 4;
    const QLane* L[QL];
    QFOR(q) L[q] = base + q_leg(q);
    QOut out[QL];
    q_solve(m, L, I32(B_WARM)[b], out);
    QFOR(q) {
        const int leg = q_leg(q);
        for (int a = 0; a < 3; ++a) {
            const int d = 6 + 4 * a + leg, l = 1 + 4 * a + leg;
            F32(B_SC_QACC)[b * NV + d] = out[q].qacc_l[a];
            F32(B_QACC_WS)[b * NV + d] = out[q].qacc_l[a];
            float* o = F32(B_CONTACT_FORCE) + ((size_t)b * NLINK + l) * 3;
            o[0] = out[q].cf_leg[a].x; o[1] = out[q].cf_leg[a].y; o[2] = out[q].cf_leg[a].z;
        }
        if (leg == 0) {
            for (int d = 0; d < 6; ++d) { F32(B_SC_QACC)[b * NV + d] = out[q].qacc_b[d]; F32(B_QACC_WS)[b * NV + d] = out[q].qacc_b[d]; }
            float* o = F32(B_CONTACT_FORCE) + (size_t)b * NLINK * 3;
            o[0] = out[q].cf_base.x; o[1] = out[q].cf_base.y; o[2] = out[q].cf_base.z;
            I32(B_WARM)[b] = 1;
            si[SCI_NNEWTON] = out[q].n_newton;
            if (out[q].bad) si[SCI_ERR] |= ERR_NAN_FORCE;
        }
    }
}

#undef F32
#undef I32
#undef U8

}  // namespace go2
