// go2_b200 - one fused environment step for one env (host+device; the CUDA kernel calls this per thread).
//
// Covers, in the reference's order (examples/locomotion/final/go2_env_walk.py:985-1109 and
// go2_env_base.py:123-192): action clip + delay ring, PLS / fixed-gain / solver-side PD torque law with
// torque clamp, push forces, `substeps` rigid substeps, state read-back, foot contacts, command
// resampling, termination, every reward term (+episode sums), reset with domain-randomisation draws,
// observation + privileged observation assembly with noise.  Walk/stair: termination -> rewards ->
// reset -> obs.  Base: termination -> reset -> rewards -> obs.
#pragma once
#include "go2_buffers.cuh"
#include "go2_solve.cuh"

namespace go2 {

// -DG2_NO_LIDAR: A/B build without the lidar variant's code paths (profiles/r02_experiments.md)
#ifdef G2_NO_LIDAR
#define G2_LIDAR_ON(cfg) false
#else
#define G2_LIDAR_ON(cfg) ((cfg).lidar_on != 0)
#endif
#define F32(id) ((float*)bf.p[id])
#define I32(id) ((int*)bf.p[id])
#define U8(id) ((uint8_t*)bf.p[id])

#ifdef __CUDA_ARCH__
#define G2_LDCG(p) __ldcg(p)
#else
#define G2_LDCG(p) (*(p))
#endif
// Per-env rows in global memory are touched once per tick: streaming (evict-first) accesses keep them from pushing the thread-local
// working set out of L2 (-DG2_NO_STREAM_HINTS: plain accesses, for A/B)
#if defined(__CUDA_ARCH__) && !defined(G2_NO_STREAM_HINTS)
#define G2_ST(p, v) __stcs((p), (v))
#define G2_LD(p) __ldcs(p)
#else
#define G2_ST(p, v) (*(p) = (v))
#define G2_LD(p) (*(p))
#endif
enum RngSlot { SLOT_ACT_NOISE = 0, SLOT_OBS_NOISE = 8, SLOT_PUSH = 32, SLOT_CMD = 33, SLOT_CMD_RESET = 34, SLOT_RESET = 40, SLOT_LIDAR = 56 };

struct EnvStepOut {
    int reset, timeout;
    float tracking_per_sec;
    float ep_stat[MAX_REW];
};

struct Rng {
    uint32_t env, step, k0, k1;
};
G2_HD u4 rng4(const Rng& r, uint32_t slot) { return philox(r.env, r.step, slot, 0u, r.k0, r.k1); }
G2_HD float urange(float lo, float hi, uint32_t x) { return (hi - lo) * u01(x) + lo; }

// Stair difficulty row of a reset env (go2_env_stair.py:809-854).  The reference splits the n envs reset in a tick into
// int(0.4 n) frontier rows, int(0.3 n) near-frontier rows and the rest easy rows, then shuffles: a reset env is on the frontier
// with probability int(0.4 n) / n etc.  (n = 1 or 2: always an easy row; large n: 40 / 30 / 30).  Drawn per env with exactly
// those probabilities: same expected counts for every n, counts multinomial instead of fixed.  n <= 0: the large-n limit.
G2_HD int stair_pick_row(u4 rr, int mr, int n) {
    float pf = 0.4f, pn = 0.3f;
    if (n > 0) {
        pf = (float)(int)((double)n * 0.40) / (float)n;
        pn = (float)(int)((double)n * 0.30) / (float)n;
    }
    float u = u01(rr.a);
    if (u < pf) return mr;
    if (u < pf + pn) {
        if (mr >= 2) { int lo = mr - 2, hi = mr - 1; return lo + (int)(rr.b % (uint32_t)(hi - lo + 1)); }
        return mr;
    }
    int emax = mr >= 3 ? mr - 3 : 0;
    return (int)(rr.c % (uint32_t)(emax + 1));
}

// quaternion -> extrinsic roll/pitch/yaw in degrees (genesis/utils/geom.py:718-775 with rpy=True)
G2_HD v3 quat_to_rpy_deg(q4 q) {
    float ww = q.w * q.w, wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z;
    float xx = q.x * q.x, xy = q.x * q.y, xz = q.x * q.z, yy = q.y * q.y, yz = q.y * q.z, zz = q.z * q.z;
    float sinp = wy - xz, sinrcosp = wx + yz, sinycosp = wz + xy;
    float cosrcosp = (ww - xx - yy + zz) * 0.5f, cosycosp = (ww + xx - yy - zz) * 0.5f;
    float cosp = sqrtf(cosycosp * cosycosp + sinycosp * sinycosp);
    float r = atan2f(sinrcosp, cosrcosp), pch = atan2f(sinp, cosp), y = atan2f(sinycosp, cosycosp);
    if (cosp < EPSF) {
        r = 0.0f;
        y = atan2f(wz - xy, (ww - xx + yy - zz) * 0.5f);
    }
    const float k = 57.29577951308232f;
    return mk3(r * k, pch * k, y * k);
}

template <class P>
G2_HD v3 link_vel_world(const P& p, int l) { return p.cd_v[l] + cross(p.cd_a[l], p.lpos[l] - p.root_com); }

// `rp` != null: the three command values the reference drew in the replayed tick (B_REPLAY) instead of Philox draws
G2_HD void resample_commands(const EnvConfig& cfg, const StepParams& sp, const Rng& rng, uint32_t slot, int gid, float* cmd,
                             const float* rp) {
    u4 r = rng4(rng, slot);
    if (rp) {
        cmd[0] = rp[0]; cmd[1] = rp[1]; cmd[2] = rp[2];
    } else if (cfg.compound_commands) {
        cmd[0] = urange(sp.cmd_lo[0], sp.cmd_hi[0], r.a);
        cmd[1] = urange(sp.cmd_lo[1], sp.cmd_hi[1], r.b);
        cmd[2] = urange(sp.cmd_lo[2], sp.cmd_hi[2], r.c);
    } else {
        int choice = (int)(r.d % 3u);
        cmd[0] = cmd[1] = cmd[2] = 0.0f;
        uint32_t rr = choice == 0 ? r.a : (choice == 1 ? r.b : r.c);
        cmd[choice] = urange(sp.cmd_lo[choice], sp.cmd_hi[choice], rr);
    }
    if (gid < cfg.n_standing) cmd[0] = cmd[1] = cmd[2] = 0.0f;
}

// Everything the step needs from the env-level buffers, held in registers / local memory.
struct EnvLocal {
    float act[NACT_MAX], last_act[NACT_MAX], applied[NACT_MAX];
    float dof_pos[NHINGE], dof_vel[NHINGE], last_dof_vel[NHINGE], target[NHINGE], torque[NHINGE];
    float cmd[3];
    v3 base_pos, lin_vel, ang_vel, proj_g, euler, vel_world;
    q4 base_quat;
    float air[4];
    int fc[4], lfc[4];
    int ep_len;
    float last_x;     // stair: base x at the previous tick
    int terrain_row;  // stair: difficulty row
    float ground_h;   // heightmap walk variant: ground height under the base (terrain_manager.get_height_at_robot)
};

// Terrain height under world (x, y) - go2_env_stair.py:758-770 (truncating index, clamped to the field)
G2_HD float terrain_height(const Go2Model& m, float x, float y) {
    if (!m.terrain_on) return 0.0f;
    int col = (int)((x - m.terrain_org[0]) / m.terrain_sh), row = (int)((y - m.terrain_org[1]) / m.terrain_sh);
    col = col < 0 ? 0 : (col > m.terrain_r - 1 ? m.terrain_r - 1 : col);
    row = row < 0 ? 0 : (row > m.terrain_c - 1 ? m.terrain_c - 1 : row);
    return m.terrain_hf[(size_t)col * m.terrain_c + row];
}

// Bilinear ground height at world (x, y) - terrain_manager.py:354-392 `_query_height_batch`: numpy arithmetic in float64 on the
// float32 height field, clipped to the field, result rounded to float32.  The model keeps the field in metres (field x vertical
// scale, exact for the shipped vertical_scale = 1.0; the reference scales after interpolating).
G2_HD float ground_height_bilinear(const Go2Model& m, const EnvConfig& cfg, float x, float y) {
    if (!m.terrain_on || cfg.tm_blind) return 0.0f;
    const int R = m.terrain_r, C = m.terrain_c;
    double lx = ((double)x - cfg.tm_ox) / cfg.tm_hs, ly = ((double)y - cfg.tm_oy) / cfg.tm_hs;
    lx = lx < 0.0 ? 0.0 : (lx > (double)(R - 2) ? (double)(R - 2) : lx);
    ly = ly < 0.0 ? 0.0 : (ly > (double)(C - 2) ? (double)(C - 2) : ly);
    const int ix = (int)lx, iy = (int)ly;
    const double fx = lx - ix, fy = ly - iy;
    const int ix1 = ix + 1 < R - 1 ? ix + 1 : R - 1, iy1 = iy + 1 < C - 1 ? iy + 1 : C - 1;
    const float* hf = m.terrain_hf;
    const double h00 = hf[(size_t)ix * C + iy], h10 = hf[(size_t)ix1 * C + iy], h01 = hf[(size_t)ix * C + iy1], h11 = hf[(size_t)ix1 * C + iy1];
    return (float)(h00 * (1 - fx) * (1 - fy) + h10 * fx * (1 - fy) + h01 * (1 - fx) * fy + h11 * fx * fy);
}

template <class P>
G2_HDN float reward_term(int id, const Go2Model& m, const EnvConfig& cfg, const P& p, EnvLocal& e) {
    const float dt = 0.02f;
    switch (id) {
        case R_TRACKING_LIN_VEL: {
            float ex = e.cmd[0] - e.lin_vel.x, ey = e.cmd[1] - e.lin_vel.y;
            return expf(-(ex * ex + ey * ey) / cfg.tracking_sigma);
        }
        case R_TRACKING_ANG_VEL: {
            float ez = e.cmd[2] - e.ang_vel.z;
            return expf(-(ez * ez) / cfg.tracking_sigma);
        }
        case R_LIN_VEL_Z: {  // stair: optional dead zone (go2_env_stair.py:1615-1626)
            if (cfg.lin_vel_z_deadzone > 0.0f) { float x = fmaxf(fabsf(e.lin_vel.z) - cfg.lin_vel_z_deadzone, 0.0f); return x * x; }
            return e.lin_vel.z * e.lin_vel.z;
        }
        case R_BASE_HEIGHT: {  // stair: height above the local terrain (go2_env_stair.py:1634-1648)
            float h = e.base_pos.z;
            if (cfg.has_terrain) h = h - terrain_height(m, e.base_pos.x, e.base_pos.y);
            else if (cfg.tm_on) h = h - e.ground_h;  // go2_env_Omni_kplearn_varied_terrain.py: height above the interpolated ground
            float d = h - cfg.base_height_target;
            return d * d;
        }
        case R_ACTION_RATE: {
            float s = 0.0f;
            for (int i = 0; i < cfg.num_actions; ++i) { float d = e.last_act[i] - e.act[i]; s += d * d; }
            return s;
        }
        case R_SIMILAR_TO_DEFAULT: {
            float s = 0.0f;
            for (int i = 0; i < NHINGE; ++i) s += fabsf(e.dof_pos[i] - cfg.default_dof_pos[i]);
            return s;
        }
        case R_ORIENTATION_PENALTY: return e.proj_g.x * e.proj_g.x + e.proj_g.y * e.proj_g.y;
        case R_DOF_ACC: {
            float s = 0.0f;
            for (int i = 0; i < NHINGE; ++i) { float d = (e.dof_vel[i] - e.last_dof_vel[i]) / dt; s += d * d; }
            return s;
        }
        case R_DOF_VEL: {
            float s = 0.0f;
            for (int i = 0; i < NHINGE; ++i) s += e.dof_vel[i] * e.dof_vel[i];
            return s;
        }
        case R_ANG_VEL_XY: return e.ang_vel.x * e.ang_vel.x + e.ang_vel.y * e.ang_vel.y;
        case R_FEET_AIR_TIME: {
            float rew = 0.0f;
            for (int i = 0; i < 4; ++i) {
                bool first = e.air[i] > 0.0f && e.fc[i];
                e.air[i] += dt;
                e.air[i] *= e.fc[i] ? 0.0f : 1.0f;
                rew += (e.air[i] - cfg.feet_air_time_target) * (first ? 1.0f : 0.0f);
            }
            float mv = sqrtf(e.cmd[0] * e.cmd[0] + e.cmd[1] * e.cmd[1]) > 0.1f ? 1.0f : 0.0f;
            return rew * mv;
        }
        case R_FOOT_SLIP: {
            float s = 0.0f;
            for (int i = 0; i < 4; ++i) {
                v3 v = link_vel_world(p, cfg.foot_link[i]);
                s += (e.fc[i] ? 1.0f : 0.0f) * (v.x * v.x + v.y * v.y);
            }
            return s;
        }
        case R_FOOT_CLEARANCE: {
            float s = 0.0f;
            for (int i = 0; i < 4; ++i) {
                int l = cfg.foot_link[i];
                v3 v = link_vel_world(p, l);
                float fz = p.lpos[l].z;
                if (cfg.has_terrain) fz = fz - terrain_height(m, p.lpos[l].x, p.lpos[l].y);  // go2_env_stair.py:1745-1750
                else if (cfg.tm_on && cfg.tm_variant) fz = fz - e.ground_h;  // go2_env_terrain.py:1355: above the ground under the BASE
                float he = cfg.feet_height_target - fz;
                s += (e.fc[i] ? 0.0f : 1.0f) * (he * he) * sqrtf(v.x * v.x + v.y * v.y);
            }
            float mv = sqrtf(e.cmd[0] * e.cmd[0] + e.cmd[1] * e.cmd[1]) > 0.1f ? 1.0f : 0.0f;
            return s * mv;
        }
        case R_STUMBLE: {  // go2_env_terrain.py:1380-1394: contact force on the non-foot links above a threshold
            float s = 0.0f;
            for (int l = 0; l < NLINK; ++l) {
                bool foot = false;
                for (int i = 0; i < 4; ++i) foot = foot || (cfg.has_foot_tracking && cfg.foot_link[i] == l);
                if (foot) continue;
                v3 f = p.cforce[l];
                float d = sqrtf(f.x * f.x + f.y * f.y + f.z * f.z) - cfg.stumble_threshold;
                s += d > 0.0f ? d : 0.0f;
            }
            return cfg.has_foot_tracking ? s : 0.0f;
        }
        case R_BODY_CONTACT: {  // go2_env_stair_lidar_2.py:1865-1880: how many of the listed links carry a contact force above the threshold
            float s = 0.0f;
            for (int i = 0; i < cfg.n_body_contact; ++i) {
                v3 f = p.cforce[cfg.body_contact_link[i]];
                s += sqrtf(f.x * f.x + f.y * f.y + f.z * f.z) > cfg.body_contact_threshold ? 1.0f : 0.0f;
            }
            return s;
        }
        case R_JOINT_TRACKING: {
            float s = 0.0f;
            for (int i = 0; i < NHINGE; ++i) { float d = e.target[i] - e.dof_pos[i]; s += d * d; }
            return s;
        }
        case R_ENERGY: {
            float s = 0.0f;
            for (int i = 0; i < NHINGE; ++i) s += fabsf(e.torque[i] * e.dof_vel[i]);
            return s;
        }
        case R_TORQUE_LOAD: {
            float s = 0.0f;
            for (int i = 0; i < NHINGE; ++i) s += fabsf(e.torque[i]);
            return cfg.kind == ENV_BASE ? -0.001f * s : s;
        }
        case R_STAND_STILL: {
            float cn = sqrtf(e.cmd[0] * e.cmd[0] + e.cmd[1] * e.cmd[1] + e.cmd[2] * e.cmd[2]);
            float s = 0.0f;
            for (int i = 0; i < NHINGE; ++i) s += fabsf(e.dof_pos[i] - cfg.default_dof_pos[i]);
            return s * (cn < 0.1f ? 1.0f : 0.0f);
        }
        case R_STAND_STILL_VEL: {
            float cn = sqrtf(e.cmd[0] * e.cmd[0] + e.cmd[1] * e.cmd[1] + e.cmd[2] * e.cmd[2]);
            float lv = e.lin_vel.x * e.lin_vel.x + e.lin_vel.y * e.lin_vel.y;
            return (lv + 0.5f * e.ang_vel.z * e.ang_vel.z) * (cn < 0.1f ? 1.0f : 0.0f);
        }
        case R_FEET_STANCE: {
            float cn = sqrtf(e.cmd[0] * e.cmd[0] + e.cmd[1] * e.cmd[1] + e.cmd[2] * e.cmd[2]);
            float s = 0.0f;
            for (int i = 0; i < 4; ++i) s += e.air[i];
            for (int i = 0; i < 4; ++i) s += e.fc[i] ? 0.0f : 1.0f;
            return s * (cn < 0.1f ? 1.0f : 0.0f);
        }
        // ---- base env (crouch / jump), go2_env_base.py:246-390
        case R_JUMP_IMPULSE: return (e.base_pos.z < 0.50f ? 1.0f : 0.0f) * fmaxf(e.lin_vel.z, 0.0f);
        case R_JUMP_APEX: { float d = (e.base_pos.z - cfg.rp[0]) / cfg.rp[1]; return expf(-(d * d)); }
        case R_XY_STABILITY: return -(e.vel_world.x * e.vel_world.x + e.vel_world.y * e.vel_world.y);
        case R_ORIENTATION: return -e.proj_g.z;
        case R_NO_SHAKE: return -(e.ang_vel.x * e.ang_vel.x + e.ang_vel.y * e.ang_vel.y + e.ang_vel.z * e.ang_vel.z) / 1.0f;
        case R_CROUCH: return e.base_pos.z < 0.25f ? 1.0f : 0.0f;
        case R_CROUCH_2: return (e.base_pos.z <= 0.30f && e.base_pos.z >= 0.20f) ? 1.0f : 0.0f;
        case R_GROUND_PENALTY: { float x = clampf((0.15f - e.base_pos.z) / (0.15f - 0.05f), 0.0f, 1.0f); return -(x * x); }
        case R_CROUCH_TARGET: { float d = (e.base_pos.z - 0.15f) / 0.03f; return expf(-(d * d)); }
        case R_NO_FALL: { float dn = fmaxf(-e.lin_vel.z - 0.5f, 0.0f); return -(dn * dn); }
        case R_Y_STABILITY: return -(e.vel_world.y * e.vel_world.y);
        case R_CROUCH_PROGRESS: return fmaxf(0.35f - e.base_pos.z, 0.0f);
        case R_CROUCH_SPEED: return -(e.lin_vel.z * e.lin_vel.z);
        // ---- stair env, go2_env_stair.py:1659-1669, 1763-1772
        case R_ORIENTATION_ROLL_ONLY: return e.proj_g.y * e.proj_g.y;
        case R_FORWARD_PROGRESS: { float dx = e.base_pos.x - e.last_x; e.last_x = e.base_pos.x; return dx; }
        default: return 0.0f;
    }
}

// Reset of one env (walk: go2_env_walk.py:1156-1240; base: go2_env_base.py:201-238).
template <class P>
G2_HDN void reset_env(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, int b, const Rng& rng,
                      P& p, EnvLocal& e, EnvStepOut& out) {
    const float dt = 0.02f;
    int A = cfg.num_actions;
    const float* rp = sp.replay_on ? F32(B_REPLAY) + (size_t)b * REPLAY_LEN : nullptr;  // fixture replay of the draws
    if (cfg.kind != ENV_BASE) {
        // per-episode domain randomisation (per-env draws; the global ones are host-side scalars)
        if (cfg.ctrl == CTRL_MANUAL_PLS || cfg.ctrl == CTRL_MANUAL_FIXED) {
            for (int k = 0; k < 3; ++k) {
                u4 r1 = rng4(rng, SLOT_RESET + k), r2 = rng4(rng, SLOT_RESET + 3 + k);
                uint32_t a1[4] = {r1.a, r1.b, r1.c, r1.d}, a2[4] = {r2.a, r2.b, r2.c, r2.d};
                for (int j = 0; j < 4; ++j) {
                    if (cfg.has_kp_factor) G2_ST(F32(B_KP_FACTORS) + (b * NHINGE + 4 * k + j), rp ? rp[RP_KPF + 4 * k + j] : urange(sp.kpf_lo, sp.kpf_hi, a1[j]));
                    if (cfg.has_kd_factor) G2_ST(F32(B_KD_FACTORS) + (b * NHINGE + 4 * k + j), rp ? rp[RP_KDF + 4 * k + j] : urange(sp.kdf_lo, sp.kdf_hi, a2[j]));
                }
            }
        }
        if (cfg.ctrl == CTRL_MANUAL_FIXED && cfg.has_kp_range) {  // go2_env_walk.py:775-795
            u4 r = rng4(rng, SLOT_RESET + 11);
            float kp = rp ? rp[RP_KP] : urange(sp.kp_lo, sp.kp_hi, r.a), kd = rp ? rp[RP_KD] : urange(sp.kd_lo, sp.kd_hi, r.b);
            for (int j = 0; j < NHINGE; ++j) {
                G2_ST(F32(B_EFF_KP) + (b * NHINGE + j), kp * F32(B_KP_FACTORS)[b * NHINGE + j]);
                G2_ST(F32(B_EFF_KD) + (b * NHINGE + j), kd * F32(B_KD_FACTORS)[b * NHINGE + j]);
            }
        }
        if (cfg.has_motor_strength) {
            for (int k = 0; k < 3; ++k) {
                u4 r = rng4(rng, SLOT_RESET + 6 + k);
                uint32_t a[4] = {r.a, r.b, r.c, r.d};
                for (int j = 0; j < 4; ++j) G2_ST(F32(B_MOTOR_STRENGTH) + (b * NHINGE + 4 * k + j), rp ? rp[RP_MS + 4 * k + j] : urange(sp.ms_lo, sp.ms_hi, a[j]));
            }
        }
        u4 rg = rng4(rng, SLOT_RESET + 9);
        if (cfg.has_gravity_offset) {
            G2_ST(F32(B_GRAV_OFF) + (b * 3 + 0), rp ? rp[RP_GO + 0] : urange(sp.go_lo, sp.go_hi, rg.a));
            G2_ST(F32(B_GRAV_OFF) + (b * 3 + 1), rp ? rp[RP_GO + 1] : urange(sp.go_lo, sp.go_hi, rg.b));
            G2_ST(F32(B_GRAV_OFF) + (b * 3 + 2), rp ? rp[RP_GO + 2] : urange(sp.go_lo, sp.go_hi, rg.c));
        }
        G2_ST(I32(B_DELAY) + (b), rp ? (int)rp[RP_DELAY] : sp.delay_lo + (int)(rg.d % (uint32_t)(sp.delay_hi - sp.delay_lo + 1)));
    }
    // joint + base state
    for (int j = 0; j < NHINGE; ++j) {
        e.dof_pos[j] = cfg.default_dof_pos[j];
        e.dof_vel[j] = 0.0f;
        p.qpos[cfg.motors_dof[j] + 1] = cfg.default_dof_pos[j];
    }
    for (int d = 0; d < NV; ++d) p.qvel[d] = 0.0f;
    e.base_pos = ld3(cfg.base_init_pos);
    e.base_quat = ld4(cfg.base_init_quat);
    u4 rq = rng4(rng, SLOT_RESET + 10);
    if (cfg.kind == ENV_STAIR) {
        if (cfg.has_terrain && cfg.n_terrain_rows > 1 && !sp.lock_terrain_rows) {
            // provisional when the number of resets of this tick is not known yet (n_resets_hint = 0): the post-pass redraws
            // with the true count (post_stair_rows)
            int row = stair_pick_row(rng4(rng, SLOT_RESET + 12), sp.terrain_max_row, sp.n_resets_hint);
            e.terrain_row = rp ? (int)rp[RP_ROW] : row;
        }
        if (cfg.has_terrain) {  // spawn at the row centre, robot standing height on top (go2_env_stair.py:856-871)
            e.base_pos.x = cfg.row_center[e.terrain_row][0];
            e.base_pos.y = cfg.row_center[e.terrain_row][1];
            e.base_pos.z = cfg.row_center[e.terrain_row][2] + cfg.base_init_pos[2];
        }
        // reference quirk (go2_env_stair.py:1541-1546): the drawn z is ADDED to the spawn height
        if (cfg.has_init_z) e.base_pos.z = e.base_pos.z + (rp ? rp[RP_INIT_Z] : urange(cfg.init_z_lo, cfg.init_z_hi, rq.a)) - cfg.base_init_pos[2] + cfg.base_init_pos[2];
    } else if (cfg.tm_on) {
        // terrain_manager._sample_spawn_heightmap (:285-306): uniform xy inside the safe area, z = interpolated ground height + the base's
        // standing height + offset (init_pos_z_range is not used on terrain); the spawn point is kept for the boundary check
        u4 rs = rng4(rng, SLOT_RESET + 13);
        float ux = rp ? rp[RP_SPAWN] : u01(rs.a), uy = rp ? rp[RP_SPAWN + 1] : u01(rs.b);
        float cx = 0.0f, cy = 0.0f;
        if (cfg.tm_grid_rows > 0) {  // _sample_spawn_grid (:308-335): uniform tile, its centre (float32 of a float64 sum) + the jitter
            const int tr = rp ? (int)rp[RP_TILE] : (int)(rs.c % (uint32_t)cfg.tm_grid_rows);
            const int tc = rp ? (int)rp[RP_TILE + 1] : (int)(rs.d % (uint32_t)cfg.tm_grid_cols);
            cx = (float)(cfg.tm_ox + ((double)tr + 0.5) * cfg.tm_tile_x);
            cy = (float)(cfg.tm_oy + ((double)tc + 0.5) * cfg.tm_tile_y);
        }
        e.base_pos.x = cfg.tm_safe_half_x > 0.0f ? add_nf(cx, mul_nf(ux * 2.0f - 1.0f, cfg.tm_safe_half_x)) : cx;  // torch: separate multiply and add
        e.base_pos.y = cfg.tm_safe_half_y > 0.0f ? add_nf(cy, mul_nf(uy * 2.0f - 1.0f, cfg.tm_safe_half_y)) : cy;
        e.base_pos.z = ground_height_bilinear(m, cfg, e.base_pos.x, e.base_pos.y) + cfg.base_init_pos[2] + cfg.tm_spawn_offset;
        G2_ST(F32(B_SPAWN_POS) + (b * 3), e.base_pos.x); G2_ST(F32(B_SPAWN_POS) + (b * 3 + 1), e.base_pos.y); G2_ST(F32(B_SPAWN_POS) + (b * 3 + 2), e.base_pos.z);
    } else if (cfg.has_init_z) e.base_pos.z = rp ? rp[RP_INIT_Z] : urange(cfg.init_z_lo, cfg.init_z_hi, rq.a);
    if (cfg.has_init_euler) {
        float roll = rp ? rp[RP_ROLL] : urange(cfg.init_euler_lo, cfg.init_euler_hi, rq.b);
        float pitch = rp ? rp[RP_PITCH] : urange(cfg.init_euler_lo, cfg.init_euler_hi, rq.c);
        float cr = cosf(roll * 0.5f), sr = sinf(roll * 0.5f), cp = cosf(pitch * 0.5f), sp_ = sinf(pitch * 0.5f);
        e.base_quat = mk4(cr * cp, sr * cp, cr * sp_, -sr * sp_);  // yaw = 0
    }
    p.qpos[0] = e.base_pos.x; p.qpos[1] = e.base_pos.y; p.qpos[2] = e.base_pos.z;
    p.qpos[3] = e.base_quat.w; p.qpos[4] = e.base_quat.x; p.qpos[5] = e.base_quat.y; p.qpos[6] = e.base_quat.z;
    e.lin_vel = mk3(0, 0, 0);
    e.ang_vel = mk3(0, 0, 0);
    e.vel_world = mk3(0, 0, 0);
    for (int i = 0; i < A; ++i) e.last_act[i] = 0.0f;
    for (int j = 0; j < NHINGE; ++j) e.last_dof_vel[j] = 0.0f;
    if (cfg.kind != ENV_BASE) {
        for (int i = 0; i < A; ++i) e.applied[i] = 0.0f;
        for (int i = 0; i < cfg.hist_size * A; ++i) G2_ST(F32(B_HIST) + (b * cfg.hist_size * A + i), 0.0f);
        if (cfg.has_push) {
            G2_ST(F32(B_PUSH_FORCE) + (b * 3 + 0), F32(B_PUSH_FORCE)[b * 3 + 1] = F32(B_PUSH_FORCE)[b * 3 + 2] = 0.0f);
            G2_ST(I32(B_PUSH_REMAIN) + (b), 0);
        }
        for (int i = 0; i < 4; ++i) { e.air[i] = 0.0f; e.fc[i] = 0; e.lfc[i] = 0; }
        e.last_x = e.base_pos.x;  // stair: forward_progress restarts at the spawn (go2_env_stair.py:1568)
        if (G2_LIDAR_ON(cfg))  // go2_env_stair_lidar.py: reset_idx zeroes the scan buffers of the reset envs
            for (int i = 0; i < cfg.n_actor_scan; ++i) G2_ST(F32(B_ACTOR_SCAN) + ((size_t)b * cfg.n_actor_scan + i), 0.0f);
    }
    // episode statistics, then clear the sums
    float ep_s = fmaxf((float)e.ep_len, 1.0f) * dt;
    out.tracking_per_sec = 0.0f;
    for (int k = 0; k < cfg.n_rew; ++k) {
        float s = F32(B_EP_SUMS)[b * cfg.n_rew + k];
        out.ep_stat[k] = cfg.kind == ENV_BASE ? s : s / ep_s;
        if (cfg.rew_id[k] == R_TRACKING_LIN_VEL || cfg.rew_id[k] == R_TRACKING_ANG_VEL) out.tracking_per_sec += s / ep_s;
        G2_ST(F32(B_EP_SUMS) + (b * cfg.n_rew + k), 0.0f);
    }
    e.ep_len = 0;
    resample_commands(cfg, sp, rng, SLOT_CMD_RESET, cfg.env_offset + b, e.cmd, rp ? rp + RP_CMD_RESET : nullptr);
}

// Scalar part of the tick BEFORE the physics (executed by one lane): loads the env-level buffers, turns the
// action into joint torques / PD targets and sets up the push force.  `p` must hold fresh kinematics.
template <class P>
G2_HDN void env_pre(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, const float* actions, int b,
                    P& p, EnvLocal& e, EnvStepOut& out) {
    const int A = cfg.num_actions, H = cfg.hist_size;
    const float* rp = sp.replay_on ? F32(B_REPLAY) + (size_t)b * REPLAY_LEN : nullptr;
    Rng rng;
    rng.env = (uint32_t)(cfg.env_offset + b); rng.step = sp.step_count; rng.k0 = cfg.seed_lo; rng.k1 = cfg.seed_hi;
    out.reset = 0; out.timeout = 0; out.tracking_per_sec = 0.0f;
    p.pd_mode = 0; p.pd_kp = 0.0f; p.pd_kv = 0.0f;
    for (int i = 0; i < A; ++i) e.last_act[i] = F32(B_LAST_ACTIONS)[b * A + i];
    for (int j = 0; j < NHINGE; ++j) {
        e.dof_pos[j] = F32(B_DOF_POS)[b * NHINGE + j];
        e.dof_vel[j] = F32(B_DOF_VEL)[b * NHINGE + j];
        e.last_dof_vel[j] = F32(B_LAST_DOF_VEL)[b * NHINGE + j];
        e.target[j] = 0.0f;
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
    p.ext_link = -1;
    p.ext_f = mk3(0, 0, 0);
    p.ext_t = mk3(0, 0, 0);
    p.ext2_link = -1;
    p.ext2_f = mk3(0, 0, 0);
    p.ext2_t = mk3(0, 0, 0);
    if (sp.ext_on && sp.do_physics) {
        // apply_links_external_force (abd/misc.py:655-715): world-frame force at the link origin -> force + torque about the
        // root COM on that link for both substeps of this tick; the host clears the buffer afterwards (simulator.py:283)
        const float* ex = F32(B_EXT_FORCE) + (size_t)b * 4;
        int l = (int)ex[0];
        if (l >= 0 && l < NLINK) {
            p.ext2_link = l;
            p.ext2_f = mk3(ex[1], ex[2], ex[3]);
            p.ext2_t = cross(p.lpos[l] - p.root_com, p.ext2_f);
        }
    }
    for (int d = 0; d < NV; ++d) { p.ctrl_force[d] = 0.0f; p.ctrl_pos[d] = 0.0f; }

    if (sp.do_physics) {
        // (L2 loads: with go2_step_host the rows are written by the copy engine while this kernel runs - no read-only-cache path)
        for (int i = 0; i < A; ++i) e.act[i] = clampf(G2_LDCG(actions + b * A + i), -cfg.clip_actions, cfg.clip_actions);
        if (cfg.kind != ENV_BASE) {
            // ---- action delay ring (shared write index, per-env delay)
            float* hist = F32(B_HIST) + (size_t)b * H * A;
            for (int i = 0; i < A; ++i) hist[sp.hist_write_idx * A + i] = e.act[i];
            int delay = I32(B_DELAY)[b];
            int rd = ((sp.hist_write_idx - delay) % H + H) % H;
            for (int i = 0; i < A; ++i) e.applied[i] = hist[rd * A + i];
            // ---- position targets (+ action noise)
            float noise[NHINGE];
            for (int j = 0; j < NHINGE; ++j) noise[j] = 0.0f;
            if (sp.action_noise_std > 0.0f) {
                for (int k = 0; k < 3; ++k) {
                    u4 r = rng4(rng, SLOT_ACT_NOISE + k);
                    boxmuller(r.a, r.b, noise[4 * k], noise[4 * k + 1]);
                    boxmuller(r.c, r.d, noise[4 * k + 2], noise[4 * k + 3]);
                }
                if (rp) for (int j = 0; j < NHINGE; ++j) noise[j] = rp[RP_ACT_NOISE + j];
            }
            for (int j = 0; j < NHINGE; ++j)
                e.target[j] = e.applied[j] * cfg.action_scale + cfg.default_dof_pos[j] + noise[j] * sp.action_noise_std;
            if (cfg.ctrl == CTRL_GENESIS_PD) {
                p.pd_mode = 1; p.pd_kp = sp.kp_genesis; p.pd_kv = sp.kd_genesis;
                for (int j = 0; j < NHINGE; ++j) p.ctrl_pos[cfg.motors_dof[j]] = e.target[j];
            } else {
                p.pd_mode = 0;
                for (int j = 0; j < NHINGE; ++j) {
                    float ekp, ekd;
                    if (cfg.ctrl == CTRL_MANUAL_PLS) {
                        float kp = clampf(cfg.pls_kp_default + e.applied[cfg.num_pos_actions + j / 3] * cfg.pls_kp_scale, cfg.pls_kp_lo, cfg.pls_kp_hi);
                        float kd = 0.2f * sqrtf(kp);
                        ekp = kp * F32(B_KP_FACTORS)[b * NHINGE + j] * F32(B_MOTOR_STRENGTH)[b * NHINGE + j];
                        ekd = kd * F32(B_KD_FACTORS)[b * NHINGE + j];
                        G2_ST(F32(B_EFF_KP) + (b * NHINGE + j), ekp);
                        G2_ST(F32(B_EFF_KD) + (b * NHINGE + j), ekd);
                    } else {
                        ekp = F32(B_EFF_KP)[b * NHINGE + j];
                        ekd = F32(B_EFF_KD)[b * NHINGE + j];
                    }
                    float tq = ekp * (e.target[j] - e.dof_pos[j]) - ekd * e.dof_vel[j];
                    tq = clampf(tq, -cfg.torque_limit[j], cfg.torque_limit[j]);
                    p.ctrl_force[cfg.motors_dof[j]] = tq;
                }
            }
            // ---- pushes (go2_env_walk.py:872-906): force at the link origin, world frame
            v3 cur = mk3(0, 0, 0);
            if (cfg.has_push && sp.push_enable) {
                if (sp.push_resample) {
                    u4 r = rng4(rng, SLOT_PUSH);
                    G2_ST(F32(B_PUSH_FORCE) + (b * 3 + 0), rp ? rp[RP_PUSH] : urange(sp.push_lo, sp.push_hi, r.a));
                    G2_ST(F32(B_PUSH_FORCE) + (b * 3 + 1), rp ? rp[RP_PUSH + 1] : urange(sp.push_lo, sp.push_hi, r.b));
                    G2_ST(F32(B_PUSH_FORCE) + (b * 3 + 2), 0.0f);
                    G2_ST(I32(B_PUSH_REMAIN) + (b), rp ? (int)rp[RP_PUSH + 2] : cfg.push_dur_lo + (int)(r.c % (uint32_t)(cfg.push_dur_hi - cfg.push_dur_lo + 1)));
                }
                int rem = I32(B_PUSH_REMAIN)[b];
                float on = rem > 0 ? 1.0f : 0.0f;
                cur = mk3(F32(B_PUSH_FORCE)[b * 3] * on, F32(B_PUSH_FORCE)[b * 3 + 1] * on, F32(B_PUSH_FORCE)[b * 3 + 2] * on);
                G2_ST(I32(B_PUSH_REMAIN) + (b), rem - 1 > 0 ? rem - 1 : 0);
                p.ext_link = cfg.push_link;
                p.ext_f = cur;
                p.ext_t = cross(p.lpos[cfg.push_link] - p.root_com, cur);
            }
            if (cfg.has_push) { G2_ST(F32(B_CUR_PUSH) + (b * 3), cur.x); G2_ST(F32(B_CUR_PUSH) + (b * 3 + 1), cur.y); G2_ST(F32(B_CUR_PUSH) + (b * 3 + 2), cur.z); }
        } else {
            // base env: 1-step latency through last_actions, PD evaluated inside every substep
            p.pd_mode = 1; p.pd_kp = sp.kp_genesis; p.pd_kv = sp.kd_genesis;
            for (int j = 0; j < NHINGE; ++j) {
                float ex = cfg.simulate_action_latency ? e.last_act[j] : e.act[j];
                e.target[j] = ex * cfg.action_scale + cfg.default_dof_pos[j];
                p.ctrl_pos[cfg.motors_dof[j]] = e.target[j];
            }
        }
    }
}

// Scalar part of the tick AFTER the physics: read-back, foot contacts, command resampling, termination,
// rewards, reset (+ DR draws), observations, write-back of the env-level buffers.
// one simulated-lidar sample: terrain height under the body-frame point (lx, ly) of a base at `pos` with yaw (cy, sy), relative to
// the base (go2_env_stair_lidar.py:895-937 _compute_lidar_scan; zeros without terrain)
G2_HD float lidar_point(const Go2Model& m, const EnvConfig& cfg, v3 pos, float cy, float sy, float lx, float ly) {
    if (!cfg.has_terrain) return 0.0f;
    return terrain_height(m, pos.x + cy * lx - sy * ly, pos.y + sy * lx + cy * ly) - pos.z;
}
G2_HD void yaw_cs(q4 q, float& cy, float& sy) {
    float yaw = atan2f(2.0f * (q.w * q.z + q.x * q.y), 1.0f - 2.0f * (q.y * q.y + q.z * q.z));
    cy = cosf(yaw); sy = sinf(yaw);
}
// Actor lidar scan of the tick with the sim-to-real noise model, in the reference's order (go2_env_stair_lidar.py:939-1002):
// per-cell Gaussian noise, per-cell dropout (-> 0 = base height), per-env vertical offset, per-env latency (the previous tick's scan
// is replayed), per-env blackout (-> 0), clip, scale.  All probabilities / stds scale with the DR level (lidar_noise_level).
// B_ACTOR_SCAN holds the previous scan on entry and the new one on exit.
G2_HDN void lidar_update(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, int b, const Rng& rng,
                         const float* rp, v3 pos, q4 q) {
    const int n = cfg.n_actor_scan;
    if (n <= 0) return;
    float* buf = F32(B_ACTOR_SCAN) + (size_t)b * n;
    const float lvl = sp.lidar_noise_level;
    float cy, sy;
    yaw_cs(q, cy, sy);
    u4 rv = rng4(rng, SLOT_LIDAR + 16);
    float zv, zv2;
    boxmuller(rv.a, rv.b, zv, zv2);
    const float v = rp ? rp[RP_LIDAR_V] : zv;
    const float ulat = rp ? rp[RP_LIDAR_V + 1] : u01(rv.c), ublk = rp ? rp[RP_LIDAR_V + 2] : u01(rv.d);
    const bool on = lvl > 0.0f;
    const bool lat = on && cfg.lidar_latency > 0.0f && ulat < cfg.lidar_latency * lvl;
    const bool blk = on && cfg.lidar_blackout > 0.0f && ublk < cfg.lidar_blackout * lvl;
    for (int i0 = 0; i0 < n; i0 += 4) {
        u4 rn = rng4(rng, SLOT_LIDAR + (i0 >> 2)), ru = rng4(rng, SLOT_LIDAR + 8 + (i0 >> 2));
        float z[4];
        boxmuller(rn.a, rn.b, z[0], z[1]);
        boxmuller(rn.c, rn.d, z[2], z[3]);
        const float u[4] = {u01(ru.a), u01(ru.b), u01(ru.c), u01(ru.d)};
        for (int j = 0; j < 4 && i0 + j < n; ++j) {
            const int i = i0 + j;
            float raw = lidar_point(m, cfg, pos, cy, sy, cfg.actor_lx[i], cfg.actor_ly[i]);
            if (on && cfg.lidar_noise_std > 0.0f) raw = raw + ((rp ? rp[RP_LIDAR_N + i] : z[j]) * cfg.lidar_noise_std) * lvl;
            if (on && cfg.lidar_dropout > 0.0f && (rp ? rp[RP_LIDAR_U + i] : u[j]) < cfg.lidar_dropout * lvl) raw = 0.0f;
            if (on && cfg.lidar_voffset_std > 0.0f) raw = raw + (v * cfg.lidar_voffset_std) * lvl;
            if (lat) raw = buf[i];
            if (blk) raw = 0.0f;
            buf[i] = clampf(raw, -cfg.lidar_clip, cfg.lidar_clip) * cfg.lidar_scale;
        }
    }
}

// ---- 4-value stair info (go2_env_stair_lidar_2.py, cfg.lidar_on == 2) -------------------------------------------------------------
// Edge detector (:848-906 _compute_edge_detection): terrain heights at si_n_samples points on the forward axis of a base at `pos`
// with yaw of `q`; the first consecutive pair whose heights differ by more than the threshold gives the distance (midpoint of the
// pair) and the direction (sign of the change); (max range, 0) if there is none or no terrain.
G2_HD void stair_edge(const Go2Model& m, const EnvConfig& cfg, v3 pos, q4 q, float& dist, float& dir) {
    dist = cfg.si_max_range;
    dir = 0.0f;
    if (!cfg.has_terrain) return;
    float cy, sy;
    yaw_cs(q, cy, sy);
    float hprev = 0.0f;
    bool found = false;
    for (int i = 0; i < cfg.si_n_samples; ++i) {
        const float lx = cfg.si_edge_x[i];
        const float h = terrain_height(m, pos.x + cy * lx, pos.y + sy * lx);
        if (i > 0 && !found) {
            const float dh = h - hprev;
            if (fabsf(dh) > cfg.si_threshold) {
                found = true;
                dist = cfg.si_edge_mid[i - 1];
                dir = dh > 0.0f ? 1.0f : (dh < 0.0f ? -1.0f : 0.0f);
            }
        }
        hprev = h;
    }
}
// Noisy actor copy of the tick (:908-975 _update_stair_info), in the reference's order: Gaussian noise on the edge distance (then
// clamp to [0, max range]), on the riser height, on the tread depth; direction flip; latency (edge distance and direction of the
// previous tick's info); blackout (edge -> max range, direction -> 0); scale.  B_ACTOR_SCAN holds the previous info on entry.
G2_HDN void stair_info_update(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, int b, const Rng& rng,
                              const float* rp, v3 pos, q4 q, int row) {
    float* buf = F32(B_ACTOR_SCAN) + (size_t)b * 4;
    const float lvl = sp.lidar_noise_level;
    const bool on = lvl > 0.0f;
    float edge, dir;
    stair_edge(m, cfg, pos, q, edge, dir);
    float h = cfg.si_row_height[row], d = cfg.si_row_depth[row];
    u4 rn = rng4(rng, SLOT_LIDAR), ru = rng4(rng, SLOT_LIDAR + 8);
    float z[4];
    boxmuller(rn.a, rn.b, z[0], z[1]);
    boxmuller(rn.c, rn.d, z[2], z[3]);
    if (rp) { z[0] = rp[RP_LIDAR_N]; z[1] = rp[RP_LIDAR_N + 1]; z[2] = rp[RP_LIDAR_N + 2]; }
    const float uf = rp ? rp[RP_LIDAR_U] : u01(ru.a), ul = rp ? rp[RP_LIDAR_U + 1] : u01(ru.b), ub = rp ? rp[RP_LIDAR_U + 2] : u01(ru.c);
    if (on && cfg.si_edge_noise_std > 0.0f) edge = clampf(edge + (z[0] * cfg.si_edge_noise_std) * lvl, 0.0f, cfg.si_max_range);
    if (on && cfg.si_height_noise_std > 0.0f) h = h + (z[1] * cfg.si_height_noise_std) * lvl;
    if (on && cfg.si_depth_noise_std > 0.0f) d = d + (z[2] * cfg.si_depth_noise_std) * lvl;
    if (on && cfg.si_flip_prob > 0.0f && uf < cfg.si_flip_prob * lvl) dir = -dir;
    if (on && cfg.si_latency_prob > 0.0f && ul < cfg.si_latency_prob * lvl) {
        edge = buf[0] / cfg.si_scale[0];
        dir = buf[3] / cfg.si_scale[3];
    }
    if (on && cfg.si_blackout_prob > 0.0f && ub < cfg.si_blackout_prob * lvl) { edge = cfg.si_max_range; dir = 0.0f; }
    buf[0] = edge * cfg.si_scale[0];
    buf[1] = h * cfg.si_scale[1];
    buf[2] = d * cfg.si_scale[2];
    buf[3] = dir * cfg.si_scale[3];
}

// tail of the stair env's privileged obs (go2_env_stair.py:1466-1484): terrain difficulty of the env's row, then the body-frame
// height scan around the base - or, in the lidar variant, the clean critic lidar scan of the tick's pre-reset pose (zeros for an
// env reset in this tick, go2_env_stair_lidar.py:1004-1015,1689-1692); `k` = first index of the tail (num_obs + 55)
G2_HD void stair_priv_tail(const Go2Model& m, const EnvConfig& cfg, float* pr, int k, int terrain_row, v3 base_pos, q4 q, bool was_reset) {
    if (k < cfg.num_priv) {
        int den = cfg.n_terrain_rows - 1 > 1 ? cfg.n_terrain_rows - 1 : 1;
        G2_ST(pr + (k++), cfg.has_terrain ? (float)terrain_row / (float)den : 0.0f);
    }
    float cy, sy;
    yaw_cs(q, cy, sy);
    if (G2_LIDAR_ON(cfg) && cfg.lidar_on == 2) {  // clean stair info of the tick's pre-reset pose (go2_env_stair_lidar_2.py:930-934,1621-1624)
        if (k + 4 <= cfg.num_priv) {
            float edge = 0.0f, dir = 0.0f;
            if (!was_reset) stair_edge(m, cfg, base_pos, q, edge, dir);
            G2_ST(pr + (k++), was_reset ? 0.0f : edge * cfg.si_scale[0]);
            G2_ST(pr + (k++), was_reset ? 0.0f : cfg.si_row_height[terrain_row] * cfg.si_scale[1]);
            G2_ST(pr + (k++), was_reset ? 0.0f : cfg.si_row_depth[terrain_row] * cfg.si_scale[2]);
            G2_ST(pr + (k++), was_reset ? 0.0f : dir * cfg.si_scale[3]);
        }
        return;
    }
    if (G2_LIDAR_ON(cfg)) {
        if (cfg.n_critic_scan > 0 && k + cfg.n_critic_scan <= cfg.num_priv)
            for (int i = 0; i < cfg.n_critic_scan; ++i)
                G2_ST(pr + (k++), was_reset ? 0.0f
                                    : clampf(lidar_point(m, cfg, base_pos, cy, sy, cfg.critic_lx[i], cfg.critic_ly[i]), -cfg.lidar_clip, cfg.lidar_clip) * cfg.lidar_scale);
        return;
    }
    int ns = cfg.scan_nx * cfg.scan_ny;
    if (ns > 0 && k + ns <= cfg.num_priv) {
        for (int ix = 0; ix < cfg.scan_nx; ++ix) {
            float lx = cfg.scan_nx > 1 ? cfg.scan_x0 + (cfg.scan_x1 - cfg.scan_x0) * (float)ix / (float)(cfg.scan_nx - 1) : cfg.scan_x0;
            for (int iy = 0; iy < cfg.scan_ny; ++iy) {
                float ly = cfg.scan_ny > 1 ? cfg.scan_y0 + (cfg.scan_y1 - cfg.scan_y0) * (float)iy / (float)(cfg.scan_ny - 1) : cfg.scan_y0;
                float wx = base_pos.x + cy * lx - sy * ly, wy = base_pos.y + sy * lx + cy * ly;
                G2_ST(pr + (k++), cfg.has_terrain ? terrain_height(m, wx, wy) - base_pos.z : 0.0f);
            }
        }
    }
}

template <class P>
G2_HDN void env_post(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, int b, P& p, EnvLocal& e,
                     EnvStepOut& out) {
    const int A = cfg.num_actions;
    const float* rp = sp.replay_on ? F32(B_REPLAY) + (size_t)b * REPLAY_LEN : nullptr;
    const float* gl = F32(B_GLOBALS) + sp.parity * GLOBALS_LEN;
    Rng rng;
    rng.env = (uint32_t)(cfg.env_offset + b); rng.step = sp.step_count; rng.k0 = cfg.seed_lo; rng.k1 = cfg.seed_hi;
    if (sp.do_physics) {
        e.ep_len += 1;
        // ---- read-back
        e.base_pos = p.lpos[0];
        e.base_quat = p.lquat[0];
        e.euler = quat_to_rpy_deg(qnormalize(qmul(e.base_quat, ld4(cfg.inv_base_init_quat))));
        q4 inv = qconj(e.base_quat);
        e.vel_world = link_vel_world(p, 0);
        e.lin_vel = qrot(e.vel_world, inv);
        e.ang_vel = qrot(p.cd_a[0], inv);
        e.proj_g = qrot(mk3(0.0f, 0.0f, -1.0f), inv);
        for (int j = 0; j < NHINGE; ++j) {
            e.dof_pos[j] = p.qpos[cfg.motors_dof[j] + 1];
            e.dof_vel[j] = p.qvel[cfg.motors_dof[j]];
            e.torque[j] = applied_force(m, p, cfg.motors_dof[j]);
        }
        if (cfg.has_foot_tracking) {
            for (int i = 0; i < 4; ++i) {
                e.lfc[i] = e.fc[i];
                e.fc[i] = fabsf(p.cforce[cfg.foot_link[i]].z) > cfg.foot_contact_threshold ? 1 : 0;
            }
        }
        if (G2_LIDAR_ON(cfg)) {
            if (cfg.lidar_on == 2) stair_info_update(m, cfg, sp, bf, b, rng, rp, e.base_pos, e.base_quat, e.terrain_row);
            else lidar_update(m, cfg, sp, bf, b, rng, rp, e.base_pos, e.base_quat);
        }
        if (cfg.tm_on) {  // go2_env_Omni_kplearn_varied_terrain.py: _ground_height[:] = terrain.get_height_at_robot(base_pos[:, :2])
            e.ground_h = ground_height_bilinear(m, cfg, e.base_pos.x, e.base_pos.y);
            G2_ST(F32(B_GROUND_H) + (b), e.ground_h);
        }
        if (e.ep_len % cfg.resample_steps == 0) resample_commands(cfg, sp, rng, SLOT_CMD, cfg.env_offset + b, e.cmd, rp ? rp + RP_CMD : nullptr);
    } else {
        for (int i = 0; i < A; ++i) { e.act[i] = 0.0f; e.applied[i] = F32(B_APPLIED)[b * A + i]; }
        e.base_pos = ld3(F32(B_BASE_POS) + b * 3);
        e.base_quat = ld4(F32(B_BASE_QUAT) + b * 4);
        e.lin_vel = ld3(F32(B_BASE_LIN_VEL) + b * 3);
        e.ang_vel = ld3(F32(B_BASE_ANG_VEL) + b * 3);
        e.proj_g = ld3(F32(B_PROJ_G) + b * 3);
        e.euler = ld3(F32(B_BASE_EULER) + b * 3);
        e.vel_world = mk3(0, 0, 0);
    }

    const v3 lid_pos = e.base_pos;  // pose the lidar scans of this tick were taken from (before a reset moves the base)
    const q4 lid_q = e.base_quat;
    // ---- termination
    bool tout = e.ep_len > cfg.max_episode_length;
    bool rst = tout;
    if (sp.do_physics) {
        rst = rst || fabsf(e.euler.y) > cfg.term_pitch || fabsf(e.euler.x) > cfg.term_roll;
        rst = rst || fabsf(e.lin_vel.z) > cfg.term_zvel || fabsf(e.lin_vel.y) > cfg.term_yvel;
        if (cfg.tm_on) {  // terrain_manager.check_boundary: wandered further than max_wander from the spawn point
            float dx = e.base_pos.x - F32(B_SPAWN_POS)[b * 3], dy = e.base_pos.y - F32(B_SPAWN_POS)[b * 3 + 1];
            rst = rst || sqrtf(dx * dx + dy * dy) > cfg.tm_max_wander;
        }
    }
    if (sp.force_reset_all) rst = true;
    // an env whose physics went non-finite is force-reset (stricter than the reference, which raises)
    if (p.err & (ERR_NAN_ACC | ERR_NAN_FORCE)) rst = true;

    float rew = 0.0f;
    if (cfg.kind == ENV_BASE) {
        if (rst) reset_env(m, cfg, sp, bf, b, rng, p, e, out);
        if (sp.do_physics) {
            if (rst) {  // live getters see the post-reset state
                for (int j = 0; j < NHINGE; ++j) e.torque[j] = applied_force(m, p, cfg.motors_dof[j]);
            }
            for (int k = 0; k < cfg.n_rew; ++k) {
                float r = reward_term(cfg.rew_id[k], m, cfg, p, e) * cfg.rew_scale[k];
                rew += r;
                F32(B_EP_SUMS)[b * cfg.n_rew + k] += r;
                if (cfg.export_rew_terms) G2_ST(F32(B_REW_TERMS) + (b * cfg.n_rew + k), r);
            }
        }
    } else {
        if (sp.do_physics) {
            for (int k = 0; k < cfg.n_rew; ++k) {
                float r = reward_term(cfg.rew_id[k], m, cfg, p, e) * cfg.rew_scale[k];
                rew += r;
                F32(B_EP_SUMS)[b * cfg.n_rew + k] += r;
                if (cfg.export_rew_terms) G2_ST(F32(B_REW_TERMS) + (b * cfg.n_rew + k), r);
            }
        }
        if (rst) reset_env(m, cfg, sp, bf, b, rng, p, e, out);
    }
    out.reset = rst ? 1 : 0;
    out.timeout = (rst && tout) ? 1 : 0;

    // ---- observations
    if (sp.do_physics) {
        float* obs = F32(B_OBS) + (size_t)b * cfg.num_obs;
        float o[NOBS_MAX];
        o[0] = e.ang_vel.x * cfg.obs_scale_ang_vel; o[1] = e.ang_vel.y * cfg.obs_scale_ang_vel; o[2] = e.ang_vel.z * cfg.obs_scale_ang_vel;
        float go[3] = {0.0f, 0.0f, 0.0f};
        if (cfg.kind != ENV_BASE) { go[0] = F32(B_GRAV_OFF)[b * 3]; go[1] = F32(B_GRAV_OFF)[b * 3 + 1]; go[2] = F32(B_GRAV_OFF)[b * 3 + 2]; }
        o[3] = e.proj_g.x + go[0]; o[4] = e.proj_g.y + go[1]; o[5] = e.proj_g.z + go[2];
        o[6] = e.cmd[0] * cfg.obs_scale_lin_vel; o[7] = e.cmd[1] * cfg.obs_scale_lin_vel; o[8] = e.cmd[2] * cfg.obs_scale_ang_vel;
        for (int j = 0; j < NHINGE; ++j) {
            o[9 + j] = (e.dof_pos[j] - cfg.default_dof_pos[j]) * cfg.obs_scale_dof_pos;
            o[21 + j] = e.dof_vel[j] * cfg.obs_scale_dof_vel;
        }
        for (int i = 0; i < A; ++i) o[33 + i] = cfg.kind == ENV_BASE ? e.act[i] : e.applied[i];
        if (sp.obs_noise_on) {
            for (int k = 0; k < 9; ++k) {  // entries 0..35 can carry noise (33.. are always zero-std)
                u4 r = rng4(rng, SLOT_OBS_NOISE + k);
                float z[4];
                boxmuller(r.a, r.b, z[0], z[1]);
                boxmuller(r.c, r.d, z[2], z[3]);
                for (int j = 0; j < 4; ++j) {
                    int idx = 4 * k + j;
                    if (idx < cfg.num_obs && idx < NOBS_MAX) o[idx] += (rp ? rp[RP_OBS_NOISE + idx] : z[j]) * sp.obs_noise[idx];
                }
            }
        }
        const int n_scan = G2_LIDAR_ON(cfg) ? cfg.n_actor_scan : 0, n_prop = cfg.num_obs - n_scan;
        for (int i = 0; i < n_prop; ++i) G2_ST(obs + i, o[i]);
        for (int i = 0; i < n_scan; ++i) G2_ST(obs + n_prop + i, F32(B_ACTOR_SCAN)[(size_t)b * n_scan + i]);  // go2_env_stair_lidar.py:1626-1627
        if (cfg.num_priv > 0) {  // go2_env_walk.py:1115-1141
            float* pr = F32(B_PRIV) + (size_t)b * cfg.num_priv;
            int k = 0;
            for (int i = 0; i < cfg.num_obs; ++i) G2_ST(pr + (k++), obs[i]);
            G2_ST(pr + (k++), e.lin_vel.x * cfg.obs_scale_lin_vel); G2_ST(pr + (k++), e.lin_vel.y * cfg.obs_scale_lin_vel); G2_ST(pr + (k++), e.lin_vel.z * cfg.obs_scale_lin_vel);
            G2_ST(pr + (k++), gl[G_FRIC_GROUND]);  // patched by the post-pass kernel when this tick's resets change the global DR
            for (int j = 0; j < NHINGE; ++j) G2_ST(pr + (k++), F32(B_KP_FACTORS)[b * NHINGE + j]);
            for (int j = 0; j < NHINGE; ++j) G2_ST(pr + (k++), F32(B_KD_FACTORS)[b * NHINGE + j]);
            for (int j = 0; j < NHINGE; ++j) G2_ST(pr + (k++), F32(B_MOTOR_STRENGTH)[b * NHINGE + j]);
            G2_ST(pr + (k++), gl[G_MASS_SHIFT]);
            G2_ST(pr + (k++), gl[G_COM_SHIFT]); G2_ST(pr + (k++), gl[G_COM_SHIFT + 1]); G2_ST(pr + (k++), gl[G_COM_SHIFT + 2]);
            // leg mass shifts in cfg order FR, FL, RR, RL (robot links 2, 1, 4, 3)
            for (int i = 0; i < 4; ++i) G2_ST(pr + (k++), cfg.has_leg_mass_dr ? gl[G_MASS_SHIFT + cfg.hip_link[i]] : 0.0f);
            G2_ST(pr + (k++), go[0]); G2_ST(pr + (k++), go[1]); G2_ST(pr + (k++), go[2]);
            G2_ST(pr + (k++), F32(B_CUR_PUSH)[b * 3]); G2_ST(pr + (k++), F32(B_CUR_PUSH)[b * 3 + 1]); G2_ST(pr + (k++), F32(B_CUR_PUSH)[b * 3 + 2]);
            G2_ST(pr + (k++), cfg.max_delay > 0 ? (float)I32(B_DELAY)[b] / (float)cfg.max_delay : 0.0f);
            if (cfg.tm_variant && k < cfg.num_priv) G2_ST(pr + (k++), cfg.tm_on ? e.ground_h : 0.0f);  // go2_env_terrain.py:1139-1141
            if (cfg.kind == ENV_STAIR)
                stair_priv_tail(m, cfg, pr, k, e.terrain_row, G2_LIDAR_ON(cfg) ? lid_pos : e.base_pos, G2_LIDAR_ON(cfg) ? lid_q : e.base_quat, rst);
        }
        G2_ST(F32(B_REW) + (b), rew);
        for (int i = 0; i < A; ++i) e.last_act[i] = e.act[i];
        for (int j = 0; j < NHINGE; ++j) e.last_dof_vel[j] = e.dof_vel[j];
    }

    // ---- write back
    G2_ST(U8(B_RESET) + (b), rst ? 1 : 0);
    U8(B_FK_FRESH)[b] = rst ? 1 : 0;
    G2_ST(F32(B_TIMEOUT) + (b), tout ? 1.0f : 0.0f);
    G2_ST(I32(B_EP_LEN) + (b), e.ep_len);
    G2_ST(F32(B_LAST_X) + (b), e.last_x);
    G2_ST(I32(B_TERRAIN_ROW) + (b), e.terrain_row);
    I32(B_ERRNO)[b] |= p.err;
    G2_ST(I32(B_NCONTACT) + (b), p.n_contacts);
    G2_ST(I32(B_NNEWTON) + (b), p.n_newton);
    for (int i = 0; i < 3; ++i) G2_ST(F32(B_COMMANDS) + (b * 3 + i), e.cmd[i]);
    for (int i = 0; i < A; ++i) {
        G2_ST(F32(B_LAST_ACTIONS) + (b * A + i), e.last_act[i]);
        if (cfg.kind != ENV_BASE) G2_ST(F32(B_APPLIED) + (b * A + i), e.applied[i]);
    }
    for (int j = 0; j < NHINGE; ++j) {
        G2_ST(F32(B_DOF_POS) + (b * NHINGE + j), e.dof_pos[j]);
        G2_ST(F32(B_DOF_VEL) + (b * NHINGE + j), e.dof_vel[j]);
        G2_ST(F32(B_LAST_DOF_VEL) + (b * NHINGE + j), e.last_dof_vel[j]);
        G2_ST(F32(B_TARGET) + (b * NHINGE + j), e.target[j]);
        G2_ST(F32(B_TORQUE) + (b * NHINGE + j), e.torque[j]);
    }
    G2_ST(F32(B_BASE_POS) + (b * 3), e.base_pos.x); G2_ST(F32(B_BASE_POS) + (b * 3 + 1), e.base_pos.y); G2_ST(F32(B_BASE_POS) + (b * 3 + 2), e.base_pos.z);
    G2_ST(F32(B_BASE_QUAT) + (b * 4), e.base_quat.w); G2_ST(F32(B_BASE_QUAT) + (b * 4 + 1), e.base_quat.x); G2_ST(F32(B_BASE_QUAT) + (b * 4 + 2), e.base_quat.y); G2_ST(F32(B_BASE_QUAT) + (b * 4 + 3), e.base_quat.z);
    G2_ST(F32(B_BASE_LIN_VEL) + (b * 3), e.lin_vel.x); G2_ST(F32(B_BASE_LIN_VEL) + (b * 3 + 1), e.lin_vel.y); G2_ST(F32(B_BASE_LIN_VEL) + (b * 3 + 2), e.lin_vel.z);
    G2_ST(F32(B_BASE_ANG_VEL) + (b * 3), e.ang_vel.x); G2_ST(F32(B_BASE_ANG_VEL) + (b * 3 + 1), e.ang_vel.y); G2_ST(F32(B_BASE_ANG_VEL) + (b * 3 + 2), e.ang_vel.z);
    G2_ST(F32(B_PROJ_G) + (b * 3), e.proj_g.x); G2_ST(F32(B_PROJ_G) + (b * 3 + 1), e.proj_g.y); G2_ST(F32(B_PROJ_G) + (b * 3 + 2), e.proj_g.z);
    G2_ST(F32(B_BASE_EULER) + (b * 3), e.euler.x); G2_ST(F32(B_BASE_EULER) + (b * 3 + 1), e.euler.y); G2_ST(F32(B_BASE_EULER) + (b * 3 + 2), e.euler.z);
    if (cfg.has_foot_tracking) {
        for (int i = 0; i < 4; ++i) {
            G2_ST(F32(B_AIR_TIME) + (b * 4 + i), e.air[i]);
            G2_ST(U8(B_FOOT_CONTACT) + (b * 4 + i), (uint8_t)e.fc[i]);
            G2_ST(U8(B_LAST_FOOT_CONTACT) + (b * 4 + i), (uint8_t)e.lfc[i]);
        }
    }
    for (int l = 0; l < NLINK; ++l) {
        G2_ST(F32(B_CONTACT_FORCE) + (b * NLINK + l) * 3, p.cforce[l].x);
        G2_ST(F32(B_CONTACT_FORCE) + (b * NLINK + l) * 3 + 1, p.cforce[l].y);
        G2_ST(F32(B_CONTACT_FORCE) + (b * NLINK + l) * 3 + 2, p.cforce[l].z);
    }
}

// ---------------------------------------------------------------------------------------------------
// One THREAD per env.  The whole control tick of env b runs in this thread; `p` is the thread's
// private working set (registers + thread-local memory).  Persistent physics state is env-major in HBM.
// ---------------------------------------------------------------------------------------------------
G2_HD void load_phys(const Buffers& bf, int b, Phys& p) {
    for (int i = 0; i < NQ; ++i) p.qpos[i] = G2_LD(F32(B_QPOS) + b * NQ + i);
    for (int i = 0; i < NV; ++i) p.qvel[i] = G2_LD(F32(B_QVEL) + b * NV + i);
    for (int i = 0; i < NV; ++i) p.qacc_ws[i] = G2_LD(F32(B_QACC_WS) + b * NV + i);
    p.warm = I32(B_WARM)[b];
    int nc = I32(B_NCACHE)[b];
    p.n_cache = nc;
    for (int k = 0; k < nc; ++k) {
        p.cache_pair[k] = I32(B_CACHE_PAIR)[b * MAX_CPAIRS + k];
        const float* cn = F32(B_CACHE_N) + (size_t)b * MAX_CPAIRS * 3 + 3 * k;
        p.cache_n[k] = mk3(cn[0], cn[1], cn[2]);
    }
    p.err = 0;
    p.cost = 0.0f;
    p.prev_cost = 0.0f;
    p.pd_mode = 0; p.pd_kp = 0.0f; p.pd_kv = 0.0f;
    for (int l = 0; l < NLINK; ++l) p.cforce[l] = mk3(0, 0, 0);
    p.n_contacts = 0;
    p.n_rows = 0;
    p.n_newton = 0;
}
G2_HD void store_phys(const Buffers& bf, int b, const Phys& p) {
    for (int i = 0; i < NQ; ++i) G2_ST(F32(B_QPOS) + b * NQ + i, p.qpos[i]);
    for (int i = 0; i < NV; ++i) G2_ST(F32(B_QVEL) + b * NV + i, p.qvel[i]);
    for (int i = 0; i < NV; ++i) G2_ST(F32(B_QACC_WS) + b * NV + i, p.qacc_ws[i]);
    G2_ST(I32(B_WARM) + (b), p.warm);
    G2_ST(I32(B_NCACHE) + (b), p.n_cache);
    for (int k = 0; k < p.n_cache; ++k) {
        G2_ST(I32(B_CACHE_PAIR) + (b * MAX_CPAIRS + k), p.cache_pair[k]);
        float* cn = F32(B_CACHE_N) + (size_t)b * MAX_CPAIRS * 3 + 3 * k;
        cn[0] = p.cache_n[k].x; cn[1] = p.cache_n[k].y; cn[2] = p.cache_n[k].z;
    }
}

// Block-wide phase alignment: the kernel is far larger than the SM's instruction cache (ncu: icc hit rate 59 %,
// no-instruction stalls ~ long-scoreboard stalls), so the warps of a block are kept in the same phase and share
// their instruction fetches.  `valid` = this thread owns an env (threads past B only take part in the barriers).
#if defined(__CUDA_ARCH__) && !defined(G2_WARP_ASYNC)
#define G2_PHASE_SYNC() __syncthreads()
#else
#define G2_PHASE_SYNC()
#endif
// Profiling build (-DG2_PHASE_CLOCKS): thread 0 of every block adds the SM clock ticks it spent in each phase to a global
// table (0 load+FK+env_pre, 1 smooth, 2 collide, 3 assemble, 4 solve, 5 integrate+FK, 6 env_post+store); read by
// profiles/tools/phase_clocks.py through go2_phase_clocks().
#if defined(G2_PHASE_CLOCKS) && defined(__CUDACC__)
__device__ unsigned long long g2_phase_clk[8];
#endif
#if defined(G2_PHASE_CLOCKS) && defined(__CUDA_ARCH__)
#define G2_CLK(i) do { if (threadIdx.x == 0) { long long t1_ = clock64(); atomicAdd(&g2_phase_clk[i], (unsigned long long)(t1_ - t0_)); t0_ = t1_; } } while (0)
#define G2_CLK_INIT() long long t0_ = clock64()
#else
#define G2_CLK(i)
#define G2_CLK_INIT()
#endif
#if defined(__CUDA_ARCH__) && defined(G2_FINE_SYNC2_ON)
#define G2_FINE_SYNC2() __syncthreads()
#else
#define G2_FINE_SYNC2()
#endif
G2_HD void t_env_step(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, const float* actions, int b,
                      bool valid, Phys& p, EnvLocal& e, EnvStepOut& out) {
    G2_CLK_INIT();
    const float* gl = F32(B_GLOBALS) + sp.parity * GLOBALS_LEN;  // global physics DR in force
    if (valid) {
        load_phys(bf, b, p);
        // The reference's substep starts from the link frames / inertias of the PREVIOUS substep's closing FK
        // (rigid_solver.py:3155-3180), and set_mass_shift / set_COM_shift refresh nothing (rigid_solver.py:2051-2071): after a
        // global mass / COM change the first substep still runs on inertias of the OLD values - except for envs whose state
        // was set afterwards (reset_idx draws the global DR before set_dofs_position / set_pos, whose FK refresh sees the new
        // values).  Same here: the tick-opening FK uses the other parity slot (= the previous tick's values) unless the env
        // was refreshed since.
        const float* gl0 = U8(B_FK_FRESH)[b] ? gl : F32(B_GLOBALS) + (sp.parity ^ 1) * GLOBALS_LEN;
        kinematics(m, gl0, p);
        velocities(m, p);
    }
    G2_FINE_SYNC2();
    bool act_late = false;
#ifdef __CUDA_ARCH__
    if (sp.act_flag != nullptr) {  // go2_step_host: the action rows are still on their way (copy stream); see StepParams.act_flag
        if (threadIdx.x == 0) {
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            for (;;) {
                unsigned int v;
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(sp.act_flag) : "memory");
                if (v == sp.act_seq) break;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 2000000000ull) { act_late = true; break; }  // never hang the GPU: flag the env and go on
                __nanosleep(200);
            }
        }
        act_late = __syncthreads_or(act_late ? 1 : 0) != 0;
    }
#endif
    if (valid) {
        if (act_late) p.err |= ERR_ACTION_WAIT;
        env_pre(m, cfg, sp, bf, actions, b, p, e, out);
    }
    if (sp.do_physics) {
        for (int s = 0; s < 2; ++s) {
            G2_PHASE_SYNC();
            G2_CLK(s == 0 ? 0 : 5);
            if (valid) smooth_dynamics(m, p);
            G2_PHASE_SYNC();
            G2_CLK(1);
            if (valid) detect_collisions(m, p, gl[G_FRIC_GROUND], gl[G_FRIC_ROBOT]);
            G2_PHASE_SYNC();
            G2_CLK(2);
            if (valid) assemble_constraints(m, p, cfg.solver_dense);
            G2_CLK(3);
            solve_constraints(m, p, valid);  // block-synchronous iteration loop: every thread calls it
            G2_PHASE_SYNC();
            G2_CLK(4);
            if (valid) integrate(m, p);
            G2_FINE_SYNC2();
            if (valid) {
                kinematics(m, gl, p);
                velocities(m, p);
            }
        }
    }
    G2_PHASE_SYNC();
    G2_CLK(5);
    if (valid) {
        env_post(m, cfg, sp, bf, b, p, e, out);
        store_phys(bf, b, p);
    }
    G2_CLK(6);
}

// ---------------------------------------------------------------------------------------------------
// Post-pass of a tick (one tiny kernel after the step kernel, csrc/go2_kernels.cu: go2_post_kernel).  The reference draws
// the GLOBAL physics DR (Genesis keeps one friction / mass shift / COM shift for all envs) inside reset_idx, i.e. whenever
// at least one env reset in the tick (go2_env_walk.py:737-756 friction with its reset-count throttle, :803-848 mass, COM,
// per-leg mass), and the privileged obs of that same tick already show the new values (:1115-1141).  "Did any env
// reset" is a grid-wide fact, so it is resolved here: the host passes this tick's candidate values in StepParams, the
// step kernel counts the resets into B_STATS, and the post-pass puts the candidates in force (B_GLOBALS, double-buffered
// by tick parity), patches the 9 global-DR columns of every env's privileged obs, and folds the tick's statistics into
// the accumulator row that the host reads every `stats_sync_interval` ticks.
// ---------------------------------------------------------------------------------------------------
G2_HD bool post_globals(const EnvConfig& cfg, const StepParams& sp, float n_resets, const float* cur, float* next) {
    for (int i = 0; i < GLOBALS_LEN; ++i) next[i] = cur[i];
    if (cfg.kind == ENV_BASE || n_resets < 0.5f) return false;
    bool changed = false;
    if (cfg.has_friction_dr) {
        float cnt = cur[G_DR_COUNTER] + n_resets;
        if (cnt >= (float)cfg.global_dr_interval) {
            cnt = 0.0f;
            next[G_FRIC_GROUND] = next[G_FRIC_ROBOT] = sp.cand_friction;
            changed = true;
        }
        next[G_DR_COUNTER] = cnt;
    }
    if (cfg.has_mass_dr) { next[G_MASS_SHIFT] = sp.cand_mass; changed = true; }
    if (cfg.has_com_dr) {
        for (int j = 0; j < 3; ++j) next[G_COM_SHIFT + j] = sp.cand_com[j];
        changed = true;
    }
    if (cfg.has_leg_mass_dr) {
        for (int i = 0; i < 4; ++i) next[G_MASS_SHIFT + cfg.hip_link[i]] = sp.cand_leg[i];
        changed = true;
    }
    return changed;
}
// the global-DR entries of one env's privileged obs row (layout of env_post / go2_env_walk.py:1115-1141)
G2_HD void post_patch_priv(const EnvConfig& cfg, const float* g, float* pr) {
    int k = cfg.num_obs + 3;
    pr[k] = g[G_FRIC_GROUND];
    k += 1 + 3 * NHINGE;
    G2_ST(pr + (k++), g[G_MASS_SHIFT]);
    G2_ST(pr + (k++), g[G_COM_SHIFT]); G2_ST(pr + (k++), g[G_COM_SHIFT + 1]); G2_ST(pr + (k++), g[G_COM_SHIFT + 2]);
    for (int i = 0; i < 4; ++i) G2_ST(pr + (k++), cfg.has_leg_mass_dr ? g[G_MASS_SHIFT + cfg.hip_link[i]] : 0.0f);
}

// Post-pass, stair env: the row of every env reset in this tick is redrawn with the tick's true reset count n (all ranks) -
// the step kernel could only draw provisionally (stair_pick_row) - and everything the spawn point feeds is redone exactly as
// reset_env does it: qpos xyz, the base_pos / last_x / terrain_row buffers and the tail of the privileged obs.
G2_HD bool post_stair_rows_due(const EnvConfig& cfg, const StepParams& sp, float n_resets) {
    return cfg.kind == ENV_STAIR && cfg.has_terrain && cfg.n_terrain_rows > 1 && !sp.lock_terrain_rows && !sp.replay_on &&
           !sp.force_reset_all && sp.n_resets_hint == 0 && n_resets >= 0.5f;
}
G2_HD void post_stair_rows(const Go2Model& m, const EnvConfig& cfg, const StepParams& sp, const Buffers& bf, int b, int n) {
    if (!U8(B_RESET)[b]) return;
    Rng rng;
    rng.env = (uint32_t)(cfg.env_offset + b); rng.step = sp.step_count; rng.k0 = cfg.seed_lo; rng.k1 = cfg.seed_hi;
    int row = stair_pick_row(rng4(rng, SLOT_RESET + 12), sp.terrain_max_row, n);
    if (row == I32(B_TERRAIN_ROW)[b]) return;
    v3 pos = mk3(cfg.row_center[row][0], cfg.row_center[row][1], cfg.row_center[row][2] + cfg.base_init_pos[2]);
    if (cfg.has_init_z) {
        u4 rq = rng4(rng, SLOT_RESET + 10);
        pos.z = pos.z + urange(cfg.init_z_lo, cfg.init_z_hi, rq.a) - cfg.base_init_pos[2] + cfg.base_init_pos[2];
    }
    G2_ST(I32(B_TERRAIN_ROW) + (b), row);
    G2_ST(F32(B_QPOS) + (b * NQ), pos.x); G2_ST(F32(B_QPOS) + (b * NQ + 1), pos.y); G2_ST(F32(B_QPOS) + (b * NQ + 2), pos.z);
    G2_ST(F32(B_BASE_POS) + (b * 3), pos.x); G2_ST(F32(B_BASE_POS) + (b * 3 + 1), pos.y); G2_ST(F32(B_BASE_POS) + (b * 3 + 2), pos.z);
    G2_ST(F32(B_LAST_X) + (b), pos.x);
    if (cfg.num_priv > 0)
        stair_priv_tail(m, cfg, F32(B_PRIV) + (size_t)b * cfg.num_priv, cfg.num_obs + 55, row, pos, ld4(F32(B_BASE_QUAT) + b * 4), true);
}

#undef F32
#undef I32
#undef U8

}  // namespace go2
