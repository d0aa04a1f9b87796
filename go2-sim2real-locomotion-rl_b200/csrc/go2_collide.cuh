// go2_b200 - per-environment collision detection for the Go2 scene (28 convex primitives).
//
// Same contact model as the reference collider (MPR with warm-started centres + 4 perturbed
// re-detections for non-sphere pairs; collider/mpr.py, collider/narrowphase.py:515-961,
// collider/contact.py:165-370) with a GPU-friendly front end:
//   * no sweep-and-prune: the valid pair list is a compile-time table; each pair gets a direct AABB test
//     (same candidate set as collider/broadphase.py, canonical order instead of sort order);
//   * the per-pair contact-normal cache (319 x vec3 per env in the reference) is stored compactly as the
//     list of pairs that were in contact at the previous detection - equivalent because the reference
//     zeroes a pair's cache entry whenever it is not in contact;
//   * geom poses are derived on demand from link poses (never stored).
#pragma once
#include "go2_phys.cuh"

namespace go2 {

struct GeomPose {
    v3 pos;
    q4 quat;
};

template <class P>
G2_HD GeomPose geom_pose(const Go2Model& m, const P& p, int g) {
    GeomPose r;
    int l = m.g_link[g];
    if (l < 0) {
        r.pos = ld3(m.g_pos[g]);
        r.quat = qnormalize(ld4(m.g_quat[g]));
    } else {
        r.pos = p.lpos[l] + qrot(ld3(m.g_pos[g]), p.lquat[l]);
        r.quat = qcompose(ld4(m.g_quat[g]), p.lquat[l]);
    }
    return r;
}

// World AABBs from the oriented local box: centre +- |R| h  (== min/max over the 8 rotated corners).
G2_HD void link_box_add(Phys& p, int k, v3 c, v3 h) {
    v3 lo = c - h, hi = c + h;
    p.lk_lo[k] = mk3(fminf(p.lk_lo[k].x, lo.x), fminf(p.lk_lo[k].y, lo.y), fminf(p.lk_lo[k].z, lo.z));
    p.lk_hi[k] = mk3(fmaxf(p.lk_hi[k].x, hi.x), fmaxf(p.lk_hi[k].y, hi.y), fmaxf(p.lk_hi[k].z, hi.z));
}
G2_HDN void geom_aabbs(const Go2Model& m, Phys& p) {
    for (int k = 0; k <= NLINK; ++k) {
        p.lk_lo[k] = mk3(1e30f, 1e30f, 1e30f);
        p.lk_hi[k] = mk3(-1e30f, -1e30f, -1e30f);
    }
    for (int g = 0; g < NGEOM; ++g) {
        if (g == 0 && m.terrain_on) {  // whole field; the terrain narrow phase repeats the exact extent test
            v3 h = mk3(0.5f * m.terrain_r * m.terrain_sh, 0.5f * m.terrain_c * m.terrain_sh, 0.5f * (m.terrain_zmax - m.terrain_zbot));
            p.g_c[0] = ld3(m.terrain_org) + mk3(h.x, h.y, m.terrain_zbot + h.z);
            p.g_h[0] = h;
            link_box_add(p, 0, p.g_c[0], h);
            continue;
        }
        GeomPose gp = geom_pose(m, p, g);
        m3 R = qmat(gp.quat);
        v3 h = ld3(m.g_half[g]);
        p.g_c[g] = gp.pos;
        p.g_h[g] = mk3(fabsf(R.a[0][0]) * h.x + fabsf(R.a[0][1]) * h.y + fabsf(R.a[0][2]) * h.z,
                       fabsf(R.a[1][0]) * h.x + fabsf(R.a[1][1]) * h.y + fabsf(R.a[1][2]) * h.z,
                       fabsf(R.a[2][0]) * h.x + fabsf(R.a[2][1]) * h.y + fabsf(R.a[2][2]) * h.z);
        link_box_add(p, m.g_link[g] + 1, p.g_c[g], p.g_h[g]);
    }
}

// Support point of geom g (type/size from the model) posed at (pos, R) in world direction d.
G2_HD v3 support(const Go2Model& m, int g, v3 pos, const m3& R, v3 d) {
    int t = m.g_type[g];
    if (t == GEOM_SPHERE) return pos + d * m.g_size[g][0];
    v3 dl = mtmul(R, d);
    v3 v;
    if (t == GEOM_BOX) {
        v = mk3(dl.x < 0.0f ? -m.g_half[g][0] : m.g_half[g][0], dl.y < 0.0f ? -m.g_half[g][1] : m.g_half[g][1],
                dl.z < 0.0f ? -m.g_half[g][2] : m.g_half[g][2]);
    } else {
        // cylinder collision mesh = 32-gon prism: nearest ring vertex to the direction's azimuth
        const float step = 6.283185307179586f / 32.0f;
        float th = atan2f(dl.y, dl.x);
        float k = floorf(th / step + 0.5f);
        float s, c;
#ifdef __CUDA_ARCH__
        sincosf(k * step, &s, &c);
#else
        s = sinf(k * step); c = cosf(k * step);
#endif
        float r = m.g_size[g][0];
        v = mk3(r * c, r * s, dl.z < 0.0f ? -m.g_half[g][2] : m.g_half[g][2]);
    }
    return pos + mmul(R, v);
}

struct PairCtx {
    int ga, gb;
    v3 pa, pb;
    m3 Ra, Rb;
    const v3* prism;  // non-null: geom b is the current terrain prism (6 vertices, bottom 0..2, top 3..5)
};
struct Portal {
    v3 v[4], v1[4], v2[4];
};

// support point of the terrain prism (collider/support_field.py:265-282): top or bottom triangle by the sign of d.z
G2_HD v3 prism_support(const v3* prism, v3 d) {
    int i0 = d.z < 0.0f ? 0 : 3;
    v3 best = prism[i0];
    float bd = dot(best, d);
    for (int i = 1; i < 3; ++i) {
        float t = dot(prism[i0 + i], d);
        if (t > bd) { bd = t; best = prism[i0 + i]; }
    }
    return best;
}

G2_HD void mink_support(const Go2Model& m, const PairCtx& c, v3 d, v3& v, v3& v1, v3& v2) {
    v1 = support(m, c.ga, c.pa, c.Ra, d);
    v2 = c.prism ? prism_support(c.prism, -d) : support(m, c.gb, c.pb, c.Rb, -d);
    v = v1 - v2;
}
G2_HD v3 portal_dir(const Portal& s) { return normalize3(cross(s.v[2] - s.v[1], s.v[3] - s.v[1])); }
G2_HD bool reach_tol(const Portal& s, v3 v, v3 d) {
    float dv4 = dot(v, d);
    float d1 = fminf3(dv4 - dot(s.v[1], d), dv4 - dot(s.v[2], d), dv4 - dot(s.v[3], d));
    return d1 < CCD_TOL + CCD_EPS * fmaxf(1.0f, d1);
}
G2_HD void expand_portal(Portal& s, v3 v, v3 v1, v3 v2) {
    v3 v4v0 = cross(v, s.v[0]);
    int i;
    if (dot(s.v[1], v4v0) > 0.0f) i = dot(s.v[2], v4v0) > 0.0f ? 1 : 3;
    else i = dot(s.v[3], v4v0) > 0.0f ? 2 : 1;
    s.v[i] = v; s.v1[i] = v1; s.v2[i] = v2;
}
G2_HD bool tiny3(v3 a) { return fabsf(a.x) < CCD_EPS && fabsf(a.y) < CCD_EPS && fabsf(a.z) < CCD_EPS; }

// Minkowski portal refinement.  Returns true on collision with (normal: b -> a, penetration, position).
G2_HDN bool mpr(const Go2Model& m, const PairCtx& c, v3 ca, v3 cb, v3& normal, float& pen, v3& pos) {
    Portal s;
    s.v1[0] = ca; s.v2[0] = cb; s.v[0] = ca - cb;
    if (tiny3(s.v[0])) s.v[0].x += 10.0f * CCD_EPS;
    v3 d = -normalize3(s.v[0]);
    v3 v, v1, v2;
    mink_support(m, c, d, v, v1, v2);
    s.v[1] = v; s.v1[1] = v1; s.v2[1] = v2;
    if (dot(v, d) < CCD_EPS) return false;
    d = cross(s.v[0], s.v[1]);
    if (dot(d, d) < CCD_EPS) {
        if (tiny3(s.v[1])) {  // touching at the origin
            pen = 0.0f;
            normal = -normalize3(s.v[0]);
        } else {              // origin on the v0-v1 segment
            pen = norm3(s.v[1]);
            normal = -normalize3(s.v[1]);
        }
        pos = (s.v1[1] + s.v2[1]) * 0.5f;
        return true;
    }
    d = normalize3(d);
    mink_support(m, c, d, v, v1, v2);
    if (dot(v, d) < CCD_EPS) return false;
    s.v[2] = v; s.v1[2] = v1; s.v2[2] = v2;
    d = normalize3(cross(s.v[1] - s.v[0], s.v[2] - s.v[0]));
    if (dot(d, s.v[0]) > 0.0f) {
        v3 t;
        t = s.v[1]; s.v[1] = s.v[2]; s.v[2] = t;
        t = s.v1[1]; s.v1[1] = s.v1[2]; s.v1[2] = t;
        t = s.v2[1]; s.v2[1] = s.v2[2]; s.v2[2] = t;
        d = -d;
    }
    // discover the portal
    int trials = 0;
    for (;;) {
        mink_support(m, c, d, v, v1, v2);
        if (dot(v, d) < CCD_EPS) return false;
        bool again = false;
        if (dot(cross(s.v[1], v), s.v[0]) < -CCD_EPS) {
            s.v[2] = v; s.v1[2] = v1; s.v2[2] = v2;
            again = true;
        } else if (dot(cross(v, s.v[2]), s.v[0]) < -CCD_EPS) {
            s.v[1] = v; s.v1[1] = v1; s.v2[1] = v2;
            again = true;
        }
        if (!again) {
            s.v[3] = v; s.v1[3] = v1; s.v2[3] = v2;
            break;
        }
        d = normalize3(cross(s.v[1] - s.v[0], s.v[2] - s.v[0]));
        if (++trials == 15) return false;
    }
    // refine until the portal encloses the origin (the reference loop is unbounded; the cap only guards the GPU
    // against a non-terminating thread and is never reached in the oracle's runs)
    for (int guard = 0;; ++guard) {
        if (guard > 4 * CCD_ITERS) return false;
        d = portal_dir(s);
        if (dot(s.v[1], d) > -CCD_EPS) break;
        mink_support(m, c, d, v, v1, v2);
        if (!(dot(v, d) > -CCD_EPS) || reach_tol(s, v, d)) return false;
        expand_portal(s, v, v1, v2);
    }
    // penetration along the final portal normal
    for (int it = 0;; ++it) {
        d = portal_dir(s);
        mink_support(m, c, d, v, v1, v2);
        if (reach_tol(s, v, d) || it > CCD_ITERS) {
            pen = dot(d, s.v[1]);
            normal = -d;
            float b1 = dot(cross(s.v[2], s.v[3]), d), b2 = dot(cross(s.v[3], s.v[1]), d), b3 = dot(cross(s.v[1], s.v[2]), d);
            float inv = 0.5f / (b1 + b2 + b3);
            pos = ((b1 * s.v1[1] + b2 * s.v1[2] + b3 * s.v1[3]) + (b1 * s.v2[1] + b2 * s.v2[2] + b3 * s.v2[3])) * inv;
            return true;
        }
        expand_portal(s, v, v1, v2);
    }
}

// Geometry centres used to seed MPR; with a cached normal the centres are slid inside their bounding
// boxes so that the initial ray is aligned with it (collider/mpr.py:602-684).
G2_HD void seed_centres(const Go2Model& m, const PairCtx& c, v3 nws, bool have, v3& ca, v3& cb) {
    ca = c.pa;  // primitive centres coincide with the geom origin (local AABB is symmetric)
    cb = c.pb;
    if (have) {
        v3 delta = ca - cb;
        v3 nrm = normalize3(delta);
        if (norm3(cross(nws, nrm)) > 0.01f) {
            v3 off = dot(delta, nws) * nws - delta;
            float on = norm3(off);
            if (on > EPSF) {
                v3 dir = off * (1.0f / on);
                v3 da = mtmul(c.Ra, dir), db = mtmul(c.Rb, dir);
                float la = 2.0f * (m.g_half[c.ga][0] * fabsf(da.x) + m.g_half[c.ga][1] * fabsf(da.y) + m.g_half[c.ga][2] * fabsf(da.z));
                float lb = 2.0f * (m.g_half[c.gb][0] * fabsf(db.x) + m.g_half[c.gb][1] * fabsf(db.y) + m.g_half[c.gb][2] * fabsf(db.z));
                float ratio = fminf(on / (la + lb), 0.5f);
                ca = ca + dir * (la * ratio);
                cb = cb - dir * (lb * ratio);
            }
        }
    }
}

G2_HD void push_contact(const Go2Model& m, Phys& p, int ga, int gb, v3 n, v3 pos, float pen, float mu) {
    if (p.n_contacts < MAX_CONTACTS) {
        Contact& c = p.con[p.n_contacts++];
        c.pos = pos; c.normal = n; c.pen = pen; c.mu = mu;
        c.la = m.g_link[ga]; c.lb = m.g_link[gb];
    } else {
        p.err |= ERR_CONTACT_OVERFLOW;
    }
}

G2_HD float aabb_diag(const Go2Model& m, int g) {
    return 2.0f * sqrtf(m.g_half[g][0] * m.g_half[g][0] + m.g_half[g][1] * m.g_half[g][1] + m.g_half[g][2] * m.g_half[g][2]);
}

// One geom pair: first detection (+ up to 4 perturbed detections for non-sphere pairs).
// `nws`/`have`: cached normal from the previous detection.  Returns whether the pair is in contact and
// the normal to cache.
G2_HDN bool pair_contacts(const Go2Model& m, Phys& p, int ga, int gb, v3 nws, bool have, float mu, v3& n_out) {
    bool multi = m.g_type[ga] != GEOM_SPHERE && m.g_type[gb] != GEOM_SPHERE;
    float tol = 0.5f * MC_TOL * fminf(aabb_diag(m, ga), aabb_diag(m, gb));
    GeomPose A = geom_pose(m, p, ga), B = geom_pose(m, p, gb);
    PairCtx c;
    c.ga = ga; c.gb = gb; c.pa = A.pos; c.pb = B.pos; c.Ra = qmat(A.quat); c.Rb = qmat(B.quat); c.prism = nullptr;
    v3 ca, cb, n0, pos0;
    float pen0;
    seed_centres(m, c, nws, have, ca, cb);
    bool col0 = mpr(m, c, ca, cb, n0, pen0, pos0);
    if (!col0 && have) {  // retry without the warm start (narrowphase.py:690-700)
        seed_centres(m, c, nws, false, ca, cb);
        col0 = mpr(m, c, ca, cb, n0, pen0, pos0);
        have = false;
    }
    if (!col0) return false;
    n_out = n0;
    int first = p.n_contacts;
    push_contact(m, p, ga, gb, n0, pos0, pen0, mu);
    // The reference re-runs a detection with GJK/EPA when MPR's penetration exceeds the multi-contact tolerance and no cached
    // direction exists or the overlap is large (narrowphase.py:724-736); not built here.  For sphere pairs (the feet) both
    // algorithms converge to the analytic sphere / face contact (tests/test_oracle_closed_form.py); for the other pairs MPR's
    // answer can differ from GJK's multi-contact set, so the env raises an errno bit (measured: 0 of 202 k collisions in the
    // walk fixture, 5 of 4.7 k in the crouch one - DESIGN.md 2).
    if (multi && pen0 > tol && (!have || MC_TOL * pen0 >= 0.25f * tol)) p.err |= ERR_GJK_WANTED;
    if (!multi) return true;
    // perturbation axes: mix the principal axes of the smaller geom's link inertia with the normal
    v3 ax0, ax1;
    {
        float va = m.g_half[ga][0] * m.g_half[ga][1] * m.g_half[ga][2], vb = m.g_half[gb][0] * m.g_half[gb][1] * m.g_half[gb][2];
        int l = m.g_link[va < vb ? ga : gb];
        m3 R;
        if (l < 0) { q4 id = mk4(1, 0, 0, 0); R = qmat(id); }
        else R = qmat(qcompose(ld4(m.inertial_quat[l]), p.lquat[l]));
        int best = 0;
        float amax = 0.0f;
        for (int i = 0; i < 3; ++i) {
            float a = fabsf(dot(mcol(R, i), n0));
            if (a > amax) { amax = a; best = i; }
        }
        ax0 = mcol(R, (best + 1) % 3);
        ax0 = normalize3(ax0 - dot(n0, ax0) * n0);
        ax1 = cross(n0, ax0);
    }
    int n_pair = 1;
    for (int k = 1; k < 5; ++k) {
        v3 axis = (float)(2 * (k % 2) - 1) * ax0 + (float)(1 - 2 * ((k / 2) % 2)) * ax1;
        q4 qr = rotvec_quat(axis * MC_PERTURB);
        q4 qri = qconj(qr);
        // rotate both geoms about the first contact point, in opposite directions
        PairCtx cp = c;
        {
            q4 qa = qcompose(A.quat, qr);
            v3 rel = pos0 - A.pos;
            cp.pa = A.pos - (qrot(rel, qr) - rel);
            cp.Ra = qmat(qa);
            q4 qb = qcompose(B.quat, qri);
            rel = pos0 - B.pos;
            cp.pb = B.pos - (qrot(rel, qri) - rel);
            cp.Rb = qmat(qb);
        }
        v3 n, pos;
        float pen;
        seed_centres(m, cp, n0, true, ca, cb);  // the cache already holds this detection's first normal
        if (!mpr(m, cp, ca, cb, n, pen, pos)) continue;
        if (pen > tol && MC_TOL * pen >= 0.25f * tol) p.err |= ERR_GJK_WANTED;
        // undo the perturbation on the two surface points, first-order normal correction
        v3 pA = qrot((pos - 0.5f * pen * n) - pos0, qri) + pos0;
        v3 pB = qrot((pos + 0.5f * pen * n) - pos0, qr) + pos0;
        pos = 0.5f * (pA + pB);
        v3 tw = cross(n, n0);
        tw = mk3(clampf(tw.x, -MC_PERTURB, MC_PERTURB), clampf(tw.y, -MC_PERTURB, MC_PERTURB), clampf(tw.z, -MC_PERTURB, MC_PERTURB));
        n = n + cross(tw, n);
        pen = dot(n, pB - pA);
        bool rep = false;
        for (int i = 0; i < n_pair && !rep; ++i) {
            int idx = p.n_contacts - 1 - i;
            if (idx >= first && norm3(pos - p.con[idx].pos) < tol) rep = true;
        }
        if (!rep && pen > -tol) {
            push_contact(m, p, ga, gb, n, pos, fmaxf(pen, 0.0f), mu);
            n_pair++;
        }
    }
    return true;
}

// squared distance in the xy plane from point q to triangle (a, b, c); 0 inside
#ifndef G2_SPHERE_CULL_3D
#define G2_SPHERE_CULL_3D 1  // 0: only the horizontal footprint test (A/B, and the cross-check that the contacts are unchanged)
#endif
G2_HD float seg_dist2_xy(v3 q, v3 a, v3 b) {
    float ex = b.x - a.x, ey = b.y - a.y, px = q.x - a.x, py = q.y - a.y;
    float l2 = ex * ex + ey * ey;
    float t = l2 > 0.0f ? clampf((px * ex + py * ey) / l2, 0.0f, 1.0f) : 0.0f;
    float dx = px - t * ex, dy = py - t * ey;
    return dx * dx + dy * dy;
}
G2_HD float tri_dist2_xy(v3 q, v3 a, v3 b, v3 c) {
    float s0 = (b.x - a.x) * (q.y - a.y) - (b.y - a.y) * (q.x - a.x);
    float s1 = (c.x - b.x) * (q.y - b.y) - (c.y - b.y) * (q.x - b.x);
    float s2 = (a.x - c.x) * (q.y - c.y) - (a.y - c.y) * (q.x - c.x);
    if ((s0 >= 0.0f && s1 >= 0.0f && s2 >= 0.0f) || (s0 <= 0.0f && s1 <= 0.0f && s2 <= 0.0f)) return 0.0f;
    return fminf(seg_dist2_xy(q, a, b), fminf(seg_dist2_xy(q, b, c), seg_dist2_xy(q, c, a)));
}

// Squared distance from point q to the triangle (a, b, c) in 3-D (closest point by Voronoi region)
G2_HD float tri_dist2_3d(v3 q, v3 a, v3 b, v3 c) {
    const v3 ab = b - a, ac = c - a, aq = q - a;
    const float d1 = dot(ab, aq), d2 = dot(ac, aq);
    v3 cp;
    if (d1 <= 0.0f && d2 <= 0.0f) cp = a;
    else {
        const v3 bq = q - b;
        const float d3 = dot(ab, bq), d4 = dot(ac, bq);
        if (d3 >= 0.0f && d4 <= d3) cp = b;
        else {
            const float vc = d1 * d4 - d3 * d2;
            if (vc <= 0.0f && d1 >= 0.0f && d3 <= 0.0f) cp = a + ab * (d1 / (d1 - d3));
            else {
                const v3 cq = q - c;
                const float d5 = dot(ab, cq), d6 = dot(ac, cq);
                if (d6 >= 0.0f && d5 <= d6) cp = c;
                else {
                    const float vb = d5 * d2 - d1 * d6;
                    if (vb <= 0.0f && d2 >= 0.0f && d6 <= 0.0f) cp = a + ac * (d2 / (d2 - d6));
                    else {
                        const float va = d3 * d6 - d5 * d4;
                        if (va <= 0.0f && (d4 - d3) >= 0.0f && (d5 - d6) >= 0.0f) cp = b + (c - b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)));
                        else {
                            const float den = 1.0f / (va + vb + vc);
                            cp = a + ab * (vb * den) + ac * (vc * den);
                        }
                    }
                }
            }
        }
    }
    const v3 e = q - cp;
    return dot(e, e);
}
// Squared distance from point q to the prism under the triangle (a, b, c) (its top face; the prism goes straight down, far below q):
// 0 inside; else the nearest boundary point is on the top face or on a side wall.  On the wall under edge (u, v) the nearest point
// is level with q when q is not above the wall's top there - then the distance is the horizontal one.
G2_HD float prism_wall_dist2(v3 q, v3 u, v3 v) {
    const float ex = v.x - u.x, ey = v.y - u.y, px = q.x - u.x, py = q.y - u.y;
    const float l2 = ex * ex + ey * ey;
    const float t = l2 > 0.0f ? clampf((px * ex + py * ey) / l2, 0.0f, 1.0f) : 0.0f;
    if (q.z > u.z + t * (v.z - u.z)) return 1e30f;  // above the wall's top: the top face / its edges are nearer
    const float dx = px - t * ex, dy = py - t * ey;
    return dx * dx + dy * dy;
}
G2_HD float prism_dist2(v3 q, v3 a, v3 b, v3 c) {
    const float s0 = (b.x - a.x) * (q.y - a.y) - (b.y - a.y) * (q.x - a.x);
    const float s1 = (c.x - b.x) * (q.y - b.y) - (c.y - b.y) * (q.x - b.x);
    const float s2 = (a.x - c.x) * (q.y - c.y) - (a.y - c.y) * (q.x - c.x);
    const bool inside_xy = (s0 >= 0.0f && s1 >= 0.0f && s2 >= 0.0f) || (s0 <= 0.0f && s1 <= 0.0f && s2 <= 0.0f);
    if (inside_xy) {
        const v3 nrm = cross(b - a, c - a);  // height of the top face above / below q along z
        if (fabsf(nrm.z) > 1e-12f) {
            const float ztop = a.z - (nrm.x * (q.x - a.x) + nrm.y * (q.y - a.y)) / nrm.z;
            if (q.z <= ztop) return 0.0f;
        }
        return tri_dist2_3d(q, a, b, c);
    }
    return fminf(tri_dist2_3d(q, a, b, c), fminf(prism_wall_dist2(q, a, b), fminf(prism_wall_dist2(q, b, c), prism_wall_dist2(q, c, a))));
}

// Geom g against the heightfield terrain (collider/narrowphase.py:346-512 func_contact_mpr_terrain): the cells under
// the geom's extent are swept row by row as a strip of triangular prisms (each new vertex pair forms a prism with
// the previous two), MPR runs against every prism whose top reaches the geom's lowest point, at most
// TERRAIN_PAIR_CONTACTS contacts are kept per geom and near-duplicates are dropped.  The terrain frame is the
// world frame shifted by terrain_org (the reference builds it with an identity quaternion).
G2_HDN void terrain_contacts(const Go2Model& m, Phys& p, int g, float mu) {
    v3 org = ld3(m.terrain_org);
    if (m.terrain_mip) {
        // Conservative cull before any support evaluation: if the geom's box floats above the highest cell of every 4 x 4
        // block under it, no prism top can reach the geom's lowest point and the sweep below would find no candidate.
        // On the stair cfg all 27 geoms pass the broad phase (the terrain's box is the whole field) but only the feet and
        // calves are near the surface.
        const float sh = m.terrain_sh;
        const v3 lo = p.g_c[g] - p.g_h[g] - org, hi = p.g_c[g] + p.g_h[g] - org;
        int r0 = (int)floorf(lo.x / sh) - 1, r1 = (int)ceilf(hi.x / sh) + 1, c0 = (int)floorf(lo.y / sh) - 1, c1 = (int)ceilf(hi.y / sh) + 1;
        r0 = r0 > 0 ? r0 : 0; c0 = c0 > 0 ? c0 : 0;
        r1 = r1 < m.terrain_r - 1 ? r1 : m.terrain_r - 1; c1 = c1 < m.terrain_c - 1 ? c1 : m.terrain_c - 1;
        float top = m.terrain_zbot;
        for (int br = r0 >> TERRAIN_MIP_SHIFT; br <= (r1 >> TERRAIN_MIP_SHIFT); ++br)
            for (int bc = c0 >> TERRAIN_MIP_SHIFT; bc <= (c1 >> TERRAIN_MIP_SHIFT); ++bc)
                top = fmaxf(top, m.terrain_mip[(size_t)br * m.terrain_mip_c + bc]);
        if (r1 >= r0 && c1 >= c0 && lo.z - 1e-4f > top) return;
    }
    GeomPose G = geom_pose(m, p, g);
    PairCtx c;
    c.ga = g; c.gb = 0; c.pa = G.pos - org; c.pb = mk3(0, 0, 0); c.Ra = qmat(G.quat); c.Rb = c.Ra;
    v3 prism[6];
    c.prism = prism;
    // extent of the geom along the terrain axes from its own support function
    float hi[3], lo[3];
    for (int ax = 0; ax < 3; ++ax) {
        v3 d = mk3(ax == 0 ? 1.0f : 0.0f, ax == 1 ? 1.0f : 0.0f, ax == 2 ? 1.0f : 0.0f);
        hi[ax] = comp(support(m, g, c.pa, c.Ra, d), ax);
        lo[ax] = comp(support(m, g, c.pa, c.Ra, -d), ax);
    }
    float tmax[3] = {m.terrain_r * m.terrain_sh, m.terrain_c * m.terrain_sh, m.terrain_zmax};
    float tmin[3] = {0.0f, 0.0f, m.terrain_zbot};
    for (int i = 0; i < 3; ++i)
        if (tmax[i] < lo[i] || tmin[i] > hi[i]) return;
    for (int i = 0; i < 6; ++i) prism[i] = mk3(0.0f, 0.0f, m.terrain_zbot);
    float sh = m.terrain_sh;
    int r_min = (int)floorf(lo[0] / sh), r_max = (int)ceilf(hi[0] / sh);
    int c_min = (int)floorf(lo[1] / sh), c_max = (int)ceilf(hi[1] / sh);
    r_min = r_min > 0 ? r_min : 0;
    c_min = c_min > 0 ? c_min : 0;
    r_max = r_max < m.terrain_r - 1 ? r_max : m.terrain_r - 1;
    c_max = c_max < m.terrain_c - 1 ? c_max : m.terrain_c - 1;
    float tol = 0.5f * MC_TOL * aabb_diag(m, g);  // the terrain's own AABB diagonal is never the smaller one
    int n_con = 0, first = p.n_contacts;
    const bool sphere = m.g_type[g] == GEOM_SPHERE;
    const float sph_r2 = (m.g_size[g][0] + 1e-5f) * (m.g_size[g][0] + 1e-5f);
    const float sph_r3 = (m.g_size[g][0] + 1e-4f) * (m.g_size[g][0] + 1e-4f);
    // Two passes per batch of candidates so that the lanes of a warp run MPR together: (1) walk the vertex strip of each row
    // (vertex n of row r = cell column c_min + (n-1)/2, row r + (n-1)%2) and record the prisms whose top reaches the geom's
    // lowest point; (2) rebuild each recorded prism and run MPR on it, in strip order (so duplicates and the contact cap
    // resolve exactly as in a single sweep).  ncu on the single-sweep version: MPR ran with 3 of 32 lanes.
    constexpr int CAP = 48;
    int cand[CAP];
    const int nmax = 2 * (c_max - c_min + 1);
    int r = r_min, n = 1;
    float hA = 0.0f, hB = 0.0f;
    while (r < r_max && nmax >= 3 && n_con < TERRAIN_PAIR_CONTACTS) {
        int ncand = 0;
        while (r < r_max && ncand < CAP) {
            int cc = c_min + ((n - 1) >> 1), i = (n - 1) & 1;
            float h = m.terrain_hf[(size_t)(r + i) * m.terrain_c + cc];
            if (n > 2 && (hA >= lo[2] || hB >= lo[2] || h >= lo[2])) {
                bool keep = true;
                if (sphere) {
                    // a sphere can only touch the prism if its centre is within one radius of the prism's triangle in the
                    // xy plane (MPR answers "no collision" for shapes separated along a horizontal axis): skips ~2/3 of the
                    // foot candidates, whose 2 x 2-cell footprint holds 8-12 prisms of which the foot touches 1-3
                    v3 t[3];
                    const float hz[3] = {hA, hB, h};
                    for (int j = 0; j < 3; ++j) {
                        int vn = n - 2 + j;
                        t[j] = mk3(sh * (float)(r + ((vn - 1) & 1)), sh * (float)(c_min + ((vn - 1) >> 1)), G2_SPHERE_CULL_3D ? hz[j] : 0.0f);
                    }
                    // ... and, exactly: only if the centre is within one radius (+ 0.1 mm) of the PRISM in 3-D.  A foot standing on
                    // a tread is ~21 mm above it, so of the triangles inside its 22 mm footprint only those within ~7 mm of the
                    // contact point can touch it.
                    keep = G2_SPHERE_CULL_3D ? prism_dist2(c.pa, t[0], t[1], t[2]) <= sph_r3 : tri_dist2_xy(c.pa, t[0], t[1], t[2]) <= sph_r2;
                }
                if (keep) cand[ncand++] = (r << 12) | n;
            }
            hA = hB;
            hB = h;
            if (++n > nmax) { n = 1; ++r; }
        }
#ifndef G2_NO_MPR_SM
        // MPR over the recorded prisms as a per-lane state machine with ONE Minkowski-support site per trip: a lane fetches
        // its next candidate as soon as the current one is decided, so the lanes of a warp stay on the support evaluation
        // together whatever the depth at which their prisms leave the algorithm (ncu on the call-per-candidate version: the
        // MPR code ran with 2-5 of 32 lanes).  Per lane the operations and their order are exactly those of mpr().
        {
            enum { ST_FETCH = 0, ST_S1, ST_S2, ST_DISC, ST_REFINE, ST_PEN };
            int k = 0, stage = ST_FETCH, cnt = 0;
            Portal s;
            v3 d = mk3(0, 0, 0);
            for (;;) {
                if (stage == ST_FETCH) {
                    if (k >= ncand || n_con >= TERRAIN_PAIR_CONTACTS) break;
                    const int cr = cand[k] >> 12, cn = cand[k] & 4095;
                    ++k;
                    v3 cb = mk3(0, 0, 0);
                    for (int j = 0; j < 3; ++j) {
                        int vn = cn - 2 + j;
                        int cc = c_min + ((vn - 1) >> 1), i = (vn - 1) & 1;
                        float x = sh * (float)(cr + i), y = sh * (float)cc;
                        prism[j] = mk3(x, y, m.terrain_zbot);
                        prism[3 + j] = mk3(x, y, m.terrain_hf[(size_t)(cr + i) * m.terrain_c + cc]);
                    }
                    for (int j = 0; j < 6; ++j) cb = cb + prism[j];
                    cb = cb * (1.0f / 6.0f);
                    s.v1[0] = c.pa; s.v2[0] = cb; s.v[0] = c.pa - cb;
                    if (tiny3(s.v[0])) s.v[0].x += 10.0f * CCD_EPS;
                    d = -normalize3(s.v[0]);
                    stage = ST_S1;
                }
                v3 v, v1, v2;
                mink_support(m, c, d, v, v1, v2);
                bool hit = false;
                v3 nrm = mk3(0, 0, 0), pos = mk3(0, 0, 0);
                float pen = 0.0f;
                if (stage == ST_S1) {
                    s.v[1] = v; s.v1[1] = v1; s.v2[1] = v2;
                    if (dot(v, d) < CCD_EPS) { stage = ST_FETCH; continue; }
                    d = cross(s.v[0], s.v[1]);
                    if (dot(d, d) < CCD_EPS) {
                        if (tiny3(s.v[1])) {  // touching at the origin
                            pen = 0.0f;
                            nrm = -normalize3(s.v[0]);
                        } else {              // origin on the v0-v1 segment
                            pen = norm3(s.v[1]);
                            nrm = -normalize3(s.v[1]);
                        }
                        pos = (s.v1[1] + s.v2[1]) * 0.5f;
                        hit = true;
                    } else {
                        d = normalize3(d);
                        stage = ST_S2;
                        continue;
                    }
                } else if (stage == ST_S2) {
                    if (dot(v, d) < CCD_EPS) { stage = ST_FETCH; continue; }
                    s.v[2] = v; s.v1[2] = v1; s.v2[2] = v2;
                    d = normalize3(cross(s.v[1] - s.v[0], s.v[2] - s.v[0]));
                    if (dot(d, s.v[0]) > 0.0f) {
                        v3 t;
                        t = s.v[1]; s.v[1] = s.v[2]; s.v[2] = t;
                        t = s.v1[1]; s.v1[1] = s.v1[2]; s.v1[2] = t;
                        t = s.v2[1]; s.v2[1] = s.v2[2]; s.v2[2] = t;
                        d = -d;
                    }
                    cnt = 0;
                    stage = ST_DISC;
                    continue;
                } else if (stage == ST_DISC) {
                    if (dot(v, d) < CCD_EPS) { stage = ST_FETCH; continue; }
                    bool again = false;
                    if (dot(cross(s.v[1], v), s.v[0]) < -CCD_EPS) {
                        s.v[2] = v; s.v1[2] = v1; s.v2[2] = v2;
                        again = true;
                    } else if (dot(cross(v, s.v[2]), s.v[0]) < -CCD_EPS) {
                        s.v[1] = v; s.v1[1] = v1; s.v2[1] = v2;
                        again = true;
                    }
                    if (again) {
                        d = normalize3(cross(s.v[1] - s.v[0], s.v[2] - s.v[0]));
                        if (++cnt == 15) stage = ST_FETCH;
                        continue;
                    }
                    s.v[3] = v; s.v1[3] = v1; s.v2[3] = v2;
                    // head of the refinement loop (guard = 0)
                    cnt = 0;
                    d = portal_dir(s);
                    stage = dot(s.v[1], d) > -CCD_EPS ? ST_PEN : ST_REFINE;
                    continue;
                } else {  // ST_REFINE / ST_PEN: same helpers, other exits
                    bool rt = reach_tol(s, v, d);
                    if (stage == ST_REFINE) {
                        if (!(dot(v, d) > -CCD_EPS) || rt) { stage = ST_FETCH; continue; }
                    } else if (rt || cnt > CCD_ITERS) {
                        pen = dot(d, s.v[1]);
                        nrm = -d;
                        float b1 = dot(cross(s.v[2], s.v[3]), d), b2 = dot(cross(s.v[3], s.v[1]), d), b3 = dot(cross(s.v[1], s.v[2]), d);
                        float inv = 0.5f / (b1 + b2 + b3);
                        pos = ((b1 * s.v1[1] + b2 * s.v1[2] + b3 * s.v1[3]) + (b1 * s.v2[1] + b2 * s.v2[2] + b3 * s.v2[3])) * inv;
                        hit = true;
                    }
                    if (!hit) {
                        expand_portal(s, v, v1, v2);
                        ++cnt;
                        if (stage == ST_REFINE && cnt > 4 * CCD_ITERS) { stage = ST_FETCH; continue; }
                        d = portal_dir(s);
                        if (stage == ST_REFINE && dot(s.v[1], d) > -CCD_EPS) { stage = ST_PEN; cnt = 0; }
                        continue;
                    }
                }
                // decided: collision
                stage = ST_FETCH;
                pos = pos + org;
                bool dup = false;
                for (int j = 0; j < n_con && !dup; ++j) {
                    int idx = p.n_contacts - 1 - j;
                    if (idx >= first && norm3(pos - p.con[idx].pos) < tol) dup = true;
                }
                if (!dup) {
                    push_contact(m, p, g, 0, nrm, pos, pen, mu);
                    n_con++;
                }
            }
        }
#else
        for (int k = 0; k < ncand && n_con < TERRAIN_PAIR_CONTACTS; ++k) {
            const int cr = cand[k] >> 12, cn = cand[k] & 4095;
            v3 cb = mk3(0, 0, 0);
            for (int j = 0; j < 3; ++j) {
                int vn = cn - 2 + j;
                int cc = c_min + ((vn - 1) >> 1), i = (vn - 1) & 1;
                float x = sh * (float)(cr + i), y = sh * (float)cc;
                prism[j] = mk3(x, y, m.terrain_zbot);
                prism[3 + j] = mk3(x, y, m.terrain_hf[(size_t)(cr + i) * m.terrain_c + cc]);
            }
            for (int j = 0; j < 6; ++j) cb = cb + prism[j];
            cb = cb * (1.0f / 6.0f);
            v3 nrm, pos;
            float pen;
            if (mpr(m, c, c.pa, cb, nrm, pen, pos)) {
                pos = pos + org;
                bool dup = false;
                for (int j = 0; j < n_con && !dup; ++j) {
                    int idx = p.n_contacts - 1 - j;
                    if (idx >= first && norm3(pos - p.con[idx].pos) < tol) dup = true;
                }
                if (!dup) {
                    push_contact(m, p, g, 0, nrm, pos, pen, mu);
                    n_con++;
                }
            }
        }
#endif
    }
}

// Full detection for one env: AABBs -> static pair table -> narrow phase; rebuilds the normal cache.
// `fric_ground` / `fric_robot`: friction of the ground entity's and the robot entity's geoms (set_friction is per entity,
// rigid_geom.py:327-336); a contact uses max(mu_a, mu_b, 0.01) (narrowphase.py).
G2_HDN void detect_collisions(const Go2Model& m, Phys& p, float fric_ground, float fric_robot) {
    geom_aabbs(m, p);
    p.n_contacts = 0;
    int n_new = 0;
    int new_pair[MAX_CPAIRS];
    v3 new_n[MAX_CPAIRS];
    const float mu_rr = fmaxf(fric_robot, 1e-2f), mu_gr = fmaxf(fmaxf(fric_ground, fric_robot), 1e-2f);
    // Two-level broad phase: 79 (link, link) box tests first; a run of geom pairs is only walked if the boxes of its two links
    // overlap (a walking Go2's legs rarely do).  Conservative (a link box contains its geom boxes), so the candidate set and
    // its order - hence every contact - are those of the flat sweep over the 319 geom pairs; ~4x fewer thread-local loads.
    uint32_t ov0 = 0u, ov1 = 0u, ov2 = 0u;
    for (int k = 0; k < m.n_lp; ++k) {
        const int a = m.lp_a[k], b = m.lp_b[k];
        const v3 alo = p.lk_lo[a], ahi = p.lk_hi[a], blo = p.lk_lo[b], bhi = p.lk_hi[b];
        const bool ov = alo.x < bhi.x && blo.x < ahi.x && alo.y < bhi.y && blo.y < ahi.y && alo.z < bhi.z && blo.z < ahi.z;
        const uint32_t bit = ov ? (1u << (k & 31)) : 0u;
        if (k < 32) ov0 |= bit;
        else if (k < 64) ov1 |= bit;
        else ov2 |= bit;
    }
    // The sweep only COLLECTS the pairs whose boxes overlap (up to NCAND at a time); the narrow phase then runs on "the k-th
    // candidate of every lane" together.  Inside the sweep the lanes of a warp met their candidates at different pair indices and
    // ran MPR 8 lanes wide (ncu: support / mpr / mtmul at 7.5-9.3 active lanes); same candidates in the same order per env, so
    // every contact is unchanged.
    constexpr int NCAND = 16;
    int ip = 0, r = 0;
    while (r < m.n_run) {
        uint16_t cand[NCAND];
        int nc = 0;
        while (r < m.n_run && nc < NCAND) {
            const int end = m.run_end[r], k = m.run_lp[r];
            const uint32_t w = k < 32 ? ov0 : (k < 64 ? ov1 : ov2);
            if (!((w >> (k & 31)) & 1u)) { ip = end; ++r; continue; }
            for (; ip < end && nc < NCAND; ++ip) {
                const int a = m.pair_a[ip], b = m.pair_b[ip];
                v3 dc = p.g_c[a] - p.g_c[b], hs = p.g_h[a] + p.g_h[b];
                if (fabsf(dc.x) < hs.x && fabsf(dc.y) < hs.y && fabsf(dc.z) < hs.z) cand[nc++] = (uint16_t)ip;
            }
            if (ip >= end) ++r;
        }
        for (int ic = 0; ic < nc; ++ic) {
            const int jp = cand[ic];
            int a = m.pair_a[jp], b = m.pair_b[jp];
            const float mu = (a == 0 || b == 0) ? mu_gr : mu_rr;
            if (m.terrain_on && (a == 0 || b == 0)) {  // geom 0 is the terrain: no normal cache, own narrow phase
                terrain_contacts(m, p, a == 0 ? b : a, mu);
                continue;
            }
            // narrow phase wants the lower geom type first (sphere < cylinder < box)
            int ga = a, gb = b;
            if (m.g_type[ga] > m.g_type[gb]) { ga = b; gb = a; }
            v3 nws = mk3(0, 0, 0);
            bool have = false;
            for (int k = 0; k < p.n_cache; ++k)
                if (p.cache_pair[k] == jp) { nws = p.cache_n[k]; have = true; }
            v3 n;
            if (pair_contacts(m, p, ga, gb, nws, have, mu, n)) {
                if (n_new < MAX_CPAIRS) {
                    new_pair[n_new] = jp;
                    new_n[n_new] = n;
                    n_new++;
                } else {
                    p.err |= ERR_CPAIR_OVERFLOW;
                }
            }
        }
    }
    p.n_cache = n_new;
    for (int k = 0; k < n_new; ++k) {
        p.cache_pair[k] = new_pair[k];
        p.cache_n[k] = new_n[k];
    }
}

}  // namespace go2
