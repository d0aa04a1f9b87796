// go2_b200 - small vector / quaternion helpers (float, host+device).
// Conventions follow genesis/utils/geom.py (w,x,y,z quaternions; transform_quat_by_quat(v,u) = normalize(u*v)).
#pragma once
#include "go2_types.cuh"
#include <math.h>

namespace go2 {

#if defined(__CUDACC__) && defined(G2_ROLLED)
#define G2_UNROLL _Pragma("unroll 1")
#elif defined(__CUDACC__)
#define G2_UNROLL _Pragma("unroll")
#else
#define G2_UNROLL
#endif
#if defined(__CUDACC__) && defined(G2_ROLLED) && !defined(G2_ROLLED_ALL)
#define G2_UNROLL_H _Pragma("unroll")
#else
#define G2_UNROLL_H G2_UNROLL
#endif

G2_HD int tri(int i, int j) { return (i * (i + 1)) / 2 + j; }  // packed lower triangle, j <= i
// M[i][j] is structurally nonzero iff one of the dofs is a base dof or both hinges belong to the same leg
G2_HD bool m_nz(int i, int j) { return i < 6 || j < 6 || (((i - 6) & 3) == ((j - 6) & 3)); }
G2_HD int trs(int i, int j) { return i >= j ? tri(i, j) : tri(j, i); }
// Fixed Go2 tree: links 1-4 hips (parent = base), 5-8 thighs, 9-12 calves (parent = l - 4); hinge of link l = dof 5 + l.
// Checked against the loaded model at handle creation (go2_model_host.hpp).
G2_HD int par_link(int l) { return l == 0 ? -1 : (l <= 4 ? 0 : l - 4); }


struct v3 {
    float x, y, z;
};
struct q4 {
    float w, x, y, z;
};

G2_HD v3 mk3(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
G2_HD v3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }
G2_HD q4 mk4(float w, float x, float y, float z) { q4 r; r.w = w; r.x = x; r.y = y; r.z = z; return r; }
G2_HD q4 ld4(const float* p) { return mk4(p[0], p[1], p[2], p[3]); }
G2_HD v3 operator+(v3 a, v3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
G2_HD v3 operator-(v3 a, v3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
G2_HD v3 operator-(v3 a) { return mk3(-a.x, -a.y, -a.z); }
G2_HD v3 operator*(v3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
G2_HD v3 operator*(float s, v3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
G2_HD float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
G2_HD v3 cross(v3 a, v3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
G2_HD float norm3(v3 a) { return sqrtf(dot(a, a)); }
G2_HD v3 normalize3(v3 a) { return a * (1.0f / norm3(a)); }
G2_HD float comp(v3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
G2_HD float fminf3(float a, float b, float c) { return fminf(fminf(a, b), c); }
G2_HD float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

G2_HD q4 qmul(q4 u, q4 v) {
    return mk4(u.w * v.w - u.x * v.x - u.y * v.y - u.z * v.z, u.w * v.x + u.x * v.w + u.y * v.z - u.z * v.y,
               u.w * v.y - u.x * v.z + u.y * v.w + u.z * v.x, u.w * v.z + u.x * v.y - u.y * v.x + u.z * v.w);
}
G2_HD q4 qnormalize(q4 q) {
    float s = 1.0f / sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
    return mk4(q.w * s, q.x * s, q.y * s, q.z * s);
}
// compose child orientation v (local) under parent orientation u
G2_HD q4 qcompose(q4 v, q4 u) { return qnormalize(qmul(u, v)); }
G2_HD q4 qconj(q4 q) { return mk4(q.w, -q.x, -q.y, -q.z); }

// rotation matrix (rows) of a quaternion that is not assumed to be exactly unit
struct m3 {
    float a[3][3];
};
G2_HD m3 qmat(q4 q) {
    m3 R;
    float d = q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z;
    float s = 2.0f / d;
    float xs = q.x * s, ys = q.y * s, zs = q.z * s;
    float wx = q.w * xs, wy = q.w * ys, wz = q.w * zs, xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
    float yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
    R.a[0][0] = 1.0f - (yy + zz); R.a[0][1] = xy - wz;          R.a[0][2] = xz + wy;
    R.a[1][0] = xy + wz;          R.a[1][1] = 1.0f - (xx + zz); R.a[1][2] = yz - wx;
    R.a[2][0] = xz - wy;          R.a[2][1] = yz + wx;          R.a[2][2] = 1.0f - (xx + yy);
    return R;
}
G2_HD v3 mcol(const m3& R, int i) { return mk3(R.a[0][i], R.a[1][i], R.a[2][i]); }
G2_HD v3 mmul(const m3& R, v3 v) {
    return mk3(R.a[0][0] * v.x + R.a[0][1] * v.y + R.a[0][2] * v.z, R.a[1][0] * v.x + R.a[1][1] * v.y + R.a[1][2] * v.z,
               R.a[2][0] * v.x + R.a[2][1] * v.y + R.a[2][2] * v.z);
}
G2_HD v3 mtmul(const m3& R, v3 v) {  // R^T v
    return mk3(R.a[0][0] * v.x + R.a[1][0] * v.y + R.a[2][0] * v.z, R.a[0][1] * v.x + R.a[1][1] * v.y + R.a[2][1] * v.z,
               R.a[0][2] * v.x + R.a[1][2] * v.y + R.a[2][2] * v.z);
}
// rotate v by q (q need not be unit: result is divided by |q|^2, as geom.py:246-262 does)
G2_HD v3 qrot(v3 v, q4 q) { return mmul(qmat(q), v); }
G2_HD v3 qrot_inv(v3 v, q4 q) { return mtmul(qmat(q), v); }

// rotation vector -> quaternion with the reference's first-order renormalisation (geom.py:108-127)
G2_HD q4 rotvec_quat(v3 r) {
    float t2 = dot(r, r);
    if (t2 > EPSF * EPSF) {
        float t = sqrtf(t2), s, c;
#ifdef __CUDA_ARCH__
        sincosf(0.5f * t, &s, &c);
#else
        s = sinf(0.5f * t); c = cosf(0.5f * t);
#endif
        float k = s / t;
        q4 q = mk4(c, r.x * k, r.y * k, r.z * k);
        float f = 0.5f * (3.0f - (q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z));
        return mk4(q.w * f, q.x * f, q.y * f, q.z * f);
    }
    return mk4(1.0f, 0.0f, 0.0f, 0.0f);
}
// rotation about a unit axis by angle (same first-order renorm)
G2_HD q4 axis_angle_quat(v3 axis, float ang) { return rotvec_quat(axis * ang); }

// two unit vectors orthogonal to a (geom.py:391-405)
G2_HD void orthobasis(v3 a, v3& b, v3& c) {
    if (fabsf(a.y) < 0.5f) b = mk3(-a.x * a.y, 1.0f - a.y * a.y, -a.z * a.y);
    else b = mk3(-a.x * a.z, -a.y * a.z, 1.0f - a.z * a.z);
    b = normalize3(b);
    c = cross(a, b);
}

// constraint impedance / reference acceleration (geom.py:408-424); power==2 fast path
G2_HD void imp_aref(const float* p, float neg_pen, float vel, float pos, float& imp, float& aref) {
    float timeconst = p[0], dampratio = p[1], dmin = p[2], dmax = p[3], width = p[4], mid = p[5], power = p[6];
    float x = fabsf(neg_pen) / width;
    float ia, ib;
    if (power == 2.0f) {
        ia = x * x / mid;
        ib = 1.0f - (1.0f - x) * (1.0f - x) / (1.0f - mid);
    } else {
        ia = (1.0f / powf(mid, power - 1.0f)) * powf(x, power);
        ib = 1.0f - (1.0f / powf(1.0f - mid, power - 1.0f)) * powf(1.0f - x, power);
    }
    float iy = x < mid ? ia : ib;
    imp = clampf(dmin + iy * (dmax - dmin), dmin, dmax);
    if (x > 1.0f) imp = dmax;
    float b = 2.0f / (dmax * timeconst);
    float k = 1.0f / (dmax * dmax * timeconst * timeconst * dampratio * dampratio);
    aref = -b * vel - k * imp * pos;
}


// ---- counter-based RNG: Philox4x32-10 ---------------------------------------------------------------
struct u4 {
    uint32_t a, b, c, d;
};
G2_HD void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
    uint64_t p = (uint64_t)a * (uint64_t)b;
    hi = (uint32_t)(p >> 32);
    lo = (uint32_t)p;
}
G2_HD u4 philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0, lo0, hi1, lo1;
        mulhilo(0xD2511F53u, c0, hi0, lo0);
        mulhilo(0xCD9E8D57u, c2, hi1, lo1);
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    u4 o; o.a = c0; o.b = c1; o.c = c2; o.d = c3;
    return o;
}
G2_HD float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }  // [0,1)
G2_HD void boxmuller(uint32_t a, uint32_t b, float& z0, float& z1) {
    float u1 = ((float)(a >> 8) + 1.0f) * (1.0f / 16777216.0f);  // (0,1]
    float u2 = u01(b);
    float r = sqrtf(-2.0f * logf(u1));
    float s, c;
#ifdef __CUDA_ARCH__
    sincosf(6.283185307179586f * u2, &s, &c);
#else
    s = sinf(6.283185307179586f * u2); c = cosf(6.283185307179586f * u2);
#endif
    z0 = r * c;
    z1 = r * s;
}

}  // namespace go2
