/* sfp_math.h — the arithmetic of the physics substep, written once.
 *
 * Every function here is `__host__ __device__` and is compiled twice: by g++ into the CPU
 * oracle (oracle/sfp_ref.cpp) and by nvcc into the sm_100a kernels (starframe_b200/csrc).
 * Both compilations switch FMA contraction OFF (g++ -ffp-contract=off, nvcc --fmad=false) and
 * use only IEEE-754 correctly rounded operations (+ - * / sqrtf) plus exact ones (floorf,
 * fabsf, fminf, fmaxf), so that every DISCRETE decision (cell index, AABB overlap, contact /
 * no contact, which feature pair is closest) and every float result is the same bit pattern
 * on both sides.  No sinf/cosf/atan2f in here: their
 * results differ between glibc and the CUDA math library.
 *
 * Algorithm source: the reference tree is not mounted; the step follows
 * north_star's stage list and the public XPBD rigid-body algorithm (Müller et al. 2020,
 * "Detailed Rigid Body Simulation with Extended Position Based Dynamics"), specialised to
 * 2D: scalar angle / inertia, rotation stored as (cos, sin).
 */
#ifndef SFP_MATH_H_
#define SFP_MATH_H_

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define SFP_HD __host__ __device__ __forceinline__
/* for the rarely taken heavy paths (8-vertex polygons): kept out of line so that their
 * registers and code do not weigh on the kernels' common path */
#define SFP_HD_COLD static __host__ __device__ __noinline__
#else
#define SFP_HD static inline
#define SFP_HD_COLD static inline
#endif

namespace sfp {

struct v2 {
    float x, y;
};

SFP_HD v2 mk(float x, float y) {
    v2 r;
    r.x = x;
    r.y = y;
    return r;
}
SFP_HD v2 operator+(v2 a, v2 b) { return mk(a.x + b.x, a.y + b.y); }
SFP_HD v2 operator-(v2 a, v2 b) { return mk(a.x - b.x, a.y - b.y); }
SFP_HD v2 operator*(float s, v2 a) { return mk(s * a.x, s * a.y); }
SFP_HD v2 neg(v2 a) { return mk(-a.x, -a.y); }
SFP_HD float dot(v2 a, v2 b) { return a.x * b.x + a.y * b.y; }
SFP_HD float cross(v2 a, v2 b) { return a.x * b.y - a.y * b.x; }
/* omega x r for a scalar angular velocity */
SFP_HD v2 crossz(float w, v2 r) { return mk(-w * r.y, w * r.x); }
SFP_HD float len(v2 a) { return sqrtf(a.x * a.x + a.y * a.y); }
/* rotate v by the rotor q = (cos, sin) */
SFP_HD v2 rot(v2 q, v2 v) { return mk(q.x * v.x - q.y * v.y, q.y * v.x + q.x * v.y); }
SFP_HD v2 rot_inv(v2 q, v2 v) { return mk(q.x * v.x + q.y * v.y, q.x * v.y - q.y * v.x); }

/* q <- normalise(q + dtheta * perp(q)).  A rotation by atan(dtheta): first order in dtheta,
 * exactly unit length, built from + * / sqrtf only.  rot_delta_tan() is its exact inverse, so
 * an unconstrained body keeps its angular velocity through integrate -> derive. */
SFP_HD v2 rot_advance(v2 q, float dtheta) {
    float c = q.x - dtheta * q.y;
    float s = q.y + dtheta * q.x;
    float inv = 1.0f / sqrtf(c * c + s * s);
    return mk(c * inv, s * inv);
}
/* The same rotation for the positional corrections of the solver, where it sits on every
 * constraint's dependent chain: renormalised with one Newton step of 1/sqrt (q * (1.5 - 0.5 |q|^2))
 * instead of sqrtf and a division.  |q + dtheta perp q|^2 = 1 + dtheta^2, so the length error
 * left behind is 3/8 dtheta^4 < 4e-7 for the |dtheta| <= 0.03 this branch accepts (the step is
 * self-correcting: it measures the length it is given), and integrate_body() renormalises
 * exactly once per substep anyway.  Larger corrections (light bodies on a taut joint) take the
 * exact routine: the Newton step diverges beyond |dtheta| ~ 0.8. */
SFP_HD v2 rot_nudge(v2 q, float dtheta) {
#ifdef SFP_EXACT_NUDGE
    return rot_advance(q, dtheta);
#else
    if (!(dtheta * dtheta <= 1.0e-3f)) return rot_advance(q, dtheta);
    float c = q.x - dtheta * q.y;
    float s = q.y + dtheta * q.x;
    float k = 1.5f - 0.5f * (c * c + s * s);
    return mk(c * k, s * k);
#endif
}
/* tan of the angle from qp to q */
SFP_HD float rot_delta_tan(v2 q, v2 qp) {
    float sd = q.y * qp.x - q.x * qp.y;
    float cd = q.x * qp.x + q.y * qp.y;
    return sd / cd;
}

/* atan / atan2 from + - * / only (Cephes' atanf reduction and polynomial, |error| < 2e-7 rad):
 * the library's and the device's atan2f differ in the last bits, and the angle reported to the
 * scene graph must be the same number on both sides. */
SFP_HD float atan_det(float x) {
    float sign = 1.0f;
    if (x < 0.0f) {
        sign = -1.0f;
        x = -x;
    }
    float y;
    if (x > 2.414213562373095f) { /* tan(3 pi / 8) */
        y = 1.5707963267948966f;
        x = -(1.0f / x);
    } else if (x > 0.4142135623730950f) { /* tan(pi / 8) */
        y = 0.7853981633974483f;
        x = (x - 1.0f) / (x + 1.0f);
    } else {
        y = 0.0f;
    }
    float z = x * x;
    float p = (((8.05374449538e-2f * z - 1.38776856032e-1f) * z + 1.99777106478e-1f) * z -
               3.33329491539e-1f) * z * x + x;
    return sign * (y + p);
}
SFP_HD float atan2_det(float y, float x) {
    const float pi = 3.14159265358979323846f;
    if (x > 0.0f) return atan_det(y / x);
    if (x < 0.0f) return (y >= 0.0f) ? atan_det(y / x) + pi : atan_det(y / x) - pi;
    if (y > 0.0f) return 0.5f * pi;
    if (y < 0.0f) return -0.5f * pi;
    return 0.0f;
}

/* ------------------------------------------------------------------------------------- */
/* State records (SoA: one array per record; SURVEY.md §8a A0, 88 B/body)                */
/* ------------------------------------------------------------------------------------- */
struct f4 {
    float x, y, z, w;
}; /* layout-compatible with CUDA float4 */
struct f2 {
    float x, y;
};

SFP_HD uint32_t f2u(float f) {
    union {
        float f;
        uint32_t u;
    } c;
    c.f = f;
    return c.u;
}
SFP_HD float u2f(uint32_t u) {
    union {
        float f;
        uint32_t u;
    } c;
    c.u = u;
    return c.f;
}

/* pose  = (x, y, cos, sin)      vel = (vx, vy, omega, 0)
 * prev  = (dx, dy, cos0, sin0): displacement accumulated in the current substep and the
 *         rotation at its start.  Velocities are derived from (dx, dy), never from the
 *         difference of two absolute positions: at |x| ~ 10..100 m an FP32 position is
 *         quantised to ~1e-6..1e-5 m, which divided by h = 1/480 s would inject up to
 *         ~5e-3 m/s of noise per substep (measured: 1 % gravity error in half a second).
 * minv  = (1/m, 1/I)            shape = (hx, hy, r, bits(type))
 * mat   = (mu_s, mu_d, e, bits(category | mask<<16))                                    */
enum { SHAPE_CIRCLE = 0, SHAPE_CAPSULE = 1, SHAPE_RECT = 2 };

SFP_HD uint32_t shape_type(f4 shape) { return f2u(shape.w) & 3u; }

/* what the scene graph exchanges per body (layout of SfpBodyState in sfp.h) */
struct StateRec {
    float x, y, c, s, vx, vy, omega, angle;
};
SFP_HD StateRec state_pack(f4 pose, f4 vel) {
    StateRec r;
    r.x = pose.x;
    r.y = pose.y;
    r.c = pose.z;
    r.s = pose.w;
    r.vx = vel.x;
    r.vy = vel.y;
    r.omega = vel.z;
    r.angle = atan2_det(pose.w, pose.z);
    return r;
}
/* sin and cos from + - * only (Cephes' sinf / cosf: octant reduction with a three-part pi/4,
 * two polynomials; |error| < 2e-7 for |x| < 8192, larger arguments are first brought into that
 * range and lose accuracy accordingly).  Same reason as atan_det: the library's and the
 * device's functions differ in the last bits. */
SFP_HD void sincos_det(float x, float* sn, float* cs) {
    float ax = fabsf(x);
    if (!(ax < 8192.0f)) {
        if (!(ax - ax == 0.0f)) { /* not finite */
            *sn = 0.0f;
            *cs = 1.0f;
            return;
        }
        ax = ax - 6.283185307179586f * floorf(ax * 0.15915494309189535f);
    }
    int j = (int)(ax * 1.27323954473516f); /* 4 / pi */
    float y = (float)j;
    if (j & 1) {
        j += 1;
        y += 1.0f;
    }
    j &= 7;
    float r = ((ax - y * 0.78515625f) - y * 2.4187564849853515625e-4f) - y * 3.77489497744594108e-8f;
    float z = r * r;
    float ps = ((-1.9515295891e-4f * z + 8.3321608736e-3f) * z - 1.6666654611e-1f) * z * r + r;
    float pc = ((2.443315711809948e-5f * z - 1.388731625493765e-3f) * z + 4.166664568298827e-2f) * z * z -
               0.5f * z + 1.0f;
    /* octant pairs 0: (s, c)  2: (c, -s)  4: (-s, -c)  6: (-c, s) */
    float s0 = ps, c0 = pc;
    if (j == 2) {
        s0 = pc;
        c0 = -ps;
    } else if (j == 4) {
        s0 = -ps;
        c0 = -pc;
    } else if (j == 6) {
        s0 = -pc;
        c0 = ps;
    }
    *sn = (x < 0.0f) ? -s0 : s0;
    *cs = c0;
}

/* a state whose rotor is (0, 0) carries its rotation in `angle` only */
SFP_HD void state_unpack(StateRec r, f2 minv, f4* pose, f4* vel) {
    pose->x = r.x;
    pose->y = r.y;
    if (!(r.c * r.c + r.s * r.s > 0.0f)) sincos_det(r.angle, &r.s, &r.c);
    float inv = 1.0f / sqrtf(r.c * r.c + r.s * r.s);
    pose->z = r.c * inv;
    pose->w = r.s * inv;
    bool is_static = !(minv.x > 0.0f || minv.y > 0.0f);
    vel->x = is_static ? 0.0f : r.vx;
    vel->y = is_static ? 0.0f : r.vy;
    vel->z = is_static ? 0.0f : r.omega;
    vel->w = 0.0f;
}

/* ------------------------------------------------------------------------------------- */
/* Broadphase arithmetic (A2)                                                            */
/* ------------------------------------------------------------------------------------- */
#define SFP_CELL_MIN (-32768)
#define SFP_CELL_MAX (32767)

/* swept AABB (lo.x, lo.y, hi.x, hi.y) of a collider for one frame of length dt: the tight box
 * at the current pose, grown by `margin` plus the spin allowance on every side, and extended
 * by vel_expand*dt*v only on the side the body is moving towards.  (A symmetric speed-based
 * expansion makes every body of a free-falling pile a candidate partner of dozens of
 * neighbours, which costs pairs and, worse, constraint colours.) */
SFP_HD f4 body_aabb(f4 pose, f4 vel, f4 shape, float margin, float vel_expand, float dt,
                    const f4* table = nullptr) {
    float k = vel_expand * dt;
    float sx = k * vel.x, sy = k * vel.y;
    f4 r;
    if (table != nullptr && (f2u(shape.w) & 3u) == 3u) {
        /* general polygon with its vertex table at hand (sfp_math.h "shape.w bits"): the box of
         * the rotated vertices.  Without the table — callers that only need SOME bound — the
         * polygon falls through to the box of its bounding circle below. */
        const f4* rec = table + (size_t)(f2u(shape.w) >> 8) * 8u;
        const int count = (int)((f2u(shape.w) >> 2) & 15u);
        float lox = 3.0e38f, loy = 3.0e38f, hix = -3.0e38f, hiy = -3.0e38f;
        for (int j = 0; j < 4; ++j) {
            f4 vv = rec[j];
            if (2 * j < count) {
                float ox = pose.z * vv.x - pose.w * vv.y, oy = pose.w * vv.x + pose.z * vv.y;
                lox = fminf(lox, ox), hix = fmaxf(hix, ox);
                loy = fminf(loy, oy), hiy = fmaxf(hiy, oy);
            }
            if (2 * j + 1 < count) {
                float ox = pose.z * vv.z - pose.w * vv.w, oy = pose.w * vv.z + pose.z * vv.w;
                lox = fminf(lox, ox), hix = fmaxf(hix, ox);
                loy = fminf(loy, oy), hiy = fmaxf(hiy, oy);
            }
        }
        float e = shape.z + margin + k * (fabsf(vel.z) * shape.x); /* shape.x: the core's radius */
        r.x = (pose.x + lox - e) + fminf(sx, 0.0f);
        r.y = (pose.y + loy - e) + fminf(sy, 0.0f);
        r.z = (pose.x + hix + e) + fmaxf(sx, 0.0f);
        r.w = (pose.y + hiy + e) + fmaxf(sy, 0.0f);
        return r;
    }
    float ac = fabsf(pose.z), as = fabsf(pose.w);
    float ex = ac * shape.x + as * shape.y + shape.z;
    float ey = as * shape.x + ac * shape.y + shape.z;
    float bound = sqrtf(shape.x * shape.x + shape.y * shape.y); /* core only: r is spin-invariant */
    float e = margin + k * (fabsf(vel.z) * bound);
    r.x = (pose.x - ex - e) + fminf(sx, 0.0f);
    r.y = (pose.y - ey - e) + fminf(sy, 0.0f);
    r.z = (pose.x + ex + e) + fmaxf(sx, 0.0f);
    r.w = (pose.y + ey + e) + fmaxf(sy, 0.0f);
    return r;
}

SFP_HD int cell_coord(float x, float inv_cell) {
    float f = floorf(x * inv_cell);
    f = fmaxf(f, (float)SFP_CELL_MIN);
    f = fminf(f, (float)SFP_CELL_MAX);
    return (int)f;
}

/* cell range of an AABB.  Returns false — and an EMPTY range — when the box is not finite or
 * covers more than SFP_MAX_CELLS_PER_BODY cells (a runaway body; the caller raises the
 * overflow flag): such a body takes no part in the broadphase. */
#define SFP_MAX_CELLS_PER_BODY 16384
#define SFP_KEY_BITS 12
#define SFP_KEY_SPAN (1 << SFP_KEY_BITS)
SFP_HD bool cell_range(f4 bb, float inv_cell, int* x0, int* y0, int* x1, int* y1) {
    bool finite = (bb.x - bb.x == 0.0f) && (bb.y - bb.y == 0.0f) && (bb.z - bb.z == 0.0f) &&
                  (bb.w - bb.w == 0.0f);
    *x0 = cell_coord(bb.x, inv_cell);
    *y0 = cell_coord(bb.y, inv_cell);
    *x1 = cell_coord(bb.z, inv_cell);
    *y1 = cell_coord(bb.w, inv_cell);
    long long cells = (long long)(*x1 - *x0 + 1) * (long long)(*y1 - *y0 + 1);
    if (!finite || cells > SFP_MAX_CELLS_PER_BODY || cells <= 0 ||
        *x1 - *x0 >= SFP_KEY_SPAN || *y1 - *y0 >= SFP_KEY_SPAN) {
        *x1 = *x0 - 1;
        *y1 = *y0 - 1;
        return false;
    }
    return true;
}

/* 24-bit cell key (three 8-bit radix passes): the cell coordinates modulo SFP_KEY_SPAN = 4096,
 * row-major.  Cells 4096 apart share a key, which only adds candidates that fail the AABB test;
 * a body's own cells never alias (cell_range refuses spans of 4096 cells or more).  The pair
 * rule "reported from the cell holding the larger of the two lower corners" is evaluated
 * modulo the span as well (same_cell). */
SFP_HD uint32_t cell_key(int cx, int cy) {
    return (((uint32_t)cy & (SFP_KEY_SPAN - 1u)) << SFP_KEY_BITS) | ((uint32_t)cx & (SFP_KEY_SPAN - 1u));
}
SFP_HD int key_cx(uint32_t key) { return (int)(key & (SFP_KEY_SPAN - 1u)); }
SFP_HD int key_cy(uint32_t key) { return (int)(key >> SFP_KEY_BITS); }
/* is the cell with true coordinates (x, y) the one a key with (key_cx, key_cy) stands for? */
SFP_HD bool same_cell(int x, int y, int kcx, int kcy) {
    return (((uint32_t)x & (SFP_KEY_SPAN - 1u)) == (uint32_t)kcx) &&
           (((uint32_t)y & (SFP_KEY_SPAN - 1u)) == (uint32_t)kcy);
}

SFP_HD bool aabb_overlap(f4 a, f4 b) {
    return a.x <= b.z && b.x <= a.z && a.y <= b.w && b.y <= a.w;
}

/* layer filter: pair kept iff each body's category is in the other's mask */
SFP_HD bool layers_collide(uint32_t la, uint32_t lb) {
    uint32_t cat_a = la & 0xffffu, mask_a = la >> 16;
    uint32_t cat_b = lb & 0xffffu, mask_b = lb >> 16;
    return (cat_a & mask_b) != 0u && (cat_b & mask_a) != 0u;
}

/* ------------------------------------------------------------------------------------- */
/* Narrowphase (A4): convex cores with <= 4 vertices, each inflated by a radius            */
/* ------------------------------------------------------------------------------------- */
#define SFP_LINEAR_SLOP 0.005f
#define SFP_EPS 1.0e-9f

/* Convex core with 2 (capsule), 4 (rectangle) or up to 8 (general convex polygon) vertices.
 * One runtime-sized type and ONE copy of each routine per vertex capacity serve every shape
 * pair: per-(NA,NB) template instances made the narrowphase 84 KB of SASS, and the batched
 * kernel (three CTAs per SM in different stages) spent most of its stall cycles waiting for
 * instruction fetch.  Loops run to MV with an `i < count` predicate so the vertex arrays stay
 * in registers.  MV = 4 serves capsules and rectangles; MV = 8 is instantiated only for pairs
 * that involve a general polygon. */
#define SFP_POLY_VERTS 8
enum { SHAPE_POLYGON = 3 };
template <int MV>
struct PolyT {
    v2 v[MV]; /* CCW vertices, relative to the frame origin */
    v2 n[MV]; /* outward normal of edge v[i] -> v[(i+1)%count]  */
    int count;
    float r;
};
typedef PolyT<4> Poly;

/* shape.w bits of a general polygon: type (2) | vertex count (4) << 2 | table index << 8;
 * the table holds, per polygon, 4 x f4 of vertices (x0,y0,x1,y1 ...) then 4 x f4 of edge
 * normals, all in the body frame */
SFP_HD uint32_t poly_count(f4 shape) { return (f2u(shape.w) >> 2) & 15u; }
SFP_HD uint32_t poly_index(f4 shape) { return f2u(shape.w) >> 8; }

/* world-space core polygon of a capsule / rectangle / polygon collider, translated by -origin */
template <int MV>
SFP_HD PolyT<MV> make_poly_t(v2 pos, v2 q, f4 shape, v2 origin, const f4* table) {
    PolyT<MV> p;
    v2 c = pos - origin;
    v2 ax = mk(q.x * shape.x, q.y * shape.x);  /* R * (hx, 0) */
    p.r = shape.z;
    const uint32_t type = shape_type(shape);
    if (MV > 4 && type == SHAPE_POLYGON) {
        const f4* rec = table + (size_t)poly_index(shape) * 8u;
        p.count = (int)poly_count(shape);
#pragma unroll
        for (int k = 0; k < MV / 2; ++k) {
            f4 vv = rec[k], nn = rec[4 + k];
            p.v[2 * k] = c + rot(q, mk(vv.x, vv.y));
            p.v[2 * k + 1] = c + rot(q, mk(vv.z, vv.w));
            p.n[2 * k] = rot(q, mk(nn.x, nn.y));
            p.n[2 * k + 1] = rot(q, mk(nn.z, nn.w));
        }
        return p;
    }
    if (type == SHAPE_CAPSULE) {
        p.count = 2;
        p.v[0] = c - ax;
        p.v[1] = c + ax;
        p.n[0] = mk(q.y, -q.x);
        p.n[1] = mk(-q.y, q.x);
#pragma unroll
        for (int k = 2; k < MV; ++k) {
            p.v[k] = p.v[k & 1];
            p.n[k] = p.n[k & 1];
        }
    } else {
        v2 ay = mk(-q.y * shape.y, q.x * shape.y); /* R * (0, hy) */
        p.count = 4;
        p.v[0] = (c - ax) - ay;
        p.v[1] = (c + ax) - ay;
        p.v[2] = (c + ax) + ay;
        p.v[3] = (c - ax) + ay;
        p.n[0] = mk(q.y, -q.x);
        p.n[1] = mk(q.x, q.y);
        p.n[2] = mk(-q.y, q.x);
        p.n[3] = mk(-q.x, -q.y);
#pragma unroll
        for (int k = 4; k < MV; ++k) {
            p.v[k] = p.v[k & 3];
            p.n[k] = p.n[k & 3];
        }
    }
    return p;
}

/* up to two contact points; everything relative to the frame origin (= position of A) */
struct Manifold {
    int count;
    v2 n;      /* unit normal from A to B */
    v2 pa[2];  /* point on A's surface */
    v2 pb[2];  /* point on B's surface */
    float sep[2]; /* dot(pb - pa, n): negative when penetrating */
};

SFP_HD Manifold manifold_none() {
    Manifold m;
    m.count = 0;
    m.n = mk(0.0f, 1.0f);
    m.pa[0] = m.pa[1] = m.pb[0] = m.pb[1] = mk(0.0f, 0.0f);
    m.sep[0] = m.sep[1] = 0.0f;
    return m;
}

/* point core vs point core */
SFP_HD Manifold collide_points(v2 ca, float ra, v2 cb, float rb, float margin) {
    Manifold m = manifold_none();
    v2 d = cb - ca;
    float dist = len(d);
    float sep = dist - (ra + rb);
    if (sep >= margin) return m;
    v2 n = mk(0.0f, 1.0f);
    if (dist > SFP_EPS) n = (1.0f / dist) * d;
    m.count = 1;
    m.n = n;
    m.pa[0] = ca + ra * n;
    m.pb[0] = cb - rb * n;
    m.sep[0] = sep;
    return m;
}

/* edge e of P: its two vertices and its normal (register-resident selection) */
template <int MV>
SFP_HD void poly_edge(const PolyT<MV>& P, int e, v2* a, v2* b, v2* n) {
    int e2 = (e + 1 == P.count) ? 0 : e + 1;
    *a = P.v[0];
    *b = P.v[0];
    *n = P.n[0];
#pragma unroll
    for (int i = 0; i < MV; ++i) {
        if (i == e) {
            *a = P.v[i];
            *n = P.n[i];
        }
        if (i == e2) *b = P.v[i];
    }
}

/* polygon core P vs point core c.  Normal returned points from P to the point. */
template <int MV>
SFP_HD Manifold collide_poly_point(const PolyT<MV>& P, v2 c, float rc, float margin) {
    Manifold m = manifold_none();
    float best = -3.0e38f;
    int bi = 0;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
        if (i < P.count) {
            float s = dot(P.n[i], c - P.v[i]);
            if (s > best) {
                best = s;
                bi = i;
            }
        }
    }
    float rsum = P.r + rc;
    if (best - rsum >= margin) return m;
    v2 v1, v2_, nf;
    poly_edge(P, bi, &v1, &v2_, &nf);
    float u1 = dot(c - v1, v2_ - v1);
    float u2 = dot(c - v2_, v1 - v2_);
    v2 n = nf;
    float dist = best;
    v2 closest;
    if (best > SFP_EPS && u1 < 0.0f) {
        v2 d = c - v1;
        dist = len(d);
        if (dist - rsum >= margin) return m;
        if (dist > SFP_EPS) n = (1.0f / dist) * d;
        closest = v1;
    } else if (best > SFP_EPS && u2 < 0.0f) {
        v2 d = c - v2_;
        dist = len(d);
        if (dist - rsum >= margin) return m;
        if (dist > SFP_EPS) n = (1.0f / dist) * d;
        closest = v2_;
    } else {
        closest = c - best * nf;
    }
    m.count = 1;
    m.n = n;
    m.pa[0] = closest + P.r * n;
    m.pb[0] = c - rc * n;
    m.sep[0] = dist - rsum;
    return m;
}

/* max over edges of P1 of the min over vertices of P2 of the signed distance */
template <int MV>
SFP_HD float max_separation(const PolyT<MV>& P1, const PolyT<MV>& P2, int* edge) {
    float best = -3.0e38f;
    int bi = 0;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
        if (i < P1.count) {
            float si = 3.0e38f;
#pragma unroll
            for (int j = 0; j < MV; ++j) {
                if (j < P2.count) {
                    float sij = dot(P1.n[i], P2.v[j] - P1.v[i]);
                    si = fminf(si, sij);
                }
            }
            if (si > best) {
                best = si;
                bi = i;
            }
        }
    }
    *edge = bi;
    return best;
}

/* closest points of segments p1-q1 and p2-q2 as fractions along each (Ericson 5.1.9) */
SFP_HD void segment_fractions(v2 p1, v2 q1, v2 p2, v2 q2, float* f1, float* f2) {
    v2 d1 = q1 - p1, d2 = q2 - p2, r = p1 - p2;
    float dd1 = dot(d1, d1), dd2 = dot(d2, d2), rd1 = dot(r, d1), rd2 = dot(r, d2);
    float s = 0.0f, t = 0.0f;
    if (dd1 < SFP_EPS || dd2 < SFP_EPS) {
        if (dd1 >= SFP_EPS) s = fminf(fmaxf(-rd1 / dd1, 0.0f), 1.0f);
        else if (dd2 >= SFP_EPS) t = fminf(fmaxf(rd2 / dd2, 0.0f), 1.0f);
    } else {
        float d12 = dot(d1, d2);
        float denom = dd1 * dd2 - d12 * d12;
        if (denom != 0.0f) s = fminf(fmaxf((d12 * rd2 - rd1 * dd2) / denom, 0.0f), 1.0f);
        t = (d12 * s + rd2) / dd2;
        if (t < 0.0f) {
            t = 0.0f;
            s = fminf(fmaxf(-rd1 / dd1, 0.0f), 1.0f);
        } else if (t > 1.0f) {
            t = 1.0f;
            s = fminf(fmaxf((d12 - rd1) / dd1, 0.0f), 1.0f);
        }
    }
    *f1 = s;
    *f2 = t;
}

/* edge of P whose normal is most anti-parallel to n */
template <int MV>
SFP_HD int incident_edge(const PolyT<MV>& P, v2 n) {
    float best = 3.0e38f;
    int bi = 0;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
        if (i < P.count) {
            float d = dot(P.n[i], n);
            if (d < best) {
                best = d;
                bi = i;
            }
        }
    }
    return bi;
}

/* polygon core vs polygon core, both with >= 2 vertices.  SAT picks the reference face; if
 * the cores are disjoint and the closest features are two vertices the result is one point
 * with a vertex-vertex normal, otherwise the incident edge is clipped against the reference
 * face's side planes (1 or 2 points). */
template <int MV>
SFP_HD Manifold collide_polys(const PolyT<MV>& A, const PolyT<MV>& B, float margin) {
    Manifold m = manifold_none();
    int ea = 0, eb = 0;
    float sa = max_separation(A, B, &ea);
    float sb = max_separation(B, A, &eb);
    float rsum = A.r + B.r;
    if (sa - rsum >= margin || sb - rsum >= margin) return m;

    bool flip = sb > sa + 0.1f * SFP_LINEAR_SLOP;
    v2 v11, v12, n1, v21, v22, n2;
    float r1, r2;
    if (flip) {
        poly_edge(B, eb, &v11, &v12, &n1);
        int ie = incident_edge(A, n1);
        poly_edge(A, ie, &v21, &v22, &n2);
        r1 = B.r;
        r2 = A.r;
    } else {
        poly_edge(A, ea, &v11, &v12, &n1);
        int ie = incident_edge(B, n1);
        poly_edge(B, ie, &v21, &v22, &n2);
        r1 = A.r;
        r2 = B.r;
    }
    (void)n2;
    float smax = flip ? sb : sa;

    if (smax > 0.1f * SFP_LINEAR_SLOP) {
        /* cores disjoint: closest features decide */
        float f1, f2;
        segment_fractions(v11, v12, v21, v22, &f1, &f2);
        bool end1 = (f1 == 0.0f) || (f1 == 1.0f);
        bool end2 = (f2 == 0.0f) || (f2 == 1.0f);
        if (end1 && end2) {
            v2 p1 = (f1 == 0.0f) ? v11 : v12;
            v2 p2 = (f2 == 0.0f) ? v21 : v22;
            v2 d = p2 - p1;
            float dist = len(d);
            if (dist - rsum >= margin) return m;
            /* a vertex-vertex pair whose connecting direction is (nearly) the face normal is
             * a face contact seen through rounding; let the clipper handle it */
            v2 n = n1;
            if (dist > SFP_EPS) n = (1.0f / dist) * d;
            if (dot(n, n1) < 0.999f) {
                v2 s1 = p1 + r1 * n; /* on poly1's surface */
                v2 s2 = p2 - r2 * n; /* on poly2's surface */
                m.count = 1;
                m.sep[0] = dist - rsum;
                if (flip) {
                    m.n = neg(n);
                    m.pa[0] = s2;
                    m.pb[0] = s1;
                } else {
                    m.n = n;
                    m.pa[0] = s1;
                    m.pb[0] = s2;
                }
                return m;
            }
        }
    }

    /* clip the incident edge v21->v22 to the slab of the reference edge v11->v12 */
    v2 tangent = mk(-n1.y, n1.x); /* direction v11 -> v12 for a CCW polygon */
    float lower1 = 0.0f;
    float upper1 = dot(v12 - v11, tangent);
    float upper2 = dot(v21 - v11, tangent);
    float lower2 = dot(v22 - v11, tangent);
    float span = upper2 - lower2;
    if (fmaxf(lower2, upper2) < lower1 || fminf(lower2, upper2) > upper1) {
        /* the incident edge lies entirely beside the reference face: nothing to clip (clipping
         * would extrapolate along the incident edge and invent a huge depth).  The contact is
         * between the closest points of the two edges. */
        float f1, f2;
        segment_fractions(v11, v12, v21, v22, &f1, &f2);
        v2 p1 = v11 + f1 * (v12 - v11);
        v2 p2 = v21 + f2 * (v22 - v21);
        v2 d = p2 - p1;
        float dist = len(d);
        v2 n = n1;
        if (dist > SFP_EPS) n = (1.0f / dist) * d;
        if (dot(n, n1) <= 0.0f) {
            n = n1;
            dist = dot(d, n1);
        }
        if (dist - rsum >= margin) return m;
        v2 s1 = p1 + r1 * n, s2 = p2 - r2 * n;
        m.count = 1;
        m.sep[0] = dist - rsum;
        if (flip) {
            m.n = neg(n);
            m.pa[0] = s2;
            m.pb[0] = s1;
        } else {
            m.n = n;
            m.pa[0] = s1;
            m.pb[0] = s2;
        }
        return m;
    }
    v2 vlo = v22, vup = v21;
    if (lower2 < lower1 && span > SFP_EPS) {
        float t = fminf((lower1 - lower2) / span, 1.0f);
        vlo = v22 + t * (v21 - v22);
    }
    if (upper2 > upper1 && span > SFP_EPS) {
        float t = fmaxf((upper1 - lower2) / span, 0.0f);
        vup = v22 + t * (v21 - v22);
    }
    float seplo = dot(vlo - v11, n1);
    float sepup = dot(vup - v11, n1);

    int k = 0;
    v2 cand[2];
    float csep[2];
    cand[0] = vlo;
    csep[0] = seplo;
    cand[1] = vup;
    csep[1] = sepup;
    /* a degenerate incident edge (zero span after clipping) yields one point */
    bool same = (fabsf(vup.x - vlo.x) + fabsf(vup.y - vlo.y)) < 1.0e-6f;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        if (i == 1 && same) continue;
        float s = csep[i] - rsum;
        if (s >= margin) continue;
        v2 s2 = cand[i] - r2 * n1;                 /* on poly2's surface */
        v2 s1 = (cand[i] - csep[i] * n1) + r1 * n1; /* on poly1's surface */
        if (flip) {
            m.pa[k] = s2;
            m.pb[k] = s1;
        } else {
            m.pa[k] = s1;
            m.pb[k] = s2;
        }
        m.sep[k] = s;
        ++k;
    }
    m.count = k;
    m.n = flip ? neg(n1) : n1;
    return m;
}

/* swap the roles of A and B in a manifold */
SFP_HD Manifold manifold_flip(Manifold m) {
    Manifold r = m;
    r.n = neg(m.n);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        r.pa[i] = m.pb[i];
        r.pb[i] = m.pa[i];
    }
    return r;
}

/* dispatch on the two shape types; everything is expressed relative to A's position */
template <int MV>
SFP_HD Manifold collide_cores(v2 xa, v2 qa, f4 shape_a, v2 xb, v2 qb, f4 shape_b, float margin,
                              const f4* table) {
    const uint32_t ta = shape_type(shape_a), tb = shape_type(shape_b);
    v2 cb = xb - xa;
    v2 zero = mk(0.0f, 0.0f);
    if (ta == SHAPE_CIRCLE || tb == SHAPE_CIRCLE) {
        /* one call site: the polygon is whichever body is not the circle */
        const bool a_round = ta == SHAPE_CIRCLE;
        PolyT<MV> P = make_poly_t<MV>(a_round ? xb : xa, a_round ? qb : qa,
                                      a_round ? shape_b : shape_a, xa, table);
        Manifold m = collide_poly_point<MV>(P, a_round ? zero : cb,
                                            a_round ? shape_a.z : shape_b.z, margin);
        return a_round ? manifold_flip(m) : m;
    }
    return collide_polys<MV>(make_poly_t<MV>(xa, qa, shape_a, xa, table),
                             make_poly_t<MV>(xb, qb, shape_b, xa, table), margin);
}

SFP_HD_COLD Manifold collide_cores_polygon(v2 xa, v2 qa, f4 shape_a, v2 xb, v2 qb, f4 shape_b,
                                           float margin, const f4* table) {
    return collide_cores<SFP_POLY_VERTS>(xa, qa, shape_a, xb, qb, shape_b, margin, table);
}

/* POLYS = false compiles the general-polygon path out (kernels of worlds that hold none) */
template <bool POLYS>
SFP_HD Manifold collide(f4 pose_a, f4 shape_a, f4 pose_b, f4 shape_b, float margin,
                        const f4* table) {
    v2 xa = mk(pose_a.x, pose_a.y), qa = mk(pose_a.z, pose_a.w);
    v2 xb = mk(pose_b.x, pose_b.y), qb = mk(pose_b.z, pose_b.w);
    uint32_t ta = shape_type(shape_a), tb = shape_type(shape_b);
    if (ta == SHAPE_CIRCLE && tb == SHAPE_CIRCLE)
        return collide_points(mk(0.0f, 0.0f), shape_a.z, xb - xa, shape_b.z, margin);
    if (POLYS && table != nullptr && (ta == SHAPE_POLYGON || tb == SHAPE_POLYGON))
        return collide_cores_polygon(xa, qa, shape_a, xb, qb, shape_b, margin, table);
    return collide_cores<4>(xa, qa, shape_a, xb, qb, shape_b, margin, table);
}

/* Do two colliders overlap?  The narrowphase with a zero margin: true when it reports a point of
 * negative separation (what the solver would push apart).  A circle of radius 0 is a point. */
SFP_HD bool colliders_overlap(f4 pose_a, f4 shape_a, f4 pose_b, f4 shape_b, const f4* table) {
    float ba = sqrtf(shape_a.x * shape_a.x + shape_a.y * shape_a.y) + shape_a.z;
    float bb = sqrtf(shape_b.x * shape_b.x + shape_b.y * shape_b.y) + shape_b.z;
    float dx = pose_b.x - pose_a.x, dy = pose_b.y - pose_a.y, reach = ba + bb;
    if (dx * dx + dy * dy >= reach * reach) return false;
    return collide<true>(pose_a, shape_a, pose_b, shape_b, 0.0f, table).count > 0;
}

/* ------------------------------------------------------------------------------------- */
/* Contact record (per candidate pair; 5 x 16 B)                                         */
/* ------------------------------------------------------------------------------------- */
struct ContactRec {
    f4 a;  /* n.x, n.y, mu_s, mu_d                              */
    f4 b;  /* e, bits(count), vn_pre[0], vn_pre[1]              */
    f4 p0; /* la0.x, la0.y, lb0.x, lb0.y   (body-frame anchors) */
    f4 p1; /* la1.x, la1.y, lb1.x, lb1.y                        */
    f4 s;  /* sep0, sep1 at narrowphase time, 0, 0              */
};

/* 0..3: which narrowphase routine a pair runs — both circles / one circle / neither / a general
 * polygon involved.  Pairs are bucketed by (colour, this class) so that a warp runs one routine
 * instead of several; colour (6 bits) and class (SFP_TYPE_BITS) make an 8-bit sort key, one
 * radix pass. */
#define SFP_TYPE_BITS 2
SFP_HD uint32_t pair_type_key(f4 shape_a, f4 shape_b) {
    uint32_t ta = shape_type(shape_a), tb = shape_type(shape_b);
    if (ta == 3u || tb == 3u) return 3u;
    return (ta != 0u ? 1u : 0u) + (tb != 0u ? 1u : 0u);
}

/* bounding-circle reject: most candidate pairs of a frame are not touching in a given substep,
 * and this costs a dozen flops instead of a SAT.  The reach of a pair (sum of the two bounding
 * radii plus the margin) does not change within a frame: the device computes it once per frame
 * per pair (narrowphase_reach) and tests with narrowphase_far_reach; same operations, same bits. */
SFP_HD float narrowphase_reach(f4 shape_a, f4 shape_b, float margin) {
    float ba = sqrtf(shape_a.x * shape_a.x + shape_a.y * shape_a.y) + shape_a.z;
    float bb = sqrtf(shape_b.x * shape_b.x + shape_b.y * shape_b.y) + shape_b.z;
    return (ba + bb) + margin;
}
SFP_HD bool narrowphase_far_reach(f4 pose_a, f4 pose_b, float reach) {
    float dx = pose_b.x - pose_a.x, dy = pose_b.y - pose_a.y;
    return dx * dx + dy * dy >= reach * reach;
}
SFP_HD bool narrowphase_far(f4 pose_a, f4 shape_a, f4 pose_b, f4 shape_b, float margin) {
    return narrowphase_far_reach(pose_a, pose_b, narrowphase_reach(shape_a, shape_b, margin));
}
SFP_HD ContactRec empty_contact() {
    ContactRec c;
    f4 z;
    z.x = z.y = z.z = z.w = 0.0f;
    c.a = c.b = c.p0 = c.p1 = c.s = z;
    return c;
}

/* material of a pair (mu_s, mu_d, e): mean friction, larger restitution */
SFP_HD void pair_material(f4 mat_a, f4 mat_b, float* mu_s, float* mu_d, float* e) {
    *mu_s = 0.5f * (mat_a.x + mat_b.x);
    *mu_d = 0.5f * (mat_a.y + mat_b.y);
    *e = fmaxf(mat_a.z, mat_b.z);
}

/* the contact record of a pair that passed narrowphase_far(), from the pair's material */
template <bool POLYS>
SFP_HD ContactRec narrowphase_near_mat(f4 pose_a, f4 vel_a, f4 shape_a, f4 pose_b, f4 vel_b,
                                       f4 shape_b, float mu_s, float mu_d, float e, float margin,
                                       const f4* table) {
    ContactRec c;
    Manifold m = collide<POLYS>(pose_a, shape_a, pose_b, shape_b, margin, table);
    v2 qa = mk(pose_a.z, pose_a.w), qb = mk(pose_b.z, pose_b.w);
    v2 dab = mk(pose_b.x - pose_a.x, pose_b.y - pose_a.y);
    c.a.x = m.n.x;
    c.a.y = m.n.y;
    c.a.z = mu_s;
    c.a.w = mu_d;
    c.b.x = e;
    c.b.y = u2f((uint32_t)m.count);
    float vn[2];
    f4 pp[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        v2 ra = m.pa[k];
        v2 rb = m.pb[k] - dab;
        v2 va = mk(vel_a.x, vel_a.y) + crossz(vel_a.z, ra);
        v2 vb = mk(vel_b.x, vel_b.y) + crossz(vel_b.z, rb);
        vn[k] = dot(vb - va, m.n);
        v2 la = rot_inv(qa, ra);
        v2 lb = rot_inv(qb, rb);
        pp[k].x = la.x;
        pp[k].y = la.y;
        pp[k].z = lb.x;
        pp[k].w = lb.y;
    }
    c.b.z = vn[0];
    c.b.w = vn[1];
    c.p0 = pp[0];
    c.p1 = pp[1];
    c.s.x = m.sep[0];
    c.s.y = m.sep[1];
    c.s.z = 0.0f;
    c.s.w = 0.0f;
    return c;
}

template <bool POLYS>
SFP_HD ContactRec narrowphase_near(f4 pose_a, f4 vel_a, f4 shape_a, f4 mat_a, f4 pose_b, f4 vel_b,
                                   f4 shape_b, f4 mat_b, float margin, const f4* table) {
    float mu_s, mu_d, e;
    pair_material(mat_a, mat_b, &mu_s, &mu_d, &e);
    return narrowphase_near_mat<POLYS>(pose_a, vel_a, shape_a, pose_b, vel_b, shape_b, mu_s, mu_d, e,
                                       margin, table);
}

template <bool POLYS>
SFP_HD ContactRec narrowphase(f4 pose_a, f4 vel_a, f4 shape_a, f4 mat_a, f4 pose_b, f4 vel_b,
                              f4 shape_b, f4 mat_b, float margin, const f4* table) {
    if (narrowphase_far(pose_a, shape_a, pose_b, shape_b, margin)) return empty_contact();
    return narrowphase_near<POLYS>(pose_a, vel_a, shape_a, mat_a, pose_b, vel_b, shape_b, mat_b, margin,
                                   table);
}

/* contact event of one pair from its records (layout of SfpContactEvent in sfp.h) */
struct EventRec {
    uint32_t a, b;
    float nx, ny, px, py, impulse;
    uint32_t points;
};
SFP_HD bool pair_event(uint32_t a, uint32_t b, f4 ca, f4 cb, f4 p0, f4 p1, f4 lam, f4 pose_a,
                       f4 pose_b, float inv_h, EventRec* e) {
    uint32_t cnt = f2u(cb.y);
    bool on0 = cnt > 0u && lam.x > 0.0f, on1 = cnt > 1u && lam.z > 0.0f;
    if (!on0 && !on1) return false;
    f4 pp = on0 ? p0 : p1;
    v2 wa = mk(pose_a.x, pose_a.y) + rot(mk(pose_a.z, pose_a.w), mk(pp.x, pp.y));
    v2 wb = mk(pose_b.x, pose_b.y) + rot(mk(pose_b.z, pose_b.w), mk(pp.z, pp.w));
    e->a = a;
    e->b = b;
    e->nx = ca.x;
    e->ny = ca.y;
    e->px = 0.5f * (wa.x + wb.x);
    e->py = 0.5f * (wa.y + wb.y);
    e->impulse = ((on0 ? lam.x : 0.0f) + (on1 ? lam.z : 0.0f)) * inv_h;
    e->points = (on0 ? 1u : 0u) + (on1 ? 1u : 0u);
    return true;
}

/* ------------------------------------------------------------------------------------- */
/* Integrate (A3) and velocity derivation (A7 i)                                         */
/* ------------------------------------------------------------------------------------- */
SFP_HD void integrate_body(f4* pose, f4* prev, f4* vel, f2 minv, float gx, float gy, float h) {
    prev->x = 0.0f;
    prev->y = 0.0f;
    prev->z = pose->z;
    prev->w = pose->w;
    if (minv.x > 0.0f) {
        vel->x = vel->x + h * gx;
        vel->y = vel->y + h * gy;
    }
    if (minv.x > 0.0f || minv.y > 0.0f) {
        prev->x = h * vel->x;
        prev->y = h * vel->y;
        pose->x = pose->x + prev->x;
        pose->y = pose->y + prev->y;
        v2 q = rot_advance(mk(pose->z, pose->w), h * vel->z);
        pose->z = q.x;
        pose->w = q.y;
    }
}

SFP_HD f4 derive_velocity(f4 pose, f4 prev, f4 vel, f2 minv, float inv_h) {
    f4 v = vel;
    if (minv.x > 0.0f || minv.y > 0.0f) {
        v.x = prev.x * inv_h;
        v.y = prev.y * inv_h;
        v.z = rot_delta_tan(mk(pose.z, pose.w), mk(prev.z, prev.w)) * inv_h;
    }
    return v;
}

/* ------------------------------------------------------------------------------------- */
/* XPBD position projection (A5, A6)                                                     */
/* ------------------------------------------------------------------------------------- */
struct BodyPos {
    v2 x, q; /* current pose */
    v2 d;    /* displacement since the substep started (kept separately: see the state notes) */
    v2 qp;   /* rotation at substep start */
    float im, ii;
};

SFP_HD float gen_inv_mass(float im, float ii, v2 r, v2 n) {
    float rn = cross(r, n);
    return im + ii * (rn * rn);
}

/* apply the positional impulse +P at arm rb to B and -P at arm ra to A */
SFP_HD void apply_pos(BodyPos* A, BodyPos* B, v2 ra, v2 rb, v2 P) {
    v2 ua = A->im * P, ub = B->im * P;
    A->x = A->x - ua;
    A->d = A->d - ua;
    B->x = B->x + ub;
    B->d = B->d + ub;
    float da = -(A->ii * cross(ra, P));
    float db = B->ii * cross(rb, P);
    if (A->ii > 0.0f) A->q = rot_nudge(A->q, da);
    if (B->ii > 0.0f) B->q = rot_nudge(B->q, db);
}

/* 2x2 (or 1x1) non-penetration solve of a manifold at the bodies' current poses; adds the
 * multipliers it applied to *l0, *l1 */
SFP_HD void solve_manifold_normal(BodyPos* A, BodyPos* B, v2 n, uint32_t count, v2 la0, v2 lb0,
                                  v2 la1, v2 lb1, float* l0_acc, float* l1_acc) {
    v2 ra0 = rot(A->q, la0), rb0 = rot(B->q, lb0);
    float d0 = -dot((B->x + rb0) - (A->x + ra0), n); /* penetration depth, > 0 when overlapping */
    float ka0 = cross(ra0, n), kb0 = cross(rb0, n);
    float ims = A->im + B->im;
    float a00 = ims + A->ii * (ka0 * ka0) + B->ii * (kb0 * kb0);
    if (a00 <= 0.0f) return;
    float l0 = 0.0f, l1 = 0.0f;
    float ka1 = 0.0f, kb1 = 0.0f;
    if (count > 1u) {
        v2 ra1 = rot(A->q, la1), rb1 = rot(B->q, lb1);
        float d1 = -dot((B->x + rb1) - (A->x + ra1), n);
        ka1 = cross(ra1, n);
        kb1 = cross(rb1, n);
        float a11 = ims + A->ii * (ka1 * ka1) + B->ii * (kb1 * kb1);
        float a01 = ims + A->ii * (ka0 * ka1) + B->ii * (kb0 * kb1);
        float det = a00 * a11 - a01 * a01;
        if (det > 1.0e-3f * (a00 * a11)) {
            /* both active */
            float inv = 1.0f / det; /* one division on the chain, not two */
            float x0 = (a11 * d0 - a01 * d1) * inv;
            float x1 = (a00 * d1 - a01 * d0) * inv;
            if (x0 >= 0.0f && x1 >= 0.0f) {
                l0 = x0;
                l1 = x1;
            } else {
                /* only point 0 / only point 1 active, if that leaves the other one clear */
                float y0 = d0 / a00, y1 = d1 / a11;
                if (y0 >= 0.0f && a01 * y0 - d1 >= 0.0f) l0 = y0;
                else if (y1 >= 0.0f && a01 * y1 - d0 >= 0.0f) l1 = y1;
                else l0 = fmaxf(y0, 0.0f);
            }
        } else {
            /* (nearly) coincident points: treat the deeper one */
            if (d1 > d0) l1 = fmaxf(d1 / a11, 0.0f);
            else l0 = fmaxf(d0 / a00, 0.0f);
        }
    } else {
        l0 = fmaxf(d0 / a00, 0.0f);
    }
    if (l0 > 0.0f || l1 > 0.0f) {
        v2 P = (l0 + l1) * n;
        v2 ua = A->im * P, ub = B->im * P;
        A->x = A->x - ua;
        A->d = A->d - ua;
        B->x = B->x + ub;
        B->d = B->d + ub;
        float da = -(A->ii * (ka0 * l0 + ka1 * l1));
        float db = B->ii * (kb0 * l0 + kb1 * l1);
        if (A->ii > 0.0f) A->q = rot_nudge(A->q, da);
        if (B->ii > 0.0f) B->q = rot_nudge(B->q, db);
    }
    *l0_acc = *l0_acc + l0;
    *l1_acc = *l1_acc + l1;
}

/* Position pass of one manifold (1 or 2 points sharing one normal).
 *
 * Non-penetration: two points are solved TOGETHER as a 2x2 complementarity problem; solving
 * them one after the other leaves a net rotation every substep.
 *
 * Static friction (Müller et al. §3.5 cancels the tangential drift of the contact points at
 * position level if that takes less than mu_s * lambda_n).  Measured here: with one
 * Gauss-Seidel sweep per substep, applying that correction through the contact arms (i.e. with
 * its rotational part) makes resting towers of 3+ boxes rock with growing amplitude until
 * they topple, in every constraint order tried.  What is kept at position level is the part
 * that is needed against creep and is harmless: for a FACE contact (two active points — no
 * rolling possible) the tangential drift of the two centres of mass relative to each other
 * over the substep is cancelled by pure translation when inside the friction cone.  Everything
 * else about friction (stick/slip of single-point contacts, rolling, the rotational part) is
 * handled at velocity level in solve_manifold_velocity().
 * Returns (ln0, lt0, ln1, lt1). */
SFP_HD f4 solve_contact_manifold(BodyPos* A, BodyPos* B, v2 n, uint32_t count, v2 la0, v2 lb0,
                                 v2 la1, v2 lb1, float mu_s) {
    f4 lam;
    lam.x = lam.y = lam.z = lam.w = 0.0f;
    if (count == 0u) return lam;
    solve_manifold_normal(A, B, n, count, la0, lb0, la1, lb1, &lam.x, &lam.z);
    if (count > 1u && lam.x > 0.0f && lam.z > 0.0f) {
        v2 dp = B->d - A->d;
        v2 dpt = dp - dot(dp, n) * n;
        float wl = A->im + B->im;
#ifdef SFP_EXACT_NUDGE
        float tl = len(dpt);
        if (tl > SFP_EPS && wl > 0.0f) {
            float dlt = tl / wl;
            if (dlt < mu_s * (lam.x + lam.z)) {
                v2 P = (-dlt / tl) * dpt;
                v2 ua = A->im * P, ub = B->im * P;
                A->x = A->x - ua;
                A->d = A->d - ua;
                B->x = B->x + ub;
                B->d = B->d + ub;
                lam.y = 0.5f * dlt;
                lam.w = 0.5f * dlt;
            }
        }
#else
        /* the multiplier is |dpt| / wl and the correction P = -dpt / wl: the bodies only need
         * 1 / wl (known before the normal solve) and the cone test compares squares, so no
         * square root or second division stands between the normal solve and the body update */
        float tl2 = dot(dpt, dpt);
        float lim = mu_s * (lam.x + lam.z) * wl;
        if (tl2 > SFP_EPS * SFP_EPS && wl > 0.0f && tl2 < lim * lim) {
            v2 P = (-1.0f / wl) * dpt;
            v2 ua = A->im * P, ub = B->im * P;
            A->x = A->x - ua;
            A->d = A->d - ua;
            B->x = B->x + ub;
            B->d = B->d + ub;
            float dlt = sqrtf(tl2) / wl;
            lam.y = 0.5f * dlt;
            lam.w = 0.5f * dlt;
        }
#endif
    }
    return lam;
}

/* distance (bilateral) / rope (unilateral) joint, one XPBD iteration with lambda = 0 */
SFP_HD void solve_joint(BodyPos* A, BodyPos* B, v2 la, v2 lb, float rest, float compliance,
                        uint32_t kind, float inv_h2) {
    v2 ra = rot(A->q, la), rb = rot(B->q, lb);
    v2 d = (B->x + rb) - (A->x + ra);
    float l = len(d);
    float c = l - rest;
    if (l <= SFP_EPS) return;
    if (kind == 1u && c <= 0.0f) return;
    v2 n = (1.0f / l) * d;
    float at = compliance * inv_h2;
    float w = gen_inv_mass(A->im, A->ii, ra, n) + gen_inv_mass(B->im, B->ii, rb, n) + at;
    if (w <= 0.0f) return;
    float dl = -c / w;
    apply_pos(A, B, ra, rb, dl * n);
}

/* ------------------------------------------------------------------------------------- */
/* Velocity pass (A7 ii): dynamic friction, then restitution (Müller et al. §3.6)        */
/* ------------------------------------------------------------------------------------- */
struct BodyVel {
    v2 v;
    float w;
    float im, ii;
};

/* apply the impulse +p at arm rb to B and -p at arm ra to A */
SFP_HD void apply_impulse(BodyVel* A, BodyVel* B, v2 ra, v2 rb, v2 p) {
    A->v = A->v - A->im * p;
    B->v = B->v + B->im * p;
    A->w = A->w - A->ii * cross(ra, p);
    B->w = B->w + B->ii * cross(rb, p);
}

/* Coulomb friction impulse for a required stopping impulse p_stop and normal multiplier
 * lambda_n (f_n = lambda_n / h^2, so the limit impulse is mu * lambda_n / h): stick if inside
 * the static cone, otherwise slide with the dynamic coefficient.  (The paper's eq. 30 clamps
 * the velocity change by that impulse-valued quantity; clamping the impulse is the
 * dimensionally consistent reading and gives the textbook deceleration mu_d * g.) */
SFP_HD float coulomb(float p_stop, float lambda_n, float mu_s, float mu_d, float h) {
    float lim_s = mu_s * (lambda_n / h), lim_d = mu_d * (lambda_n / h);
    float a = fabsf(p_stop);
    if (a <= lim_s) return p_stop;
    float m = fminf(a, lim_d);
    return p_stop < 0.0f ? -m : m;
}

/* Velocity pass of one manifold (Müller et al. §3.6): friction, and restitution — the normal
 * velocity of every active point is driven to max(-e * vn_pre, 0), with e = 0 below the resting
 * threshold.
 *   two active points: normals and ONE tangential unknown at the midpoint are solved together
 *     (3x3 linear system); if the tangential impulse leaves the friction cone it is clamped
 *     and the normals are re-solved (2x2).  Sequential point-by-point solving lets the
 *     rotational cross-talk eat most of the friction impulse (measured: a box that should
 *     hold on a 20 degree incline crept at 17 mm/s);
 *   one active point: friction, then the normal, sequentially. */
SFP_HD void solve_manifold_velocity(BodyVel* A, BodyVel* B, v2 qa, v2 qb, v2 n, uint32_t count,
                                    v2 la0, v2 lb0, v2 la1, v2 lb1, f4 lam, float vn_pre0,
                                    float vn_pre1, float mu_s, float mu_d, float e, float h,
                                    float rest_threshold) {
    bool on0 = lam.x > 0.0f, on1 = count > 1u && lam.z > 0.0f;
    if (!on0 && !on1) return;
    v2 t = mk(-n.y, n.x);
    float ims = A->im + B->im;
    float tg0 = fmaxf(-((fabsf(vn_pre0) <= rest_threshold) ? 0.0f : e) * vn_pre0, 0.0f);
    float tg1 = fmaxf(-((fabsf(vn_pre1) <= rest_threshold) ? 0.0f : e) * vn_pre1, 0.0f);
    if (on0 && on1) {
        v2 ra0 = rot(qa, la0), rb0 = rot(qb, lb0), ra1 = rot(qa, la1), rb1 = rot(qb, lb1);
        v2 ram = 0.5f * (ra0 + ra1), rbm = 0.5f * (rb0 + rb1);
        float ka0 = cross(ra0, n), kb0 = cross(rb0, n), ka1 = cross(ra1, n), kb1 = cross(rb1, n);
        float kat = cross(ram, t), kbt = cross(rbm, t);
        float k00 = ims + A->ii * (ka0 * ka0) + B->ii * (kb0 * kb0);
        float k11 = ims + A->ii * (ka1 * ka1) + B->ii * (kb1 * kb1);
        float k01 = ims + A->ii * (ka0 * ka1) + B->ii * (kb0 * kb1);
        float k0t = A->ii * (ka0 * kat) + B->ii * (kb0 * kbt);
        float k1t = A->ii * (ka1 * kat) + B->ii * (kb1 * kbt);
        float ktt = ims + A->ii * (kat * kat) + B->ii * (kbt * kbt);
        float b0 = tg0 - dot((B->v + crossz(B->w, rb0)) - (A->v + crossz(A->w, ra0)), n);
        float b1 = tg1 - dot((B->v + crossz(B->w, rb1)) - (A->v + crossz(A->w, ra1)), n);
        float bt = -dot((B->v + crossz(B->w, rbm)) - (A->v + crossz(A->w, ram)), t);
        float det2 = k00 * k11 - k01 * k01;
        float p0 = 0.0f, p1 = 0.0f, pt = 0.0f;
        if (det2 > 1.0e-3f * (k00 * k11)) {
            /* 3x3 by cofactors; the two reciprocals are independent of each other and of the
             * right-hand sides, so no division waits for another one */
            float inv2 = 1.0f / det2;
            float c00 = k11 * ktt - k1t * k1t, c01 = k0t * k1t - k01 * ktt;
            float c02 = k01 * k1t - k0t * k11;
            float det3 = k00 * c00 + k01 * c01 + k0t * c02;
            if (det3 > 1.0e-4f * (k00 * k11 * ktt)) {
                float inv3 = 1.0f / det3;
                float c11 = k00 * ktt - k0t * k0t, c12 = k01 * k0t - k00 * k1t;
                float c22 = det2;
                float ptf = (c02 * b0 + c12 * b1 + c22 * bt) * inv3;
                pt = coulomb(ptf, lam.x + lam.z, mu_s, mu_d, h);
                if (pt == ptf) {
                    p0 = (c00 * b0 + c01 * b1 + c02 * bt) * inv3;
                    p1 = (c01 * b0 + c11 * b1 + c12 * bt) * inv3;
                } else {
                    float r0 = b0 - k0t * pt, r1 = b1 - k1t * pt;
                    p0 = (k11 * r0 - k01 * r1) * inv2;
                    p1 = (k00 * r1 - k01 * r0) * inv2;
                }
            } else {
                p0 = (k11 * b0 - k01 * b1) * inv2;
                p1 = (k00 * b1 - k01 * b0) * inv2;
            }
        } else if (k00 > 0.0f) {
            p0 = b0 / k00;
        }
        v2 P = (p0 + p1) * n + pt * t;
        A->v = A->v - A->im * P;
        B->v = B->v + B->im * P;
        A->w = A->w - A->ii * (ka0 * p0 + ka1 * p1 + kat * pt);
        B->w = B->w + B->ii * (kb0 * p0 + kb1 * p1 + kbt * pt);
        return;
    }
    /* single active point */
    v2 la = on0 ? la0 : la1, lb = on0 ? lb0 : lb1;
    float ln = on0 ? lam.x : lam.z, tg = on0 ? tg0 : tg1;
    v2 ra = rot(qa, la), rb = rot(qb, lb);
    float kat = cross(ra, t), kbt = cross(rb, t);
    float ktt = ims + A->ii * (kat * kat) + B->ii * (kbt * kbt);
    if (ktt > 0.0f) {
        float vt = dot((B->v + crossz(B->w, rb)) - (A->v + crossz(A->w, ra)), t);
        float pt = coulomb(-vt / ktt, ln, mu_s, mu_d, h);
        apply_impulse(A, B, ra, rb, pt * t);
    }
    float kan = cross(ra, n), kbn = cross(rb, n);
    float knn = ims + A->ii * (kan * kan) + B->ii * (kbn * kbn);
    if (knn > 0.0f) {
        float vn = dot((B->v + crossz(B->w, rb)) - (A->v + crossz(A->w, ra)), n);
        apply_impulse(A, B, ra, rb, ((tg - vn) / knn) * n);
    }
}

/* ------------------------------------------------------------------------------------- */
/* Spatial queries: ray vs collider, exact for the rounded shapes     */
/* ------------------------------------------------------------------------------------- */
/* layout of SfpRay / SfpRayHit in sfp.h */
struct RayRec {
    float ox, oy, dx, dy; /* origin, direction (any non-zero length; t is in units of it) */
    float max_t;
    uint32_t mask;        /* category bits the ray can hit */
    uint32_t pad[2];
};
struct RayHitRec {
    uint32_t body; /* 0xffffffff: no hit */
    float t, nx, ny;
};

/* ray o + t d vs the disc of radius r at c: smallest t in [0, tmax] where the ray ENTERS */
SFP_HD bool ray_disc(v2 o, v2 d, v2 c, float r, float tmax, float* t, v2* n) {
    v2 m = o - c;
    float a = dot(d, d), b = dot(m, d), cc = dot(m, m) - r * r;
    if (cc <= 0.0f) return false; /* origin inside: handled by the caller's inside test */
    if (b >= 0.0f) return false;  /* pointing away */
    float disc = b * b - a * cc;
    if (disc < 0.0f) return false;
    float tt = (-b - sqrtf(disc)) / a;
    if (tt < 0.0f || tt > tmax) return false;
    v2 hp = m + tt * d;
    float l = len(hp);
    *t = tt;
    *n = (l > SFP_EPS) ? (1.0f / l) * hp : neg((1.0f / sqrtf(a)) * d);
    return true;
}

/* ray vs the segment p-q seen from its outward side `nrm` (one offset edge of a rounded shape) */
SFP_HD bool ray_edge(v2 o, v2 d, v2 p, v2 q, v2 nrm, float tmax, float* t) {
    float dn = dot(d, nrm);
    if (dn >= 0.0f) return false; /* parallel or leaving */
    float tt = dot(p - o, nrm) / dn;
    if (tt < 0.0f || tt > tmax) return false;
    v2 hp = o + tt * d;
    v2 e = q - p;
    float u = dot(hp - p, e);
    if (u < 0.0f || u > dot(e, e)) return false;
    *t = tt;
    return true;
}

/* closest hit of a ray with an inflated convex core */
template <int MV>
SFP_HD bool ray_poly(v2 o, v2 d, float tmax, const PolyT<MV>& P, v2 back, float* t_out, v2* n_out) {
    const float r = P.r;
    /* origin inside the inflated core? */
    float smax = -3.0e38f, dmin2 = 3.0e38f;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
        if (i < P.count) {
            v2 a, b, nn;
            poly_edge<MV>(P, i, &a, &b, &nn);
            smax = fmaxf(smax, dot(nn, o - a));
            v2 e = b - a;
            float ee = dot(e, e);
            float u = (ee > 0.0f) ? fminf(fmaxf(dot(o - a, e) / ee, 0.0f), 1.0f) : 0.0f;
            v2 w = o - (a + u * e);
            dmin2 = fminf(dmin2, dot(w, w));
        }
    }
    bool inside_core = (P.count > 2) && smax <= 0.0f;
    if (inside_core || dmin2 <= r * r) {
        *t_out = 0.0f;
        *n_out = back;
        return true;
    }
    bool hit = false;
    float best = tmax;
    v2 bn = back;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
        if (i < P.count) {
            v2 a, b, nn;
            poly_edge<MV>(P, i, &a, &b, &nn);
            float tt;
            v2 off = r * nn;
            if (ray_edge(o, d, a + off, b + off, nn, best, &tt)) {
                hit = true;
                best = tt;
                bn = nn;
            }
            v2 cn;
            if (r > 0.0f && ray_disc(o, d, a, r, best, &tt, &cn)) {
                hit = true;
                best = tt;
                bn = cn;
            }
        }
    }
    *t_out = best;
    *n_out = bn;
    return hit;
}

SFP_HD_COLD bool ray_polygon(v2 o, v2 d, float tmax, v2 x, v2 q, f4 shape, const f4* table, v2 back,
                             float* t_out, v2* n_out) {
    return ray_poly<SFP_POLY_VERTS>(
        o, d, tmax, make_poly_t<SFP_POLY_VERTS>(x, q, shape, mk(0.0f, 0.0f), table), back, t_out,
        n_out);
}

/* closest hit of a ray with one collider; true if hit within [0, tmax].
 * A ray that starts far from the collider (kilometres away: the quadratic of a disc then loses
 * every digit to cancellation, |o - c|^2 against r^2) is first moved along itself to a point two
 * bounding radii in front of the collider's centre; the hit parameter is the sum.  Near rays
 * — every ray whose origin is within four bounding radii, measured along the ray — take the
 * direct route, bit for bit as before. */
SFP_HD bool ray_collider(v2 o, v2 d, float tmax, f4 pose, f4 shape, const f4* table, float* t_out,
                         v2* n_out);
SFP_HD bool ray_collider_near(v2 o, v2 d, float tmax, f4 pose, f4 shape, const f4* table,
                              float* t_out, v2* n_out);
SFP_HD bool ray_collider(v2 o, v2 d, float tmax, f4 pose, f4 shape, const f4* table, float* t_out,
                         v2* n_out) {
    float dd = dot(d, d);
    if (!(dd > 0.0f)) return false;
    float bound = sqrtf(shape.x * shape.x + shape.y * shape.y) + shape.z;
    float reach = bound / sqrtf(dd);                      /* the bounding radius in units of t */
    float tc = dot(mk(pose.x, pose.y) - o, d) / dd;       /* closest approach to the centre */
    if (tc > 4.0f * reach && reach > 0.0f) {
        float tb = tc - 2.0f * reach;
        if (tb >= tmax) return false;
        float t;
        if (!ray_collider_near(o + tb * d, d, tmax - tb, pose, shape, table, &t, n_out)) return false;
        *t_out = tb + t;
        return true;
    }
    return ray_collider_near(o, d, tmax, pose, shape, table, t_out, n_out);
}
SFP_HD bool ray_collider_near(v2 o, v2 d, float tmax, f4 pose, f4 shape, const f4* table,
                              float* t_out, v2* n_out) {
    v2 x = mk(pose.x, pose.y), q = mk(pose.z, pose.w);
    float r = shape.z;
    uint32_t type = shape_type(shape);
    float dl = sqrtf(dot(d, d));
    if (!(dl > 0.0f)) return false;
    v2 back = neg((1.0f / dl) * d);
    if (type == SHAPE_CIRCLE) {
        v2 m = o - x;
        if (dot(m, m) <= r * r) {
            *t_out = 0.0f;
            *n_out = back;
            return true;
        }
        return ray_disc(o, d, x, r, tmax, t_out, n_out);
    }
    if (type == SHAPE_POLYGON) return ray_polygon(o, d, tmax, x, q, shape, table, back, t_out, n_out);
    return ray_poly<4>(o, d, tmax, make_poly_t<4>(x, q, shape, mk(0.0f, 0.0f), table), back, t_out,
                       n_out);
}

/* tight AABB of a collider at its current pose (no margin, no sweep) */
SFP_HD f4 tight_aabb(f4 pose, f4 shape, const f4* table = nullptr) {
    f4 zero;
    zero.x = zero.y = zero.z = zero.w = 0.0f;
    return body_aabb(pose, zero, shape, 0.0f, 0.0f, 0.0f, table);
}

/* ------------------------------------------------------------------------------------- */
/* Colouring (A8) — integer only                                                         */
/* ------------------------------------------------------------------------------------- */
#define SFP_MAX_COLOURS 64
#define SFP_MAX_BODIES (1u << 22) /* body ids are packed into 22 bits of the priority key */
#define SFP_NO_CLAIM 0xffffffffffffffffull
/* Colours 0..62 are proper colours (a body's mask bit is set, no two constraints of a colour
 * share a dynamic body).  A constraint that wins its bodies but finds all 63 taken — a body with
 * more than 63 partners, e.g. a long dynamic plank carrying a crowd — joins the SERIAL group,
 * "colour" 63: its members may share bodies, so they are solved one after another, after every
 * proper colour, joints in upload order and then pairs in (a, b) order.  Slow but correct;
 * at most SFP_SERIAL_MAX members (beyond that the step reports SFP_E_CAPACITY). */
#define SFP_SERIAL_COLOUR 63
#define SFP_SERIAL_MAX 1024u
/* Solve order inside a substep: the colours with more than SFP_TAIL_ITEMS constraints in
 * ascending colour index, then the remaining (small) colours in ascending index.  (The CUDA
 * path runs every small colour inside one grid-wide phase; fixing the order here keeps the
 * oracle and the device on the same Gauss-Seidel sequence.) */
#define SFP_TAIL_ITEMS 512u
/* Colours are carried over from the previous frame for the pairs that persist; every
 * SFP_RECOLOUR_PERIOD-th colouring starts from scratch so that stale high colours cannot pile
 * up.  The count restarts at every body / joint upload. */
#define SFP_RECOLOUR_PERIOD 32u
/* previous-frame colour table: open-addressing hash of (a, b) -> colour, one u64 per slot:
 * (colour + 1) << 44 | a << 22 | b;  0 = empty */
#define SFP_CTAB_KEYMASK ((1ull << 44) - 1ull)

SFP_HD uint32_t hash32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

/* unique, order-independent 64-bit priority of a constraint: 19-bit hash | kind | id(44) */
SFP_HD uint64_t pair_priority(uint32_t a, uint32_t b) {
    uint32_t h = hash32(a * 0x9e3779b9u + hash32(b));
    uint64_t id = ((uint64_t)a << 22) | (uint64_t)b;
    return ((uint64_t)(h >> 13) << 45) | id;
}
SFP_HD uint64_t joint_priority(uint32_t j) {
    uint32_t h = hash32(j ^ 0x5bd1e995u);
    return ((uint64_t)(h >> 13) << 45) | (1ull << 44) | (uint64_t)j;
}

SFP_HD uint64_t ctab_key(uint32_t a, uint32_t b) { return ((uint64_t)a << 22) | (uint64_t)b; }
SFP_HD uint64_t ctab_pack(uint32_t a, uint32_t b, uint32_t colour) {
    return ((uint64_t)(colour + 1u) << 44) | ctab_key(a, b);
}
SFP_HD uint32_t ctab_hash(uint32_t a, uint32_t b) { return hash32(a * 0x9e3779b9u ^ hash32(b + 0x7f4a7c15u)); }

/* ------------------------------------------------------------------------------------- */
/* Tiles (solve order) — integer only                                                    */
/* ------------------------------------------------------------------------------------- */
/* The dynamic bodies of a world are cut into TILES: spatially compact groups of SFP_TILE_MIN ..
 * SFP_TILE_MAX bodies (consecutive runs of the bodies sorted by the Morton code of their
 * broadphase cell, ties by body index).  A constraint whose dynamic bodies all lie in one tile
 * is INTERIOR to it, every other one is a BOUNDARY constraint (static bodies belong to no tile
 * and never make a constraint a boundary one).  The two classes are coloured in separate
 * colour namespaces (a body has one colour mask per class), and a solver pass runs
 *     the interior constraints, colour 0, 1, 2, ... (tiles are independent of each other),
 *     then the boundary constraints: colours with more than SFP_TAIL_ITEMS of them ascending,
 *          then the smaller ones ascending,
 *     then the serial group.
 * Why: on the device one thread block keeps a tile's bodies in shared memory and runs all its
 * interior colours between block-wide barriers; only the boundary constraints (a few per cent)
 * need grid-wide phases.  The assignment is renewed whenever the colouring starts from scratch
 * (every SFP_RECOLOUR_PERIOD-th colouring), from the poses at that moment, and is part of the
 * algorithm's definition: the oracle computes the same tiles.  SFP_FLAG_NO_TILES turns every
 * constraint into a boundary constraint (the plain global colour order). */
#define SFP_TILE_TARGET 148u /* tiles aimed for: one per SM of a B200 */
#define SFP_TILE_MIN 256u
#define SFP_TILE_MAX 1024u
#define SFP_NO_TILE 0xffffffffu
#define SFP_CLASS_INTERIOR 64u /* get_colours reports an interior constraint's colour + 64 */

SFP_HD uint32_t tile_size(uint32_t n_dynamic) {
    uint32_t b = (n_dynamic + SFP_TILE_TARGET - 1u) / SFP_TILE_TARGET;
    if (b < SFP_TILE_MIN) b = SFP_TILE_MIN;
    if (b > SFP_TILE_MAX) b = SFP_TILE_MAX;
    return b;
}
SFP_HD uint32_t morton_spread16(uint32_t v) {
    v &= 0xffffu;
    v = (v | (v << 8)) & 0x00ff00ffu;
    v = (v | (v << 4)) & 0x0f0f0f0fu;
    v = (v | (v << 2)) & 0x33333333u;
    v = (v | (v << 1)) & 0x55555555u;
    return v;
}
/* sort key of a body for the tile order: 31 bits of the Morton code of its cell (the lowest x
 * bit dropped), statics behind every dynamic body */
SFP_HD uint32_t tile_sort_key(float x, float y, float inv_cell, bool dynamic) {
    if (!dynamic) return 0x80000000u;
    uint32_t cx = (uint32_t)(cell_coord(x, inv_cell) - SFP_CELL_MIN);
    uint32_t cy = (uint32_t)(cell_coord(y, inv_cell) - SFP_CELL_MIN);
    return (morton_spread16(cx) | (morton_spread16(cy) << 1)) >> 1;
}
/* class of a constraint between bodies with tiles ta, tb (SFP_NO_TILE for a static body):
 * the tile it is interior to, or SFP_NO_TILE for a boundary constraint */
SFP_HD uint32_t constraint_tile(uint32_t ta, uint32_t tb) {
    if (ta == SFP_NO_TILE) return tb;
    if (tb == SFP_NO_TILE) return ta;
    return ta == tb ? ta : SFP_NO_TILE;
}

/* index of the lowest zero bit of a 64-bit mask (64 if none) */
SFP_HD int lowest_free(uint64_t mask);
/* colour for a constraint whose dynamic bodies use the colours in `mask` */
SFP_HD int pick_colour(uint64_t mask) {
    int c = lowest_free(mask | (1ull << SFP_SERIAL_COLOUR));
    return c >= SFP_SERIAL_COLOUR ? SFP_SERIAL_COLOUR : c;
}

SFP_HD int lowest_free(uint64_t mask) {
    uint64_t inv = ~mask;
    if (inv == 0ull) return 64;
    int i = 0;
    if ((inv & 0xffffffffull) == 0ull) {
        i += 32;
        inv >>= 32;
    }
    if ((inv & 0xffffull) == 0ull) {
        i += 16;
        inv >>= 16;
    }
    if ((inv & 0xffull) == 0ull) {
        i += 8;
        inv >>= 8;
    }
    if ((inv & 0xfull) == 0ull) {
        i += 4;
        inv >>= 4;
    }
    if ((inv & 0x3ull) == 0ull) {
        i += 2;
        inv >>= 2;
    }
    if ((inv & 0x1ull) == 0ull) i += 1;
    return i;
}

} /* namespace sfp */

#endif /* SFP_MATH_H_ */
