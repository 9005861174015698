/* sfp_host.h — host-side conversions shared by the CUDA library's host code and the oracle:
 * scene-graph records (AoS, sfp.h) <-> solver records (SoA, sfp_math.h), validation and
 * mass properties.  Host only (uses libm cosf/sinf/atan2f); both libraries run it on the
 * same CPU, so both start from identical bits.
 *
 * Reference: the mass/inertia formulas of the reference's collider shapes are unavailable
 * (tree not mounted); these are the textbook solid-shape formulas.
 */
#ifndef SFP_HOST_H_
#define SFP_HOST_H_

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "sfp.h"
#include "sfp_math.h"

namespace sfp {

inline void default_params(SfpParams* p) {
    std::memset(p, 0, sizeof(*p));
    p->gravity_x = 0.0f;
    p->gravity_y = -9.81f;
    p->cell_size = 0.5f;
    p->aabb_margin = 0.02f;
    p->vel_expand = 1.0f;
    p->contact_margin = 0.01f;
    p->solver_order = SFP_ORDER_COLOUR;
    p->flags = 0;
    p->device = 0;
}

inline bool params_valid(const SfpParams& p) {
    return std::isfinite(p.gravity_x) && std::isfinite(p.gravity_y) && p.cell_size > 0.0f &&
           std::isfinite(p.cell_size) && p.aabb_margin >= 0.0f && p.vel_expand >= 0.0f &&
           p.contact_margin >= 0.0f && p.contact_margin <= p.aabb_margin + 1.0f &&
           p.solver_order <= SFP_ORDER_REFERENCE;
}

/* host copy of the polygon table: validated input + derived records */
struct PolyTable {
    std::vector<SfpPolygon> polys;
    std::vector<f4> recs;      /* 8 f4 per polygon: vertices then edge normals (device layout) */
    std::vector<float> bound;  /* largest vertex distance from the origin */
    std::vector<double> area, perimeter, inertia_per_density; /* of the core polygon */
};

/* validates and derives; returns the index of the first bad polygon or -1 */
inline long build_poly_table(const SfpPolygon* polys, uint32_t n, PolyTable* t) {
    t->polys.assign(polys, polys + n);
    t->recs.assign((size_t)n * 8u, f4{0.0f, 0.0f, 0.0f, 0.0f});
    t->bound.assign(n, 0.0f);
    t->area.assign(n, 0.0);
    t->perimeter.assign(n, 0.0);
    t->inertia_per_density.assign(n, 0.0);
    for (uint32_t k = 0; k < n; ++k) {
        const SfpPolygon& p = polys[k];
        if (p.count < 3 || p.count > SFP_POLYGON_MAX_VERTS) return k;
        double area2 = 0.0, cx = 0.0, cy = 0.0, inert = 0.0, per = 0.0;
        float bound = 0.0f;
        for (uint32_t i = 0; i < p.count; ++i) {
            uint32_t j = (i + 1) % p.count, l = (i + 2) % p.count;
            if (!std::isfinite(p.x[i]) || !std::isfinite(p.y[i])) return k;
            double ex = (double)p.x[j] - p.x[i], ey = (double)p.y[j] - p.y[i];
            double fx = (double)p.x[l] - p.x[j], fy = (double)p.y[l] - p.y[j];
            if (!(ex * fy - ey * fx > 0.0)) return k; /* not CCW / not strictly convex */
            double cr = (double)p.x[i] * p.y[j] - (double)p.x[j] * p.y[i];
            area2 += cr;
            cx += (p.x[i] + p.x[j]) * cr;
            cy += (p.y[i] + p.y[j]) * cr;
            inert += cr * ((double)p.x[i] * p.x[i] + (double)p.x[i] * p.x[j] + (double)p.x[j] * p.x[j] +
                           (double)p.y[i] * p.y[i] + (double)p.y[i] * p.y[j] + (double)p.y[j] * p.y[j]);
            per += std::sqrt(ex * ex + ey * ey);
            bound = std::max(bound, sqrtf(p.x[i] * p.x[i] + p.y[i] * p.y[i]));
        }
        if (!(area2 > 0.0)) return k;
        cx /= 3.0 * area2;
        cy /= 3.0 * area2;
        if (std::sqrt(cx * cx + cy * cy) > 1.0e-3 * bound) return k; /* centroid must be the origin */
        t->bound[k] = bound;
        t->area[k] = 0.5 * area2;
        t->perimeter[k] = per;
        t->inertia_per_density[k] = inert / 12.0;
        /* device records: vertices (pairs), then unit edge normals; unused slots repeat */
        float vx[8], vy[8], nx[8], ny[8];
        for (uint32_t i = 0; i < 8; ++i) {
            uint32_t a = i % p.count, b = (a + 1) % p.count;
            vx[i] = p.x[a];
            vy[i] = p.y[a];
            float ex = p.x[b] - p.x[a], ey = p.y[b] - p.y[a];
            float inv = 1.0f / sqrtf(ex * ex + ey * ey);
            nx[i] = ey * inv;
            ny[i] = -ex * inv;
        }
        for (uint32_t i = 0; i < 4; ++i) {
            t->recs[(size_t)k * 8 + i] = f4{vx[2 * i], vy[2 * i], vx[2 * i + 1], vy[2 * i + 1]};
            t->recs[(size_t)k * 8 + 4 + i] = f4{nx[2 * i], ny[2 * i], nx[2 * i + 1], ny[2 * i + 1]};
        }
    }
    return -1;
}

inline bool desc_valid(const SfpBodyDesc& d, const PolyTable* polys = nullptr) {
    const uint32_t type = d.shape & 0xffu;
    if (type > SFP_SHAPE_POLYGON) return false;
    if (type == SFP_SHAPE_POLYGON) {
        if (!polys || (d.shape >> 8) >= polys->polys.size()) return false;
    } else if (d.shape >> 8) {
        return false;
    }
    const float f[] = {d.x, d.y, d.angle, d.vx, d.vy, d.omega, d.hx, d.hy, d.radius,
                       d.density, d.mu_s, d.mu_d, d.restitution};
    for (float v : f)
        if (!std::isfinite(v)) return false;
    if (d.hx < 0.0f || d.hy < 0.0f || d.radius < 0.0f) return false;
    if (d.mu_s < 0.0f || d.mu_d < 0.0f || d.restitution < 0.0f) return false;
    bool is_static = (d.flags & SFP_BODY_STATIC) != 0u;
    if (!is_static && !(d.density > 0.0f)) return false;
    /* a dynamic collider needs a positive area */
    if (!is_static) {
        if (type == SFP_SHAPE_CIRCLE && !(d.radius > 0.0f)) return false;
        if (type == SFP_SHAPE_CAPSULE && !(d.radius > 0.0f)) return false;
        if (type == SFP_SHAPE_RECT && !((d.hx + d.radius) > 0.0f && (d.hy + d.radius) > 0.0f))
            return false;
    }
    return true;
}

/* solid-shape mass and moment of inertia about the centre (double, then rounded once) */
inline void mass_properties(const SfpBodyDesc& d, float* inv_mass, float* inv_inertia,
                            const PolyTable* polys = nullptr) {
    if (d.flags & SFP_BODY_STATIC) {
        *inv_mass = 0.0f;
        *inv_inertia = 0.0f;
        return;
    }
    const double pi = 3.14159265358979323846;
    double rho = d.density, r = d.radius, m = 0.0, I = 0.0;
    if ((d.shape & 0xffu) == SFP_SHAPE_POLYGON) {
        /* core polygon exactly; the rounding rim (perimeter * r + pi r^2) adds mass, and the
         * inertia is scaled by the same ratio */
        const uint32_t k = d.shape >> 8;
        double core = polys->area[k], rim = polys->perimeter[k] * r + pi * r * r;
        m = rho * (core + rim);
        I = rho * polys->inertia_per_density[k] * (core + rim) / core;
    } else if (d.shape == SFP_SHAPE_CIRCLE) {
        m = rho * pi * r * r;
        I = 0.5 * m * r * r;
    } else if (d.shape == SFP_SHAPE_CAPSULE) {
        double h = d.hx, L = 2.0 * h;
        double mb = rho * L * 2.0 * r, mc = rho * pi * r * r;
        double lc = 4.0 * r / (3.0 * pi);
        m = mb + mc;
        I = mb * (L * L + 4.0 * r * r) / 12.0 + mc * (0.5 * r * r + h * h + 2.0 * h * lc);
    } else {
        /* rounded rectangle approximated by its bounding rectangle */
        double w = 2.0 * (d.hx + r), h = 2.0 * (d.hy + r);
        m = rho * w * h;
        I = m * (w * w + h * h) / 12.0;
    }
    *inv_mass = (float)(1.0 / m);
    *inv_inertia = (float)(1.0 / I);
}

inline void body_from_desc(const SfpBodyDesc& d, f4* pose, f4* vel, f2* minv, f4* shape, f4* mat,
                           const PolyTable* polys = nullptr) {
    pose->x = d.x;
    pose->y = d.y;
    pose->z = cosf(d.angle);
    pose->w = sinf(d.angle);
    /* renormalise so the rotor is unit to the last bit the solver would produce */
    float inv = 1.0f / sqrtf(pose->z * pose->z + pose->w * pose->w);
    pose->z *= inv;
    pose->w *= inv;
    bool is_static = (d.flags & SFP_BODY_STATIC) != 0u;
    vel->x = is_static ? 0.0f : d.vx;
    vel->y = is_static ? 0.0f : d.vy;
    vel->z = is_static ? 0.0f : d.omega;
    vel->w = 0.0f;
    mass_properties(d, &minv->x, &minv->y, polys);
    if ((d.shape & 0xffu) == SFP_SHAPE_POLYGON) {
        const uint32_t k = d.shape >> 8;
        shape->x = polys->bound[k]; /* half extents = bounding radius: conservative boxes */
        shape->y = polys->bound[k];
        shape->z = d.radius;
        shape->w = u2f((uint32_t)SFP_SHAPE_POLYGON | (polys->polys[k].count << 2) | (k << 8));
    } else {
        shape->x = (d.shape == SFP_SHAPE_CIRCLE) ? 0.0f : d.hx;
        shape->y = (d.shape == SFP_SHAPE_RECT) ? d.hy : 0.0f;
        shape->z = d.radius;
        shape->w = u2f(d.shape);
    }
    mat->x = d.mu_s;
    mat->y = d.mu_d;
    mat->z = d.restitution;
    mat->w = u2f((uint32_t)d.category | ((uint32_t)d.mask << 16));
}

/* probe collider of sfp_query_shape: pose and shape of `d` as a body would get them; only the
 * geometry fields are looked at (the probe is treated as static: no mass, no material) */
inline bool probe_from_desc(const SfpBodyDesc& d, const PolyTable* polys, f4* pose, f4* shape) {
    SfpBodyDesc q = d;
    q.flags = SFP_BODY_STATIC;
    q.vx = q.vy = q.omega = 0.0f;
    q.density = 1.0f;
    q.mu_s = q.mu_d = q.restitution = 0.0f;
    if (!desc_valid(q, polys)) return false;
    f4 vel, mat;
    f2 minv;
    body_from_desc(q, pose, &vel, &minv, shape, &mat, polys);
    return true;
}

/* Does the collider, as uploaded, fit the broadphase grid?  A body whose box (tight box +
 * aabb_margin) covers more than SFP_MAX_CELLS_PER_BODY cells, or 4096 or more cells along one
 * axis, would be left out of the broadphase (cell_range() refuses it: it would produce no pairs
 * and bodies would fall through it).  For a floor or a wall that is a property of the UPLOAD —
 * statics do not move — so the upload rejects it with a message instead of letting the step
 * raise the runaway flag later.  At the default 0.5 m cell: 8192 m^2 of box area at most, and
 * less than 2048 m along either axis. */
inline bool body_fits_grid(f4 pose, f4 shape, const SfpParams& p) {
    f4 zero = {0.0f, 0.0f, 0.0f, 0.0f};
    f4 bb = body_aabb(pose, zero, shape, p.aabb_margin, 0.0f, 0.0f);
    int x0, y0, x1, y1;
    return cell_range(bb, 1.0f / p.cell_size, &x0, &y0, &x1, &y1);
}

static_assert(sizeof(StateRec) == sizeof(SfpBodyState), "StateRec mirrors SfpBodyState");

inline SfpBodyState state_from_records(f4 pose, f4 vel) {
    StateRec r = state_pack(pose, vel);
    SfpBodyState s;
    std::memcpy(&s, &r, sizeof(s));
    return s;
}

/* (a state whose rotor is (0, 0) takes its rotation from `angle`: state_unpack, sfp_math.h) */
inline void records_from_state(const SfpBodyState& in, f2 minv, f4* pose, f4* vel) {
    StateRec r;
    std::memcpy(&r, &in, sizeof(r));
    state_unpack(r, minv, pose, vel);
}

inline bool joint_valid(const SfpJointDesc& j, uint32_t n_bodies) {
    return j.body_a < n_bodies && j.body_b < n_bodies && j.body_a != j.body_b &&
           j.kind <= SFP_JOINT_ROPE && std::isfinite(j.rest) && j.rest >= 0.0f &&
           std::isfinite(j.compliance) && j.compliance >= 0.0f && std::isfinite(j.ax) &&
           std::isfinite(j.ay) && std::isfinite(j.bx) && std::isfinite(j.by);
}

/* API-side contact manifold from the solver's records (poses = end of the substep) */
inline SfpContact contact_from_records(uint32_t a, uint32_t b, const ContactRec& c, f4 lam,
                                       f4 pose_a, f4 pose_b) {
    SfpContact o;
    std::memset(&o, 0, sizeof(o));
    o.a = a;
    o.b = b;
    o.count = f2u(c.b.y);
    o.nx = c.a.x;
    o.ny = c.a.y;
    const f4 pp[2] = {c.p0, c.p1};
    const float sep[2] = {c.s.x, c.s.y};
    const float ln[2] = {lam.x, lam.z}, lt[2] = {lam.y, lam.w};
    for (uint32_t k = 0; k < o.count && k < 2; ++k) {
        v2 pa = mk(pose_a.x, pose_a.y) + rot(mk(pose_a.z, pose_a.w), mk(pp[k].x, pp[k].y));
        v2 pb = mk(pose_b.x, pose_b.y) + rot(mk(pose_b.z, pose_b.w), mk(pp[k].z, pp[k].w));
        o.px[k] = 0.5f * (pa.x + pb.x);
        o.py[k] = 0.5f * (pa.y + pb.y);
        o.depth[k] = -sep[k];
        o.lambda_n[k] = ln[k];
        o.lambda_t[k] = lt[k];
    }
    return o;
}

inline void energies(const std::vector<f4>& pose, const std::vector<f4>& vel,
                     const std::vector<f2>& minv, float gx, float gy, float* kinetic,
                     float* potential) {
    double ke = 0.0, pe = 0.0;
    for (size_t i = 0; i < pose.size(); ++i) {
        if (minv[i].x > 0.0f) {
            double m = 1.0 / minv[i].x;
            ke += 0.5 * m * ((double)vel[i].x * vel[i].x + (double)vel[i].y * vel[i].y);
            pe -= m * ((double)gx * pose[i].x + (double)gy * pose[i].y);
        }
        if (minv[i].y > 0.0f) ke += 0.5 / minv[i].y * (double)vel[i].z * vel[i].z;
    }
    *kinetic = (float)ke;
    *potential = (float)pe;
}

} /* namespace sfp */

#endif /* SFP_HOST_H_ */
