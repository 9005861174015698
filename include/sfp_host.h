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

inline bool desc_valid(const SfpBodyDesc& d) {
    if (d.shape > SFP_SHAPE_RECT) return false;
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
        if (d.shape == SFP_SHAPE_CIRCLE && !(d.radius > 0.0f)) return false;
        if (d.shape == SFP_SHAPE_CAPSULE && !(d.radius > 0.0f)) return false;
        if (d.shape == SFP_SHAPE_RECT && !((d.hx + d.radius) > 0.0f && (d.hy + d.radius) > 0.0f))
            return false;
    }
    return true;
}

/* solid-shape mass and moment of inertia about the centre (double, then rounded once) */
inline void mass_properties(const SfpBodyDesc& d, float* inv_mass, float* inv_inertia) {
    if (d.flags & SFP_BODY_STATIC) {
        *inv_mass = 0.0f;
        *inv_inertia = 0.0f;
        return;
    }
    const double pi = 3.14159265358979323846;
    double rho = d.density, r = d.radius, m = 0.0, I = 0.0;
    if (d.shape == SFP_SHAPE_CIRCLE) {
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

inline void body_from_desc(const SfpBodyDesc& d, f4* pose, f4* vel, f2* minv, f4* shape,
                           f4* mat) {
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
    mass_properties(d, &minv->x, &minv->y);
    shape->x = (d.shape == SFP_SHAPE_CIRCLE) ? 0.0f : d.hx;
    shape->y = (d.shape == SFP_SHAPE_RECT) ? d.hy : 0.0f;
    shape->z = d.radius;
    shape->w = u2f(d.shape);
    mat->x = d.mu_s;
    mat->y = d.mu_d;
    mat->z = d.restitution;
    mat->w = u2f((uint32_t)d.category | ((uint32_t)d.mask << 16));
}

static_assert(sizeof(StateRec) == sizeof(SfpBodyState), "StateRec mirrors SfpBodyState");

inline SfpBodyState state_from_records(f4 pose, f4 vel) {
    StateRec r = state_pack(pose, vel);
    SfpBodyState s;
    std::memcpy(&s, &r, sizeof(s));
    return s;
}

/* a state whose rotor is (0, 0) carries its rotation in `angle` only: give it a rotor */
inline SfpBodyState state_with_rotor(const SfpBodyState& in) {
    SfpBodyState s = in;
    if (!(s.c * s.c + s.s * s.s > 0.0f)) {
        s.c = cosf(s.angle);
        s.s = sinf(s.angle);
    }
    return s;
}

inline void records_from_state(const SfpBodyState& in, f2 minv, f4* pose, f4* vel) {
    SfpBodyState s = state_with_rotor(in);
    StateRec r;
    std::memcpy(&r, &s, sizeof(r));
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
