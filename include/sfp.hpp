// sfp.hpp — header-only C++ mirror of the reference's Physics / Body / Collider surface over the
// C ABI of sfp.h.  The reference is compiled (Rust) code; its exact signatures are unavailable
// (tree not mounted), so names and argument meaning follow BASELINE.json's
// north_star vocabulary.  The Rust-side equivalent is sketched in INTEGRATION.md.
//
// Link against libsfp_b200.so (product) — or, in tests only, compile with -DSFP_USE_ORACLE and
// link libsfp_ref.so to drive the CPU oracle through the same classes.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "sfp.h"

namespace starframe {

#ifdef SFP_USE_ORACLE
#define SFP_FN(name) sfp_ref_##name
#else
#define SFP_FN(name) sfp_##name
#endif

struct Material {
    float static_friction = 0.5f, dynamic_friction = 0.3f, restitution = 0.0f;
};

struct Collider {
    uint32_t shape = SFP_SHAPE_CIRCLE;
    float hx = 0.0f, hy = 0.0f, radius = 0.5f;
    Material material;
    uint16_t layer = 1, mask = 0xffff;
    static Collider new_circle(float r) { return {SFP_SHAPE_CIRCLE, 0.0f, 0.0f, r}; }
    static Collider new_capsule(float length, float r) {
        return {SFP_SHAPE_CAPSULE, 0.5f * length, 0.0f, r};
    }
    static Collider new_rect(float w, float h) { return {SFP_SHAPE_RECT, 0.5f * w, 0.5f * h, 0.0f}; }
    static Collider new_rounded_rect(float w, float h, float r) {
        return {SFP_SHAPE_RECT, 0.5f * w - r, 0.5f * h - r, r};
    }
    /* general convex polygon: `index` into the table given to Physics::set_polygons */
    static Collider new_polygon(uint32_t index, float rounding = 0.0f) {
        return {SFP_SHAPE_POLYGON | (index << 8), 0.0f, 0.0f, rounding};
    }
    Collider with_material(Material m) const {
        Collider c = *this;
        c.material = m;
        return c;
    }
};

struct Body {
    float x = 0.0f, y = 0.0f, angle = 0.0f, vx = 0.0f, vy = 0.0f, omega = 0.0f, density = 1.0f;
    bool is_static = false;
    static Body new_static(float x, float y, float angle = 0.0f) {
        Body b;
        b.x = x, b.y = y, b.angle = angle, b.is_static = true;
        return b;
    }
    static Body new_dynamic(float x, float y, float angle = 0.0f, float density = 1.0f) {
        Body b;
        b.x = x, b.y = y, b.angle = angle, b.density = density;
        return b;
    }
};

class Physics {
   public:
    explicit Physics(float gx = 0.0f, float gy = -9.81f, uint32_t substeps = 8)
        : substeps_(substeps) {
        SfpParams p;
        SFP_FN(default_params)(&p);
        p.gravity_x = gx;
        p.gravity_y = gy;
        check(SFP_FN(create)(&p, &w_), "create");
    }
    ~Physics() { SFP_FN(destroy)(w_); }
    Physics(const Physics&) = delete;
    Physics& operator=(const Physics&) = delete;

    uint32_t add_body(const Body& b, const Collider& c) {
        SfpBodyDesc d{};
        d.x = b.x, d.y = b.y, d.angle = b.angle, d.vx = b.vx, d.vy = b.vy, d.omega = b.omega;
        d.hx = c.hx, d.hy = c.hy, d.radius = c.radius, d.shape = c.shape, d.density = b.density;
        d.mu_s = c.material.static_friction, d.mu_d = c.material.dynamic_friction;
        d.restitution = c.material.restitution;
        d.category = c.layer, d.mask = c.mask, d.flags = b.is_static ? SFP_BODY_STATIC : 0u;
        bodies_.push_back(d);
        dirty_ = true;
        return (uint32_t)bodies_.size() - 1u;
    }
    uint32_t add_distance_constraint(uint32_t a, uint32_t b, float rest, float compliance = 0.0f) {
        return add_joint(a, b, rest, compliance, SFP_JOINT_DISTANCE);
    }
    uint32_t add_rope_constraint(uint32_t a, uint32_t b, float max_len, float compliance = 0.0f) {
        return add_joint(a, b, max_len, compliance, SFP_JOINT_ROPE);
    }
    /* vertex table of the polygon colliders (CCW, strictly convex, centroid at the origin) */
    void set_polygons(const std::vector<SfpPolygon>& polys) {
        check(SFP_FN(set_polygons)(w_, polys.data(), (uint32_t)polys.size()), "set_polygons");
        dirty_ = true;
    }
    // the reference's per-frame step: frame dt, substep count
    void tick(float dt = 1.0f / 60.0f) {
        flush();
        check(SFP_FN(step)(w_, dt, substeps_), "step");
    }
    /* one call per frame for engines that keep body state on the host: states in, states out */
    void tick_host(const std::vector<SfpBodyState>& in, std::vector<SfpBodyState>& out,
                   float dt = 1.0f / 60.0f) {
        flush();
        out.resize(in.size());
        check(SFP_FN(step_host)(w_, in.data(), out.data(), (uint32_t)in.size(), dt, substeps_), "step_host");
    }
    /* collisions of the last substep, for game code */
    std::vector<SfpContactEvent> contact_events() {
        uint32_t n = 0;
        check(SFP_FN(get_contact_events)(w_, nullptr, 0, &n), "get_contact_events");
        std::vector<SfpContactEvent> out(n);
        if (n) check(SFP_FN(get_contact_events)(w_, out.data(), n, &n), "get_contact_events");
        return out;
    }
    std::vector<SfpRayHit> ray_cast(const std::vector<SfpRay>& rays) {
        flush();
        std::vector<SfpRayHit> hits(rays.size());
        if (!rays.empty())
            check(SFP_FN(ray_cast)(w_, rays.data(), (uint32_t)rays.size(), hits.data()), "ray_cast");
        return hits;
    }
    /* bodies whose tight box overlaps [lo, hi] */
    std::vector<uint32_t> query_aabb(float lo_x, float lo_y, float hi_x, float hi_y, uint32_t mask = 0xffff) {
        flush();
        std::vector<uint32_t> out(bodies_.size() + 1);
        uint32_t n = 0;
        check(SFP_FN(query_aabb)(w_, lo_x, lo_y, hi_x, hi_y, mask, out.data(), (uint32_t)out.size(), &n),
              "query_aabb");
        out.resize(n);
        return out;
    }
    /* bodies whose collider overlaps `c` placed at (x, y, angle): the narrowphase with zero margin */
    std::vector<uint32_t> query_shape(const Collider& c, float x, float y, float angle = 0.0f,
                                      uint32_t mask = 0xffff) {
        flush();
        SfpBodyDesc q = {};
        q.shape = c.shape, q.hx = c.hx, q.hy = c.hy, q.radius = c.radius;
        q.x = x, q.y = y, q.angle = angle;
        std::vector<uint32_t> out(bodies_.size() + 1);
        uint32_t n = 0;
        check(SFP_FN(query_shape)(w_, &q, mask, out.data(), (uint32_t)out.size(), &n), "query_shape");
        out.resize(n);
        return out;
    }
    /* bodies whose collider contains the point (picking) */
    std::vector<uint32_t> query_point(float x, float y, uint32_t mask = 0xffff) {
        return query_shape(Collider::new_circle(0.0f), x, y, 0.0f, mask);
    }
    /* teleport / kinematic control of individual bodies between ticks */
    void set_body_states(const std::vector<uint32_t>& idx, const std::vector<SfpBodyState>& states) {
        flush();
        check(SFP_FN(update_bodies)(w_, idx.data(), states.data(), (uint32_t)idx.size()), "update_bodies");
    }
    std::vector<SfpBodyState> body_states() {
        flush();
        std::vector<SfpBodyState> out(bodies_.size());
        check(SFP_FN(download_bodies)(w_, out.data(), (uint32_t)out.size()), "download_bodies");
        return out;
    }
    SfpStats stats() {
        SfpStats s;
        check(SFP_FN(get_stats)(w_, &s), "get_stats");
        return s;
    }

   private:
    uint32_t add_joint(uint32_t a, uint32_t b, float rest, float compliance, uint32_t kind) {
        SfpJointDesc j{};
        j.body_a = a, j.body_b = b, j.rest = rest, j.compliance = compliance, j.kind = kind;
        joints_.push_back(j);
        dirty_ = true;
        return (uint32_t)joints_.size() - 1u;
    }
    void flush() {
        if (!dirty_) return;
        check(SFP_FN(set_joints)(w_, nullptr, 0), "set_joints");
        check(SFP_FN(upload_bodies)(w_, bodies_.data(), (uint32_t)bodies_.size()), "upload_bodies");
        check(SFP_FN(set_joints)(w_, joints_.data(), (uint32_t)joints_.size()), "set_joints");
        dirty_ = false;
    }
    void check(int32_t rc, const char* what) {
        if (rc != SFP_OK)
            throw std::runtime_error(std::string(what) + " failed (" + std::to_string(rc) +
                                     "): " + (w_ ? SFP_FN(last_error)(w_) : ""));
    }
    SfpWorld* w_ = nullptr;
    uint32_t substeps_;
    std::vector<SfpBodyDesc> bodies_;
    std::vector<SfpJointDesc> joints_;
    bool dirty_ = true;
};

}  // namespace starframe
