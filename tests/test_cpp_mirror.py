"""The C++ mirror of the Physics / Body / Collider surface (include/sfp.hpp) compiles, links
against the C ABI and steps a scene — here against the oracle library, so it runs without a GPU."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SRC = r'''
#include <cmath>
#include <cstdio>
#include "sfp.hpp"
using namespace starframe;
int main() {
    Physics ph;
    ph.add_body(Body::new_static(0.f, -0.5f), Collider::new_rect(10.f, 1.f));
    uint32_t ball = ph.add_body(Body::new_dynamic(0.f, 2.f), Collider::new_circle(0.5f));
    uint32_t box = ph.add_body(Body::new_dynamic(2.f, 0.26f), Collider::new_rounded_rect(1.f, 0.5f, 0.05f));
    uint32_t hook = ph.add_body(Body::new_static(-3.f, 3.f), Collider::new_circle(0.05f));
    uint32_t bob = ph.add_body(Body::new_dynamic(-2.f, 3.f), Collider::new_circle(0.1f));
    ph.add_distance_constraint(hook, bob, 1.0f);
    // a hexagon from the polygon table, resting on the ground
    SfpPolygon hex{};
    hex.count = 6;
    for (int k = 0; k < 6; ++k) {
        hex.x[k] = 0.3f * std::cos(6.283185307f * k / 6.f - 1.570796327f + 0.523598776f);
        hex.y[k] = 0.3f * std::sin(6.283185307f * k / 6.f - 1.570796327f + 0.523598776f);
    }
    ph.set_polygons({hex});
    uint32_t nut = ph.add_body(Body::new_dynamic(4.f, 0.5f), Collider::new_polygon(0));
    for (int f = 0; f < 180; ++f) ph.tick();
    auto ev = ph.contact_events();
    bool ball_on_ground = false;
    for (auto& e : ev) ball_on_ground |= (e.a == 0u && e.b == ball && e.impulse > 0.f);
    SfpRay ray{};
    ray.ox = 4.f, ray.oy = 5.f, ray.dx = 0.f, ray.dy = -1.f, ray.max_t = 10.f, ray.mask = 0xffffffffu;
    auto hit = ph.ray_cast({ray});
    if (!ball_on_ground || hit[0].body != nut) return 2;
    // picking: the point at the nut's centre is inside the nut only; a wide flat probe on the ground
    // line touches the ground and what rests on it
    auto s_now = ph.body_states();
    auto pick = ph.query_point(s_now[nut].x, s_now[nut].y);
    if (pick.size() != 1 || pick[0] != nut) return 4;
    auto under = ph.query_shape(Collider::new_rect(20.f, 0.2f), 0.f, 0.05f);
    bool has_ground = false, has_ball = false;
    for (uint32_t i : under) has_ground |= (i == 0u), has_ball |= (i == ball);
    if (!has_ground || !has_ball) return 5;
    if (ph.query_aabb(-50.f, -5.f, 50.f, 50.f).size() != s_now.size()) return 6;
    auto s0 = ph.body_states();
    std::vector<SfpBodyState> s1;
    ph.tick_host(s0, s1);
    if (s1.size() != s0.size()) return 3;
    auto s = ph.body_states();
    float len = std::hypot(s[bob].x - s[hook].x, s[bob].y - s[hook].y);
    std::printf("%f %f %f\n", s[ball].y, s[box].y, len);
    return (std::fabs(s[ball].y - 0.5f) < 0.01f && std::fabs(s[box].y - 0.25f) < 0.01f &&
            std::fabs(len - 1.0f) < 0.01f) ? 0 : 1;
}
'''


def test_cpp_mirror_steps_a_scene_against_the_oracle(tmp_path, ref_api):
    src = tmp_path / "mirror.cpp"
    src.write_text(SRC)
    exe = tmp_path / "mirror"
    subprocess.run(["g++", "-std=c++17", "-DSFP_USE_ORACLE", "-I", os.path.join(ROOT, "include"),
                    str(src), "-o", str(exe), "-L", os.path.join(ROOT, "oracle"), "-lsfp_ref",
                    "-Wl,-rpath," + os.path.join(ROOT, "oracle")], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


def test_cpp_mirror_compiles_against_the_product_abi(tmp_path):
    src = tmp_path / "mirror.cpp"
    src.write_text(SRC)
    subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "include"),
                    str(src)], check=True)
