"""Host-side mirror of the reference's `Physics` / `Body` / `Collider` surface for the hot path.

north_star: "It keeps the reference's Physics/Collider/Body API surface and scene graph so it is
a drop-in for that path".  The reference's exact Rust signatures are unavailable (tree not
mounted); the names and argument meanings below follow north_star's vocabulary
(Physics step with a frame dt and a substep count; Body with pose / velocity / mass; Collider
of circle / capsule / rounded-rect shape with a friction + restitution material; distance and
rope constraints).  The Rust-side binding a maintainer would add is in INTEGRATION.md.

Everything here is bookkeeping: bodies are collected into an AoS array (what a scene-graph walk
yields), handed to `sfp_upload_bodies`, and each `tick` is one `sfp_step` on the device.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

import numpy as np

from . import _ffi
from ._ffi import (BODY_DESC_DTYPE, BODY_STATIC, JOINT_DESC_DTYPE, JOINT_DISTANCE, JOINT_ROPE,
                   POLYGON_DTYPE, SHAPE_CAPSULE, SHAPE_CIRCLE, SHAPE_POLYGON, SHAPE_RECT, Api, World)

LIB_PATH = os.environ.get(
    "SFP_B200_LIB",
    os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libsfp_b200.so"))
_api: Api | None = None


def product_api() -> Api:
    """Binds libsfp_b200.so.  Raises ImportError when it has not been built: the product has no
    CPU fallback."""
    global _api
    if _api is None:
        _api = Api(LIB_PATH, "sfp_")
    return _api


@dataclass
class Material:
    static_friction: float = 0.5
    dynamic_friction: float = 0.3
    restitution: float = 0.0


@dataclass
class Collider:
    shape: int = SHAPE_CIRCLE
    hx: float = 0.0
    hy: float = 0.0
    radius: float = 0.5
    material: Material = field(default_factory=Material)
    layer: int = 1
    mask: int = 0xFFFF
    points: tuple | None = None  # general convex polygon: CCW body-frame vertices (3..8)

    @staticmethod
    def new_polygon(points, r: float = 0.0) -> "Collider":
        """Convex polygon (counter-clockwise, 3..8 vertices) with rounding radius r.  The
        vertices are re-centred on the polygon's centroid: the body's position is its centre of
        mass."""
        return Collider(SHAPE_POLYGON, 0.0, 0.0, r, points=tuple(map(tuple, points)))

    @staticmethod
    def new_circle(r: float) -> "Collider":
        return Collider(SHAPE_CIRCLE, 0.0, 0.0, r)

    @staticmethod
    def new_capsule(length: float, r: float) -> "Collider":
        """`length` is the distance between the two cap centres."""
        return Collider(SHAPE_CAPSULE, 0.5 * length, 0.0, r)

    @staticmethod
    def new_rect(width: float, height: float) -> "Collider":
        return Collider(SHAPE_RECT, 0.5 * width, 0.5 * height, 0.0)

    @staticmethod
    def new_rounded_rect(width: float, height: float, r: float) -> "Collider":
        """Outer dimensions width x height with corner radius r."""
        return Collider(SHAPE_RECT, 0.5 * width - r, 0.5 * height - r, r)

    def with_material(self, m: Material) -> "Collider":
        self.material = m
        return self

    def with_layer(self, layer_bits: int, mask: int = 0xFFFF) -> "Collider":
        self.layer, self.mask = layer_bits, mask
        return self


@dataclass
class Body:
    x: float = 0.0
    y: float = 0.0
    angle: float = 0.0
    vx: float = 0.0
    vy: float = 0.0
    omega: float = 0.0
    density: float = 1.0
    is_static: bool = False

    @staticmethod
    def new_static(x: float = 0.0, y: float = 0.0, angle: float = 0.0) -> "Body":
        return Body(x, y, angle, is_static=True)

    @staticmethod
    def new_dynamic(x: float = 0.0, y: float = 0.0, angle: float = 0.0,
                    density: float = 1.0) -> "Body":
        return Body(x, y, angle, density=density)

    def with_velocity(self, vx: float, vy: float, omega: float = 0.0) -> "Body":
        self.vx, self.vy, self.omega = vx, vy, omega
        return self


class Physics:
    """The physics world.  `tick(dt, substeps)` is the reference's per-frame step."""

    def __init__(self, gravity=(0.0, -9.81), substeps: int = 8, api: Api | None = None, **params):
        self._api = api if api is not None else product_api()
        self.substeps = substeps
        self._world = World(self._api, gravity_x=gravity[0], gravity_y=gravity[1], **params)
        self._bodies: list[np.void] = []
        self._joints: list[np.void] = []
        self._polygons: list[np.void] = []
        self._dirty = True

    # -- scene building ---------------------------------------------------------------
    def add_body(self, body: Body, collider: Collider) -> int:
        d = np.zeros(1, dtype=BODY_DESC_DTYPE)[0]
        d["x"], d["y"], d["angle"] = body.x, body.y, body.angle
        d["vx"], d["vy"], d["omega"] = body.vx, body.vy, body.omega
        d["hx"], d["hy"], d["radius"] = collider.hx, collider.hy, collider.radius
        d["shape"] = collider.shape
        if collider.shape == SHAPE_POLYGON:
            from .scenes import polygon

            self._polygons.append(polygon(collider.points)[0])
            d["shape"] = SHAPE_POLYGON | ((len(self._polygons) - 1) << 8)
        d["density"] = body.density
        d["mu_s"] = collider.material.static_friction
        d["mu_d"] = collider.material.dynamic_friction
        d["restitution"] = collider.material.restitution
        d["category"], d["mask"] = collider.layer, collider.mask
        d["flags"] = BODY_STATIC if body.is_static else 0
        self._bodies.append(d)
        self._dirty = True
        return len(self._bodies) - 1

    def add_bodies(self, descs: np.ndarray) -> range:
        start = len(self._bodies)
        self._bodies.extend(np.asarray(descs, dtype=BODY_DESC_DTYPE))
        self._dirty = True
        return range(start, len(self._bodies))

    def _joint(self, a, b, anchor_a, anchor_b, rest, compliance, kind) -> int:
        j = np.zeros(1, dtype=JOINT_DESC_DTYPE)[0]
        j["body_a"], j["body_b"] = a, b
        j["ax"], j["ay"] = anchor_a
        j["bx"], j["by"] = anchor_b
        j["rest"], j["compliance"], j["kind"] = rest, compliance, kind
        self._joints.append(j)
        self._dirty = True
        return len(self._joints) - 1

    def add_distance_constraint(self, a: int, b: int, rest: float, anchor_a=(0.0, 0.0),
                                anchor_b=(0.0, 0.0), compliance: float = 0.0) -> int:
        return self._joint(a, b, anchor_a, anchor_b, rest, compliance, JOINT_DISTANCE)

    def add_rope_constraint(self, a: int, b: int, max_length: float, anchor_a=(0.0, 0.0),
                            anchor_b=(0.0, 0.0), compliance: float = 0.0) -> int:
        return self._joint(a, b, anchor_a, anchor_b, max_length, compliance, JOINT_ROPE)

    def add_rope(self, start, links: int, spacing: float, radius: float, anchor_body: int | None,
                 direction=(0.0, -1.0), density: float = 1.0) -> range:
        """A chain of `links` circle particles joined by distance constraints."""
        ids = []
        prev = anchor_body
        for k in range(links):
            x = start[0] + direction[0] * spacing * (k + 1)
            y = start[1] + direction[1] * spacing * (k + 1)
            i = self.add_body(Body.new_dynamic(x, y, density=density), Collider.new_circle(radius))
            if prev is not None:
                self.add_distance_constraint(prev, i, spacing)
            prev = i
            ids.append(i)
        return range(ids[0], ids[-1] + 1)

    # -- stepping ---------------------------------------------------------------------
    def _flush(self):
        if self._dirty:
            bodies = (np.array(self._bodies, dtype=BODY_DESC_DTYPE) if self._bodies
                      else np.zeros(0, dtype=BODY_DESC_DTYPE))
            # joints first cleared so a shrinking body set never leaves dangling references
            self._world.set_joints(np.zeros(0, dtype=JOINT_DESC_DTYPE))
            if self._polygons:
                self._world.set_polygons(np.array(self._polygons, dtype=POLYGON_DTYPE))
            self._world.upload_bodies(bodies)
            joints = (np.array(self._joints, dtype=JOINT_DESC_DTYPE) if self._joints
                      else np.zeros(0, dtype=JOINT_DESC_DTYPE))
            self._world.set_joints(joints)
            self._dirty = False

    def tick(self, dt: float = 1.0 / 60.0, substeps: int | None = None):
        self._flush()
        self._world.step(dt, self.substeps if substeps is None else substeps)

    def sync_to_scene(self):
        """Reads poses/velocities back and folds them into the stored descs (scene graph)."""
        self._flush()
        st = self._world.download_bodies()
        for d, s in zip(self._bodies, st):
            d["x"], d["y"], d["angle"] = s["x"], s["y"], s["angle"]
            d["vx"], d["vy"], d["omega"] = s["vx"], s["vy"], s["omega"]
        return st

    def body_states(self) -> np.ndarray:
        self._flush()
        return self._world.download_bodies()

    def contacts(self) -> np.ndarray:
        return self._world.get_contacts()

    def contact_events(self) -> np.ndarray:
        """collisions of the last substep (body pair, normal, point, impulse) for game code"""
        return self._world.get_contact_events()

    def ray_cast(self, origins, directions, max_t: float = 1.0e9, mask: int = 0xFFFF) -> np.ndarray:
        """closest hit per ray (body index 0xffffffff = none) against the current poses"""
        self._flush()
        from ._ffi import RAY_DTYPE
        o = np.atleast_2d(np.asarray(origins, dtype=np.float32))
        d = np.atleast_2d(np.asarray(directions, dtype=np.float32))
        rays = np.zeros(len(o), dtype=RAY_DTYPE)
        rays["ox"], rays["oy"], rays["dx"], rays["dy"] = o[:, 0], o[:, 1], d[:, 0], d[:, 1]
        rays["max_t"], rays["mask"] = max_t, mask
        return self._world.ray_cast(rays)

    def query_aabb(self, lo, hi, mask: int = 0xFFFF) -> np.ndarray:
        """indices of the bodies whose tight box overlaps [lo, hi]"""
        self._flush()
        return self._world.query_aabb(lo, hi, mask)

    def query_point(self, x: float, y: float, mask: int = 0xFFFF) -> np.ndarray:
        """indices of the bodies whose collider contains the point (picking)"""
        self._flush()
        return self._world.query_point(x, y, mask)

    def query_shape(self, collider: "Collider", x: float, y: float, angle: float = 0.0,
                    mask: int = 0xFFFF) -> np.ndarray:
        """indices of the bodies whose collider overlaps `collider` placed at (x, y, angle)"""
        if collider.points is not None:
            raise ValueError("query_shape: general-polygon probes go through World.query_shape "
                             "with an index into the world's polygon table")
        self._flush()
        from ._ffi import BODY_DESC_DTYPE
        q = np.zeros(1, dtype=BODY_DESC_DTYPE)
        q["shape"], q["hx"], q["hy"], q["radius"] = collider.shape, collider.hx, collider.hy, collider.radius
        q["x"], q["y"], q["angle"] = x, y, angle
        return self._world.query_shape(q, mask)

    def set_body_states(self, idx, states: np.ndarray):
        """teleport / kinematic control of individual bodies between ticks"""
        self._flush()
        self._world.update_bodies(np.asarray(idx, dtype=np.uint32), states)

    def tick_host(self, states: np.ndarray, dt: float = 1.0 / 60.0, out: np.ndarray | None = None):
        """states in from the host, one frame, states out (the engine's per-frame round trip)"""
        self._flush()
        return self._world.step_host(states, dt, self.substeps, out=out)

    def stats(self) -> dict:
        return self._world.stats()

    @property
    def world(self) -> World:
        self._flush()
        return self._world

    def __len__(self):
        return len(self._bodies)
