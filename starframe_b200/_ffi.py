"""ctypes binding of the C ABI declared in include/sfp.h.

`Api(lib_path, prefix)` binds one shared library.  The product package only ever binds
`starframe_b200/csrc/libsfp_b200.so` (prefix ``sfp_``) and raises if it is missing — there is no
CPU fallback.  `oracle/loader.py` (tests / bench cpu_baseline / smoke only) binds the oracle with
the prefix ``sfp_ref_`` through the same class, so parity tests are backend-generic.

Reference: the reference's `Physics` API cannot be cited by file:line (tree not mounted,
BASELINE.md §0); this mirrors SURVEY.md §8b.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

SFP_OK = 0
SFP_E_INVALID = -1
SFP_E_CUDA = -2
SFP_E_CAPACITY = -3
SFP_E_NOMEM = -5

SHAPE_CIRCLE, SHAPE_CAPSULE, SHAPE_RECT, SHAPE_POLYGON = 0, 1, 2, 3
BODY_STATIC = 1
ORDER_COLOUR, ORDER_REFERENCE = 0, 1
JOINT_DISTANCE, JOINT_ROPE = 0, 1
FLAG_BROADPHASE_EVERY_SUBSTEP = 1
FLAG_MULTI_KERNEL = 2
FLAG_FLOW_SOLVE = 4
FLAG_TILES = 8
FLAG_BATCH_WHOLE_WORLDS = 16

# numpy mirrors of the PODs (sizes asserted against the C side's static_asserts)
PARAMS_DTYPE = np.dtype(
    [
        ("gravity_x", "<f4"), ("gravity_y", "<f4"), ("cell_size", "<f4"),
        ("aabb_margin", "<f4"), ("vel_expand", "<f4"), ("contact_margin", "<f4"),
        ("solver_order", "<u4"), ("flags", "<u4"), ("max_bodies", "<u4"),
        ("max_joints", "<u4"), ("max_entries", "<u4"), ("max_pairs", "<u4"),
        ("device", "<i4"), ("reserved", "<u4", (3,)),
    ]
)
BODY_DESC_DTYPE = np.dtype(
    [
        ("x", "<f4"), ("y", "<f4"), ("angle", "<f4"),
        ("vx", "<f4"), ("vy", "<f4"), ("omega", "<f4"),
        ("hx", "<f4"), ("hy", "<f4"), ("radius", "<f4"),
        ("shape", "<u4"), ("density", "<f4"),
        ("mu_s", "<f4"), ("mu_d", "<f4"), ("restitution", "<f4"),
        ("category", "<u2"), ("mask", "<u2"), ("flags", "<u4"),
    ]
)
JOINT_DESC_DTYPE = np.dtype(
    [
        ("body_a", "<u4"), ("body_b", "<u4"), ("ax", "<f4"), ("ay", "<f4"),
        ("bx", "<f4"), ("by", "<f4"), ("rest", "<f4"), ("compliance", "<f4"),
        ("kind", "<u4"), ("reserved", "<u4", (3,)),
    ]
)
BODY_STATE_DTYPE = np.dtype(
    [
        ("x", "<f4"), ("y", "<f4"), ("c", "<f4"), ("s", "<f4"),
        ("vx", "<f4"), ("vy", "<f4"), ("omega", "<f4"), ("angle", "<f4"),
    ]
)
PAIR_DTYPE = np.dtype([("a", "<u4"), ("b", "<u4")])
CONTACT_DTYPE = np.dtype(
    [
        ("a", "<u4"), ("b", "<u4"), ("count", "<u4"), ("nx", "<f4"), ("ny", "<f4"),
        ("depth", "<f4", (2,)), ("px", "<f4", (2,)), ("py", "<f4", (2,)),
        ("lambda_n", "<f4", (2,)), ("lambda_t", "<f4", (2,)),
    ]
)
EVENT_DTYPE = np.dtype(
    [
        ("a", "<u4"), ("b", "<u4"), ("nx", "<f4"), ("ny", "<f4"), ("px", "<f4"), ("py", "<f4"),
        ("impulse", "<f4"), ("points", "<u4"),
    ]
)
POLYGON_DTYPE = np.dtype([("count", "<u4"), ("reserved", "<u4"), ("x", "<f4", (8,)),
                          ("y", "<f4", (8,))])
RAY_DTYPE = np.dtype([("ox", "<f4"), ("oy", "<f4"), ("dx", "<f4"), ("dy", "<f4"), ("max_t", "<f4"),
                      ("mask", "<u4"), ("reserved", "<u4", (2,))])
RAY_HIT_DTYPE = np.dtype([("body", "<u4"), ("t", "<f4"), ("nx", "<f4"), ("ny", "<f4")])
STATS_DTYPE = np.dtype(
    [
        ("frames", "<u8"), ("body_substeps", "<u8"),
        ("n_bodies", "<u4"), ("n_dynamic", "<u4"), ("n_joints", "<u4"),
        ("n_entries", "<u4"), ("n_pairs", "<u4"), ("n_contacts", "<u4"),
        ("n_manifolds", "<u4"), ("n_colours", "<u4"), ("colour_rounds", "<u4"),
        ("overflow", "<u4"), ("max_penetration", "<f4"), ("kinetic_energy", "<f4"),
        ("potential_energy", "<f4"), ("launches", "<u4"), ("reserved", "<u4", (2,)),
    ]
)
assert PARAMS_DTYPE.itemsize == 64
assert BODY_DESC_DTYPE.itemsize == 64
assert JOINT_DESC_DTYPE.itemsize == 48
assert BODY_STATE_DTYPE.itemsize == 32
assert PAIR_DTYPE.itemsize == 8
assert CONTACT_DTYPE.itemsize == 60
assert STATS_DTYPE.itemsize == 80
assert EVENT_DTYPE.itemsize == 32
assert RAY_DTYPE.itemsize == 32 and RAY_HIT_DTYPE.itemsize == 16
assert POLYGON_DTYPE.itemsize == 72

# every symbol include/sfp.h declares for the product library (checked by the CPU tests)
PRODUCT_SYMBOLS = [
    "default_params", "create", "destroy", "upload_bodies", "set_joints", "set_polygons", "step",
    "sync",
    "download_bodies", "get_pairs", "get_contacts", "get_colours", "get_stats", "last_error",
    "get_contact_events", "update_bodies", "ray_cast", "query_aabb", "query_shape", "step_host",
    "alloc_host_states", "free_host_states",
    "batch_create", "batch_destroy", "batch_upload", "batch_step", "batch_sync",
    "batch_download", "batch_step_host", "batch_set_joints", "batch_set_polygons", "batch_stats", "batch_last_error", "last_step_ms", "batch_last_step_ms",
    "last_step_breakdown_ms", "last_step_stage_ns", "backend", "step_small_noalloc",
    "batch_create_multi", "batch_multi_destroy", "batch_multi_partition", "batch_multi_upload",
    "batch_multi_set_polygons", "batch_multi_set_joints", "batch_multi_step", "batch_multi_sync",
    "batch_multi_download", "batch_multi_step_host", "batch_multi_stats", "batch_multi_last_step_ms",
    "batch_multi_last_error",
]
ORACLE_SYMBOLS = [
    "default_params", "create", "destroy", "upload_bodies", "set_joints", "set_polygons", "step",
    "sync",
    "download_bodies", "get_pairs", "get_contacts", "get_colours", "get_stats", "last_error",
    "get_contact_events", "update_bodies", "ray_cast", "query_aabb", "query_shape", "step_host",
    "last_step_ms", "backend", "brute_pairs",
]


class SfpError(RuntimeError):
    def __init__(self, code: int, what: str, detail: str = ""):
        self.code = code
        names = {-1: "SFP_E_INVALID", -2: "SFP_E_CUDA", -3: "SFP_E_CAPACITY",
                 -5: "SFP_E_NOMEM"}
        super().__init__(f"{what}: {names.get(code, code)}" + (f" ({detail})" if detail else ""))


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Api:
    """One bound shared library exporting `<prefix>*`."""

    def __init__(self, lib_path: str, prefix: str):
        if not os.path.exists(lib_path):
            raise ImportError(
                f"{lib_path} is missing — build it first (python -c 'import __graft_entry__ as g; "
                "g.build()').  There is no fallback path."
            )
        self.path = lib_path
        self.prefix = prefix
        self.lib = C.CDLL(lib_path)
        self._bind()

    def fn(self, name: str):
        return getattr(self.lib, self.prefix + name)

    def has(self, name: str) -> bool:
        try:
            self.fn(name)
            return True
        except AttributeError:
            return False

    def _bind(self):
        vp, u32, f32, i32 = C.c_void_p, C.c_uint32, C.c_float, C.c_int32
        sig = {
            "default_params": [vp],
            "create": [vp, C.POINTER(vp)],
            "destroy": [vp],
            "upload_bodies": [vp, vp, u32],
            "set_joints": [vp, vp, u32],
            "set_polygons": [vp, vp, u32],
            "step": [vp, f32, u32],
            "sync": [vp],
            "download_bodies": [vp, vp, u32],
            "get_pairs": [vp, vp, u32, C.POINTER(u32)],
            "get_contacts": [vp, vp, u32, C.POINTER(u32)],
            "get_contact_events": [vp, vp, u32, C.POINTER(u32)],
            "update_bodies": [vp, vp, vp, u32],
            "ray_cast": [vp, vp, u32, vp],
            "query_aabb": [vp, f32, f32, f32, f32, u32, vp, u32, C.POINTER(u32)],
            "query_shape": [vp, vp, u32, vp, u32, C.POINTER(u32)],
            "alloc_host_states": [u32, C.POINTER(vp)],
            "free_host_states": [vp],
            "get_colours": [vp, vp, u32, vp, u32],
            "get_stats": [vp, vp],
            "step_host": [vp, vp, vp, u32, f32, u32],
            "last_step_ms": [vp, C.POINTER(f32)],
            "last_step_breakdown_ms": [vp, vp],
            "last_step_stage_ns": [vp, vp],
            "brute_pairs": [vp, f32, vp, u32, C.POINTER(u32)],
            "batch_create": [vp, u32, u32, C.POINTER(vp)],
            "batch_destroy": [vp],
            "batch_upload": [vp, vp],
            "batch_step": [vp, f32, u32],
            "batch_sync": [vp],
            "batch_download": [vp, vp],
            "batch_step_host": [vp, vp, vp, f32, u32],
            "batch_set_joints": [vp, vp, u32, vp],
            "batch_set_polygons": [vp, vp, u32],
            "batch_stats": [vp, vp],
            "batch_last_step_ms": [vp, C.POINTER(f32)],
            "step_small_noalloc": [vp, vp, u32, f32, u32, u32, vp, C.c_char_p, u32],
            "batch_create_multi": [vp, u32, u32, vp, u32, C.POINTER(vp)],
            "batch_multi_destroy": [vp],
            "batch_multi_partition": [vp, u32, C.POINTER(i32), C.POINTER(u32), C.POINTER(u32)],
            "batch_multi_upload": [vp, vp],
            "batch_multi_set_polygons": [vp, vp, u32],
            "batch_multi_set_joints": [vp, vp, u32, vp],
            "batch_multi_step": [vp, f32, u32],
            "batch_multi_sync": [vp],
            "batch_multi_download": [vp, vp],
            "batch_multi_step_host": [vp, vp, vp, f32, u32],
            "batch_multi_stats": [vp, vp],
            "batch_multi_last_step_ms": [vp, C.POINTER(f32)],
        }
        for name, args in sig.items():
            if not self.has(name):
                continue
            f = self.fn(name)
            f.argtypes = args
            f.restype = i32
        for name in ("last_error", "batch_last_error", "batch_multi_last_error"):
            if self.has(name):
                f = self.fn(name)
                f.argtypes = [vp]
                f.restype = C.c_char_p
        if self.has("backend"):
            self.fn("backend").argtypes = []
            self.fn("backend").restype = C.c_char_p

    def alloc_host_states(self, n: int) -> np.ndarray:
        """page-locked SfpBodyState array (freed when the returned array's owner is collected)"""
        p = C.c_void_p()
        rc = self.fn("alloc_host_states")(n, C.byref(p))
        if rc != SFP_OK:
            raise SfpError(rc, "alloc_host_states")
        buf = (C.c_char * (n * BODY_STATE_DTYPE.itemsize)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=BODY_STATE_DTYPE)
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(p)
        return arr

    def backend(self) -> str:
        return self.fn("backend")().decode()

    def default_params(self) -> np.ndarray:
        p = np.zeros(1, dtype=PARAMS_DTYPE)
        rc = self.fn("default_params")(_ptr(p))
        if rc != SFP_OK:
            raise SfpError(rc, "default_params")
        return p


def step_small_noalloc(api: "Api", bodies: np.ndarray, frames: int = 1, dt: float = 1.0 / 60.0,
                       substeps: int = 8, params: np.ndarray | None = None, **overrides) -> np.ndarray:
    """sfp_step_small_noalloc: one small world stepped inside the library's fixed scratch
    array — for hosts that may not allocate device memory while running (include/sfp.h)."""
    b = np.ascontiguousarray(bodies, dtype=BODY_DESC_DTYPE)
    out = np.zeros(b.shape[0], dtype=BODY_STATE_DTYPE)
    err = C.create_string_buffer(512)
    p = api.default_params() if params is None else np.array(params, dtype=PARAMS_DTYPE).reshape(1).copy()
    for k, v in overrides.items():
        p[k] = v
    rc = api.fn("step_small_noalloc")(_ptr(p), _ptr(b), b.shape[0], dt, substeps, frames,
                                      _ptr(out), err, 512)
    if rc != SFP_OK:
        raise SfpError(rc, "step_small_noalloc", err.value.decode())
    return out


class World:
    """Thin RAII wrapper over SfpWorld* for one bound library."""

    def __init__(self, api: Api, params: np.ndarray | None = None, **overrides):
        self.api = api
        p = api.default_params() if params is None else np.array(params, dtype=PARAMS_DTYPE).reshape(1).copy()
        for k, v in overrides.items():
            p[k] = v
        self.params = p
        h = C.c_void_p()
        rc = api.fn("create")(_ptr(p), C.byref(h))
        if rc != SFP_OK:
            raise SfpError(rc, "create")
        self.h = h
        self.n = 0
        self.n_joints = 0

    def _check(self, rc: int, what: str):
        if rc != SFP_OK:
            detail = self.api.fn("last_error")(self.h)
            raise SfpError(rc, what, detail.decode() if detail else "")

    def close(self):
        if getattr(self, "h", None):
            self.api.fn("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def upload_bodies(self, bodies: np.ndarray):
        b = np.ascontiguousarray(bodies, dtype=BODY_DESC_DTYPE)
        self._check(self.api.fn("upload_bodies")(self.h, _ptr(b), b.shape[0]), "upload_bodies")
        self.n = int(b.shape[0])

    def set_polygons(self, polys: np.ndarray):
        p = np.ascontiguousarray(polys, dtype=POLYGON_DTYPE)
        self._check(self.api.fn("set_polygons")(self.h, _ptr(p), p.shape[0]), "set_polygons")

    def set_joints(self, joints: np.ndarray):
        j = np.ascontiguousarray(joints, dtype=JOINT_DESC_DTYPE)
        self._check(self.api.fn("set_joints")(self.h, _ptr(j), j.shape[0]), "set_joints")
        self.n_joints = int(j.shape[0])

    def step(self, dt: float = 1.0 / 60.0, substeps: int = 8):
        self._check(self.api.fn("step")(self.h, dt, substeps), "step")

    def sync(self):
        self._check(self.api.fn("sync")(self.h), "sync")

    def download_bodies(self) -> np.ndarray:
        out = np.zeros(self.n, dtype=BODY_STATE_DTYPE)
        self._check(self.api.fn("download_bodies")(self.h, _ptr(out), self.n), "download_bodies")
        return out

    def step_host(self, states: np.ndarray, dt: float = 1.0 / 60.0, substeps: int = 8,
                  out: np.ndarray | None = None) -> np.ndarray:
        s = np.ascontiguousarray(states, dtype=BODY_STATE_DTYPE)
        if out is None:
            out = np.zeros(self.n, dtype=BODY_STATE_DTYPE)
        self._check(self.api.fn("step_host")(self.h, _ptr(s), _ptr(out), self.n, dt, substeps),
                    "step_host")
        return out

    def get_pairs(self) -> np.ndarray:
        n = C.c_uint32(0)
        self._check(self.api.fn("get_pairs")(self.h, None, 0, C.byref(n)), "get_pairs")
        out = np.zeros(n.value, dtype=PAIR_DTYPE)
        if n.value:
            self._check(self.api.fn("get_pairs")(self.h, _ptr(out), n.value, C.byref(n)), "get_pairs")
        return out

    def get_contacts(self) -> np.ndarray:
        n = C.c_uint32(0)
        self._check(self.api.fn("get_contacts")(self.h, None, 0, C.byref(n)), "get_contacts")
        out = np.zeros(n.value, dtype=CONTACT_DTYPE)
        if n.value:
            self._check(self.api.fn("get_contacts")(self.h, _ptr(out), n.value, C.byref(n)),
                        "get_contacts")
        return out

    def get_contact_events(self) -> np.ndarray:
        n = C.c_uint32(0)
        self._check(self.api.fn("get_contact_events")(self.h, None, 0, C.byref(n)), "events")
        out = np.zeros(n.value, dtype=EVENT_DTYPE)
        if n.value:
            self._check(self.api.fn("get_contact_events")(self.h, _ptr(out), n.value, C.byref(n)),
                        "events")
        return out

    def ray_cast(self, rays: np.ndarray) -> np.ndarray:
        r = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.zeros(r.shape[0], dtype=RAY_HIT_DTYPE)
        self._check(self.api.fn("ray_cast")(self.h, _ptr(r), r.shape[0], _ptr(hits)), "ray_cast")
        return hits

    def query_aabb(self, lo, hi, mask: int = 0xFFFF) -> np.ndarray:
        n = C.c_uint32(0)
        out = np.zeros(max(self.n, 1), dtype=np.uint32)
        self._check(self.api.fn("query_aabb")(self.h, lo[0], lo[1], hi[0], hi[1], mask, _ptr(out),
                                              out.shape[0], C.byref(n)), "query_aabb")
        return out[: n.value].copy()

    def query_shape(self, probe: np.ndarray, mask: int = 0xFFFF) -> np.ndarray:
        """bodies whose collider overlaps the probe collider (one BODY_DESC record: shape, x, y,
        angle, hx, hy, radius)"""
        q = np.ascontiguousarray(probe, dtype=BODY_DESC_DTYPE).reshape(1)
        n = C.c_uint32(0)
        out = np.zeros(max(self.n, 1), dtype=np.uint32)
        self._check(self.api.fn("query_shape")(self.h, _ptr(q), mask, _ptr(out), out.shape[0],
                                               C.byref(n)), "query_shape")
        return out[: n.value].copy()

    def query_point(self, x: float, y: float, mask: int = 0xFFFF) -> np.ndarray:
        """bodies whose collider contains the point: a circle probe of radius 0"""
        q = np.zeros(1, dtype=BODY_DESC_DTYPE)
        q["x"], q["y"] = x, y
        return self.query_shape(q, mask)

    def update_bodies(self, idx: np.ndarray, states: np.ndarray):
        i = np.ascontiguousarray(idx, dtype=np.uint32)
        s = np.ascontiguousarray(states, dtype=BODY_STATE_DTYPE)
        assert i.shape[0] == s.shape[0]
        self._check(self.api.fn("update_bodies")(self.h, _ptr(i), _ptr(s), i.shape[0]),
                    "update_bodies")

    def get_colours(self):
        pairs = self.get_pairs()
        pc = np.zeros(pairs.shape[0], dtype=np.uint32)
        jc = np.zeros(self.n_joints, dtype=np.uint32)
        self._check(
            self.api.fn("get_colours")(self.h, _ptr(pc), pc.shape[0], _ptr(jc), jc.shape[0]),
            "get_colours",
        )
        return pairs, pc, jc

    def stats(self) -> dict:
        s = np.zeros(1, dtype=STATS_DTYPE)
        self._check(self.api.fn("get_stats")(self.h, _ptr(s)), "get_stats")
        return {k: (s[k][0].item() if s[k][0].shape == () else s[k][0].tolist())
                for k in STATS_DTYPE.names if k != "reserved"}

    def last_step_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self.api.fn("last_step_ms")(self.h, C.byref(ms)), "last_step_ms")
        return float(ms.value)

    def last_step_breakdown_ms(self) -> dict:
        out = np.zeros(4, dtype=np.float32)
        self._check(self.api.fn("last_step_breakdown_ms")(self.h, _ptr(out)), "breakdown")
        return {"step": float(out[0]), "broadphase": float(out[1]), "colour": float(out[2]),
                "substeps": float(out[3])}

    def last_step_stage_ns(self) -> dict:
        """ns per stage inside the persistent substep kernel of the last step (whole frame)"""
        out = np.zeros(6, dtype=np.uint64)
        self._check(self.api.fn("last_step_stage_ns")(self.h, _ptr(out)), "stage_ns")
        names = ("integrate", "narrowphase", "position", "derive", "velocity", "small_colours")
        return {k: int(v) for k, v in zip(names, out)}

    def brute_pairs(self, dt: float) -> np.ndarray:
        n = C.c_uint32(0)
        self._check(self.api.fn("brute_pairs")(self.h, dt, None, 0, C.byref(n)), "brute_pairs")
        out = np.zeros(n.value, dtype=PAIR_DTYPE)
        if n.value:
            self._check(self.api.fn("brute_pairs")(self.h, dt, _ptr(out), n.value, C.byref(n)),
                        "brute_pairs")
        return out


class Batch:
    """RAII wrapper over SfpBatch* (product library only)."""

    def __init__(self, api: Api, n_worlds: int, bodies_per_world: int,
                 params: np.ndarray | None = None, **overrides):
        self.api = api
        p = api.default_params() if params is None else np.array(params, dtype=PARAMS_DTYPE).reshape(1).copy()
        for k, v in overrides.items():
            p[k] = v
        self.params = p
        self.n_worlds = int(n_worlds)
        self.bpw = int(bodies_per_world)
        h = C.c_void_p()
        rc = api.fn("batch_create")(_ptr(p), self.n_worlds, self.bpw, C.byref(h))
        if rc != SFP_OK:
            raise SfpError(rc, "batch_create")
        self.h = h

    def _check(self, rc: int, what: str):
        if rc != SFP_OK:
            detail = self.api.fn("batch_last_error")(self.h)
            raise SfpError(rc, what, detail.decode() if detail else "")

    def close(self):
        if getattr(self, "h", None):
            self.api.fn("batch_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, bodies: np.ndarray):
        b = np.ascontiguousarray(bodies, dtype=BODY_DESC_DTYPE).reshape(-1)
        assert b.shape[0] == self.n_worlds * self.bpw
        self._check(self.api.fn("batch_upload")(self.h, _ptr(b)), "batch_upload")

    def set_polygons(self, polys: np.ndarray):
        p = np.ascontiguousarray(polys, dtype=POLYGON_DTYPE)
        self._check(self.api.fn("batch_set_polygons")(self.h, _ptr(p) if p.size else None, p.shape[0]),
                    "batch_set_polygons")

    def set_joints(self, joints: np.ndarray, counts: np.ndarray | None = None):
        """joints: (n_worlds, joints_per_world) descs with world-local body indices"""
        j = np.ascontiguousarray(joints, dtype=JOINT_DESC_DTYPE).reshape(self.n_worlds, -1)
        cnt = None if counts is None else np.ascontiguousarray(counts, dtype=np.uint32)
        self._check(self.api.fn("batch_set_joints")(self.h, _ptr(j) if j.size else None, j.shape[1],
                                                    None if cnt is None else _ptr(cnt)),
                    "batch_set_joints")

    def step(self, dt: float = 1.0 / 60.0, substeps: int = 8):
        self._check(self.api.fn("batch_step")(self.h, dt, substeps), "batch_step")

    def sync(self):
        self._check(self.api.fn("batch_sync")(self.h), "batch_sync")

    def download(self) -> np.ndarray:
        out = np.zeros(self.n_worlds * self.bpw, dtype=BODY_STATE_DTYPE)
        self._check(self.api.fn("batch_download")(self.h, _ptr(out)), "batch_download")
        return out.reshape(self.n_worlds, self.bpw)

    def step_host(self, states: np.ndarray, dt: float = 1.0 / 60.0, substeps: int = 8,
                  out: np.ndarray | None = None) -> np.ndarray:
        """host states in -> one frame -> host states out (n_worlds * bpw, world-major)"""
        s = np.ascontiguousarray(states, dtype=BODY_STATE_DTYPE).reshape(-1)
        if out is None:
            out = np.zeros(self.n_worlds * self.bpw, dtype=BODY_STATE_DTYPE)
        self._check(self.api.fn("batch_step_host")(self.h, _ptr(s), _ptr(out), dt, substeps),
                    "batch_step_host")
        return out

    def stats(self) -> np.ndarray:
        out = np.zeros(self.n_worlds, dtype=STATS_DTYPE)
        self._check(self.api.fn("batch_stats")(self.h, _ptr(out)), "batch_stats")
        return out

    def last_step_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self.api.fn("batch_last_step_ms")(self.h, C.byref(ms)), "batch_last_step_ms")
        return float(ms.value)


class BatchMulti:
    """n_worlds worlds spread over several GPUs by the library itself, from this one process
    (sfp_batch_create_multi): world-major arrays over ALL worlds, as for Batch."""

    def __init__(self, api: Api, n_worlds: int, bodies_per_world: int, devices, params: np.ndarray | None = None,
                 **overrides):
        self.api = api
        p = api.default_params() if params is None else np.array(params, dtype=PARAMS_DTYPE).reshape(1).copy()
        for k, v in overrides.items():
            p[k] = v
        dev = np.ascontiguousarray(devices, dtype=np.int32)
        h = C.c_void_p()
        rc = api.fn("batch_create_multi")(_ptr(p), n_worlds, bodies_per_world, _ptr(dev), len(dev), C.byref(h))
        if rc != SFP_OK:
            raise SfpError(rc, "batch_create_multi")
        self.h = h
        self.n_worlds, self.bpw, self.n_dev = n_worlds, bodies_per_world, len(dev)

    def _check(self, rc: int, what: str):
        if rc != SFP_OK:
            detail = self.api.fn("batch_multi_last_error")(self.h)
            raise SfpError(rc, what, detail.decode() if detail else "")

    def close(self):
        if getattr(self, "h", None):
            self.api.fn("batch_multi_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def partition(self):
        """[(device, first_world, n_worlds)] per block"""
        out = []
        for d in range(self.n_dev):
            dev, first, n = C.c_int32(0), C.c_uint32(0), C.c_uint32(0)
            self._check(self.api.fn("batch_multi_partition")(self.h, d, C.byref(dev), C.byref(first),
                                                             C.byref(n)), "partition")
            out.append((dev.value, first.value, n.value))
        return out

    def upload(self, bodies: np.ndarray):
        b = np.ascontiguousarray(bodies, dtype=BODY_DESC_DTYPE).reshape(-1)
        assert b.shape[0] == self.n_worlds * self.bpw
        self._check(self.api.fn("batch_multi_upload")(self.h, _ptr(b)), "batch_multi_upload")

    def step(self, dt: float = 1.0 / 60.0, substeps: int = 8):
        self._check(self.api.fn("batch_multi_step")(self.h, dt, substeps), "batch_multi_step")

    def sync(self):
        self._check(self.api.fn("batch_multi_sync")(self.h), "batch_multi_sync")

    def download(self) -> np.ndarray:
        out = np.zeros(self.n_worlds * self.bpw, dtype=BODY_STATE_DTYPE)
        self._check(self.api.fn("batch_multi_download")(self.h, _ptr(out)), "batch_multi_download")
        return out.reshape(self.n_worlds, self.bpw)

    def step_host(self, states: np.ndarray, dt: float = 1.0 / 60.0, substeps: int = 8,
                  out: np.ndarray | None = None) -> np.ndarray:
        s = np.ascontiguousarray(states, dtype=BODY_STATE_DTYPE).reshape(-1)
        if out is None:
            out = np.zeros(self.n_worlds * self.bpw, dtype=BODY_STATE_DTYPE)
        self._check(self.api.fn("batch_multi_step_host")(self.h, _ptr(s), _ptr(out), dt, substeps),
                    "batch_multi_step_host")
        return out

    def stats(self) -> np.ndarray:
        out = np.zeros(self.n_worlds, dtype=STATS_DTYPE)
        self._check(self.api.fn("batch_multi_stats")(self.h, _ptr(out)), "batch_multi_stats")
        return out

    def last_step_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self.api.fn("batch_multi_last_step_ms")(self.h, C.byref(ms)), "batch_multi_last_step_ms")
        return float(ms.value)
