"""Dataflow solve against the barrier solve (the default; the dataflow solve is SFP_FLAG_FLOW_SOLVE) on one workload: frame time,
stage times seen by thread 0, and whether every state bit agrees (development aid).
  python scripts/flow_ab.py [workload] [settle_frames] [timed_frames]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import starframe_b200 as sf
from bench import make_scene

name = sys.argv[1] if len(sys.argv) > 1 else "pile100k"
settle = int(sys.argv[2]) if len(sys.argv) > 2 else 240
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 30
b, j = make_scene(name)
api = sf.product_api()
fn = api.lib.sfp_debug_stage_ns
fn.argtypes = [C.c_void_p, C.c_void_p]
fn.restype = C.c_int32
names = {4: "integrate", 5: "narrow", 6: "pos", 7: "derive", 8: "vel", 9: "tails", 10: "tile_pos", 11: "tile_vel"}
states = {}
for label, flags in (("flow", 4), ("barrier", 0)):
    w = sf.World(api, flags=flags)
    w.upload_bodies(b)
    w.set_joints(j)
    for _ in range(settle):
        w.step()
    w.sync()
    acc = {}
    for _ in range(frames):
        w.step()
        w.sync()
        for k, v in w.last_step_breakdown_ms().items():
            acc[k] = acc.get(k, 0) + v / frames
    out = np.zeros(12, dtype=np.uint64)
    fn(w.h, out.ctypes.data_as(C.c_void_p))
    st = w.stats()
    states[label] = w.download_bodies()
    print(label, name, {k: round(v, 4) for k, v in acc.items()},
          "colours", st["n_colours"], "overflow", st["overflow"],
          {names[k]: round(float(out[k]) / 1e3) for k in names}, flush=True)
    del w
a, c = states["flow"], states["barrier"]
same = all(np.array_equal(a[f].view(np.uint32), c[f].view(np.uint32)) for f in a.dtype.names)
print("bit-identical states after", settle + frames, "frames:", same)
