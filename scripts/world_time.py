import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene
name = sys.argv[1] if len(sys.argv) > 1 else "pile100k"
b, j = make_scene(name)
w = sf.World(sf.product_api()); w.upload_bodies(b); w.set_joints(j)
for _ in range(240): w.step()
w.sync()
acc = {}
for _ in range(30):
    w.step(); w.sync()
    for k, v in w.last_step_breakdown_ms().items(): acc[k] = acc.get(k, 0) + v / 30
fn = w.api.lib.sfp_debug_stage_ns; fn.argtypes = [C.c_void_p, C.c_void_p]; fn.restype = C.c_int32
out = np.zeros(12, dtype=np.uint64); fn(w.h, out.ctypes.data_as(C.c_void_p))
names = {4: "integrate", 5: "narrow", 6: "pos", 7: "derive", 8: "vel", 9: "tails", 10: "tile_pos", 11: "tile_vel"}
print(os.environ.get("SFP_B200_LIB", "default"), name, {k: round(v, 3) for k, v in acc.items()}, "colour kernel us (init+warm, rounds, record)", [round(float(out[k]) / 1e3, 1) for k in (1, 2, 3)], "big", int(out[0]), {names[k]: round(float(out[k]) / 1e3) for k in names})
