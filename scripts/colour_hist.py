import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene
name = sys.argv[1] if len(sys.argv) > 1 else "pile100k"
b, j = make_scene(name)
w = sf.World(sf.product_api()); w.upload_bodies(b); w.set_joints(j)
for f in range(300): w.step()
pairs, pc, jc = w.get_colours()
print("pairs", len(pairs), "colour sizes", np.bincount(pc).tolist())
c = w.get_contacts(); print("manifolds", len(c), "points", int(c["count"].sum()))
print(w.last_step_breakdown_ms())
import ctypes as C
fn = w.api.lib.sfp_debug_stage_ns; fn.argtypes = [C.c_void_p, C.c_void_p]; fn.restype = C.c_int32
out = np.zeros(12, dtype=np.uint64); fn(w.h, out.ctypes.data_as(C.c_void_p))
names = {4: "integrate", 5: "narrowphase", 6: "pos head", 7: "derive", 8: "vel head", 9: "tails"}
print("first tail colour", int(out[0]), {names[k]: round(float(out[k]) / 1e3, 1) for k in names}, "us per frame")
