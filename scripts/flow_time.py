"""frame time and in-kernel stage times of the product path on one workload (development aid);
SFP_B200_LIB selects the library build"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene
name = sys.argv[1] if len(sys.argv) > 1 else "pile100k"
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
scene = make_scene(name)
b, j = scene[0], scene[1]
api = sf.product_api()
w = sf.World(api, flags=flags)
if len(scene) > 2: w.set_polygons(scene[2])
w.upload_bodies(b); w.set_joints(j)
for _ in range(240): w.step()
w.sync()
acc = {}
for _ in range(20):
    w.step(); w.sync()
    for k, v in w.last_step_breakdown_ms().items(): acc[k] = acc.get(k, 0) + v / 20
fn = api.lib.sfp_debug_stage_ns; fn.argtypes = [C.c_void_p, C.c_void_p]; fn.restype = C.c_int32
out = np.zeros(12, dtype=np.uint64); fn(w.h, out.ctypes.data_as(C.c_void_p))
names = {4: "integrate", 5: "narrow", 6: "pos", 7: "derive", 8: "vel", 9: "tails", 10: "tile_pos", 11: "tile_vel"}
st = w.stats()
print(os.path.basename(os.environ.get("SFP_B200_LIB", "default")), name, "flags", flags, {k: round(v, 4) for k, v in acc.items()},
      "overflow", st["overflow"], "pairs/body", round(st["n_pairs"] / len(b), 2), "contacts/body", round(st["n_contacts"] / len(b), 2), "colours", st["n_colours"], {names[k]: round(float(out[k]) / 1e3) for k in names}, flush=True)
