"""colours, big colours and colour sizes over one warm-start cycle (development aid)"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene
scene = make_scene(sys.argv[1] if len(sys.argv) > 1 else "pile100k")
b, j = scene[0], scene[1]
api = sf.product_api()
w = sf.World(api)
if len(scene) > 2: w.set_polygons(scene[2])
w.upload_bodies(b); w.set_joints(j)
fs = api.lib.sfp_debug_stage_ns; fs.argtypes = [C.c_void_p, C.c_void_p]; fs.restype = C.c_int32
for f in range(300):
    w.step()
    if f >= 254 and (f - 254) % 4 == 0 or f in (256, 257):
        w.sync(); out = np.zeros(12, dtype=np.uint64); fs(w.h, out.ctypes.data_as(C.c_void_p))
        st = w.stats(); _, pc, _ = w.get_colours()
        sizes = np.bincount(pc, minlength=st["n_colours"])
        print(f, "colours", st["n_colours"], "big", int(out[0]), "rounds", st["colour_rounds"], "sizes", sizes.tolist(), "substeps ms", round(w.last_step_breakdown_ms()["substeps"], 3), flush=True)
