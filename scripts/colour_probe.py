import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene
b, j = make_scene("pile100k")[:2]
api = sf.product_api()
w = sf.World(api); w.upload_bodies(b); w.set_joints(j)
fs = api.lib.sfp_debug_stage_ns; fs.argtypes = [C.c_void_p, C.c_void_p]; fs.restype = C.c_int32
for f in range(250): w.step()
for f in range(8):
    w.step(); w.sync()
    out = np.zeros(12, dtype=np.uint64); fs(w.h, out.ctypes.data_as(C.c_void_p))
    st = w.stats()
    print("frame", 250 + f, "colour kernel us: init+warm", int(out[1]) / 1e3, "rounds", int(out[2]) / 1e3, "record", int(out[3]) / 1e3, "rounds#", st["colour_rounds"], "colours", st["n_colours"], w.last_step_breakdown_ms())
