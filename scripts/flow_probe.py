"""reads the SFP_FLOW_PROBE counters of one frame (development aid; library built with
-DSFP_FLOW_PROBE, selected through SFP_B200_LIB)"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene
name = sys.argv[1] if len(sys.argv) > 1 else "pile100k"
b, j = make_scene(name)
api = sf.product_api()
w = sf.World(api); w.upload_bodies(b); w.set_joints(j)
for _ in range(245): w.step()
w.sync()
fn = api.lib.sfp_debug_round_flags; fn.argtypes = [C.c_void_p, C.c_void_p]; fn.restype = C.c_int32
f = np.zeros(512, dtype=np.uint32); fn(w.h, f.ctypes.data_as(C.c_void_p))
fs = api.lib.sfp_debug_stage_ns; fs.argtypes = [C.c_void_p, C.c_void_p]; fs.restype = C.c_int32
out = np.zeros(12, dtype=np.uint64); fs(w.h, out.ctypes.data_as(C.c_void_p))
runs = int(f[488])
print("runs with work (8 position passes)", runs, "pos pass us/frame", float(out[6]) / 1e3)
seg = ["record loads", "wait", "gather+solve+store", "release"]
for i, s in enumerate(seg):
    print(f"  {s:22s} mean {16.0 * f[480 + i] / max(runs, 1):8.0f} cycles   max {int(f[484 + i])}")
print(f"  whole run mean {16.0 * f[489] / max(runs, 1):.0f} cycles")
n = int(out[0])
print("per colour of the solve order: runs (8 passes), mean wait, mean run cycles")
for oi in range(n):
    r = max(int(f[64 + oi]), 1)
    print(f"  {oi:2d}: {int(f[64 + oi]):6d} {16.0 * f[oi] / r:8.0f} {16.0 * f[128 + oi] / r:8.0f}")
