"""reads the SFP_TILE_PROBE counters of one frame (development aid)"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene
b, j = make_scene("pile100k")
api = sf.product_api()
w = sf.World(api, flags=8); w.upload_bodies(b); w.set_joints(j)
for _ in range(245): w.step()
w.sync()
fn = api.lib.sfp_debug_round_flags; fn.argtypes = [C.c_void_p, C.c_void_p]; fn.restype = C.c_int32
f = np.zeros(512, dtype=np.uint32); fn(w.h, f.ctypes.data_as(C.c_void_p))
n = max(int(f[488]), 1)
print("sampled active interior pairs of one tile, 8 position passes:", n)
for i, s in enumerate(("record load", "wait", "gather+solve+store")):
    print(f"  {s:20s} mean {16.0 * f[480 + i] / n:8.0f} cycles   max {int(f[484 + i])}")
print("per colour: samples, mean wait, mean solve")
for col in range(40):
    if f[64 + col]:
        print(f"  {col:2d}: {int(f[64 + col]):5d} {16.0 * f[col] / f[64 + col]:8.0f} {16.0 * f[128 + col] / f[64 + col]:8.0f}")
