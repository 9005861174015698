"""Per-stage SM-clock cycles of the batched kernel (development aid)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes
nw = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
bpw = 512
api = sf.product_api()
b = sf.Batch(api, nw, bpw); b.upload(scenes.worlds(min(nw, 64), bpw).repeat((nw + 63) // 64, axis=0)[:nw])
for _ in range(240): b.step()
b.sync(); b.step(); b.sync()
print("frame ms", b.last_step_ms())
fn = api.lib.sfp_batch_debug_stage_cycles; fn.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]; fn.restype = C.c_int32
names = ["load/store", "broadphase", "colouring", "bucket", "integrate", "narrowphase", "solve_pos", "derive", "solve_vel", "-"]
tot = np.zeros(10)
for w in (0, 1, nw // 2, nw - 1):
    out = np.zeros(10, dtype=np.uint64); fn(b.h, w, out.ctypes.data_as(C.c_void_p)); tot += out
tot /= 4
st = b.stats()
print("pairs", st["n_pairs"][:4], "colours", st["n_colours"][:4], "rounds", st["colour_rounds"][:4])
for n, c in zip(names, tot): print(f"{n:12s} {c/1e3:10.1f} kcyc  {100*c/tot.sum():5.1f}%")
print("total kcyc", tot.sum() / 1e3, " = us at 1.9 GHz:", tot.sum() / 1.9e3)
