"""Small driver for profiling the batched frame kernel: settle, then 3 more frames.
  python scripts/prof_batch.py [n_worlds] [settle_frames] [flags]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes
nw = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
settle = int(sys.argv[2]) if len(sys.argv) > 2 else 120
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
bodies = scenes.worlds(min(nw, 64), 512)
bodies = np.concatenate([bodies] * ((nw + len(bodies) - 1) // len(bodies)), axis=0)[:nw]
b = sf.Batch(sf.product_api(), nw, 512, flags=flags)
b.upload(bodies)
for _ in range(settle + 3): b.step()
b.sync()
st = b.stats()
print("worlds", nw, "frame ms", b.last_step_ms(), "pairs/world", st["n_pairs"].mean(), "colours", st["n_colours"].mean())
import ctypes as C
fn = b.api.lib.sfp_batch_debug_stage_cycles
fn.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
names = ["load/store", "broadphase", "colouring", "bucket", "integrate", "narrowphase", "pos solve", "derive", "vel solve", "-"]
for w in (0, nw // 2):
    out = np.zeros(10, dtype=np.uint64)
    fn(b.h, w, out.ctypes.data_as(C.c_void_p))
    tot = float(out.sum())
    print("world", w, "cycles", int(tot), {n: f"{100 * float(v) / tot:.1f}%" for n, v in zip(names, out) if v})
