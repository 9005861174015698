"""Reads the SFP_PHASE_PROBE build's split of the big position phases: slowest / fastest block's
work time per phase and the phase duration (cycles)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from bench import make_scene
b, j = make_scene(sys.argv[1] if len(sys.argv) > 1 else "pile100k")
w = sf.World(sf.product_api()); w.upload_bodies(b); w.set_joints(j)
for _ in range(240): w.step()
w.sync(); w.step(); w.sync()
fn = w.api.lib.sfp_debug_stage_ns; fn.argtypes = [C.c_void_p, C.c_void_p]; fn.restype = C.c_int32
out = np.zeros(12, dtype=np.uint64); fn(w.h, out.ctypes.data_as(C.c_void_p))
n = float(out[3])
print(f"big position phases probed: {int(n)}; per phase (cycles): slowest block's work {out[1]/n:.0f}, "
      f"fastest block's work {out[2]/n:.0f}, phase duration (block 0) {out[4]/n:.0f}, block 0 waits at the barrier {out[5]/n:.0f}")
fn2 = w.api.lib.sfp_debug_round_flags; fn2.argtypes = [C.c_void_p, C.c_void_p]; fn2.restype = C.c_int32
rf = np.zeros(512, dtype=np.uint32); fn2(w.h, rf.ctypes.data_as(C.c_void_p))
nb = int(out[0])
print("big colours", nb, "-> slowest / fastest block work per colour (cycles), last substep of the frame:")
last = int(n) - nb
for i in range(nb):
    print(f"  phase {i:2d}: slowest {int(rf[last + i]):6d}  fastest {int(0xffffffff - rf[256 + last + i]):6d}")
cnt = float(rf[488])
print(f"busy warps x phases: {int(cnt)}; per busy warp (cycles, mean / max):")
for k, name in enumerate(("gather", "compute", "store", "rest + prefetch issue")):
    print(f"  {name:22s} {16.0 * rf[480 + k] / cnt:7.0f} / {int(rf[484 + k]):6d}")
