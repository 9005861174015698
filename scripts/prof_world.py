"""Small driver for profiling the single-world kernels: settle, then 3 more frames.
  python scripts/prof_world.py [workload] [settle_frames] [flags]   (flags 2: one kernel per stage)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import starframe_b200 as sf
from bench import make_scene
name = sys.argv[1] if len(sys.argv) > 1 else "pile100k"
settle = int(sys.argv[2]) if len(sys.argv) > 2 else 240
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
scene = make_scene(name)
b, j = scene[0], scene[1]
w = sf.World(sf.product_api(), flags=flags)
if len(scene) > 2: w.set_polygons(scene[2])
w.upload_bodies(b); w.set_joints(j)
for _ in range(settle + 3): w.step()
w.sync()
st = w.stats()
print(name, "frame ms", w.last_step_ms(), "pairs", st["n_pairs"], "entries", st["n_entries"], "contacts", st["n_contacts"])
