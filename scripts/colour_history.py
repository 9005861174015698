"""colours and frame time over the first frames of a scene (while it settles): how far does the
warm-started colouring inflate between two colourings from scratch?
  python scripts/colour_history.py [workload] [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import starframe_b200 as sf
from bench import make_scene
scene = make_scene(sys.argv[1] if len(sys.argv) > 1 else "pile100k")
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 200
w = sf.World(sf.product_api())
if len(scene) > 2: w.set_polygons(scene[2])
w.upload_bodies(scene[0]); w.set_joints(scene[1])
row = []
for f in range(frames):
    w.step(); w.sync()
    st = w.stats()
    row.append(f"{f}:{st['n_colours']}c/{w.last_step_ms():.2f}ms/{st['n_pairs'] // 1000}kp")
    if len(row) == 8:
        print("  ".join(row), flush=True); row = []
