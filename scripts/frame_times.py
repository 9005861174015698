"""per-frame stage times around a full-recolour frame (development aid)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import starframe_b200 as sf
from bench import make_scene
name = sys.argv[1] if len(sys.argv) > 1 else "pile100k"
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
b, j = make_scene(name)
w = sf.World(sf.product_api(), flags=flags); w.upload_bodies(b); w.set_joints(j)
for f in range(262):
    w.step()
    if f >= 252:
        w.sync(); br = w.last_step_breakdown_ms(); st = w.stats()
        print(f, {k: round(v, 3) for k, v in br.items()}, "rounds", st["colour_rounds"], "colours", st["n_colours"], flush=True)
