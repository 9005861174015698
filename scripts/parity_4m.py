"""One-off: a large scene (default: the 4 M-body scale point; polysN: a polygon pile), frames from a settled state, CUDA vs the CPU oracle (all
host threads) — pair set, counts and every state bit.  Too slow for the test suite (the oracle
needs ~20 s per frame at this size).  Optional third argument "shuffle": the bodies are uploaded
in a random order."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SFP_REF_THREADS"] = str(os.cpu_count())
import numpy as np
import starframe_b200 as sf
from bench import make_scene, shuffle_scene
from oracle import loader
scene = make_scene(sys.argv[1] if len(sys.argv) > 1 else "shelves4000000")
b, j = scene[0], scene[1]
polys = scene[2] if len(scene) > 2 else None
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 2
if len(sys.argv) > 3 and sys.argv[3] == "shuffle":
    b, j = shuffle_scene(b, j)
    print("bodies uploaded in a random order")
settle = sf.World(sf.product_api())
if polys is not None: settle.set_polygons(polys)
settle.upload_bodies(b)
for _ in range(200):
    settle.step()
st = settle.download_bodies()
del settle
g, r = sf.World(sf.product_api()), sf.World(loader.api())
for w in (g, r):
    if polys is not None: w.set_polygons(polys)
    w.upload_bodies(b)
t0 = time.time()
sg_, sr_ = st, st
for f in range(frames):
    sg_ = g.step_host(sg_)
    sr_ = r.step_host(sr_)
print(f"{len(b)} bodies settled 200 frames on the GPU, then {frames} frames on both sides; "
      f"oracle + GPU {time.time() - t0:.1f} s")
sg, sr = g.stats(), r.stats()
for k in ("n_entries", "n_pairs", "n_contacts", "n_manifolds", "n_colours", "colour_rounds"):
    print(k, sg[k], sr[k], "OK" if sg[k] == sr[k] else "MISMATCH")
a, c = g.download_bodies(), r.download_bodies()
bad = sum(int(np.count_nonzero(a[f].view(np.uint32) != c[f].view(np.uint32))) for f in ("x", "y", "c", "s", "vx", "vy", "omega", "angle"))
print("state words that differ:", bad)
print("pairs identical:", np.array_equal(g.get_pairs(), r.get_pairs()))
