"""pairs that appear / disappear from frame to frame in the settled 100 k pile (what the warm-started
colouring has to colour afresh every frame): ~1.5 k of 347 k (development aid)"""
import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, starframe_b200 as sf
from bench import make_scene
b, j = make_scene("pile100k")[:2]
w = sf.World(sf.product_api()); w.upload_bodies(b); w.set_joints(j)
for _ in range(260): w.step()
prev = None
for f in range(4):
    w.step(); p = w.get_pairs()
    key = p["a"].astype(np.int64) << 32 | p["b"].astype(np.int64)
    if prev is not None:
        print("frame", 261 + f, "pairs", len(key), "new since last frame", len(np.setdiff1d(key, prev)), "gone", len(np.setdiff1d(prev, key)), "rounds", w.stats()["colour_rounds"])
    prev = key
