"""CPU oracle, colour-order step with 1..N threads (SFP_REF_THREADS) on a settled pile: speed and
bit-identity of the results (development aid; the container this repository is developed in has
about one core's worth of CPU, so this is run on the GPU box's host)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes
from oracle import loader
api = loader.api()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30000
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 6
b, j = scenes.mixed_pile(n)
g = sf.World(sf.product_api()); g.upload_bodies(b)
for _ in range(240): g.step()
st0 = g.download_bodies()
res = {}
for t in (1, 2, 4, 8, 16, os.cpu_count()):
    os.environ["SFP_REF_THREADS"] = str(t)
    w = sf.World(api); w.upload_bodies(b); w.step_host(st0)
    t0 = time.time()
    for _ in range(frames): w.step()
    dt = time.time() - t0
    res[t] = w.download_bodies()
    same = all(np.array_equal(res[1][f].view(np.uint32), res[t][f].view(np.uint32)) for f in ("x", "y", "c", "s", "vx", "vy", "omega"))
    print(f"{n} bodies, {t:3d} threads: {len(b) * 8 * frames / dt / 1e6:6.2f} M body-substeps/s, same bits as 1 thread: {same}")
    del w
