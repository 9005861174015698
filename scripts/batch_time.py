"""Frame time of the batched path against the number of worlds (512 bodies each), with the
default scheduling (worlds from a queue) and with one CTA per world (SFP_FLAG_BATCH_WHOLE_WORLDS); the two must end in
the same state bits.
  python scripts/batch_time.py [world counts ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes

counts = [int(a) for a in sys.argv[1:]] or [256, 444, 512, 1024, 2048, 8192]
api = sf.product_api()
for nw in counts:
    bodies = scenes.worlds(1 if os.environ.get("IDENTICAL") else min(nw, 64), 512)
    bodies = np.concatenate([bodies] * ((nw + len(bodies) - 1) // len(bodies)), axis=0)[:nw]
    res = {}
    for label, flags in (("queue", 0), ("whole", sf.FLAG_BATCH_WHOLE_WORLDS)):
        b = sf.Batch(api, nw, 512, flags=flags)
        b.upload(bodies)
        for _ in range(240):
            b.step()
        b.sync()
        t = []
        for _ in range(20):
            b.step(); b.sync(); t.append(b.last_step_ms())
        st = b.download()
        res[label] = (float(np.median(t)), st)
        del b
    same = all(np.array_equal(res["queue"][1][f].view(np.uint32), res["whole"][1][f].view(np.uint32))
               for f in res["queue"][1].dtype.names)
    c, w = res["queue"][0], res["whole"][0]
    print(f"worlds {nw:5d}  queue {c:7.3f} ms ({nw * 512 * 8 / c / 1e6:7.1f} G/s)  whole {w:7.3f} ms "
          f"({nw * 512 * 8 / w / 1e6:7.1f} G/s)  ratio {w / c:.3f}  same bits {same}", flush=True)
