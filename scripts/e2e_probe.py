"""sfp_batch_step_host round trip (pinned host arrays) of n worlds of 512 bodies: wall clock and the
library's own event time per frame.
  python scripts/e2e_probe.py [n_worlds]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes
nw = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
api = sf.product_api()
bodies = scenes.worlds(64, 512); bodies = np.concatenate([bodies] * (nw // 64), axis=0)
b = sf.Batch(api, nw, 512); b.upload(bodies)
for _ in range(200): b.step()
b.sync()
bufs = [api.alloc_host_states(nw * 512), api.alloc_host_states(nw * 512)]
bufs[0][:] = b.download().reshape(-1)
for f in range(3): b.step_host(bufs[f & 1], out=bufs[(f + 1) & 1])
t = time.perf_counter()
for f in range(3, 13): b.step_host(bufs[f & 1], out=bufs[(f + 1) & 1])
print("worlds", nw, "step_host wall ms", (time.perf_counter() - t) * 100, "event ms", b.last_step_ms())
