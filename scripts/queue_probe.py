"""Per-CTA start / end times of one frame of the queue schedule (needs a library built with -DSFP_QUEUE_PROBE,
SFP_B200_LIB=var/libsfp_probe.so): how evenly do the CTAs finish?"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes
nw = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
api = sf.product_api()
fn = api.lib.sfp_debug_batch_scratch
fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
bodies = scenes.worlds(1 if os.environ.get("IDENTICAL") else 64, 512)
bodies = np.concatenate([bodies] * ((nw + len(bodies) - 1) // len(bodies)), axis=0)[:nw]
b = sf.Batch(api, nw, 512)
b.upload(bodies)
for _ in range(240): b.step()
b.sync()
b.step(); b.sync()
raw = np.zeros(4 * 8192 + 1024, dtype=np.uint32)
fn(b.h, raw.ctypes.data_as(C.c_void_p), raw.nbytes)
t = raw[:4 * 8192].view(np.uint64).reshape(-1, 2).astype(np.int64)
ncta = int(np.count_nonzero(t[:, 1]))   # CTAs of the launch (they all record their end)
t = t[:ncta]
sm = raw[4 * 8192:4 * 8192 + ncta]
t0 = t[:, 0].min()
start, end = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
dur = end - start
print("frame ms", b.last_step_ms(), "CTAs", len(dur))
print("start us  min/med/max", start.min(), np.median(start), start.max())
print("end   us  min/p10/med/p90/max", end.min(), np.percentile(end, 10), np.median(end), np.percentile(end, 90), end.max())
print("dur   us  min/p10/med/p90/max", dur.min(), np.percentile(dur, 10), np.median(dur), np.percentile(dur, 90), dur.max())
# per SM: mean duration of its CTAs
per_sm = {}
for s, d in zip(sm, dur): per_sm.setdefault(int(s), []).append(d)
m = np.array([np.mean(v) for v in per_sm.values()])
print("SMs", len(per_sm), "CTAs per SM", sorted(set(len(v) for v in per_sm.values())), "mean dur per SM min/med/max", m.min(), np.median(m), m.max())
order = np.argsort(end)
print("last 10 CTAs to finish (blockIdx, sm, end us):", [(int(i), int(sm[i]), round(float(end[i]))) for i in order[-10:]])
print("first 10:", [(int(i), int(sm[i]), round(float(end[i]))) for i in order[:10]])
