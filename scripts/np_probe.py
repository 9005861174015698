"""One settled pile100k frame through the per-stage kernels (MULTI_KERNEL flag), so that ncu can
capture k_narrowphase alone (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import starframe_b200 as sf
from starframe_b200 import scenes, _ffi
b, j = scenes.mixed_pile(100_000)
w = sf.World(sf.product_api()); w.upload_bodies(b)
for _ in range(240): w.step()
st = w.download_bodies()
w2 = sf.World(sf.product_api(), flags=_ffi.FLAG_MULTI_KERNEL); w2.upload_bodies(b)
w2.step_host(st); w2.step(); w2.sync()
print("ok")
