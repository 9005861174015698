import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import starframe_b200 as sf
from starframe_b200 import scenes
nw = 1024
b = sf.Batch(sf.product_api(), nw, 512)
b.upload(scenes.worlds(64, 512).repeat(nw // 64, axis=0))
for _ in range(240): b.step()
b.sync()
t = 0.0
for _ in range(20):
    b.step(); b.sync(); t += b.last_step_ms()
print(os.environ.get("SFP_B200_LIB", "default"), "frame ms", t / 20)
