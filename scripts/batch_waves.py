"""Batched frame time against the number of worlds: shows the wave quantisation of one CTA per
world at 3 CTAs per SM (444 resident CTAs on 148 SMs). Development aid."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import starframe_b200 as sf
from starframe_b200 import scenes
base = scenes.worlds(64, 512)
for nw in (148, 296, 444, 592, 888, 1024, 1332):
    b = sf.Batch(sf.product_api(), nw, 512)
    b.upload(base.repeat((nw + 63) // 64, axis=0)[:nw])
    for _ in range(240): b.step()
    b.sync()
    t = 0.0
    for _ in range(10):
        b.step(); b.sync(); t += b.last_step_ms()
    print(f"worlds {nw:5d}  frame ms {t/10:.3f}  per-world us {t/10/nw*1e3:.2f}")
    del b
