"""tile solve order on the GPU against the oracle (bit for bit) on a few scenes, then its timing
against the plain colour order on pile100k (development aid)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import starframe_b200 as sf
from starframe_b200 import scenes
from oracle import loader

gpu, ref = sf.product_api(), loader.api()
F = ("x", "y", "c", "s", "vx", "vy", "omega")
def compare(name, b, j, frames, polys=None, shuffle=False):
    if shuffle:
        rng = np.random.default_rng(5)
        perm = rng.permutation(len(b)); inv = np.empty_like(perm); inv[perm] = np.arange(len(b))
        b = b[perm]
        j = j.copy(); j["body_a"] = inv[j["body_a"]]; j["body_b"] = inv[j["body_b"]]
    ws = []
    for api in (gpu, ref):
        w = sf.World(api, flags=8)
        if polys is not None: w.set_polygons(polys)
        w.upload_bodies(b); w.set_joints(j); ws.append(w)
    g, r = ws
    ok = True
    for f in range(frames):
        g.step(); r.step()
        if f % 10 == 9 or f == frames - 1:
            a, c = g.download_bodies(), r.download_bodies()
            same = all(np.array_equal(a[k].view(np.uint32), c[k].view(np.uint32)) for k in F)
            pg, pr = g.get_pairs(), r.get_pairs()
            sp = pg.shape == pr.shape and np.array_equal(pg, pr)
            cg, cr = g.get_colours(), r.get_colours()
            sc = np.array_equal(cg[1], cr[1]) and np.array_equal(cg[2], cr[2])
            sg, sr = g.stats(), r.stats()
            ss = all(sg[k] == sr[k] for k in ("n_pairs", "n_contacts", "n_colours", "colour_rounds"))
            if not (same and sp and sc and ss):
                ok = False
                print(name, "frame", f, "state", same, "pairs", sp, "colours", sc, "stats", ss,
                      {k: (sg[k], sr[k]) for k in ("n_pairs", "n_contacts", "n_colours", "colour_rounds", "overflow")})
                break
    st = g.stats()
    print(name, "shuffled" if shuffle else "", "OK" if ok else "MISMATCH", "bodies", len(b), "pairs", st["n_pairs"], "colours", st["n_colours"],
          "interior share", float(np.mean(g.get_colours()[1] >= 64)) if st["n_pairs"] else 0, flush=True)
    return ok

ok = True
ok &= compare("pile3000", *scenes.mixed_pile(3000), 60)
ok &= compare("pile3000", *scenes.mixed_pile(3000), 40, shuffle=True)
ok &= compare("circles2000", *scenes.circles_container(2000), 45)
b, j = scenes.ropes(n_ropes=40, links=32, n_free=800)
ok &= compare("ropes40", b, j, 45)
ok &= compare("ropes40", b, j, 40, shuffle=True)
b, j, polys = scenes.polygon_pile(1500)
ok &= compare("polys1500", b, j, 45, polys=polys)
ok &= compare("plank", *scenes.crowded_plank(100), 30)
print("ALL OK" if ok else "FAILED")
