import sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT / "metal-renderer_b200"), str(ROOT / "oracle"), str(ROOT / "tests")]
import numpy as np, mrt_b200 as m
from helpers import bits_equal, mismatch_report, setup_pair
scene = m.Scene.procedural("cornell", 0)
tiles = m.noise_tiles(2, 0, 3).copy()
tiles[:, 3::4][:, ::5] = 0.0
tiles[:, 2::4][:, ::7] = 0.0
L = 4
r, o = setup_pair(scene, 64, 64, max_path_length=L)
names = {0: "rays", 1: "light", 2: "isect", 3: "image"}
def cmp(tag):
    for b in (0, 1, 2, 3):
        x, y = r.read_buffer(b), o.read_buffer(b)
        if not bits_equal(x, y):
            print("MISMATCH", tag, names[b], mismatch_report(x, y) if x.dtype.fields else "")
            if x.dtype.fields:
                for nme in x.dtype.names:
                    if nme.startswith("_"): continue
                    bad = np.argwhere(~((x[nme] == y[nme]) | (np.isnan(x[nme]) & np.isnan(y[nme]))) if x[nme].dtype.kind == 'f' else x[nme] != y[nme])
                    if len(bad):
                        i = bad[0][0]; print("  idx", i, "gpu", x[i], "\n  cpu", y[i]); break
            else:
                bad = np.argwhere(~((x == y) | (np.isnan(x) & np.isnan(y)))); print("  ", bad[:5], x[bad[0][0]], y[bad[0][0]])
            return False
    return True
ok = True
for f in range(3):
    r.update_noise(f % 3, tiles[f]); o.update_noise(f % 3, tiles[f])
    r.raygen(f); o.raygen(f); ok = ok and cmp(f"f{f} raygen")
    for i in range(L):
        r.intersect(0); o.intersect(0); ok = ok and cmp(f"f{f} i{i} isect")
        r.shade(f, i); o.shade(f, i); ok = ok and cmp(f"f{f} i{i} shade")
        r.intersect(1); o.intersect(1); ok = ok and cmp(f"f{f} i{i} sisect")
        r.shadow_resolve(); o.shadow_resolve(); ok = ok and cmp(f"f{f} i{i} resolve")
        if not ok: break
    r.accumulate(f); o.accumulate(f); ok = ok and cmp(f"f{f} accum")
    if not ok: break
print("done ok=", ok)
