#!/usr/bin/env python
"""Device time per frame with the optics extension (mrt_set_extensions, beyond the reference) off and on: the
extension has its own shade-kernel instantiation, so flags = 0 must cost what it cost before, and the other rows show
what refraction (longer paths inside glass), dispersion and alias light selection cost.
usage: ext_cost.py [scene seed width height frames]"""
import json, pathlib, sys
root = pathlib.Path(__file__).resolve().parents[1]
sys.path[:0] = [str(root / "metal-renderer_b200")]
import mrt_b200 as m

scene_name, seed, w, h, n = (sys.argv[1:6] + ["spheres-1m", "3", "1920", "1080", "32"][len(sys.argv) - 1:])[:5]
seed, w, h, n = int(seed), int(w), int(h), int(n)
scene = m.Scene.procedural(scene_name, seed)
r = m.Renderer(max_path_length=8)
r.init_raytracing(scene)
tiles = m.noise_tiles(seed, 0, n)
out = {"scene": scene_name, "size": [w, h], "frames": n}
R, D, A, C = m.EXT_REFRACTION, m.EXT_DISPERSION, m.EXT_ALIAS_LIGHTS, m.EXT_CONTINUOUS_SPECTRUM
for label, flags, disp in (("off", 0, 0.0), ("refraction", R, 0.0), ("dispersion", R | D, 0.02), ("alias", A, 0.0), ("spectrum", R | D | C, 0.02), ("all", R | D | A | C, 0.02), ("off again", 0, 0.0)):
    r.set_extensions(flags, disp)
    r.drawable_size_will_change(w, h)
    for slot in range(3):
        r.update_noise(slot, m.noise_tile(seed, 0xFFFFFFFF))
    for _ in range(2):
        r.draw(n, tiles, first_frame=0, sync=False)
    r.sync(); r.reset_stats()
    r.event_record(0)
    for _ in range(3):
        r.draw(n, tiles, first_frame=0, sync=False)
    r.event_record(1)
    ms = r.event_elapsed_ms(0, 1)
    st = r.stats()
    rays = st["primary_rays"] + st["shadow_rays"]
    out[label] = {"ms_per_frame": round(ms / (3 * n), 4), "mrays_per_s": round(rays / ms / 1e3, 1), "rays_per_sample": round(rays / (3.0 * n * w * h), 3)}
print(json.dumps(out))
