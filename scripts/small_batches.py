#!/usr/bin/env python
"""Device time per frame of mrt_render_frames for small batches (the reference's interactive loop renders ONE frame per
-drawInMTKView: call, Renderer.mm:587-638): n frames per call, repeated; CUDA events on the ctx stream.
usage: small_batches.py [scene seed width height]"""
import json, pathlib, sys
root = pathlib.Path(__file__).resolve().parents[1]
sys.path[:0] = [str(root / "metal-renderer_b200")]
import mrt_b200 as m

scene_name, seed, w, h = (sys.argv[1:5] + ["spheres-1m", "3", "1920", "1080"][len(sys.argv) - 1:])[:4]
seed, w, h = int(seed), int(w), int(h)
scene = m.Scene.procedural(scene_name, seed)
r = m.Renderer(max_path_length=8)
r.init_raytracing(scene); r.drawable_size_will_change(w, h)
init = m.noise_tile(seed, 0xFFFFFFFF)
for s in range(3):
    r.update_noise(s, init)
tiles = m.noise_tiles(seed, 0, 256)
out = {"scene": scene_name, "size": [w, h]}
for n in (1, 2, 4, 8, 16, 64):
    calls = max(4, 128 // n)
    frame = 0
    for _ in range(3):                                   # warm-up (first call builds the wave graph for this n)
        r.draw(n, tiles[frame % 128:frame % 128 + n], first_frame=frame, sync=False); frame += n
    r.sync(); r.reset_stats()
    r.event_record(0)
    for _ in range(calls):
        r.draw(n, tiles[frame % 128:frame % 128 + n], first_frame=frame, sync=False); frame += n
    r.event_record(1)
    ms = r.event_elapsed_ms(0, 1)
    st = r.stats()
    rays = st["primary_rays"] + st["shadow_rays"]
    out[f"n={n}"] = {"ms_per_frame": ms / (calls * n), "mrays_per_s": rays / ms / 1e3, "waves_per_call": st["waves_traced"],
                     "kernels_per_call": st["kernel_launches"] / calls}
print(json.dumps(out))
