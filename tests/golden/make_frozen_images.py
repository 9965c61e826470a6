#!/usr/bin/env python
"""Writes tests/golden/frozen_images.json: SHA-256 of the accumulated image the CPU checker (oracle/) renders for a few
small cases, NaNs canonicalised.  Nothing here comes from the reference (its Metal path cannot run on this box): the
hashes FREEZE today's agreed output, so that a change which moves the oracle and the CUDA path together — a shared
misreading introduced later, a compiler flag, a libm difference on another machine — shows up as a diff in a committed
file instead of passing silently.  tests/test_frozen_images.py holds the oracle (CPU) and the CUDA path (GPU) to them.
Regenerate only on purpose: python tests/golden/make_frozen_images.py"""
import hashlib
import json
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[2]
sys.path[:0] = [str(ROOT / "metal-renderer_b200"), str(ROOT / "oracle")]
import mrt_b200 as m          # noqa: E402  (host-side scene generators only)
import oracle as orc          # noqa: E402

ALL = m.EXT_REFRACTION | m.EXT_DISPERSION | m.EXT_ALIAS_LIGHTS | m.EXT_CONTINUOUS_SPECTRUM
CASES = [   # name, scene, scene seed, width, height, depth, frames, noise seed, extension flags, spread
    ("cornell-64x64-L4-f4", "cornell", 0, 64, 64, 4, 4, 1, 0, 0.0),
    ("cornell-water-96x54-L8-f3", "cornell-water", 0, 96, 54, 8, 3, 2, 0, 0.0),
    ("many-lights-64x36-L8-f2", "many-lights", 5, 64, 36, 8, 2, 3, 0, 0.0),
    ("spheres-64k-64x36-L8-f2", "spheres-64k", 3, 64, 36, 8, 2, 4, 0, 0.0),
    ("own-box-obj-64x36-L8-f2", "obj:own_box.obj", 0, 64, 36, 8, 2, 5, 0, 0.0),
    ("spheres-64k-64x36-L8-f2-optics", "spheres-64k", 3, 64, 36, 8, 2, 4, ALL, 0.02),
]


def load_scene(scene, seed):
    if scene.startswith("obj:"):
        return m.Scene.load_obj(str(pathlib.Path(__file__).resolve().parent / scene[4:]))
    return m.Scene.procedural(scene, seed)


def image_hash(img):
    a = np.ascontiguousarray(img, dtype=np.float32).copy()
    a[np.isnan(a)] = np.float32(np.nan)                    # one NaN pattern whatever produced it
    bits = a.view(np.uint32).copy()
    bits[np.isnan(a)] = 0x7FC00000
    return hashlib.sha256(bits.tobytes()).hexdigest()


def setup(renderer, scene, w, h, noise_seed):
    renderer.init_raytracing(scene); renderer.drawable_size_will_change(w, h)
    init = m.noise_tile(noise_seed, 0xFFFFFFFF)
    for s in range(3):
        renderer.update_noise(s, init)


def oracle_image(case):
    name, scene_name, sseed, w, h, depth, frames, nseed, flags, spread = case
    scene = load_scene(scene_name, sseed)
    o = orc.Oracle(max_path_length=depth)
    setup(o, scene, w, h, nseed)
    if flags:
        prob, alias = m.build_alias_table(scene.lights["pdf"][: scene.light_count])
        o.set_extensions(flags, spread, prob, alias)
    o.draw(frames, m.noise_tiles(nseed, 0, frames))
    return o.save_current_image()


if __name__ == "__main__":
    out = {c[0]: image_hash(oracle_image(c)) for c in CASES}
    path = pathlib.Path(__file__).resolve().parent / "frozen_images.json"
    path.write_text(json.dumps(out, indent=1) + "\n")
    print(f"wrote {path}: {len(out)} hashes")
