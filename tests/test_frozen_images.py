"""Frozen outputs (tests/golden/frozen_images.json, written by make_frozen_images.py): the image hashes of a few small
renders, committed.  The CPU checker must reproduce them on whatever machine the suite runs (CPU test), and the CUDA path
must reproduce THE SAME committed hashes (GPU test) — so the two are tied to a file in the history, not only to each
other: a later change that moved both together would have to change the file too.  These hashes are not the reference's
(its Metal path cannot run here): they freeze, they do not pin."""
import json
import pathlib
import runpy

import pytest

import mrt_b200 as m

GOLDEN = pathlib.Path(__file__).resolve().parent / "golden"
G = runpy.run_path(str(GOLDEN / "make_frozen_images.py"), run_name="frozen")
WANT = json.loads((GOLDEN / "frozen_images.json").read_text())
CASES = [pytest.param(c, id=c[0]) for c in G["CASES"]]


def test_every_case_has_a_committed_hash():
    assert sorted(WANT) == sorted(c[0] for c in G["CASES"]) and all(len(h) == 64 for h in WANT.values())


@pytest.mark.parametrize("case", CASES)
def test_oracle_reproduces_the_frozen_image(case):
    assert G["image_hash"](G["oracle_image"](case)) == WANT[case[0]]


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [pytest.param(m.PIPELINE_WAVEFRONT, id="wavefront"), pytest.param(m.PIPELINE_STAGED, id="staged")])
@pytest.mark.parametrize("case", CASES)
def test_cuda_path_reproduces_the_frozen_image(case, pipeline):
    name, scene_name, sseed, w, h, depth, frames, nseed, flags, spread = case
    scene = G["load_scene"](scene_name, sseed)
    r = m.Renderer(max_path_length=depth, pipeline=pipeline)
    G["setup"](r, scene, w, h, nseed)
    if flags:
        r.set_extensions(flags, spread)
    r.draw(frames, m.noise_tiles(nseed, 0, frames))
    assert G["image_hash"](r.save_current_image()) == WANT[name]
    r.close()
