"""Presentation pass (blitFragment, renderer/Shaders.metal:33-69; toSRGB, renderer/Raytracing.h:130-135) and the
accumulation checkpoint — the callers' side of the hot path.

CPU part: known answers of the oracle's restatement.  GPU part: the CUDA kernel through the C-ABI (mrt_display) against
the oracle on a rendered image — bit-exact wherever no exp/pow is involved, within ATOL/RTOL where libm (CPU) and
libdevice (GPU) evaluate exp/pow; checkpoint save -> load in a fresh context -> continue equals the uninterrupted
render bit for bit."""
import numpy as np
import pytest

import mrt_b200 as m
from helpers import bits_equal, setup_pair

# exp/pow: both libraries are within 2 ulp of the exact value.  Worst case for the absolute error: 1 - exp(-c) for tiny c
# carries the rounding of exp near 1.0 (2 ulp = 2.4e-7), the sRGB toe multiplies it by 12.92 and a comparison mode by the
# scale (10): 3e-5.
ATOL, RTOL = 4e-5, 2e-6


def _image(h=5, w=7, seed=0):
    rng = np.random.default_rng(seed)
    img = rng.uniform(0.0, 2.5, size=(h, w, 4)).astype(np.float32)
    img[..., 3] = 1.0
    img[0, 0, :3] = 0.0; img[0, 1, :3] = 1.0; img[0, 2, :3] = 0.001; img[0, 3, :3] = 37.5
    return img


def test_oracle_srgb_and_tone_known_answers():
    import oracle as orc
    f = orc.lib().orc_toSRGB
    assert f(0.0) == 0.0 and f(-3.0) == 0.0 and f(1.0) == 1.0 and f(7.0) == 1.0
    assert f(0.001) == np.float32(12.92) * np.float32(0.001)                        # linear toe below 0.0031308
    assert abs(f(0.5) - 0.735357) < 1e-5 and abs(f(0.2) - 0.484529) < 1e-5          # IEC 61966-2-1 values
    assert abs(f(0.0031308) - 12.92 * 0.0031308) < 1e-6                             # the two branches meet
    img = _image()
    out = orc.display(img)                                                          # everything off: identity
    assert bits_equal(out, img)
    tone = orc.display(img, tone_mapping=True)
    assert np.allclose(tone, 1.0 - np.exp(-img.astype(np.float64)), atol=1e-6)
    assert np.allclose(tone[..., 3], 1.0 - np.exp(-1.0), atol=1e-6)                 # alpha goes through the curve too
    both = orc.display(img, tone_mapping=True, manual_srgb=True)
    assert np.array_equal(both[..., 3], tone[..., 3])                               # sRGB touches rgb only
    assert (both[..., :3] >= tone[..., :3] - 1e-7).all()                            # the encode brightens [0, 1]


def test_oracle_comparison_modes_known_answers():
    import oracle as orc
    img = np.zeros((2, 2, 4), np.float32); ref = np.zeros((2, 2, 4), np.float32)
    img[..., 3] = 1.0; ref[..., 3] = 1.0
    img[0, 0, :3] = (0.5, 0.25, 1.0); ref[0, 0, :3] = (0.25, 0.5, 1.0)
    a = orc.display(img, comparison_mode=m.COMPARE_ABSOLUTE_VALUE, reference=ref)
    assert np.array_equal(a[0, 0], np.float32([2.5, 2.5, 0.0, 0.0]))
    b = orc.display(img, comparison_mode=m.COMPARE_REF_TO_COLOR, reference=ref)
    assert np.array_equal(b[0, 0], np.float32([0.0, 2.5, 0.0, 0.0]))                # visible where the output is darker
    c = orc.display(img, comparison_mode=m.COMPARE_COLOR_TO_REF, reference=ref, comparison_scale=4.0)
    assert np.array_equal(c[0, 0], np.float32([1.0, 0.0, 0.0, 0.0]))
    img[1, 1, :3] = (0.9, 0.9, 0.9); ref[1, 1, :3] = (0.3, 0.3, 0.3)
    d = orc.display(img, comparison_mode=m.COMPARE_LUMINANCE, reference=ref)
    assert abs(d[1, 1, 0] - 6.0) < 1e-5 and d[1, 1, 1] == 0.0 and d[1, 1, 2] == 0.0 and d[1, 1, 3] == 10.0
    assert d[0, 0, 0] == 0.0 and d[0, 0, 1] == 0.0                                  # equal luminance
    # without a reference every mode is the plain image
    assert bits_equal(orc.display(img, comparison_mode=m.COMPARE_LUMINANCE), img)
    # a reference of another size is read at the nearest texel (Shaders.metal:40-41: filter::nearest)
    big = np.zeros((4, 6, 4), np.float32); big[..., 0] = np.arange(6, dtype=np.float32)[None, :]; big[..., 1] = np.arange(4, dtype=np.float32)[:, None]
    flat = np.zeros((2, 3, 4), np.float32)
    e = orc.display(flat, comparison_mode=m.COMPARE_REF_TO_COLOR, reference=big, comparison_scale=1.0)
    assert np.array_equal(e[..., 0], np.float32([[1, 3, 5], [1, 3, 5]])) and np.array_equal(e[..., 1], np.float32([[1, 1, 1], [3, 3, 3]]))


@pytest.mark.gpu
def test_display_matches_oracle_on_a_rendered_image():
    import oracle as orc
    scene = m.Scene.procedural("cornell-water", 0)
    r, _ = setup_pair(scene, 96, 54, max_path_length=5)
    r.draw(3, m.noise_tiles(4, 0, 3))
    img = r.save_current_image()
    rng = np.random.default_rng(1)
    ref_same = (img * rng.uniform(0.5, 1.5, size=img.shape)).astype(np.float32)
    ref_small = rng.uniform(0.0, 1.0, size=(17, 31, 4)).astype(np.float32)
    for tone in (False, True):
        for srgb in (False, True):
            for mode in range(5):
                for ref in (None, ref_same, ref_small):
                    got = r.display(tone, srgb, mode, reference=ref)
                    want = orc.display(img, tone, srgb, mode, reference=ref)
                    what = f"tone={tone} srgb={srgb} mode={mode} ref={'none' if ref is None else ref.shape}"
                    if not tone and not srgb:
                        assert bits_equal(got, want), what                          # no transcendental involved
                    else:
                        assert np.allclose(got, want, atol=ATOL, rtol=RTOL), what
    assert bits_equal(r.save_current_image(), img)                                  # the pass does not touch the image
    with pytest.raises(m.MrtError):
        r.display(comparison_mode=9)


@pytest.mark.gpu
def test_checkpoint_resume_is_bit_exact(tmp_path):
    scene = m.Scene.procedural("cornell-water", 0)
    w, h, L = 80, 60, 6
    tiles = m.noise_tiles(12, 0, 7)
    whole, o = setup_pair(scene, w, h, max_path_length=L)
    whole.draw(7, tiles); o.draw(7, tiles)
    want = whole.save_current_image()
    assert bits_equal(want, o.save_current_image())
    first, _ = setup_pair(scene, w, h, max_path_length=L)
    first.draw(3, tiles[:3])
    path = tmp_path / "render.ckpt"
    first.checkpoint_save(path)
    assert path.stat().st_size == 8 + 16 + w * h * 16 + 3 * m.NOISE_TILE_FLOATS * 4
    second = m.Renderer(max_path_length=L)                                          # a fresh context: scene only
    second.init_raytracing(scene)
    assert second.checkpoint_load(path) == 3
    assert (second.width, second.height, second.frame_index) == (w, h, 3)
    assert bits_equal(second.save_current_image(), first.save_current_image())
    second.draw(4, tiles[3:])                                                       # continues at frame 3
    assert bits_equal(second.save_current_image(), want)
    bad = tmp_path / "bad.ckpt"
    bad.write_bytes(path.read_bytes()[:1000])
    with pytest.raises(m.MrtError) as e:
        second.checkpoint_load(bad)
    assert e.value.code == -4                                                       # MRT_E_IO
    with pytest.raises(m.MrtError):
        second.checkpoint_load(tmp_path / "missing.ckpt")
