"""mrt_peer_reduce (csrc/peer.cu): the sample-sharded framebuffer reduction as one kernel per rank over peer memory.
On this one-GPU box the "peers" are several contexts on device 0 (same kernel, same flags protocol, local stores instead of
NVLink ones; both kinds of exchange block: cudaMalloc + IPC handle, and the driver's virtual memory management); the two- and eight-GPU runs are in profiles/ (mrt_render --reduce peer, bench.py --reduce peer).
The answer is defined exactly: scale * (((img0 + img1) + img2) ...) in fp32, alpha = 1, identical on every rank."""
import numpy as np
import pytest

import mrt_b200 as m

pytestmark = pytest.mark.gpu


def make_ranks(world, w, h, seed, sharing=m.PEER_SHARE_IPC):
    scene = m.Scene.procedural("cornell", 0)
    rng = np.random.default_rng(seed)
    ranks, images = [], []
    for g in range(world):
        r = m.Renderer(device=0, max_path_length=2)
        r.init_raytracing(scene); r.drawable_size_will_change(w, h)
        img = rng.uniform(0.0, 50.0, (h, w, 4)).astype(np.float32)
        r.write_buffer(m.BUF_IMAGE, img)
        ranks.append(r); images.append(img)
    handles = [r.peer_export(sharing) for r in ranks]
    for g, r in enumerate(ranks):
        r.peer_connect(g, world, handles)
    return ranks, images


def expected(images, scale):
    total = images[0].copy()
    for img in images[1:]:
        total = total + img                                  # fp32, rank order
    out = total * np.float32(scale)
    out[..., 3] = 1.0
    return out


@pytest.mark.parametrize("sharing", [pytest.param(m.PEER_SHARE_IPC, id="cudaMalloc-block"), pytest.param(m.PEER_SHARE_PROCESS, id="vmm-block")])
@pytest.mark.parametrize("world,w,h", [(1, 17, 5), (2, 64, 36), (3, 33, 7), (5, 200, 101), (8, 320, 180), (2, 1920, 1080)])
def test_peer_reduce_is_the_rank_ordered_sum_on_every_rank(world, w, h, sharing):
    ranks, images = make_ranks(world, w, h, seed=world, sharing=sharing)
    for step, scale in enumerate((0.25, 1.0 / 3.0)):
        if step == 1:                                        # the SUM images keep growing between steps
            for g, r in enumerate(ranks):
                images[g] = images[g] + np.float32(g + 1)
                r.write_buffer(m.BUF_IMAGE, images[g])
        for r in ranks:
            r.peer_reduce(scale)                             # asynchronous: every rank's kernel is in flight before any is waited for
        want = expected(images, scale)
        for g, r in enumerate(ranks):
            got = r.peer_result()
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), f"rank {g}, step {step}"
            assert np.array_equal(r.read_buffer(m.BUF_IMAGE).reshape(h, w, 4), images[g])      # the SUM image is only read
    for r in ranks:
        r.peer_disconnect()
    for r in ranks:
        r.close()


def test_peer_reduce_argument_and_state_errors():
    scene = m.Scene.procedural("cornell", 0)
    a = m.Renderer(device=0, max_path_length=2); a.init_raytracing(scene)
    with pytest.raises(m.MrtError, match="mrt_resize"):
        a.peer_export()
    with pytest.raises(m.MrtError, match="sharing must be"):
        a.peer_export(7)
    a.drawable_size_will_change(32, 16)
    with pytest.raises(m.MrtError, match="mrt_peer_connect first"):
        a.peer_reduce(1.0)
    b = m.Renderer(device=0, max_path_length=2); b.init_raytracing(scene); b.drawable_size_will_change(32, 8)
    ha, hb = a.peer_export(), b.peer_export()
    with pytest.raises(m.MrtError, match="renders 32x8"):
        a.peer_connect(0, 2, [ha, hb])
    with pytest.raises(m.MrtError, match="not this context's handle"):
        a.peer_connect(1, 2, [ha, hb])
    with pytest.raises(m.MrtError, match="rank < world"):
        a.peer_connect(2, 2, [ha, hb])
    b.drawable_size_will_change(32, 16)                      # a resize drops the exchange block: export again
    hb = b.peer_export()
    a.peer_connect(0, 2, [ha, hb]); b.peer_connect(1, 2, [ha, hb])
    a.drawable_size_will_change(32, 16)                      # ... and the connection
    with pytest.raises(m.MrtError, match="mrt_peer_connect first"):
        a.peer_reduce(1.0)
    a.close(); b.close()


def test_a_rank_that_never_shows_up_is_reported_not_waited_for_forever():
    ranks, _ = make_ranks(2, 16, 8, seed=9)
    ranks[0].peer_reduce(1.0)                                # rank 1 never calls: both waits give up after ~2 s each
    with pytest.raises(m.MrtError, match="gave up waiting for rank 1"):
        ranks[0].peer_status()
    ranks[0].peer_status()                                   # reported once
    for r in ranks:
        r.peer_disconnect(); r.close()
