"""The multi-GPU step of mrt_b200/distributed.py on the CUDA renderer, on ONE GPU: two ranks = two contexts driven by
two threads, an in-process collective standing in for NCCL (tests/sharding_helpers.py).  What bench.py times at N > 1
is this very `ShardedRenderer.step` — lead-in noise slots, MRT_ACCUM_SUM, copy + reduce + mrt_scale_buffer, and for row
shards mrt_pack_rows / mrt_unpack_rows — here held against the oracle's sequential render over TWO steps:
rows mode bit for bit, samples mode within fp32 re-association (relative RMSE <= 1e-5)."""
import threading

import numpy as np
import pytest

import mrt_b200 as m
from mrt_b200 import distributed as D
from helpers import bits_equal
from sharding_helpers import LocalGroup, sequential_oracle_image

pytestmark = pytest.mark.gpu

W, H, L, SEED, STEPS = 200, 101, 6, 13, 2          # 101 rows over 3 ranks: 34 / 34 / 33 rows


def _rank(rank, world, group, mode, frames_per_rank, scene, out, errors):
    try:
        r = m.Renderer(device=0, max_path_length=L)
        r.init_raytracing(scene)
        r.drawable_size_will_change(W, H)
        init = m.noise_tile(SEED, 0xFFFFFFFF)
        for s in range(3):
            r.update_noise(s, init)
        ep = D.CudaEndpoint(r, 0)
        reduce = "peer" if mode == "samples-peer" else "collective"
        mode = "samples" if mode == "samples-peer" else mode
        sr = D.ShardedRenderer(ep, rank, world, group.collective(rank, ep), mode=mode, reduce=reduce)
        for step in range(STEPS):
            lead = None
            if mode == "samples":
                plan = D.plan_sample_shard(sr.next_frame, frames_per_rank, rank, world)
                tiles = m.noise_tiles(SEED, plan.first_frame, plan.span)
                lead = {f: m.noise_tile(SEED, f) for f in D.lead_in_frames(plan.first_frame)}
            else:
                tiles = m.noise_tiles(SEED, sr.next_frame, frames_per_rank)
            sr.step(frames_per_rank, tiles, lead_in=lead, sync=True)
            if reduce == "peer":
                r.peer_status()                             # a wait that gave up is an error, not an image
            out[(rank, step)] = ep.read(sr.output()).copy()
        group.barrier.wait()                                # nobody frees buffers a peer may still map
        r.close()
    except Exception as e:              # noqa: BLE001 — reported by the main thread
        errors.append((rank, repr(e)))
        group.barrier.abort()


@pytest.mark.parametrize("mode,world", [("rows", 2), ("rows", 3), ("samples", 2), ("samples", 3), ("samples-peer", 2), ("samples-peer", 3)])
def test_sharded_step_on_one_gpu_matches_the_sequential_oracle(mode, world):
    scene = m.Scene.procedural("cornell-water", 0)
    frames_per_rank = 3
    group, out, errors = LocalGroup(world), {}, []
    threads = [threading.Thread(target=_rank, args=(g, world, group, mode, frames_per_rank, scene, out, errors)) for g in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=600)
    assert not errors, errors
    for step in range(STEPS):
        frames = (step + 1) * frames_per_rank * (world if mode != "rows" else 1)
        want = sequential_oracle_image(scene, W, H, L, SEED, frames)
        for g in range(world):
            got = out[(g, step)]
            assert bits_equal(got, out[(0, step)]), f"rank {g} differs from rank 0 at step {step}"
        got = out[(0, step)]
        if mode == "rows":
            assert bits_equal(got, want), f"rows mode, step {step}"
        else:
            ok = np.isfinite(want[..., :3]).all(axis=-1) & np.isfinite(got[..., :3]).all(axis=-1)
            rgb, ref = got[..., :3][ok].astype(np.float64), want[..., :3][ok].astype(np.float64)
            rel_rmse = np.sqrt(np.mean((rgb - ref) ** 2)) / max(np.sqrt(np.mean(ref ** 2)), 1e-12)
            assert ok.mean() > 0.999 and rel_rmse <= 1e-5, (step, rel_rmse)
            assert (got[..., 3] == 1).all()


def test_pack_and_unpack_rows_round_trip():
    import torch
    scene = m.Scene.procedural("cornell", 0)
    r = m.Renderer(device=0, max_path_length=3)
    r.init_raytracing(scene); r.drawable_size_will_change(64, 23)
    img = np.random.default_rng(3).random((23, 64, 4), dtype=np.float32)
    r.write_buffer(m.BUF_IMAGE, img)
    shards = 4
    rows_max, nbytes = r.shard_layout(shards)
    assert rows_max == 6 and nbytes == 6 * 64 * 16
    recv = torch.zeros(shards * nbytes // 4, dtype=torch.float32, device="cuda:0")
    for g in range(shards):
        r.set_row_shard(g, shards)
        r.pack_rows(recv.data_ptr() + g * nbytes)
    r.sync()
    blocks = recv.cpu().numpy().reshape(shards, rows_max, 64, 4)
    for g in range(shards):
        own = img[g::shards]
        assert np.array_equal(blocks[g, :own.shape[0]], own)
    r.write_buffer(m.BUF_IMAGE, np.zeros_like(img))
    r.unpack_rows(recv.data_ptr(), shards)
    assert np.array_equal(r.read_buffer(m.BUF_IMAGE).reshape(23, 64, 4), img)
    r.scale_buffer(recv.data_ptr(), 10, 0.5)
    r.sync()
    head = recv.cpu().numpy()[:40].reshape(10, 4)
    assert np.array_equal(head[:, :3], blocks[0].reshape(-1, 4)[:10, :3] * np.float32(0.5)) and (head[:, 3] == 1).all()
