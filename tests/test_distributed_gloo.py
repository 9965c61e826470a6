"""World-size-2 test of the multi-GPU host logic (mrt_b200/distributed.py) on CPU: two processes, `gloo` backend, the
CPU checker standing in for the per-rank renderer.  Sample sharding must equal the sequential render within fp32
rounding (sum-then-scale vs the running mean of Shaders.metal:241-247); row sharding must equal it bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

import mrt_b200 as m
from mrt_b200 import distributed as D

W, H, L, FRAMES_PER_RANK, SEED = 48, 36, 4, 3, 9


class CheckerAdapter:
    """Gives the CPU checker the few extra methods ShardedRenderer needs (image clear / scale live in the CUDA library)."""

    def __init__(self, scene):
        import oracle as orc
        self.o = orc.Oracle(max_path_length=L, threads=1)
        self.o.init_raytracing(scene)
        self.o.drawable_size_will_change(W, H)
        init = m.noise_tile(SEED, 0xFFFFFFFF)
        for s in range(3):
            self.o.update_noise(s, init)

    def draw(self, n, tiles, first_frame=0, frame_stride=1):
        self.o.draw(n, tiles, first_frame=first_frame, frame_stride=frame_stride)

    def update_noise(self, slot, tile): self.o.update_noise(slot, tile)
    def set_accumulate_mode(self, mode): self.o.set_accumulate_mode(mode)
    def set_row_shard(self, first, stride): self.o.set_row_shard(first, stride)
    def image(self): return self.o.read_buffer(3)
    def set_image(self, img): self.o.write_buffer(3, np.ascontiguousarray(img, dtype=np.float32))
    def clear_image(self): self.set_image(np.zeros((W * H, 4), np.float32))

    def scale_image(self, s):
        img = self.image()
        img[:, :3] *= np.float32(s); img[:, 3] = 1.0
        self.set_image(img)


def _worker(rank, world, port, mode, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    scene = m.Scene.procedural("cornell", 0)
    r = CheckerAdapter(scene)

    def all_reduce(rr):
        t = torch.from_numpy(rr.image())
        dist.all_reduce(t)
        rr.set_image(t.numpy())

    sr = D.ShardedRenderer(r, rank, world, all_reduce, mode=mode)
    lead = None
    if mode == "samples":
        plan = D.plan_sample_shard(0, FRAMES_PER_RANK, rank, world)
        tiles = m.noise_tiles(SEED, plan.first_frame, plan.span)
        lead = {f: m.noise_tile(SEED, f) for f in D.lead_in_frames(plan.first_frame)}
    else:
        tiles = m.noise_tiles(SEED, 0, FRAMES_PER_RANK)
    used = sr.step(FRAMES_PER_RANK, tiles, lead_in=lead)
    np.save(os.path.join(out_dir, f"{mode}_{rank}.npy"), r.image())
    assert used.total_frames == (FRAMES_PER_RANK * world if mode == "samples" else FRAMES_PER_RANK)
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _sequential(frames):
    r = CheckerAdapter(m.Scene.procedural("cornell", 0))
    r.draw(frames, m.noise_tiles(SEED, 0, frames))
    return r.image()


@pytest.mark.parametrize("mode", ["samples", "rows"])
def test_two_rank_sharding_over_gloo(mode, tmp_path):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), mode, str(tmp_path)), nprocs=world, join=True)
    a, b = (np.load(tmp_path / f"{mode}_{k}.npy") for k in range(world))
    assert np.array_equal(a, b)                                    # every rank holds the reduced image
    if mode == "rows":
        assert np.array_equal(a.view(np.uint32), _sequential(FRAMES_PER_RANK).view(np.uint32))     # bit-identical
    else:
        want = _sequential(FRAMES_PER_RANK * world)
        rgb, ref = a[:, :3].astype(np.float64), want[:, :3].astype(np.float64)
        rel_rmse = np.sqrt(np.mean((rgb - ref) ** 2)) / max(np.sqrt(np.mean(ref ** 2)), 1e-12)
        assert rel_rmse <= 1e-5, rel_rmse                          # stated tolerance: fp32 re-association only
        assert (a[:, 3] == 1).all()


def test_frame_plans():
    p = D.plan_sample_shard(10, 4, 1, 4)
    assert (p.first_frame, p.n_frames, p.frame_stride, p.span, p.total_frames) == (11, 4, 4, 13, 16)
    covered = sorted(f for g in range(4) for f in range(10 + g, 10 + g + 4 * 4, 4))
    assert covered == list(range(10, 26))                          # the ranks' frames tile the step exactly
    assert list(D.rows_owned(7, 2, 3)) == [2, 5]
    assert D.lead_in_frames(0) == [] and D.lead_in_frames(1) == [0] and D.lead_in_frames(9) == [7, 8]
    with pytest.raises(ValueError):
        D.plan_sample_shard(0, 1, 2, 2)
