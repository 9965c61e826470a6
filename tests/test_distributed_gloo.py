"""World-size-2 test of the multi-GPU host logic (mrt_b200/distributed.py) on CPU: two processes, `gloo` backend, the
CPU checker standing in for the per-rank renderer, TWO steps (the image must keep accumulating across steps).
Sample sharding must equal the sequential render within fp32 rounding (sum-then-scale vs the running mean of
Shaders.metal:241-247); row sharding must equal it bit for bit."""
import os
import socket

import numpy as np
import pytest

import mrt_b200 as m
from mrt_b200 import distributed as D
from sharding_helpers import CheckerEndpoint, sequential_oracle_image

W, H, L, FRAMES_PER_RANK, STEPS, SEED = 48, 37, 4, 3, 2, 9      # 37 rows: the two ranks own 19 and 18 of them


def _worker(rank, world, port, mode, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ep = CheckerEndpoint(m.Scene.procedural("cornell", 0), W, H, L, SEED)
    reduce = "peer" if mode.startswith("samples-peer") else "collective"
    if mode == "samples-peer-one-rank-cannot-connect":
        ep.r.fail_connect_on_rank = 1
        with pytest.raises(RuntimeError, match="peer-memory reduction unavailable: rank 1: cannot map"):     # on EVERY rank, together
            D.ShardedRenderer(ep, rank, world, D.TorchCollective(), mode="samples", reduce="peer")
        np.save(os.path.join(out_dir, f"{mode}_{rank}_raised.npy"), np.ones(1))
        dist.barrier()
        dist.destroy_process_group()
        return
    mode = "samples" if reduce == "peer" else mode
    sr = D.ShardedRenderer(ep, rank, world, D.TorchCollective(), mode=mode, reduce=reduce)
    for step in range(STEPS):
        lead = None
        if mode == "samples":
            plan = D.plan_sample_shard(sr.next_frame, FRAMES_PER_RANK, rank, world)
            tiles = m.noise_tiles(SEED, plan.first_frame, plan.span)
            lead = {f: m.noise_tile(SEED, f) for f in D.lead_in_frames(plan.first_frame)}
        else:
            tiles = m.noise_tiles(SEED, sr.next_frame, FRAMES_PER_RANK)
        used = sr.step(FRAMES_PER_RANK, tiles, lead_in=lead)
        assert used.total_frames == (FRAMES_PER_RANK * world if mode == "samples" else FRAMES_PER_RANK)
        np.save(os.path.join(out_dir, f"{mode}{'-peer' if reduce == 'peer' else ''}_{rank}_{step}.npy"), ep.read(sr.output()))
    sr.close()                                                          # peer mode: let go of the peers' blocks, then a barrier
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("mode", ["samples", "rows", "samples-peer"])
def test_two_rank_sharding_over_gloo_two_steps(mode, tmp_path):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), mode, str(tmp_path)), nprocs=world, join=True)
    for step in range(STEPS):
        a, b = (np.load(tmp_path / f"{mode}_{k}_{step}.npy") for k in range(world))
        assert np.array_equal(a, b)                                    # every rank holds the combined image
        frames = (step + 1) * FRAMES_PER_RANK * (world if mode != "rows" else 1)
        want = sequential_oracle_image(m.Scene.procedural("cornell", 0), W, H, L, SEED, frames)
        if mode == "rows":
            assert np.array_equal(a.view(np.uint32), want.view(np.uint32)), f"step {step}"     # bit-identical
        else:
            rgb, ref = a[..., :3].astype(np.float64), want[..., :3].astype(np.float64)
            rel_rmse = np.sqrt(np.mean((rgb - ref) ** 2)) / max(np.sqrt(np.mean(ref ** 2)), 1e-12)
            assert rel_rmse <= 1e-5, (step, rel_rmse)                  # stated tolerance: fp32 re-association only
            assert (a[..., 3] == 1).all()


def test_all_ranks_fail_together_when_one_cannot_map_its_peers(tmp_path):
    """ShardedRenderer(reduce="peer"): the connect result is exchanged, so no rank is left waiting in a barrier."""
    import torch.multiprocessing as mp
    mode = "samples-peer-one-rank-cannot-connect"
    mp.spawn(_worker, args=(2, _free_port(), mode, str(tmp_path)), nprocs=2, join=True)
    assert all((tmp_path / f"{mode}_{k}_raised.npy").exists() for k in range(2))


def test_frame_plans():
    p = D.plan_sample_shard(10, 4, 1, 4)
    assert (p.first_frame, p.n_frames, p.frame_stride, p.span, p.total_frames) == (11, 4, 4, 13, 16)
    covered = sorted(f for g in range(4) for f in range(10 + g, 10 + g + 4 * 4, 4))
    assert covered == list(range(10, 26))                          # the ranks' frames tile the step exactly
    assert list(D.rows_owned(7, 2, 3)) == [2, 5]
    assert D.lead_in_frames(0) == [] and D.lead_in_frames(1) == [0] and D.lead_in_frames(9) == [7, 8]
    with pytest.raises(ValueError):
        D.plan_sample_shard(0, 1, 2, 2)
