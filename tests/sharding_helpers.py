"""Test-side pieces for mrt_b200.distributed.ShardedRenderer: the CPU checker behind the endpoint interface (so that the
gloo tests run the product's step logic without a GPU), an in-process collective for several ranks on ONE GPU, and the
sequential reference render."""
import contextlib
import threading

import numpy as np

import mrt_b200 as m


def sequential_oracle_image(scene, w, h, depth, seed, frames, threads=0):
    import oracle as orc
    o = orc.Oracle(max_path_length=depth, threads=threads)
    o.init_raytracing(scene)
    o.drawable_size_will_change(w, h)
    init = m.noise_tile(seed, 0xFFFFFFFF)
    for s in range(3):
        o.update_noise(s, init)
    o.draw(frames, m.noise_tiles(seed, 0, frames))
    return o.save_current_image()


class _CheckerRenderer:
    """The oracle with the method names ShardedRenderer calls on a renderer."""

    def __init__(self, scene, w, h, depth, seed):
        import oracle as orc
        self.o = orc.Oracle(max_path_length=depth, threads=1)
        self.o.init_raytracing(scene)
        self.o.drawable_size_will_change(w, h)
        self.width, self.height = w, h
        init = m.noise_tile(seed, 0xFFFFFFFF)
        for s in range(3):
            self.o.update_noise(s, init)

    def draw(self, n, tiles, first_frame=0, frame_stride=1, sync=True):
        self.o.draw(n, tiles, first_frame=first_frame, frame_stride=frame_stride)

    def update_noise(self, slot, tile): self.o.update_noise(slot, tile)
    def set_accumulate_mode(self, mode): self.o.set_accumulate_mode(mode)
    def sync(self): pass

    def set_row_shard(self, first, stride):
        self.o.set_row_shard(first, stride)
        self.first, self.stride = first, stride

    # ---- stand-in for mrt_peer_* (csrc/peer.cu): same call sequence, the exchange done with torch.distributed ----------
    fail_connect_on_rank = None          # set by a test: that rank's peer_connect raises

    def peer_export(self, sharing=0):
        self._exported = True
        return b"checker-handle"

    def peer_connect(self, rank, world, handles):
        assert getattr(self, "_exported", False) and len(handles) == world and all(h == b"checker-handle" for h in handles)
        if rank == self.fail_connect_on_rank:
            raise RuntimeError("cannot map the memory of a peer (simulated)")
        self._rank, self._world, self._connected = rank, world, True

    def peer_reduce(self, scale):
        import torch
        import torch.distributed as dist
        assert self._connected
        mine = torch.from_numpy(self.o.read_buffer(3).reshape(-1).copy())
        parts = [torch.empty_like(mine) for _ in range(self._world)]
        dist.all_gather(parts, mine)
        total = parts[0]
        for p in parts[1:]:
            total = total + p                                   # rank order, like k_peer_reduce
        out = (total * np.float32(scale)).reshape(-1, 4)
        out[:, 3] = 1.0
        self._peer_result.copy_(out.reshape(-1))

    def peer_status(self): pass

    def peer_disconnect(self):
        self._connected = False


class CheckerEndpoint:
    """Endpoint over host tensors: image copies, pack / unpack and scale in numpy (the CUDA library's versions are
    compared with these in tests/test_gpu_sharded.py)."""

    def __init__(self, scene, w, h, depth, seed):
        self.r = _CheckerRenderer(scene, w, h, depth, seed)
        self.w, self.h = w, h

    def _img(self): return self.r.o.read_buffer(3).reshape(self.h, self.w, 4)
    def _set(self, img): self.r.o.write_buffer(3, np.ascontiguousarray(img.reshape(-1, 4), dtype=np.float32))
    def stream(self): return contextlib.nullcontext()

    def new_buffer(self, floats):
        import torch
        return torch.zeros(floats, dtype=torch.float32)

    @property
    def image(self):
        import torch
        return torch.from_numpy(self._img().reshape(-1).copy())

    def image_to(self, buf): buf.copy_(self.image)

    def scale(self, buf, scale):
        v = buf.numpy().reshape(-1, 4)
        v[:, :3] *= np.float32(scale); v[:, 3] = 1.0

    def shard_layout(self, shards):
        rows = (self.h + shards - 1) // shards
        return rows, rows * self.w * 4

    def pack_rows(self, buf):
        own = self._img()[self.r.first::self.r.stride]
        buf.numpy().reshape(-1, self.w, 4)[:own.shape[0]] = own

    def unpack_rows(self, buf, shards):
        img = self._img().copy()
        blocks = buf.numpy().reshape(shards, -1, self.w, 4)
        for g in range(shards):
            n = len(range(g, self.h, shards))
            img[g::shards] = blocks[g, :n]
        self._set(img)

    def peer_result(self):
        import torch
        self.r._peer_result = torch.zeros(self.w * self.h * 4, dtype=torch.float32)
        return self.r._peer_result

    def load_noise(self, slot, tile): self.r.update_noise(slot, tile)
    def read(self, buf, host_tensor=None): return buf.numpy().reshape(self.h, self.w, 4).copy()


class LocalGroup:
    """Collective between `world` threads of one process (several contexts on ONE GPU): each rank deposits its tensor,
    rank 0 combines.  Host-synchronising by construction — a test stand-in for NCCL, nothing more."""

    def __init__(self, world):
        self.world, self.slots, self.barrier = world, [None] * world, threading.Barrier(world)

    def collective(self, rank, endpoint):
        return _LocalCollective(self, rank, endpoint)


class _LocalCollective:
    def __init__(self, group, rank, endpoint):
        self.g, self.rank, self.ep = group, rank, endpoint

    def _exchange(self, payload):
        self.ep.r.sync()                                   # the tensor is written on the ctx stream
        self.g.slots[self.rank] = payload
        self.g.barrier.wait()
        got = list(self.g.slots)
        self.g.barrier.wait()
        return got

    def all_reduce(self, t):
        import torch
        parts = [p.clone() for p in self._exchange(t)]
        self.g.barrier.wait()                              # everyone has its copies before anyone overwrites its tensor
        total = parts[0]
        for p in parts[1:]:                                # same order on every rank: identical bits
            total = total + p
        t.copy_(total)
        torch.cuda.synchronize()

    def all_gather_object(self, obj):
        return self._exchange(obj)

    def barrier(self):
        self.g.barrier.wait()

    def all_gather(self, out, inp):
        import torch
        parts = [p.clone() for p in self._exchange(inp)]
        self.g.barrier.wait()
        out.copy_(torch.cat(parts))
        torch.cuda.synchronize()
