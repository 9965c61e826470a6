"""Multi-GPU host logic: one process per GPU, scene + BVH replicated, work sharded, ONE collective per step.

The reference is single-device (one MTLCommandQueue, renderer/Renderer.mm:153); every pixel-sample depends only on
(x, y, frameIndex) and the shared noise slots, so the path shards without any exchange inside a frame:

  * sample (spp) sharding — rank g of N renders frames first+g, first+g+N, ... into its own SUM image (MRT_ACCUM_SUM,
    kept across steps), then ONE all-reduce(sum) of a copy of that image and a scale by 1 / (all frames so far).  The
    result is the mean over every frame rendered since the start, whatever the number of steps; it differs from the
    sequential running mean of Shaders.metal:241-247 only by fp32 re-association.
  * row sharding — rank g owns rows y with y % N == g and runs every frame with the reference's running mean; the owned
    rows are packed, all-gathered and scattered back into every rank's image ("bit-exact mode").  The
    owned rows are never overwritten by the collective, so any number of steps stays bit-identical to one GPU.

The same `ShardedRenderer.step` is what bench.py times on NCCL, what tests/test_distributed_gloo.py runs on gloo with
the CPU checker standing in for the renderer, and what tests/test_gpu_sharded.py runs with two contexts on one GPU.
It talks to the renderer through an *endpoint* (CudaEndpoint below; the tests have their own for the checker) and to
the other ranks through a *collective* (TorchCollective = torch.distributed: NCCL over NVLink on the device buffers,
gloo on host tensors).
"""
import contextlib
import dataclasses

import numpy as np

from . import _lib as L


@dataclasses.dataclass(frozen=True)
class FramePlan:
    """Frames one rank renders in one step of `frames_per_rank` frames per rank."""
    first_frame: int        # first frame this rank renders
    n_frames: int
    frame_stride: int       # = world size
    span: int               # noise tiles needed: one per frame first_frame .. last frame (inclusive)
    total_frames: int       # frames rendered by all ranks in this step


def plan_sample_shard(step_first_frame, frames_per_rank, rank, world):
    if not (0 <= rank < world) or frames_per_rank < 0:
        raise ValueError("need 0 <= rank < world and frames_per_rank >= 0")
    span = (frames_per_rank - 1) * world + 1 if frames_per_rank else 0
    return FramePlan(step_first_frame + rank, frames_per_rank, world, span, frames_per_rank * world)


def lead_in_frames(first_frame):
    """Frames whose noise tiles must already sit in their slots when `first_frame` is rendered: the reference refreshes
    slot f % 3 at frame f (Renderer.mm:492-496) and bounce i of frame f reads slot (f + i) % 3 (:538), so frame f sees
    the tiles of f-1 and f-2 in the other two slots.  A rank that starts in the middle of the sequence loads them first."""
    return [f for f in (first_frame - 2, first_frame - 1) if f >= 0]


def rows_owned(height, rank, world):
    """Rows of a `height`-row image that rank owns under row sharding (y % world == rank)."""
    return range(rank, height, world)


class TorchCollective:
    """torch.distributed on the default process group: NCCL for device tensors, gloo for host tensors."""

    def all_reduce(self, t):
        import torch.distributed as dist
        dist.all_reduce(t)

    def all_gather(self, out, inp):
        import torch.distributed as dist
        dist.all_gather_into_tensor(out, inp)

    def all_gather_object(self, obj):
        import torch.distributed as dist
        out = [None] * dist.get_world_size()
        dist.all_gather_object(out, obj)
        return out

    def barrier(self):
        import torch.distributed as dist
        dist.barrier()


class CudaEndpoint:
    """The CUDA Renderer as ShardedRenderer sees it.  torch is plumbing here: it owns the reduce / gather buffers and
    issues the collective; every copy, pack, scale and the collective itself are enqueued on the ctx's own stream
    (mrt_stream_handle), so a step has no host synchronisation until the caller reads the result."""

    def __init__(self, renderer, device):
        import torch
        self.r, self.device, self.torch = renderer, device, torch
        self._stream = torch.cuda.ExternalStream(renderer.stream_handle(), device=torch.device("cuda", device))
        ptr, nbytes = renderer.image_device_ptr()
        h, w = renderer.height, renderer.width

        class _Img:
            __cuda_array_interface__ = {"shape": (h * w * 4,), "typestr": "<f4", "data": (ptr, False), "version": 2}
        self.image = torch.as_tensor(_Img(), device=f"cuda:{device}")      # aliases the accumulation image
        self.pixels = h * w

    def stream(self):
        return self.torch.cuda.stream(self._stream)

    def new_buffer(self, floats):
        return self.torch.empty(floats, dtype=self.torch.float32, device=f"cuda:{self.device}")

    def peer_result(self):
        """The result buffer of the peer-memory reduction (csrc/peer.cu) as a tensor aliasing it."""
        ptr, nbytes = self.r.peer_result_ptr()

        class _Buf:
            __cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4", "data": (ptr, False), "version": 2}
        return self.torch.as_tensor(_Buf(), device=f"cuda:{self.device}")

    def image_to(self, buf):
        with self.stream():
            buf.copy_(self.image, non_blocking=True)

    def scale(self, buf, scale):
        self.r.scale_buffer(buf.data_ptr(), self.pixels, scale)

    def shard_layout(self, shards):
        rows, nbytes = self.r.shard_layout(shards)
        return rows, nbytes // 4

    def pack_rows(self, buf):
        self.r.pack_rows(buf.data_ptr())

    def unpack_rows(self, buf, shards):
        self.r.unpack_rows(buf.data_ptr(), shards)

    def load_noise(self, slot, tile):
        if isinstance(tile, np.ndarray):
            self.r.update_noise(slot, tile)
        else:                               # raw host or device address; stays valid until the step is synchronised
            self.r.update_noise_from(slot, tile)

    def read(self, buf, host_tensor=None):
        """Synchronise the ctx stream and return the (H, W, 4) result as numpy (through `host_tensor` if given)."""
        with self.stream():
            if host_tensor is None:
                host_tensor = self.torch.empty(buf.shape, dtype=self.torch.float32, pin_memory=True)
            host_tensor.view(-1).copy_(buf, non_blocking=True)
        self.r.sync()
        return host_tensor.numpy().reshape(self.r.height, self.r.width, 4)


class ShardedRenderer:
    """Drives one rank's renderer and the per-step collective; see the module docstring for the two modes."""

    def __init__(self, endpoint, rank, world, collective=None, mode="samples", reduce="collective"):
        """reduce (samples mode): "collective" = copy + all-reduce (NCCL) + scale; "peer" = ONE kernel per rank that sums
        the ranks' images over NVLink peer memory (mrt_peer_reduce, csrc/peer.cu) — the collective is then only used
        once, to exchange the handles."""
        if mode not in ("samples", "rows"):
            raise ValueError("mode must be 'samples' or 'rows'")
        reduce = "collective" if reduce == "nccl" else reduce
        if reduce not in ("collective", "peer"):
            raise ValueError("reduce must be 'collective' or 'peer'")
        self.ep, self.rank, self.world, self.mode = endpoint, rank, world, mode
        self.reduce = reduce if (world > 1 and mode == "samples") else "collective"
        self.coll = collective or TorchCollective()
        self.next_frame = 0
        self.total_frames = 0
        self.result = None              # samples mode: the reduced, normalised image of all steps so far (device buffer)
        self._send = self._recv = None  # rows mode: packed rows of this rank / of every rank
        self.time_events = None         # (a, b): record the renderer's event markers a / b on the ctx stream around the collective
        if world > 1:
            if mode == "samples":
                endpoint.r.set_accumulate_mode(L.ACCUM_SUM)
                if self.reduce == "peer":
                    handles = self.coll.all_gather_object(endpoint.r.peer_export())
                    try:
                        endpoint.r.peer_connect(rank, world, handles)
                        err = None
                    except Exception as e:                  # noqa: BLE001 — every rank must reach the exchange below
                        err = f"rank {rank}: {e}"
                    # doubles as the barrier: nobody reduces before everybody has connected, and all ranks fail together
                    errs = [e for e in self.coll.all_gather_object(err) if e]
                    if errs:
                        raise RuntimeError("peer-memory reduction unavailable: " + "; ".join(errs))
                    self.result = endpoint.peer_result()
                else:
                    self.result = endpoint.new_buffer(endpoint.r.width * endpoint.r.height * 4)
            else:
                endpoint.r.set_row_shard(rank, world)
                _, floats = endpoint.shard_layout(world)
                self._send = endpoint.new_buffer(floats)
                self._recv = endpoint.new_buffer(floats * world)

    def step(self, frames_per_rank, noise_tiles=None, lead_in=None, sync=False):
        """Render one step and run its collective.  noise_tiles: tiles for frames next_frame+rank ... (FramePlan.span of
        them) in 'samples' mode, tiles for next_frame ... next_frame+frames_per_rank-1 in 'rows' mode (array or raw
        address).  lead_in: {frame: tile or address} for lead_in_frames(first frame of this rank) — needed in 'samples'
        mode, where a rank skips frames.  Asynchronous unless sync; returns the FramePlan used."""
        ep, r = self.ep, self.ep.r
        if self.world == 1:
            plan = FramePlan(self.next_frame, frames_per_rank, 1, frames_per_rank, frames_per_rank)
            r.draw(frames_per_rank, noise_tiles, first_frame=plan.first_frame, frame_stride=1, sync=False)
        elif self.mode == "samples":
            plan = plan_sample_shard(self.next_frame, frames_per_rank, self.rank, self.world)
            for f in lead_in_frames(plan.first_frame):
                if lead_in is not None and f in lead_in:
                    ep.load_noise(f % 3, lead_in[f])
            r.draw(plan.n_frames, noise_tiles, first_frame=plan.first_frame, frame_stride=plan.frame_stride, sync=False)
            if self.reduce == "peer":                     # one kernel: wait for the peers, sum over NVLink, scale, publish
                if self.time_events:
                    r.event_record(self.time_events[0])
                r.peer_reduce(1.0 / (self.total_frames + plan.total_frames))
                if self.time_events:
                    r.event_record(self.time_events[1])
            else:
                ep.image_to(self.result)                  # the rank's own sum keeps growing; the reduce works on a copy
                if self.time_events:
                    r.event_record(self.time_events[0])
                with ep.stream():
                    self.coll.all_reduce(self.result)
                if self.time_events:
                    r.event_record(self.time_events[1])
                ep.scale(self.result, 1.0 / (self.total_frames + plan.total_frames))
        else:
            plan = FramePlan(self.next_frame, frames_per_rank, 1, frames_per_rank, frames_per_rank)
            r.draw(frames_per_rank, noise_tiles, first_frame=plan.first_frame, frame_stride=1, sync=False)
            ep.pack_rows(self._send)
            if self.time_events:
                r.event_record(self.time_events[0])
            with ep.stream():
                self.coll.all_gather(self._recv, self._send)
            if self.time_events:
                r.event_record(self.time_events[1])
            ep.unpack_rows(self._recv, self.world)
        self.next_frame += plan.total_frames
        self.total_frames += plan.total_frames
        if sync:
            r.sync()
        return plan

    def close(self):
        """Peer reduction between processes: every rank lets go of its peers' exchange blocks, then a barrier, so that no
        rank frees (resize, destroy, exit) memory another process still has open."""
        if self.reduce == "peer":
            self.ep.r.peer_disconnect()
            self.coll.barrier()
            self.reduce = "closed"

    def output(self):
        """Device buffer holding the current image of the whole job: the reduced mean ('samples', world > 1) or the
        renderer's own image."""
        return self.result if (self.world > 1 and self.mode == "samples") else self.ep.image
