"""Multi-GPU host logic: one process per GPU, scene + BVH replicated, work sharded, ONE collective per step.

The reference is single-device (one MTLCommandQueue, renderer/Renderer.mm:153); every pixel-sample depends only on
(x, y, frameIndex) and the shared noise slots, so the path shards without any exchange inside a frame:

  * sample (spp) sharding — rank g of N renders frames first+g, first+g+N, ... into a SUM image
    (MRT_ACCUM_SUM), one all-reduce(sum) of the W*H*float4 image, then scale by 1/total frames.  Differs from the
    sequential running mean of Shaders.metal:241-247 only by fp32 rounding.
  * row sharding — rank g owns rows y with y % N == g and runs every frame with the reference's running mean; the
    images are combined with one all-reduce(sum) as well (rows not owned stay zero), bit-identical to one GPU.

`backend` only needs torch.distributed-style all_reduce on a tensor viewing the image; on the GPU box that is NCCL over
NVLink on the device image (no host copy), in the CPU tests it is gloo on a host copy.  The renderer is duck-typed
(draw / clear_image / scale_image / set_row_shard / set_accumulate_mode), so the same code drives the CUDA Renderer
and, inside tests/, the CPU checker.
"""
import dataclasses

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


class ShardedRenderer:
    """Drives one rank's renderer and the per-step collective.

    all_reduce(renderer) must sum the renderer's image over all ranks in place (see bench.py for the NCCL version on
    the device pointer, tests/test_distributed_gloo.py for the gloo version)."""

    def __init__(self, renderer, rank, world, all_reduce, mode="samples"):
        if mode not in ("samples", "rows"):
            raise ValueError("mode must be 'samples' or 'rows'")
        self.r, self.rank, self.world, self.all_reduce, self.mode = renderer, rank, world, all_reduce, mode
        self.next_frame = 0
        if world > 1:
            if mode == "samples":
                renderer.set_accumulate_mode(L.ACCUM_SUM)
            else:
                renderer.set_row_shard(rank, world)

    def step(self, frames_per_rank, noise_tiles=None, lead_in=None):
        """Render one step.  noise_tiles: tiles for frames next_frame+rank ... (FramePlan.span of them) in 'samples'
        mode, tiles for next_frame ... next_frame+frames_per_rank-1 in 'rows' mode.  lead_in: {frame: tile} for
        lead_in_frames(first frame of this rank) — needed in 'samples' mode, where a rank skips frames.
        Returns the FramePlan used."""
        if self.world == 1:
            plan = FramePlan(self.next_frame, frames_per_rank, 1, frames_per_rank, frames_per_rank)
            self.r.draw(frames_per_rank, noise_tiles, first_frame=plan.first_frame, frame_stride=1)
            self.next_frame += frames_per_rank
            return plan
        if self.mode == "samples":
            plan = plan_sample_shard(self.next_frame, frames_per_rank, self.rank, self.world)
            for f in lead_in_frames(plan.first_frame):
                if lead_in is not None and f in lead_in:
                    self.r.update_noise(f % 3, lead_in[f])
            self.r.clear_image()
            self.r.draw(plan.n_frames, noise_tiles, first_frame=plan.first_frame, frame_stride=plan.frame_stride)
            self.all_reduce(self.r)
            self.r.scale_image(1.0 / plan.total_frames)
            self.next_frame += plan.total_frames
        else:
            plan = FramePlan(self.next_frame, frames_per_rank, 1, frames_per_rank, frames_per_rank)
            self.r.draw(frames_per_rank, noise_tiles, first_frame=plan.first_frame, frame_stride=1)
            self.all_reduce(self.r)             # rows this rank does not own are zero
            self.next_frame += frames_per_rank
        return plan
