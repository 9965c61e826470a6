"""Host-side mirror of the reference's `Renderer` object for the path-tracing path
(reference renderer/Renderer.h:5-6, renderer/Renderer.mm).  Method names follow the reference's:

    Renderer(...)                       -initWithMetalKitView: / -loadMetalWithView:   Renderer.mm:72-154
    init_raytracing(scene)              -initRaytracing                               Renderer.mm:255-470
    drawable_size_will_change(w, h)     -mtkView:drawableSizeWillChange:              Renderer.mm:640-657
    update_noise(slot, tile)            noise part of -updateSharedData               Renderer.mm:485-497
    perform_raytracing(frame)           -performRaytracing: (34 dispatches)           Renderer.mm:500-585
    draw(n_frames, noise)               -drawInMTKView: loop, headless, fused kernel  Renderer.mm:587-638
    save_current_image()                -saveCurrentImage (a stub in the reference)   Renderer.mm:659-662

All computation happens in libmrt_b200.so on the GPU; this file only marshals arguments.
"""
import ctypes as C

import numpy as np

from . import _lib as L
from .scene import Scene


class Renderer:
    def __init__(self, device=0, max_path_length=8, accumulate_mode=L.ACCUM_MEAN, pipeline=L.PIPELINE_AUTO,
                 animate_noise=True, count_rays=True, bvh_max_leaf=4, bvh_threads=0):
        cfg = L.Config()
        L.lib().mrt_config_default(C.byref(cfg))
        cfg.device = device
        cfg.max_path_length = max_path_length
        cfg.accumulate_mode = accumulate_mode
        cfg.pipeline = pipeline
        cfg.animate_noise = int(animate_noise)
        cfg.count_rays = int(count_rays)
        cfg.bvh_max_leaf = bvh_max_leaf
        cfg.bvh_threads = bvh_threads
        self._ctx = C.c_void_p()
        L.check(L.lib().mrt_create(C.byref(cfg), C.byref(self._ctx)))
        self.width = self.height = 0
        self.frame_index = 0
        self.max_path_length = max_path_length

    def close(self):
        if getattr(self, "_ctx", None):
            L.lib().mrt_destroy(self._ctx)
            self._ctx = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- setup -----------------------------------------------------------------------------------------
    def init_raytracing(self, scene: Scene):
        L.check(L.lib().mrt_set_scene(self._ctx, L.ptr(scene.vertices), len(scene.vertices), L.ptr(scene.indices),
                                      scene.triangle_count, L.ptr(scene.references), L.ptr(scene.materials),
                                      len(scene.materials), L.ptr(scene.lights), scene.light_count))

    def drawable_size_will_change(self, width, height):
        L.check(L.lib().mrt_resize(self._ctx, width, height))
        self.width, self.height = width, height
        self.frame_index = 0

    def update_noise(self, slot, tile):
        tile = np.ascontiguousarray(tile, dtype=np.float32).reshape(-1)
        assert tile.size == L.NOISE_TILE_FLOATS
        L.check(L.lib().mrt_set_noise(self._ctx, slot, L.ptr(tile)))
        L.check(L.lib().mrt_sync(self._ctx))          # tile is borrowed only for this call

    def update_noise_from(self, slot, address):
        """Same, from a raw host or device address (asynchronous: the memory must stay valid until sync())."""
        L.check(L.lib().mrt_set_noise(self._ctx, slot, C.c_void_p(int(address))))

    def set_row_shard(self, first_row, row_stride):
        L.check(L.lib().mrt_set_row_shard(self._ctx, first_row, row_stride))

    def set_accumulate_mode(self, mode):
        L.check(L.lib().mrt_set_accumulate_mode(self._ctx, mode))

    def set_profile(self, flags):
        """flags: OR of PROFILE_TRACE_TIME (event pair per trace launch) and PROFILE_VISITS (node / triangle test counters)."""
        L.check(L.lib().mrt_set_profile(self._ctx, int(flags)))

    def set_wave_schedule(self, schedule):
        """SCHEDULE_GRAPH (default: one CUDA graph per batch, device-side stop) or SCHEDULE_HOST_LOOP."""
        L.check(L.lib().mrt_set_wave_schedule(self._ctx, schedule))

    def set_extensions(self, flags, dispersion=0.0):
        """Optics extension, beyond the reference (include/mrt.h): OR of EXT_REFRACTION, EXT_DISPERSION, EXT_ALIAS_LIGHTS;
        0 restores the reference's behaviour."""
        L.check(L.lib().mrt_set_extensions(self._ctx, int(flags), float(dispersion)))

    def set_pipeline(self, pipeline):
        L.check(L.lib().mrt_set_pipeline(self._ctx, pipeline))

    # ---- the six dispatch kinds ------------------------------------------------------------------------
    def raygen(self, frame):
        L.check(L.lib().mrt_raygen(self._ctx, frame))

    def intersect(self, which=L.BUF_RAYS):
        L.check(L.lib().mrt_intersect(self._ctx, which))

    def shade(self, frame, bounce_iteration):
        L.check(L.lib().mrt_shade(self._ctx, frame, bounce_iteration))

    def shadow_resolve(self):
        L.check(L.lib().mrt_shadow_resolve(self._ctx))

    def accumulate(self, frame):
        L.check(L.lib().mrt_accumulate(self._ctx, frame))

    def perform_raytracing(self, frame):
        L.check(L.lib().mrt_render_frame_staged(self._ctx, frame))

    # ---- production path -------------------------------------------------------------------------------
    def draw(self, n_frames, noise_tiles=None, first_frame=None, frame_stride=1, sync=True):
        """Render n_frames frames (first_frame, first_frame+stride, ...).  noise_tiles: float32 array of
        (n_frames-1)*stride+1 tiles (tile k refreshes slot (first_frame+k)%3 before that frame), or None."""
        first = self.frame_index if first_frame is None else first_frame
        p = None
        if noise_tiles is not None:
            need = ((n_frames - 1) * frame_stride + 1) * L.NOISE_TILE_FLOATS if n_frames else 0
            if isinstance(noise_tiles, np.ndarray):
                assert noise_tiles.dtype == np.float32 and noise_tiles.flags["C_CONTIGUOUS"] and noise_tiles.size >= need
                p = L.ptr(noise_tiles)
            else:                                   # raw host or device address (e.g. a torch tensor's data_ptr())
                p = C.c_void_p(int(noise_tiles))
        L.check(L.lib().mrt_render_frames(self._ctx, first, n_frames, frame_stride, p))
        if n_frames:
            self.frame_index = first + (n_frames - 1) * frame_stride + 1
        if sync:
            self.sync()

    def sync(self):
        L.check(L.lib().mrt_sync(self._ctx))

    def event_record(self, slot):
        L.check(L.lib().mrt_event_record(self._ctx, slot))

    def event_elapsed_ms(self, begin, end):
        ms = C.c_float()
        L.check(L.lib().mrt_event_elapsed_ms(self._ctx, begin, end, C.byref(ms)))
        return ms.value

    # ---- results ---------------------------------------------------------------------------------------
    def save_current_image(self, out=None):
        """RGBA32F image, shape (H, W, 4), row 0 = bottom row (as the reference's texture)."""
        img = out if out is not None else np.empty((self.height, self.width, 4), dtype=np.float32)
        L.check(L.lib().mrt_read_image(self._ctx, L.ptr(img)))
        return img

    def display(self, tone_mapping=False, manual_srgb=False, comparison_mode=L.COMPARE_DISABLED, comparison_scale=None,
                reference=None):
        """Presentation pass (blitFragment, Shaders.metal:33-69) of the accumulated image; reference = (RH, RW, 4)
        float32, row 0 = bottom.  Returns (H, W, 4) float32."""
        p = L.DisplayParams()
        L.lib().mrt_display_params_default(C.byref(p))
        p.tone_mapping, p.manual_srgb, p.comparison_mode = int(bool(tone_mapping)), int(bool(manual_srgb)), int(comparison_mode)
        if comparison_scale is not None:
            p.comparison_scale = float(comparison_scale)
        out = np.empty((self.height, self.width, 4), dtype=np.float32)
        if reference is not None:
            reference = np.ascontiguousarray(reference, dtype=np.float32)
            rh, rw = reference.shape[:2]
            L.check(L.lib().mrt_display(self._ctx, C.byref(p), L.ptr(reference), rw, rh, L.ptr(out)))
        else:
            L.check(L.lib().mrt_display(self._ctx, C.byref(p), None, 0, 0, L.ptr(out)))
        return out

    def checkpoint_save(self, path):
        """Image + size + next frame index + accumulate mode + the three noise slots (the reference's
        -saveCurrentImage is a stub, Renderer.mm:659-662)."""
        L.check(L.lib().mrt_checkpoint_save(self._ctx, str(path).encode()))

    def checkpoint_load(self, path):
        """Restore a checkpoint (resizes to the stored size); returns the frame index to continue from."""
        n = C.c_uint32()
        L.check(L.lib().mrt_checkpoint_load(self._ctx, str(path).encode(), C.byref(n)))
        w, h = C.c_uint32(), C.c_uint32()
        L.check(L.lib().mrt_get_size(self._ctx, C.byref(w), C.byref(h)))
        self.width, self.height, self.frame_index = w.value, h.value, n.value
        return n.value

    def read_image_into(self, host_address):
        L.check(L.lib().mrt_read_image(self._ctx, C.c_void_p(int(host_address))))

    def scale_image(self, scale):
        L.check(L.lib().mrt_scale_image(self._ctx, scale))

    def clear_image(self):
        L.check(L.lib().mrt_clear_image(self._ctx))

    def image_device_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        L.check(L.lib().mrt_image_device_ptr(self._ctx, C.byref(p), C.byref(n)))
        return p.value, n.value

    # ---- multi-GPU plumbing (all enqueued on the ctx stream) -------------------------------------------
    def stream_handle(self):
        """The ctx's cudaStream_t as an integer (for torch.cuda.ExternalStream / ncclAllReduce)."""
        p = C.c_void_p()
        L.check(L.lib().mrt_stream_handle(self._ctx, C.byref(p)))
        return p.value or 0

    # ---- framebuffer reduction over NVLink peer memory (csrc/peer.cu; no NCCL) ----------------------------
    def peer_export(self, sharing=L.PEER_SHARE_IPC):
        """This ctx's handle as bytes (ctypes.sizeof(PeerHandle)); give every rank the handles of all ranks.
        sharing: PEER_SHARE_IPC (ranks are processes) or PEER_SHARE_PROCESS (all ranks are contexts of this process)."""
        h = L.PeerHandle()
        L.check(L.lib().mrt_peer_export(self._ctx, sharing, C.byref(h)))
        return bytes(h)

    def peer_connect(self, rank, world, handles):
        """handles: `world` byte strings from peer_export in rank order.  Barrier between ranks before the first reduce."""
        arr = (L.PeerHandle * world)()
        for i, b in enumerate(handles):
            C.memmove(C.byref(arr[i]), bytes(b), C.sizeof(L.PeerHandle))
        L.check(L.lib().mrt_peer_connect(self._ctx, rank, world, arr))

    def peer_reduce(self, scale):
        """Asynchronous: result = scale * sum over ranks of the SUM images, in every rank's result buffer."""
        L.check(L.lib().mrt_peer_reduce(self._ctx, scale))

    def peer_result_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        L.check(L.lib().mrt_peer_result(self._ctx, C.byref(p), C.byref(n)))
        return p.value, n.value

    def peer_result(self):
        """Synchronises and returns the reduced image as (H, W, 4) float32."""
        out = np.empty((self.height, self.width, 4), dtype=np.float32)
        L.check(L.lib().mrt_peer_read_result(self._ctx, L.ptr(out)))
        return out

    def peer_status(self):
        L.check(L.lib().mrt_peer_status(self._ctx))

    def peer_disconnect(self):
        L.check(L.lib().mrt_peer_disconnect(self._ctx))

    def scale_buffer(self, device_address, pixels, scale):
        """rgb *= scale, alpha = 1 on a device buffer of `pixels` float4 (the reduced sum image)."""
        L.check(L.lib().mrt_scale_buffer(self._ctx, C.c_void_p(int(device_address)), pixels, scale))

    def shard_layout(self, shards):
        rows, nbytes = C.c_uint32(), C.c_size_t()
        L.check(L.lib().mrt_shard_layout(self._ctx, shards, C.byref(rows), C.byref(nbytes)))
        return rows.value, nbytes.value

    def pack_rows(self, device_address):
        L.check(L.lib().mrt_pack_rows(self._ctx, C.c_void_p(int(device_address))))

    def unpack_rows(self, device_address, shards):
        L.check(L.lib().mrt_unpack_rows(self._ctx, C.c_void_p(int(device_address)), shards))

    def read_buffer(self, which):
        n = self.width * self.height
        if which == L.BUF_IMAGE:
            out = np.empty((n, 4), dtype=np.float32)
        else:
            out = np.empty(n, dtype=L.BUFFER_DTYPES[which])
        L.check(L.lib().mrt_read_buffer(self._ctx, which, L.ptr(out), out.nbytes))
        return out

    def write_buffer(self, which, array):
        array = np.ascontiguousarray(array)
        L.check(L.lib().mrt_write_buffer(self._ctx, which, L.ptr(array), array.nbytes))

    def stats(self):
        s = L.Stats()
        L.check(L.lib().mrt_get_stats(self._ctx, C.byref(s)))
        return {name: getattr(s, name) for name, _ in L.Stats._fields_ if name != "reserved"}

    def reset_stats(self):
        L.check(L.lib().mrt_reset_stats(self._ctx))

    def intersect_host(self, rays, stride=None):
        """One MPSRayIntersector-style nearest-hit pass over host rays (structured array or raw bytes)."""
        rays = np.ascontiguousarray(rays)
        stride = stride or rays.dtype.itemsize
        count = rays.nbytes // stride
        out = np.empty(count, dtype=L.INTERSECTION)
        L.check(L.lib().mrt_intersect_host(self._ctx, L.ptr(rays), stride, count, L.ptr(out)))
        return out

    def intersect_host_noalloc(self, rays, stride=None, out=None):
        """intersect_host without any device allocation (static scratch array; <= NOALLOC_MAX_RAYS rays of stride
        <= NOALLOC_MAX_STRIDE per call; calls that overlap on one GPU race)."""
        rays = np.ascontiguousarray(rays)
        stride = stride or rays.dtype.itemsize
        count = rays.nbytes // stride
        out = np.empty(count, dtype=L.INTERSECTION) if out is None else out
        L.check(L.lib().mrt_intersect_host_noalloc(self._ctx, L.ptr(rays), stride, count, L.ptr(out)))
        return out

    def profile_traversal(self, rays, stride=None):
        rays = np.ascontiguousarray(rays)
        stride = stride or rays.dtype.itemsize
        L.check(L.lib().mrt_profile_traversal(self._ctx, L.ptr(rays), stride, rays.nbytes // stride))
