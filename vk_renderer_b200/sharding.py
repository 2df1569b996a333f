"""Sort-first screen sharding of one frame over the ranks of a torch.distributed group
(BASELINE config 4).

Every rank holds the whole scene.  Rank r renders rows [y0, y0+h) of the frame — `VkrFrameParams.scissor`;
the cluster test of the geometry pass drops everything outside the band before any vertex work — and rows
[s0, s0+sh) of every shadow layer (`VkrFrameParams.shadowScissor`).  Two exchange steps, both all-gathers
of row bands:

  * the shadow layers, enqueued by the shadow-exchange hook on the renderer's shadow stream between the
    shadow raster and the shading pass (so it overlaps the main-pass rasterisation);
  * the presented RGBA8 image, after the frame, optionally asynchronous so that frame i's gather overlaps
    frame i+1's rendering (the renderer double-buffers the presented image, FRAME_OVERLAP = 2).

The same class drives the CUDA library over NCCL and, in the CPU test-suite, the host emulation over gloo:
only `tensor_from_ptr` differs.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from . import capi


def row_band(rank: int, world: int, rows: int) -> tuple[int, int]:
    """Rows [y0, y0+h) owned by `rank`: equal bands of ceil(rows/world); the last ones may be short or empty."""
    per = (rows + world - 1) // world
    y0 = min(rows, rank * per)
    return y0, max(0, min(rows, y0 + per) - y0)


def cuda_tensor_from_ptr(ptr: int, nbytes: int) -> torch.Tensor:
    class _P:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}
    return torch.as_tensor(_P(), device="cuda")


def host_tensor_from_ptr(ptr: int, nbytes: int) -> torch.Tensor:
    return torch.frombuffer((C.c_ubyte * nbytes).from_address(ptr), dtype=torch.uint8)


class TileSharder:
    def __init__(self, renderer, rank: int, world: int, cuda: bool, group=None, shard_shadow: bool = True, shadow_group=None):
        """`shadow_group`: a second process group (its own communicator, hence its own NCCL stream) for the shadow
        exchange, so that frame i+1's shadow bands do not queue behind frame i's image gather.

        On CUDA the image gather is issued on torch's CURRENT stream and reads the renderer's presented image through its raw
        device pointer: it is ordered behind the frame only if the renderer submits on that same stream.  Create the renderer
        with `stream=torch.cuda.current_stream().cuda_stream` (a non-default stream: handle 0 means "own stream" to vkr_create)
        and keep that stream current while gathering; anything else is refused here rather than left to race."""
        if cuda and world > 1:
            cur = torch.cuda.current_stream().cuda_stream
            if getattr(renderer, "stream", None) in (None, 0) or renderer.stream != cur:
                raise ValueError("TileSharder: the renderer must submit on torch's current CUDA stream "
                                 "(Renderer(..., stream=torch.cuda.current_stream().cuda_stream)); see the class docstring")
        self.r, self.rank, self.world, self.cuda, self.group = renderer, rank, world, cuda, group
        self.shadow_group = shadow_group if shadow_group is not None else group
        self.W, self.H, self.S = renderer.width, renderer.height, renderer.shadow_extent
        self.y0, self.h = row_band(rank, world, self.H)
        self.rows_per = (self.H + world - 1) // world
        self.even = self.H % world == 0
        self.shard_shadow = shard_shadow and world > 1 and self.S % world == 0
        self.s0, self.sh = row_band(rank, world, self.S) if self.shard_shadow else (0, self.S)
        self._from_ptr = cuda_tensor_from_ptr if cuda else host_tensor_from_ptr
        dev = "cuda" if cuda else "cpu"
        rows = self.H if self.even else self.rows_per * world
        self.gathered = [torch.empty((rows, self.W, 4), dtype=torch.uint8, device=dev) for _ in range(2)]
        self._stage = None if self.even else torch.zeros((self.rows_per, self.W, 4), dtype=torch.uint8, device=dev)
        self._pending = [None, None]
        self._frame = 0
        if world > 1 and self.shard_shadow:
            renderer.set_shadow_exchange(self._exchange_shadow)

    # ---- per frame
    def frame_params(self, fp: capi.FrameParams) -> capi.FrameParams:
        if self.world > 1:
            fp.scissor[:] = (0, self.y0, self.W, max(self.h, 0))
            if self.h == 0:
                fp.scissor[:] = (0, 0, 1, 1)          # an empty band cannot be expressed (0 = full); render one pixel
            if self.shard_shadow:
                fp.shadowScissor[:] = (0, self.s0, self.S, self.sh)
        return fp

    def _exchange_shadow(self, shadow_ptr, layer_bytes, layers, stream):
        """All-gather the row bands of every shadow layer in place, ordered on the renderer's shadow stream."""
        full = self._from_ptr(shadow_ptr, layer_bytes * layers).view(layers, self.S, self.S * 4)
        ctx = torch.cuda.stream(torch.cuda.ExternalStream(stream)) if self.cuda else _Null()
        with ctx:
            for layer in range(layers):
                out = full[layer]
                dist.all_gather_into_tensor(out.view(-1), out[self.s0:self.s0 + self.sh].reshape(-1), group=self.shadow_group)

    def gather(self, async_op: bool = False) -> torch.Tensor:
        """All-gather the presented image of the frame just drawn.  Returns the buffer it lands in (valid once
        `wait()` has been called for it, or immediately when async_op is False)."""
        k = self._frame % 2
        self._frame += 1
        out = self.gathered[k]
        if self.world == 1:
            ptr, nbytes = self._rgba_ptr()
            out.view(-1).copy_(self._from_ptr(ptr, nbytes))
            return out
        # The renderer double-buffers the presented image: the NEXT draw writes the buffer the PREVIOUS frame's
        # gather reads, so that gather must be complete (stream-ordered) before this call returns; the gather issued
        # here then overlaps the next draw.  (Its own landing buffer was released by the same rule one frame ago.)
        self.wait()
        ptr, nbytes = self._rgba_ptr()
        img = self._from_ptr(ptr, nbytes).view(self.H, self.W, 4)
        if self.even:
            band = img[self.y0:self.y0 + self.h].reshape(-1)
        else:
            self._stage[:self.h].copy_(img[self.y0:self.y0 + self.h])
            band = self._stage.view(-1)
        work = dist.all_gather_into_tensor(out.view(-1), band, group=self.group, async_op=async_op)
        if async_op:
            self._pending[k] = work
        return out

    def wait(self):
        for k in range(2):
            if self._pending[k] is not None:
                self._pending[k].wait()
                self._pending[k] = None

    def image(self, buf: torch.Tensor) -> torch.Tensor:
        return buf[:self.H]

    def _rgba_ptr(self):
        if hasattr(self.r, "device_ptr") and self.r.prefix == "vkr":
            return self.r.device_ptr(capi.TARGET_RGBA8)
        return self.r.host_ptr(capi.TARGET_RGBA8)


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


# ----------------------------------------------------------------------------------------------------------------
# Frame-parallel multi-view rendering (BASELINE config 5)
def view_slice(rank: int, world: int, n_views: int) -> range:
    """Views owned by `rank`: view i -> rank i mod world (the camera paths of a dataset are smooth, so dealing the
    views round-robin gives every rank the same mix of cheap and expensive ones)."""
    return range(rank, n_views, world)


class ViewSharder:
    """A batch of camera views of one replicated scene, view i on rank i mod world.  Views are independent: there is
    no exchange step — every rank renders its views and reads them back to its own host memory.  On the CUDA library
    the read-back of view k overlaps the rendering of view k+1 (two pinned landing buffers = the renderer's
    FRAME_OVERLAP), on the host emulation it is a plain copy."""

    def __init__(self, renderer, rank: int, world: int, cuda: bool, rgb8: bool = False):
        """rgb8: read the views back as packed R,G,B (VKR_TARGET_RGB8, three quarters of the bytes over the bus; CUDA library
        only) instead of RGBA8."""
        self.r, self.rank, self.world, self.cuda = renderer, rank, world, cuda
        self.W, self.H = renderer.width, renderer.height
        self.channels = 3 if (rgb8 and cuda) else 4
        self.target = capi.TARGET_RGB8 if self.channels == 3 else capi.TARGET_RGBA8
        self._pinned = None
        if cuda:
            self._pinned = [torch.empty((self.H, self.W, self.channels), dtype=torch.uint8).pin_memory() for _ in range(2)]

    def render(self, scenes, objects, fp, sink=None):
        """`scenes[i]` is view i's GPUSceneData (camera matrices, light, cascades).  Calls `sink(view, image)` for every
        view of this rank, `image` an (H, W, 4) uint8 array valid during the call; returns how many views it rendered.
        Without a sink the images are still produced and read back (throughput measurements)."""
        mine = view_slice(self.rank, self.world, len(scenes))
        if not self.cuda:
            for v in mine:
                self.r.draw(scenes[v], objects, fp)
                img = self.r.readback(capi.TARGET_RGBA8)
                if sink is not None:
                    sink(v, img)
            return len(mine)
        nbytes = self.W * self.H * self.channels
        prev = None
        for k, v in enumerate(mine):
            self.r.draw(scenes[v], objects, fp)
            self.r.readback_async(self.target, self._pinned[k % 2].data_ptr(), nbytes)
            if prev is not None and sink is not None:
                self.r.sync()          # a consumer on the host serialises the pipeline; throughput runs pass no sink
                sink(prev[0], self._pinned[prev[1] % 2].numpy())
            prev = (v, k)
        self.r.sync()
        if prev is not None and sink is not None:
            sink(prev[0], self._pinned[prev[1] % 2].numpy())
        return len(mine)
