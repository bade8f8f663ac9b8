"""
Multi-GPU sharding of the DVR path. The reference is single-GPU; rays are independent, so a
frame or a multi-view batch shards with no exchange during rendering: one process per GPU (torchrun), the network
replicated, work partitioned by screen bands (one big frame, BASELINE C4) or by view (camera sweeps, C5), and ONE
collective per frame / batch: the finished RGBA pieces are gathered on rank 0.

Everything here is backend-agnostic (NCCL on the GPUs, gloo in the CPU tests): the partition functions are pure, the
gather works on whatever device the tensors live on, and the renderer is passed in as a callable.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

Rect = Tuple[int, int, int, int]   # x0, y0, w, h  (the tile rectangle of fvsrn_dvr_params)


def world_info() -> Tuple[int, int]:
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def partition_views(num_views: int, world: int, rank: int) -> List[int]:
    """Views of a camera sweep owned by `rank`: contiguous blocks, sizes differing by at most one."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, rem = divmod(num_views, world)
    start = rank * base + min(rank, rem)
    return list(range(start, start + base + (1 if rank < rem else 0)))


def partition_bands(width: int, height: int, world: int, rank: int, band: int = 64) -> List[Rect]:
    """Screen-tile sharding of one frame: full-width bands of `band` rows dealt round-robin, so that every rank gets
    rows from all over the image (empty rays at the borders finish early; contiguous blocks would be unbalanced)."""
    if band <= 0 or band % 8:
        raise ValueError("band height must be a positive multiple of 8 (the kernel's pixel-tile height)")
    rects = []
    for i, y0 in enumerate(range(0, height, band)):
        if i % world == rank:
            rects.append((0, y0, width, min(band, height - y0)))
    return rects


def partition_rows_by_cost(row_cost: Sequence[float], world: int, align: int = 16) -> List[Tuple[int, int]]:
    """One contiguous band of rows per rank, (y0, h) for rank 0..world-1, with (nearly) equal summed cost: rows are cut
    where the running cost passes k/world of the total, rounded to `align` rows (the kernels' pixel-tile height). A rank
    may get an empty band (h = 0) when the image has fewer aligned row groups than ranks. `row_cost` is any per-row
    estimate of the march work - scenes.estimate_row_cost sums the box-intersection lengths of a subsampled ray grid."""
    import numpy as np
    H = len(row_cost)
    c = np.maximum(np.asarray(row_cost, np.float64), 0.0) + 1e-9         # empty rows still cost their ray setup
    cum = np.concatenate([[0.0], np.cumsum(c)])
    cuts = [0]
    for k in range(1, world):
        y = int(np.searchsorted(cum, cum[-1] * k / world))
        y = int(round(y / align)) * align
        cuts.append(min(max(y, cuts[-1]), H))
    cuts.append(H)
    return [(cuts[k], cuts[k + 1] - cuts[k]) for k in range(world)]


def rebalance_rows_by_time(row_cost: Sequence[float], bands: Sequence[Tuple[int, int]], band_ms: Sequence[float], align: int = 16):
    """Feedback step for partition_rows_by_cost: `band_ms[r]` is the measured render time of rank r's band (y0, h). The estimate
    of every band is rescaled so that its rows sum to the measured time - what the estimate misses (early termination, empty
    space, TF-dependent density) is constant enough within a band - and the frame is cut again. Returns (bands, row_cost): feed the
    returned row_cost to the next call for a second iteration. Pure host arithmetic, identical on every rank given the same inputs
    (all-gather the times first)."""
    import numpy as np
    c = np.maximum(np.asarray(row_cost, np.float64), 0.0) + 1e-9
    for (y0, h), ms in zip(bands, band_ms):
        if h > 0 and ms > 0:
            c[y0:y0 + h] *= float(ms) / c[y0:y0 + h].sum()
    return partition_rows_by_cost(c, len(bands), align), c


def gather_row_bands(local: Optional[torch.Tensor], bands: Sequence[Tuple[int, int]], out: Optional[torch.Tensor], dst: int = 0,
                     async_op: bool = False):
    """The finished frame on `dst`: every rank owns one contiguous band of rows (partition_rows_by_cost) of ONE packed-pixel
    frame - (1,H,W) int32 RGBA8 or (1,H,W,4) fp16 from the fused colour output (fvsrn_dvr_params.output_format). `local` is
    this rank's band (1,h,W[,4]); `out` the full frame on `dst`, whose own band is already in place (rank `dst` renders
    straight into its slice). The other bands are received directly into their slices of `out`: no padding, no staging
    copy, one point-to-point transfer per rank. Returns the work handles when async_op (wait on them before reusing the
    buffers), else waits."""
    rank, world = world_info()
    if world == 1:
        return []
    ops = []
    if rank == dst:
        if out.shape[0] != 1:
            raise ValueError("gather_row_bands assembles one frame (batch 1): row bands of a batch are not contiguous")
        for r, (y0, h) in enumerate(bands):
            if r != dst and h > 0:
                ops.append(dist.P2POp(dist.irecv, out[:, y0:y0 + h], r))
    elif bands[rank][1] > 0:
        ops.append(dist.P2POp(dist.isend, local.contiguous(), dst))
    if not ops:
        return []
    works = dist.batch_isend_irecv(ops)
    if async_op:
        return works
    for w in works:
        w.wait()
    return []


def band_rows(rects: Sequence[Rect]) -> int:
    return sum(r[3] for r in rects)


def pack_bands(img: torch.Tensor, rects: Sequence[Rect], channels: int = 4) -> torch.Tensor:
    """(B,C>=channels,H,W) -> (B,channels,rows,W): this rank's bands stacked along y (contiguous, ready to send)."""
    if not rects:
        return img.new_zeros((img.shape[0], channels, 0, img.shape[3]))
    return torch.cat([img[:, :channels, y0:y0 + h, :] for (_, y0, _, h) in rects], dim=2).contiguous()


def gather_bands(local: torch.Tensor, width: int, height: int, band: int = 64, dst: int = 0) -> Optional[torch.Tensor]:
    """Gathers every rank's packed bands on `dst` and reassembles the (B,C,H,W) frame there (None elsewhere).
    Ranks may own different numbers of rows, so pieces are padded to the largest and sent with one gather."""
    rank, world = world_info()
    B, Cc = local.shape[0], local.shape[1]
    if world == 1:
        rects = partition_bands(width, height, 1, 0, band)
        out = local.new_empty((B, Cc, height, width)); _scatter_rows(out, local, rects)
        return out
    rows = [band_rows(partition_bands(width, height, world, r, band)) for r in range(world)]
    pad = max(rows)
    buf = local.new_zeros((B, Cc, pad, width))
    buf[:, :, :local.shape[2]] = local
    pieces = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
    dist.gather(buf, pieces, dst=dst)
    if rank != dst:
        return None
    out = local.new_empty((B, Cc, height, width))
    for r in range(world):
        _scatter_rows(out, pieces[r][:, :, :rows[r]], partition_bands(width, height, world, r, band))
    return out


def _scatter_rows(out: torch.Tensor, packed: torch.Tensor, rects: Sequence[Rect]) -> None:
    y = 0
    for (_, y0, _, h) in rects:
        out[:, :, y0:y0 + h, :] = packed[:, :, y:y + h, :]
        y += h


def gather_views(local: torch.Tensor, num_views: int, dst: int = 0, out: Optional[torch.Tensor] = None,
                 collective: str = "gather", async_op: bool = False):
    """local (v_rank,C,H,W) = this rank's views (partition_views order) -> (num_views,C,H,W) on `dst` (None elsewhere).

    When every rank owns the same number of views the pieces land directly in `out` (allocated here if None; pass a
    persistent buffer to avoid the allocation per frame) with no staging copy. `collective`:
      "gather"      ncclSend/Recv fan-in to `dst` - least traffic, but NCCL gives each peer pair only a few channels;
      "all_gather"  every rank receives the whole batch (needs `out` on every rank) - rides NCCL's ring / NVLS
                    all-gather at full NVSwitch bandwidth.
    Uneven view counts fall back to a padded gather. With async_op (equal view counts only) the collective is only enqueued and
    (result, work handle) is returned: the caller renders the next batch into another buffer meanwhile and waits on the
    handle before it reuses `local` / reads `out`."""
    rank, world = world_info()
    if world == 1:
        return local
    counts = [len(partition_views(num_views, world, r)) for r in range(world)]
    shape = (num_views,) + tuple(local.shape[1:])
    if min(counts) == max(counts):
        local = local.contiguous()
        if collective == "all_gather":
            if out is None:
                out = local.new_empty(shape)
            work = dist.all_gather_into_tensor(out, local, async_op=async_op)
            res = out if rank == dst else None
            return (res, work) if async_op else res
        if rank == dst and out is None:
            out = local.new_empty(shape)
        pieces = list(out.chunk(world, dim=0)) if rank == dst else None     # contiguous slices of `out`
        work = dist.gather(local, pieces, dst=dst, async_op=async_op)
        res = out if rank == dst else None
        return (res, work) if async_op else res
    if async_op:
        raise ValueError("async gather_views needs the same number of views on every rank")
    pad = max(counts)
    buf = local.new_zeros((pad,) + tuple(local.shape[1:]))
    buf[:local.shape[0]] = local
    pieces = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
    dist.gather(buf, pieces, dst=dst)
    if rank != dst:
        return None
    return torch.cat([pieces[r][:counts[r]] for r in range(world)], dim=0)


class _DevicePointer:
    """Raw device memory as an object torch.as_tensor understands (__cuda_array_interface__ v2)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(int(x) for x in shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class PeerFrames:
    """Fused render + gather over peer memory (NVLink / NVSwitch), one process per GPU. Rank `dst` owns `slots` frame buffers of
    `shape` (e.g. (views, H, W) int32 for RGBA8); every other rank maps them (CUDA IPC, fvsrn_peer_buffer_*) and passes
    `address(slot, index)` - the device address of element `index` along dim 0 - as the output pointer of its DVR launch, so its
    kernel's pixel writes land in rank dst's frame while the march is still running: no local image, no separate transfer.
    `signal(slot)` enqueues a 4-byte all-reduce behind the render on every rank; when it has completed on `dst`, all writers'
    kernels have finished and `tensor(slot)` (rank dst only) holds the frame. NCCL backend with peer access; the CPU (gloo)
    tests cover gather_views / gather_row_bands, which move the same pixels through the collective instead."""

    _TYPESTR = {torch.int32: "<i4", torch.float16: "<f2", torch.float32: "<f4", torch.uint8: "|u1"}

    def __init__(self, shape, dtype=torch.int32, dst: int = 0, slots: int = 2):
        import ctypes as C
        from . import _lib
        self.rank, self.world = world_info()
        self.dst, self.shape, self.dtype = dst, tuple(int(x) for x in shape), dtype
        self.itemsize = torch.empty((), dtype=dtype).element_size()
        self.stride0 = self.itemsize * int(torch.tensor(self.shape[1:]).prod().item()) if len(self.shape) > 1 else self.itemsize
        nbytes = self.stride0 * self.shape[0]
        L = _lib.lib()
        self._L, self._ptrs, self._owner = L, [], self.rank == dst
        handles = []
        if self._owner:
            for _ in range(slots):
                p = C.c_void_p(); h = C.create_string_buffer(64)
                _lib.check(L.fvsrn_peer_buffer_create(nbytes, C.byref(p), h))
                self._ptrs.append(int(p.value)); handles.append(bytes(h.raw))
        box = [handles]
        if self.world > 1:
            dist.broadcast_object_list(box, src=dst)
        if not self._owner:
            for h in box[0]:
                p = C.c_void_p()
                _lib.check(L.fvsrn_peer_buffer_open(C.create_string_buffer(h, 64), C.byref(p)))
                self._ptrs.append(int(p.value))
        self._flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        self._tensors = [torch.as_tensor(_DevicePointer(p, self.shape, self._TYPESTR[dtype]), device="cuda") for p in self._ptrs] if self._owner else None

    def address(self, slot: int, index: int = 0) -> int:
        return self._ptrs[slot] + index * self.stride0

    def tensor(self, slot: int) -> Optional[torch.Tensor]:
        return self._tensors[slot] if self._owner else None

    def signal(self, slot: int, async_op: bool = True):
        if self.world == 1:
            return None
        return dist.all_reduce(self._flag, async_op=async_op)

    def close(self):
        import ctypes as C
        from . import _lib
        if self.world > 1:
            dist.barrier()
        self._tensors = None
        for p in self._ptrs:
            _lib.check((self._L.fvsrn_peer_buffer_free if self._owner else self._L.fvsrn_peer_buffer_close)(C.c_void_p(p)))
        self._ptrs = []


def render_frame_sharded(render_rect: Callable[[Rect], torch.Tensor], width: int, height: int, band: int = 64,
                         dst: int = 0) -> Optional[torch.Tensor]:
    """One frame across all ranks. `render_rect((x0,y0,w,h))` renders that rectangle into a (B,8,H,W) (or (B,4,H,W))
    image on the local device (ImageEvaluatorSimple with `.tile` set); returns the gathered RGBA frame on `dst`."""
    rank, world = world_info()
    rects = partition_bands(width, height, world, rank, band)
    parts = []
    for r in rects:
        img = render_rect(r)
        parts.append(img[:, :4, r[1]:r[1] + r[3], :])
    local = torch.cat(parts, dim=2).contiguous() if parts else None
    if local is None:   # a rank without bands still takes part in the collective
        probe = render_rect((0, 0, width, 8))
        local = probe.new_zeros((probe.shape[0], 4, 0, width))
    return gather_bands(local, width, height, band, dst)


def render_views_sharded(render_view: Callable[[int], torch.Tensor], num_views: int, dst: int = 0) -> Optional[torch.Tensor]:
    """A camera sweep across all ranks. `render_view(i)` -> (1,C,H,W) of view i on the local device."""
    rank, world = world_info()
    mine = partition_views(num_views, world, rank)
    if not mine:
        raise ValueError("more ranks than views")
    local = torch.cat([render_view(i)[:, :4] for i in mine], dim=0).contiguous()
    return gather_views(local, num_views, dst)
