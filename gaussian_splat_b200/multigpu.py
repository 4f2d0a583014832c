"""Multi-GPU plumbing for the splat path: one process per GPU over torch.distributed.

* Single frame: contiguous bands of tile rows per rank, splats replicated, ONE exchange per frame
  (framebuffer bands gathered on rank 0). Every pixel's list is unchanged, so the composed frame
  is bit-identical to the single-GPU frame.
* Camera batches: views sharded across ranks, no data-path collective.
Works with the nccl backend on device tensors and with gloo on CPU tensors (the CPU tests).
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple

import numpy as np


def band_rows(tiles_y: int, world: int, weights: Sequence[float] | None = None) -> List[Tuple[int, int]]:
    """Split tile rows [0, tiles_y) into `world` contiguous bands [begin, end) (the C ABI's splat_band_rows).

    weights: per-row cost (e.g. pairs per tile row of a previous frame). Without weights rows are
    split evenly (4K: 135 rows -> 17/17/.../16 at 8 ranks). With weights the boundaries minimise the
    heaviest band (exact, by binary search on the bottleneck), never leaving a rank without a row while
    rows remain."""
    from . import capi
    if world <= 0 or tiles_y < 0:
        raise ValueError("bad arguments")
    if weights is not None:
        wt = np.asarray(weights, np.float64)
        if wt.shape != (tiles_y,) or np.any(wt < 0):
            raise ValueError("weights must be one non-negative number per tile row")
        weights = wt
    return capi.band_rows(tiles_y, world, weights)


def reweight_rows(weights: Sequence[float], bands: Sequence[Tuple[int, int]], band_times: Sequence[float],
                  damping: float = 0.7) -> np.ndarray:
    """One step of measured-time balancing: rows keep their relative weights inside a band, and each band's
    rows are rescaled so the band's weight moves towards the time that band actually took (the part of the
    frame that scales with the band: sort, ranges, blend). Pairs per row are a good first guess but not the
    cost: rows of large, near splats blend slower per pair. Feed the result to band_rows() again."""
    wt = np.asarray(weights, np.float64).copy()
    t = np.asarray(band_times, np.float64)
    if len(bands) != t.shape[0] or np.any(t < 0):
        raise ValueError("one non-negative time per band")
    total_w, total_t = wt.sum(), t.sum()
    if total_w <= 0 or total_t <= 0:
        return wt
    for (b, e), tb in zip(bands, t):
        wb = wt[b:e].sum()
        if e > b and wb > 0:
            target = tb / total_t * total_w                       # the band's share of the frame, in weight units
            wt[b:e] *= (1.0 - damping) + damping * target / wb
    return wt


def bands_to_pixel_rows(bands: Sequence[Tuple[int, int]], height: int, tile: int = 16) -> List[Tuple[int, int]]:
    return [(min(b * tile, height), min(e * tile, height)) for b, e in bands]


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """Block-sharded view indices of this rank (config 4: 64 orbit views, 8 per GPU on 8 GPUs)."""
    return list(range(n_views * rank // world, n_views * (rank + 1) // world))


def _nccl_comm_ptr(device):
    """ncclComm_t of the default process group (PyTorch's ProcessGroupNCCL), or None when the backend is not NCCL."""
    import torch.distributed as dist
    try:
        pg = dist.distributed_c10d._get_default_group()
        be = pg._get_backend(device)
        return int(be._comm_ptr())
    except Exception:      # noqa: BLE001 - gloo (CPU tests) or a PyTorch without the accessor
        return None


def gather_bands(fb, pixel_rows: Sequence[Tuple[int, int]], rank: int, world: int, dst: int = 0):
    """The path's one exchange step: every rank sends its band fb[a:b] to `dst`, which receives each
    band in place. fb: [H, W, C] tensor on every rank (same shape). On CUDA tensors under the nccl backend this is the
    C ABI's splat_gather_bands (grouped ncclSend / ncclRecv on the current stream, PyTorch's communicator); otherwise
    (gloo, the CPU tests) torch.distributed's grouped isend / irecv."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return
    if fb.is_cuda and dist.get_backend() == "nccl":
        comm = _nccl_comm_ptr(fb.device)
        if comm:
            from . import capi
            capi.gather_bands(comm, torch.cuda.current_stream(fb.device).cuda_stream, fb.data_ptr(), list(pixel_rows),
                              fb.shape[1], fb.shape[2] * fb.element_size(), rank, world, dst)
            return
    ops = []
    if rank == dst:
        for r in range(world):
            a, b = pixel_rows[r]
            if r != dst and b > a:
                ops.append(dist.P2POp(dist.irecv, fb[a:b], r))
    else:
        a, b = pixel_rows[rank]
        if b > a:
            ops.append(dist.P2POp(dist.isend, fb[a:b], dst))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()


def gather_bands_batch(fbs, pixel_rows: Sequence[Tuple[int, int]], rank: int, world: int, dst: int = 0):
    """gather_bands for several framebuffers of the same partition (consecutive frames of a pipelined loop): under nccl one
    fused exchange (the C ABI's splat_gather_bands_batch), otherwise one gather_bands per framebuffer."""
    import torch
    import torch.distributed as dist
    fbs = list(fbs)
    if world == 1 or not fbs:
        return
    if fbs[0].is_cuda and dist.get_backend() == "nccl":
        comm = _nccl_comm_ptr(fbs[0].device)
        if comm:
            from . import capi
            capi.gather_bands_batch(comm, torch.cuda.current_stream(fbs[0].device).cuda_stream, [f.data_ptr() for f in fbs],
                                    list(pixel_rows), fbs[0].shape[1], fbs[0].shape[2] * fbs[0].element_size(), rank, world, dst)
            return
    for f in fbs:
        gather_bands(f, pixel_rows, rank, world, dst)


class _DevView:
    """A device buffer that is not torch's, presented through __cuda_array_interface__ (torch.as_tensor aliases it)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class PeerFrames:
    """The banded frame's exchange as stores instead of a gather: `count` framebuffers [H, W, 4] half live on rank `dst`
    (splat_peer_alloc) and are mapped by every other rank (CUDA IPC, splat_peer_open). A rank renders its band with
    `ptr(i)` as the framebuffer, so its blend kernel writes the band straight into rank dst's memory over NVLink; nothing
    is gathered afterwards. `signal()` enqueues the completion signal on the current stream (a 4-byte all-reduce: when it
    has completed on rank dst, every rank's stream has passed its render). Call on every rank (collective)."""

    def __init__(self, count: int, height: int, width: int, rank: int, world: int, device, dst: int = 0):
        import torch
        import torch.distributed as dist
        from . import capi
        self.rank, self.world, self.dst, self.h, self.w = rank, world, dst, height, width
        self.nbytes = height * width * 8
        self._own, self._mapped = [], []
        handles = [None] * count
        if rank == dst:
            for k in range(count):
                p, hd = capi.peer_alloc(self.nbytes)
                self._own.append(p); handles[k] = hd
        if world > 1:
            dist.broadcast_object_list(handles, src=dst)
        if rank == dst:
            self._ptrs = list(self._own)
        else:
            self._mapped = [capi.peer_open(hd) for hd in handles]
            self._ptrs = list(self._mapped)
        self._flag = torch.zeros(1, dtype=torch.int32, device=device)

    def ptr(self, i: int) -> int:
        return self._ptrs[i]

    def tensor(self, i: int):
        """Rank dst only: framebuffer i as a torch tensor [H, W, 4] float16 aliasing the device memory."""
        import torch
        assert self.rank == self.dst
        return torch.as_tensor(_DevView(self._own[i], (self.h, self.w, 4), "<f2"), device=self._flag.device)

    def signal(self):
        import torch.distributed as dist
        if self.world > 1:
            dist.all_reduce(self._flag)

    def close(self):
        from . import capi
        for p in self._mapped:
            capi.peer_close(p)
        for p in self._own:
            capi.peer_free(p)
        self._mapped, self._own, self._ptrs = [], [], []


class SharedHostFrame:
    """One host framebuffer shared by every rank of the node: a POSIX shared-memory file mapped by all processes
    and page-locked in each (cudaHostRegister), so a rank's device-to-host copy lands its band in the final frame
    directly, over its own PCIe link. Delivering a frame to the HOST this way needs no collective at all (the NCCL
    gather is for a frame that has to end up on one DEVICE): 8 ranks move 1/8 of the 66 MB 4K frame each, in
    parallel, instead of rank 0 moving all of it after the gather.

    Call on every rank with the same name (a barrier follows the creation). `tensor` is [H, W, C] (float16 for the
    half4 framebuffer, uint8 for BGRA8)."""

    def __init__(self, name: str, height: int, width: int, rank: int, world: int, dtype="float16", channels: int = 4,
                 register: bool = True):
        import torch
        import torch.distributed as dist
        self.path = os.path.join("/dev/shm", name)
        self.rank = rank
        np_dtype = np.dtype(dtype)
        nbytes = height * width * channels * np_dtype.itemsize
        if rank == 0:
            with open(self.path, "wb") as f:
                f.truncate(nbytes)
        if world > 1:
            dist.barrier()
        self._map = np.memmap(self.path, dtype=np_dtype, mode="r+", shape=(height, width, channels))
        self.tensor = torch.from_numpy(self._map)
        self._registered = False
        if register and torch.cuda.is_available():
            rc = torch.cuda.cudart().cudaHostRegister(self.tensor.data_ptr(), nbytes, 0)
            if int(rc) != 0:
                raise RuntimeError(f"cudaHostRegister failed ({rc})")
            self._registered = True
        if world > 1:
            dist.barrier()

    def deliver(self, fb, pixel_rows: Tuple[int, int]):
        """Enqueue the copy of this rank's band fb[a:b] (device) into the shared frame on the current stream."""
        a, b = pixel_rows
        if b > a:
            self.tensor[a:b].copy_(fb[a:b], non_blocking=True)

    def close(self):
        import torch
        if self._registered:
            torch.cuda.cudart().cudaHostUnregister(self.tensor.data_ptr())
            self._registered = False
        self.tensor = None
        self._map = None
        if self.rank == 0 and os.path.exists(self.path):
            os.unlink(self.path)


def bind_to_gpu_numa_node(device_index: int) -> dict:
    """Pin this process to the CPU cores of the NUMA node its GPU hangs off, BEFORE any page-locked host buffer is
    allocated (first-touch places the pages on that node). With one process per GPU and framebuffers streaming
    to host memory, buffers on the far socket put every copy on the inter-socket link. Best effort: returns what
    it found and did; never raises (containers may hide sysfs)."""
    info = {"device": device_index, "bound": False}
    try:
        import torch
        pr = torch.cuda.get_device_properties(device_index)
        bus = "%04x:%02x:%02x.0" % (getattr(pr, "pci_domain_id", 0), pr.pci_bus_id, pr.pci_device_id)
        info["pci"] = bus
        with open("/sys/bus/pci/devices/%s/numa_node" % bus) as f:
            node = int(f.read().strip())
        info["numa_node"] = node
        if node < 0:
            return info
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if use:
            os.sched_setaffinity(0, use)
            info["bound"] = True
            info["cpus"] = len(use)
    except Exception as e:      # noqa: BLE001 - diagnostic only
        info["error"] = repr(e)
    return info
