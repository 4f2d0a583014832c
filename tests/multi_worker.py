"""Worker of tests/test_gpu_multi.py (launched by torch.distributed.run, one process per GPU): a small frame over tile-row bands,
composed on rank 0 (a) by the NCCL gather (splat_gather_bands, splat_gather_bands_batch) and (b) by the fused exchange - every
rank's blend writing its band straight into rank 0's framebuffer through a CUDA IPC mapping (splat_peer_alloc / splat_peer_open) -
and compared bit for bit with the frame rank 0 renders alone and with Oracle-B."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaussian_splat_b200 import capi, multigpu, scene  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    w, h = 640, 360
    tiles_y = (h + 15) // 16
    inst, _ = scene.make_instances(20000, 0x5EED0001)
    V, P = O.app_camera(w, h)
    ctx = capi.Context(local)
    ctx.upload(inst)
    s = torch.cuda.current_stream().cuda_stream
    for mode in (capi.MODE_TILE, capi.MODE_LAYER50):
        single = torch.zeros((h, w, 4), dtype=torch.float16, device=dev)
        ctx.render(s, single.data_ptr(), V, P, w, h, capi.make_opts(mode=mode))
        torch.cuda.synchronize()
        weights = np.linspace(1.0, 3.0, tiles_y)                      # uneven bands on purpose
        bands = multigpu.band_rows(tiles_y, world, weights)
        rows_px = multigpu.bands_to_pixel_rows(bands, h)
        a, b = bands[rank]
        opts = capi.make_opts(mode=mode, tile_row_begin=a, tile_row_end=b)
        # (a) NCCL gather, one framebuffer and a batch of two
        fbs = [torch.zeros((h, w, 4), dtype=torch.float16, device=dev) for _ in range(3)]
        for f in fbs:
            if b > a:
                ctx.render(s, f.data_ptr(), V, P, w, h, opts)
        multigpu.gather_bands(fbs[0], rows_px, rank, world, dst=0)
        multigpu.gather_bands_batch(fbs[1:], rows_px, rank, world, dst=0)
        torch.cuda.synchronize(); dist.barrier()
        if rank == 0:
            for f in fbs:
                assert torch.equal(f, single), "NCCL-gathered frame differs from the single-GPU frame (mode %d)" % mode
        # (b) fused exchange: the blend writes into rank 0's memory
        peer = multigpu.PeerFrames(2, h, w, rank, world, dev, dst=0)
        for k in range(2):
            if b > a:
                ctx.render(s, peer.ptr(k), V, P, w, h, opts)
        peer.signal()
        torch.cuda.synchronize(); dist.barrier()
        if rank == 0:
            for k in range(2):
                assert torch.equal(peer.tensor(k), single), "peer-written frame differs from the single-GPU frame (mode %d)" % mode
            ref, _ = O.Frame(inst, V, P, w, h, mode=mode).blend()
            assert np.array_equal(single.cpu().numpy().view(np.uint16), ref)
        dist.barrier()
        peer.close()
    ctx.close()
    dist.barrier()
    if rank == 0:
        print("MULTI_OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
