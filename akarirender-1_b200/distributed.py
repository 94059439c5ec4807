"""Multi-GPU frame split: one process per GPU, scene replicated, interleaved image tiles
(owner = tile_id % world_size), and ONE collective per frame to combine the partial films.

Every rank renders only the pixels it owns (`akb_render_params.rank/world_size/tile_size`); pixels it does not own
stay exactly 0 in its film, so a sum-reduce to rank 0 reproduces the single-GPU film bit for bit (x + 0 == x).
torch.distributed is plumbing here (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
import os

import torch
import torch.distributed as dist


def init_from_env(device_index=None):
    """Join the process group torchrun describes (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*). Returns (rank, world)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        use_cuda = torch.cuda.is_available()
        if use_cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")) if device_index is None else device_index)
        dist.init_process_group(backend="nccl" if use_cuda else "gloo")
    return rank, world


def combine_film(film, dst=0, group=None):
    """Sum-reduce the per-rank partial films (float32 tensor, H*W*4) onto rank `dst`. No-op for a single rank."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return film


def render_sharded(backend, film, rank, world, shard_tile=32, params=None, stream=0, split=0):
    """Render this rank's share (split 0: interleaved tiles, 1: a share of every pixel's samples; akb_split) into
    `film` (a CUDA float32 tensor of H*W*4) and combine on rank 0. Either way the combine is one sum-reduce."""
    p = params if params is not None else backend.scene.integrator.params(rank=rank, world_size=world,
                                                                          shard_tile=shard_tile)
    p.rank, p.world_size, p.tile_size, p.split = rank, world, shard_tile, split
    stats = backend.render_device(film.data_ptr(), p, stream)
    combine_film(film)
    return stats
