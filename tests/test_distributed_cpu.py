"""world_size-2 gloo test (CPU) of the multi-GPU frame split: per-rank partial films (interleaved tiles) are combined
with the package's single collective and must reproduce the single-rank film bit for bit. The per-rank films come from
tests/hostcheck (the device functions compiled for the host), so no GPU is needed."""
import os
import socket
import sys

import numpy as np
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), LOCAL_RANK=str(rank))
    import importlib
    pkg = importlib.import_module("akarirender-1_b200")
    dmod = importlib.import_module("akarirender-1_b200.distributed")
    from oracle_libs import HostCheck
    r, w = dmod.init_from_env()
    assert (r, w) == (rank, world)
    sc = pkg.scenes.mixed_box((72, 40), 3, 4)
    part = HostCheck(sc, threads=2).render(sc.integrator.params(rank=rank, world_size=world, shard_tile=16), threads=2)
    film = torch.from_numpy(part.reshape(-1).copy())
    dmod.combine_film(film)
    if rank == 0:
        np.save(out_path, film.numpy().reshape(part.shape))
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_film_combine(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from common import same_bits, scenes
    from oracle_libs import HostCheck, Oracle
    out = str(tmp_path / "film.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    sc = scenes.mixed_box((72, 40), 3, 4)
    full = Oracle(sc).render(sc.integrator.params())
    assert same_bits(np.load(out), full)
