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


def _worker(rank, world, port, out_path, split=0):
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
    part = HostCheck(sc, threads=2).render(sc.integrator.params(rank=rank, world_size=world, shard_tile=16, split=split),
                                           threads=2)
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


def test_two_rank_sample_split_combine(tmp_path):
    """AKB_SPLIT_SAMPLE: each rank renders every pixel with its share of the samples from its own seeds; the combined
    film is the sum of the per-rank films the oracle computes, carries exactly spp samples per pixel, and rank 0's
    share is the single-process render at that sample count."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from common import abi, same_bits, scenes
    from oracle_libs import Oracle
    out = str(tmp_path / "film.npy")
    mp.spawn(_worker, args=(2, _free_port(), out, 1), nprocs=2, join=True)
    sc = scenes.mixed_box((72, 40), 3, 4)
    o = Oracle(sc)
    parts = [o.render(sc.integrator.params(rank=r, world_size=2, split=abi.SPLIT_SAMPLE)) for r in range(2)]
    got = np.load(out)
    assert same_bits(got, parts[0] + parts[1])
    assert (got[..., 3] == 3.0).all() and (parts[0][..., 3] == 2.0).all() and (parts[1][..., 3] == 1.0).all()
    sc2 = scenes.mixed_box((72, 40), 2, 4)
    assert same_bits(parts[0], o.render(sc2.integrator.params()))
    assert not same_bits(got, o.render(sc.integrator.params()))  # a different, equally valid 3-sample estimate
