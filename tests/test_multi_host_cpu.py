"""world_size-2 gloo test (CPU) of the host-side part of the multi-GPU path: the slab
partition every rank applies to decide which atoms it uploads, and the id broadcast pattern
bench.py uses for the NCCL unique id."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from imd_b200 import host, synth
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    obj = [os.urandom(128) if rank == 0 else None]            # stands in for imdcu_comm_unique_id
    dist.broadcast_object_list(obj, src=0)
    pos, box = synth.film_lattice((20, 4, 4), synth.AL.a, 10.0, 12.0)
    rng = np.random.default_rng(1)
    pos = pos + rng.normal(0, 0.05, pos.shape)
    owner = host.slab_rank_of_atoms(pos, box, 0, 16, world)
    mine = np.nonzero(owner == rank)[0]
    counts = [None] * world
    dist.all_gather_object(counts, (len(mine), int(mine.sum()), obj[0]))
    dist.destroy_process_group()
    q.put((rank, counts, pos.shape[0], int(np.arange(pos.shape[0]).sum())))


def test_slab_partition_two_ranks_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, counts, n, idsum in res:
        assert sum(c[0] for c in counts) == n                 # complete
        assert sum(c[1] for c in counts) == idsum             # disjoint (index sums add up)
        assert counts[0][2] == counts[1][2] and len(counts[0][2]) == 128   # same id everywhere
        assert all(c[0] > 0 for c in counts)


def test_slab_partition_matches_layer_arithmetic():
    sys.path.insert(0, ROOT)
    from imd_b200 import host
    box = np.diag([40.0, 10.0, 10.0])
    x = np.array([0.0, 4.999, 5.0, 19.99, 20.0, 39.999, -0.001, 40.0])
    pos = np.stack([x, np.full_like(x, 1.0), np.full_like(x, 1.0)], axis=1)
    # non-periodic: clamp into the edge layers; 8 layers of 5 A, 4 ranks of 2 layers
    assert host.slab_rank_of_atoms(pos, box, 0, 8, 4).tolist() == [0, 0, 0, 1, 2, 3, 0, 3]
    # periodic: wrap
    assert host.slab_rank_of_atoms(pos, box, 1, 8, 4).tolist() == [0, 0, 0, 1, 2, 3, 3, 0]
