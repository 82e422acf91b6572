"""world_size-2 gloo test (CPU) of the host-side part of the multi-GPU path: the slab
partition every rank applies to decide which atoms it uploads, and the id broadcast pattern
bench.py uses for the NCCL unique id."""
import os
import socket
import sys

import numpy as np
import pytest
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


def test_balance_cuts_film_with_vacuum():
    """imdcu_balance_cuts (host arithmetic of the library): a film with 30 % vacuum cut into slabs of equal
    atom count on cell-layer boundaries -- spread of atoms per rank <= 3 %, where equal-width slabs leave the end
    ranks short; degenerate inputs."""
    from imd_b200 import capi, host
    from imd_b200 import synth
    spec = synth.AL
    pos, box = synth.film_lattice((400, 2, 2), spec.a, 0.15 * 400 * spec.a / 0.7, 0.15 * 400 * spec.a / 0.7)   # 30 % vacuum
    g = 320                                                     # cell layers of ~7.2 A
    layer = host.x_layer_of_atoms(pos, box, 0, g)
    hist = np.bincount(layer, minlength=g)
    for nranks in (2, 4, 8):
        xcut = capi.balance_cuts(nranks, hist, 2)
        assert xcut[0] == 0 and xcut[-1] == g and np.all(np.diff(xcut) >= 2)
        per = np.array([hist[a:b].sum() for a, b in zip(xcut[:-1], xcut[1:])])
        assert per.sum() == pos.shape[0]
        spread = (per.max() - per.min()) / per.mean()
        equal = np.array([hist[r * g // nranks:(r + 1) * g // nranks].sum() for r in range(nranks)])
        print("ranks %d: cuts %s atoms %s spread %.3f (equal-width slabs: %.3f)"
              % (nranks, [int(x) for x in xcut], [int(x) for x in per], spread, (equal.max() - equal.min()) / equal.mean()))
        # cuts sit on layer boundaries: the spread cannot be asked to be finer than one layer of atoms
        assert spread <= 1.01 * hist.max() / per.mean()
        if nranks <= 4:
            assert spread <= 0.03                                # >= 56 material layers per rank here
        # the owner of every atom by these cuts is the slab its layer falls into
        owner = host.slab_rank_of_atoms(pos, box, 0, g, nranks, xcut)
        assert np.array_equal(np.bincount(owner, minlength=nranks), per)
    # a uniform system keeps the reference's equal slabs
    assert list(capi.balance_cuts(4, np.full(16, 100), 1)) == [0, 4, 8, 12, 16]
    # everything in one layer: the minimum slab width still holds
    one = np.zeros(12, np.int64); one[5] = 1000
    x = capi.balance_cuts(3, one, 2)
    assert x[0] == 0 and x[-1] == 12 and np.all(np.diff(x) >= 2)
    with pytest.raises(ValueError):
        capi.balance_cuts(8, np.ones(10, np.int64), 2)           # 8 slabs of 2 layers do not fit into 10
