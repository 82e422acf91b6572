#!/usr/bin/env python
"""Multi-GPU parity check of the x-slab decomposition (run on a box with >= 2 GPUs):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/multi_gpu_check.py [--ttm]

Every rank uploads the atoms of its own slab, the run is advanced over several neighbour-list
rebuilds (atoms migrate), and per-atom results gathered by atom id must equal a 1-GPU run of
the same system (done on rank 0 with a second context): neighbour pair set bit-exact,
numneighs bit-exact, forces/energies/positions to 1e-10, and with --ttm the FD lattice to
the TTM tolerances.  tests/test_multi_gpu.py wraps this for pytest.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def build_system(args):
    from imd_b200 import synth
    spec = synth.AL if args.ttm else synth.CU
    if args.ttm:
        pos, box = synth.film_lattice((48, 6, 6), spec.a, 20.0, 24.0)
        pbc = (0, 1, 1)
    else:
        pos, box = synth.fcc_lattice((14, 6, 6), spec.a)
        pbc = (1, 1, 1)
    args.nrb_state = None
    if args.nrb:                                   # init_nrb on the perfect lattice, atom numbers from 1
        from imd_b200 import nrb
        args.nrb_state = nrb.init_nrb(pos, np.arange(1, pos.shape[0] + 1, dtype=np.int32), box, pbc, spec.a, 0.3)
    rng = np.random.default_rng(3)
    pos = pos + rng.normal(0, 0.04, pos.shape)
    n = pos.shape[0]
    mom = synth.maxwell_momenta(n, spec.mass, 0.12, seed=17)     # hot: rebuilds + migration within few steps
    return spec, pos, box, pbc, mom


def make_imd(args, device, comm, spec, pos, box, pbc, mom, eos, tp):
    from imd_b200 import host, synth
    imd = host.Imd(device, comm)
    imd.setup_potentials(synth.make_eam_tables(spec))
    imd.make_box(box, pbc, 0.4 if args.ttm else 0.5)
    imd.set_params(1, 0.1)
    if args.cuts and comm is not None:
        # slabs by atom count (imdcu_balance_cuts on the histogram of atoms per x cell layer) for the film, whose
        # vacuum makes equal-width slabs uneven; a deliberately lopsided cut for the uniform periodic bulk
        from imd_b200 import capi
        g = int(imd.global_cell_dim[0])
        hist = np.bincount(host.x_layer_of_atoms(pos, box, pbc[0], g), minlength=g)
        xcut = capi.balance_cuts(comm[2], hist, 1) if args.ttm else np.array([0, max(1, g // 3), g], np.int32)
        imd.set_cuts(xcut)
        args.cut_info = "cuts %s of %d layers" % (list(int(x) for x in xcut), g)
    mask = imd.owned_mask(pos)
    ids = np.arange(pos.shape[0], dtype=np.int32) + (1 if args.nrb else 0)
    n = int(mask.sum())
    imd.set_atoms(pos[mask], mom[mask], np.full(n, spec.mass), np.zeros(n, np.int32), ids[mask])
    if args.ttm:
        imd.init_ttm(tp, eos, 2)
    if args.nrb:
        imd.set_nrb(1.2, args.nrb_state, startup_images=True)
    return imd


def check_ttm3d(capi, dist, uid, rank, world, local):
    """x-slab sharded 3-D sweep == the numpy restatement of the whole lattice, bit for bit."""
    from oracle import ttm3d_np as t3
    ok = True
    ctx = capi.Context(local)
    ctx.comm_init(uid, rank, world)              # one communicator (a unique id is good for one init)
    for pbc in ((0, 1, 1), (1, 1, 1)):
        nxl, ny, nz = 9, 21, 40
        shape = (nxl * world, ny, nz)
        f = t3.make_case(shape, seed=31, hole=True, pbc=pbc)
        want = {k: v.copy() for k, v in f.items()}
        e_want = t3.run(want, (16.0, 15.0, 17.0), 0.1 / 250, 20, 250, 1, 5, pbc)
        ctx.ttm3d_init((nxl, ny, nz), (16.0, 15.0, 17.0), pbc, 20)
        sl = slice(rank * nxl, (rank + 1) * nxl + 2)
        ctx.ttm3d_upload(**{k: np.ascontiguousarray(v[sl]) for k, v in f.items()})
        e = ctx.ttm3d_diff(0.1 / 250, 250, 1, 5)
        out = ctx.ttm3d_download()
        parts = [None] * world if rank == 0 else None
        dist.gather_object((e, {k: v[1:-1] for k, v in out.items()}), parts, dst=0)
        if rank == 0:
            e_tot = sum(p[0] for p in parts)
            for k in ("natoms", "temp", "U", "xi"):
                got = np.concatenate([p[1][k] for p in parts])[:, 1:-1, 1:-1]
                assert np.array_equal(got, want[k][1:-1, 1:-1, 1:-1]), (pbc, k)
            assert abs(e_tot - e_want) <= 1e-12 * abs(e_want)
            print("ttm3d pbc %s ok: %d ranks x %d planes, Eabs %.12e" % (pbc, world, nxl, e_tot))
        dist.barrier()
    ctx.close()
    if rank == 0:
        print("MULTI_GPU_CHECK PASSED world=%d ttm3d=1" % world)
    dist.destroy_process_group()
    return 0 if ok else 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ttm", action="store_true")
    ap.add_argument("--nrb", action="store_true", help="--ttm plus the non-reflecting rear boundary")
    ap.add_argument("--ttm3d", action="store_true", help="3-D FD lattice sharded in x slabs")
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--cuts", action="store_true", help="x slabs of equal atom count / uneven width (imdcu_comm_set_cuts)")
    args = ap.parse_args()
    args.cut_info = ""
    args.ttm = args.ttm or args.nrb
    import torch
    import torch.distributed as dist
    from imd_b200 import capi, synth
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    obj = [capi.Context.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    if args.ttm3d:
        return check_ttm3d(capi, dist, obj[0], rank, world, local)
    spec, pos, box, pbc, mom = build_system(args)
    eos, tp = None, None
    if args.ttm:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from tests.ttm_util import TtmGolden, TTM_CASES, compare_cells
        eos = TtmGolden(TTM_CASES[0]).eos
        tp = synth.TtmParams(ttmdimx=12, fd_n_timesteps=20, fd_min_atoms=15, vlatdim=6, vlatbuffer=1, I0=3.0e16)
    imd = make_imd(args, local, (obj[0], rank, world), spec, pos, box, pbc, mom, eos, tp)
    ref = make_imd(args, local, None, spec, pos, box, pbc, mom, eos, tp) if rank == 0 else None

    def gather(fields):
        d = imd.atoms(fields=("id",) + fields, sort_by_id=False)
        out = [None] * world if rank == 0 else None
        dist.gather_object(d, out, dst=0)
        if rank != 0:
            return None
        cat = {k: np.concatenate([o[k] for o in out]) for k in d}
        order = np.argsort(cat["id"], kind="stable")
        return {k: v[order] for k, v in cat.items()}

    ok = True
    checks = [0, 1, args.steps // 2, args.steps]
    s = 0
    for target in checks:
        if target > s:
            (imd.main_loop_ttm if args.ttm else imd.main_loop)(target - s)
            if ref:
                (ref.main_loop_ttm if args.ttm else ref.main_loop)(target - s)
            s = target
        e = imd.calc_forces(s)
        pairs = imd.nblist_pairs()
        allp = [None] * world if rank == 0 else None
        dist.gather_object(pairs, allp, dst=0)
        if args.ttm:
            sc = imd.calc_ttm(s)
        got = gather(("pos", "mom", "force", "epot", "rho", "numneighs"))
        if rank == 0:
            er = ref.calc_forces(s)
            want = ref.atoms(fields=("id", "pos", "mom", "force", "epot", "rho", "numneighs"))
            assert np.array_equal(got["id"], want["id"]), "atom ids lost or duplicated"
            from tests.util import canon_pairs
            assert np.array_equal(canon_pairs(np.concatenate(allp)), canon_pairs(ref.nblist_pairs())), "pair sets differ"
            assert np.array_equal(got["numneighs"], want["numneighs"])
            bx = np.diag(box)
            d = got["pos"] - want["pos"]; d -= np.round(d / bx) * bx * np.array(pbc)
            assert np.abs(d).max() < 1e-10, np.abs(d).max()
            fs = np.abs(want["force"]).max()
            assert np.abs(got["force"] - want["force"]).max() < 1e-10 * fs
            assert np.abs(got["mom"] - want["mom"]).max() < 1e-10 * np.abs(want["mom"]).max()
            if args.nrb:                           # the boundary layer took the NRB momenta on both sides
                b = args.nrb_state.bnd > 0
                assert b.sum() > 0 and (s <= 1 or np.abs(got["mom"][b]).max() > 0)
            np.testing.assert_allclose(got["rho"], want["rho"], rtol=1e-11)
            np.testing.assert_allclose(e[0], er[0], rtol=1e-11)
            np.testing.assert_allclose(e[1], er[1], rtol=1e-9, atol=1e-9 * abs(er[0]))
            if args.ttm:
                scr = ref.calc_ttm(s)
                assert int(sc["diff_substeps"]) == int(scr["diff_substeps"])
                compare_cells(imd.ttm_lattice()["fields"], ref.ttm_lattice()["fields"], s,
                              int(scr["last_active_cell_global"]), "2gpu")
            print("step %3d ok: Epot %.12e  nbl rebuilds %d  atoms/rank %s" % (s, e[0], imd.nbl_count,
                                                                            [len(p) for p in allp]))
        if args.ttm:
            imd.move_atoms_ttm()
            if ref:
                ref.move_atoms_ttm()
        else:
            imd.move_atoms()
            if ref:
                ref.move_atoms()
        s += 1
        imd.steps = s
        if ref:
            ref.steps = s
    n_local = imd.natoms
    cnt = [None] * world if rank == 0 else None
    dist.gather_object(n_local, cnt, dst=0)
    if rank == 0:
        assert sum(cnt) == pos.shape[0]
        print("MULTI_GPU_CHECK PASSED world=%d ttm=%d atoms per rank %s rebuilds %d %s"
              % (world, args.ttm + args.nrb, cnt, imd.nbl_count, args.cut_info))
    imd.close()
    if ref:
        ref.close()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
