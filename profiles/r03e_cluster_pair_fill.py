#!/usr/bin/env python
"""How full would cluster-pair tiles be on the bench's system?  ("next" 4a: one evaluation per pair
with 8 x 4 cluster-pair tiles, one warp per tile.)  Pure geometry, runs on the CPU.

Al fcc (a = 4.05 A), periodic, 300 K displacements, list radius r_c + margin = 6.3 + 0.4 A -- the bench's film
interior.  For three atom orders, clusters of 8 consecutive atoms (i side) and of 4 consecutive atoms (j side) are
formed, and every (i-cluster, j-cluster) tile that holds at least one listed pair is counted.  A tile costs a warp one
iteration of 32 candidate pairs whatever its filling.

    python profiles/r03e_cluster_pair_fill.py > profiles/r03e_cluster_pair_fill.txt
"""
import numpy as np
from scipy.spatial import cKDTree

A, RC, MARGIN = 4.05, 6.3, 0.4
NC = 20                                   # 20^3 fcc cells = 32 000 atoms
RL = RC + MARGIN


def fcc(nc, a, rng):
    base = np.array([[0, 0, 0], [.5, .5, 0], [.5, 0, .5], [0, .5, .5]])
    g = np.stack(np.meshgrid(*[np.arange(nc)] * 3, indexing="ij"), -1).reshape(-1, 1, 3)
    pos = ((g + base).reshape(-1, 3) + 0.25) * a
    return np.mod(pos + rng.normal(0.0, 0.09, pos.shape), nc * a)     # ~300 K rms displacement per direction


def order_cells(pos, box):
    """the device's order: linked cells of side >= RL, x-major; inside a cell the input (lattice) order"""
    n = int(box / RL)
    c = np.minimum((pos / (box / n)).astype(int), n - 1)
    return np.argsort((c[:, 0] * n + c[:, 1]) * n + c[:, 2], kind="stable")


def morton3(c):
    out = np.zeros(c.shape[0], np.int64)
    for b in range(10):
        for d in range(3):
            out |= ((c[:, d] >> b) & 1) << (3 * b + d)
    return out


def order_cells_morton(pos, box):
    """the same cells, atoms inside a cell along a Morton curve of an 8^3 sub-grid"""
    n = int(box / RL)
    size = box / n
    c = np.minimum((pos / size).astype(int), n - 1)
    sub = np.minimum(((pos - c * size) / (size / 8)).astype(int), 7)
    key = ((c[:, 0] * n + c[:, 1]) * n + c[:, 2]) * 512 + morton3(sub)
    return np.argsort(key, kind="stable")


def order_columns(pos, box):
    """NBNXM-like: xy columns of side (8 / density)^(1/3), atoms of a column sorted by z, clusters = consecutive 8"""
    dens = pos.shape[0] / box ** 3
    n = max(1, int(round(box / (8.0 / dens) ** (1.0 / 3.0))))
    c = np.minimum((pos[:, :2] / (box / n)).astype(int), n - 1)
    col = c[:, 0] * n + c[:, 1]
    return np.lexsort((pos[:, 2], col))


def tiles(pos, box, order, radius):
    p = pos[order]
    pairs = cKDTree(p, boxsize=box).query_pairs(radius, output_type="ndarray")     # i < j, each pair once
    n = p.shape[0]
    per_atom_full = 2.0 * pairs.shape[0] / n
    i8, j4 = pairs[:, 0] // 8, pairs[:, 1] // 4
    i8b, j4b = pairs[:, 1] // 8, pairs[:, 0] // 4
    # half list: the pair (i < j) belongs to the tile (cluster8 of i, cluster4 of j)
    half = np.unique(i8.astype(np.int64) * (n // 4 + 1) + j4).shape[0]
    # full list over tiles: both directions
    full = np.unique(np.concatenate([i8.astype(np.int64) * (n // 4 + 1) + j4,
                                     i8b.astype(np.int64) * (n // 4 + 1) + j4b])).shape[0]
    return per_atom_full, half, full, n


def main():
    rng = np.random.default_rng(7)
    pos = fcc(NC, A, rng)
    box = NC * A
    print("Al fcc %d^3 cells, %d atoms, periodic, list radius %.1f A (r_c %.1f + margin %.1f)" % (NC, pos.shape[0], RL, RC, MARGIN))
    in_rc = 2.0 * cKDTree(pos, boxsize=box).query_pairs(RC, output_type="ndarray").shape[0] / pos.shape[0]
    for name, fn in (("cells of side >= r_list, lattice order inside (the device's order)", order_cells),
                     ("the same cells, Morton order of an 8^3 sub-grid inside", order_cells_morton),
                     ("xy columns of side (8/density)^(1/3), z-sorted (NBNXM-like)", order_columns)):
        full_pairs, half, full, n = tiles(pos, box, fn(pos, box), RL)
        print("\n%s" % name)
        print("  listed pairs per atom (full list)            %6.1f   of which inside r_c %5.1f" % (full_pairs, in_rc))
        print("  8x4 tiles per atom, half list                %6.3f   -> candidate pairs evaluated per atom %6.1f, filling %4.2f "
              "(listed) %4.2f (inside r_c)" % (half / n, 32.0 * half / n, (full_pairs / 2) / (32.0 * half / n),
                                             (in_rc / 2) / (32.0 * half / n)))
        print("  8x4 tiles per atom, full list over tiles     %6.3f   -> candidate pairs evaluated per atom %6.1f"
              % (full / n, 32.0 * full / n))
        print("  warp iterations per 32 atoms: now (full list in 32-atom slices) %5.1f, half-list tiles %5.1f"
              % (full_pairs, 32.0 * half / n))


if __name__ == "__main__":
    main()
