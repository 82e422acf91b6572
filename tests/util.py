"""Shared helpers for the tests: golden fixtures (dumps of the unmodified reference,
tests/golden/make_golden.py) and the synthetic tables they were produced with."""
import os

import numpy as np

from imd_b200 import synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = ["cu_6x6x6_pbc", "cu_8x6x7_pbc", "cu_7x6x6_free_x", "al_5x5x5_pbc", "ni3al_6x6x6_pbc",
                "ni3al_7x5x6_free_x"]
INTERP_CASES = ["cu_6x6x6_pbc_4point", "cu_6x6x6_pbc_spline", "ni3al_6x5x5_free_x_4point",
                "ni3al_6x5x5_free_x_spline"]     # potaccess.h FOURPOINT / SPLINE modes
SPECS = {"cu": synth.CU, "al": synth.AL, "ni3al": synth.NI3AL}


class Golden:
    def __init__(self, name):
        self.name = name
        self.z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        self.spec = SPECS[str(self.z["spec_name"])]
        self.binary = isinstance(self.spec, synth.BinarySpec)
        self.ntypes = 2 if self.binary else 1
        self.interp = str(self.z["interp"]) if "interp" in self.z.files else "3point"
        self.tables = (synth.make_eam_tables_binary(self.spec, interp=self.interp) if self.binary
                       else synth.make_eam_tables(self.spec, interp=self.interp))
        self.pbc = self.z["pbc"]
        self.box = np.diag(self.z["box_param"].astype(np.float64) * self.spec.a)
        self.margin = float(self.z["nbl_margin"])
        self.timestep = float(self.z["timestep"])
        self.steps = [int(s) for s in self.z["dump_steps"]]

    def step(self, s, key):
        return self.z["s%d_%s" % (s, key)]

    def by_id(self, s, key, tag="post"):
        """Array ``key`` of step s reordered by atom id (ids are 0..n-1 or 1..n)."""
        ids = self.step(s, tag + "_id")
        order = np.argsort(ids, kind="stable")
        return self.step(s, key)[order]

    def ids_sorted(self, s, tag="post"):
        return np.sort(self.step(s, tag + "_id"))


def canon_pairs(pairs):
    """Unordered id pairs -> sorted unique rows (multiplicity kept)."""
    p = np.sort(np.asarray(pairs, dtype=np.int64), axis=1)
    order = np.lexsort((p[:, 1], p[:, 0]))
    return p[order]
