"""Helpers shared by the TTM tests: fixtures dumped from the UNMODIFIED reference's
-DMPI -DTTM build (tests/golden/make_golden_ttm.py) and the tolerances they are compared with.

Tolerances (DESIGN.md "TTM parity"):
* step 0 has no root-finding on the path: every FD-cell field must agree to 1e-10 relative.
* from step 1 on, do_ADV sets Te = EOS_te_from_r_ee(rho, U) with the reference's Brent
  minimiser (fminbnd3.c) whose hard-coded tolerance is 1e-5 K; its iteration path is sensitive
  to last-bit changes of its inputs (|ee(T)-U| is V-shaped on a piecewise-linear table, so the
  parabolic step divides by a difference of nearly equal numbers).  Te is therefore only
  defined to ~1e-5 K per MD step by the reference itself; fields that depend on Te carry the
  same uncertainty.  TE_ABS_K below is that bound; REL_TE is used for the derived fields.
"""
import os

import numpy as np

from imd_b200 import synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TTM_CASES = ["ttm_al_56x6x6_p2", "ttm_al_56x6x6_p4", "ttm_al_48x5x5_p2_hot"]
TMM_CASES = ["ttm_tmm_al_96x5x5_p2", "ttm_tmm_al_96x5x5_p4"]     # do_tmm of the reference itself (imd_tmm.c)
NRB_CASES = ["ttm_nrb_al_40x5x5_p2", "ttm_nrb_al_56x5x6_p4"]     # reference built with -DNRB -DREFPOS
TTM_FIELDS = ["natoms", "natoms_old", "temp", "md_temp", "xi", "dens", "fd_k", "fd_g", "Z", "ne", "Ce", "U",
              "source"]
TE_ABS_K = 5e-5             # K: a few times fminbnd's tolx
REL_STEP0 = 1e-10
REL_TE = 2e-7               # fields downstream of the Brent inverse (xi amplifies by Te/(Te-Ti))
REL_LATTICE = 1e-9          # md_temp, dens, vcom: pure atom sums


class TtmGolden:
    def __init__(self, name):
        self.name = name
        self.z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        z = self.z
        self.tp = synth.TtmParams(**{k[3:]: z[k].item() for k in z.files if k.startswith("tp_")})
        self.eos = synth.make_eos_tables()
        self.eos["cve"].diag = z["eos_cve_diag"]
        self.eos["ee"].diag = z["eos_ee_diag"]
        self.spec = synth.AL
        self.tables = synth.make_eam_tables(self.spec)
        self.box = np.diag(z["box"])
        self.nranks = int(z["nranks"])
        self.steps = [int(s) for s in z["dump_steps"]]
        self.nsteps = int(z["nsteps"])
        self.timestep = float(z["timestep"])
        self.margin = float(z["nbl_margin"])
        self.pdecay = None
        extra = str(z["extra"])
        kv = dict(ln.split() for ln in extra.strip().splitlines())
        if "xipdecay" in kv:
            self.pdecay = (1, float(kv["xipdecay"]), float(kv["ramp_start"]), float(kv["ramp_end"]))
        self.nrb = None
        if "nrb_k" in kv:
            self.nrb = dict(nrb_k=float(kv["nrb_k"]), nrb_alat=float(kv["nrb_alat"]), nrb_eps=float(kv["nrb_eps"]))

    def nrb_state(self):
        """NRB state of init_nrb (static), rows in atom-number order like atoms(s)."""
        from imd_b200.nrb import NrbState
        z = self.z
        return NrbState(z["s0_post_id"], z["nrb_nrbbnd"], z["nrb_nrbneigh"], z["nrb_nrbi"], z["nrb_refpos"])

    def pre_mom(self, s):
        """momenta BEFORE calc_forces(s) (NRB cases: nrb_forces changes them inside calc_forces)."""
        return self.z["s%d_pre_mom" % s]

    def atoms(self, s):
        z = self.z
        return dict(id=z["s%d_post_id" % s], pos=z["s%d_post_pos" % s], mom=z["s%d_post_mom" % s],
                    type=z["s%d_type" % s], mass=z["s%d_mass" % s], force=z["s%d_force" % s],
                    numneighs=z["s%d_numneighs" % s])

    def ttm(self, s):
        z = self.z
        pre = "s%d_ttm_" % s
        d = {k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}
        d["fields"] = {k: d["cells"][:, i] for i, k in enumerate(TTM_FIELDS)}
        return d


# observed errors of every compare_cells call of the session, printed by tests/conftest.py at the end of the run
# ("TTM parity, observed"): {(label, kind): max error}; kinds: Te_K (|dTe| in K, steps > 0), step0_rel, te_rel
# (fields downstream of the Brent inverse, steps > 0), xi_rel (relative to max |xi|), lattice_rel (dens, md_temp)
OBSERVED = {}


def _observe(label, kind, value):
    key = (label, kind)
    OBSERVED[key] = max(OBSERVED.get(key, 0.0), float(value))


def _relmax(a, b, scale=None):
    m = b != 0
    if scale is not None:
        return float(np.abs(a - b).max() / scale) if a.size else 0.0
    return float(np.abs(a[m] / b[m] - 1.0).max()) if m.any() else 0.0


def compare_cells(got, ref, step, last_active, label=""):
    """got/ref: dict field -> [G].  Cells beyond last_active hold uninitialised memory in the
    reference for some fields (FILLMESH skips them, imd_ttm.c:649-650): compare natoms/dens/md_temp
    there and nothing else."""
    G = ref["natoms"].shape[0]
    act = np.arange(G) <= last_active
    for k in TTM_FIELDS:
        a, b = np.asarray(got[k], dtype=np.float64), np.asarray(ref[k], dtype=np.float64)
        if k in ("natoms", "natoms_old"):
            assert np.array_equal(a, b), "%s step %d %s differs: %s vs %s" % (label, step, k, a, b)
            continue
        if k in ("dens", "md_temp"):
            _observe(label, "lattice_rel", _relmax(a, b))
            np.testing.assert_allclose(a, b, rtol=REL_LATTICE, atol=0, err_msg="%s step %d %s" % (label, step, k))
            continue
        a, b = a[act], b[act]
        if step == 0:
            _observe(label, "step0_rel", _relmax(a, b))
        elif k == "temp":
            _observe(label, "Te_K", np.abs(a - b).max() * 11604.5)
        elif k == "xi":
            _observe(label, "xi_rel", _relmax(a, b, np.abs(b).max()))
        else:
            _observe(label, "te_rel", _relmax(a, b))
        if step == 0:
            np.testing.assert_allclose(a, b, rtol=REL_STEP0, atol=0, err_msg="%s step %d %s" % (label, step, k))
        elif k == "temp":
            np.testing.assert_allclose(a * 11604.5, b * 11604.5, rtol=0, atol=TE_ABS_K,
                                       err_msg="%s step %d Te [K]" % (label, step))
        elif k == "xi":
            # xi ~ (Te' - Ti): where the two nearly cancel, the 1e-5 K of Te are no longer small against the
            # difference itself -> the bound is absolute, on the scale of the cells that do not cancel
            np.testing.assert_allclose(a, b, rtol=REL_TE, atol=REL_TE * np.abs(b).max(),
                                       err_msg="%s step %d %s" % (label, step, k))
        else:
            np.testing.assert_allclose(a, b, rtol=REL_TE, atol=0, err_msg="%s step %d %s" % (label, step, k))
