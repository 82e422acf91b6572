import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_terminal_summary(terminalreporter):
    """Observed TTM parity errors of this session (tests/ttm_util.py: OBSERVED), so that the tolerances in
    that file can be checked against what the device actually does."""
    try:
        from tests import ttm_util
    except Exception:
        return
    if not ttm_util.OBSERVED:
        return
    tr = terminalreporter
    tr.write_line("TTM parity, observed (tolerances: Te %.1e K, downstream %.1e, step 0 %.1e, lattice sums %.1e)"
                  % (ttm_util.TE_ABS_K, ttm_util.REL_TE, ttm_util.REL_STEP0, ttm_util.REL_LATTICE))
    for (label, kind), v in sorted(ttm_util.OBSERVED.items()):
        tr.write_line("  %-28s %-12s %.3e" % (label, kind, v))
