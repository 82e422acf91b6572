"""imdcu_eval_table_noalloc (include/imdcu.h): the table probe without cudaMalloc / cudaFree, device staging in a
file-scope static array of the library.  No counterpart in the reference.
Checked through the C ABI against the existing imdcu_eval_table (same kernel, same inputs: the results
must be bit-identical, which is inside the 1e-13 / 1e-12 the table test asks of the oracle) and against the oracle.

This file is named to be collected LAST: a failure here hides nothing.
It makes its own context, so it passes alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

from imd_b200 import capi, host, synth

pytestmark = pytest.mark.gpu


def make_imd(tables):
    imd = host.Imd(0)
    imd.setup_potentials(tables)
    imd.make_box(np.diag([30.0, 30, 30]), (1, 1, 1), 0.5)
    imd.set_params(1, 0.1)
    return imd


def arguments(pt, n, seed):
    """inside the table, below begin (is_short clamp), above end (clamped), exact knots"""
    rng = np.random.default_rng(seed)
    knots = pt.begin[0] + pt.step[0] * np.arange(0, pt.len[0], 37)
    arg = np.concatenate([rng.uniform(pt.begin[0] - 0.5, pt.end[0] + 1.0, n - knots.shape[0] - 2), knots,
                          [pt.begin[0], pt.end[0]]])
    assert arg.shape[0] == n
    return arg


def test_noalloc_equals_eval_table_on_two_calls_in_a_row():
    from oracle.binding import Table
    tables = synth.make_eam_tables(synth.CU)
    imd = make_imd(tables)
    c = imd.ctx
    # call 1: 30 000 arguments on the pair table; call 2, right behind it: 1 000 other arguments on the rho table
    # (the array still holds call 1's values beyond the first 1 000 elements of each third)
    for which, key, n, seed in ((0, "pair", 30000, 1), (1, "rho", 1000, 2)):
        pt = tables[key]
        arg = arguments(pt, n, seed)
        val, grad = c.eval_table_noalloc(which, arg)
        rval, rgrad = c.eval_table(which, arg)
        assert np.array_equal(val, rval) and np.array_equal(grad, rgrad), (key, n)
        oval, ograd = Table(pt).pair_int(arg)
        assert np.max(np.abs(val - oval)) <= 1e-13 * np.max(np.abs(oval))
        assert np.max(np.abs(grad - ograd)) <= 1e-12 * np.max(np.abs(ograd))
    # the same two calls again, the other way round, interleaved with the allocating entry point
    for which, key, n, seed in ((1, "rho", 1000, 3), (0, "pair", 30000, 4)):
        arg = arguments(tables[key], n, seed)
        rval, rgrad = c.eval_table(which, arg)
        val, grad = c.eval_table_noalloc(which, arg)
        assert np.array_equal(val, rval) and np.array_equal(grad, rgrad), (key, n)
    imd.close()


def test_noalloc_limit_and_empty_call():
    tables = synth.make_eam_tables(synth.CU)
    imd = make_imd(tables)
    c = imd.ctx
    nmax = capi.EVAL_NOALLOC_MAX
    pt = tables["emb"]
    arg = arguments(pt, nmax, 5)                               # exactly the maximum: accepted
    val, grad = c.eval_table_noalloc(2, arg)
    rval, rgrad = c.eval_table(2, arg)
    assert np.array_equal(val, rval) and np.array_equal(grad, rgrad)
    with pytest.raises(Exception) as e:                        # one more: rejected, the message names the limit
        c.eval_table_noalloc(2, np.concatenate([arg, arg[:1]]))
    assert "IMDCU_EVAL_NOALLOC_MAX" in str(e.value) and str(nmax) in str(e.value), str(e.value)
    val, grad = c.eval_table_noalloc(2, np.zeros(0))           # nothing to do
    assert val.shape == (0,) and grad.shape == (0,)
    val, grad = c.eval_table_noalloc(2, arg[:777])             # still usable after the rejected call
    assert np.array_equal(val, rval[:777]) and np.array_equal(grad, rgrad[:777])
    imd.close()
