"""Tie-aware top-k comparators shared by the parity tests.

The bar (BASELINE.json north_star): ids bit-exact "except for ties within fp32 tolerance", values within 1e-4
relative.  A mismatch is therefore only legal INSIDE a group of reference values that agree to fp32 rounding (or
at the k-th place, where the group may extend past the list); inside such a group the id SETS must agree."""
import numpy as np


def check_topk(V, I, Vref, Iref, rtol=1e-5, atol=1e-6, max_tied_rows=None):
    """V/I: values and ids under test [Q,k]; Vref/Iref: the reference's.  Works for ascending (distances) and
    descending (similarities) lists alike.  Returns the number of rows that differed inside a tie group."""
    V, I, Vref, Iref = (np.asarray(a) for a in (V, I, Vref, Iref))
    assert V.shape == Vref.shape and I.shape == Iref.shape
    fin = np.isfinite(Vref)
    assert np.array_equal(fin, np.isfinite(V)), "padding differs"
    np.testing.assert_allclose(V[fin], Vref[fin], rtol=rtol, atol=atol)
    tied_rows = 0
    k = I.shape[1]
    for r in np.nonzero((I != Iref).any(axis=1))[0]:
        tied_rows += 1
        for j in np.nonzero(I[r] != Iref[r])[0]:
            tol = rtol * abs(Vref[r, j]) + atol
            near = np.abs(Vref[r] - Vref[r, j]) <= 2 * tol
            last = near[k - 1]                      # the group reaches the end of the list: members may be cut off
            assert near.sum() > 1 or last, f"row {r} col {j}: id {I[r, j]} != {Iref[r, j]} without a value tie"
            if not last:
                assert set(I[r][near].tolist()) == set(Iref[r][near].tolist()), f"row {r}: tie group differs"
    if max_tied_rows is not None:
        assert tied_rows <= max_tied_rows, f"{tied_rows} rows differ by ties (allowed {max_tied_rows})"
    return tied_rows
