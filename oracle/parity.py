"""Parity tolerance shared by tests/ and __graft_entry__.smoke() (test infrastructure, like the rest of oracle/).

BASELINE.json asks for <= 1e-12 relative error per nonzero.  A purely relative bound is meaningless for entries that
vanish in exact arithmetic (orthogonal gradients on right-angled tets give |a_ij| ~ 1e-17 of pure rounding noise), so the
bound carries the natural absolute floor of a Gram-type matrix: the rounding error of a^e_ij scales with
sqrt(a^e_ii a^e_jj).  With S = sum_e |a^e| (oracle ABS_FLAG run):

    |a_ij - ref_ij| <= 1e-12 * |ref_ij| + 1e-14 * sqrt(S_ii * S_jj)
"""
import numpy as np

RTOL, FLOOR = 1e-12, 1e-14


def tolerance(ref, scale, ia, ja):
    n = len(ia) - 1
    rows = np.repeat(np.arange(n), np.diff(ia))
    diag = np.zeros(n)
    on = rows == ja
    diag[rows[on]] = scale[on]
    return RTOL * np.abs(ref) + FLOOR * np.sqrt(diag[rows] * diag[ja])


def check(a, ref, scale, ia, ja, what="values"):
    err = np.abs(a - ref)
    tol = tolerance(ref, scale, ia, ja)
    bad = err > tol
    assert not bad.any(), "%s: %d/%d nonzeros out of tolerance, worst err %.3e at |ref| %.3e (tol %.3e)" % (
        what, bad.sum(), len(ref), err[bad].max(), np.abs(ref[bad][err[bad].argmax()]), tol[bad][err[bad].argmax()])
    return float((err / np.maximum(tol, 1e-300)).max())
