"""CPU tests of the nonadiabatic-coupling oracle (oracle.coupling_Mij*; Electron_ion_coupling_Mij, Nonadiabatic.f90:782-945):
the statement-by-statement loop restatement against the matrix-product form, and the properties the reference's
expression has by construction (zero diagonal, symmetric copy of an antisymmetric quantity, vanishing for identical
eigenvector sets in Tully's form)."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def two_steps(wk, orc):
    wl = wk.au_fcc((1, 1, 1), sigma=0.1, seed=3)            # 4 atoms, N = 36: the loop form is O(N^3) in Python
    o0 = orc.nrl_step(wl, with_pressure=False)
    w1 = dict(wl); w1["S"] = wl["S"] + np.random.default_rng(0).normal(0, 2e-3, wl["S"].shape)
    o1 = orc.nrl_step(w1, with_pressure=False)
    return o0, o1


@pytest.mark.parametrize("kind_M", [0, -1])
@pytest.mark.parametrize("with_S", [True, False])
def test_loops_equal_matrix_form(two_steps, orc, kind_M, with_S):
    o0, o1 = two_steps
    S = o1["S"] if with_S else None
    A = orc.coupling_Mij_loops(o1["Ha"], o0["Ha"], S, kind_M)
    B = orc.coupling_Mij(o1["Ha"], o0["Ha"], S, kind_M)
    assert np.abs(A - B).max() <= 1e-14
    assert np.array_equal(B, B.T) and not np.diag(B).any()


def test_tully_vanishes_for_identical_sets(two_steps, orc):
    o0, _ = two_steps
    M = orc.coupling_Mij(o0["Ha"], o0["Ha"], o0["S"], 0)    # psi^T S psi is symmetric: the antisymmetric part is zero
    assert np.abs(M).max() < 1e-13


def test_zero_norm_columns_are_dropped(two_steps, orc):
    o0, o1 = two_steps
    Ha = o1["Ha"].copy(); Ha[:, 5] = 0.0
    A = orc.coupling_Mij_loops(Ha, o0["Ha"], o1["S"], 0)
    B = orc.coupling_Mij(Ha, o0["Ha"], o1["S"], 0)
    assert np.abs(A - B).max() <= 1e-14 and np.isfinite(B).all()
