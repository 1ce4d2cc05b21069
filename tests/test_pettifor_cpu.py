"""CPU tests of the Pettifor (orthogonal sp3) oracle, oracle/pettifor_oracle.c.

The reference pins nothing for this path, so the oracle is pinned by
  * agreement of its verbatim form (dense per-atom derivative matrices, TB_Pettifor.f90:141-223,463-521) with the
    sparse-equivalent form;
  * invariants (H symmetric, N_e, sum of forces = 0);
  * finite differences of the band energy: dHopping_s is the exact derivative of Hopping, so
    F_k = sum_n f_n dE_n/ds_k  (Hellmann-Feynman at fixed occupations) must hold to the FD truncation error;
  * the virial-type identity between the supercell force and the strain derivative of the band energy.
"""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def d64(wk, orc):
    wl = wk.diamond64()
    return wl, orc.pettifor_step(wl, verbatim=True), orc.pettifor_step(wl)


def test_verbatim_equals_sparse(d64):
    wl, ov, os_ = d64
    assert np.abs(ov["F"] - os_["F"]).max() <= 1e-12 * max(1.0, np.abs(ov["F"]).max())
    assert np.abs(ov["rho"] - os_["rho"]).max() <= 1e-13


def test_invariants(d64, orc):
    wl, ov, os_ = d64
    H = ov["H"]
    assert np.array_equal(H, H.T)                                   # mirrored minimum images are exact negatives
    assert abs(ov["fe"].sum() - wl["n_electrons"]) < 1e-9
    A = ov["Aij"]
    assert abs(np.trace(A) - wl["n_electrons"]) < 1e-9              # tr(rho) = N_e in an orthogonal basis
    assert np.abs(ov["F"].sum(axis=0)).max() < 1e-10                # translation invariance
    # on-site blocks: diag(Es, Ep, Ep, Ep)
    p = wl["params"][0]
    assert np.allclose(np.diag(H)[:4], [p["Es"], p["Ep"], p["Ep"], p["Ep"]], rtol=0, atol=0)


def _band_energy(orc, wl, fe):
    nb = orc.neighbours(wl)
    H = orc.pet_assemble_H(wl, nb)
    w = np.linalg.eigvalsh(H)
    return float((fe * w).sum())


def test_forces_are_band_energy_derivatives(d64, orc):
    wl, ov, os_ = d64
    fe = ov["fe"]
    # degenerate levels move between each other under a perturbation; sum_n f_n E_n is still differentiable when
    # f is constant across a degenerate multiplet (Fermi function of the unperturbed level) -- true here
    eps = 1e-5
    for (k, al) in ((0, 0), (17, 1), (40, 2)):
        e = []
        for sgn in (+1, -1):
            w2 = dict(wl); S = wl["S"].copy(); S[k, al] += sgn * eps; w2["S"] = S
            e.append(_band_energy(orc, w2, fe))
        fd = (e[0] - e[1]) / (2 * eps)
        assert abs(fd - ov["F"][k, al]) <= 2e-6 * max(1.0, abs(fd)), (k, al, fd, ov["F"][k, al])


def test_supercell_force_is_strain_derivative(d64, orc):
    wl, ov, os_ = d64
    fe = ov["fe"]
    P = os_["P"]                                                    # att(k,i) = dE/dh with h(k? ,i?) -- check both index orders
    eps = 1e-6
    h0 = wl["supce"]
    got = np.zeros((3, 3))
    for a in range(3):
        for b in range(3):
            e = []
            for sgn in (+1, -1):
                w2 = dict(wl); h = h0.copy(); h[a, b] += sgn * eps; w2["supce"] = h
                e.append(_band_energy(orc, w2, fe))
            got[a, b] = (e[0] - e[1]) / (2 * eps)
    # SCforce%att(delta, gamma) = sum dE/dr_delta * s_gamma  = dE/dh(gamma, delta)   (x_delta = sum_gamma s_gamma h(gamma,delta))
    assert np.abs(got.T - P).max() <= 1e-5 * np.abs(P).max(), (got, P)


def test_vs_dvs_branch_quirk(orc, wk):
    """Vs switches to the cubic tail at r >= r1, dVs at r > r1 (TB_Pettifor.f90:1967 vs :2010): both branches reachable."""
    wl = wk.diamond64()
    p = wl["params"][0]
    assert p["r1"] < p["rm"]
    nb = orc.neighbours(wl)
    r = nb.dist[:, :, 3][np.arange(nb.cap)[None, :] < nb.size[:, None]]
    assert (r < p["r1"]).any() and (r >= p["r1"]).any()             # the fixture exercises both radial branches
