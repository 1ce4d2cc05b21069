"""Edge cases of the hot path on the GPU (through the C ABI), each against the oracle:
isolated atoms (empty neighbour lists), a single atom, a ragged cell where the neighbour capacity has to grow between
two calls on the same context, a very small cell where an atom meets several periodic images (the reference keeps the
single minimum image, overwrite semantics), and an empty owned range in a partitioned run."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _step(c, o, wl, orc, tolF=1e-9):
    g = c.get_near_neighbours()
    np.testing.assert_array_equal(g["size"], o["nb"].size)
    H, S = c.assemble_HS()
    assert np.abs(H - o["H"]).max() <= 1e-12 * max(1.0, np.abs(o["H"]).max())
    if S is not None:
        assert np.abs(S - o["S"]).max() <= 1e-12
    c.set_eigen(o["Ha"], o["Ei"], o["fe"])
    c.density()
    F = c.forces()
    dF = np.abs(orc.forces_cartesian(F, wl["supce"]) - orc.forces_cartesian(o["F"], wl["supce"])).max()
    assert dF <= tolF, dF
    return F


def test_isolated_atoms_have_empty_lists(wk, orc, api):
    S = np.array([[0.1, 0.1, 0.1], [0.6, 0.6, 0.6], [0.1, 0.6, 0.35]])
    wl = wk._wl("iso3", "nrl", "Au", S, np.diag([40.0, 40.0, 40.0]), 11, 0.1)
    o = orc.nrl_step(wl, with_pressure=True)
    assert o["nb"].size.sum() == 0
    c = api.TBContext(0); c.set_workload(wl)
    F = _step(c, o, wl, orc)
    assert not F.any()
    assert np.abs(c.pressure() - o["P"]).max() <= 1e-12
    c.close()


def test_single_atom(wk, orc, api):
    wl = wk._wl("one", "pettifor", "C", np.array([[0.3, 0.4, 0.5]]), np.diag([30.0, 30.0, 30.0]), 4, 0.1)
    o = orc.pettifor_step(wl)
    c = api.TBContext(0); c.set_workload(wl)
    F = _step(c, o, wl, orc)
    assert not F.any()
    E, Epot = c.get_Erep_s()
    Eo, _ = orc.pet_erep(wl, o["nb"])
    assert E == Eo
    c.close()


def test_capacity_grows_on_the_same_context(wk, orc, api):
    """dilute cell first (small capacity guess), then a dense one with the same atom count"""
    c = api.TBContext(0)
    S, h = wk.lattice(wk._FCC, 9.0, (2, 2, 2), 0.0, 0)
    wl0 = wk._wl("dilute", "nrl", "Au", S, h, 11, 0.1)
    c.set_workload(wl0)
    g = c.get_near_neighbours()
    assert g["size"].max() == 0 or g["size"].max() < 20
    wl1 = wk.au_fcc((2, 2, 2), sigma=0.1, seed=4)
    assert wl1["na"] == wl0["na"]
    o = orc.nrl_step(wl1, with_pressure=False)
    c.set_system(wl1["koa"], wl1["S"], wl1["supce"])
    _step(c, o, wl1, orc)
    c.close()


def test_tiny_cell_many_images(wk, orc, api):
    """1x1x1 fcc Au (4 atoms, a = 4.08 A < cut-off 6.6 A): every pair has several images inside the cut-off; the reference
    keeps one (Atomic_tools.f90:4080-4131) and so must the library"""
    wl = wk.au_fcc((1, 1, 1), sigma=0.1, seed=3)
    o = orc.nrl_step(wl, with_pressure=False)
    assert o["nb"].size.max() == wl["na"] - 1
    c = api.TBContext(0); c.set_workload(wl)
    _step(c, o, wl, orc)
    c.close()


def test_empty_owned_range(wk, api):
    wl = wk.au_fcc((2, 2, 2), sigma=0.1, seed=4)
    c = api.TBContext(0); c.set_workload(wl)
    c.get_near_neighbours_device()
    c.set_partition(5, 4)                       # first = last + 1: this rank owns nothing
    c.assemble_blocks()
    C_, E, fe = wk.synthetic_eigen(wl["na"] * 9, seed=2)
    c.set_eigen(C_, E, fe)
    c.density()
    F = c.forces()
    assert not F.any()
    assert not c.pressure().any()
    c.close()


def test_split_work_units_give_the_same_blocks(wk, api, monkeypatch):
    """density kernel with the owned neighbours of an atom split over two CTAs (XTB_DENSITY_SPLIT, used for atoms that need
    two sub-tasks when the persistent grid has few rounds): same rho / rho_E blocks"""
    import importlib
    abi = importlib.import_module("xtant-3_b200._abi"); interop = importlib.import_module("xtant-3_b200.interop")
    wl = wk.au108()
    C_, E, fe = wk.synthetic_eigen(wl["na"] * 9, seed=5)
    res = []
    for sp in ("1", "2", "3"):
        monkeypatch.setenv("XTB_DENSITY_SPLIT", sp)
        c = api.TBContext(0); c.set_workload(wl); c.get_near_neighbours_device(); c.set_eigen(C_, E, fe)
        c.density()
        res.append((interop.buffer(c, abi.BUF_RHO).cpu().numpy().copy(), interop.buffer(c, abi.BUF_RHOE).cpu().numpy().copy()))
        c.close()
    for r, e in res[1:]:
        assert np.abs(r - res[0][0]).max() <= 1e-14 * np.abs(res[0][0]).max()
        assert np.abs(e - res[0][1]).max() <= 1e-14 * np.abs(res[0][1]).max()


def test_split_tail_round_gives_the_same_blocks(wk, api, monkeypatch):
    """last, partly filled round of the persistent density grid: its atoms are given to two CTAs each (automatic when
    they fill at most half the grid, e.g. 512 owned atoms on 148 SMs; forced here from owned atom 50 / 0 / 107 on):
    bit-identical rho / rho_E blocks (the order of the state sum of a block does not depend on the unit that owns it)"""
    import importlib
    abi = importlib.import_module("xtant-3_b200._abi"); interop = importlib.import_module("xtant-3_b200.interop")
    wl = wk.au108()
    C_, E, fe = wk.synthetic_eigen(wl["na"] * 9, seed=6)
    res = []
    for tail in ("-1", "50", "0", "107"):
        monkeypatch.setenv("XTB_DENSITY_TAIL", tail)
        c = api.TBContext(0); c.set_workload(wl); c.get_near_neighbours_device(); c.set_eigen(C_, E, fe)
        c.density()
        res.append((interop.buffer(c, abi.BUF_RHO).cpu().numpy().copy(), interop.buffer(c, abi.BUF_RHOE).cpu().numpy().copy()))
        c.close()
    assert np.abs(res[0][0]).max() > 0
    for r, e in res[1:]:
        assert np.array_equal(r, res[0][0]) and np.array_equal(e, res[0][1])


def test_pair_once_forces_are_bit_identical_to_the_every_entry_form(wk, api, monkeypatch):
    """nrl_force_pairs_kernel differentiates every unordered pair once and hands the partner the negated value;
    XTB_FORCES_MODE=0 evaluates every list entry from its own atom (the earlier single-kernel form).  Every operation on the
    path is odd in (delta_ik - delta_jk), so the two give the same bits -- on the whole cell and on an owned range with
    partners on both sides outside it (those pairs are evaluated by the owned atom; the context of the range is handed the
    complete rho / rho_E stores of the whole-cell context, as the halo exchange of a partitioned run would)."""
    import importlib
    abi = importlib.import_module("xtant-3_b200._abi"); interop = importlib.import_module("xtant-3_b200.interop")
    wl = wk.au108()
    C_, E, fe = wk.synthetic_eigen(wl["na"] * 9, seed=11)
    full = api.TBContext(0); full.set_workload(wl); full.get_near_neighbours_device(); full.assemble_blocks()
    full.set_eigen(C_, E, fe); full.density()
    res, prs = [], []
    for mode in ("0", "1"):
        monkeypatch.setenv("XTB_FORCES_MODE", mode)
        res.append(full.forces().copy())
        prs.append(full.pressure().copy())
    assert np.abs(res[0]).max() > 1e-3 and np.abs(prs[0]).max() > 1e-3
    assert np.array_equal(res[0], res[1]), np.abs(res[0] - res[1]).max()
    assert np.array_equal(prs[0], prs[1]), np.abs(prs[0] - prs[1]).max()      # supercell force: the same pair-once rule
    first, last = 31, 77
    part = api.TBContext(0); part.set_workload(wl); part.get_near_neighbours_device(); part.set_partition(first, last)
    part.assemble_blocks(); part.set_eigen(C_, E, fe); part.density()
    for which in (abi.BUF_RHO, abi.BUF_RHOE):
        src = interop.buffer(full, which); dst = interop.buffer(part, which)
        n = min(src.numel(), dst.numel())
        dst[:n].copy_(src[:n])
    import torch
    torch.cuda.synchronize()
    for mode in ("0", "1"):
        monkeypatch.setenv("XTB_FORCES_MODE", mode)
        F = part.forces()
        assert np.array_equal(F[first - 1:last], res[0][first - 1:last]), (mode, np.abs(F[first - 1:last] - res[0][first - 1:last]).max())
        prs.append(part.pressure().copy())
    assert np.array_equal(prs[2], prs[3]) and np.abs(prs[2]).max() > 1e-3       # owned range: partial sums of the two forms
    part.close(); full.close()


def test_pair_once_blocks_of_an_owned_range_equal_the_full_cell(wk, api):
    """nrl_blocks_kernel evaluates an unordered pair once and writes both transposed blocks when both atoms are owned; with an
    owned range the pairs that reach outside it are evaluated by the owned atom: same H / S block rows either way."""
    import importlib
    abi = importlib.import_module("xtant-3_b200._abi"); interop = importlib.import_module("xtant-3_b200.interop")
    import torch
    wl = wk.au108()
    c = api.TBContext(0); c.set_workload(wl); c.get_near_neighbours_device(); c.assemble_blocks()
    rp = interop.buffer(c, abi.BUF_ROWPTR, torch.int32).cpu().numpy().copy()
    Hb = interop.buffer(c, abi.BUF_HBLK).cpu().numpy().copy(); Sb = interop.buffer(c, abi.BUF_SBLK).cpu().numpy().copy()
    c.close()
    assert np.abs(Hb).max() > 0
    first, last = 31, 77
    c = api.TBContext(0); c.set_workload(wl); c.get_near_neighbours_device(); c.set_partition(first, last); c.assemble_blocks()
    rp2 = interop.buffer(c, abi.BUF_ROWPTR, torch.int32).cpu().numpy().copy()
    Hp = interop.buffer(c, abi.BUF_HBLK).cpu().numpy().copy(); Sp = interop.buffer(c, abi.BUF_SBLK).cpu().numpy().copy()
    c.close()
    assert np.array_equal(rp, rp2)
    lo, hi = rp[first - 1] * 81, rp[last] * 81
    assert np.array_equal(Hb.ravel()[lo:hi], Hp.ravel()[lo:hi]) and np.array_equal(Sb.ravel()[lo:hi], Sp.ravel()[lo:hi])


@pytest.mark.parametrize("name", ["c_3x3x3", "au_5x5x6", "si216"])
def test_single_image_neighbour_path_is_bit_identical_to_the_27_image_scan(wk, api, monkeypatch, name):
    """orthorhombic cells with every edge > 2 rm take the one-image path (neighbours.cu, min_image_ortho); XTB_NN_SCAN27=1
    forces the reference's 27-image scan: same sizes, lists, distances and fractional shifts, bit for bit."""
    wl = {"c_3x3x3": lambda: wk.c_dia((3, 3, 3), sigma=0.1, seed=5), "si216": wk.si216,
          "au_5x5x6": lambda: wk.au_fcc((5, 5, 6), sigma=0.08, seed=11)}[name]()
    got = []
    for scan in ("1", "0"):
        monkeypatch.setenv("XTB_NN_SCAN27", scan)
        c = api.TBContext(0); c.set_workload(wl)
        g = c.get_near_neighbours()
        got.append(g)
        c.close()
    a, b = got
    assert a["size"].sum() > 0
    np.testing.assert_array_equal(a["size"], b["size"])
    m = int(a["size"].max())
    mask = np.arange(m)[None, :] < a["size"][:, None]
    np.testing.assert_array_equal(a["list"][:, :m][mask], b["list"][:, :m][mask])
    assert np.array_equal(a["dist"][:, :m][mask], b["dist"][:, :m][mask])
    assert np.array_equal(a["dist_s"][:, :m][mask], b["dist_s"][:, :m][mask])


def test_segmented_block_layout_gives_identical_results(wk, api):
    """xtb_set_block_layout(parts): block stores in equal-capacity segments per atom range (one in-place all-gather in a
    multi-GPU run).  Dense H/S, dense rho / rho_E, forces and pressure are bit-identical to the compact layout; the pair
    count is unchanged; the segments start at multiples of the largest range's block count."""
    import importlib
    abi = importlib.import_module("xtant-3_b200._abi"); interop = importlib.import_module("xtant-3_b200.interop")
    import torch
    wl = wk.au108()
    C_, E, fe = wk.synthetic_eigen(wl["na"] * 9, seed=9)
    res = []
    for parts in (1, 3, 8):
        c = api.TBContext(0); c.set_workload(wl); c.set_block_layout(parts)
        c.get_near_neighbours_device()
        H, S = c.assemble_HS()
        c.set_eigen(C_, E, fe)
        A, AE = c.density(dense=True)
        F = c.forces(); P = c.pressure()
        rp = interop.buffer(c, abi.BUF_ROWPTR, torch.int32).cpu().numpy()
        res.append((H, S, A, AE, F, P, c.counter(5), c.counter(4), rp))
        c.close()
    na = wl["na"]
    for parts, r in zip((3, 8), res[1:]):
        for a, b in zip(r[:6], res[0][:6]):
            assert np.array_equal(a, b)
        assert r[6] == res[0][6]                         # real pairs
        per = (na + parts - 1) // parts
        mx = r[8][na] // parts
        assert r[8][na] == parts * mx and r[7] == parts * mx >= res[0][7]
        for k in range(parts):
            if k * per < na:
                assert r[8][k * per] == k * mx
    assert res[0][8][na] == res[0][6] + na               # compact: n_blocks = pairs + atoms


def test_reference_procedure_names_on_the_host_mirror(wk, api):
    """the reference-named methods are the same calls as the generic ones, and refuse the wrong parametrisation arm"""
    wl = wk.au_fcc((2, 2, 2))
    c = api.TBContext(0); c.set_workload(wl)
    c.get_near_neighbours()
    H, S = c.construct_TB_H_NRL()
    H2, S2 = c.assemble_HS()
    assert np.array_equal(H, H2) and np.array_equal(S, S2)
    t = c.Construct_Vij_NRL()
    assert t["M_Vij"].shape == (wl["na"], wl["na"], 10)
    C_, E, fe = wk.synthetic_eigen(wl["na"] * 9, seed=4)
    c.set_eigen(C_, E, fe)
    A = c.Construct_Aij(); AE = c.Construct_Aij_x_En()
    A2, AE2 = c.density(dense=True)
    assert np.array_equal(A, A2) and np.array_equal(AE, AE2)
    assert np.array_equal(c.get_dHij_drij_NRL(), c.forces())
    assert np.array_equal(c.Attract_TB_Forces_Press_NRL(), c.pressure())
    for wrong in (c.dHij_s, c.construct_TB_H_Pettifor, c.get_dHij_drij_DFTB, c.Construct_Vij_DFTB):
        with pytest.raises(api.XtbError):
            wrong()
    c.close()
    wlp = wk.diamond64()
    c = api.TBContext(0); c.set_workload(wlp)
    c.get_near_neighbours()
    Hp = c.construct_TB_H_Pettifor()
    assert np.array_equal(Hp, c.assemble_HS()[0])
    with pytest.raises(api.XtbError):
        c.construct_TB_H_NRL()
    c.close()
