"""Readers for the reference's own input files on this path.  Host-side, pure Python.

Each function restates one reader of Source_files/Read_input_data.f90 (list-directed `read(FN,*)`: the first
numeric token(s) of a line, Fortran `d` exponents) and returns the dict the C-ABI structs are filled from
(`_abi.*_from_dict`):

    read_Mehl_TB_Hamiltonian        Read_input_data.f90:3963-4158  ->  xtb_nrl_pair           (NRL)
    read_Pettifor_TB_Hamiltonian    Read_input_data.f90:3377-3661  ->  xtb_pettifor_pair
    read_Pettifor_TB_repulsive      Read_input_data.f90:2908-3056  ->  xtb_pettifor_rep_pair
    read_Exponential_wall_TB        Read_input_data.f90:1920-2040,2313-2356 -> xtb_wall_pair (class line must read exactly
                                    'Simple_wall'; e.g. the shipped 'Simple_wall:exclude' disables the wall, as in the reference)
    read_Short_Rep_TB               Read_input_data.f90:2021-2071,2110-2312 -> xtb_short_pair ('General' class of *_TB_short.txt / *_TB_wall.txt),
                                    with make_cubic_splines (Algebra_tools.f90:115-190) and read_NLH_coefficients (ZBL_potential.f90:79-147)
    read_vdW_TB                     Read_input_data.f90:1660-1840 + read_vdW_Girifalco_TB :2392-2574, read_vdW_LJ_TB :2576-2652,
                                    read_vdW_ILJ_TB :2654-2716 -> (class, xtb_vdw_pair); the shipped 'Girifalco:exclude' disables it
    read_Coulomb_TB                 Read_input_data.f90:1846-1917 + read_Coulomb_cut_TB :2360-2386 -> xtb_coulomb_pair ('Coulomb_cut' only)
    read_SAVE                       Dealing_with_output_files.f90:2120-2158 (writer) / Initial_configuration.f90:1594 (reader)
                                    SAVE_atoms.dat + SAVE_supercell.dat -> koa, S, supce
    read_unit_cell                  Unit_cell_equilibrium.txt + Unit_cell_atom_relative_coordinates.txt
                                    (Initial_configuration.f90:  set_initial_configuration) -> replicated supercell
The matching writers (`write_*`) exist so that tests can round-trip without the reference tree being present.
"""
import numpy as np


def _num(tok):
    return float(tok.lower().replace("d", "e"))


def _lines(path):
    with open(path) as f:
        return f.read().splitlines()


def read_Mehl_TB_Hamiltonian(path):
    """NRL parameter file.  Line 1 name, line 2 `ind_split ind_overlap ind`, line 3 `Rc lden`, then lambda, the on-site
    sets (a,b,c,d) x {s,p,t2g[,eg]}, then 10 channels x (e,f,[g,]h) for H and for S (file order ss,sp,pp_s,pp_p,sd,pd_s,pd_p,
    dd_s,dd_p,dd_d)."""
    lines = _lines(path)
    if lines[0].split()[0] not in ("Mehl", "NRL"):
        raise ValueError(f"{path}: not an NRL parameter file ({lines[0]!r})")
    it = iter(lines[1:])
    t = next(it).split()
    ind_split, ind_overlap, ind = int(t[0]), int(t[1]), int(t[2])
    t = next(it).split()
    p = {"ind_split": ind_split, "ind_overlap": ind_overlap, "Rc": _num(t[0]), "lden": _num(t[1])}

    def nxt():
        return _num(next(it).split()[0])

    p["lambd"] = nxt()
    al, bl, cl, dl = [0.0] * 4, [0.0] * 4, [0.0] * 4, [0.0] * 4
    for s in range(4 if ind_split == 1 else 3):
        al[s] = nxt(); bl[s] = nxt(); cl[s] = nxt(); dl[s] = nxt()
    p.update(al=al, bl=bl, cl=cl, dl=dl)
    for names in (("ellm", "fllm", "gllm", "hllm"), ("pllm", "qllm", "rllm", "sllm")):
        arrs = [[0.0] * 10 for _ in range(4)]
        for c in range(10):
            arrs[0][c] = nxt()
            arrs[1][c] = nxt()
            arrs[2][c] = nxt() if ind > 0 else 0.0
            arrs[3][c] = nxt()
        for n, a in zip(names, arrs):
            p[n] = a
    return p


def read_Pettifor_TB_Hamiltonian(path):
    lines = _lines(path)
    if lines[0].split()[0] != "Pettifor":
        raise ValueError(f"{path}: not a Pettifor parameter file ({lines[0]!r})")
    v = [_num(l.split()[0]) for l in lines[1:] if l.strip()]
    p = {"Es": v[0], "Ep": v[1], "V0": v[2:6], "r0": v[6], "n": v[7], "r1": v[8], "rm": v[9],
         "nc": v[10:14], "rc": v[14:18]}
    c = v[18:34]
    for q, name in enumerate(("c0", "c1", "c2", "c3")):
        p[name] = [c[4 * i + q] for i in range(4)]
    return p


_REP_ORDER = ("E0_TB", "phi0", "m", "mc", "d0", "d1", "dm", "dc")


def read_Pettifor_TB_repulsive(path):
    """file order E0_TB, phi0, m, mc, d0, d1, dm, dc, c0(1:4), a0(1:5)  (Read_input_data.f90:2919-3048)"""
    lines = _lines(path)
    if lines[0].split()[0] != "Pettifor":
        raise ValueError(f"{path}: not a Pettifor repulsive file ({lines[0]!r})")
    v = [_num(l.split()[0]) for l in lines[1:] if l.strip()]
    if len(v) < 17:
        raise ValueError(f"{path}: {len(v)} numbers, 17 expected")
    p = dict(zip(_REP_ORDER, v[:8]))
    p["c0"] = v[8:12]
    p["a0"] = v[12:17]
    return p


def read_Exponential_wall_TB(path):
    """<El>_<El>_TB_wall.txt (Read_input_data.f90:1920-2040): the first list-directed token names the class.  'Simple_wall'
    (any of the three spellings the reference accepts) -> dict(C, r0, d0, dd) read by read_Exponential_wall_TB (:2313-2356).
    Any other token -- including the 'Simple_wall:exclude' / 'Simple_wall:exclde' the reference ships in INPUT_DATA/Aluminum,
    /Silicon and /Diamond -- falls into `case default`: the reference prints "Unknown short-range ... class" and proceeds
    WITHOUT the wall; this reader returns None in that case, and for the 'General' class (read_Short_Rep_TB reads that one)."""
    lines = _lines(path)
    cls = lines[0].replace(",", " ").split()[0]
    if cls not in ("Simple_wall", "SIMPLE_WALL", "simple_wall"):
        return None
    v = [_num(lines[k].split()[0]) for k in range(1, 5)]
    return {"C": v[0], "r0": v[1], "d0": v[2], "dd": v[3]}


def write_Exponential_wall_TB(path, p, enabled=True):
    with open(path, "w") as f:
        f.write("Simple_wall\n" if enabled else "Simple_wall:exclude\n")
        for k, c in (("C", "C [eV]"), ("r0", "r0 [A]"), ("d0", "d0 [A]"), ("dd", "dd [A]")):
            f.write(f"{_f(p[k])}\t\t! {c}\n")


_SR_KEYS = {
    "cut": ('CUTOFF', 'Cutoff', 'cutoff', 'CUT_OFF', 'Cut_off', 'cut_off', 'CUT-OFF', 'Cut-off', 'cut-off', 'FERMI', 'Fermi', 'fermi'),
    "cut_inv": ('CUTOFF_SHORT', 'Cutoff_short', 'cutoff_short', 'CUT_OFF_SHORT', 'Cut_off_short', 'cut_off_short', 'CUT-OFF_short',
                'Cut-off_short', 'cut-off_short', 'FERMI_INV', 'Fermi_inv', 'fermi_inv'),
    "exp": ('EXP', 'Exp', 'exp', 'Exponential', 'exponential', 'EXPONENTIAL'),
    "inv_exp": ('INVEXP', 'InvExp', 'invexp', 'Invexp', 'INV_EXP', 'Inv_Exp', 'Inv_exp', 'inv_exp'),
    "pow": ('POWER', 'Power', 'power', 'POW', 'Pow', 'pow'),
    "zbl": ('ZBL', 'zbl'),
    "nlh": ('NLH', 'nlh'),
    "tab": ('TAB', 'Tab', 'tab', 'TABLE', 'Table', 'table', 'TABULATED', 'Tabulated', 'tabulated'),
}


def make_cubic_splines(x, y):
    """make_cubic_splines, Algebra_tools.f90:115-190 (one-sided first / second differences, not the natural spline) -> a, b, c, d"""
    x = np.asarray(x, dtype=np.float64); y = np.asarray(y, dtype=np.float64)
    n = len(x)
    if n < 2:
        raise ValueError("a table of fewer than two points cannot be splined")
    h = np.empty(n); h[:-1] = x[1:] - x[:-1]; h[-1] = h[-2]
    h_inv = 1.0 / h
    fp = np.empty(n); fp2 = np.empty(n)
    for i in range(1, n + 1):                        # 1-based, as get_first_derivative / get_second_derivative
        j = 1 if i <= 1 else (n - 1 if i >= n else i)
        fp[i - 1] = (y[j] - y[j - 1]) / (x[j] - x[j - 1])
        j = 2 if i <= 2 else (n - 1 if i >= n - 1 else i)
        if n >= 3:
            xnn1 = x[j] - x[j - 1]; xn1n2 = x[j - 1] - x[j - 2]; xnn2 = x[j] - x[j - 2]
            fp2[i - 1] = 2.0 * (y[j] * xn1n2 - y[j - 1] * xnn2 + y[j - 2] * xnn1) / (xnn2 * xn1n2 * xnn1)
        else:
            fp2[i - 1] = 0.0
    a = y.copy(); b = np.zeros(n); c = np.zeros(n); d = np.zeros(n)
    b[n - 1] = fp[n - 1]; b[n - 2] = fp[n - 2]
    for i in range(n - 2, 0, -1):
        b[i - 1] = fp[i] + 0.75 * fp2[i] * h[i - 1]
    for i in range(n - 2, 0, -1):
        c[i - 1] = h_inv[i - 1] * (fp[i] - b[i - 1] - 0.5 * h[i - 1] * fp2[i])
    for i in range(n - 2, 0, -1):
        d[i - 1] = (1.0 / 6.0) * h_inv[i - 1] * (fp2[i] - 2.0 * c[i - 1])
    return a, b, c, d


def read_NLH_coefficients(path, Z1, Z2):
    """get_NLH_coefficients, ZBL_potential.f90:79-147: nlh_coeffs.dat holds the upper triangle Z1 <= Z2 (one comment line, then
    `Z1 Z2 a1 b1 a2 b2 a3 b3 ...` for 92 x 92 elements).  As written, the lower triangle is filled with a(:) = a(j,i) AND
    b(:) = a(j,i) (`NLH_data(i,j)%b(:) = NLH_data(j,i)%a(:)`, :133), so a pair asked for with Z1 > Z2 gets b = a; reproduced."""
    Z1 = int(round(Z1)); Z2 = int(round(Z2))
    lo, hi = min(Z1, Z2), max(Z1, Z2)
    row = None
    for line in _lines(path)[1:]:
        t = line.split()
        if len(t) >= 8 and int(t[0]) == lo and int(t[1]) == hi:
            row = [_num(x) for x in t[2:8]]
            break
    if row is None:
        raise ValueError(f"{path}: no NLH coefficients for Z = {lo}, {hi}")
    a = (row[0], row[2], row[4]); b = (row[1], row[3], row[5])
    return (a, b) if Z1 <= Z2 else (a, a)


def read_Short_Rep_TB(path, Z=None, nlh_file=None):
    """'General' class of <El>_<El>_TB_short.txt / _TB_wall.txt (class line 'General' | 'general' | 'GENERAL', Read_input_data.f90:2021;
    anything else, e.g. the shipped 'General:exclude', -> None).  read_Short_Rep_TB (:2110-2146) then reads list-directed keyword
    records until the end of the file (interpret_short_range_data :2149-2235): Cut-off `d0 dd`; Exp `Phi r0 a`; Inv_exp `C r0`;
    Pow `n` + n lines `Phi r0 m`; ZBL; NLH (coefficients from nlh_file for Z = (Z1, Z2)); Tab `n [diff|spline]` + n lines `R E`
    (Process_tabulated_potential :2238-2312; the splines are made in either case).  Unknown first tokens are skipped, as there.
    Defaults: d0 = 0, dd = 0.01 (no repulsion).  Cut-off_short is read and, as in the reference, never used."""
    lines = _lines(path)
    if not lines or not lines[0].strip():
        return None
    if lines[0].replace(",", " ").split()[0] not in ("General", "general", "GENERAL"):
        return None
    p = {"d0": 0.0, "dd": 0.01, "exp": None, "inv_exp": None, "pow": [], "zbl": False, "nlh": False, "tab": None}
    k = 1

    def next_record(k):                 # list-directed read: blank records are skipped
        while k < len(lines) and not lines[k].replace(",", " ").split():
            k += 1
        return k

    while True:
        k = next_record(k)
        if k >= len(lines):
            break
        key = lines[k].replace(",", " ").split()[0]
        k += 1
        if key in _SR_KEYS["cut"]:
            k = next_record(k); p["d0"], p["dd"] = _leading_numbers(lines[k], 2); k += 1
        elif key in _SR_KEYS["cut_inv"]:
            k = next_record(k); p["cut_inv"] = tuple(_leading_numbers(lines[k], 2)); k += 1
        elif key in _SR_KEYS["exp"]:
            k = next_record(k); p["exp"] = tuple(_leading_numbers(lines[k], 3)); k += 1
        elif key in _SR_KEYS["inv_exp"]:
            k = next_record(k); p["inv_exp"] = tuple(_leading_numbers(lines[k], 2)); k += 1
        elif key in _SR_KEYS["pow"]:
            k = next_record(k); n = int(_leading_numbers(lines[k], 1)[0]); k += 1
            p["pow"] = []
            for _ in range(n):
                k = next_record(k); p["pow"].append(tuple(_leading_numbers(lines[k], 3))); k += 1
        elif key in _SR_KEYS["zbl"]:
            p["zbl"] = True
        elif key in _SR_KEYS["nlh"]:
            p["nlh"] = True
        elif key in _SR_KEYS["tab"]:
            t = lines[k].replace(",", " ").split(); k += 1          # formatted read: the very next line
            n = int(_num(t[0]))
            spline = len(t) > 1 and t[1] in ("spline", "SPLINE")
            R, E = [], []
            for _ in range(n):
                k = next_record(k); r, e = _leading_numbers(lines[k], 2); R.append(r); E.append(e); k += 1
            a, b, c, d = make_cubic_splines(R, E)
            p["tab"] = {"R": np.array(R), "E": np.array(E), "spline": bool(spline), "a": a, "b": b, "c": c, "d": d}
    if p["nlh"]:
        if Z is None or nlh_file is None:
            raise ValueError(f"{path}: the NLH term needs the atomic numbers of the pair and nlh_coeffs.dat")
        p["nlh_a"], p["nlh_b"] = read_NLH_coefficients(nlh_file, Z[0], Z[1])
    return p


_GIRIFALCO_ORDER = ("C12", "C6", "dm", "d_cut", "a", "b", "c", "d", "dsm", "ds_cut", "as", "bs", "cs", "ds", "es", "fs")
_LJ_NAMES = ("LJ", "Lennar-Jones", "Lennard_Jones", "lj", "Lj", "LENNARD-JONES", "LENNARD_JONES", "lennard-jones", "lennard_jones")
_ILJ_NAMES = ("ILJ", "ilj", "Improved_LJ", "Ilj", "improved_lj", "I_LJ", "i_lj")


def _leading_numbers(line, most):
    """what a list-directed `read(line,*) a, b, ...` would take: numeric tokens until the first that is not a number"""
    out = []
    for tok in line.replace(",", " ").split():
        if len(out) == most:
            break
        try:
            out.append(_num(tok))
        except ValueError:
            break
    return out


def read_vdW_TB(path):
    """<El>_<El>_TB_vdW.txt (Read_input_data.f90:1709-1840): the first list-directed token names the class.
      'Girifalco'                   -> ('Girifalco', dict of the 16 numbers in file order; d0_cut = d_cut, dd_cut = 0 as :1799-1800)
      'LJ' (and spellings :1732)    -> ('LJ', eps, r0, n from the 'SE' / 'AB' / 'n-exp' form (:2611-2623), d0_short dd_short, d0_cut dd_cut)
      'ILJ' (and spellings :1750)   -> ('ILJ', eps, r0, n, m (defaults 12, 6), d0_short dd_short, d0_cut dd_cut)
    Any other token -- including the 'Girifalco:exclude' the reference ships for diamond, graphite, ... -- falls into
    `case default`: the reference prints "Unknown TB-vdW parametrization class" and proceeds WITHOUT van der Waals forces; this
    reader returns None.  'Dumitrica' ("UNFINISHED, DO NOT USE", returns zero energy and force) is refused."""
    lines = _lines(path)
    cls = lines[0].replace(",", " ").split()[0]
    if cls == "Dumitrica":
        raise NotImplementedError(f"{path}: class TB_vdW_Dumitrica is unfinished in the reference and not part of this graft")
    if cls == "Girifalco":
        v = [_num(lines[k].split()[0]) for k in range(1, 17)]
        p = dict(zip(_GIRIFALCO_ORDER, v))
        p["d0_cut"] = p["d_cut"]; p["dd_cut"] = 0.0
        return "Girifalco", p
    if cls in _LJ_NAMES:
        form = lines[1].split()[0]
        ab = _leading_numbers(lines[2], 3)
        if len(ab) < 2:
            raise ValueError(f"{path}: could not read the LJ coefficients")
        A, B = ab[0], ab[1]
        n = ab[2] if len(ab) == 3 else 6.0
        p = {"eps": 0.0, "r0": 0.0, "n": n}                         # defaults :1742-1744 stay when the form is not recognised
        if form in ("SE", "Se", "se", "sE", "Segma-E", "sigma-e", "SIGMA-E"):
            p["eps"] = A / 8.0; p["r0"] = 2.0 ** 0.5 * B
        elif form in ("AB", "Ab", "ab", "aB"):
            p["eps"] = B ** 2 / (4.0 * A); p["r0"] = (A / B) ** (1.0 / 6.0)
        elif form in ("n-exp", "N-exp", "N-Exp", "N-EXP"):
            p["eps"] = A; p["r0"] = B
        p["d0_short"], p["dd_short"] = _leading_numbers(lines[3], 2)
        p["d0_cut"], p["dd_cut"] = _leading_numbers(lines[4], 2)
        return "LJ", p
    if cls in _ILJ_NAMES:
        ab = _leading_numbers(lines[1], 4)
        if len(ab) < 2:
            raise ValueError(f"{path}: could not read the ILJ coefficients")
        p = {"eps": ab[0], "r0": ab[1], "n": ab[2] if len(ab) > 2 else 12.0, "m": ab[3] if len(ab) > 3 else 6.0}
        p["d0_short"], p["dd_short"] = _leading_numbers(lines[2], 2)
        p["d0_cut"], p["dd_cut"] = _leading_numbers(lines[3], 2)
        return "ILJ", p
    return None


def write_vdW_TB(path, kind, p, enabled=True):
    with open(path, "w") as f:
        if kind == "Girifalco":
            f.write("Girifalco\n" if enabled else "Girifalco:exclude\n")
            for k in _GIRIFALCO_ORDER:
                f.write(f"{_f(p[k])}\t\t! {k}\n")
        elif kind == "LJ":
            f.write(("LJ\n" if enabled else "LJ:exclude\n") + "n-exp\n")
            f.write(f"{_f(p['eps'])}\t{_f(p['r0'])}\t{_f(p['n'])}\t! epsylon, r0, n\n")
            f.write(f"{_f(p['d0_short'])}\t{_f(p['dd_short'])}\t! d0_short, dd_short\n")
            f.write(f"{_f(p['d0_cut'])}\t{_f(p['dd_cut'])}\t! d0_cut, dd_cut\n")
        elif kind == "ILJ":
            f.write("ILJ\n" if enabled else "ILJ:exclude\n")
            f.write(f"{_f(p['eps'])}\t{_f(p['r0'])}\t{_f(p['n'])}\t{_f(p['m'])}\t! epsylon, r0, n, m\n")
            f.write(f"{_f(p['d0_short'])}\t{_f(p['dd_short'])}\t! d0_short, dd_short\n")
            f.write(f"{_f(p['d0_cut'])}\t{_f(p['dd_cut'])}\t! d0_cut, dd_cut\n")
        else:
            raise ValueError(kind)


COULOMB_K = 1.602176634e-19 / (4.0 * 3.1415926535897932384626433832795 * 8.854187817620e-12) * 1.0e10     # g_e/(4 g_Pi g_e0)*1d10, :1907


def read_Coulomb_TB(path):
    """<El>_<El>_TB_Coulomb.txt (Read_input_data.f90:1846-1917): class line 'Coulomb_cut' -> dict(k, dm, dd), k set by the reader
    (:1907).  Any other first token -- the shipped 'Coulomb_cut:exclude' of INPUT_DATA/Diamond, /C, ... -- falls into
    `case default` ("Proceeding without Coulomb forces from unballanced charge"): None.  'Cutie' (testing only) is refused."""
    lines = _lines(path)
    cls = lines[0].replace(",", " ").split()[0]
    if cls == "Cutie":
        raise NotImplementedError(f"{path}: class Cutie is a test stub of the reference")
    if cls != "Coulomb_cut":
        return None
    return {"k": COULOMB_K, "dm": _num(lines[1].split()[0]), "dd": _num(lines[2].split()[0])}


def write_Coulomb_TB(path, p, enabled=True):
    with open(path, "w") as f:
        f.write("Coulomb_cut\n" if enabled else "Coulomb_cut:exclude\n")
        f.write(f"{_f(p['dm'])}\t! dm\n{_f(p['dd'])}\t! dd\n")


def read_SAVE(atoms_path, cell_path):
    """SAVE_atoms.dat: `KOA S(1:3) ...` per atom; SAVE_supercell.dat: file line i = COLUMN i of supce
    (`Scell%supce(:,i) = temp_vec`, Initial_configuration.f90:2409-2418)."""
    koa, S = [], []
    for line in _lines(atoms_path):
        t = line.split()
        if not t:
            continue
        koa.append(int(t[0]))
        S.append([_num(x) for x in t[1:4]])
    rows = [l.split() for l in _lines(cell_path)]
    rows = [r for r in rows if r][:3]
    cols = [[_num(x) for x in r] for r in rows]
    supce = [[cols[b][a] for b in range(3)] for a in range(3)]
    return {"koa": koa, "S": S, "supce": supce}


def read_unit_cell(cell_path, coords_path, reps):
    """Unit_cell_equilibrium.txt (first three lines; line i = column i of the unit cell) and
    Unit_cell_atom_relative_coordinates.txt (`KOA s1 s2 s3`), replicated as set_initial_coords does
    (Initial_configuration.f90:2241-2285: loops nx, ny, nz outermost-first, unit-cell atoms innermost;
    S = (s + n)/ncell; the reference's +-1e-10 random shift is left out) and wrapped into [0,1)
    (check_periodic_boundaries).  supce(:,i) = ncell_i * unit_cell(:,i) (:2546-2548)."""
    rows = [l.split() for l in _lines(cell_path)]
    rows = [r for r in rows if r][:3]
    unit = np.array([[_num(x) for x in r[:3]] for r in rows]).T          # unit[:, i] = file line i
    koa0, rel = [], []
    for line in _lines(coords_path):
        t = line.split()
        if len(t) < 4:
            continue
        koa0.append(int(t[0]))
        rel.append([_num(x) for x in t[1:4]])
    rel = np.array(rel)
    ncell = np.array(reps, dtype=np.float64)
    koa, S = [], []
    for nx in range(reps[0]):
        for ny in range(reps[1]):
            for nz in range(reps[2]):
                for k in range(len(koa0)):
                    S.append((rel[k] + np.array([nx, ny, nz], dtype=np.float64)) / ncell)
                    koa.append(koa0[k])
    S = np.mod(np.array(S), 1.0)
    supce = unit * ncell[None, :]
    return {"koa": koa, "S": S.tolist(), "supce": supce.tolist()}


# ---- writers in the reference's file formats (tests round-trip through them) ----------------------------------
def _f(x):
    return f"{x:.16e}".replace("e", "d")


def write_Pettifor_TB_repulsive(path, p):
    vals = [p[k] for k in _REP_ORDER] + list(p["c0"]) + list(p["a0"])
    with open(path, "w") as f:
        f.write("Pettifor\n" + "".join(f"{_f(v)}\t! \n" for v in vals))


def write_Pettifor_TB_Hamiltonian(path, p):
    vals = [p["Es"], p["Ep"]] + list(p["V0"]) + [p["r0"], p["n"], p["r1"], p["rm"]] + list(p["nc"]) + list(p["rc"])
    for i in range(4):
        vals += [p["c0"][i], p["c1"][i], p["c2"][i], p["c3"][i]]
    with open(path, "w") as f:
        f.write("Pettifor\n" + "".join(f"{_f(v)}\n" for v in vals))


def write_Mehl_TB_Hamiltonian(path, p):
    out = ["NRL", f"{p['ind_split']} {p['ind_overlap']} 1", f"{_f(p['Rc'])} {_f(p['lden'])}", _f(p["lambd"])]
    for s in range(4 if p["ind_split"] == 1 else 3):
        out += [_f(p[k][s]) for k in ("al", "bl", "cl", "dl")]
    for names in (("ellm", "fllm", "gllm", "hllm"), ("pllm", "qllm", "rllm", "sllm")):
        for c in range(10):
            out += [_f(p[n][c]) for n in names]
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")


def write_SAVE(atoms_path, cell_path, cell):
    with open(atoms_path, "w") as f:
        for k, s in zip(cell["koa"], cell["S"]):
            f.write(f"  {int(k)}  " + "  ".join(f"{x:.16E}" for x in s) + "\n")
    h = np.array(cell["supce"])
    with open(cell_path, "w") as f:
        for i in range(3):
            f.write(" " + " ".join(f"{h[a, i]:.16E}" for a in range(3)) + "\n")
