#!/usr/bin/env python
"""Generate the INPUT fixtures of the parity tests from the reference's own data files.

Run once in the build container (needs /root/reference, which does not exist on the GPU box):
    python tests/golden/make_fixtures.py
It writes
    xtant-3_b200/data/params_nrl_{Au,Al,Si}.json    NRL parameter sets
    xtant-3_b200/data/params_pettifor_C.json        Pettifor parameter set
    xtant-3_b200/data/cell_{diamond64,al500,au108}.json   relative coords + supercell
    xtant-3_b200/data/dftb_{matsci_SiAl,mio_CH}.npz  DFTB Slater-Koster tables (INPUT_DATA/DFTB/<set>/<A>-<B>.skf, read by
                                                     xtant-3_b200/skf.py = read_skf_file, Dealing_with_DFTB.f90:48-149)
The parsers restate the reference readers:
    read_Mehl_TB_Hamiltonian      Source_files/Read_input_data.f90:3963-4158
    read_Pettifor_TB_Hamiltonian  Source_files/Read_input_data.f90:3377-3661
    SAVE_atoms.dat / SAVE_supercell.dat  Source_files/Dealing_with_output_files.f90:2120-2158
      (file line i of SAVE_supercell.dat = column i of supce)
Only numbers are copied (they are model parameters / coordinates, not code).
"""
import importlib
import json
import os
import re
import sys

REF = "/root/reference/INPUT_DATA"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "..", "xtant-3_b200", "data")


def _num(tok):
    return float(tok.lower().replace("d", "e"))


def _first_numbers(path):
    """first numeric token(s) of each line after the name line"""
    with open(path) as f:
        lines = f.read().splitlines()
    return lines


def parse_nrl(path):
    lines = _first_numbers(path)
    assert lines[0].split()[0] in ("Mehl", "NRL"), lines[0]
    it = iter(lines[1:])
    t = next(it).split()
    ind_split, ind_overlap, ind = int(t[0]), int(t[1]), int(t[2])
    t = next(it).split()
    Rc, lden = _num(t[0]), _num(t[1])

    def nxt():
        return _num(next(it).split()[0])

    p = {"ind_split": ind_split, "ind_overlap": ind_overlap, "Rc": Rc, "lden": lden}
    p["lambd"] = nxt()
    al, bl, cl, dl = [0.0] * 4, [0.0] * 4, [0.0] * 4, [0.0] * 4
    for s in range(3):
        al[s] = nxt(); bl[s] = nxt(); cl[s] = nxt(); dl[s] = nxt()
    if ind_split == 1:
        al[3] = nxt(); bl[3] = nxt(); cl[3] = nxt(); dl[3] = nxt()
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


def parse_pettifor(path):
    lines = _first_numbers(path)
    assert lines[0].split()[0] == "Pettifor"
    v = [_num(l.split()[0]) for l in lines[1:] if l.strip()]
    p = {"Es": v[0], "Ep": v[1], "V0": v[2:6], "r0": v[6], "n": v[7], "r1": v[8], "rm": v[9],
         "nc": v[10:14], "rc": v[14:18]}
    c = v[18:34]
    p["c0"] = [c[4 * i + 0] for i in range(4)]
    p["c1"] = [c[4 * i + 1] for i in range(4)]
    p["c2"] = [c[4 * i + 2] for i in range(4)]
    p["c3"] = [c[4 * i + 3] for i in range(4)]
    return p


def parse_cell(atoms_path, cell_path):
    koa, S = [], []
    with open(atoms_path) as f:
        for line in f:
            t = line.split()
            if not t:
                continue
            koa.append(int(t[0]))
            S.append([_num(x) for x in t[1:4]])
    with open(cell_path) as f:
        rows = [l.split() for l in f.read().splitlines()]
    rows = [r for r in rows if r][:3]
    # file line i = column i of supce  ->  supce[a][b] = rows[b][a]
    cols = [[_num(x) for x in r] for r in rows]
    supce = [[cols[b][a] for b in range(3)] for a in range(3)]
    return {"koa": koa, "S": S, "supce": supce}


def main():
    os.makedirs(OUT, exist_ok=True)
    jobs = {
        "params_nrl_Au.json": lambda: parse_nrl(f"{REF}/Gold/NRL/Au_Au_TB_Hamiltonian_parameters.txt"),
        "params_nrl_Al.json": lambda: parse_nrl(f"{REF}/Aluminum/NRL/Al_Al_TB_Hamiltonian_parameters.txt"),
        "params_nrl_Si.json": lambda: parse_nrl(f"{REF}/Silicon/NRL/Si_Si_TB_Hamiltonian_parameters.txt"),
        "params_pettifor_C.json": lambda: parse_pettifor(f"{REF}/Diamond/C_C_TB_Hamiltonian_parameters.txt"),
        "cell_diamond64.json": lambda: parse_cell(f"{REF}/Diamond/64_atoms/SAVE_atoms.dat",
                                                  f"{REF}/Diamond/64_atoms/SAVE_supercell.dat"),
        "cell_al500.json": lambda: parse_cell(f"{REF}/Aluminum/SAVE_atoms_500atoms.dat",
                                              f"{REF}/Aluminum/SAVE_supercell_500atoms.dat"),
        "cell_au108.json": lambda: parse_cell(f"{REF}/Gold/SAVE_atoms.dat", f"{REF}/Gold/SAVE_supercell.dat"),
    }
    for name, fn in jobs.items():
        d = fn()
        with open(os.path.join(OUT, name), "w") as f:
            json.dump(d, f, indent=0, separators=(",", ":"))
        print("wrote", name, {k: (len(v) if isinstance(v, list) else v) for k, v in d.items() if k in ("koa", "Rc", "rm")})
    # DFTB table sets: pair index (KOA1-1) + n*(KOA2-1) <- file "<El(KOA1)>-<El(KOA2)>.skf" (Read_input_data.f90:4224)
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    skf = importlib.import_module("xtant-3_b200.skf")
    for name, folder, els in (("matsci_SiAl", "matsci-0-3", ["Si", "Al"]), ("mio_CH", "mio-1-1", ["C", "H"])):
        pairs = [skf.parse_skf(f"{REF}/DFTB/{folder}/{a}-{b}.skf", a == b) for b in els for a in els]
        skf.save_set(name, els, pairs)
        print("wrote", f"dftb_{name}.npz", els, [p["Vr"].shape for p in pairs])


if __name__ == "__main__":
    sys.exit(main())
