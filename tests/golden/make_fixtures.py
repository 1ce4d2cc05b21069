#!/usr/bin/env python
"""Generate the INPUT fixtures of the parity tests from the reference's own data files.

Run once in the build container (needs /root/reference, which does not exist on the GPU box):
    python tests/golden/make_fixtures.py
It writes
    xtant-3_b200/data/params_nrl_{Au,Al,Si}.json    NRL parameter sets
    xtant-3_b200/data/params_pettifor{,_rep}_Si.json   second Pettifor-type set (Silicon/Pettifor): physics pin of the oracle
    xtant-3_b200/data/params_pettifor_C.json        Pettifor parameter set
    xtant-3_b200/data/params_pettifor_rep_C.json    Pettifor repulsive parameter set
    xtant-3_b200/data/params_vdw_{C_amorphous,I_lj}.json   van der Waals sets: Girifalco (INPUT_DATA/Amorphous_Carbon, the only
                                                    carbon folder of the configurations whose class line is active) and LJ (PbI2)
    xtant-3_b200/data/params_short_{I_I,Pb_I,Pb_Pb,Fe_Fe,O_O_tab}.json   'General' short-range sets: Pow + ZBL (PbI2), Exp (Fe), Tab (PbO2)
    xtant-3_b200/data/cell_ac216.json               the reference's own 216-atom amorphous-carbon cell
    xtant-3_b200/data/cell_{diamond64,al500,au108}.json   relative coords + supercell
    xtant-3_b200/data/dftb_{matsci_SiAl,mio_CH}.npz  DFTB Slater-Koster tables (INPUT_DATA/DFTB/<set>/<A>-<B>.skf, read by
                                                     xtant-3_b200/skf.py = read_skf_file, Dealing_with_DFTB.f90:48-149)
The parsers (xtant-3_b200/readers.py) restate the reference readers:
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


sys.path.insert(0, os.path.join(HERE, "..", ".."))
_rd = importlib.import_module("xtant-3_b200.readers")      # the parsers live in the package (host mirror of Read_input_data.f90)
parse_nrl, parse_pettifor, parse_pettifor_rep, parse_cell = (_rd.read_Mehl_TB_Hamiltonian, _rd.read_Pettifor_TB_Hamiltonian,
                                                             _rd.read_Pettifor_TB_repulsive, _rd.read_SAVE)


def short(path):
    p = _rd.read_Short_Rep_TB(path)
    if p["tab"] is not None:            # the splines are remade on load (workloads.load_short)
        p["tab"] = {"R": [float(x) for x in p["tab"]["R"]], "E": [float(x) for x in p["tab"]["E"]], "spline": p["tab"]["spline"]}
    return p


def vdw(path):
    kind, p = _rd.read_vdW_TB(path)
    return dict(p, kind=kind)


def main():
    os.makedirs(OUT, exist_ok=True)
    jobs = {
        "params_nrl_Au.json": lambda: parse_nrl(f"{REF}/Gold/NRL/Au_Au_TB_Hamiltonian_parameters.txt"),
        "params_nrl_Al.json": lambda: parse_nrl(f"{REF}/Aluminum/NRL/Al_Al_TB_Hamiltonian_parameters.txt"),
        "params_nrl_Si.json": lambda: parse_nrl(f"{REF}/Silicon/NRL/Si_Si_TB_Hamiltonian_parameters.txt"),
        "params_pettifor_C.json": lambda: parse_pettifor(f"{REF}/Diamond/C_C_TB_Hamiltonian_parameters.txt"),
        "params_pettifor_rep_C.json": lambda: parse_pettifor_rep(f"{REF}/Diamond/C_C_TB_Repulsive_parameters.txt"),
        "params_pettifor_Si.json": lambda: parse_pettifor(f"{REF}/Silicon/Pettifor/Si_Si_TB_Hamiltonian_parameters.txt"),
        "params_pettifor_rep_Si.json": lambda: parse_pettifor_rep(f"{REF}/Silicon/Pettifor/Si_Si_TB_Repulsive_parameters.txt"),
        "cell_diamond64.json": lambda: parse_cell(f"{REF}/Diamond/64_atoms/SAVE_atoms.dat",
                                                  f"{REF}/Diamond/64_atoms/SAVE_supercell.dat"),
        "cell_al500.json": lambda: parse_cell(f"{REF}/Aluminum/SAVE_atoms_500atoms.dat",
                                              f"{REF}/Aluminum/SAVE_supercell_500atoms.dat"),
        "params_vdw_C_amorphous.json": lambda: vdw(f"{REF}/Amorphous_Carbon/C_C_TB_vdW.txt"),
        "params_vdw_I_lj.json": lambda: vdw(f"{REF}/PbI2/I_I_TB_vdW.txt"),
        "params_short_I_I.json": lambda: short(f"{REF}/PbI2/I_I_TB_short.txt"),
        "params_short_Pb_I.json": lambda: short(f"{REF}/PbI2/Pb_I_TB_short.txt"),
        "params_short_Pb_Pb.json": lambda: short(f"{REF}/PbI2/Pb_Pb_TB_short.txt"),
        "params_short_Fe_Fe.json": lambda: short(f"{REF}/Fe/Fe_Fe_TB_short.txt"),
        "params_short_O_O_tab.json": lambda: short(f"{REF}/PbO2/O_O_TB_short.txt"),
        "cell_ac216.json": lambda: parse_cell(f"{REF}/Amorphous_Carbon/SAVE_atoms.dat", f"{REF}/Amorphous_Carbon/SAVE_supercell.dat"),
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
