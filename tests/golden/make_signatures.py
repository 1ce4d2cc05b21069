#!/usr/bin/env python
"""Records the argument lists of the reference procedures the Fortran shim re-exposes (interface facts only: procedure
name, home file, dummy-argument names in order), read from /root/reference/Source_files in the build container.
Output: tests/golden/reference_signatures.json, compared with the shim by tests/test_abi_cpu.py.

    python tests/golden/make_signatures.py
"""
import glob
import json
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
NAMES = ["get_near_neighbours", "Construct_Vij_NRL", "construct_TB_H_NRL", "Construct_Aij", "Construct_Aij_x_En",
         "get_dHij_drij_NRL", "Attract_TB_Forces_Press_NRL", "construct_TB_H_Pettifor", "dHij_s", "Attract_TB_Forces_Press",
         "get_Erep_s", "dErdr_s", "dErdr_Pressure_s", "Construct_Vij_DFTB", "construct_TB_H_DFTB", "get_dHij_drij_DFTB",
         "Attract_TB_Forces_Press_DFTB", "Electron_ion_coupling_Mij",
         "get_Exp_wall_s", "d_Exp_wall_pot_s", "d_Exp_wall_Pressure_s", "get_vdW_s", "vdW_forces", "get_mirror_cell_num",
         "get_Coulomb_s", "Coulomb_forces",
         "get_short_range_rep_s", "d_Short_range_pot_s", "d_Short_range_Pressure_s"]


def subroutines(text):
    text = re.sub(r"&\s*\n\s*&?", " ", text)
    out = {}
    for m in re.finditer(r"^\s*subroutine\s+(\w+)\s*\(([^)]*)\)", text, flags=re.M | re.I):
        out.setdefault(m.group(1).lower(), [a.strip().lower() for a in m.group(2).split(",") if a.strip()])
    return out


def main():
    found = {}
    for f in sorted(glob.glob("/root/reference/Source_files/*.f90")):
        for k, v in subroutines(open(f, errors="ignore").read()).items():
            found.setdefault(k, {"file": os.path.basename(f), "args": v})
    out = {n: found[n.lower()] for n in NAMES}
    json.dump(out, open(os.path.join(HERE, "reference_signatures.json"), "w"), indent=1)
    print(len(out), "signatures")


if __name__ == "__main__":
    main()
