"""xtant-3_b200 -- B200-native tight-binding hot path of XTANT-3 (H/S assembly, rho/rho_E,
Hellmann-Feynman forces and supercell force) behind a C ABI (include/xtb_b200.h).

The directory name is not a Python identifier; import it with
    importlib.import_module("xtant-3_b200")
"""
__version__ = "0.1.0"
