"""two steps of the ac1000 (Pettifor sp3) configuration: target of `ncu -k regex:'density_warp_kernel|rows_by_state_kernel'`"""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wk = importlib.import_module("xtant-3_b200.workloads"); api = importlib.import_module("xtant-3_b200.api")
wl = wk.ac1000(); c = api.TBContext(0); c.set_workload(wl)
for _ in range(2):
    c.get_near_neighbours_device(); c.assemble_HS(to_host=False); Ei = c.eigensolve()[0]
    fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"]); c.set_occupations(fe); c.density(); c.forces_device()
print("ok", c.stage_ms(3), c.stage_ms(7))
