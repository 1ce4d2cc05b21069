"""Why does density_dmma_kernel take 3.5 ms under ncu for the 32-atom smoke() cell?  Stage times of repeated calls, no ncu."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
wk = importlib.import_module("xtant-3_b200.workloads"); api = importlib.import_module("xtant-3_b200.api")
for name, wl in (("au32", wk.au_fcc((2, 2, 2), sigma=0.1, seed=21)), ("au108", wk.au108()), ("si216", wk.si216())):
    C_, E, fe = wk.synthetic_eigen(wl["na"] * 9, seed=5)
    c = api.TBContext(0); c.set_workload(wl); c.get_near_neighbours_device(); c.set_eigen(C_, E, fe)
    t = []
    for _ in range(6):
        c.density(); t.append((round(c.stage_ms(3), 4), round(c.stage_ms(7), 4), round(c.stage_ms(16), 4)))
    print(name, "density stage / contraction kernel / retile [ms] per call:", t, "slots,stages:", c.counter(9), c.counter(10), flush=True)
    for lead in ("0",):
        os.environ["XTB_DENSITY_LEAD"] = lead
        c2 = api.TBContext(0); c2.set_workload(wl); c2.get_near_neighbours_device(); c2.set_eigen(C_, E, fe)
        t = []
        for _ in range(4):
            c2.density(); t.append(round(c2.stage_ms(7), 4))
        print(name, "lead=0 kernel ms:", t, flush=True)
        del os.environ["XTB_DENSITY_LEAD"]; c2.close()
    c.close()
