#!/usr/bin/env python
"""S*Ha part of xtb_coupling at au4096: tiled kernel (default) vs one column atom per CTA (XTB_COUPLING_TILE=0), same box."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import torch
wk = importlib.import_module("xtant-3_b200.workloads"); api = importlib.import_module("xtant-3_b200.api")
wl = wk.CONFIGS["au4096"](); N = wl["na"] * 9
ctx = api.TBContext(0); ctx.set_stream(torch.cuda.current_stream().cuda_stream); ctx.set_workload(wl)
ctx.get_near_neighbours_device(); ctx.assemble_blocks()
Cm, dE, dfe, E, fe = bench.make_eigen_device(torch, wk, wl, N)
ctx.set_eigen_device(Cm.data_ptr(), dE.data_ptr(), dfe.data_ptr()); ctx.save_eigen_prev()
out = {"N": N, "n_blocks": ctx.counter(4)}
for tag, env in (("tiled", "1"), ("per_atom", "0"), ("tiled_again", "1")):
    os.environ["XTB_COUPLING_TILE"] = env
    ctx.Electron_ion_coupling_Mij(0, to_host=False)
    out[tag] = {"total_ms": ctx.stage_ms(11), "norms_plus_SHa_ms": ctx.stage_ms(12), "dgemm_ms": ctx.stage_ms(13)}
    print(json.dumps(out), flush=True)
