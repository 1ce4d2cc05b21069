#!/usr/bin/env python
"""bench.py -- ms per MD step of the tight-binding hot path (neighbour list + H/S build + rho/rho_E + HF forces)
on the BASELINE.json north-star workload (synthetic 4096-atom Au fcc supercell, NRL spd TB), FP64.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload au4096]

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for what each key means.
  value      device-timed ms/step with every input resident in HBM (eigenvectors included)
  e2e        the same step through the C ABI with HOST buffers, copies inside the timed region: S in, Ei out, fe in,
             forces out (eigensolve on the device, the north-star flow); e2e.host_eigensolver is the reference's own
             flow with H_non + Sij out to host LAPACK and Ha + Ei + fe back in
  roofline   dominant kernel (rho/rho_E contraction) against the FP64 tensor peak measured in this run
  cpu_baseline  the oracle (CPU restatement of the reference) on a bounded sample of the same workload
The eigensolve between assembly and density is outside the graft (BASELINE.json) and is not timed:
eigenvectors are a synthetic resident input.
--impl reference: the reference's CPU path (the oracle port; the Fortran cannot be built here) on the host cores.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "ms/MD-step (H/S build + rho + HF forces)"


def _density_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/), or None"""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_density_traffic.json")))
    if not files:
        return None
    try:
        return json.load(open(files[-1]))["dram_bytes_per_launch"]
    except Exception:
        return None


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            t = [x.strip() for x in line.split(",")]
            if len(t) < 7:
                continue
            try:
                sm.append(float(t[0])); mx.append(float(t[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), t[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        return out


def fp64_peak_tflops(torch, ctx):
    """FP64 tensor denominators, measured in this run (MEASURED_PEAKS.json carries HBM and bf16 only):
    cuBLAS DGEMM 8192^3 through torch.matmul (burst, best of 4) and the DMMA.8x8x4 issue-rate probe."""
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    best = 0.0
    for i in range(5):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
        if i:
            best = max(best, 2 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) * 1e-12)
    del a, b, c
    torch.cuda.empty_cache()
    # the DMMA issue-rate probe lives in its own library (not in the product .so)
    probe = C.CDLL(os.path.join(ROOT, "xtant-3_b200", "libxtb_b200_probe.so"), mode=C.RTLD_GLOBAL)
    f = probe.xtb_debug_fp64_peak
    f.restype = C.c_double; f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    return best, float(f(ctx.h, 0, 16, 3))


def make_eigen_device(torch, wk, wl, N):
    """synthetic eigen-solution resident in HBM: C ~ N(0,1)/sqrt(N), E sorted uniform(-10,40),
    f = Fermi at 61 % filling.  An (N, N) block of i.i.d. numbers: read as Fortran Ha(N,N) (state n contiguous at offset n*N,
    xtb_set_eigen_device) or as the state-minor HaT (row r contiguous at offset r*N, xtb_set_eigen_device_t -- the layout
    xtb_eigensolve leaves on the device) it is the same kind of synthetic eigenvector matrix."""
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    Cm = torch.empty((N, N), dtype=torch.float64, device="cuda")
    step = 2048
    for s in range(0, N, step):
        Cm[s:s + step].normal_(0.0, 1.0, generator=g)
    Cm.mul_(1.0 / np.sqrt(N))
    rng = np.random.default_rng(7)
    E = np.sort(rng.uniform(-10.0, 40.0, size=N))
    fe, mu = wk.occupations(E, 2.0 * 0.61 * N, wl["Te_eV"])
    return Cm, torch.from_numpy(E).cuda(), torch.from_numpy(fe).cuda(), E, fe


def run_own(args):
    import torch
    import torch.distributed as dist
    wk = importlib.import_module("xtant-3_b200.workloads")
    api = importlib.import_module("xtant-3_b200.api")
    abi = importlib.import_module("xtant-3_b200._abi")
    interop = importlib.import_module("xtant-3_b200.interop")
    mg = importlib.import_module("xtant-3_b200.multigpu")
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = wk.CONFIGS[args.workload]()
    if wl["tb"] != "nrl":
        raise SystemExit("bench.py times the NRL branch (north star); other configs are parity-test cases")
    na, N = wl["na"], wl["na"] * 9
    ctx = api.TBContext(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    ctx.set_workload(wl)
    a0, a1 = mg.partition(na, world, rank)      # = the range xtb_comm_init gives this rank
    Cm, dE, dfe, E, fe = make_eigen_device(torch, wk, wl, N)
    # headline: eigenvectors resident in the layout the on-device eigensolve leaves them in (state-minor; its last GEMM writes
    # it at no extra cost), which the density kernel streams through a tensor map -- no re-tiled copy.  XTB_BENCH_EIGEN_LAYOUT=
    # fortran: Fortran-layout Ha(N,N) as a host eigensolver would upload it (re-tiling pass inside the density stage); that
    # flow is also timed below as `variant_fortran_layout_eigenvectors`.
    state_minor = os.environ.get("XTB_BENCH_EIGEN_LAYOUT", "state_minor") != "fortran"
    if state_minor:
        ctx.set_eigen_device_t(Cm.data_ptr(), dE.data_ptr(), dfe.data_ptr())
    else:
        ctx.set_eigen_device(Cm.data_ptr(), dE.data_ptr(), dfe.data_ptr())

    # N > 1: one NCCL communicator INSIDE the library (xtb_comm_init): rank 0 draws the id, torch.distributed only carries its
    # 128 bytes to the other ranks (what MPI_Bcast does for the Fortran caller).  Every exchange of the step then runs in
    # libxtb_b200.so: H/S block segments all-gathered, rho / rho_E halo only, forces all-gathered.
    if world > 1:
        box = [ctx.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        ctx.comm_init(box[0], rank, world)
    # "replicated" (default; north star: H/S row blocks all-gathered over NVLink to a replicated eigensolve) or
    # "panels" (variant: every rank writes only the dense column panels of its atoms -- 1-D column-block layout)
    replicate_dense = os.environ.get("XTB_BENCH_DENSE_LAYOUT", "replicated") != "panels"
    layout = {"replicated": replicate_dense}
    STEP_ALL = abi.STEP_ALL

    def step():
        if world == 1 or layout["replicated"]:
            # the whole step as ONE library call (xtb_step): list -> H/S blocks [-> all-gather] -> dense H_non, Sij ->
            # rho, rho_E [-> halo exchange -> mirror] -> forces [-> all-gather]; one host synchronisation at its end
            ctx._ck(ctx.lib.xtb_step(ctx.h, STEP_ALL, None))
        else:
            # variant (not the headline): every rank writes only the column panels of its atoms, so that across the job every
            # dense element is written once (the input layout of a multi-GPU eigensolver such as cusolverMgSyevd); no H/S exchange
            ctx.get_near_neighbours_device()
            ctx.assemble_blocks()
            ctx.expand_dense(a0 + 1, a1)
            ctx.density()
            ctx.exchange_blocks(abi.XCHG_RHO)
            ctx.density_mirror()
            ctx.forces_device()
            ctx.exchange_forces()

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    sync_all()
    launches0 = ctx.counter(6)
    sampler = ClockSampler(local) if rank == 0 else None
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    stage = np.zeros(24)
    e0.record(stream)
    for _ in range(args.steps):
        step()
        for s in range(1, 24):
            stage[s] += ctx.stage_ms(s)
    e1.record(stream)
    sync_all()
    ms = e0.elapsed_time(e1) / args.steps
    clocks = sampler.stop() if sampler else None
    launches = (ctx.counter(6) - launches0) // max(args.steps, 1)
    stage /= args.steps
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    n_blocks, n_pairs, k_kept = ctx.counter(4), ctx.counter(5), ctx.counter(7)

    out = None
    if rank == 0:
        peaks = _peaks()
        dgemm_tf, dmma_tf = fp64_peak_tflops(torch, ctx)
        own_frac = (a1 - a0) / na
        flops = 2 * 2 * 81 * k_kept * (n_pairs / 2 + na) * own_frac
        ach = flops / (stage[7] * 1e-3) * 1e-12 if stage[7] > 0 else None
        hbm = peaks["hbm_gbs"] if peaks else 6650.0
        # HBM-bound kernel of the step: the eigenvector re-tiling (reads the kept columns of the atom groups a rank touches,
        # writes them tiled).  Algorithmic bytes at N = 1: 2 * 8 * N * k_kept; partitioned runs re-tile own rows + halo only,
        # so the fraction is reported at N = 1.  (The dense H_non / Sij used to be the HBM stage; they now stay resident and
        # only changed blocks are rewritten -- `dense_update` below; XTB_DENSE_FULL=1 gives the old full write.)
        retile_bytes = 2 * 8 * N * k_kept
        out = {
            "metric": METRIC, "value": ms, "unit": "ms", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            # `config` names the WORKLOAD only (identical in the reference arm); how this arm runs it is in `implementation`
            "config": {"workload": f"{wl['name']}: synthetic Au fcc 8x8x16 supercell ({na} atoms, N={N} orbitals), NRL spd TB",
                       "n_pairs": int(n_pairs), "states_kept": int(k_kept),
                       "eigensolve": "outside the graft; synthetic eigenvectors (N(0,1)/sqrt(N)), E sorted uniform(-10,40) eV, Fermi occupations at 61 % filling"},
            "implementation": {
                       "n_blocks": n_blocks, "l2_policy": "inputs larger than L2 (eigenvectors 10.9 GB per step)",
                       "eigenvectors": "resident in HBM, " + ("state-minor (the layout xtb_eigensolve leaves on the device)" if state_minor else "Fortran layout Ha(N,N)"),
                       "density_reads": "eigenvectors through a tensor map, no re-tiled copy" if ctx.counter(17) == 1 else "re-tiled copy of the eigenvectors",
                       "partition": "atom rows" if world > 1 else "single GPU",
                       "step_call": "xtb_step (one library call per step; steady-state neighbour list without a size read-back at N = 1)",
                       "block_exchange": (None if world == 1 else "inside libxtb_b200.so over NCCL (xtb_comm_init): H/S block segments in-place "
                                          "all-gather; rho/rho_E halo only (ncclSend/ncclRecv of packed blocks); forces all-gather"),
                       "dense_layout": ("single GPU" if world == 1 else "replicated on every rank" if replicate_dense
                                        else "column panels of the owned atoms (each dense element written once in the job)")},
            "stage_ms": {"neighbours": stage[1], "assembly": stage[2], "density": stage[3], "forces": stage[4],
                         "dense_update": stage[6], "density_kernel": stage[7], "retile_kernel": stage[16],
                         "exchange_HS_blocks": stage[14] if world > 1 else None, "exchange_rho_blocks": stage[15] if world > 1 else None},
            "exchange": (None if world == 1 else {
                "HS_blocks": {"ms": stage[14], "bytes_received_per_rank": ctx.counter(abi.CNT_XCHG_HS_BYTES),
                              "GB_per_s": ctx.counter(abi.CNT_XCHG_HS_BYTES) / (stage[14] * 1e-3) * 1e-9 if stage[14] > 0 else None},
                "rho_blocks": {"ms": stage[15], "bytes_received_per_rank": ctx.counter(abi.CNT_XCHG_RHO_BYTES),
                               "GB_per_s": ctx.counter(abi.CNT_XCHG_RHO_BYTES) / (stage[15] * 1e-3) * 1e-9 if stage[15] > 0 else None,
                               "halo_blocks_sent": ctx.counter(abi.CNT_HALO_SEND), "halo_blocks_received": ctx.counter(abi.CNT_HALO_RECV),
                               "all_gather_would_receive_bytes": int(2 * 8 * 81 * n_blocks * (world - 1) / world)}}),
            "pair_blocks_per_s": n_blocks * own_frac / (stage[2] * 1e-3) if stage[2] > 0 else None,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "density_dmma_kernel (rho and rho_E on the neighbour pattern, FP64 DMMA.8x8x4)",
                         "achieved": ach, "peak": dgemm_tf, "unit": "TFLOP/s", "frac": (ach / dgemm_tf) if ach else None,
                         "traffic": _density_traffic() if world == 1 else None, "traffic_unit": "bytes (dram read + write per launch, ncu)",
                         "flops_per_launch": flops,
                         "peak_source": "in-run torch.matmul fp64 8192^3 (cuBLAS DGEMM, burst); MEASURED_PEAKS.json has no FP64 entry",
                         "peak_pipe_dmma884": dmma_tf},
            # the two HBM-side stages of the step: algorithmic bytes / measured stage time
            "roofline_hbm_stage": (lambda by, t: {
                "bound": "hbm", "kernel": "force stage: nrl_onsite_w_kernel + nrl_force_pairs_kernel + nrl_force_sum_kernel",
                "achieved": by / (t * 1e-3) * 1e-9 if t > 0 else None, "peak": hbm, "unit": "GB/s",
                "frac": (by / (t * 1e-3) * 1e-9 / hbm) if t > 0 else None, "ms": t, "bytes": by,
                "bytes_note": "2 * 8 * 81 per block (rho, rho_E read once) + 36 per directed pair (list entry + geometry) + 24 per atom out",
                "fp64_floor_ms": 3.0 * 2 * 1500 * (n_pairs / 2) * own_frac / 32 / (148 * 4 / 2 * 1.965e9) * 1e3,
                "fp64_note": "each unordered pair is differentiated once: 3 coordinates x 2 matrices x ~1500 FP64 instructions per thread "
                             "(radial functions with 11 exp, 45 dual-number SK entries, contraction) at one FP64 warp instruction per 2 cycles "
                             "per SM sub-partition -- the stage is bound by the FP64 pipe, not by HBM",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)"})(
                    (2 * 8 * 81 * n_blocks + 36 * n_pairs + 24 * na) * own_frac, stage[4]),
            "roofline_hbm_assembly": (lambda by, t: {
                "bound": "hbm", "kernel": "assembly stage: nrl_dens_kernel + dense_zero_prev_kernel + dense_remember_kernel + nrl_blocks_kernel "
                                          "(block stores of H and S and the same values into the resident dense H_non, Sij)",
                "achieved": by / (t * 1e-3) * 1e-9 if t > 0 else None, "peak": hbm, "unit": "GB/s",
                "frac": (by / (t * 1e-3) * 1e-9 / hbm) if t > 0 else None, "ms": t, "bytes": by,
                "bytes_note": "2 * 8 * 81 per block into the block stores + the same into the dense matrices (72-byte runs: partial "
                              "sectors, read-modify-write in L2) + 36 per directed pair read"})(
                    (2 * 2 * 8 * 81 * n_blocks + 36 * n_pairs) * own_frac, stage[2]),
            "dense_update": {"ms": stage[6], "ms_is": "preparation only (list comparison + pattern record): the values are written by nrl_blocks_kernel in the same pass as the block stores (XTB_DENSE_FUSED=0: separate scatter kernel)", "bytes_touched": 2 * 8 * 81 * n_blocks,
                             "note": "dense H_non, Sij (2 x 10.9 GB) stay resident in HBM; a step writes the blocks of the new neighbour "
                                     "pattern and clears only the blocks the pattern dropped (none here: same positions every step); "
                                     "every element ends up exactly as after a full write.  bytes_touched = the block writes"},
        }
    # ---- informational variant (not the headline): occupation cut-off f_n <= 1e-30 in the rho / rho_E contraction.
    # Dropping those states changes rho by < 1e-30 * (C C^T)_ii, far below the FP64 rounding of the sum itself, but it is
    # not bit-identical to summing them, so the default (and `value`) keeps every state with f_n > 0.
    if rank == 0 and world == 1:
        try:
            ctx.set_density_cutoff(1e-30)
            for _ in range(2):
                ctx.density()
            out["variant_density_fcut_1e-30"] = {"density_kernel_ms": ctx.stage_ms(7), "density_ms": ctx.stage_ms(3),
                                                 "states_kept": ctx.counter(7), "of_states": N}
        finally:
            ctx.set_density_cutoff(0.0)
            ctx.density()
    # ---- variant: Fortran-layout eigenvectors (what a host eigensolver uploads): the density stage then re-tiles them first
    if rank == 0 and world == 1 and state_minor:
        ctx.set_eigen_device(Cm.data_ptr(), dE.data_ptr(), dfe.data_ptr())
        for _ in range(3):
            ctx.density()
        t3, t7, t16 = ctx.stage_ms(3), ctx.stage_ms(7), ctx.stage_ms(16)
        out["variant_fortran_layout_eigenvectors"] = {
            "density_ms": t3, "density_kernel_ms": t7, "retile_kernel_ms": t16, "ms_per_step_estimate": ms - stage[3] + t3,
            "note": "xtb_set_eigen_device with Ha(N,N) column-major: ct_tile_kernel writes the tiled copy (10.9 GB in, 10.8 GB out) "
                    "before the contraction; same rho / rho_E bits (tests/test_density_tm_gpu.py)"}
        out["variant_fortran_layout_eigenvectors"]["roofline_hbm"] = {
            "bound": "hbm", "kernel": "ct_tile_kernel (Fortran-layout variant only: eigenvector columns re-tiled into TMA / DMMA-fragment order)",
            "achieved": retile_bytes / (t16 * 1e-3) * 1e-9 if t16 > 0 else None, "peak": hbm, "unit": "GB/s",
            "frac": (retile_bytes / (t16 * 1e-3) * 1e-9 / hbm) if t16 > 0 else None, "ms": t16,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)"}
        ctx.set_eigen_device_t(Cm.data_ptr(), dE.data_ptr(), dfe.data_ptr())
        ctx.density()
    # ---- informational variant: the dense H_non / Sij written in full (every element) as up to round 2
    if rank == 0 and world == 1:
        os.environ["XTB_DENSE_FULL"] = "1"
        try:
            for _ in range(2):
                ctx.assemble_HS(to_host=False)
            ms_full = ctx.stage_ms(6)
            out["variant_dense_full_write"] = {"dense_panel_kernel_ms": ms_full, "GB_per_s": 2 * 8 * N * N / (ms_full * 1e-3) * 1e-9,
                                               "frac_of_hbm_peak": 2 * 8 * N * N / (ms_full * 1e-3) * 1e-9 / hbm}
        finally:
            del os.environ["XTB_DENSE_FULL"]
            ctx.assemble_HS(to_host=False)
    # ---- informational variant at N > 1 (not the headline): dense H_non / Sij as column panels of the owned atoms, no
    # H/S exchange (see step()).  Same timing rules: warm-up, K steps between barriers, CUDA events, max over ranks.
    if world > 1 and replicate_dense:
        layout["replicated"] = False
        for _ in range(args.warmup):
            step()
        sync_all()
        v0 = torch.cuda.Event(enable_timing=True); v1 = torch.cuda.Event(enable_timing=True)
        v0.record(stream)
        for _ in range(args.steps):
            step()
        v1.record(stream)
        sync_all()
        t = torch.tensor([v0.elapsed_time(v1) / args.steps], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        layout["replicated"] = True
        if out is not None:
            out["variant_dense_column_panels"] = {
                "ms_per_step": float(t.item()), "dense_update_ms": ctx.stage_ms(6),
                "note": "every rank writes only the dense columns of its atoms (each element once in the job; the 1-D "
                        "column-block input layout of a multi-GPU eigensolver); no H/S block all-gather"}
    # ---- N > 1: the partitioned, exchanged result against the same step computed by rank 0 alone on its GPU ----
    # forces and H/S blocks of ALL atoms (all-gathered); rho / rho_E of rank 0's own rows (the halo exchange delivers a rank
    # what ITS rows need, not the other ranks' rows)
    if world > 1:
        layout["replicated"] = True
        step(); sync_all()
        if rank == 0:
            rp = interop.buffer(ctx, abi.BUF_ROWPTR, torch.int32).cpu().numpy()
            nn_last = int(interop.buffer(ctx, abi.BUF_NNSIZE, torch.int32)[a1 - 1].item())
            own = slice(int(rp[a0]) * 81, (int(rp[a1 - 1]) + 1 + nn_last) * 81)      # blocks of the rows a0 .. a1-1
            bufs = (abi.BUF_FORCES, abi.BUF_RHO, abi.BUF_RHOE, abi.BUF_HBLK, abi.BUF_SBLK)
            sel = lambda b, t: t[own] if b in (abi.BUF_RHO, abi.BUF_RHOE) else (t[:3 * na] if b == abi.BUF_FORCES else t)
            got = [sel(b, interop.buffer(ctx, b)).clone() for b in bufs]
            ctx.set_partition(1, na)
            ctx.get_near_neighbours_device(); ctx.assemble_HS(to_host=False); ctx.density(); ctx.forces_device()
            torch.cuda.synchronize()
            names = ("forces_eV_per_A(frac. coords as returned)", "rho(rows of rank 0)", "rho_E(rows of rank 0)", "H_blocks", "S_blocks")
            ref = [sel(b, interop.buffer(ctx, b)) for b in bufs]
            out["multi_gpu_vs_single_gpu_max_abs_diff"] = {n: float((g - r).abs().max().item()) for n, g, r in zip(names, got, ref)}
            out["multi_gpu_vs_single_gpu_max_abs_diff"]["forces_scale"] = float(ref[0].abs().max().item())
            del got, ref
            ctx.set_partition(a0 + 1, a1)
        sync_all()
    # ---- e2e through the C ABI with host buffers (rank-local share of the work at N > 1) ----
    if not args.no_e2e:
        e2e = run_e2e(torch, ctx, wl, Cm, E, fe, world, rank, a0, a1, args)
        if world > 1:
            t = torch.tensor([e2e["value"]], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e["value"] = float(t.item())
        if out is not None:
            out["e2e"] = e2e
    if rank == 0 and world == 1 and not args.no_configs:
        out["configs"] = small_configs(torch, api, abi, wk)
    if rank == 0 and world == 1 and not args.no_cpu:
        rp = interop.buffer(ctx, abi.BUF_ROWPTR, torch.int32).cpu().numpy()
        chk = (rp, interop.buffer(ctx, abi.BUF_RHO).cpu().numpy(), interop.buffer(ctx, abi.BUF_RHOE).cpu().numpy(),
               interop.buffer(ctx, abi.BUF_FORCES).cpu().numpy().reshape(na, 3))
        out["cpu_baseline"] = cpu_sample(wl, args, check=chk)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))


def small_configs(torch, api, abi, wk, reps=40):
    """BASELINE.json configs 1-4 (parity-test cases, not the headline): ms per step of the same four stages, host clock
    around the calls a caller makes (these sizes are launch- / synchronisation-bound, so the wall time is the honest figure):
      staged   xtb_neighbours, xtb_assemble_HS, xtb_density, xtb_forces one after the other (each waits for the device)
      step     xtb_step, eager (one wait per step, steady-state list)
      graph    xtb_step | XTB_STEP_GRAPH (the step's launch sequence replayed as one CUDA graph)
      graph_with_pair_terms   the same plus XTB_STEP_PAIR_TERMS where the configuration has pair terms (`pair_terms`: the Pettifor
               repulsive part; config 4 also the Girifalco van der Waals term its material folder ships) -- the complete force
    Eigenvectors: the configuration's own, from xtb_eigensolve once (outside the metric); occupations from the host."""
    out = {}
    for name in ("diamond64", "si216", "al500", "ac1000"):
        wl = wk.CONFIGS[name]()
        c = api.TBContext(0)            # own (capturable) stream
        try:
            c.set_workload(wl)
            c.get_near_neighbours_device(); c.assemble_HS(to_host=False)
            Ei = c.eigensolve()[0]
            fe, mu = wk.occupations(Ei, wl["n_electrons"], wl["Te_eV"])
            c.set_occupations(fe)

            def staged():
                c.get_near_neighbours_device(); c.assemble_HS(to_host=False); c.density(); c.forces_device()

            def fused(flags):
                c._ck(c.lib.xtb_step(c.h, flags, None))

            res = {"na": wl["na"], "N": wl["na"] * c.n_orb, "tb": wl["tb"]}
            legs = [("staged", staged), ("step", lambda: fused(abi.STEP_ALL)), ("graph", lambda: fused(abi.STEP_ALL | abi.STEP_GRAPH))]
            res["pair_terms"] = [k for k, on in (("repulsive", bool(wl.get("rep_params"))), ("wall", bool(wl.get("wall_params"))),
                                                 ("vdw_" + str(wl.get("vdw_kind")), bool(wl.get("vdw_params")))) if on]
            if res["pair_terms"]:
                legs.append(("graph_with_pair_terms", lambda: fused(abi.STEP_ALL | abi.STEP_PAIR_TERMS | abi.STEP_GRAPH)))
            for key, fn in legs:
                for _ in range(4):
                    fn()
                c.sync()
                l0 = c.counter(6)
                t0 = time.perf_counter()
                for _ in range(reps):
                    fn()
                c.sync()
                res[key + "_ms"] = (time.perf_counter() - t0) * 1e3 / reps
                res[key + "_launches"] = (c.counter(6) - l0) / reps
            res["graph_device_ms"] = c.stage_ms(0)
            if wl.get("vdw_params"):
                res["vdw_ms"] = c.stage_ms(17) if c.stage_ms(17) > 0 else None      # eager legs only: no events inside a graph
            res["graphs_captured"] = c.counter(16)
            try:      # BASELINE.json's second metric at this atom count: pair-blocks/s of the assembly stage (one staged pass, device-timed)
                staged(); c.sync()
                t2 = c.stage_ms(2)
                res["n_blocks"] = int(c.counter(4)); res["assembly_ms"] = t2
                res["pair_blocks_per_s"] = c.counter(4) / (t2 * 1e-3) if t2 > 0 else None
            except Exception:
                pass
            out[name] = res
        except Exception as ex:      # a parity-test case must not take the headline line down
            out[name] = {"error": str(ex)[:200]}
        finally:
            c.close()
    return out


def run_e2e(torch, ctx, wl, Cm, E, fe, world, rank, a0, a1, args):
    """The step through the reference-facing C ABI with HOST buffers, copies inside the timed region, timed with the
    host clock around the calls a Fortran caller would make.  Two flavours:
      e2e (headline)         the north-star flow: the eigensolve runs on the device (xtb_eigensolve, cuSOLVER), so only
                             MDAtoms%S goes in and Ei comes out / fe goes in / forces come out; H_non, Sij and Ha never
                             cross PCIe.  The eigensolve itself is outside the metric and is not executed here (the
                             eigenvectors are the synthetic resident ones; no solver can diagonalise N = 36864 inside a bench run).
      host_eigensolver       the reference's own flow kept verbatim: H_non + Sij out to host LAPACK, Ha + Ei + fe back in.
    """
    api = importlib.import_module("xtant-3_b200.api")
    abi = importlib.import_module("xtant-3_b200._abi")
    na, N = wl["na"], wl["na"] * 9
    if world > 1:
        return run_e2e_multi(torch, ctx, wl, E, fe, world, rank, a0, a1, args)
    Eh = np.ascontiguousarray(E); fh = np.ascontiguousarray(fe)
    Ei_host = np.zeros(N)
    Fh = np.zeros((na, 3))
    koa = np.ascontiguousarray(wl["koa"], dtype=np.int32)
    S = np.ascontiguousarray(wl["S"]); h = np.asfortranarray(wl["supce"])
    lib, hnd = ctx.lib, ctx.h
    n = max(1, args.e2e_steps)

    io = abi.StepIO()
    io.S = api._dp(S); io.fe = api._dp(fh); io.Ei = api._dp(Ei_host); io.F_att = api._dp(Fh)

    def step_resident():
        # ONE call: H2D of S and fe, the whole device step, D2H of Ei and the forces (occupations are a host job in the
        # reference's flow: Ei out, fe in -- here both travel with the same call since the eigen-solution is resident)
        ctx._ck(lib.xtb_step(hnd, abi.STEP_ALL, C.byref(io)))

    for _ in range(2):
        step_resident()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(max(n, 3)):
        step_resident()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / max(n, 3)
    out = {"value": ms, "unit": "ms", "h2d_bytes_per_step": 8 * (3 * na + 9 + N) + 4 * na, "d2h_bytes_per_step": 8 * (N + 3 * na),
           "steps": max(n, 3), "flow": "device-resident eigenvectors (xtb_eigensolve flow; the eigensolve is outside the metric)"}

    # ---- the reference's own flow: dense matrices through host memory ----
    try:
        Hh = torch.empty(N * N, dtype=torch.float64, pin_memory=True)
        Sh = torch.empty(N * N, dtype=torch.float64, pin_memory=True)
        Ch = torch.empty(N * N, dtype=torch.float64, pin_memory=True)
        Ch.copy_(Cm.view(-1))
    except Exception as ex:      # host memory
        out["host_eigensolver"] = {"value": None, "unit": "ms", "error": f"pinned host buffers: {ex}"}
        return out
    dp = lambda t: C.cast(t.data_ptr(), C.POINTER(C.c_double))

    def step_host():
        ctx._ck(lib.xtb_set_system(hnd, na, api._ip(koa), api._dp(S), api._dp(h)))
        ctx._ck(lib.xtb_neighbours(hnd, 0.0, None, None, None, None))
        ctx._ck(lib.xtb_assemble_HS(hnd, dp(Hh), dp(Sh)))
        ctx._ck(lib.xtb_set_eigen(hnd, dp(Ch), api._dp(Eh), api._dp(fh)))
        ctx._ck(lib.xtb_density(hnd, None, None))
        ctx._ck(lib.xtb_forces(hnd, api._dp(Fh)))

    step_host()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        step_host()
    torch.cuda.synchronize()
    ms_h = (time.perf_counter() - t0) * 1e3 / n
    out["host_eigensolver"] = {"value": ms_h, "unit": "ms", "h2d_bytes_per_step": 8 * (3 * na + N * N + 2 * N) + 4 * na,
                               "d2h_bytes_per_step": int(8 * (2 * N * N + 3 * na)), "steps": n,
                               "note": "reference flow: H_non, Sij to host LAPACK and Ha, Ei, fe back (32.6 GB of PCIe per step); "
                                       "the LAPACK time itself is outside the metric"}
    return out


def run_e2e_multi(torch, ctx, wl, E, fe, world, rank, a0, a1, args):
    """e2e at N > 1: every rank drives its context through the C ABI with HOST buffers exactly as the reference's MPI
    build calls the path (every rank holds the replicated host arrays): S in, owned H/S blocks -> NVLink all-gather ->
    dense H_non / Sij expanded on every GPU (the eigensolve boundary), Ei out, fe in, owner rho blocks -> all-gather ->
    mirror, owned forces out to the host and all-gathered.  Host clock between barriers, max over ranks (the caller
    reduces)."""
    import torch.distributed as dist
    api = importlib.import_module("xtant-3_b200.api")
    abi = importlib.import_module("xtant-3_b200._abi")
    interop = importlib.import_module("xtant-3_b200.interop")
    mg = importlib.import_module("xtant-3_b200.multigpu")
    na, N = wl["na"], wl["na"] * 9
    fh = np.ascontiguousarray(fe)
    Ei_host = np.zeros(N); Fh = np.zeros((na, 3))
    koa = np.ascontiguousarray(wl["koa"], dtype=np.int32)
    S = np.ascontiguousarray(wl["S"]); h = np.asfortranarray(wl["supce"])
    lib, hnd = ctx.lib, ctx.h
    n = max(args.e2e_steps, 3)
    replicate = os.environ.get("XTB_BENCH_DENSE_LAYOUT", "replicated") != "panels"

    io = abi.StepIO()
    io.S = api._dp(S); io.fe = api._dp(fh); io.Ei = api._dp(Ei_host); io.F_att = api._dp(Fh)

    def step():
        if replicate:
            ctx._ck(lib.xtb_step(hnd, abi.STEP_ALL, C.byref(io)))      # S, fe in; Ei, ALL forces out; exchanges inside the library
        else:
            ctx._ck(lib.xtb_set_system(hnd, na, api._ip(koa), api._dp(S), api._dp(h)))
            ctx._ck(lib.xtb_neighbours(hnd, 0.0, None, None, None, None))
            ctx._ck(lib.xtb_assemble_blocks(hnd))
            ctx._ck(lib.xtb_expand_dense(hnd, a0 + 1, a1, None, None))
            ctx._ck(lib.xtb_get_eigenvalues(hnd, api._dp(Ei_host)))
            ctx._ck(lib.xtb_set_occupations(hnd, api._dp(fh)))
            ctx._ck(lib.xtb_density(hnd, None, None))
            ctx._ck(lib.xtb_exchange_blocks(hnd, abi.XCHG_RHO))
            ctx._ck(lib.xtb_density_mirror(hnd))
            ctx._ck(lib.xtb_forces(hnd, None))
            ctx._ck(lib.xtb_exchange_forces(hnd, api._dp(Fh)))

    for _ in range(2):
        step()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / n
    return {"value": ms, "unit": "ms", "h2d_bytes_per_step": world * (8 * (3 * na + 9 + N) + 4 * na),
            "d2h_bytes_per_step": world * (8 * N + 8 * 3 * na), "steps": n,
            "flow": "per-rank xtb_step with host buffers (S, fe in; Ei, all forces out on every rank, as the MPI build replicates them); "
                    + ("H/S blocks all-gathered, " if replicate else "dense H_non / Sij as owned column panels; ")
                    + "rho blocks (halo only) and forces exchanged over NVLink inside the library; dense H_non / Sij and the eigenvectors never cross PCIe"}


def cpu_sample(wl, args, check=None, n_dens=256, threads=None):
    """oracle (CPU restatement of the reference algorithm, sparse-equivalent form) on the SAME workload, host cores only (no
    GPU anywhere in this function).  The cheap stages run IN FULL: neighbour list, radial tables, H/S assembly of every
    column panel (in slabs of 128 atoms, each slab's dense panels written and dropped) and the forces of every atom.  Only
    the rho / rho_E contraction is sampled: the blocks of the first `n_dens` atoms (all their neighbours' eigenvector rows
    gathered), scaled by na / n_dens -- every atom of the cell has the same number of blocks within 3 %.
    Eigenvector rows / rho blocks are synthetic of the same shape and distribution (timing does not depend on the values).
    check = (rowptr, rho, rhoE, F) numpy copies of a device step: the forces then run on the device's own rho blocks
    and double as a full-size parity check.  threads: OpenMP threads to pin (default: all host cores)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc
    wk = importlib.import_module("xtant-3_b200.workloads")
    cores = orc.set_num_threads(threads or os.cpu_count() or 1)
    na, N = wl["na"], wl["na"] * 9
    n_dens = min(n_dens, na)
    t = {}
    t0 = time.perf_counter(); nb = orc.neighbours(wl, cap=128); t["neighbours"] = time.perf_counter() - t0
    t0 = time.perf_counter(); lmn, xrr, x1 = orc.geometry(wl, nb); rad = orc.nrl_radial(wl, nb); t["radial"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    slab = 128
    for j0 in range(0, na, slab):
        orc.nrl_assemble_panels(wl, nb, lmn, rad, j0, min(j0 + slab, na))
    t["assembly"] = time.perf_counter() - t0
    # eigenvector rows of the sampled atoms and their neighbours
    rng = np.random.default_rng(7)
    E = np.sort(rng.uniform(-10.0, 40.0, size=N))
    fe, mu = wk.occupations(E, 2.0 * 0.61 * N, wl["Te_eV"])
    keep = fe > 0
    kk = int(keep.sum())
    atoms = sorted(set(range(n_dens)) | {int(j) for i in range(n_dens) for j in nb.list[i, :nb.size[i]]})
    rowmap = -np.ones(na, dtype=np.int32); rowmap[atoms] = np.arange(len(atoms), dtype=np.int32)
    Ct = rng.standard_normal((len(atoms) * 9, kk)) / np.sqrt(N)
    t0 = time.perf_counter()
    orc.density_blocks_rows(nb, 9, Ct, rowmap, fe[keep], E[keep], 0, n_dens)
    dens_sample = time.perf_counter() - t0
    t["density"] = dens_sample * na / n_dens
    del Ct
    # what the reference itself does for rho and rho_E: two dense dgemm('N','T') of order N (TB.f90:4386,4465);
    # sampled as a 256-row panel with numpy's BLAS, scaled to N rows.  Reported beside, not inside, `value`.
    m = 256
    A = rng.standard_normal((m, kk)); B = rng.standard_normal((2304, kk))
    try:                                   # numpy's BLAS took its thread count from the environment at import: pin it as well
        from threadpoolctl import threadpool_limits
        with threadpool_limits(limits=cores):
            (A[:8] @ B.T)
            t0 = time.perf_counter(); (A @ B.T); td = time.perf_counter() - t0
    except ImportError:
        t0 = time.perf_counter(); (A @ B.T); td = time.perf_counter() - t0
    dense_ms = 1e3 * 2 * td * (N / m) * (N / 2304)
    del A, B
    if check is not None:
        rp, rho_d, rhoE_d, Fg = check
        rho = np.zeros((na, nb.cap + 1, 81)); rhoE = np.zeros_like(rho)
        for i in range(na):
            mm = nb.size[i] + 1
            rho[i, :mm] = rho_d[rp[i] * 81:(rp[i] + mm) * 81].reshape(mm, 81)
            rhoE[i, :mm] = rhoE_d[rp[i] * 81:(rp[i] + mm) * 81].reshape(mm, 81)
    else:
        rho = rng.standard_normal((na, nb.cap + 1, 81)) * 0.05; rhoE = rng.standard_normal((na, nb.cap + 1, 81)) * 0.05
    t0 = time.perf_counter()
    Fo = orc.nrl_forces_sparse(wl, nb, lmn, rad, rho, rhoE)
    t["forces"] = time.perf_counter() - t0
    total_ms = 1e3 * sum(t.values())
    measured_ms = 1e3 * (t["neighbours"] + t["radial"] + t["assembly"] + t["forces"] + dens_sample)
    out = {"value": total_ms, "unit": "ms", "cores": cores, "kind": "port", "sampled": True, "extrapolated": True,
           "n_pairs": int(nb.size.sum()), "states_kept": kk,
           "scale_factors": {"neighbours": 1.0, "radial": 1.0, "assembly": 1.0, "forces": 1.0, "density": na / n_dens},
           "measured_ms_before_scaling": measured_ms,
           "sample": f"neighbour list, radial tables, H/S assembly (all {na} column panels) and forces of all {na} atoms in full; "
                     f"rho/rho_E blocks of {n_dens} atoms scaled by {na / n_dens:g}; sparse-equivalent form of the "
                     "reference algorithm (its verbatim force loop needs 2 x 32.6 GB per thread at this size)",
           "stage_ms": {k: 1e3 * v for k, v in t.items()},
           "variant_reference_dense_dgemm": {"density_ms": dense_ms, "value": total_ms - 1e3 * t["density"] + dense_ms,
                                             "note": "rho and rho_E as the reference forms them: two dense dgemm('N','T') of order N "
                                                     "(TB.f90:4386,4465), numpy BLAS on a 256-row panel scaled to N rows; the other stages as above"}}
    if check is not None:
        out["full_size_force_check_eV_per_A"] = float(np.abs(orc.forces_cartesian(Fg, wl["supce"]) -
                                                            orc.forces_cartesian(Fo, wl["supce"])).max())
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The reference is Fortran and no Fortran
    compiler exists on either machine, so this is the oracle port (kind 'port') on ALL host cores -- the thread count is
    pinned explicitly (torchrun exports OMP_NUM_THREADS=1) and reported; each step = cpu_sample (cheap stages in full, the
    rho contraction sampled on 256 atoms and scaled: `sampled` / `extrapolated` / `scale_factors` in the line).  Under
    torchrun rank 0 alone runs it.  No GPU, none of the library's kernels."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wk = importlib.import_module("xtant-3_b200.workloads")
    wl = wk.CONFIGS[args.workload]()
    N = wl["na"] * 9
    vals = []
    for s in range(args.warmup + args.steps):
        cb = cpu_sample(wl, args)
        if s >= args.warmup:
            vals.append(cb)
    ms = float(np.mean([v["value"] for v in vals]))
    cb = vals[-1]; cb["value"] = ms
    out = {"impl": "reference", "metric": METRIC, "value": ms, "unit": "ms", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic", "sampled": True, "extrapolated": True, "cores": cb["cores"],
           "config": {"workload": f"{wl['name']}: synthetic Au fcc 8x8x16 supercell ({wl['na']} atoms, N={N} orbitals), NRL spd TB",
                      "n_pairs": int(cb["n_pairs"]), "states_kept": int(cb["states_kept"]),
                      "eigensolve": "outside the graft; synthetic eigenvectors (N(0,1)/sqrt(N)), E sorted uniform(-10,40) eV, Fermi occupations at 61 % filling"},
           "cpu_baseline": cb,
           "e2e": {"value": ms, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--workload", default="au4096")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
