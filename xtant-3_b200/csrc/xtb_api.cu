// C ABI of libxtb_b200 (include/xtb_b200.h): argument checking, host<->device staging in the
// reference's array shapes, and dispatch on the TB parametrisation (the `select type` arms of
// Source_files/TB.f90:247-317,459-487 collapse to ctx->tb).  All arithmetic is in the stage files.
#include "xtb_ctx.h"
#include "dftb_device.cuh"
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <new>

int xtb_reserve(xtb_ctx* c, DevBuf& b, size_t bytes) {
    if (bytes == 0) bytes = 16;
    if (b.bytes >= bytes) return XTB_OK;
    if (b.p) cudaFree(b.p);
    b.p = nullptr; b.bytes = 0;
    cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        XTB_FAIL(c, XTB_ENOMEM, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
    }
    b.bytes = bytes;
    c->alloc_epoch++;        // a captured graph holds the old address: it must not be replayed (step.cu, GraphKey)
    // zero once at (re)allocation: padding that no kernel addresses (ELL slack, the gaps of xtb_set_block_layout) then holds
    // zeros instead of stale memory, so whole-buffer copies and exchanges are deterministic
    cudaMemsetAsync(b.p, 0, bytes, c->stream);
    return XTB_OK;
}
void xtb_free(DevBuf& b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr; b.bytes = 0;
}

int xtb_t0(xtb_ctx* c, int s) {
    if (c->capturing) return XTB_OK;
    for (int k = 0; k < 2; ++k)
        if (!c->tev[s][k]) XTB_CUDA(c, cudaEventCreate(&c->tev[s][k]));
    XTB_CUDA(c, cudaEventRecord(c->tev[s][0], c->stream));
    c->t_started |= 1u << s;
    return XTB_OK;
}
int xtb_t1(xtb_ctx* c, int s) {
    if (c->capturing || !(c->t_started & (1u << s))) return XTB_OK;       // an end without a start is not a measurement
    XTB_CUDA(c, cudaEventRecord(c->tev[s][1], c->stream));
    c->t_pending |= 1u << s;
    return XTB_OK;
}
int xtb_collect_timing(xtb_ctx* c) {
    for (int s = 0; s < 24; ++s)
        if (c->t_pending & (1u << s)) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, c->tev[s][0], c->tev[s][1]) == cudaSuccess) c->stage_ms[s] = ms;
            else cudaGetLastError();
        }
    c->t_pending = 0; c->t_started = 0;
    return XTB_OK;
}
int xtb_stage_finish(xtb_ctx* c) {
    if (c->deferred || c->capturing) return XTB_OK;
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return xtb_collect_timing(c);
}

// the `select type` arms of TB.f90:247-317,459-487
int xtb_impl_assemble(xtb_ctx* c, bool dense) {
    if (c->tb == XTB_TB_NRL) return xtb_impl_nrl_assemble(c, dense);
    if (c->tb == XTB_TB_PETTIFOR) return xtb_impl_pettifor_assemble(c, dense);
    if (c->tb == XTB_TB_DFTB) return xtb_impl_dftb_assemble(c, dense);
    XTB_FAIL(c, XTB_EINVAL, "assembly: no TB parametrisation set");
}
int xtb_impl_forces(xtb_ctx* c) {
    if (c->tb == XTB_TB_NRL) return xtb_impl_nrl_forces(c);
    if (c->tb == XTB_TB_PETTIFOR) return xtb_impl_pettifor_forces(c);
    if (c->tb == XTB_TB_DFTB) return xtb_impl_dftb_forces(c);
    XTB_FAIL(c, XTB_EINVAL, "forces: no TB parametrisation set");
}
int xtb_impl_pressure(xtb_ctx* c) {
    if (c->tb == XTB_TB_NRL) return xtb_impl_nrl_pressure(c);
    if (c->tb == XTB_TB_PETTIFOR) return xtb_impl_pettifor_pressure(c);
    if (c->tb == XTB_TB_DFTB) return xtb_impl_dftb_pressure(c);
    XTB_FAIL(c, XTB_EINVAL, "pressure: no TB parametrisation set");
}

// pair terms of a fused step: Pettifor-type repulsive part, short-range wall, van der Waals term, whichever the context holds
int xtb_impl_pair_terms(xtb_ctx* c, const xtb_step_io* io, bool pressure) {
    const bool list_terms = (c->tb == XTB_TB_PETTIFOR && !c->h_pet_rep.empty()) || xtb_has_wall(c);
    if (list_terms && !c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "pair terms: no neighbour list");
    int rc;
    const size_t fb = sizeof(double) * 3 * c->na;
    const bool multi = c->comm && c->nranks > 1;
    if (c->tb == XTB_TB_PETTIFOR && !c->h_pet_rep.empty()) {
        if ((rc = xtb_impl_pettifor_rep_energy(c)) || (rc = xtb_impl_pettifor_rep_forces(c))) return rc;
        if (multi && (rc = xtb_impl_exchange_forces(c, c->d_Frep))) return rc;
        if (io && io->E_rep) XTB_CUDA(c, cudaMemcpyAsync(io->E_rep, c->d_Erep.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        if (io && io->F_rep) XTB_CUDA(c, cudaMemcpyAsync(io->F_rep, c->d_Frep.p, fb, cudaMemcpyDeviceToHost, c->stream));
        if (pressure) {
            if ((rc = xtb_impl_pettifor_rep_pressure(c))) return rc;
            if (multi && (rc = xtb_impl_allreduce9(c, (double*)c->d_SCrep.p))) return rc;
            if (io && io->SCforce_rep) XTB_CUDA(c, cudaMemcpyAsync(io->SCforce_rep, c->d_SCrep.p, 72, cudaMemcpyDeviceToHost, c->stream));
        }
    }
    if (xtb_has_wall(c)) {
        if ((rc = xtb_impl_wall_energy(c)) || (rc = xtb_impl_wall_forces(c))) return rc;
        if (multi && (rc = xtb_impl_exchange_forces(c, c->d_Fwall))) return rc;
        if (io && io->E_wall) XTB_CUDA(c, cudaMemcpyAsync(io->E_wall, c->d_Ewall.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        if (io && io->F_wall) XTB_CUDA(c, cudaMemcpyAsync(io->F_wall, c->d_Fwall.p, fb, cudaMemcpyDeviceToHost, c->stream));
        if (pressure) {
            if ((rc = xtb_impl_wall_pressure(c))) return rc;
            if (multi && (rc = xtb_impl_allreduce9(c, (double*)c->d_SCwall.p))) return rc;
            if (io && io->SCforce_wall) XTB_CUDA(c, cudaMemcpyAsync(io->SCforce_wall, c->d_SCwall.p, 72, cudaMemcpyDeviceToHost, c->stream));
        }
    }
    if (!c->h_vdw.empty()) {
        if ((rc = xtb_impl_vdw(c, pressure ? 7u : 3u))) return rc;          // energy, forces [, supercell force] in one pass
        if (multi && (rc = xtb_impl_exchange_forces(c, c->d_Fvdw))) return rc;
        if (io && io->E_vdw) XTB_CUDA(c, cudaMemcpyAsync(io->E_vdw, c->d_Evdw.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        if (io && io->F_vdw) XTB_CUDA(c, cudaMemcpyAsync(io->F_vdw, c->d_Fvdw.p, fb, cudaMemcpyDeviceToHost, c->stream));
        if (pressure) {
            if (multi && (rc = xtb_impl_allreduce9(c, (double*)c->d_SCvdw.p))) return rc;
            if (io && io->SCforce_vdw) XTB_CUDA(c, cudaMemcpyAsync(io->SCforce_vdw, c->d_SCvdw.p, 72, cudaMemcpyDeviceToHost, c->stream));
        }
    }
    if (!c->h_coul.empty()) {
        if ((rc = xtb_impl_coulomb(c, pressure ? 7u : 3u))) return rc;
        if (multi && (rc = xtb_impl_exchange_forces(c, c->d_Fcoul))) return rc;
        if (io && io->E_coul) XTB_CUDA(c, cudaMemcpyAsync(io->E_coul, c->d_Ecoul.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        if (io && io->F_coul) XTB_CUDA(c, cudaMemcpyAsync(io->F_coul, c->d_Fcoul.p, fb, cudaMemcpyDeviceToHost, c->stream));
        if (pressure) {
            if (multi && (rc = xtb_impl_allreduce9(c, (double*)c->d_SCcoul.p))) return rc;
            if (io && io->SCforce_coul) XTB_CUDA(c, cudaMemcpyAsync(io->SCforce_coul, c->d_SCcoul.p, 72, cudaMemcpyDeviceToHost, c->stream));
        }
    }
    return XTB_OK;
}

extern "C" {

const char* xtb_version(void) { return "xtb_b200 0.1 (sm_100a, FP64)"; }

int xtb_create(xtb_ctx** out, int device) {
    if (!out) return XTB_EINVAL;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { cudaGetLastError(); return XTB_ENODEV; }
    if (device < 0 || device >= n) return XTB_ENODEV;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return XTB_ENODEV; }
    xtb_ctx* c = new (std::nothrow) xtb_ctx();
    if (!c) return XTB_ENOMEM;
    c->device = device;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return XTB_ECUDA; }
    c->own_stream = true;
    if (cudaMallocHost((void**)&c->h_meta, 64) != cudaSuccess) { cudaStreamDestroy(c->stream); delete c; return XTB_ENOMEM; }
    memset(c->h_meta, 0, 64);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
    const char* impl = getenv("XTB_DENSITY_IMPL");
    if (impl) c->density_impl = atoi(impl);
    const char* ncw = getenv("XTB_DENSITY_NCW");
    if (ncw) c->density_ncw = atoi(ncw);
    const char* lead = getenv("XTB_DENSITY_LEAD");
    if (lead) c->density_lead = atoi(lead);
    const char* pairing = getenv("XTB_DENSITY_PAIRING");
    if (pairing) c->density_pairing = atoi(pairing);
    const char* dbg = getenv("XTB_DENSITY_DBG");
    if (dbg) c->density_dbg = atoi(dbg);
    *out = c;
    return XTB_OK;
}

int xtb_destroy(xtb_ctx* c) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->comm) xtb_comm_destroy(c);
    xtb_release_lib_handles(c);
    DevBuf* all[] = {&c->d_params, &c->d_koa, &c->d_S, &c->d_nn_size, &c->d_nn_list, &c->d_nn_dist, &c->d_nn_dist_s, &c->d_rev, &c->d_own,
                     &c->d_rowptr, &c->d_maxnn, &c->d_dens, &c->d_onsite_w, &c->d_Hblk, &c->d_Sblk, &c->d_H, &c->d_S2, &c->own_Ha,
                     &c->own_Ei, &c->own_fe, &c->own_HaT, &c->d_tmap, &c->d_keep, &c->d_wf, &c->d_wfe, &c->d_Ct, &c->d_nkeep, &c->d_rho, &c->d_rhoE, &c->d_F,
                     &c->d_SC, &c->d_scratch, &c->d_hdev, &c->d_ppart, &c->d_fpair, &c->d_tab, &c->d_dftb_tab, &c->d_prog, &c->d_eig_work, &c->d_eig_X, &c->d_eig_T, &c->d_eigS,
                     &c->d_ctneed, &c->d_units, &c->d_pairw, &c->d_Ha0, &c->d_cpl_T, &c->d_cpl_M, &c->d_cpl_inv, &c->d_cpl_union, &c->d_rep_params, &c->d_Epot, &c->d_dfpsi, &c->d_Erep, &c->d_Frep, &c->d_SCrep,
                     &c->d_halo_tr, &c->d_halo_cnt, &c->d_halo_slots, &c->d_halo_send, &c->d_halo_recv, &c->d_wall_params, &c->d_Ewall_at, &c->d_Ewall,
                     &c->d_Fwall, &c->d_SCwall, &c->d_prev_size, &c->d_prev_list, &c->d_dense_out, &c->d_fermi, &c->d_S0, &c->d_SV, &c->d_SV0, &c->d_mdpar,
                     &c->d_vdw_params, &c->d_Evdw_at, &c->d_Evdw, &c->d_Fvdw, &c->d_SCvdw,
                     &c->d_coul_params, &c->d_Ecoul_at, &c->d_Ecoul, &c->d_Fcoul, &c->d_SCcoul, &c->d_short_params, &c->d_short_tab};
    for (DevBuf* b : all) xtb_free(*b);
    for (int s = 0; s < 24; ++s) for (int k = 0; k < 2; ++k) if (c->tev[s][k]) cudaEventDestroy(c->tev[s][k]);
    if (c->h_meta) cudaFreeHost(c->h_meta);
    if (c->graph_exec) cudaGraphExecDestroy((cudaGraphExec_t)c->graph_exec);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
    return XTB_OK;
}

const char* xtb_last_error(const xtb_ctx* c) { return c ? c->err.c_str() : "null context"; }

int xtb_set_stream(xtb_ctx* c, void* s) {
    if (!c) return XTB_EINVAL;
    if (c->own_stream) { cudaStreamSynchronize(c->stream); cudaStreamDestroy(c->stream); c->own_stream = false; }
    c->stream = (cudaStream_t)s;
    return XTB_OK;
}

int xtb_sync(xtb_ctx* c) {
    if (!c) return XTB_EINVAL;
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

static int upload_params(xtb_ctx* c, const void* p, size_t bytes) {
    int rc = xtb_reserve(c, c->d_params, bytes);
    if (rc) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_params.p, p, bytes, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_neigh = c->have_HS = c->have_blocks = c->have_density = c->have_dens = c->have_shape = false;
    c->rm = 0.0;
    return XTB_OK;
}

int xtb_set_params_nrl(xtb_ctx* c, int n_kinds, const xtb_nrl_pair* p) {
    if (!c) return XTB_EINVAL;
    if (n_kinds <= 0 || !p) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_nrl: n_kinds=%d", n_kinds);
    cudaSetDevice(c->device);
    c->h_nrl.assign(p, p + (size_t)n_kinds * n_kinds);
    c->tb = XTB_TB_NRL; c->n_kinds = n_kinds; c->n_orb = 9;      // NRL is always sp3d5 (TB_NRL.f90:255)
    return upload_params(c, p, sizeof(xtb_nrl_pair) * (size_t)n_kinds * n_kinds);
}

int xtb_set_params_pettifor(xtb_ctx* c, int n_kinds, const xtb_pettifor_pair* p) {
    if (!c) return XTB_EINVAL;
    if (n_kinds <= 0 || !p) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_pettifor: n_kinds=%d", n_kinds);
    cudaSetDevice(c->device);
    c->h_pet.assign(p, p + (size_t)n_kinds * n_kinds);
    c->h_pet_rep.clear();                                      // a new attractive set drops the repulsive one
    c->tb = XTB_TB_PETTIFOR; c->n_kinds = n_kinds; c->n_orb = 4;
    return upload_params(c, p, sizeof(xtb_pettifor_pair) * (size_t)n_kinds * n_kinds);
}

int xtb_set_params_pettifor_rep(xtb_ctx* c, int n_kinds, const xtb_pettifor_rep_pair* p) {
    if (!c) return XTB_EINVAL;
    if (c->tb != XTB_TB_PETTIFOR) XTB_FAIL(c, XTB_ESTATE, "xtb_set_params_pettifor_rep: call xtb_set_params_pettifor first");
    if (n_kinds != c->n_kinds || !p) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_pettifor_rep: n_kinds=%d (attractive part has %d)", n_kinds, c->n_kinds);
    cudaSetDevice(c->device);
    const size_t cnt = (size_t)n_kinds * n_kinds;
    c->h_pet_rep.assign(p, p + cnt);
    int rc = xtb_reserve(c, c->d_rep_params, sizeof(xtb_pettifor_rep_pair) * cnt);
    if (rc) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_rep_params.p, p, sizeof(xtb_pettifor_rep_pair) * cnt, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_neigh = c->have_HS = c->have_blocks = c->have_density = c->have_dens = c->have_shape = false;   // the cut-off may have grown
    c->rm = 0.0;
    return XTB_OK;
}

int xtb_set_params_dftb(xtb_ctx* c, int n_kinds, int n_orb, const xtb_dftb_pair* p) {
    if (!c) return XTB_EINVAL;
    if (n_kinds <= 0 || !p) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_dftb: n_kinds=%d", n_kinds);
    if (n_orb != 1 && n_orb != 4 && n_orb != 9) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_dftb: n_orb=%d (1 = s, 4 = sp3, 9 = sp3d5)", n_orb);
    cudaSetDevice(c->device);
    // flatten the tables: one buffer of doubles, per pair Rr[n_r], Vr[n_r*10], Sr[n_r*10]
    const int np = n_kinds * n_kinds;
    std::vector<DftbDev> dev(np);
    std::vector<double> tab;
    double rcmax = 0.0, dmax = 0.0;
    for (int q = 0; q < np; ++q) {
        const xtb_dftb_pair& s = p[q];
        if (s.n_r < 2 || !s.Rr || !s.Vr || !s.Sr) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_dftb: pair %d has no tables (n_r=%d)", q, s.n_r);
        if (!(s.d > 0.0)) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_dftb: pair %d has smoothing width d=%g", q, s.d);
        DftbDev& d = dev[q];
        d.rcut = s.rcut; d.d = s.d; d.Ed = s.Ed; d.Ep = s.Ep; d.Es = s.Es; d.n_r = s.n_r; d.pad_ = 0;
        d.off_R = (long long)tab.size(); tab.insert(tab.end(), s.Rr, s.Rr + s.n_r);
        d.off_V = (long long)tab.size(); tab.insert(tab.end(), s.Vr, s.Vr + (size_t)s.n_r * 10);
        d.off_S = (long long)tab.size(); tab.insert(tab.end(), s.Sr, s.Sr + (size_t)s.n_r * 10);
        rcmax = s.rcut > rcmax ? s.rcut : rcmax; dmax = s.d > dmax ? s.d : dmax;
    }
    int rc;
    if ((rc = xtb_reserve(c, c->d_dftb_tab, sizeof(double) * tab.size()))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_dftb_tab.p, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice, c->stream));
    c->tb = XTB_TB_DFTB; c->n_kinds = n_kinds; c->n_orb = n_orb;
    c->dftb_rm = rcmax + 10.0 * dmax;
    return upload_params(c, dev.data(), sizeof(DftbDev) * (size_t)np);
}

int xtb_set_system(xtb_ctx* c, int na, const int32_t* koa, const double* S, const double* supce) {
    if (!c) return XTB_EINVAL;
    if (na <= 0 || !koa || !S || !supce) XTB_FAIL(c, XTB_EINVAL, "xtb_set_system: bad arguments (na=%d)", na);
    if (c->tb == XTB_TB_NONE) XTB_FAIL(c, XTB_ESTATE, "xtb_set_system: set the TB parameters first");
    cudaSetDevice(c->device);
    std::vector<int32_t> k0(na);
    for (int i = 0; i < na; ++i) {
        if (koa[i] < 1 || koa[i] > c->n_kinds) XTB_FAIL(c, XTB_EINVAL, "atom %d has kind %d outside 1..%d", i + 1, koa[i], c->n_kinds);
        k0[i] = koa[i] - 1;
    }
    int rc;
    if ((rc = xtb_reserve(c, c->d_koa, sizeof(int32_t) * na))) return rc;
    if ((rc = xtb_reserve(c, c->d_S, sizeof(double) * 3 * na))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_koa.p, k0.data(), sizeof(int32_t) * na, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaMemcpyAsync(c->d_S.p, S, sizeof(double) * 3 * na, cudaMemcpyHostToDevice, c->stream));
    // Scell%supce lives on the device from here on (every stage reads it there; no per-stage upload)
    if ((rc = xtb_reserve(c, c->d_hdev, 256))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_hdev.p, supce, sizeof(double) * 9, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    memcpy(c->h_supce, supce, sizeof(double) * 9);
    if (na != c->na) {
        c->cap = 0; c->a0 = 0; c->a1 = na; c->have_shape = false; c->have_md = false;
        if (c->comm && c->nranks > 1) {         // the owned range follows the rank
            const int per = (na + c->nranks - 1) / c->nranks;
            c->a0 = c->rank * per < na ? c->rank * per : na;
            c->a1 = c->a0 + per < na ? c->a0 + per : na;
        }
    }
    c->na = na;
    c->have_neigh = c->have_HS = c->have_blocks = c->have_density = c->have_dens = false;
    return XTB_OK;
}

int xtb_set_partition(xtb_ctx* c, int first, int last) {
    if (!c) return XTB_EINVAL;
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "xtb_set_partition: no system set");
    if (first < 1 || last > c->na || last < first - 1) XTB_FAIL(c, XTB_EINVAL, "xtb_set_partition: bad range %d..%d", first, last);
    c->a0 = first - 1; c->a1 = last;
    c->have_HS = c->have_density = false;
    return XTB_OK;
}

int xtb_set_block_layout(xtb_ctx* c, int parts) {
    if (!c) return XTB_EINVAL;
    if (parts < 1 || parts > 1024) XTB_FAIL(c, XTB_EINVAL, "xtb_set_block_layout: parts = %d (1..1024)", parts);
    if (parts != c->layout_parts) {
        c->layout_parts = parts;
        c->have_neigh = c->have_HS = c->have_blocks = c->have_density = c->have_dens = false;     // rowptr changes: rebuild the list
    }
    return XTB_OK;
}

// cut-off as get_near_neighbours derives it (Atomic_tools.f90:574-586)
extern "C++" double xtb_default_cutoff(const xtb_ctx* c) {
    double rm = 0.0;
    if (c->tb == XTB_TB_NRL) { for (auto& p : c->h_nrl) rm = p.Rc > rm ? p.Rc : rm; rm *= XTB_G_AU2A; }
    else if (c->tb == XTB_TB_PETTIFOR) {
        for (auto& p : c->h_pet) rm = p.rm > rm ? p.rm : rm;
        for (auto& p : c->h_pet_rep) rm = p.dm > rm ? p.dm : rm;      // `if (rm < r(1)) rm = r(1)`, Atomic_tools.f90:611-613
    }
    else if (c->tb == XTB_TB_DFTB) rm = c->dftb_rm;
    return rm;
}

int xtb_neighbours_compact(xtb_ctx* c, int cap, int32_t* nn_size, int32_t* nn_list, double* nn_dist, double* nn_dist_s, int* max_nn_out) {
    if (!c) return XTB_EINVAL;
    if (!c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "xtb_neighbours_compact: run xtb_neighbours first");
    cudaSetDevice(c->device);
    if (max_nn_out) *max_nn_out = c->max_nn;
    if (cap < c->max_nn && (nn_list || nn_dist || nn_dist_s)) XTB_FAIL(c, XTB_ECAPACITY, "cap %d < max_nn %d", cap, c->max_nn);
    const int na = c->na, dc = c->cap;
    if (nn_size) XTB_CUDA(c, cudaMemcpyAsync(nn_size, c->d_nn_size.p, sizeof(int) * na, cudaMemcpyDeviceToHost, c->stream));
    if (nn_list) XTB_CUDA(c, cudaMemcpy2DAsync(nn_list, sizeof(int) * cap, c->d_nn_list.p, sizeof(int) * dc, sizeof(int) * (cap < dc ? cap : dc), na, cudaMemcpyDeviceToHost, c->stream));
    if (nn_dist) XTB_CUDA(c, cudaMemcpy2DAsync(nn_dist, 32 * (size_t)cap, c->d_nn_dist.p, 32 * (size_t)dc, 32 * (size_t)(cap < dc ? cap : dc), na, cudaMemcpyDeviceToHost, c->stream));
    if (nn_dist_s) XTB_CUDA(c, cudaMemcpy2DAsync(nn_dist_s, 24 * (size_t)cap, c->d_nn_dist_s.p, 24 * (size_t)dc, 24 * (size_t)(cap < dc ? cap : dc), na, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_neighbours(xtb_ctx* c, double rm, int32_t* nn_size, int32_t* nn_list, double* nn_dist, double* nn_dist_s) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "xtb_neighbours: no system set");
    if (rm <= 0.0) rm = xtb_default_cutoff(c);
    if (rm <= 0.0) XTB_FAIL(c, XTB_EINVAL, "xtb_neighbours: no cut-off radius");
    int rc = xtb_impl_neighbours(c, rm);
    if (rc) return rc;
    if (!nn_size && !nn_list && !nn_dist && !nn_dist_s) return XTB_OK;
    // expand the compact device list into the reference's dense (Na,Na[,4|3]) host arrays.
    // Reference initialisation (Atomic_tools.f90:4241-4243): size = 0, list = 0, dist = 1d8; dist_s untouched.
    const int na = c->na, mx = c->max_nn > 0 ? c->max_nn : 1;
    std::vector<int32_t> sz(na), li((size_t)na * mx);
    std::vector<double> di((size_t)na * mx * 4), ds((size_t)na * mx * 3);
    if ((rc = xtb_neighbours_compact(c, mx, sz.data(), li.data(), di.data(), ds.data(), nullptr))) return rc;
    const size_t n2 = (size_t)na * na;
    if (nn_size) memcpy(nn_size, sz.data(), sizeof(int32_t) * na);
    if (nn_list) memset(nn_list, 0, sizeof(int32_t) * n2);
    if (nn_dist) for (size_t t = 0; t < 4 * n2; ++t) nn_dist[t] = 1e8;
    for (int i = 0; i < na; ++i)
        for (int k = 0; k < sz[i]; ++k) {
            const size_t e = (size_t)i * mx + k;
            const size_t d = (size_t)i + (size_t)na * k;               // (i, k) column-major
            if (nn_list) nn_list[d] = li[e] + 1;
            if (nn_dist) for (int q = 0; q < 4; ++q) nn_dist[d + n2 * q] = di[4 * e + q];
            if (nn_dist_s) for (int q = 0; q < 3; ++q) nn_dist_s[d + n2 * q] = ds[3 * e + q];
        }
    return XTB_OK;
}

int xtb_radial(xtb_ctx* c, double* M_Vij, double* M_dVij, double* M_SVij, double* M_dSVij, double* M_lmn, double* M_x1, double* M_xrr) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "xtb_radial: run xtb_neighbours first");
    if (c->tb != XTB_TB_NRL && c->tb != XTB_TB_DFTB) XTB_FAIL(c, XTB_EINVAL, "xtb_radial: dense tables are provided for the NRL and DFTB branches");
    const int nv = c->tb == XTB_TB_NRL ? 10 : dftb::n_v(c->n_orb);     // channels of M_V*(na,na,nv)
    if (!M_Vij && !M_dVij && !M_SVij && !M_dSVij && !M_lmn && !M_x1 && !M_xrr) return XTB_OK;   // fused on the device otherwise
    const int na = c->na, cap = c->cap;
    const size_t ne = (size_t)na * cap;
    int rc;
    if ((rc = xtb_reserve(c, c->d_tab, sizeof(double) * 49 * ne))) return rc;
    double* d_rad = (double*)c->d_tab.p;
    double* d_geo = d_rad + 40 * ne;
    if ((rc = (c->tb == XTB_TB_NRL ? xtb_impl_nrl_tables(c, d_rad, d_geo) : xtb_impl_dftb_tables(c, d_rad, d_geo)))) return rc;
    std::vector<double> rad(40 * ne), geo(9 * ne);
    std::vector<int32_t> sz(na), li(ne);
    XTB_CUDA(c, cudaMemcpyAsync(rad.data(), d_rad, sizeof(double) * 40 * ne, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaMemcpyAsync(geo.data(), d_geo, sizeof(double) * 9 * ne, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaMemcpyAsync(sz.data(), c->d_nn_size.p, sizeof(int) * na, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaMemcpyAsync(li.data(), c->d_nn_list.p, sizeof(int) * ne, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    const size_t n2 = (size_t)na * na;
    double* Mv[4] = {M_Vij, M_dVij, M_SVij, M_dSVij};
    for (int t = 0; t < 4; ++t) if (Mv[t]) memset(Mv[t], 0, sizeof(double) * nv * n2);         // TB_NRL.f90:2137-2140, TB_DFTB.f90:86-89
    double* Mg[3] = {M_lmn, M_xrr, M_x1};
    for (int t = 0; t < 3; ++t) if (Mg[t]) memset(Mg[t], 0, sizeof(double) * 3 * n2);           // TB.f90:3448-3450
    for (int i = 0; i < na; ++i)
        for (int k = 0; k < sz[i]; ++k) {
            const size_t e = (size_t)i * cap + k;
            const int j = li[e];
            for (int t = 0; t < 4; ++t) if (Mv[t])
                for (int ch = 0; ch < nv; ++ch) Mv[t][(size_t)i + (size_t)na * j + n2 * ch] = rad[40 * e + 10 * t + ch];   // M(i,j,ch)
            for (int t = 0; t < 3; ++t) if (Mg[t])
                for (int a = 0; a < 3; ++a) Mg[t][a + 3 * ((size_t)i + (size_t)na * j)] = geo[9 * e + 3 * t + a];          // M(a,i,j)
        }
    return XTB_OK;
}

int xtb_assemble_HS(xtb_ctx* c, double* H_non, double* Sij) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "xtb_assemble_HS: run xtb_neighbours first");
    int rc = xtb_impl_assemble(c, true);
    if (rc) return rc;
    const size_t N = (size_t)c->na * c->n_orb;
    const size_t off = (size_t)c->a0 * c->n_orb * N, cnt = (size_t)(c->a1 - c->a0) * c->n_orb * N;
    if (H_non && cnt) XTB_CUDA(c, cudaMemcpyAsync(H_non + off, (double*)c->d_H.p + off, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
    if (Sij && cnt) {
        if (c->tb == XTB_TB_PETTIFOR) XTB_FAIL(c, XTB_EINVAL, "xtb_assemble_HS: orthogonal parametrisation has no overlap matrix");
        XTB_CUDA(c, cudaMemcpyAsync(Sij + off, (double*)c->d_S2.p + off, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
    }
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_assemble_blocks(xtb_ctx* c) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "xtb_assemble_blocks: run xtb_neighbours first");
    return xtb_impl_assemble(c, false);
}

// dense column panels of atoms first..last (1-based, inclusive) from the block stores; no host copies, no wait
extern "C++" int xtb_impl_expand_dense(xtb_ctx* c, int first, int last, bool timed) {
    if (!c->have_blocks) XTB_FAIL(c, XTB_ESTATE, "xtb_expand_dense: no block store (xtb_assemble_blocks / xtb_assemble_HS first)");
    if (first < 1 || last > c->na || last < first - 1) XTB_FAIL(c, XTB_EINVAL, "xtb_expand_dense: bad range %d..%d", first, last);
    int rc;
    if (timed) XTB_T0(c, 6);
    if ((rc = xtb_dense_update(c, first - 1, last))) return rc;
    if (timed) XTB_T1(c, 6);
    c->have_HS = true;
    return XTB_OK;
}

int xtb_expand_dense(xtb_ctx* c, int first, int last, double* H_non, double* Sij) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (Sij && c->tb == XTB_TB_PETTIFOR) XTB_FAIL(c, XTB_EINVAL, "xtb_expand_dense: orthogonal parametrisation has no overlap matrix");
    int rc;
    if ((rc = xtb_impl_expand_dense(c, first, last))) return rc;
    const size_t N = (size_t)c->na * c->n_orb;
    const size_t off = (size_t)(first - 1) * c->n_orb * N, cnt = (size_t)(last - first + 1) * c->n_orb * N;
    if (H_non && cnt) XTB_CUDA(c, cudaMemcpyAsync(H_non + off, (double*)c->d_H.p + off, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
    if (Sij && cnt) XTB_CUDA(c, cudaMemcpyAsync(Sij + off, (double*)c->d_S2.p + off, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
    return xtb_stage_finish(c);
}

int xtb_set_eigen_device(xtb_ctx* c, const double* dHa, const double* dEi, const double* dfe) {
    if (!c) return XTB_EINVAL;
    if (!dHa || !dEi || !dfe) XTB_FAIL(c, XTB_EINVAL, "xtb_set_eigen_device: null pointer");
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "xtb_set_eigen_device: no system set");
    cudaSetDevice(c->device);
    c->d_Ha = dHa; c->d_HaT = nullptr; c->d_Ei = dEi; c->d_fe = dfe;
    c->have_eigvec = false;
    c->K = (long long)c->na * c->n_orb;
    c->have_eigen = true; c->have_density = false; c->have_keep = false;
    return xtb_impl_prepare_occupations(c);      // kept states and weights are fixed from here (one small read-back, now)
}

int xtb_cutoff(xtb_ctx* c, double* rm) {
    if (!c || !rm) return XTB_EINVAL;
    if (c->tb == XTB_TB_NONE) XTB_FAIL(c, XTB_ESTATE, "xtb_cutoff: set the TB parameters first");
    *rm = xtb_default_cutoff(c);
    return XTB_OK;
}

int xtb_set_eigen_device_t(xtb_ctx* c, const double* dHaT, const double* dEi, const double* dfe) {
    if (!c) return XTB_EINVAL;
    if (!dHaT || !dEi || !dfe) XTB_FAIL(c, XTB_EINVAL, "xtb_set_eigen_device_t: null pointer");
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "xtb_set_eigen_device_t: no system set");
    if (((uintptr_t)dHaT & 15) != 0) XTB_FAIL(c, XTB_EINVAL, "xtb_set_eigen_device_t: HaT must be 16-byte aligned");
    cudaSetDevice(c->device);
    c->d_HaT = dHaT; c->d_Ha = nullptr; c->d_Ei = dEi; c->d_fe = dfe;
    c->have_eigvec = false;
    c->K = (long long)c->na * c->n_orb;
    c->have_eigen = true; c->have_density = false; c->have_keep = false;
    return xtb_impl_prepare_occupations(c);
}

int xtb_set_eigen(xtb_ctx* c, const double* Ha, const double* Ei, const double* fe) {
    if (!c) return XTB_EINVAL;
    if (!Ha || !Ei) XTB_FAIL(c, XTB_EINVAL, "xtb_set_eigen: null pointer");
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "xtb_set_eigen: no system set");
    cudaSetDevice(c->device);
    const size_t N = (size_t)c->na * c->n_orb;
    int rc;
    if ((rc = xtb_reserve(c, c->own_Ha, sizeof(double) * N * N))) return rc;
    if ((rc = xtb_reserve(c, c->own_Ei, sizeof(double) * N))) return rc;
    if ((rc = xtb_reserve(c, c->own_fe, sizeof(double) * N))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->own_Ha.p, Ha, sizeof(double) * N * N, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaMemcpyAsync(c->own_Ei.p, Ei, sizeof(double) * N, cudaMemcpyHostToDevice, c->stream));
    if (fe) XTB_CUDA(c, cudaMemcpyAsync(c->own_fe.p, fe, sizeof(double) * N, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (!fe) {      // eigenvectors and eigenvalues now, occupations later (xtb_set_occupations / xtb_fermi_occupations)
        c->d_Ha = (const double*)c->own_Ha.p; c->d_HaT = nullptr; c->d_Ei = (const double*)c->own_Ei.p; c->d_fe = nullptr;
        c->K = (long long)N;
        c->have_eigvec = false; c->have_eigen = false; c->have_density = false; c->have_keep = false;
        return XTB_OK;
    }
    return xtb_set_eigen_device(c, (const double*)c->own_Ha.p, (const double*)c->own_Ei.p, (const double*)c->own_fe.p);
}

int xtb_eigensolve(xtb_ctx* c, double* Ei, double* eigen_S, double* Ha) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_HS) XTB_FAIL(c, XTB_ESTATE, "xtb_eigensolve: no dense H_non / Sij on the device (xtb_assemble_HS or xtb_expand_dense first)");
    if (c->a0 != 0 || c->a1 != c->na) XTB_FAIL(c, XTB_EINVAL, "xtb_eigensolve: needs the complete matrices (expand all column panels first)");
    int rc = xtb_impl_eigensolve(c);
    if (rc) return rc;
    const size_t N = (size_t)c->na * c->n_orb;
    if (Ei) XTB_CUDA(c, cudaMemcpyAsync(Ei, c->own_Ei.p, sizeof(double) * N, cudaMemcpyDeviceToHost, c->stream));
    if (eigen_S && c->tb != XTB_TB_PETTIFOR) XTB_CUDA(c, cudaMemcpyAsync(eigen_S, c->d_eigS.p, sizeof(double) * N, cudaMemcpyDeviceToHost, c->stream));
    if (Ha) {
        if ((rc = xtb_ensure_Ha_std(c))) return rc;
        XTB_CUDA(c, cudaMemcpyAsync(Ha, c->d_Ha, sizeof(double) * N * N, cudaMemcpyDeviceToHost, c->stream));
    }
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_get_eigenvalues(xtb_ctx* c, double* Ei) {
    if (!c) return XTB_EINVAL;
    if (!Ei) XTB_FAIL(c, XTB_EINVAL, "xtb_get_eigenvalues: null pointer");
    if (!c->d_Ei) XTB_FAIL(c, XTB_ESTATE, "xtb_get_eigenvalues: no eigenvalues on the device");
    cudaSetDevice(c->device);
    XTB_CUDA(c, cudaMemcpyAsync(Ei, c->d_Ei, sizeof(double) * (size_t)c->na * c->n_orb, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_set_occupations(xtb_ctx* c, const double* fe) {
    if (!c) return XTB_EINVAL;
    if (!fe) XTB_FAIL(c, XTB_EINVAL, "xtb_set_occupations: null pointer");
    if ((!c->d_Ha && !c->d_HaT) || !c->d_Ei) XTB_FAIL(c, XTB_ESTATE, "xtb_set_occupations: no eigenvectors on the device (xtb_eigensolve, xtb_set_eigen or xtb_set_eigen_device first)");
    cudaSetDevice(c->device);
    const size_t N = (size_t)c->na * c->n_orb;
    int rc;
    if ((rc = xtb_reserve(c, c->own_fe, sizeof(double) * N))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->own_fe.p, fe, sizeof(double) * N, cudaMemcpyHostToDevice, c->stream));
    c->d_fe = (const double*)c->own_fe.p;       // Ha and Ei stay where they are (device eigensolve result or the caller's buffers)
    c->K = (long long)N;
    c->have_eigen = true; c->have_density = false; c->have_keep = false;
    return xtb_impl_prepare_occupations(c);      // synchronises (the host array is free when this returns)
}

int xtb_set_density_cutoff(xtb_ctx* c, double f_cut) {
    if (!c) return XTB_EINVAL;
    if (!(f_cut >= 0.0)) XTB_FAIL(c, XTB_EINVAL, "xtb_set_density_cutoff: f_cut must be >= 0");
    if (f_cut != c->f_cut) c->have_keep = false;
    c->f_cut = f_cut;
    return XTB_OK;
}

// dense (N,N) copies of the block-sparse rho / rho_E of the last density pass: zero outside the neighbour pattern
int xtb_density_get(xtb_ctx* c, double* Aij, double* Aij_x_Ei) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_density) XTB_FAIL(c, XTB_ESTATE, "xtb_density_get: run xtb_density first");
    if (!Aij && !Aij_x_Ei) return XTB_OK;
    if (c->a0 != 0 || c->a1 != c->na) XTB_FAIL(c, XTB_EINVAL, "xtb_density: dense outputs need the full (unpartitioned) context");
    const size_t N = (size_t)c->na * c->n_orb;
    int rc;
    if ((rc = xtb_reserve(c, c->d_dense_out, sizeof(double) * N * N))) return rc;     // own scratch: H_non / Sij stay resident (xtb_dense_update)
    double* outs[2] = {Aij, Aij_x_Ei};
    const double* blks[2] = {(const double*)c->d_rho.p, (const double*)c->d_rhoE.p};
    for (int t = 0; t < 2; ++t) {
        if (!outs[t]) continue;
        if ((rc = xtb_dense_from_blocks(c, blks[t], nullptr, (double*)c->d_dense_out.p, nullptr, true))) return rc;
        XTB_CUDA(c, cudaMemcpyAsync(outs[t], c->d_dense_out.p, sizeof(double) * N * N, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

int xtb_density(xtb_ctx* c, double* Aij, double* Aij_x_Ei) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "xtb_density: run xtb_neighbours first");
    if (!c->have_eigen) XTB_FAIL(c, XTB_ESTATE, "xtb_density: no eigenvectors / occupations set (xtb_set_eigen, xtb_set_occupations)");
    int rc = xtb_impl_density(c);
    if (rc) return rc;
    return xtb_density_get(c, Aij, Aij_x_Ei);
}

int xtb_density_mirror(xtb_ctx* c) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_density) XTB_FAIL(c, XTB_ESTATE, "xtb_density_mirror: run xtb_density first");
    int rc = xtb_impl_mirror_density(c);
    if (rc) return rc;
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_forces(xtb_ctx* c, double* F_att) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_density) XTB_FAIL(c, XTB_ESTATE, "xtb_forces: run xtb_density first");
    int rc = xtb_impl_forces(c);
    if (rc) return rc;
    const size_t cnt = (size_t)(c->a1 - c->a0) * 3;
    if (F_att && cnt) {
        XTB_CUDA(c, cudaMemcpyAsync(F_att + 3 * (size_t)c->a0, (double*)c->d_F.p + 3 * (size_t)c->a0, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

int xtb_pressure(xtb_ctx* c, double* SCforce_att) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_density) XTB_FAIL(c, XTB_ESTATE, "xtb_pressure: run xtb_density first");
    int rc = xtb_impl_pressure(c);
    if (rc) return rc;
    if (SCforce_att) {
        XTB_CUDA(c, cudaMemcpyAsync(SCforce_att, c->d_SC.p, sizeof(double) * 9, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

// ---- nonadiabatic coupling ----
int xtb_save_eigen_prev(xtb_ctx* c) {
    if (!c) return XTB_EINVAL;
    if (!c->d_Ha && !c->d_HaT) XTB_FAIL(c, XTB_ESTATE, "xtb_save_eigen_prev: no eigenvectors on the device");
    cudaSetDevice(c->device);
    { int rc0 = xtb_ensure_Ha_std(c); if (rc0) return rc0; }
    const size_t N = (size_t)c->na * c->n_orb;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Ha0, sizeof(double) * N * N))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_Ha0.p, c->d_Ha, sizeof(double) * N * N, cudaMemcpyDeviceToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_Ha0 = true;
    return XTB_OK;
}

int xtb_set_eigen_prev(xtb_ctx* c, const double* Ha0) {
    if (!c) return XTB_EINVAL;
    if (!Ha0) XTB_FAIL(c, XTB_EINVAL, "xtb_set_eigen_prev: null pointer");
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "xtb_set_eigen_prev: no system set");
    cudaSetDevice(c->device);
    const size_t N = (size_t)c->na * c->n_orb;
    int rc;
    if ((rc = xtb_reserve(c, c->d_Ha0, sizeof(double) * N * N))) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_Ha0.p, Ha0, sizeof(double) * N * N, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_Ha0 = true;
    return XTB_OK;
}

int xtb_coupling(xtb_ctx* c, int kind_M, double* Mij) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    if (!c->d_Ha && !c->d_HaT) XTB_FAIL(c, XTB_ESTATE, "xtb_coupling: no eigenvectors on the device (xtb_set_eigen / xtb_eigensolve first)");
    if (!c->have_Ha0) XTB_FAIL(c, XTB_ESTATE, "xtb_coupling: no previous-step eigenvectors (xtb_save_eigen_prev / xtb_set_eigen_prev first)");
    if (c->tb != XTB_TB_PETTIFOR) {
        if (!c->have_blocks) XTB_FAIL(c, XTB_ESTATE, "xtb_coupling: no overlap blocks on the device (xtb_assemble_HS first)");
        if (c->a0 != 0 || c->a1 != c->na) XTB_FAIL(c, XTB_EINVAL, "xtb_coupling: needs the overlap blocks of all atoms (not partitioned)");
    }
    int rc = xtb_impl_coupling(c, kind_M);
    if (rc) return rc;
    if (Mij) {
        const size_t N = (size_t)c->na * c->n_orb;
        XTB_CUDA(c, cudaMemcpyAsync(Mij, c->d_cpl_M.p, sizeof(double) * N * N, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

// ---- repulsive pair terms ----
static int rep_ready(xtb_ctx* c, const char* who) {
    if (c->tb != XTB_TB_PETTIFOR || c->h_pet_rep.empty()) XTB_FAIL(c, XTB_ESTATE, "%s: no repulsive parameters (xtb_set_params_pettifor_rep)", who);
    if (!c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "%s: run xtb_neighbours first", who);
    return XTB_OK;
}

int xtb_repulsive_energy(xtb_ctx* c, double* E_rep, double* Epot) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = rep_ready(c, "xtb_repulsive_energy"))) return rc;
    if ((rc = xtb_impl_pettifor_rep_energy(c))) return rc;
    if (E_rep) XTB_CUDA(c, cudaMemcpyAsync(E_rep, c->d_Erep.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (Epot) XTB_CUDA(c, cudaMemcpyAsync(Epot, c->d_Epot.p, sizeof(double) * c->na, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_repulsive_forces(xtb_ctx* c, double* F_rep) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = rep_ready(c, "xtb_repulsive_forces"))) return rc;
    if ((rc = xtb_impl_pettifor_rep_forces(c))) return rc;
    const size_t cnt = (size_t)(c->a1 - c->a0) * 3;
    if (F_rep && cnt) {
        XTB_CUDA(c, cudaMemcpyAsync(F_rep + 3 * (size_t)c->a0, (double*)c->d_Frep.p + 3 * (size_t)c->a0, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

int xtb_repulsive_pressure(xtb_ctx* c, double* SCforce_rep) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = rep_ready(c, "xtb_repulsive_pressure"))) return rc;
    if ((rc = xtb_impl_pettifor_rep_pressure(c))) return rc;
    if (SCforce_rep) {
        XTB_CUDA(c, cudaMemcpyAsync(SCforce_rep, c->d_SCrep.p, sizeof(double) * 9, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

// ---- short-range wall (second half) ----
int xtb_set_params_wall(xtb_ctx* c, int n_kinds, const xtb_wall_pair* p) {
    if (!c) return XTB_EINVAL;
    if (c->tb == XTB_TB_NONE) XTB_FAIL(c, XTB_ESTATE, "xtb_set_params_wall: set the TB parameters first");
    if (n_kinds == 0 || !p) { c->h_wall.clear(); return XTB_OK; }          // "proceeding without short-range forces"
    if (n_kinds != c->n_kinds) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_wall: n_kinds=%d (TB part has %d)", n_kinds, c->n_kinds);
    cudaSetDevice(c->device);
    const size_t cnt = (size_t)n_kinds * n_kinds;
    for (size_t q = 0; q < cnt; ++q)
        if (!(p[q].dd > 0.0)) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_wall: pair %zu has smoothing width dd=%g", q, p[q].dd);
    c->h_wall.assign(p, p + cnt);
    c->h_short.clear();                      // one class in the TB_Expwall slot
    int rc = xtb_reserve(c, c->d_wall_params, sizeof(xtb_wall_pair) * cnt);
    if (rc) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_wall_params.p, p, sizeof(xtb_wall_pair) * cnt, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

static int wall_ready(xtb_ctx* c, const char* who) {
    if (c->h_wall.empty()) XTB_FAIL(c, XTB_ESTATE, "%s: no wall parameters (xtb_set_params_wall)", who);
    if (!c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "%s: run xtb_neighbours first", who);
    return XTB_OK;
}

int xtb_wall_energy(xtb_ctx* c, double* E_wall, double* Epot) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = wall_ready(c, "xtb_wall_energy"))) return rc;
    if ((rc = xtb_impl_wall_energy(c))) return rc;
    if (E_wall) XTB_CUDA(c, cudaMemcpyAsync(E_wall, c->d_Ewall.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (Epot) XTB_CUDA(c, cudaMemcpyAsync(Epot, c->d_Ewall_at.p, sizeof(double) * c->na, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_wall_forces(xtb_ctx* c, double* F_wall) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = wall_ready(c, "xtb_wall_forces"))) return rc;
    if ((rc = xtb_impl_wall_forces(c))) return rc;
    const size_t cnt = (size_t)(c->a1 - c->a0) * 3;
    if (F_wall && cnt) {
        XTB_CUDA(c, cudaMemcpyAsync(F_wall + 3 * (size_t)c->a0, (double*)c->d_Fwall.p + 3 * (size_t)c->a0, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

int xtb_wall_pressure(xtb_ctx* c, double* SCforce_wall) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = wall_ready(c, "xtb_wall_pressure"))) return rc;
    if ((rc = xtb_impl_wall_pressure(c))) return rc;
    if (SCforce_wall) {
        XTB_CUDA(c, cudaMemcpyAsync(SCforce_wall, c->d_SCwall.p, sizeof(double) * 9, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

// ---- the 'General' short-range class (TB_Short_Rep) in the slot of the simple wall ----
int xtb_set_params_short(xtb_ctx* c, int n_kinds, const xtb_short_pair* p, const double* Z) {
    if (!c) return XTB_EINVAL;
    if (c->tb == XTB_TB_NONE) XTB_FAIL(c, XTB_ESTATE, "xtb_set_params_short: set the TB parameters first");
    if (n_kinds == 0 || !p) { c->h_short.clear(); return XTB_OK; }        // "proceeding without short-range forces"
    if (n_kinds != c->n_kinds) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: n_kinds=%d (TB part has %d)", n_kinds, c->n_kinds);
    if (!Z) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: atomic numbers Z[n_kinds] are required");
    cudaSetDevice(c->device);
    const size_t cnt = (size_t)n_kinds * n_kinds;
    size_t tab_doubles = 0;
    for (size_t q = 0; q < cnt; ++q) {
        const xtb_short_pair& a = p[q];
        if (!(a.dd > 0.0)) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: pair %zu has cut-off width dd=%g", q, a.dd);
        if (a.n_pow < 0 || a.n_pow > XTB_SHORT_MAX_POW) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: pair %zu has %d power terms (max %d)", q, a.n_pow, XTB_SHORT_MAX_POW);
        if (a.n_tab < 0 || a.n_tab == 1) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: pair %zu has a %d-point table", q, a.n_tab);
        if (a.n_tab > 0) {
            if (!a.tab_R || !a.tab_E) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: pair %zu: table without R / E", q);
            if (a.tab_spline && (!a.tab_a || !a.tab_b || !a.tab_c || !a.tab_d)) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: pair %zu: spline without coefficients", q);
            for (int k = 1; k < a.n_tab; ++k)
                if (!(a.tab_R[k] > a.tab_R[k - 1])) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: pair %zu: table radii not increasing at point %d", q, k + 1);
            tab_doubles += 6 * (size_t)a.n_tab;
        }
        if ((a.use_zbl || a.use_nlh) && !(Z[q % n_kinds] > 0.0 && Z[q / n_kinds] > 0.0)) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_short: pair %zu: ZBL / NLH need atomic numbers", q);
    }
    int rc;
    if ((rc = xtb_reserve(c, c->d_short_params, sizeof(ShortDev) * cnt))) return rc;
    if ((rc = xtb_reserve(c, c->d_short_tab, sizeof(double) * (tab_doubles ? tab_doubles : 1)))) return rc;
    std::vector<ShortDev> dev(cnt);
    std::vector<double> tab(tab_doubles ? tab_doubles : 1, 0.0);
    size_t off = 0;
    for (size_t q = 0; q < cnt; ++q) {
        const xtb_short_pair& a = p[q];
        ShortDev& d = dev[q];
        memset(&d, 0, sizeof d);
        d.d0 = a.d0; d.dd = a.dd;
        d.use_exp = a.use_exp != 0; d.use_inv_exp = a.use_inv_exp != 0; d.use_zbl = a.use_zbl != 0; d.use_nlh = a.use_nlh != 0;
        d.exp_Phi = a.exp_Phi; d.exp_r0 = a.exp_r0; d.exp_a = a.exp_a; d.inv_C = a.inv_C; d.inv_r0 = a.inv_r0;
        d.n_pow = a.n_pow; d.n_tab = a.n_tab; d.tab_spline = a.tab_spline != 0;
        for (int k = 0; k < a.n_pow; ++k) { d.pow_Phi[k] = a.pow_Phi[k]; d.pow_r0[k] = a.pow_r0[k]; d.pow_m[k] = a.pow_m[k]; }
        for (int k = 0; k < 3; ++k) { d.nlh_a[k] = a.nlh_a[k]; d.nlh_b[k] = a.nlh_b[k]; }
        d.Z1 = Z[q % n_kinds]; d.Z2 = Z[q / n_kinds];                   // TB_Expwall(KOA1,KOA2) with Z1 = Z(KOA1), Z2 = Z(KOA2)
        d.tab = (const double*)c->d_short_tab.p + off;
        if (a.n_tab > 0) {
            const double* src[6] = {a.tab_R, a.tab_E, a.tab_a, a.tab_b, a.tab_c, a.tab_d};
            for (int w = 0; w < 6; ++w)
                for (int k = 0; k < a.n_tab; ++k) tab[off + (size_t)w * a.n_tab + k] = (src[w] && (w < 2 || a.tab_spline)) ? src[w][k] : 0.0;
            off += 6 * (size_t)a.n_tab;
        }
    }
    XTB_CUDA(c, cudaMemcpyAsync(c->d_short_params.p, dev.data(), sizeof(ShortDev) * cnt, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaMemcpyAsync(c->d_short_tab.p, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->h_short = dev;
    c->h_wall.clear();                       // one class in the TB_Expwall slot
    return XTB_OK;
}

static int short_ready(xtb_ctx* c, const char* who) {
    if (c->h_short.empty()) XTB_FAIL(c, XTB_ESTATE, "%s: no short-range parameters (xtb_set_params_short)", who);
    if (!c->have_neigh) XTB_FAIL(c, XTB_ESTATE, "%s: run xtb_neighbours first", who);
    return XTB_OK;
}

int xtb_short_energy(xtb_ctx* c, double* E_short, double* Epot) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = short_ready(c, "xtb_short_energy"))) return rc;
    if ((rc = xtb_impl_wall_energy(c))) return rc;
    if (E_short) XTB_CUDA(c, cudaMemcpyAsync(E_short, c->d_Ewall.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (Epot) XTB_CUDA(c, cudaMemcpyAsync(Epot, c->d_Ewall_at.p, sizeof(double) * c->na, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_short_forces(xtb_ctx* c, double* F_short) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = short_ready(c, "xtb_short_forces"))) return rc;
    if ((rc = xtb_impl_wall_forces(c))) return rc;
    const size_t cnt = 3 * (size_t)(c->a1 - c->a0);
    if (F_short && cnt) {
        XTB_CUDA(c, cudaMemcpyAsync(F_short + 3 * (size_t)c->a0, (double*)c->d_Fwall.p + 3 * (size_t)c->a0, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

int xtb_short_pressure(xtb_ctx* c, double* SCforce_short) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = short_ready(c, "xtb_short_pressure"))) return rc;
    if ((rc = xtb_impl_wall_pressure(c))) return rc;
    if (SCforce_short) {
        XTB_CUDA(c, cudaMemcpyAsync(SCforce_short, c->d_SCwall.p, sizeof(double) * 9, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

// ---- van der Waals term (third part) ----
int xtb_set_params_vdw(xtb_ctx* c, int kind, int n_kinds, const xtb_vdw_pair* p) {
    if (!c) return XTB_EINVAL;
    if (c->tb == XTB_TB_NONE) XTB_FAIL(c, XTB_ESTATE, "xtb_set_params_vdw: set the TB parameters first");
    if (n_kinds == 0 || !p) { c->h_vdw.clear(); c->vdw_kind = 0; return XTB_OK; }          // "proceeding without van der Waals forces"
    if (kind != XTB_VDW_GIRIFALCO && kind != XTB_VDW_LJ_CUT && kind != XTB_VDW_ILJ_CUT)
        XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_vdw: class %d (1 Girifalco, 2 LJ_cut, 3 ILJ_cut)", kind);
    if (n_kinds != c->n_kinds) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_vdw: n_kinds=%d (TB part has %d)", n_kinds, c->n_kinds);
    cudaSetDevice(c->device);
    const size_t cnt = (size_t)n_kinds * n_kinds;
    c->h_vdw.assign(p, p + cnt);
    c->vdw_kind = kind;
    c->vdw_sym = true;
    for (int a = 0; a < n_kinds; ++a)
        for (int b = 0; b < a; ++b)
            if (memcmp(&p[a + n_kinds * b], &p[b + n_kinds * a], sizeof(xtb_vdw_pair)) != 0) c->vdw_sym = false;
    int rc = xtb_reserve(c, c->d_vdw_params, sizeof(xtb_vdw_pair) * cnt);
    if (rc) return rc;
    XTB_CUDA(c, cudaMemcpyAsync(c->d_vdw_params.p, p, sizeof(xtb_vdw_pair) * cnt, cudaMemcpyHostToDevice, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

static int vdw_ready(xtb_ctx* c, const char* who) {
    if (c->h_vdw.empty()) XTB_FAIL(c, XTB_ESTATE, "%s: no vdW parameters (xtb_set_params_vdw)", who);
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "%s: no system set", who);
    return XTB_OK;
}

int xtb_vdw_image_cells(xtb_ctx* c, int* N, double* R_cut) {
    if (!c || !N) return XTB_EINVAL;
    int rc;
    if ((rc = vdw_ready(c, "xtb_vdw_image_cells"))) return rc;
    return xtb_impl_vdw_cells(c, N, R_cut);
}

int xtb_vdw_energy(xtb_ctx* c, double* E_vdw, double* Epot) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = vdw_ready(c, "xtb_vdw_energy"))) return rc;
    if ((rc = xtb_impl_vdw(c, 1u))) return rc;
    if (E_vdw) XTB_CUDA(c, cudaMemcpyAsync(E_vdw, c->d_Evdw.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (Epot) XTB_CUDA(c, cudaMemcpyAsync(Epot, c->d_Evdw_at.p, sizeof(double) * c->na, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_vdw_forces(xtb_ctx* c, double* F_vdw) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = vdw_ready(c, "xtb_vdw_forces"))) return rc;
    if ((rc = xtb_impl_vdw(c, 2u))) return rc;
    const size_t cnt = 3 * (size_t)(c->a1 - c->a0);
    if (F_vdw && cnt) {
        XTB_CUDA(c, cudaMemcpyAsync(F_vdw + 3 * (size_t)c->a0, (double*)c->d_Fvdw.p + 3 * (size_t)c->a0, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

int xtb_vdw_pressure(xtb_ctx* c, double* SCforce_vdw) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = vdw_ready(c, "xtb_vdw_pressure"))) return rc;
    if ((rc = xtb_impl_vdw(c, 4u))) return rc;
    if (SCforce_vdw) {
        XTB_CUDA(c, cudaMemcpyAsync(SCforce_vdw, c->d_SCvdw.p, sizeof(double) * 9, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

// ---- Coulomb term of a charged supercell (fourth part) ----
int xtb_set_params_coulomb(xtb_ctx* c, int n_kinds, const xtb_coulomb_pair* p) {
    if (!c) return XTB_EINVAL;
    if (c->tb == XTB_TB_NONE) XTB_FAIL(c, XTB_ESTATE, "xtb_set_params_coulomb: set the TB parameters first");
    if (n_kinds == 0 || !p) { c->h_coul.clear(); return XTB_OK; }
    if (n_kinds != c->n_kinds) XTB_FAIL(c, XTB_EINVAL, "xtb_set_params_coulomb: n_kinds=%d (TB part has %d)", n_kinds, c->n_kinds);
    const size_t cnt = (size_t)n_kinds * n_kinds;
    c->h_coul.assign(p, p + cnt);
    c->coul_sym = true;
    for (int a = 0; a < n_kinds; ++a)
        for (int b = 0; b < a; ++b)
            if (memcmp(&p[a + n_kinds * b], &p[b + n_kinds * a], sizeof(xtb_coulomb_pair)) != 0) c->coul_sym = false;
    c->coul_dirty = true;
    return XTB_OK;
}

int xtb_set_charge(xtb_ctx* c, double Q) {
    if (!c) return XTB_EINVAL;
    if (Q != c->coul_Q) { c->coul_Q = Q; c->coul_dirty = true; }
    return XTB_OK;
}

static int coulomb_ready(xtb_ctx* c, const char* who) {
    if (c->h_coul.empty()) XTB_FAIL(c, XTB_ESTATE, "%s: no Coulomb parameters (xtb_set_params_coulomb)", who);
    if (c->na <= 0) XTB_FAIL(c, XTB_ESTATE, "%s: no system set", who);
    return XTB_OK;
}

int xtb_coulomb_image_cells(xtb_ctx* c, int* N) {
    if (!c || !N) return XTB_EINVAL;
    int rc;
    if ((rc = coulomb_ready(c, "xtb_coulomb_image_cells"))) return rc;
    return xtb_impl_coulomb_cells(c, N);
}

int xtb_coulomb_energy(xtb_ctx* c, double* E_coul, double* Epot) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = coulomb_ready(c, "xtb_coulomb_energy"))) return rc;
    if ((rc = xtb_impl_coulomb(c, 1u))) return rc;
    if (E_coul) XTB_CUDA(c, cudaMemcpyAsync(E_coul, c->d_Ecoul.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (Epot) XTB_CUDA(c, cudaMemcpyAsync(Epot, c->d_Ecoul_at.p, sizeof(double) * c->na, cudaMemcpyDeviceToHost, c->stream));
    XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    return XTB_OK;
}

int xtb_coulomb_forces(xtb_ctx* c, double* F_coul) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = coulomb_ready(c, "xtb_coulomb_forces"))) return rc;
    if ((rc = xtb_impl_coulomb(c, 2u))) return rc;
    const size_t cnt = 3 * (size_t)(c->a1 - c->a0);
    if (F_coul && cnt) {
        XTB_CUDA(c, cudaMemcpyAsync(F_coul + 3 * (size_t)c->a0, (double*)c->d_Fcoul.p + 3 * (size_t)c->a0, sizeof(double) * cnt, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

int xtb_coulomb_pressure(xtb_ctx* c, double* SCforce_coul) {
    if (!c) return XTB_EINVAL;
    cudaSetDevice(c->device);
    int rc;
    if ((rc = coulomb_ready(c, "xtb_coulomb_pressure"))) return rc;
    if ((rc = xtb_impl_coulomb(c, 4u))) return rc;
    if (SCforce_coul) {
        XTB_CUDA(c, cudaMemcpyAsync(SCforce_coul, c->d_SCcoul.p, sizeof(double) * 9, cudaMemcpyDeviceToHost, c->stream));
        XTB_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return XTB_OK;
}

int xtb_device_buffer(xtb_ctx* c, int which, void** dptr, size_t* bytes) {
    if (!c || !dptr) return XTB_EINVAL;
    const size_t N = (size_t)c->na * c->n_orb;
    const size_t nb2 = (size_t)c->n_orb * c->n_orb;
    DevBuf* b = nullptr; size_t used = 0;
    switch (which) {
        case XTB_BUF_H: b = &c->d_H; used = 8 * N * N; break;
        case XTB_BUF_S: b = &c->d_S2; used = 8 * N * N; break;
        case XTB_BUF_RHO: b = &c->d_rho; used = 8 * nb2 * (size_t)c->n_blocks; break;
        case XTB_BUF_RHOE: b = &c->d_rhoE; used = 8 * nb2 * (size_t)c->n_blocks; break;
        case XTB_BUF_FORCES: b = &c->d_F; used = 24 * (size_t)c->na; break;
        case XTB_BUF_ROWPTR: b = &c->d_rowptr; used = 4 * ((size_t)c->na + 1); break;
        case XTB_BUF_NNLIST: b = &c->d_nn_list; used = 4 * (size_t)c->na * c->cap; break;
        case XTB_BUF_NNSIZE: b = &c->d_nn_size; used = 4 * (size_t)c->na; break;
        case XTB_BUF_SCFORCE: b = &c->d_SC; used = 72; break;
        case XTB_BUF_HBLK: b = &c->d_Hblk; used = 8 * nb2 * (size_t)c->n_blocks; break;
        case XTB_BUF_SBLK: b = &c->d_Sblk; used = 8 * nb2 * (size_t)c->n_blocks; break;
        case XTB_BUF_FREP: b = &c->d_Frep; used = 24 * (size_t)c->na; break;
        case XTB_BUF_MIJ: b = &c->d_cpl_M; used = 8 * N * N; break;
        case XTB_BUF_FWALL: b = &c->d_Fwall; used = 24 * (size_t)c->na; break;
        case XTB_BUF_FVDW: b = &c->d_Fvdw; used = 24 * (size_t)c->na; break;
        case XTB_BUF_FCOUL: b = &c->d_Fcoul; used = 24 * (size_t)c->na; break;
        case XTB_BUF_POS: b = &c->d_S; used = 24 * (size_t)c->na; break;
        case XTB_BUF_VEL: b = &c->d_SV; used = 24 * (size_t)c->na; break;
        case XTB_BUF_EI: if (!c->d_Ei) XTB_FAIL(c, XTB_ESTATE, "xtb_device_buffer: no eigenvalues"); *dptr = (void*)c->d_Ei; if (bytes) *bytes = 8 * N; return XTB_OK;
        case XTB_BUF_FE: if (!c->d_fe) XTB_FAIL(c, XTB_ESTATE, "xtb_device_buffer: no occupations"); *dptr = (void*)c->d_fe; if (bytes) *bytes = 8 * N; return XTB_OK;
        default: XTB_FAIL(c, XTB_EINVAL, "xtb_device_buffer: unknown buffer %d", which);
    }
    if (!b->p) XTB_FAIL(c, XTB_ESTATE, "xtb_device_buffer: buffer %d not produced yet", which);
    *dptr = b->p;
    if (bytes) *bytes = used;
    return XTB_OK;
}

long long xtb_counter(const xtb_ctx* c, int which) {
    if (!c) return -1;
    switch (which) {
        case 0: return c->na;
        case 1: return c->n_orb;
        case 2: return (long long)c->na * c->n_orb;
        case 3: return c->max_nn_seen;
        case 4: return c->n_blocks;
        case 5: return c->n_pairs;
        case 6: return c->launches;
        case 7: return c->k_kept;
        case 8: return c->cap;
        case 9: return c->density_slotcap;
        case 10: return c->density_stages;
        case 11: {      // work units of the last pair-kernel launch (pairs + singles); synchronises the stream
            if (c->density_impl != 2 || !c->d_units.p || c->n_orb != 9) return -1;
            int n = -1;
            cudaStreamSynchronize(c->stream);
            if (cudaMemcpy(&n, (const int*)c->d_units.p + 2 * (size_t)(c->a1 - c->a0), sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
            return n;
        }
        case 16: return c->graph_captures;
        case 17: return c->density_tm;
        case 12: return (long long)c->xchg_bytes[XTB_XCHG_HS];
        case 13: return (long long)c->xchg_bytes[XTB_XCHG_RHO];
        case 14: return c->halo_send_blocks;
        case 15: return c->halo_recv_blocks;
        default: return -1;
    }
}

double xtb_stage_ms(const xtb_ctx* c, int stage) {
    if (!c || stage < 0 || stage > 23) return -1.0;
    return c->stage_ms[stage];
}

}  // extern "C"
