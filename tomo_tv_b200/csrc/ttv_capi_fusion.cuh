// C ABI of the multi-modal fusion engine (include/tomo_tv_b200.h, ttv_mm_*): one entry point per bound
// method of the reference's pybind11 class `multimodal` (T/chemistry/utils/multimodal.cpp:520-564).
// Included at the end of ttv_capi.cu (same translation unit, same error convention).
#pragma once
#include "ttv_fusion.cuh"

struct ttv_multimodal { ttv::MultiModal* m; };

static ttv::MultiModal& MM(ttv_mm_handle h) {
    if (!h || !h->m) throw ttv::CudaError("null multimodal handle");
    h->m->use();
    return *h->m;
}

extern "C" {

int ttv_mm_create(ttv_mm_handle* out, int ns, int nray, int nel, const float* ha, int nh, const float* ca, int nc, int device) {
    return ttv_mm_create_slab(out, ns, nray, nel, ha, nh, ca, nc, device, 0, ns);
}
int ttv_mm_create_slab(ttv_mm_handle* out, int ns, int nray, int nel, const float* ha, int nh, const float* ca, int nc, int device, int x_offset, int ns_global) {
    TTV_TRY
    *out = nullptr;
    ttv_multimodal* h = new ttv_multimodal{nullptr};
    try { h->m = new MultiModal(ns, nray, nel, ha, nh, ca, nc, device, x_offset, ns_global); }
    catch (...) { delete h; throw; }
    *out = h;
    TTV_CATCH
}
int ttv_mm_destroy(ttv_mm_handle h) { TTV_TRY if (h) { delete h->m; delete h; } TTV_CATCH }

int ttv_mm_initialize_initial_volume(ttv_mm_handle h) {
    TTV_TRY MultiModal& m = MM(h); for (int e = 0; e < m.nel; ++e) m.orig[e].alloc(m.eh->vol_elems()); TTV_CATCH
}
int ttv_mm_set_haadf_tilt_series(ttv_mm_handle h, const float* bh, int rows, int cols) {
    TTV_TRY
    MultiModal& m = MM(h);
    if (rows != m.nx || cols != m.nh * m.N) throw CudaError("set_haadf_tilt_series: expected a [Nslice, NprojHaadf*Nray] matrix");
    m.eh->upload(m.eh->b, bh, (int64_t)m.nh * m.N);
    TTV_CATCH
}
int ttv_mm_set_chem_tilt_series(ttv_mm_handle h, const float* bc, int rows, int cols) {
    TTV_TRY
    MultiModal& m = MM(h);
    const int64_t Mc = (int64_t)m.nc * m.N;
    if (rows != m.nx || cols != Mc * m.nel) throw CudaError("set_chem_tilt_series: expected a [Nslice, Nel*NprojChem*Nray] matrix");
    for (int e = 0; e < m.nel; ++e) m.ec->upload(m.bc[e], bc + (size_t)e * Mc, Mc, Mc * m.nel);
    TTV_CATCH
}
int ttv_mm_set_measureHaadf(ttv_mm_handle h, int f) { TTV_TRY MM(h).measureHaadf = f != 0; TTV_CATCH }
int ttv_mm_set_measureChem(ttv_mm_handle h, int f) { TTV_TRY MM(h).measureChem = f != 0; TTV_CATCH }
int ttv_mm_measureHaadf(ttv_mm_handle h, int* f) { TTV_TRY *f = MM(h).measureHaadf ? 1 : 0; TTV_CATCH }
int ttv_mm_measureChem(ttv_mm_handle h, int* f) { TTV_TRY *f = MM(h).measureChem ? 1 : 0; TTV_CATCH }
int ttv_mm_set_gamma(ttv_mm_handle h, float g) { TTV_TRY MM(h).gamma = g; TTV_CATCH }
int ttv_mm_gamma(ttv_mm_handle h, float* g) { TTV_TRY *g = MM(h).gamma; TTV_CATCH }
int ttv_mm_set_recon(ttv_mm_handle h, const float* s, int e, int i) { TTV_TRY MultiModal& m = MM(h); m.check_el(e); m.eh->set_slice(m.X[e], s, i); TTV_CATCH }
int ttv_mm_get_recon(ttv_mm_handle h, float* out, int e, int i) { TTV_TRY MultiModal& m = MM(h); m.check_el(e); m.eh->get_slice(m.X[e], out, i); TTV_CATCH }
int ttv_mm_set_original_volume(ttv_mm_handle h, const float* s, int e, int i) {
    TTV_TRY MultiModal& m = MM(h); m.check_el(e); if (!m.orig[e].ok()) m.orig[e].alloc(m.eh->vol_elems()); m.eh->set_slice(m.orig[e], s, i); TTV_CATCH
}
int ttv_mm_get_recon_all(ttv_mm_handle h, float* out, int e) { TTV_TRY MultiModal& m = MM(h); m.check_el(e); m.eh->download(m.X[e], out, (int64_t)m.N * m.N); TTV_CATCH }
int ttv_mm_set_recon_all(ttv_mm_handle h, const float* in, int e) { TTV_TRY MultiModal& m = MM(h); m.check_el(e); m.eh->upload(m.X[e], in, (int64_t)m.N * m.N); TTV_CATCH }
int ttv_mm_get_gpu_id(ttv_mm_handle h, int* id) { TTV_TRY *id = MM(h).gpuID; TTV_CATCH }
int ttv_mm_set_gpu(ttv_mm_handle h, int id) { TTV_TRY MM(h).gpuID = id; TTV_CATCH }
int ttv_mm_load_sigma(ttv_mm_handle h, const float* t, int nnz) { TTV_TRY MM(h).load_sigma(t, nnz); TTV_CATCH }
int ttv_mm_initialize_FP(ttv_mm_handle h) { TTV_TRY MM(h); TTV_CATCH }
int ttv_mm_initialize_BP(ttv_mm_handle h) { TTV_TRY MM(h); TTV_CATCH }
int ttv_mm_initialize_SIRT(ttv_mm_handle h) { TTV_TRY MM(h).eh->compute_weights(); TTV_CATCH }
int ttv_mm_initialize_SART(ttv_mm_handle h, const char* order) {
    TTV_TRY
    MultiModal& m = MM(h);
    m.eh->projOrder = order ? order : "sequential";
    m.eh->compute_weights(); m.eh->sart_ready = true; m.sart_ready = true;
    TTV_CATCH
}
int ttv_mm_forward_projection(ttv_mm_handle h, const float* s, float* out) { TTV_TRY MM(h).forward_projection_2d(s, out); TTV_CATCH }
int ttv_mm_estimate_lipschitz(ttv_mm_handle h) { TTV_TRY MM(h).estimate_lipschitz(); TTV_CATCH }
int ttv_mm_rescale_projections(ttv_mm_handle h) { TTV_TRY MM(h).rescale_projections(); TTV_CATCH }
int ttv_mm_rescale_tomograms(ttv_mm_handle h, float s) { TTV_TRY MM(h).rescale_tomograms(s); TTV_CATCH }
int ttv_mm_poisson_ml(ttv_mm_handle h, float lam, float* cost) { TTV_TRY *cost = MM(h).poisson_ml(lam); TTV_CATCH }
int ttv_mm_data_fusion(ttv_mm_handle h, float lh, float lc, int niter, const char* method, float* ch, float* cc) {
    TTV_TRY
    const std::string md = method ? method : "SIRT";
    MM(h).data_fusion(lh, lc, niter, md != "SIRT", ch, cc);     // anything but "SIRT" is SART (multimodal.cpp:431-432)
    TTV_CATCH
}
int ttv_mm_sirt_data_fusion(ttv_mm_handle h, float lh, float lc, int niter, float* ch, float* cc) { return ttv_mm_data_fusion(h, lh, lc, niter, "SIRT", ch, cc); }
int ttv_mm_sart_data_fusion(ttv_mm_handle h, float lh, float lc, float* ch, float* cc) { return ttv_mm_data_fusion(h, lh, lc, 1, "SART", ch, cc); }
int ttv_mm_chemical_SIRT(ttv_mm_handle h, int n) { TTV_TRY MM(h).chemical_recon(n, false); TTV_CATCH }
int ttv_mm_chemical_SART(ttv_mm_handle h, int n) { TTV_TRY MM(h).chemical_recon(n, true); TTV_CATCH }
int ttv_mm_data_distance(ttv_mm_handle h, float* v) { TTV_TRY *v = MM(h).data_distance(); TTV_CATCH }
int ttv_mm_tv_gd(ttv_mm_handle h, int ng, float d, float* tv) { TTV_TRY *tv = MM(h).tv_4d(ng, d, false); TTV_CATCH }
int ttv_mm_tv_fgp_4D(ttv_mm_handle h, int ng, float lam, float* tv) { TTV_TRY *tv = MM(h).tv_4d(ng, lam, true); TTV_CATCH }
int ttv_mm_rmse(ttv_mm_handle h, float* out) { TTV_TRY MM(h).rmse(out); TTV_CATCH }
int ttv_mm_get_haadf_projections(ttv_mm_handle h, float* out) { TTV_TRY MultiModal& m = MM(h); m.eh->download(m.eh->b, out, (int64_t)m.nh * m.N); TTV_CATCH }
int ttv_mm_get_model_projections(ttv_mm_handle h, float* out) { TTV_TRY MultiModal& m = MM(h); m.eh->download(m.eh->g, out, (int64_t)m.nh * m.N); TTV_CATCH }
int ttv_mm_get_chem_projections(ttv_mm_handle h, float* out) {
    TTV_TRY
    MultiModal& m = MM(h);
    const int64_t Mc = (int64_t)m.nc * m.N;
    for (int e = 0; e < m.nel; ++e) m.ec->download(m.bc[e], out + (size_t)e * Mc, Mc, Mc * m.nel);
    TTV_CATCH
}
int ttv_mm_restart_recon(ttv_mm_handle h) { TTV_TRY MM(h).restart_recon(); TTV_CATCH }
int ttv_mm_launch_count(ttv_mm_handle h, long long* n) { TTV_TRY *n = MM(h).launches(); TTV_CATCH }
int ttv_mm_timer_start(ttv_mm_handle h) { TTV_TRY MM(h).eh->timer_start(); TTV_CATCH }
int ttv_mm_timer_stop(ttv_mm_handle h, float* ms) { TTV_TRY *ms = MM(h).eh->timer_stop(); TTV_CATCH }
int ttv_mm_comm_init(ttv_mm_handle h, const void* id128, int rank, int nranks) {
    TTV_TRY
    MultiModal& m = MM(h);
    Nccl& n = Nccl::get();
    Nccl::uid_t id; std::memcpy(&id, id128, 128);
    n.check(n.CommInitRank(&m.eh->comm, nranks, id, rank), "ncclCommInitRank");
    m.eh->rank = rank; m.eh->nranks = nranks;
    // the chemical-geometry engine shares the communicator and the stream; it must not destroy it twice
    m.ec->comm = m.eh->comm; m.ec->rank = rank; m.ec->nranks = nranks; m.ec->own_comm = false;
    TTV_CATCH
}

}  // extern "C"
