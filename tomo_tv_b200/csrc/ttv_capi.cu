// C ABI of the engine (include/tomo_tv_b200.h).  Every entry point mirrors one bound method of the
// reference's pybind11 class (T/gpu/utils/tomoengine.cpp:487-534); errors become status codes +
// ttv_last_error() instead of the reference's process abort.
#include "../../include/tomo_tv_b200.h"
#include "ttv_engine.cuh"
#include "ttv_sart.cuh"
#include "ttv_fbp.cuh"

using namespace ttv;

struct ttv_engine { Engine* e; };

static thread_local std::string g_err;

#define TTV_TRY try {
#define TTV_CATCH                                                      \
    return 0; }                                                        \
    catch (const std::exception& ex) { g_err = ex.what(); return 1; } \
    catch (...) { g_err = "unknown error"; return 2; }

static Engine& E(ttv_handle h) {
    if (!h || !h->e) throw CudaError("null engine handle");
    h->e->use();
    return *h->e;
}

extern "C" {

const char* ttv_last_error(void) { return g_err.c_str(); }
int ttv_version(void) { return 100; }

int ttv_device_count(int* count) {
    TTV_TRY
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess) { cudaGetLastError(); c = 0; }
    *count = c;
    TTV_CATCH
}

int ttv_create(ttv_handle* out, int nslice, int nray, const float* angles, int nangles, int device) {
    return ttv_create_slab(out, nslice, nray, angles, nangles, device, 0, nslice);
}
int ttv_create_slab(ttv_handle* out, int nslice_local, int nray, const float* angles, int nangles, int device, int x_offset, int nslice_global) {
    TTV_TRY
    *out = nullptr;
    ttv_engine* h = new ttv_engine{nullptr};
    try { h->e = new Engine(nslice_local, nray, angles, nangles, device, x_offset, nslice_global); }
    catch (...) { delete h; throw; }
    *out = h;
    TTV_CATCH
}
int ttv_destroy(ttv_handle h) {
    TTV_TRY
    if (h) { delete h->e; delete h; }
    TTV_CATCH
}

// ------------------------------------------------------------------ data
int ttv_set_tilt_series(ttv_handle h, const float* b, int rows, int cols) {
    TTV_TRY
    Engine& e = E(h); e.need_angles();
    if (rows != e.nx || cols != e.Na * e.N) throw CudaError("set_tilt_series: expected a [Nslice, Nproj*Nray] matrix");
    e.upload(e.b, b, (int64_t)e.Na * e.N);
    TTV_CATCH
}
int ttv_set_model_projections(ttv_handle h, const float* g, int rows, int cols) {
    TTV_TRY
    Engine& e = E(h); e.need_angles();
    if (rows != e.nx || cols != e.Na * e.N) throw CudaError("set_model_projections: expected a [Nslice, Nproj*Nray] matrix");
    e.upload(e.g, g, (int64_t)e.Na * e.N);
    TTV_CATCH
}
int ttv_get_projections(ttv_handle h, float* out) { TTV_TRY Engine& e = E(h); e.need_angles(); e.download(e.b, out, (int64_t)e.Na * e.N); TTV_CATCH }
int ttv_get_model_projections(ttv_handle h, float* out) { TTV_TRY Engine& e = E(h); e.need_angles(); e.download(e.g, out, (int64_t)e.Na * e.N); TTV_CATCH }
int ttv_initialize_initial_volume(ttv_handle h) { TTV_TRY Engine& e = E(h); e.original.alloc(e.vol_elems()); TTV_CATCH }
int ttv_initialize_recon_copy(ttv_handle h) { TTV_TRY Engine& e = E(h); e.temp.alloc(e.vol_elems()); TTV_CATCH }
int ttv_set_original_volume(ttv_handle h, const float* s, int i) { TTV_TRY Engine& e = E(h); if (!e.original.ok()) e.original.alloc(e.vol_elems()); e.set_slice(e.original, s, i); TTV_CATCH }
int ttv_set_recon(ttv_handle h, const float* s, int i) { TTV_TRY Engine& e = E(h); e.set_slice(e.recon, s, i); TTV_CATCH }
int ttv_get_recon(ttv_handle h, float* out, int i) { TTV_TRY Engine& e = E(h); e.get_slice(e.recon, out, i); TTV_CATCH }
int ttv_get_recon_all(ttv_handle h, float* out) { TTV_TRY Engine& e = E(h); e.download(e.recon, out, (int64_t)e.N * e.N); TTV_CATCH }
int ttv_set_recon_all(ttv_handle h, const float* in) { TTV_TRY Engine& e = E(h); e.upload(e.recon, in, (int64_t)e.N * e.N); TTV_CATCH }
int ttv_set_original_volume_all(ttv_handle h, const float* in) {
    TTV_TRY Engine& e = E(h); if (!e.original.ok()) e.original.alloc(e.vol_elems()); e.upload(e.original, in, (int64_t)e.N * e.N); TTV_CATCH
}
int ttv_update_projection_angles(ttv_handle h, const float* a, int n) { TTV_TRY Engine& e = E(h); e.set_angles(a, n); TTV_CATCH }
int ttv_create_projections(ttv_handle h) {
    TTV_TRY
    Engine& e = E(h);
    if (!e.original.ok()) throw CudaError("create_projections: original volume not set");
    e.fp<FP_STORE>(e.original.p, e.b.p, nullptr, nullptr);
    e.sync();
    TTV_CATCH
}
int ttv_restart_recon(ttv_handle h) {
    TTV_TRY
    Engine& e = E(h);
    TTV_CUDA(cudaMemsetAsync(e.recon.p, 0, e.vol_elems() * sizeof(float), e.st));
    if (e.momentum) {
        TTV_CUDA(cudaMemsetAsync(e.yk.p, 0, e.vol_elems() * sizeof(float), e.st));
        TTV_CUDA(cudaMemsetAsync(e.recon_old.p, 0, e.vol_elems() * sizeof(float), e.st));
    }
    e.sync();
    TTV_CATCH
}
int ttv_get_gpu_id(ttv_handle h, int* id) { TTV_TRY *id = E(h).gpuID; TTV_CATCH }
int ttv_set_gpu(ttv_handle h, int id) { TTV_TRY E(h).gpuID = id; TTV_CATCH }   /* only records the index (tomoengine.cpp:87-89) */

// b <- Poisson(b / sum(b) * Nc * size) / (Nc*size) * sum(b)     (tomoengine.cpp:471-484); host-side simulation helper
int ttv_poisson_noise(ttv_handle h, int Nc) {
    TTV_TRY
    Engine& e = E(h); e.need_angles();
    const size_t M = (size_t)e.Na * e.N, n = (size_t)e.nx * M;
    std::vector<float> hb(n);
    e.download(e.b, hb.data(), (int64_t)M);
    double sum = 0; for (float v : hb) sum += v;
    sum = e.allreduce_sum(sum);
    const double tot = (double)Nc * (double)((size_t)e.nxg * M);
    std::default_random_engine gen((unsigned)(1 + e.rank));
    for (size_t i = 0; i < n; ++i) {
        const double mean = hb[i] / sum * tot;
        std::poisson_distribution<int> dist(mean > 0 ? mean : 0.0);
        hb[i] = (float)(dist(gen) / tot * sum);
    }
    e.upload(e.b, hb.data(), (int64_t)M);
    TTV_CATCH
}

// ------------------------------------------------------------------ algorithms
int ttv_initialize_SIRT(ttv_handle h) { TTV_TRY E(h).compute_weights(); TTV_CATCH }
int ttv_SIRT(ttv_handle h, int niter) {
    TTV_TRY
    Engine& e = E(h);
    if (!e.weights_ready) throw CudaError("SIRT called before initialize_SIRT");
    float* tgt = e.momentum ? e.yk.p : e.recon.p;           // tomoengine.cpp:194-195
    for (int i = 0; i < niter; ++i) e.sirt_step(tgt);
    /* stream-ordered: void methods do not block the host (every getter / scalar method synchronises) */
    TTV_CATCH
}
int ttv_initialize_SART(ttv_handle h, const char* order) {
    TTV_TRY
    Engine& e = E(h);
    e.projOrder = order ? order : "sequential";
    e.compute_weights();
    e.sart_ready = true;
    TTV_CATCH
}
int ttv_SART(ttv_handle h, float beta, int niter) { TTV_TRY E(h).sart_pass(beta, niter); TTV_CATCH }
int ttv_initialize_FP(ttv_handle h) { TTV_TRY E(h).need_angles(); E(h).fp_ready = true; TTV_CATCH }
int ttv_initialize_BP(ttv_handle h) { TTV_TRY E(h).need_angles(); TTV_CATCH }

int ttv_initialize_poisson_ML(ttv_handle h) {
    TTV_TRY
    Engine& e = E(h);
    e.compute_weights();
    e.L_Aml = e.L_A;                                         // max(BP(FP(1))), tomoengine.cpp:241-242
    const float m = e.max_of(e.b.p, e.sino_elems());        // b /= max(b) if max(b) > 1, :245
    if (m > 1.0f) {
        div_kernel<<<ew_blocks(e.sino_elems() / 4), 256, 0, e.st>>>(reinterpret_cast<float4*>(e.b.p), m, e.sino_elems() / 4);
        ++e.launches;
    }
    e.kl_ready = true;
    e.sync();
    TTV_CATCH
}
int ttv_poisson_ML(ttv_handle h, float lambda, float* cost) {
    TTV_TRY
    Engine& e = E(h);
    if (!e.kl_ready) throw CudaError("poisson_ML called before initialize_poisson_ML");
    if (!e.r.ok()) e.r.alloc(e.sino_elems());
    e.zero_red(3);
    e.fp<FP_KL>(e.recon.p, e.r.p, e.b.p, e.d_red + 3, -1, 1e-1f);
    e.bp<BP_STEP_UPD>(e.r.p, e.recon.p, lambda / e.L_Aml);
    *cost = (float)e.read_red(3);
    TTV_CATCH
}
int ttv_forward_projection(ttv_handle h) { TTV_TRY Engine& e = E(h); e.fp<FP_STORE>(e.recon.p, e.g.p, nullptr, nullptr); e.sync(); TTV_CATCH }
int ttv_back_projection(ttv_handle h) { TTV_TRY Engine& e = E(h); e.bp<BP_STORE>(e.g.p, e.recon.p); e.sync(); TTV_CATCH }

int ttv_initialize_fista(ttv_handle h) {
    TTV_TRY
    Engine& e = E(h);
    e.momentum = true;
    e.yk.alloc(e.vol_elems()); e.recon_old.alloc(e.vol_elems());
    TTV_CUDA(cudaMemcpyAsync(e.yk.p, e.recon.p, e.vol_elems() * sizeof(float), cudaMemcpyDeviceToDevice, e.st));
    TTV_CUDA(cudaMemcpyAsync(e.recon_old.p, e.recon.p, e.vol_elems() * sizeof(float), cudaMemcpyDeviceToDevice, e.st));
    e.compute_weights();                                     // initializeSIRT + L_A (tomoengine.cpp:364-371)
    e.sync();
    TTV_CATCH
}
int ttv_get_lipschitz(ttv_handle h, float* L) { TTV_TRY *L = E(h).L_A; TTV_CATCH }
int ttv_remove_momentum(ttv_handle h) { TTV_TRY E(h).momentum = false; TTV_CATCH }
int ttv_fista_momentum(ttv_handle h, float beta) {
    TTV_TRY
    Engine& e = E(h);
    if (!e.yk.ok()) throw CudaError("fista_momentum called before initialize_fista");
    const size_t n4 = e.vol_elems() / 4;
    {
        Engine::ProfScope ps(e, PF_MOMENTUM);
        momentum_kernel<<<ew_blocks(n4), 256, 0, e.st>>>(reinterpret_cast<float4*>(e.recon.p), reinterpret_cast<float4*>(e.yk.p),
                                                          reinterpret_cast<float4*>(e.recon_old.p), beta, n4);
        std::swap(e.yk.p, e.recon_old.p);                    // y was written over recon_old; the old yk buffer holds x
    }
    ++e.launches; TTV_LAUNCH_CHECK();
    TTV_CATCH
}

// ------------------------------------------------------------------ regularisers & reductions
int ttv_tv_fgp(ttv_handle h, int ng, float lambda, float* tv) { TTV_TRY *tv = E(h).tv_fgp(ng, lambda); TTV_CATCH }
int ttv_tv_gd(ttv_handle h, int ng, float dPOCS, float* tv) { TTV_TRY *tv = E(h).tv_gd(ng, dPOCS); TTV_CATCH }
int ttv_tv(ttv_handle h, float* tv) { TTV_TRY Engine& e = E(h); *tv = (float)e.tv_value(e.recon.p); TTV_CATCH }
int ttv_original_tv(ttv_handle h, float* tv) {
    TTV_TRY Engine& e = E(h); if (!e.original.ok()) throw CudaError("original volume not set"); *tv = (float)e.tv_value(e.original.p); TTV_CATCH
}
int ttv_soft_threshold(ttv_handle h, float lambda) {
    TTV_TRY
    Engine& e = E(h);
    float* tgt = e.momentum ? e.yk.p : e.recon.p;           // tomoengine.cpp:436-438 (threshold, then positivity)
    ew_kernel<EW_SOFT_POS><<<ew_blocks(e.vol_elems() / 4), 256, 0, e.st>>>(reinterpret_cast<float4*>(tgt), lambda, e.vol_elems() / 4);
    ++e.launches; TTV_LAUNCH_CHECK(); e.sync();
    TTV_CATCH
}
int ttv_l1_norm(ttv_handle h, float* v) { TTV_TRY Engine& e = E(h); *v = (float)e.reduce<RED_L1>(e.recon.p, nullptr, e.vol_elems()); TTV_CATCH }
int ttv_copy_recon(ttv_handle h) {
    TTV_TRY
    Engine& e = E(h);
    if (!e.temp.ok()) throw CudaError("copy_recon called before initialize_recon_copy");
    TTV_CUDA(cudaMemcpyAsync(e.temp.p, e.recon.p, e.vol_elems() * sizeof(float), cudaMemcpyDeviceToDevice, e.st));
    TTV_CATCH
}
int ttv_matrix_2norm(ttv_handle h, float* v) {
    TTV_TRY
    Engine& e = E(h);
    if (!e.temp.ok()) throw CudaError("matrix_2norm called before initialize_recon_copy");
    *v = (float)std::sqrt(e.reduce<RED_SQDIST>(e.recon.p, e.temp.p, e.vol_elems()));
    TTV_CATCH
}
int ttv_data_distance(ttv_handle h, float* v) {
    TTV_TRY
    Engine& e = E(h);
    e.zero_red(4);
    e.fp<FP_STORE_DIST>(e.recon.p, e.g.p, e.b.p, e.d_red + 4);      // g = FP(recon) and ||g-b||^2 in one pass
    *v = (float)std::sqrt(e.read_red(4));
    TTV_CATCH
}
int ttv_rmse(ttv_handle h, float* v) {
    TTV_TRY
    Engine& e = E(h);
    if (!e.original.ok()) throw CudaError("rmse: original volume not set");
    const double s = e.reduce<RED_SQDIST>(e.recon.p, e.original.p, e.vol_elems());
    *v = (float)std::sqrt(s / ((double)e.nxg * e.N * e.N));
    TTV_CATCH
}

// ------------------------------------------------------------------ CGLS / FBP (ttv_fbp.cuh)
int ttv_initialize_CGLS(ttv_handle h) { TTV_TRY E(h).need_angles(); TTV_CATCH }
int ttv_CGLS(ttv_handle h, int niter) { TTV_TRY cgls_run(E(h), niter); TTV_CATCH }
int ttv_initialize_FBP(ttv_handle h, const char* filter) { TTV_TRY Engine& e = E(h); e.need_angles(); e.fbpFilter = filter ? filter : "ram-lak"; TTV_CATCH }
int ttv_FBP(ttv_handle h, int apply_positivity) { TTV_TRY fbp_run(E(h), apply_positivity != 0); TTV_CATCH }

// ------------------------------------------------------------------ multi-GPU
int ttv_nccl_unique_id(void* id128) {
    TTV_TRY
    Nccl& n = Nccl::get();
    Nccl::uid_t id;
    n.check(n.GetUniqueId(&id), "ncclGetUniqueId");
    std::memcpy(id128, &id, 128);
    TTV_CATCH
}
int ttv_comm_init(ttv_handle h, const void* id128, int rank, int nranks) {
    TTV_TRY
    Engine& e = E(h);
    Nccl& n = Nccl::get();
    Nccl::uid_t id; std::memcpy(&id, id128, 128);
    n.check(n.CommInitRank(&e.comm, nranks, id, rank), "ncclCommInitRank");
    e.rank = rank; e.nranks = nranks;
    if (nranks > 1) e.prealloc_tv();
    TTV_CATCH
}

int ttv_p2p_export(ttv_handle h, void* handle64) { TTV_TRY E(h).p2p_export(handle64); TTV_CATCH }
int ttv_p2p_attach(ttv_handle h, const void* left64, const void* right64) { TTV_TRY E(h).p2p_attach(left64, right64); TTV_CATCH }

int ttv_p2p_attach_local(ttv_handle h, ttv_handle left, ttv_handle right) {
    TTV_TRY
    E(h).p2p_attach_local(left ? left->e : nullptr, right ? right->e : nullptr);
    TTV_CATCH
}
int ttv_set_rank(ttv_handle h, int rank, int nranks) { TTV_TRY Engine& e = E(h); e.rank = rank; e.nranks = nranks; TTV_CATCH }

// ------------------------------------------------------------------ measurement / test hooks
int ttv_sync(ttv_handle h) { TTV_TRY E(h).sync(); TTV_CATCH }
int ttv_get_trig(ttv_handle h, float* cosv, float* sinv) {
    TTV_TRY
    Engine& e = E(h); e.need_angles();
    float *dc, *ds;
    TTV_CUDA(cudaMalloc(&dc, e.Na * sizeof(float))); TTV_CUDA(cudaMalloc(&ds, e.Na * sizeof(float)));
    trig_kernel<<<(e.Na + 127) / 128, 128, 0, e.st>>>(e.d_theta, dc, ds, e.Na); ++e.launches;
    TTV_CUDA(cudaMemcpyAsync(cosv, dc, e.Na * sizeof(float), cudaMemcpyDeviceToHost, e.st));
    TTV_CUDA(cudaMemcpyAsync(sinv, ds, e.Na * sizeof(float), cudaMemcpyDeviceToHost, e.st));
    e.sync(); cudaFree(dc); cudaFree(ds);
    TTV_CATCH
}
int ttv_get_sirt_weights(ttv_handle h, float* R, float* C) {
    TTV_TRY
    Engine& e = E(h); e.compute_weights();
    TTV_CUDA(cudaMemcpyAsync(R, e.R.p, (size_t)e.Na * e.N * sizeof(float), cudaMemcpyDeviceToHost, e.st));
    TTV_CUDA(cudaMemcpyAsync(C, e.C.p, (size_t)e.N * e.N * sizeof(float), cudaMemcpyDeviceToHost, e.st));
    e.sync();
    TTV_CATCH
}
int ttv_profile(ttv_handle h, int enable) { TTV_TRY E(h).prof_enable(enable != 0); TTV_CATCH }
int ttv_profile_get(ttv_handle h, const char* name, double* total_ms, long long* count) {
    TTV_TRY
    Engine& e = E(h);
    e.prof_resolve();
    for (int i = 0; i < PF_COUNT; ++i)
        if (std::string(prof_name(i)) == name) { *total_ms = e.prof_ms[i]; *count = e.prof_n[i]; return 0; }
    throw CudaError(std::string("ttv_profile_get: unknown kernel group '") + name + "'");
    TTV_CATCH
}
int ttv_timer_start(ttv_handle h) { TTV_TRY E(h).timer_start(); TTV_CATCH }
int ttv_timer_stop(ttv_handle h, float* ms) { TTV_TRY *ms = E(h).timer_stop(); TTV_CATCH }
int ttv_set_option(ttv_handle h, const char* name, int value) {
    TTV_TRY
    Engine& e = E(h);
    const std::string nm = name ? name : "";
    if (nm == "fgp_tiled") e.fgp_tiled = value != 0;
    else if (nm == "fgp_fuse") e.fgp_fuse = value != 0;
    else if (nm == "sart_cluster") { e.sart_cluster = value != 0; e.sart_cluster_force = value == 2; }
    else if (nm == "tvgd_march") e.tvgd_march = value != 0;
    else if (nm == "fp_variant") {
        if (value < 0 || value > 2) throw CudaError("ttv_set_option: fp_variant must be 0 (by slab thickness), 1 (classic) or 2 (table-driven)");
        e.fp_variant = value;
    }
    else if (nm == "bp_variant") {
        if (value < 0 || value > 2) throw CudaError("ttv_set_option: bp_variant must be 0 (by slab thickness), 1 (classic) or 2 (table-driven)");
        e.bp_variant = value;
    }
    else if (nm == "fp_mix_n2") e.fp_mix_n2 = value;
    else if (nm == "fp_xb") {
        if (value < 0 || value > 3) throw CudaError("ttv_set_option: fp_xb must be 0 (by slab thickness and wave count), 1, 2 or 3 (both CTA sizes in one launch)");
        e.fp_xb = value;
    }
    else if (nm == "bp_xb") {
        if (value != 0 && value != 2 && value != 8) throw CudaError("ttv_set_option: bp_xb must be 0 (by slab thickness), 2 or 8");
        e.bp_xb = value;
    }
    else throw CudaError("ttv_set_option: unknown option '" + nm + "'");
    TTV_CATCH
}
int ttv_selftest_invsqrt(ttv_handle h, unsigned long long n, unsigned seed, unsigned long long* mismatches) {
    TTV_TRY
    Engine& e = E(h);
    e.use();
    unsigned long long* d = nullptr;
    TTV_CUDA(cudaMalloc(&d, sizeof(unsigned long long)));
    TTV_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long), e.st));
    invsqrt_selftest_kernel<<<592, 256, 0, e.st>>>(n, seed, d); ++e.launches;
    TTV_LAUNCH_CHECK();
    TTV_CUDA(cudaMemcpyAsync(mismatches, d, sizeof(unsigned long long), cudaMemcpyDeviceToHost, e.st));
    e.sync();
    cudaFree(d);
    TTV_CATCH
}
int ttv_launch_count(ttv_handle h, long long* n) { TTV_TRY *n = E(h).launches; TTV_CATCH }
int ttv_sizes(ttv_handle h, long long* V, long long* S, long long* samples) {
    TTV_TRY
    Engine& e = E(h);
    *V = (long long)e.nx * e.N * e.N * 4;
    *S = (long long)e.nx * e.Na * e.N * 4;
    *samples = (long long)e.nx * e.Na * e.N * e.N;
    TTV_CATCH
}

int ttv_time_kernel(ttv_handle h, const char* name, int reps, int arg_i, float arg_f, float* ms) {
    TTV_TRY
    Engine& e = E(h);
    const std::string nm = name ? name : "";
    const size_t n = e.vol_elems();
    if (!e.r.ok() && e.Na > 0) e.r.alloc(e.sino_elems());
    if (!e.scratch.ok()) e.scratch.alloc(n);
    for (int i = 0; i < 6; ++i) if (!e.P[i].ok()) e.P[i].alloc(n);
    if (!e.yk.ok()) e.yk.alloc(n);
    if (!e.recon_old.ok()) e.recon_old.alloc(n);
    if (e.Na > 0) e.compute_weights();
    FgpParams fp; fp.in = e.recon.p; fp.P1 = e.P[0].p; fp.P2 = e.P[1].p; fp.P3 = e.P[2].p; fp.Q1 = e.P[3].p; fp.Q2 = e.P[4].p; fp.Q3 = e.P[5].p;
    fp.compact = 0; fp.pitch = 0; fp.rm_gl1 = fp.rm_gr1 = fp.rm_gr2 = fp.rm_gr3 = nullptr; fp.rm_flagR = fp.rm_flagL = nullptr; fp.cnt = nullptr; fp.epoch = 0;
    fp.hl_P1 = fp.hr_in = fp.hr_P1 = fp.hr_P2 = fp.hr_P3 = nullptr; fp.lambda = 0.1f; fp.multip = 1.0f / 2.6f; fp.d = e.dims();
    dim3 gsw((e.N + TV_KSEG - 1) / TV_KSEG, (e.N + TV_WARPS - 1) / TV_WARPS, e.nxb);
    GdParams gp; gp.x = e.recon.p; gp.g = e.scratch.p; gp.hm = gp.hp = nullptr; gp.sumsq = e.d_red + 1; gp.d = e.dims();
    const int64_t rows = (int64_t)e.nxb * e.N;
    auto one = [&]() {
        if (nm == "fp") e.fp<FP_STORE>(e.recon.p, e.g.p, nullptr, nullptr);
        else if (nm == "fp_sirt") e.fp<FP_SIRT_RES>(e.recon.p, e.r.p, e.b.p, nullptr);
        else if (nm == "bp") e.bp<BP_STORE>(e.g.p, e.scratch.p);
        else if (nm == "bp_sirt") e.bp<BP_SIRT_UPD>(e.r.p, e.yk.p);
        else if (nm == "sirt") e.sirt_step(e.yk.p);
        else if (nm == "fgp_sweep_v1") { fgp_sweep_kernel<false><<<gsw, TV_WARPS * 32, 0, e.st>>>(fp); ++e.launches; }
        else if (nm == "fgp_first_v1") { fgp_sweep_kernel<true><<<gsw, TV_WARPS * 32, 0, e.st>>>(fp); ++e.launches; }
        else if (nm == "fgp_final_v1") { fgp_final_kernel<<<dim3(gsw.x, (e.N + 7) / 8, e.nxb), 256, 0, e.st>>>(fp, e.scratch.p); ++e.launches; }
        else if (nm == "fgp_sweep") e.fgp_launch(FGP_SWEEP, fp, nullptr);
        else if (nm == "fgp_first") e.fgp_launch(FGP_FIRST, fp, nullptr);
        else if (nm == "fgp_final") e.fgp_launch(FGP_FINAL, fp, e.scratch.p);
        else if (nm == "fgp2_sweep") e.fgp2_launch(FGP_SWEEP, FGP_SWEEP, fp, nullptr);
        else if (nm == "fgp2_first") e.fgp2_launch(FGP_FIRST, FGP_SWEEP, fp, nullptr);
        else if (nm == "fgp2_final") e.fgp2_launch(FGP_SWEEP, FGP_FINAL, fp, e.scratch.p);
        else if (nm == "tv_v1") { tv_value_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, e.st>>>(e.recon.p, e.dims(), nullptr, 1, e.d_red); ++e.launches; }
        else if (nm == "tv") { tv_value_tiled_kernel<<<dim3(gsw.x, (e.N + 7) / 8, e.nxb), 256, 0, e.st>>>(e.recon.p, e.dims(), nullptr, 1, e.d_red); ++e.launches; }
        else if (nm == "momentum") { momentum_kernel<<<ew_blocks(n / 4), 256, 0, e.st>>>(reinterpret_cast<float4*>(e.scratch.p), reinterpret_cast<float4*>(e.yk.p), reinterpret_cast<float4*>(e.recon_old.p), 0.5f, n / 4); ++e.launches; }
        else if (nm == "tvgd_grad") { tvgd_gradient_kernel<<<dim3(gsw.x, (e.N + 7) / 8, e.nxb), 256, 0, e.st>>>(gp); ++e.launches; }
        else if (nm == "tvgd_grad2") { tvgd_gradient_march_kernel<<<dim3(gsw.x, (e.N + 7) / 8, e.nxb), 256, 0, e.st>>>(gp); ++e.launches; }
        else if (nm == "tvgd_update") { tvgd_update_kernel<<<ew_blocks(n / 4), 256, 0, e.st>>>(e.yk.p, e.scratch.p, e.d_red + 1, 0.f, 0, n / 4); ++e.launches; }
        else if (nm == "sqdist") { reduce_kernel<RED_SQDIST><<<ew_blocks(n / 4), 256, 0, e.st>>>(reinterpret_cast<const float4*>(e.recon.p), reinterpret_cast<const float4*>(e.yk.p), n / 4, e.d_red + 2); ++e.launches; }
        else throw CudaError("ttv_time_kernel: unknown kernel name '" + nm + "'");
    };
    (void)arg_i; (void)arg_f;
    cudaEvent_t a, b;
    TTV_CUDA(cudaEventCreate(&a)); TTV_CUDA(cudaEventCreate(&b));
    one(); e.sync();
    TTV_CUDA(cudaEventRecord(a, e.st));
    for (int i = 0; i < reps; ++i) one();
    TTV_CUDA(cudaEventRecord(b, e.st));
    TTV_CUDA(cudaEventSynchronize(b));
    TTV_LAUNCH_CHECK();
    float t = 0; TTV_CUDA(cudaEventElapsedTime(&t, a, b));
    cudaEventDestroy(a); cudaEventDestroy(b);
    *ms = t / (reps > 0 ? reps : 1);
    TTV_CATCH
}

}  // extern "C"

#include "ttv_capi_fusion.cuh"

#include <mutex>
// ---------------------------------------------------------------------------------------------------------------------
// TV value WITHOUT any allocation (ttv_tv_value_noalloc, include/tomo_tv_b200.h).  No counterpart in the reference: added
// for callers that may not call cudaMalloc (CUDA-graph capture, serving loops).
//
// A variant of tv_value_kernel (ttv_tv.cuh:51; arithmetic of tv_3D_kernel, T/gpu/utils/regularizers/tv_gd.cu:27-47) on the
// caller's own array layout [nx][ny][nz].  The engine's kernel accumulates into a cudaMalloc'ed reduction slot; this one
// keeps its reduction scratch (one partial sum per CTA + a ticket) in the file-scope array below, which every GPU that
// loads the library holds until exit: TTV_NOALLOC_MAX_ROWS / 8 doubles + 256 bytes = 1 MiB + 256 B.
//   * in host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the device address is looked up once
//     per device with cudaGetSymbolAddress (noalloc_scratch()) and passed to the kernel as an argument;
//   * all calls share the array and it keeps the last call's contents: the ticket is cleared on the stream before every
//     launch, the partials are all rewritten before they are read.  Calls that overlap on the GPU (two streams, two host
//     threads) race on it -- serialise them.
// The sum is taken in a fixed order (CTA partials in index order), so the result is reproducible bit for bit.
// ---------------------------------------------------------------------------------------------------------------------
constexpr long long TTV_NOALLOC_MAX_ROWS_C = TTV_NOALLOC_MAX_ROWS;                 // rows = nx * ny
constexpr int NOALLOC_MAX_BLOCKS = (int)(TTV_NOALLOC_MAX_ROWS_C / 8);             // 8 rows (warps) per CTA
constexpr int NOALLOC_TICKET = NOALLOC_MAX_BLOCKS;                                 // element index of the ticket word
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_MAX_BLOCKS + 32];

__global__ void __launch_bounds__(256) tv_value_noalloc_kernel(const float* __restrict__ v, int nx, int ny, int nz,
                                                               double* __restrict__ scratch, float* __restrict__ out)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t rows = (int64_t)nx * ny;
    const int64_t row = (int64_t)blockIdx.x * 8 + w;          // (x, j)
    double tot = 0.0;
    if (row < rows) {
        const int x = (int)(row / ny), j = (int)(row % ny);
        const float* __restrict__ r0 = v + row * nz;
        const float* __restrict__ rx = v + ((int64_t)((x + 1) % nx) * ny + j) * nz;
        const float* __restrict__ rj = v + ((int64_t)x * ny + (j + 1) % ny) * nz;
        for (int k = lane; k < nz; k += 32) {
            const float c = r0[k];
            const float dx = c - rx[k], dy = c - rj[k], dz = c - r0[(k + 1) % nz];
            tot += (double)sqrtf(1e-6f + dx * dx + dy * dy + dz * dz);
        }
    }
    tot = warp_sum(tot);
    __shared__ double sh[8];
    __shared__ unsigned last;
    if (lane == 0) sh[w] = tot;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < 8; ++i) s += sh[i];
        scratch[blockIdx.x] = s;
        __threadfence();
        last = atomicAdd(reinterpret_cast<unsigned*>(scratch + NOALLOC_TICKET), 1u) == gridDim.x - 1 ? 1u : 0u;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    // last CTA: partials in index order, thread t takes t, t+256, ...
    double s = 0;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += 256) s += __ldcg(scratch + i);
    s = warp_sum(s);
    __syncthreads();
    if (lane == 0) sh[w] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int i = 0; i < 8; ++i) t += sh[i];
        *out = (float)t;
    }
}

// device address of g_noalloc_scratch on the CURRENT device, looked up once per device
static double* noalloc_scratch() {
    static std::mutex mu;
    static double* ptr[64] = {};
    int dev = -1;
    TTV_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw CudaError("ttv_tv_value_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[dev]) {
        void* p = nullptr;
        TTV_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) throw CudaError("ttv_tv_value_noalloc: cudaGetSymbolAddress returned a null address");
        ptr[dev] = static_cast<double*>(p);
    }
    return ptr[dev];
}

extern "C" int ttv_tv_value_noalloc(const float* d_vol, int nx, int ny, int nz, float* d_tv, void* stream) {
    TTV_TRY
    if (!d_vol || !d_tv) throw CudaError("ttv_tv_value_noalloc: null pointer");
    if (nx < 1 || ny < 1 || nz < 1) throw CudaError("ttv_tv_value_noalloc: nx, ny, nz must be >= 1");
    const long long rows = (long long)nx * ny;
    if (rows > TTV_NOALLOC_MAX_ROWS_C)
        throw CudaError("ttv_tv_value_noalloc: nx*ny = " + std::to_string(rows) + " exceeds the maximum of " +
                        std::to_string(TTV_NOALLOC_MAX_ROWS_C) + " rows (TTV_NOALLOC_MAX_ROWS) that the static scratch array holds");
    int dev = -1;
    TTV_CUDA(cudaGetDevice(&dev));
    for (const void* q : {(const void*)d_vol, (const void*)d_tv}) {
        cudaPointerAttributes at;
        TTV_CUDA(cudaPointerGetAttributes(&at, q));
        if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged)
            throw CudaError("ttv_tv_value_noalloc: d_vol and d_tv must be device pointers");
        if (at.type == cudaMemoryTypeDevice && at.device != dev)
            throw CudaError("ttv_tv_value_noalloc: pointer belongs to device " + std::to_string(at.device) +
                            ", current device is " + std::to_string(dev));
    }
    double* scratch = noalloc_scratch();
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned blocks = (unsigned)((rows + 7) / 8);
    TTV_CUDA(cudaMemsetAsync(scratch + NOALLOC_TICKET, 0, sizeof(double), st));     // the ticket must start at zero
    tv_value_noalloc_kernel<<<blocks, 256, 0, st>>>(d_vol, nx, ny, nz, scratch, d_tv);
    TTV_LAUNCH_CHECK();
    TTV_CATCH
}
