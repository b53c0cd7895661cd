// Multi-modal data fusion (HAADF + chemical maps) -- the elementwise / reduction kernels around the
// projector pair and the device-resident state machine of the reference's `multimodal` class
// (T/chemistry/utils/multimodal.cpp, multimodal.hpp).  sm_100a.
//
// The reference runs one 2-D slice at a time through host Eigen vectors: per slice an Eigen sparse
// mat-vec with the summation matrix Sigma, 2*Nel+3 ASTRA upload/compute/download cycles and host
// vector arithmetic (multimodal.cpp:452-491).  Here every element volume lives in HBM in the
// x-interleaved layout and each term is one batched launch over all slices:
//   m = Sigma x^gamma            mm_model_kernel       (Sigma has one entry per (pixel, element):
//                                                       utils/fusion_helper.py:5-32 -> per-pixel weights W_e)
//   g = A_h m, ||g - b_h||^2     fp_kernel<FP_STORE_DIST>
//   u = SIRT_h^n(m)              sino_resid_kernel (first residual re-uses g) + bp/fp SIRT kernels
//   A_c x_e, KL ratio, KL cost   fp_kernel<FP_KL>
//   x_e update (+ positivity)    bp_kernel<BP_FUSE_UPD>  (Sigma^T(u-m), gamma x^(gamma-1), both steps, clamp)
#pragma once
#include "ttv_engine.cuh"
#include <memory>

namespace ttv {

constexpr int MM_MAX_EL = 8;

struct MmModelParams {
    const float* x[MM_MAX_EL];   // element volumes (x-interleaved)
    const float* W;              // [Nel][N*N] summation weights
    float* M; float* U;          // outputs: model volume and its copy (SIRT start)
    int nel; float gamma; int64_t cells; int64_t npix;   // cells = nxb*N*N
};

// m[p] = ((0 + w_0 x_0^g) + w_1 x_1^g) + ...   -- Eigen row-major sparse * dense accumulates one rounded
// product at a time (multimodal.cpp:427-428); pow as in Eigen's array pow (std::pow)
__global__ void __launch_bounds__(256) mm_model_kernel(const MmModelParams p)
{
    const int64_t n4 = p.cells * 8;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t pix = (i >> 3) % p.npix;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int e = 0; e < p.nel; ++e) {
            float4 v = reinterpret_cast<const float4*>(p.x[e])[i];
            const float w = p.W[(int64_t)e * p.npix + pix];
            if (p.gamma != 1.0f) { v.x = powf(v.x, p.gamma); v.y = powf(v.y, p.gamma); v.z = powf(v.z, p.gamma); v.w = powf(v.w, p.gamma); }
            acc.x = __fadd_rn(acc.x, __fmul_rn(w, v.x)); acc.y = __fadd_rn(acc.y, __fmul_rn(w, v.y));
            acc.z = __fadd_rn(acc.z, __fmul_rn(w, v.z)); acc.w = __fadd_rn(acc.w, __fmul_rn(w, v.w));
        }
        reinterpret_cast<float4*>(p.M)[i] = acc;
        if (p.U) reinterpret_cast<float4*>(p.U)[i] = acc;
    }
}

// r = (b - g) * R : the residual of the first SIRT iteration from the projections already computed
__global__ void __launch_bounds__(256) sino_resid_kernel(const float4* __restrict__ b, const float4* __restrict__ g, const float* __restrict__ R,
                                                          float4* __restrict__ r, int64_t cells, int64_t nad)
{
    const int64_t n4 = cells * 8;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float w = R[(i >> 3) % nad];
        const float4 bb = b[i], gg = g[i];
        r[i] = make_float4((bb.x - gg.x) * w, (bb.y - gg.y) * w, (bb.z - gg.z) * w, (bb.w - gg.w) * w);
    }
}

// per-projection maximum over slices and detectors (multimodal.cpp:323-327): out[a] as ordered ints
__global__ void __launch_bounds__(256) angle_max_kernel(const float* __restrict__ s, int* __restrict__ out, int nx, int nxb, int Na, int Nd)
{
    const int a = blockIdx.x;
    float m = -INFINITY;
    const int64_t per = (int64_t)Nd * 32;
    for (int xb = blockIdx.y; xb < nxb; xb += gridDim.y) {
        const float* base = s + ((int64_t)xb * Na + a) * per;
        for (int64_t i = threadIdx.x; i < per; i += blockDim.x)
            if (xb * 32 + (int)(i & 31) < nx) m = fmaxf(m, base[i]);
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) atomicMax(out + a, float_ordered(m));
}
// b[:, a, :] = b / max_b[a] * max_g[a]
__global__ void __launch_bounds__(256) angle_rescale_kernel(float4* __restrict__ b, const int* __restrict__ mb, const int* __restrict__ mg,
                                                             int64_t cells, int Na, int Nd)
{
    const int64_t n4 = cells * 8;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int a = (int)(((i >> 3) / Nd) % Na);
        int ib = mb[a], ig = mg[a];
        ib = ib >= 0 ? ib : ib ^ 0x7FFFFFFF; ig = ig >= 0 ? ig : ig ^ 0x7FFFFFFF;
        const float fb = __int_as_float(ib), fg = __int_as_float(ig);
        float4 v = b[i];
        v.x = __fmul_rn(__fdiv_rn(v.x, fb), fg); v.y = __fmul_rn(__fdiv_rn(v.y, fb), fg);
        v.z = __fmul_rn(__fdiv_rn(v.z, fb), fg); v.w = __fmul_rn(__fdiv_rn(v.w, fb), fg);
        b[i] = v;
    }
}
__global__ void fill_int_kernel(int* p, int v, int n) { const int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = v; }

// -----------------------------------------------------------------------------------------------
// Device-resident mirror of the reference's `multimodal` state (multimodal.hpp:38-70)
// -----------------------------------------------------------------------------------------------
class MultiModal {
public:
    int nx, N, nel, nh, nc;
    std::unique_ptr<Engine> eh, ec;          // HAADF / chemical geometry (share one stream)
    std::vector<DBuf> X, orig, bc, gc;       // element volumes, ground truth, chemical tilt series, model projections
    DBuf W, M;                               // summation weights [Nel][N*N], model volume; U lives in eh->recon
    int* d_amax = nullptr;
    bool measureHaadf = false, measureChem = false, sigma_ready = false, lip_ready = false, sart_ready = false;
    float gamma = 1.0f, eps = 1e-1f, L_Aps = 0.f;
    int gpuID = -1;

    MultiModal(int ns, int nray, int nelements, const float* hang, int nhaadf, const float* cang, int nchem, int dev, int x_offset, int ns_global)
        : nx(ns), N(nray), nel(nelements), nh(nhaadf), nc(nchem)
    {
        if (nelements <= 0 || nelements > MM_MAX_EL) throw CudaError("multimodal: between 1 and 8 elements are supported");
        if (nhaadf <= 0 || nchem <= 0) throw CudaError("multimodal: HAADF and chemical tilt angles are required");
        eh.reset(new Engine(ns, nray, hang, nhaadf, dev, x_offset, ns_global));
        ec.reset(new Engine(ns, nray, cang, nchem, eh->device, x_offset, ns_global));
        ec->adopt_stream(eh->st);
        X.resize(nel); orig.resize(nel); bc.resize(nel); gc.resize(nel);
        for (int e = 0; e < nel; ++e) { X[e].alloc(eh->vol_elems()); bc[e].alloc(ec->sino_elems()); }
        M.alloc(eh->vol_elems());
        W.alloc((size_t)nel * N * N);
        TTV_CUDA(cudaMalloc(&d_amax, 2 * (size_t)nh * sizeof(int)));
    }
    ~MultiModal() {
        cudaSetDevice(eh->device);
        cudaStreamSynchronize(eh->st);
        for (auto* v : {&X, &orig, &bc, &gc}) for (auto& d : *v) d.release();
        W.release(); M.release(); cudaFree(d_amax);
        ec.reset();                          // borrowed stream: destroy the borrower first
        eh.reset();
    }
    void use() { eh->use(); }
    cudaStream_t st() const { return eh->st; }
    size_t npix() const { return (size_t)N * N; }
    long long launches() const { return eh->launches + ec->launches + own_launches; }
    long long own_launches = 0;
    void check_el(int e) const { if (e < 0 || e >= nel) throw CudaError("element index out of range"); }

    // Sigma as [3][nnz] float32 rows / cols / vals (multimodal.cpp:268-274; fusion_helper.py:5-32).
    // Supported structure: entry (row p, col p + Npix*e) -- one weight per (pixel, element).
    void load_sigma(const float* trip, int nnz) {
        use();
        std::vector<float> w((size_t)nel * npix(), 0.0f);
        const int64_t np = (int64_t)npix();
        for (int i = 0; i < nnz; ++i) {
            const int64_t r = (int64_t)trip[i], c = (int64_t)trip[(size_t)nnz + i];
            const float v = trip[2 * (size_t)nnz + i];
            if (r < 0 || r >= np || c < 0 || c >= np * nel || c % np != r)
                throw CudaError("load_sigma: only summation matrices with entries (p, p + Npix*e) are supported (fusion_helper.create_weighted_summation_matrix)");
            w[(size_t)(c / np) * np + r] = v;
        }
        TTV_CUDA(cudaMemcpyAsync(W.p, w.data(), w.size() * sizeof(float), cudaMemcpyHostToDevice, st()));
        eh->sync();
        sigma_ready = true;
    }

    void model(float* U) {   // M = Sigma x^gamma (and a copy in U)
        if (!sigma_ready) throw CudaError("summation matrix not loaded (load_sigma)");
        MmModelParams p;
        for (int e = 0; e < MM_MAX_EL; ++e) p.x[e] = e < nel ? X[e].p : nullptr;
        p.W = W.p; p.M = M.p; p.U = U; p.nel = nel; p.gamma = gamma; p.cells = (int64_t)eh->nxb * N * N; p.npix = (int64_t)npix();
        Engine::ProfScope ps(*eh, PF_EW);
        mm_model_kernel<<<ew_blocks((size_t)p.cells * 8), 256, 0, st()>>>(p); ++own_launches;
        TTV_LAUNCH_CHECK();
    }

    void estimate_lipschitz() { use(); ec->compute_weights(); L_Aps = ec->L_A; lip_ready = true; }   // multimodal.cpp:259-265 (L_ASig is never read)

    // multimodal.cpp:277-304
    float poisson_ml(float lamC) {
        use();
        if (!lip_ready) throw CudaError("poisson_ml called before estimate_lipschitz");
        if (!ec->r.ok()) ec->r.alloc(ec->sino_elems());
        ec->zero_red(3);
        for (int e = 0; e < nel; ++e) {
            ec->fp<FP_KL>(X[e].p, ec->r.p, bc[e].p, ec->d_red + 3, -1, eps);
            ec->bp<BP_STEP_UPD>(ec->r.p, X[e].p, lamC / L_Aps);
        }
        const double c = ec->read_red(3);
        return measureChem ? (float)c : 0.0f;
    }

    // multimodal.cpp:452-491 (+ fuse :425-441, SIRT(-1,s,n) :352-361)
    void data_fusion(float lamH, float lamC, int nIter, bool sart, float* costH, float* costC) {
        use();
        if (!lip_ready) throw CudaError("data_fusion called before estimate_lipschitz");
        if (sart) { if (!sart_ready) { eh->projOrder = "sequential"; eh->compute_weights(); eh->sart_ready = true; sart_ready = true; } }
        else eh->compute_weights();
        if (!ec->r.ok()) ec->r.alloc(ec->sino_elems());
        if (!eh->r.ok()) eh->r.alloc(eh->sino_elems());
        float* U = eh->recon.p;
        model(U);
        eh->zero_red(4);
        eh->fp<FP_STORE_DIST>(M.p, eh->g.p, eh->b.p, eh->d_red + 4);           // g = A_h m, ||g - b_h||^2
        if (sart) eh->sart_pass(1.0f, nIter);                                  // run(Nproj*nIter) on u = m
        else if (nIter >= 1) {
            const int64_t cells = (int64_t)eh->nxb * nh * N;
            sino_resid_kernel<<<ew_blocks((size_t)cells * 8), 256, 0, st()>>>(reinterpret_cast<const float4*>(eh->b.p), reinterpret_cast<const float4*>(eh->g.p),
                                                                             eh->R.p, reinterpret_cast<float4*>(eh->r.p), cells, (int64_t)nh * N);
            ++own_launches; TTV_LAUNCH_CHECK();
            eh->bp<BP_SIRT_UPD>(eh->r.p, U);
            for (int it = 1; it < nIter; ++it) eh->sirt_step(U);
        }
        ec->zero_red(3);
        BpParams ex; ex.M = M.p; ex.U = U; ex.gamma = gamma; ex.lamH = lamH;
        for (int e = 0; e < nel; ++e) {
            ec->fp<FP_KL>(X[e].p, ec->r.p, bc[e].p, ec->d_red + 3, -1, eps);
            ex.W = W.p + (size_t)e * npix();
            ec->bp<BP_FUSE_UPD>(ec->r.p, X[e].p, lamC / L_Aps, -1, &ex);
        }
        const double dh = eh->read_red(4), dc = ec->read_red(3);
        *costH = measureHaadf ? (float)std::sqrt(dh) : 0.0f;
        *costC = measureChem ? (float)dc : 0.0f;
    }

    // cuda_scalar_multiply_4D (T/chemistry/utils/container/matrix_ops.cu:625-)
    void rescale_tomograms(float scale) {
        use();
        for (int e = 0; e < nel; ++e) {
            ew_kernel<EW_SCALE><<<ew_blocks(eh->vol_elems() / 4), 256, 0, st()>>>(reinterpret_cast<float4*>(X[e].p), scale, eh->vol_elems() / 4);
            ++own_launches;
        }
        TTV_LAUNCH_CHECK();
    }

    // multimodal.cpp:312-328: g = A_h Sigma x^gamma, then per projection  b_h = b_h / max(b_h) * max(g)
    void rescale_projections() {
        use();
        model(nullptr);
        eh->fp<FP_STORE>(M.p, eh->g.p, nullptr, nullptr);
        fill_int_kernel<<<(2 * nh + 255) / 256, 256, 0, st()>>>(d_amax, float_ordered_neg_inf(), 2 * nh);
        dim3 grid(nh, std::min(eh->nxb, 64));
        angle_max_kernel<<<grid, 256, 0, st()>>>(eh->b.p, d_amax, nx, eh->nxb, nh, N);
        angle_max_kernel<<<grid, 256, 0, st()>>>(eh->g.p, d_amax + nh, nx, eh->nxb, nh, N);
        own_launches += 3;
        if (eh->nranks > 1 && eh->comm) {    // slab engines: the maxima are global over the tilt axis
            std::vector<int> h(2 * nh); std::vector<float> f(2 * nh);
            TTV_CUDA(cudaMemcpyAsync(h.data(), d_amax, h.size() * sizeof(int), cudaMemcpyDeviceToHost, st())); eh->sync();
            for (size_t i = 0; i < h.size(); ++i) { int b = h[i] >= 0 ? h[i] : h[i] ^ 0x7FFFFFFF; std::memcpy(&f[i], &b, 4); }
            float* d = reinterpret_cast<float*>(M.p);   // M is scratch here
            TTV_CUDA(cudaMemcpyAsync(d, f.data(), f.size() * sizeof(float), cudaMemcpyHostToDevice, st()));
            Nccl& nc = Nccl::get();
            nc.check(nc.AllReduce(d, d, f.size(), Nccl::Float32, Nccl::Max, eh->comm, st()), "ncclAllReduce");
            TTV_CUDA(cudaMemcpyAsync(f.data(), d, f.size() * sizeof(float), cudaMemcpyDeviceToHost, st())); eh->sync();
            for (size_t i = 0; i < h.size(); ++i) h[i] = Engine::float_ordered_host(f[i]);
            TTV_CUDA(cudaMemcpyAsync(d_amax, h.data(), h.size() * sizeof(int), cudaMemcpyHostToDevice, st())); eh->sync();
        }
        const int64_t cells = (int64_t)eh->nxb * nh * N;
        angle_rescale_kernel<<<ew_blocks((size_t)cells * 8), 256, 0, st()>>>(reinterpret_cast<float4*>(eh->b.p), d_amax, d_amax + nh, cells, nh, N);
        ++own_launches; TTV_LAUNCH_CHECK();
        eh->sync();
    }
    static int float_ordered_neg_inf() { return Engine::float_ordered_host(-INFINITY); }

    // multimodal.cpp:207-218: || A_c x - b_c ||_2 over all elements and slices
    float data_distance() {
        use();
        ec->zero_red(4);
        for (int e = 0; e < nel; ++e) {
            if (!gc[e].ok()) gc[e].alloc(ec->sino_elems());
            ec->fp<FP_STORE_DIST>(X[e].p, gc[e].p, bc[e].p, ec->d_red + 4);
        }
        return (float)std::sqrt(ec->read_red(4));
    }

    // cuda_tv_fgp_4D / cuda_tv_gd_4D: the 3-D routine per element, TV values summed in float32
    // (T/chemistry/utils/regularizers/tv_fgp.cu:192-281, tv_gd.cu:208-)
    float tv_4d(int ng, float par, bool fgp) {
        use();
        float tv = 0.0f;
        for (int e = 0; e < nel; ++e) {
            std::swap(ec->recon.p, X[e].p);
            float t;
            try { t = fgp ? ec->tv_fgp(ng, par) : ec->tv_gd(ng, par); }
            catch (...) { std::swap(ec->recon.p, X[e].p); throw; }
            std::swap(ec->recon.p, X[e].p);
            tv += t;
        }
        return tv;
    }

    // cuda_rmse_4D (container/matrix_ops.cu:568-621): sqrt(sum (x-o)^2 / Nvox) per element
    void rmse(float* out) {
        use();
        for (int e = 0; e < nel; ++e) {
            if (!orig[e].ok()) throw CudaError("rmse: original volume not set (initialize_initial_volume / set_original_volume)");
            const double s = eh->reduce<RED_SQDIST>(X[e].p, orig[e].p, eh->vol_elems());
            out[e] = (float)std::sqrt(s / ((double)eh->nxg * N * N));
        }
    }

    void restart_recon() {
        use();
        for (int e = 0; e < nel; ++e) TTV_CUDA(cudaMemsetAsync(X[e].p, 0, eh->vol_elems() * sizeof(float), st()));
        eh->sync();
    }

    // multimodal.cpp:363-371 / :415-423.  The reference's SIRT/SART objects are built on the HAADF projector and
    // only swap the data pointers (updateSlice), so these are well defined only when both tilt series share
    // their angles; anything else reads past the chemical sinogram in the reference and is rejected here.
    void chemical_recon(int nIter, bool sart) {
        use();
        if (nh != nc) throw CudaError("chemical_SIRT/SART: the reference reconstructs with the HAADF projector; HAADF and chemical angle counts must match");
        eh->compute_weights();
        for (int e = 0; e < nel; ++e) {
            if (!sart) { for (int it = 0; it < nIter; ++it) eh->sirt_step(X[e].p, bc[e].p); continue; }
            eh->sart_ready = true;
            std::swap(eh->recon.p, X[e].p); std::swap(eh->b.p, bc[e].p);
            try { eh->sart_pass(1.0f, nIter); } catch (...) { std::swap(eh->recon.p, X[e].p); std::swap(eh->b.p, bc[e].p); throw; }
            std::swap(eh->recon.p, X[e].p); std::swap(eh->b.p, bc[e].p);
        }
        eh->sync();
    }

    // multimodal.cpp:180-193: HAADF forward projection of one host slice [N*N] -> [Nproj_h*N]
    void forward_projection_2d(const float* in, float* out) {
        use();
        const size_t nv1 = (size_t)N * N * 32, ns1 = (size_t)nh * N * 32;
        DBuf v1, s1; v1.alloc(nv1); s1.alloc(ns1);
        if (eh->stage.n < std::max(npix(), (size_t)nh * N)) eh->stage.alloc(std::max(npix(), (size_t)nh * N));
        TTV_CUDA(cudaMemcpyAsync(eh->stage.p, in, npix() * sizeof(float), cudaMemcpyHostToDevice, st()));
        scatter_slice_kernel<<<256, 256, 0, st()>>>(v1.p, eh->stage.p, 0, (int64_t)npix());
        eh->fp<FP_STORE>(v1.p, s1.p, nullptr, nullptr, 1);
        gather_slice_kernel<<<256, 256, 0, st()>>>(s1.p, eh->stage.p, 0, (int64_t)nh * N);
        own_launches += 2;
        TTV_CUDA(cudaMemcpyAsync(out, eh->stage.p, (size_t)nh * N * sizeof(float), cudaMemcpyDeviceToHost, st()));
        eh->sync();
        v1.release(); s1.release();
    }
};

}  // namespace ttv
