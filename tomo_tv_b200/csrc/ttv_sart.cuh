// SART: one projection angle at a time (A/cuda/2d/sart.cu:175-227).
//   p = (b_a - FP_a x) * R_a                       callFP_SART :256-271 + MUL_SART :44-50
//   x = max(x + beta * BP_a^{clamp}(p), 0)         devBP_SART, A/cuda/2d/par_bp.cu:139-159 (Clamp addressing, :269)
// The sub-iterations are sequentially dependent, so each is a pair of whole-volume launches that
// read cells straight from global memory / L2 (a single angle has no reuse worth staging).
#pragma once
#include "ttv_engine.cuh"

namespace ttv {

struct Sart1Params {
    const float* vol; const float* b; const float* R; float* proj;   // proj: [xb][Nd][32]
    float theta, detsize, offset; int vertical;
    int a, N, Nd, Na;
};

// block = 8 warps; warp = ONE ray: 8 float4 lanes (32 slices) x 4 line quarters.  A quarter is a run of whole 64-line restart
// periods of the position recurrence (par_fp.cu:49,282-286), so every sample position is the reference's; the four partial
// sums are combined by two shuffles in a fixed order (the same for any slab thickness: sharded == whole bit for bit; it
// differs from the reference's single running sum by rounding).  One warp per ray keeps 148 SMs busy on a 64-slice slab
// (one ray per quarter warp left 32 CTAs there) and quadruples the loads in flight.
__global__ void __launch_bounds__(256) sart_fp1_kernel(const Sart1Params p)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int lx = lane & 7, lq = lane >> 3;
    const int d = blockIdx.x * 8 + w;
    const int xb = blockIdx.y;
    const int N = p.N;
    if (d >= p.Nd) return;
    const float cth = __cosf(p.theta), sth = __sinf(p.theta);
    float detStep, step, corr1;
    if (p.vertical) { detStep = p.detsize / cth; step = sth / cth; corr1 = (cth < 0.0f) ? -1.0f / cth : 1.0f / cth; }
    else { detStep = -p.detsize / sth; step = cth / sth; corr1 = (sth > 0.0f) ? 1.0f / sth : -1.0f / sth; }
    const float t1 = __fadd_rn(__fadd_rn(__fmaf_rn((float)p.Nd, -0.5f, (float)d), 0.5f), -p.offset);
    const float m = __fmul_rn(t1, detStep);
    const float* base = p.vol + (size_t)xb * N * N * 32 + lx * 4;
    // stride between consecutive cells along the interpolation axis / along the line axis
    const int64_t s_interp = p.vertical ? 32 : (int64_t)N * 32;
    const int64_t s_line = p.vertical ? (int64_t)N * 32 : 32;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float fP = 0.f;
    // four lines per trip: positions first, then all eight loads in flight, then the arithmetic in line order (the sums are
    // accumulated exactly as in a one-line loop; the kernel is bound by memory latency, not by issue slots)
    const int chunk = ((N + 255) / 256) * 64;               // lines per quarter: a multiple of the 64-line restart period
    const int la = lq * chunk, lb = min(N, la + chunk);
    for (int l4 = la; l4 < lb; l4 += 4) {
        float4 v0[4], v1[4]; float w0[4], w1[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int line = l4 + u;
            if ((line & 63) == 0) {
                const float t2 = __fadd_rn(__fmaf_rn((float)N, -0.5f, (float)line), 0.5f);
                float f = __fmaf_rn(t2, step, m);
                f = __fmaf_rn((float)N, 0.5f, f);
                f = __fadd_rn(f, -0.5f);
                fP = __fadd_rn(f, 0.5f);
            } else fP = __fadd_rn(fP, step);
            const int q = __float2int_rd(fP * 256.0f + 0.5f) - 128;   // round half up (ttv_common.cuh)
            const int i = q >> 8;
            w1[u] = (float)(q & 255); w0[u] = 256.0f - w1[u];
            v0[u] = make_float4(0.f, 0.f, 0.f, 0.f); v1[u] = v0[u];
            if (line < lb) {
                const float* lp = base + line * s_line;
                if (i >= 0 && i < N) v0[u] = *reinterpret_cast<const float4*>(lp + i * s_interp);
                if (i + 1 >= 0 && i + 1 < N) v1[u] = *reinterpret_cast<const float4*>(lp + (i + 1) * s_interp);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (l4 + u < lb) {
                acc.x = fmaf(w1[u], v1[u].x, fmaf(w0[u], v0[u].x, acc.x)); acc.y = fmaf(w1[u], v1[u].y, fmaf(w0[u], v0[u].y, acc.y));
                acc.z = fmaf(w1[u], v1[u].z, fmaf(w0[u], v0[u].z, acc.z)); acc.w = fmaf(w1[u], v1[u].w, fmaf(w0[u], v0[u].w, acc.w));
            }
        }
    }
#pragma unroll
    for (int o = 8; o <= 16; o <<= 1) {                     // (q0 + q1) + (q2 + q3)
        acc.x += __shfl_xor_sync(0xffffffffu, acc.x, o); acc.y += __shfl_xor_sync(0xffffffffu, acc.y, o);
        acc.z += __shfl_xor_sync(0xffffffffu, acc.z, o); acc.w += __shfl_xor_sync(0xffffffffu, acc.w, o);
    }
    if (lq != 0) return;
    const float cs = corr1 * (1.0f / 256.0f);
    const float4 bb = *reinterpret_cast<const float4*>(p.b + (((size_t)xb * p.Na + p.a) * p.Nd + d) * 32 + lx * 4);
    const float wgt = p.R[(size_t)p.a * p.Nd + d];
    float4 r;
    r.x = fmaf(acc.x, -cs, bb.x) * wgt; r.y = fmaf(acc.y, -cs, bb.y) * wgt;
    r.z = fmaf(acc.z, -cs, bb.z) * wgt; r.w = fmaf(acc.w, -cs, bb.w) * wgt;
    *reinterpret_cast<float4*>(p.proj + ((size_t)xb * p.Nd + d) * 32 + lx * 4) = r;
}

struct SartBpParams { float* vol; const float* proj; float sc, ss, off, relax; int N, Nd; };

__global__ void __launch_bounds__(256) sart_bp1_kernel(const SartBpParams p)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int lx = lane & 7, lr = lane >> 3;
    const int N = p.N;
    const int64_t v = ((int64_t)blockIdx.x * 8 + w) * 4 + lr;     // voxel index Y*N + X
    const int xb = blockIdx.y;
    if (v >= (int64_t)N * N) return;
    const int Y = (int)(v / N), X = (int)(v % N);
    const float fX = (float)X - 0.5f * (float)N + 0.5f, fY = (float)Y - 0.5f * (float)N + 0.5f;
    const float fT = __fadd_rn(__fmaf_rn(fX, p.sc, -__fmul_rn(fY, p.ss)), p.off);
    const int q = __float2int_rd(fT * 256.0f + 0.5f) - 128;   // round half up (ttv_common.cuh)
    const int i = q >> 8;
    const float w1 = (float)(q & 255) * (1.0f / 256.0f), w0 = 1.0f - w1;
    const int i0 = min(max(i, 0), p.Nd - 1), i1 = min(max(i + 1, 0), p.Nd - 1);     // Clamp addressing
    const float* pr = p.proj + (size_t)xb * p.Nd * 32 + lx * 4;
    const float4 v0 = *reinterpret_cast<const float4*>(pr + (size_t)i0 * 32);
    const float4 v1 = *reinterpret_cast<const float4*>(pr + (size_t)i1 * 32);
    float* o = p.vol + (((size_t)xb * N + Y) * N + X) * 32 + lx * 4;
    float4 x = *reinterpret_cast<float4*>(o);
    x.x = fmaxf(fmaf(fmaf(w1, v1.x, w0 * v0.x), p.relax, x.x), 0.f);
    x.y = fmaxf(fmaf(fmaf(w1, v1.y, w0 * v0.y), p.relax, x.y), 0.f);
    x.z = fmaxf(fmaf(fmaf(w1, v1.z, w0 * v0.z), p.relax, x.z), 0.f);
    x.w = fmaxf(fmaf(fmaf(w1, v1.w, w0 * v0.w), p.relax, x.w), 0.f);
    *reinterpret_cast<float4*>(o) = x;
}

}  // namespace ttv
#include "ttv_sartc.cuh"
namespace ttv {

inline void Engine::sart_pass(float beta, int niter) {
    need_angles();
    if (!sart_ready) throw CudaError("SART called before initialize_SART");
    if (beta != 1) sart_relax = beta;                           // sticky (tomoengine.cpp:172)
    // projection order is re-drawn on every call (tomoengine.cpp:164, CudaSartAlgorithm.cpp:109-136)
    sart_order.resize(Na);
    for (int i = 0; i < Na; ++i) sart_order[i] = i;
    if (projOrder == "random") {
        for (int i = 0; i < Na - 1; ++i) { const int k = rand() % (Na - i); std::swap(sart_order[i], sart_order[i + k]); }
    }
    // slice-resident cluster kernel (ttv_sartc.cuh): the whole pass in one launch when a slice fits in a cluster's shared memory
    if (sart_cluster && niter > 0) {
        int optin = 0;
        TTV_CUDA(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
        SartcParams q;
        size_t smem = 0;
        const int S = sartc_pick_S(N, nx, (size_t)optin, q.NCH, q.CR, smem);
        // S = 1 (one slice per cluster, N > ~360) pays the index arithmetic of every sample for a single slice: measured 128 ms per
        // pass at 512^3 x 180 against 113 ms for the per-angle kernels below, which share it between the 4 slices of a float4 lane --
        // the cluster kernel is used where it keeps at least two slices per cluster (4.0 against 11.5 ms at 256^3 x 71)
        if (S >= 2 || (S == 1 && sart_cluster_force)) {
            const int nsub = Na * niter;
            std::vector<SartcAngle> ha(nsub);
            for (int it = 0; it < nsub; ++it) {
                const int a = sart_order[it % Na];
                ha[it] = SartcAngle{geom[a].theta, geom[a].detsize, geom[a].offset, geom[a].sc, geom[a].ss, geom[a].off, geom[a].vertical, a};
            }
            if (sartc_cap < (size_t)nsub) {
                cudaFree(d_sartc);
                TTV_CUDA(cudaMalloc(&d_sartc, nsub * sizeof(SartcAngle)));
                sartc_cap = nsub;
            }
            TTV_CUDA(cudaMemcpyAsync(d_sartc, ha.data(), nsub * sizeof(SartcAngle), cudaMemcpyHostToDevice, st));
            q.vol = recon.p; q.b = b.p; q.R = R.p; q.ang = d_sartc; q.nsub = nsub; q.N = N; q.Na = Na; q.nx = nx; q.relax = sart_relax;
            q.RB = (N + SARTC_CL - 1) / SARTC_CL;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)(SARTC_CL * ((nx + S - 1) / S)));
            cfg.blockDim = dim3(SARTC_THREADS);
            cfg.dynamicSmemBytes = smem;
            cfg.stream = st;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = SARTC_CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            ProfScope ps(*this, PF_SART_FP);
            if (S == 4) {
                TTV_CUDA(cudaFuncSetAttribute(sartc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                TTV_CUDA(cudaLaunchKernelEx(&cfg, sartc_kernel<4>, q));
            } else if (S == 2) {
                TTV_CUDA(cudaFuncSetAttribute(sartc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                TTV_CUDA(cudaLaunchKernelEx(&cfg, sartc_kernel<2>, q));
            } else {
                TTV_CUDA(cudaFuncSetAttribute(sartc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                TTV_CUDA(cudaLaunchKernelEx(&cfg, sartc_kernel<1>, q));
            }
            ++launches;
            TTV_LAUNCH_CHECK();
            sync();                                             // the host copy of the angle table is reused by the next call
            return;
        }
    }
    if (scratch.n < (size_t)nxb * N * 32) scratch.alloc(vol_elems());
    float* proj = scratch.p;
    dim3 gfp((N + 7) / 8, nxb), gbp((unsigned)(((int64_t)N * N + 31) / 32), nxb);
    for (int it = 0; it < Na * niter; ++it) {
        const int a = sart_order[it % Na];
        Sart1Params f; f.vol = recon.p; f.b = b.p; f.R = R.p; f.proj = proj;
        f.theta = geom[a].theta; f.detsize = geom[a].detsize; f.offset = geom[a].offset; f.vertical = geom[a].vertical;
        f.a = a; f.N = N; f.Nd = N; f.Na = Na;
        { ProfScope ps(*this, PF_SART_FP); sart_fp1_kernel<<<gfp, 256, 0, st>>>(f); ++launches; }
        SartBpParams q; q.vol = recon.p; q.proj = proj; q.sc = geom[a].sc; q.ss = geom[a].ss; q.off = geom[a].off;
        q.relax = sart_relax; q.N = N; q.Nd = N;
        { ProfScope ps(*this, PF_SART_BP); sart_bp1_kernel<<<gbp, 256, 0, st>>>(q); ++launches; }
    }
    TTV_LAUNCH_CHECK();
    sync();
}

}  // namespace ttv
