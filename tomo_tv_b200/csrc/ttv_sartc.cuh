// SART with the slice resident in distributed shared memory: one thread-block cluster per slice group, all angles of a
// pass in ONE launch.  sm_100a (clusters of 8 CTAs, DSMEM).
//
// Arithmetic: A/cuda/2d/sart.cu:175-227 (one projection angle at a time),
//   p = (b_a - FP_a x) * R_a                       callFP_SART :256-271 + MUL_SART :44-50
//   x = max(x + beta * BP_a^{clamp}(p), 0)         devBP_SART, A/cuda/2d/par_bp.cu:139-159 (Clamp addressing, :269)
// per-sample operations as in sart_fp1_kernel / sart_bp1_kernel (ttv_sart.cuh): float32 position recurrence with a restart
// every 64 lines (par_fp.cu:49,282-286), 8-bit texture weights, Border addressing for the volume.
//
// Why: the sub-iterations are sequentially dependent, so ttv_sart.cuh runs 2 * Na whole-volume launches per pass -- about
// 3 V of HBM traffic per ANGLE (113 ms per pass at 512^3 x 180, 2.4 TB/s effective).  Slices are independent and one slice
// is small (1 MiB at N = 512): a cluster of 8 CTAs keeps S slices (S = 4, 2 or 1: what fits in 8 x ~200 KB) in its shared
// memory for the whole pass, so HBM sees every voxel once on the way in and once on the way out (2 V + 2 S per pass).
//   * The slice is stored LINE-major: the lines of the current angle's orientation (rows for "vertical" angles, columns for
//     "horizontal" ones) are dealt out to the 8 CTAs in bands, the cells of a line are contiguous.  Both taps of every Joseph
//     sample then sit in the owning CTA's own shared memory, and the position recurrence runs along the band.  When the
//     orientation changes (twice per pass in sequential order) the slice goes through its home in HBM / L2 and comes back
//     transposed.
//   * FP: thread = (ray, chunk of the band's lines): partial ray sums per CTA, left in shared memory; after ONE cluster
//     barrier per angle every CTA sums the 8 x chunks partials of every ray through DSMEM in a fixed order (deterministic, the
//     same for any sharding of the volume; it differs from the reference's single running sum by rounding only) and keeps
//     its own copy of the weighted residual p.  Partials are double-buffered by angle parity, so no second barrier.
//   * BP: every CTA updates its own band from its copy of p.  Index and weight of a sample are computed once per cell and
//     serve the S slices of the cell.
#pragma once
#include <cooperative_groups.h>
#include "ttv_engine.cuh"

namespace ttv {

namespace cg = cooperative_groups;

constexpr int SARTC_CL = 8;                 // CTAs per cluster
constexpr int SARTC_THREADS = 512;

struct SartcAngle { float theta, detsize, offset, sc, ss, off; int vertical, a; };
struct SartcParams {
    float* vol; const float* b; const float* R; const SartcAngle* ang;
    int nsub;                               // sub-iterations in this launch (Na * niter)
    int N, Na, nx; float relax;
    int RB;                                 // lines per band = ceil(N / 8)
    int NCH, CR;                            // chunks of a band per ray (FP work items = N * NCH), lines per chunk
};

// cluster-wide barrier with release / acquire at CLUSTER scope (orders shared::cluster accesses).  cooperative_groups'
// cluster.sync() adds a GPU-scope MEMBAR and an L1 invalidate (CCTL.IVALL) on either side -- microseconds per angle here.
__device__ __forceinline__ void cluster_sync_sc() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int S> struct SVec;
template <> struct SVec<1> { typedef float T; };
template <> struct SVec<2> { typedef float2 T; };
template <> struct SVec<4> { typedef float4 T; };

template <int S> __device__ __forceinline__ void sv_zero(float (&v)[S]) {
#pragma unroll
    for (int s = 0; s < S; ++s) v[s] = 0.f;
}
template <int S> __device__ __forceinline__ void sv_load(float (&v)[S], const float* p) {
    const typename SVec<S>::T t = *reinterpret_cast<const typename SVec<S>::T*>(p);
    const float* f = reinterpret_cast<const float*>(&t);
#pragma unroll
    for (int s = 0; s < S; ++s) v[s] = f[s];
}
template <int S> __device__ __forceinline__ void sv_store(float* p, const float (&v)[S]) {
    typename SVec<S>::T t;
    float* f = reinterpret_cast<float*>(&t);
#pragma unroll
    for (int s = 0; s < S; ++s) f[s] = v[s];
    *reinterpret_cast<typename SVec<S>::T*>(p) = t;
}

// smem: band [RB][N][S] | partial [2][NCH][N][S] | p [N][S]
template <int S>
__global__ void __launch_bounds__(SARTC_THREADS, 1) sartc_kernel(const SartcParams p)
{
    extern __shared__ __align__(16) float sm[];
    cg::cluster_group cluster = cg::this_cluster();
    const int cr = (int)cluster.block_rank();
    const int grp = blockIdx.x / SARTC_CL;                  // slice group
    const int x0 = grp * S, xb = x0 >> 5, xi = x0 & 31;
    const int N = p.N, RB = p.RB, tid = threadIdx.x;
    const int l0 = cr * RB, l1 = min(N, l0 + RB), nl = max(0, l1 - l0);      // my band of lines
    float* band = sm;
    float* part = band + (size_t)RB * N * S;
    float* pw = part + (size_t)2 * p.NCH * N * S;
    const int64_t cpb = (int64_t)N * N;
    float* gv = p.vol + (int64_t)xb * cpb * 32 + xi;         // element (line, pos) of a vertical layout: gv[(line*N + pos)*32]

    // band <-> HBM; vertical: line = row j, pos = k; horizontal: line = column k, pos = row j
    auto load_band = [&](int vertical) {
        for (int idx = tid; idx < nl * N; idx += SARTC_THREADS) {
            const int ll = idx / N, pos = idx - ll * N, line = l0 + ll;
            const int64_t cell = vertical ? (int64_t)line * N + pos : (int64_t)pos * N + line;
            float v[S];
            const typename SVec<S>::T t = __ldcg(reinterpret_cast<const typename SVec<S>::T*>(gv + cell * 32));
            const float* f = reinterpret_cast<const float*>(&t);
#pragma unroll
            for (int s = 0; s < S; ++s) v[s] = f[s];
            sv_store<S>(band + (size_t)idx * S, v);
        }
    };
    auto store_band = [&](int vertical) {
        for (int idx = tid; idx < nl * N; idx += SARTC_THREADS) {
            const int ll = idx / N, pos = idx - ll * N, line = l0 + ll;
            const int64_t cell = vertical ? (int64_t)line * N + pos : (int64_t)pos * N + line;
            float v[S];
            sv_load<S>(v, band + (size_t)idx * S);
            typename SVec<S>::T t;
            float* f = reinterpret_cast<float*>(&t);
#pragma unroll
            for (int s = 0; s < S; ++s) f[s] = v[s];
            __stcg(reinterpret_cast<typename SVec<S>::T*>(gv + cell * 32), t);
        }
    };

    int orient = p.ang[0].vertical;
    load_band(orient);
    __syncthreads();

    for (int it = 0; it < p.nsub; ++it) {
        const SartcAngle A = p.ang[it];
        if (A.vertical != orient) {
            // transpose through the slice's home in HBM / L2: every CTA writes its band, then reads its band of the other orientation
            store_band(orient);
            __threadfence();                                // the other CTAs read these global stores (through L2: __ldcg) after the barrier
            cluster_sync_sc();
            orient = A.vertical;
            load_band(orient);
            __syncthreads();
        }
        // ---------------- FP: partial sums of every ray over my band ----------------
        const float cth = __cosf(A.theta), sth = __sinf(A.theta);
        float detStep, step, corr1;
        if (A.vertical) { detStep = A.detsize / cth; step = sth / cth; corr1 = (cth < 0.0f) ? -1.0f / cth : 1.0f / cth; }
        else { detStep = -A.detsize / sth; step = cth / sth; corr1 = (sth > 0.0f) ? 1.0f / sth : -1.0f / sth; }
        float* mypart = part + (size_t)(it & 1) * p.NCH * N * S;
        for (int item = tid; item < N * p.NCH; item += SARTC_THREADS) {
            const int ch = item / N, d = item - ch * N;
            const int la = l0 + ch * p.CR, lb = min(l1, la + p.CR);
            float acc[S];
            sv_zero<S>(acc);
            if (la < lb) {
                const float t1 = __fadd_rn(__fadd_rn(__fmaf_rn((float)N, -0.5f, (float)d), 0.5f), -A.offset);
                const float m = __fmul_rn(t1, detStep);
                // position at line `la`: the reference restarts the recurrence every 64 lines and adds `step` per line in between
                float fP;
                {
                    const int base = la & ~63;
                    const float t2 = __fadd_rn(__fmaf_rn((float)N, -0.5f, (float)base), 0.5f);
                    float f = __fmaf_rn(t2, step, m);
                    f = __fmaf_rn((float)N, 0.5f, f);
                    f = __fadd_rn(f, -0.5f);
                    fP = __fadd_rn(f, 0.5f);
                    for (int l = base; l < la; ++l) fP = __fadd_rn(fP, step);
                }
                const float* row = band + (size_t)(la - l0) * N * S;
                for (int line = la; line < lb; ++line) {
                    if ((line & 63) == 0 && line != la) {
                        const float t2 = __fadd_rn(__fmaf_rn((float)N, -0.5f, (float)line), 0.5f);
                        float f = __fmaf_rn(t2, step, m);
                        f = __fmaf_rn((float)N, 0.5f, f);
                        f = __fadd_rn(f, -0.5f);
                        fP = __fadd_rn(f, 0.5f);
                    } else if (line != la) fP = __fadd_rn(fP, step);
                    const int q = __float2int_rd(fP * 256.0f + 0.5f) - 128;   // round half up (ttv_common.cuh)
                    const int i = q >> 8;
                    const float w1 = (float)(q & 255), w0 = 256.0f - w1;
                    float v0[S], v1[S];
                    sv_zero<S>(v0); sv_zero<S>(v1);
                    if (i >= 0 && i < N) sv_load<S>(v0, row + (size_t)i * S);
                    if (i + 1 >= 0 && i + 1 < N) sv_load<S>(v1, row + (size_t)(i + 1) * S);
#pragma unroll
                    for (int s = 0; s < S; ++s) acc[s] = fmaf(w1, v1[s], fmaf(w0, v0[s], acc[s]));
                    row += (size_t)N * S;
                }
            }
            sv_store<S>(mypart + (size_t)item * S, acc);
        }
        cluster_sync_sc();                                  // every CTA's partials of this angle are visible cluster-wide
        // ---------------- weighted residual p, summed in a fixed order through DSMEM; every CTA keeps its own copy ----------------
        {
            const float cs = corr1 * (1.0f / 256.0f);
            for (int d = tid; d < N; d += SARTC_THREADS) {
                float tot[S];
                sv_zero<S>(tot);
                for (int q = 0; q < SARTC_CL; ++q) {
                    const float* rp = cluster.map_shared_rank(mypart, q);
                    for (int ch = 0; ch < p.NCH; ++ch) {
                        float v[S];
                        sv_load<S>(v, rp + ((size_t)ch * N + d) * S);
#pragma unroll
                        for (int s = 0; s < S; ++s) tot[s] += v[s];
                    }
                }
                float bb[S];
                sv_load<S>(bb, p.b + (((int64_t)xb * p.Na + A.a) * N + d) * 32 + xi);
                const float wgt = p.R[(size_t)A.a * N + d];
                float r[S];
#pragma unroll
                for (int s = 0; s < S; ++s) r[s] = fmaf(tot[s], -cs, bb[s]) * wgt;
                sv_store<S>(pw + (size_t)d * S, r);
            }
        }
        __syncthreads();
        // ---------------- BP (Clamp addressing) + relaxation + positivity on my band ----------------
        int bl = tid / N, bpos = tid - bl * N;              // (line, pos) of idx, advanced without a division
        for (int idx = tid; idx < nl * N; idx += SARTC_THREADS) {
            const int pos = bpos, line = l0 + bl;
            bpos += SARTC_THREADS;
            while (bpos >= N) { bpos -= N; ++bl; }
            const int Y = orient ? line : pos, X = orient ? pos : line;
            const float fX = (float)X - 0.5f * (float)N + 0.5f, fY = (float)Y - 0.5f * (float)N + 0.5f;
            const float fT = __fadd_rn(__fmaf_rn(fX, A.sc, -__fmul_rn(fY, A.ss)), A.off);
            const int q = __float2int_rd(fT * 256.0f + 0.5f) - 128;
            const int i = q >> 8;
            const float w1 = (float)(q & 255) * (1.0f / 256.0f), w0 = 1.0f - w1;
            const int i0 = min(max(i, 0), N - 1), i1 = min(max(i + 1, 0), N - 1);
            float v0[S], v1[S], x[S];
            sv_load<S>(v0, pw + (size_t)i0 * S);
            sv_load<S>(v1, pw + (size_t)i1 * S);
            sv_load<S>(x, band + (size_t)idx * S);
#pragma unroll
            for (int s = 0; s < S; ++s) x[s] = fmaxf(fmaf(fmaf(w1, v1[s], w0 * v0[s]), p.relax, x[s]), 0.f);
            sv_store<S>(band + (size_t)idx * S, x);
        }
        __syncthreads();
    }
    store_band(orient);
    // no CTA may leave while another one could still read its partials through DSMEM
    cluster_sync_sc();
}

// slices per cluster that fit (0: the slice does not fit, use the per-angle kernels of ttv_sart.cuh)
inline int sartc_pick_S(int N, int nx, size_t smem_limit, int& NCH, int& CR, size_t& smem) {
    const int RB = (N + SARTC_CL - 1) / SARTC_CL;
    NCH = std::max(1, std::min(SARTC_THREADS / std::max(N, 1), RB));
    CR = (RB + NCH - 1) / NCH;
    NCH = (RB + CR - 1) / CR;
    for (int S : {4, 2, 1}) {
        if (S > 1 && nx < S) continue;
        smem = ((size_t)RB * N + (size_t)2 * NCH * N + (size_t)N) * S * sizeof(float);
        if (smem <= smem_limit) return S;
    }
    return 0;
}

}  // namespace ttv
