// Elementwise, reduction and layout kernels.  sm_100a, HBM-bound, float4 grid-stride.
//
// Restated from the reference's T/gpu/utils/container/matrix_ops.cu:
//   nesterov_momentum_kernel :77-86 (+ tomoengine::fista_nesterov_momentum, tomoengine.cpp:381-384)
//   cuda_positivity_kernel :38-49, cuda_soft_thresholding :64-75, cuda_background_kernel :51-62
//   cuda_norm/l1_norm/sum/rmse/euclidean_dist :89-264 (thrust reductions -> one fused pass here)
// Padding slices (x >= Nx in the last x-block) are zero in every volume, so whole-buffer
// reductions need no masking.
#pragma once
#include "ttv_common.cuh"

namespace ttv {

constexpr int EW_THREADS = 256;
inline int ew_blocks(size_t n4) {
    size_t b = (n4 + EW_THREADS - 1) / EW_THREADS;
    const size_t cap = 148 * 16;
    return (int)(b < cap ? (b ? b : 1) : cap);
}

// recon <- yk ; yk <- recon + beta*(recon - recon_old) ; recon_old <- recon     (tomoengine.cpp:381-384)
// 4V of traffic instead of the 5V of three stores: y is written over recon_old in place and the caller swaps
// the yk / recon_old pointers afterwards (the old yk buffer IS the new recon_old); only recon needs a copy,
// because tv_fgp overwrites it while recon_old must keep x for the next step.
__global__ void __launch_bounds__(EW_THREADS) momentum_kernel(float4* __restrict__ recon, const float4* __restrict__ yk, float4* __restrict__ old_to_y, float beta, size_t n4)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 x = yk[i], o = old_to_y[i];
        float4 y;
        y.x = fmaf(beta, x.x - o.x, x.x); y.y = fmaf(beta, x.y - o.y, x.y);
        y.z = fmaf(beta, x.z - o.z, x.z); y.w = fmaf(beta, x.w - o.w, x.w);
        recon[i] = x; old_to_y[i] = y;
    }
}

enum EwOp { EW_POS = 0, EW_SOFT = 1, EW_SCALE = 2, EW_BACKGROUND = 3, EW_SOFT_POS = 4 };

template <int OP>
__global__ void __launch_bounds__(EW_THREADS) ew_kernel(float4* __restrict__ v, float s, size_t n4)
{
    auto f = [&](float a) -> float {
        if (OP == EW_POS) return a < 0.0f ? 0.0f : a;
        if (OP == EW_SOFT || OP == EW_SOFT_POS) {
            // signbit(lambda-|x|) * copysign(|x|-lambda, x)
            const float m = fabsf(a);
            float r = (signbit(s - m) ? 1.0f : 0.0f) * copysignf(m - s, a);
            if (OP == EW_SOFT_POS) r = r < 0.0f ? 0.0f : r;
            return r;
        }
        if (OP == EW_SCALE) return a * s;
        return a == 0.0f ? s : a;
    };
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 a = v[i];
        a.x = f(a.x); a.y = f(a.y); a.z = f(a.z); a.w = f(a.w);
        v[i] = a;
    }
}

__global__ void __launch_bounds__(EW_THREADS) div_kernel(float4* __restrict__ v, float s, size_t n4)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 a = v[i];
        a.x = a.x / s; a.y = a.y / s; a.z = a.z / s; a.w = a.w / s;
        v[i] = a;
    }
}

enum RedOp { RED_SQDIST = 0, RED_SUMSQ = 1, RED_SUM = 2, RED_L1 = 3 };

template <int OP>
__global__ void __launch_bounds__(EW_THREADS) reduce_kernel(const float4* __restrict__ a, const float4* __restrict__ b, size_t n4, double* out)
{
    double acc = 0.0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 x = a[i];
        float p;
        if (OP == RED_SQDIST) {
            const float4 y = b[i];
            const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
            p = d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
        } else if (OP == RED_SUMSQ) p = x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
        else if (OP == RED_SUM) p = x.x + x.y + x.z + x.w;
        else p = fabsf(x.x) + fabsf(x.y) + fabsf(x.z) + fabsf(x.w);
        acc += (double)p;
    }
    acc = warp_sum(acc);
    __shared__ double sh[EW_THREADS / 32];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < EW_THREADS / 32; ++i) s += sh[i];
        atomicAdd(out, s);
    }
}

// max over a buffer (values may be negative): ordered-int atomicMax
__device__ __forceinline__ int float_ordered(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__global__ void __launch_bounds__(EW_THREADS) max_kernel(const float4* __restrict__ a, size_t n4, int* out)
{
    float m = -INFINITY;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 x = a[i];
        m = fmaxf(m, fmaxf(fmaxf(x.x, x.y), fmaxf(x.z, x.w)));
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) atomicMax(out, float_ordered(m));
}

// ---------------------------------------------------------------------------------------------
// Layout conversion between the reference's host layout  src[x][m]  (m = j*N+k for volumes,
// a*Nd+d for sinograms; x in [x_begin, x_begin+nx_src) of the local slab) and the
// x-interleaved device layout  dst[(xb*M + m)*32 + xi].  32x32 shared-memory transpose.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_kernel(const float* __restrict__ src, float* __restrict__ dst, int nx, int64_t M, int x_first, int nx_src)
{
    __shared__ float tile[32][33];
    const int xb = blockIdx.y;
    const int64_t m0 = (int64_t)blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // ty 0..7
    // read: tile[xi][mm] = src[x][m0+mm]   (coalesced along m)
    for (int r = ty; r < 32; r += 8) {
        const int x = xb * 32 + r;           // local slice
        const int xs = x - x_first;          // index into src
        const int64_t m = m0 + tx;
        float v = 0.0f;
        if (x < nx && xs >= 0 && xs < nx_src && m < M) v = src[(int64_t)xs * M + m];
        tile[r][tx] = v;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int64_t m = m0 + r;
        if (m < M) dst[((int64_t)xb * M + m) * 32 + tx] = tile[tx][r];
    }
}

__global__ void __launch_bounds__(256) unpack_kernel(const float* __restrict__ src, float* __restrict__ dst, int nx, int64_t M, int x_first, int nx_dst)
{
    __shared__ float tile[32][33];
    const int xb = blockIdx.y;
    const int64_t m0 = (int64_t)blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) {
        const int64_t m = m0 + r;
        tile[r][tx] = (m < M) ? src[((int64_t)xb * M + m) * 32 + tx] : 0.0f;   // tile[mm][xi]
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int x = xb * 32 + r;
        const int xd = x - x_first;
        const int64_t m = m0 + tx;
        if (x < nx && xd >= 0 && xd < nx_dst && m < M) dst[(int64_t)xd * M + m] = tile[tx][r];
    }
}

// one slice <-> plain [M] plane (halo exchange, get_recon/set_recon of a single slice)
__global__ void __launch_bounds__(256) gather_slice_kernel(const float* __restrict__ src, float* __restrict__ plane, int x, int64_t M)
{
    const int xb = x >> 5, xi = x & 31;
    for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x)
        plane[m] = src[((int64_t)xb * M + m) * 32 + xi];
}
// up to 4 slices -> planes in one launch (halo exchange of the FGP dual variables); blockIdx.y selects the job
// dst element stride ds[q] / element offset de[q]: 1/0 = plain plane, 4/e = ghost cells [m][4] (ttv_fgp.cuh)
struct GatherJobs { const float* src[4]; float* dst[4]; int x[4]; int ds[4]; int de[4]; int n; };
__global__ void __launch_bounds__(256) gather_slices_kernel(const GatherJobs g, int64_t M)
{
    const int q = blockIdx.y;
    const float* __restrict__ src = g.src[q];
    float* __restrict__ plane = g.dst[q];
    const int xb = g.x[q] >> 5, xi = g.x[q] & 31, ds = g.ds[q], de = g.de[q];
    for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x)
        plane[m * ds + de] = src[((int64_t)xb * M + m) * 32 + xi];
}
__global__ void __launch_bounds__(256) scatter_slice_kernel(float* __restrict__ dst, const float* __restrict__ plane, int x, int64_t M)
{
    const int xb = x >> 5, xi = x & 31;
    for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < M; m += (int64_t)gridDim.x * blockDim.x)
        dst[((int64_t)xb * M + m) * 32 + xi] = plane[m];
}

__global__ void __launch_bounds__(256) fill_kernel(float* __restrict__ v, float s, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) v[i] = s;
}

// fill only the valid slices (x < nx) of an interleaved buffer; padding stays zero
__global__ void __launch_bounds__(256) fill_valid_kernel(float* __restrict__ v, float s, int nx, int64_t M, int nxb)
{
    const int64_t total = (int64_t)nxb * M * 32;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int xi = (int)(i & 31);
        const int xb = (int)(i / (M * 32));
        v[i] = (xb * 32 + xi < nx) ? s : 0.0f;
    }
}

// per-angle __cosf/__sinf table exactly as the projector evaluates them (parity tests feed these
// to the CPU oracle so the comparison is not dominated by the fast-intrinsic error)
__global__ void trig_kernel(const float* __restrict__ theta, float* __restrict__ c, float* __restrict__ s, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { c[i] = __cosf(theta[i]); s[i] = __sinf(theta[i]); }
}

}  // namespace ttv
