// 3-D total-variation kernels on the x-interleaved layout.  sm_100a, HBM-bound.
//
// Arithmetic restated from the reference's tomofusion CUDA regularisers:
//   TV value   tv_3D_kernel                     T/gpu/utils/regularizers/tv_gd.cu:27-47
//   TV-FGP     Obj/Grad/Proj kernels            T/gpu/utils/regularizers/tv_fgp.cu:44-115, host loop :192-281
//   TV-GD      tv_gradient_3D_kernel (+quirks)  T/gpu/utils/regularizers/tv_gd.cu:52-122, host loop :141-218
//
// A warp owns one row segment (fixed x-block and j, a run of k); its 32 lanes are the 32 slices
// of a cell, so every global access is one 128-byte line and the x+-1 neighbours come from
// warp shuffles (only the two edge lanes touch another x-block or a halo plane).  The FGP sweep
// fuses the reference's four kernels (Obj_func, nonneg, Grad_func, Proj_func) into one pass
// that reads in,P1,P2,P3 once and writes P1,P2,P3 once.
//
// Multi-GPU: a rank holds a slab of slices; `xoff`/`nxg` give its position in the global
// volume.  Neighbour boundary slices arrive either as plain planes ([j][k], TV-GD) or as ghost
// cells ([j][k][4] floats = the 16-byte TMA box of ttv_fgp.cuh: the right neighbour's slice 0 in
// element 0, the left neighbour's last slice in element 3); see DESIGN.md.
#pragma once
#include "ttv_common.cuh"

namespace ttv {

constexpr int TV_WARPS = 8;        // rows j per CTA
constexpr int TV_KSEG = 32;        // k cells marched per warp

struct VolDims {
    int nx;        // local slices
    int nxb;       // local x-blocks = ceil(nx/32)
    int N;         // ny = nz
    int xoff;      // global index of local slice 0
    int nxg;       // global number of slices
};

// value of slice x-1 for every lane: shuffle, lane 0 reads the previous block / left halo plane.
// `line` is this lane's value at the same (j,k) cell; ptr/cell address the same array.
__device__ __forceinline__ float x_minus1(float line, const float* __restrict__ ptr, int64_t cell, int xb, int64_t cellsPerBlock,
                                           const float* __restrict__ halo_left, int64_t jk, int lane)
{
    float v = __shfl_up_sync(0xffffffffu, line, 1);
    if (lane == 0) {
        if (xb > 0) v = ptr[(cell - cellsPerBlock) * TTV_XI + 31];
        else v = halo_left ? halo_left[jk * 4 + 3] : 0.0f;
    }
    return v;
}

// ---------------------------------------------------------------------------------------------
// TV value: sum sqrt(eps + dx^2 + dy^2 + dz^2), periodic forward differences.
// halo_next: slice (x_last+1) mod nxg when the rank does not own it (NULL on one GPU).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tv_value_kernel(const float* __restrict__ v, VolDims d, const float* __restrict__ halo_next, int hs, double* out)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int N = d.N;
    const int64_t cpb = (int64_t)N * N;
    const int64_t row = (int64_t)blockIdx.x * 8 + w;      // (xb, j)
    const int xb = (int)(row / N), j = (int)(row % N);
    double tot = 0.0;
    if (xb < d.nxb) {
        const int x = xb * 32 + lane;
        const bool valid = x < d.nx;
        const int jp = (j + 1) % N;
        for (int k = 0; k < N; ++k) {
            const int64_t cell = (int64_t)xb * cpb + (int64_t)j * N + k;
            const float c = v[cell * 32 + lane];
            const float vj = v[((int64_t)xb * cpb + (int64_t)jp * N + k) * 32 + lane];
            const float vk = v[((int64_t)xb * cpb + (int64_t)j * N + (k + 1) % N) * 32 + lane];
            float vx = __shfl_down_sync(0xffffffffu, c, 1);
            if (valid) {
                if (x == d.nx - 1) {
                    // last local slice: neighbour is global (x+1) mod nxg
                    if (halo_next) vx = halo_next[((int64_t)j * N + k) * hs];
                    else vx = v[((int64_t)j * N + k) * 32 + 0];                 // wraps to local slice 0
                } else if (lane == 31) {
                    vx = v[(cell + cpb) * 32 + 0];
                }
                const float dx = c - vx, dy = c - vj, dz = c - vk;
                tot += (double)sqrtf(1e-6f + dx * dx + dy * dy + dz * dz);
            }
        }
    }
    tot = warp_sum(tot);
    __shared__ double sh[8];
    if (lane == 0) sh[w] = tot;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < 8; ++i) s += sh[i];
        atomicAdd(out, s);
    }
}

// ---------------------------------------------------------------------------------------------
// TV-FGP sweep.
//   D(p)  = max(0, in(p) - lambda*(P1+P2+P3 - P1(x-1) - P2(j-1) - P3(k-1)))        tv_fgp.cu:44-65,143-154
//   Pn(p) = proj( P + multip * (D(p) - D(p+e)) ), zero difference on the high edge  tv_fgp.cu:67-115
// FIRST: P == 0 (first sweep, no P loads).  Ghost cells for a slab (NULL on a single GPU), [j][k][4]:
//   hl_P1: P1 of global slice xoff-1 (element 3);  hr_in/hr_P1/hr_P2/hr_P3: slice xoff+nx of in and P (element 0).
// ---------------------------------------------------------------------------------------------
struct FgpParams {
    const float* in;
    const float* P1; const float* P2; const float* P3;     // old
    float* Q1; float* Q2; float* Q3;                         // new
    const float* hl_P1;
    const float* hr_in; const float* hr_P1; const float* hr_P2; const float* hr_P3;
    float lambda, multip;
    VolDims d;
    // peer-memory hand-off of a sharded volume (ttv_fgp.cuh, CUDA IPC): with compact != 0 the P ghosts
    // hl_P1 / hr_P1..3 are compact planes [N][pitch] that the NEIGHBOUR's sweep kernel wrote into this
    // GPU's memory, and this launch writes its own boundary slices into the neighbours' planes rm_*.
    int compact, pitch;
    float* rm_gl1;                                   // on the right neighbour: P1 of my last slice
    float* rm_gr1; float* rm_gr2; float* rm_gr3;     // on the left neighbour: P1,P2,P3 of my first slice
    unsigned* rm_flagR; unsigned* rm_flagL;          // epoch flags on the right / left neighbour
    unsigned* cnt;                                   // local CTA counters [2] (right, left)
    unsigned epoch;
};

template <bool FIRST>
__device__ __forceinline__ float fgp_D(const FgpParams& p, int64_t cell, int xb, int j, int k, int lane, float p1c, float p2c, float p3c, float inc)
{
    if (FIRST) return fmaxf(inc, 0.0f);
    const int N = p.d.N;
    const int64_t cpb = (int64_t)N * N;
    const float v1 = x_minus1(p1c, p.P1, cell, xb, cpb, p.hl_P1, (int64_t)j * N + k, lane);
    const float v2 = j > 0 ? p.P2[(cell - N) * 32 + lane] : 0.0f;
    const float v3 = k > 0 ? p.P3[(cell - 1) * 32 + lane] : 0.0f;
    const float s = ((((p1c + p2c) + p3c) - v1) - v2) - v3;
    return fmaxf(__fmaf_rn(-p.lambda, s, inc), 0.0f);
}

// D of the slice to the right of lane's slice at the same (j,k): only evaluated by the edge lane.
template <bool FIRST>
__device__ __forceinline__ float fgp_D_xplus(const FgpParams& p, int64_t cell, int xb, int j, int k, bool from_halo, float p1_own)
{
    const int N = p.d.N;
    const int64_t cpb = (int64_t)N * N;
    const int64_t jk = (int64_t)j * N + k;
    float inn, q1 = 0.f, q2 = 0.f, q3 = 0.f, v2 = 0.f, v3 = 0.f;
    if (from_halo) {
        inn = p.hr_in[jk * 4];
        if (!FIRST) {
            q1 = p.hr_P1[jk * 4]; q2 = p.hr_P2[jk * 4]; q3 = p.hr_P3[jk * 4];
            v2 = j > 0 ? p.hr_P2[(jk - N) * 4] : 0.f;
            v3 = k > 0 ? p.hr_P3[(jk - 1) * 4] : 0.f;
        }
    } else {
        const int64_t c2 = (cell + cpb) * 32;   // lane 0 of the next block
        inn = p.in[c2];
        if (!FIRST) {
            q1 = p.P1[c2]; q2 = p.P2[c2]; q3 = p.P3[c2];
            v2 = j > 0 ? p.P2[c2 - (int64_t)N * 32] : 0.f;
            v3 = k > 0 ? p.P3[c2 - 32] : 0.f;
        }
    }
    if (FIRST) return fmaxf(inn, 0.0f);
    const float s = ((((q1 + q2) + q3) - p1_own) - v2) - v3;
    return fmaxf(__fmaf_rn(-p.lambda, s, inn), 0.0f);
}

template <bool FIRST>
__global__ void __launch_bounds__(TV_WARPS * 32) fgp_sweep_kernel(const FgpParams p)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int N = p.d.N;
    const int64_t cpb = (int64_t)N * N;
    const int j = blockIdx.y * TV_WARPS + w;
    const int xb = blockIdx.z;
    const int k0 = blockIdx.x * TV_KSEG;
    if (j >= N) return;
    const int x = xb * 32 + lane;
    const bool valid = x < p.d.nx;
    const bool lastx = (x == p.d.nx - 1);
    const bool glast = lastx && (p.hr_in == nullptr);        // true last slice of the global volume
    const bool edge = valid && !glast && (lastx || lane == 31);   // lane whose x+1 is not lane+1
    const int kend = min(k0 + TV_KSEG, N);

    auto ld = [&](const float* a, int64_t cell) { return a[cell * 32 + lane]; };

    int64_t cell = (int64_t)xb * cpb + (int64_t)j * N + k0;
    float in_c = ld(p.in, cell), p1 = 0.f, p2 = 0.f, p3 = 0.f;
    if (!FIRST) { p1 = ld(p.P1, cell); p2 = ld(p.P2, cell); p3 = ld(p.P3, cell); }
    float Dc = fgp_D<FIRST>(p, cell, xb, j, k0, lane, p1, p2, p3, in_c);

    for (int k = k0; k < kend; ++k, ++cell) {
        // D at (k+1)
        float in_n = 0.f, n1 = 0.f, n2 = 0.f, n3 = 0.f, Dk = 0.f;
        if (k + 1 < N) {
            in_n = ld(p.in, cell + 1);
            if (!FIRST) { n1 = ld(p.P1, cell + 1); n2 = ld(p.P2, cell + 1); n3 = ld(p.P3, cell + 1); }
            Dk = fgp_D<FIRST>(p, cell + 1, xb, j, k + 1, lane, n1, n2, n3, in_n);
        }
        // D at (j+1)
        float Dj = 0.f;
        if (j + 1 < N) {
            const int64_t cj = cell + N;
            const float ij = ld(p.in, cj);
            float a1 = 0.f, a2 = 0.f, a3 = 0.f;
            if (!FIRST) { a1 = ld(p.P1, cj); a2 = ld(p.P2, cj); a3 = ld(p.P3, cj); }
            Dj = fgp_D<FIRST>(p, cj, xb, j + 1, k, lane, a1, a2, a3, ij);
        }
        // D at (x+1)
        float Dx = __shfl_down_sync(0xffffffffu, Dc, 1);
        if (edge) Dx = fgp_D_xplus<FIRST>(p, cell, xb, j, k, lastx, p1);

        const float val1 = (valid && !glast) ? Dc - Dx : 0.0f;
        const float val2 = (j < N - 1) ? Dc - Dj : 0.0f;
        const float val3 = (k < N - 1) ? Dc - Dk : 0.0f;
        float q1 = __fmaf_rn(p.multip, val1, p1);
        float q2 = __fmaf_rn(p.multip, val2, p2);
        float q3 = __fmaf_rn(p.multip, val3, p3);
        const float den = __fmaf_rn(q3, q3, __fmaf_rn(q2, q2, __fmul_rn(q1, q1)));
        if (den > 1.0f) {
            const float sq = __fdiv_rn(1.0f, __fsqrt_rn(den));
            q1 *= sq; q2 *= sq; q3 *= sq;
        }
        if (!valid) { q1 = q2 = q3 = 0.f; }
        p.Q1[cell * 32 + lane] = q1;
        p.Q2[cell * 32 + lane] = q2;
        p.Q3[cell * 32 + lane] = q3;
        // advance
        Dc = Dk; p1 = n1; p2 = n2; p3 = n3; in_c = in_n;
    }
}

// TV value, tiled: CTA = one x-block x 8 rows x TV_KSEG cells; a warp marches along k carrying the centre
// value (its k+1 neighbour becomes the next centre), the j+1 row is an L1/L2 hit from the warp below.
// Same per-voxel arithmetic as tv_value_kernel (periodic forward differences, tv_gd.cu:27-47).
__global__ void __launch_bounds__(256) tv_value_tiled_kernel(const float* __restrict__ v, VolDims d, const float* __restrict__ halo_next, int hs, double* out)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int N = d.N;
    const int64_t cpb = (int64_t)N * N;
    const int j = blockIdx.y * 8 + w;
    const int xb = blockIdx.z;
    const int k0 = blockIdx.x * TV_KSEG;
    float tot = 0.0f;
    double dtot = 0.0;
    if (j < N) {
        const int x = xb * 32 + lane;
        const bool valid = x < d.nx;
        const bool lastx = (x == d.nx - 1);
        const int jp = (j + 1 == N) ? 0 : j + 1;
        const int kend = min(k0 + TV_KSEG, N);
        const float* rowc = v + ((int64_t)xb * cpb + (int64_t)j * N) * 32 + lane;
        const float* rowj = v + ((int64_t)xb * cpb + (int64_t)jp * N) * 32 + lane;
        // x+1 neighbour of the lane whose successor is not lane+1: next x-block, halo slice, or periodic wrap
        const float* ep = nullptr; int64_t es = 0;
        if (valid) {
            if (lastx) {
                if (halo_next) { ep = halo_next + (int64_t)j * N * hs; es = hs; }
                else { ep = v + (int64_t)j * N * 32; es = 32; }
            } else if (lane == 31) { ep = v + (((int64_t)(xb + 1)) * cpb + (int64_t)j * N) * 32; es = 32; }
        }
        float c = rowc[(int64_t)k0 * 32];
        for (int kb = k0; kb < kend; kb += 8) {
            // all line loads of 8 cells in flight before any arithmetic
            float vk[8], vj[8], ve[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int k = kb + u;
                const int kp = (k + 1 >= N) ? 0 : k + 1;
                vk[u] = (k < kend) ? rowc[(int64_t)kp * 32] : 0.0f;
                vj[u] = (k < kend) ? rowj[(int64_t)k * 32] : 0.0f;
                ve[u] = (ep && k < kend) ? ep[(int64_t)k * es] : 0.0f;
            }
            float part = 0.0f;                       // 8 voxels in float, then one double add (F2F/DADD are quarter-rate)
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int k = kb + u;
                float vx = __shfl_down_sync(0xffffffffu, c, 1);
                if (ep) vx = ve[u];
                if (k < kend && valid) {
                    const float dx = c - vx, dy = c - vj[u], dz = c - vk[u];
                    part += sqrtf(1e-6f + dx * dx + dy * dy + dz * dz);
                }
                c = vk[u];
            }
            dtot += (double)part;
        }
    }
    (void)tot;
    dtot = warp_sum(dtot);
    __shared__ double sh[8];
    if (lane == 0) sh[w] = dtot;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s2 = 0;
        for (int i = 0; i < 8; ++i) s2 += sh[i];
        atomicAdd(out, s2);
    }
}

// last sweep: only D is needed (tv_fgp.cu:247,272): out = max(0, in - lambda*div P)
__global__ void __launch_bounds__(256) fgp_final_kernel(const FgpParams p, float* __restrict__ out)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int N = p.d.N;
    const int64_t cpb = (int64_t)N * N;
    const int j = blockIdx.y * 8 + w;
    const int xb = blockIdx.z;
    const int k0 = blockIdx.x * TV_KSEG;
    if (j >= N) return;
    const bool valid = xb * 32 + lane < p.d.nx;
    const int kend = min(k0 + TV_KSEG, N);
    for (int k = k0; k < kend; ++k) {
        const int64_t cell = (int64_t)xb * cpb + (int64_t)j * N + k;
        const float inc = p.in[cell * 32 + lane];
        float D;
        if (p.P1 == nullptr) D = fmaxf(inc, 0.f);
        else D = fgp_D<false>(p, cell, xb, j, k, lane, p.P1[cell * 32 + lane], p.P2[cell * 32 + lane], p.P3[cell * 32 + lane], inc);
        out[cell * 32 + lane] = valid ? D : 0.0f;
    }
}

// tv_fgp(iter = 0): the reference returns the TV *map* in place of the volume (tv_fgp.cu:232,272)
__global__ void __launch_bounds__(256) tv_map_kernel(const float* __restrict__ v, float* __restrict__ out, VolDims d, const float* __restrict__ halo_next, int hs)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int N = d.N;
    const int64_t cpb = (int64_t)N * N;
    const int64_t row = (int64_t)blockIdx.x * 8 + w;
    const int xb = (int)(row / N), j = (int)(row % N);
    if (xb >= d.nxb) return;
    const int x = xb * 32 + lane;
    const bool valid = x < d.nx;
    const int jp = (j + 1) % N;
    for (int k = 0; k < N; ++k) {
        const int64_t cell = (int64_t)xb * cpb + (int64_t)j * N + k;
        const float c = v[cell * 32 + lane];
        const float vj = v[((int64_t)xb * cpb + (int64_t)jp * N + k) * 32 + lane];
        const float vk = v[((int64_t)xb * cpb + (int64_t)j * N + (k + 1) % N) * 32 + lane];
        float vx = __shfl_down_sync(0xffffffffu, c, 1);
        float r = 0.f;
        if (valid) {
            if (x == d.nx - 1) vx = halo_next ? halo_next[((int64_t)j * N + k) * hs] : v[((int64_t)j * N + k) * 32];
            else if (lane == 31) vx = v[(cell + cpb) * 32];
            const float dx = c - vx, dy = c - vj, dz = c - vk;
            r = sqrtf(1e-6f + dx * dx + dy * dy + dz * dz);
        }
        out[cell * 32 + lane] = r;
    }
}

// ---------------------------------------------------------------------------------------------
// TV-GD gradient, with the reference's index quirks (tv_gd.cu:61-79):
//   ip_jm == (x, j-1, k);  off(k) = (k-1-nz)%nz with C truncation, so for k != 1
//   jp_km -> (x, j, k-1 mod nz) [j<ny-1] or (x-1, ny-1, k-1 mod nz) [j=ny-1, out of range -> 0 at global x=0]
//   ip_km -> (ip, j-1, k-1 mod nz) [j>=1] or (ip-1, ny-1, k-1 mod nz) [j=0; out of range -> 0 when ip wraps to 0]
//   and for k == 1 they are the intended (x, j+1, 0) / (ip, j, 0).
// Fused with the sum of squares of the gradient (the reference copies g into a thrust vector first).
// Halo planes (NULL on one GPU, where the volume wraps onto itself): hm = global slice xoff-1,
// hp = global slice xoff+nx, both modulo nxg.
// ---------------------------------------------------------------------------------------------
struct GdParams {
    const float* x; float* g; const float* hm; const float* hp; double* sumsq; VolDims d;
};

// One voxel of tv_gradient_3D_kernel (tv_gd.cu:83-103) with the floating-point contraction the reference's own build has
// (read from the SASS of oracle/_ref, nvcc default -fmad=true): every sqrt argument is the left-to-right chain
// fma(d3, d3, fma(d2, d2, fma(d1, d1, eps))), the numerator of the first term is evaluated in double (3.0 * c is a double
// product in the source), quotients are IEEE divisions added left to right.  Written with explicit intrinsics so that every
// kernel using it rounds identically whatever the surrounding code looks like.
__device__ __forceinline__ float tvgd_voxel(float c, float x_ip, float l_jp, float l_kp, float x_im, float x_imjp, float x_imkp,
                                            float l_jm, float l_jmkp, float l_km, float x_ipkm, float x_jpkm)
{
    const float eps = 1e-6f;
    auto ss3 = [&](float d1, float d2, float d3) { return __fsqrt_rn(__fmaf_rn(d3, d3, __fmaf_rn(d2, d2, __fmaf_rn(d1, d1, eps)))); };
    const float v1n = (float)(((3.0 * (double)c - (double)x_ip) - (double)l_jp) - (double)l_kp);
    const float v1d = ss3(__fsub_rn(c, x_ip), __fsub_rn(c, l_jp), __fsub_rn(c, l_kp));
    const float v2n = __fsub_rn(c, x_im);
    const float v2d = ss3(__fsub_rn(x_im, c), __fsub_rn(x_im, x_imjp), __fsub_rn(x_im, x_imkp));
    const float v3n = __fsub_rn(c, l_jm);
    const float v3d = ss3(__fsub_rn(l_jm, l_jm), __fsub_rn(l_jm, c), __fsub_rn(l_jm, l_jmkp));      // ip_jm is (x, j-1, k) itself (tv_gd.cu:73)
    const float v4n = __fsub_rn(c, l_km);
    const float v4d = ss3(__fsub_rn(l_km, x_ipkm), __fsub_rn(l_km, x_jpkm), __fsub_rn(l_km, c));
    return __fadd_rn(__fadd_rn(__fadd_rn(__fdiv_rn(v1n, v1d), __fdiv_rn(v2n, v2d)), __fdiv_rn(v3n, v3d)), __fdiv_rn(v4n, v4d));
}

__global__ void __launch_bounds__(256) tvgd_gradient_kernel(const GdParams p)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int N = p.d.N, nx = p.d.nx;
    const int64_t cpb = (int64_t)N * N;
    const int j = blockIdx.y * 8 + w;
    const int xb = blockIdx.z;
    const int k0 = blockIdx.x * TV_KSEG;
    const float eps = 1e-6f;
    double tot = 0.0;
    if (j < N) {
        const int xl = xb * 32 + lane;
        const bool valid = xl < nx;
        const int xg = p.d.xoff + xl;                       // global slice index
        const bool first_g = (xg == 0), last_g = (xg == p.d.nxg - 1);
        const int jp = (j + 1) % N, jm = (j - 1 + N) % N;
        const int kend = min(k0 + TV_KSEG, N);
        // value of array at (slice xl-1, j', k') given this lane's line value at (j',k')
        auto xm = [&](float line, int jj, int kk) -> float {
            float v = __shfl_up_sync(0xffffffffu, line, 1);
            if (lane == 0) {
                if (xb > 0) v = p.x[(((int64_t)(xb - 1)) * cpb + (int64_t)jj * N + kk) * 32 + 31];
                else if (p.hm) v = p.hm[(int64_t)jj * N + kk];
                else v = p.x[(((int64_t)((nx - 1) >> 5)) * cpb + (int64_t)jj * N + kk) * 32 + ((nx - 1) & 31)];   // wrap to last slice
            }
            return v;
        };
        auto xp = [&](float line, int jj, int kk) -> float {
            float v = __shfl_down_sync(0xffffffffu, line, 1);
            if (valid) {
                if (xl == nx - 1) v = p.hp ? p.hp[(int64_t)jj * N + kk] : p.x[((int64_t)jj * N + kk) * 32];      // wrap to slice 0
                else if (lane == 31) v = p.x[(((int64_t)(xb + 1)) * cpb + (int64_t)jj * N + kk) * 32];
            }
            return v;
        };
        auto L = [&](int jj, int kk) -> float { return p.x[((int64_t)xb * cpb + (int64_t)jj * N + kk) * 32 + lane]; };
        for (int k = k0; k < kend; ++k) {
            const int kp = (k + 1) % N, km = (k - 1 + N) % N;
            const float c = L(j, k);
            const float l_jp = L(jp, k), l_jm = L(jm, k), l_kp = L(j, kp), l_km = L(j, km), l_jmkp = L(jm, kp);
            const float x_ip = xp(c, j, k), x_im = xm(c, j, k);
            const float x_imjp = xm(l_jp, jp, k), x_imkp = xm(l_kp, j, kp);
            // jp_km
            float x_jpkm;
            if (k == 1) x_jpkm = L(jp, 0);
            else if (j < N - 1) x_jpkm = l_km;
            else { const float t = xm(l_km, j, km); x_jpkm = first_g ? 0.0f : t; }
            // ip_km
            float x_ipkm;
            if (k == 1) { const float t = xp(L(j, 0), j, 0); x_ipkm = t; }
            else if (j >= 1) { const float t = xp(L(jm, km), jm, km); x_ipkm = t; }
            else { x_ipkm = last_g ? 0.0f : L(N - 1, km); }     // (ip-1, ny-1, km) == own slice
            float g = tvgd_voxel(c, x_ip, l_jp, l_kp, x_im, x_imjp, x_imkp, l_jm, l_jmkp, l_km, x_ipkm, x_jpkm);
            if (!valid) g = 0.0f;
            p.g[((int64_t)xb * cpb + (int64_t)j * N + k) * 32 + lane] = g;
            tot += (double)(g * g);
        }
    }
    tot = warp_sum(tot);
    __shared__ double sh[8];
    if (lane == 0) sh[w] = tot;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < 8; ++i) s += sh[i];
        atomicAdd(p.sumsq, s);
    }
}

// Same gradient, marching form: a warp walks its row segment along k and keeps the three rows (j-1, j, j+1) of the three
// columns (k-1, k, k+1) plus their x-neighbour values in registers, so a step loads ONE new column (3 lines + the edge lanes'
// neighbour-block / halo values) instead of re-reading 7 lines and re-shuffling 6 values per voxel.  Every voxel is evaluated
// with exactly the expressions of tvgd_gradient_kernel (same operands, same order): the two kernels are bit-identical
// (tests/test_gpu_parity.py), and so is the comparison with the reference at realistic sizes (tests/golden C_gd10).
__global__ void __launch_bounds__(256) tvgd_gradient_march_kernel(const GdParams p)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int N = p.d.N, nx = p.d.nx;
    const int64_t cpb = (int64_t)N * N;
    const int j = blockIdx.y * 8 + w;
    const int xb = blockIdx.z;
    const int k0 = blockIdx.x * TV_KSEG;
    const float eps = 1e-6f;
    const float* __restrict__ x = p.x;
    float* __restrict__ g = p.g;
    double tot = 0.0;
    if (j < N) {
        const int xl = xb * 32 + lane;
        const bool valid = xl < nx;
        const int xg = p.d.xoff + xl;
        const bool first_g = (xg == 0), last_g = (xg == p.d.nxg - 1);
        const int jp = (j + 1) % N, jm = (j - 1 + N) % N;
        const int kend = min(k0 + TV_KSEG, N);
        // lane 0's slice x-1 and the edge lane's slice x+1 live in another x-block, a halo plane or (periodic) this volume's far end
        const float* lp; int64_t ls;
        if (xb > 0) { lp = x + ((int64_t)(xb - 1)) * cpb * 32 + 31; ls = 32; }
        else if (p.hm) { lp = p.hm; ls = 1; }
        else { lp = x + ((int64_t)((nx - 1) >> 5)) * cpb * 32 + ((nx - 1) & 31); ls = 32; }
        const bool redge = valid && (xl == nx - 1 || lane == 31);
        const float* rp; int64_t rs;
        if (xl == nx - 1) { if (p.hp) { rp = p.hp; rs = 1; } else { rp = x; rs = 32; } }
        else { rp = x + ((int64_t)(xb + 1)) * cpb * 32; rs = 32; }
        const float* rowA = x + ((int64_t)xb * cpb + (int64_t)j * N) * 32 + lane;
        const float* rowB = x + ((int64_t)xb * cpb + (int64_t)jm * N) * 32 + lane;
        const float* rowC = x + ((int64_t)xb * cpb + (int64_t)jp * N) * 32 + lane;
        const int64_t oA = (int64_t)j * N, oB = (int64_t)jm * N, oC = (int64_t)jp * N;

        // one column: the three lines and the x-neighbour values derived from them
        struct Col { float A, B, C, xmA, xpA, xpB, xmC; };
        auto load_col = [&](int kk) {
            Col q;
            q.A = rowA[(int64_t)kk * 32]; q.B = rowB[(int64_t)kk * 32]; q.C = rowC[(int64_t)kk * 32];
            q.xmA = __shfl_up_sync(0xffffffffu, q.A, 1);
            q.xmC = __shfl_up_sync(0xffffffffu, q.C, 1);
            q.xpA = __shfl_down_sync(0xffffffffu, q.A, 1);
            q.xpB = __shfl_down_sync(0xffffffffu, q.B, 1);
            if (lane == 0) { q.xmA = lp[(oA + kk) * ls]; q.xmC = lp[(oC + kk) * ls]; }
            if (redge) { q.xpA = rp[(oA + kk) * rs]; q.xpB = rp[(oB + kk) * rs]; }
            return q;
        };
        Col cm = load_col((k0 - 1 + N) % N), cc = load_col(k0);
#pragma unroll 2
        for (int k = k0; k < kend; ++k) {
            const Col cp = load_col((k + 1) % N);
            const float c = cc.A, l_jp = cc.C, l_jm = cc.B, l_kp = cp.A, l_km = cm.A, l_jmkp = cp.B;
            const float x_ip = cc.xpA, x_im = cc.xmA, x_imjp = cc.xmC, x_imkp = cp.xmA;
            float x_jpkm;
            if (k == 1) x_jpkm = cm.C;
            else if (j < N - 1) x_jpkm = l_km;
            else x_jpkm = first_g ? 0.0f : cm.xmA;
            float x_ipkm;
            if (k == 1) x_ipkm = cm.xpA;
            else if (j >= 1) x_ipkm = cm.xpB;
            else x_ipkm = last_g ? 0.0f : cm.B;
            float gg = tvgd_voxel(c, x_ip, l_jp, l_kp, x_im, x_imjp, x_imkp, l_jm, l_jmkp, l_km, x_ipkm, x_jpkm);
            if (!valid) gg = 0.0f;
            g[((int64_t)xb * cpb + (int64_t)j * N + k) * 32 + lane] = gg;
            tot += (double)(gg * gg);
            cm = cc; cc = cp;
        }
    }
    tot = warp_sum(tot);
    __shared__ double sh[8];
    if (lane == 0) sh[w] = tot;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int i = 0; i < 8; ++i) s += sh[i];
        atomicAdd(p.sumsq, s);
    }
}

// x -= dPOCS * g / ||g||   (tv_gd.cu:108-122); ||g|| read from device memory (no host sync);
// the final positivity (tv_gd.cu:125-138, applied once after the loop) is fused into the last update.
__global__ void __launch_bounds__(256) tvgd_update_kernel(float* __restrict__ x, const float* __restrict__ g, const double* __restrict__ sumsq,
                                                           float dPOCS, int clamp, size_t n4)
{
    const float nrm = sqrtf((float)(*sumsq));
    float4* x4 = reinterpret_cast<float4*>(x);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 a = x4[i]; const float4 b = g4[i];
        a.x = a.x - dPOCS * b.x / nrm; a.y = a.y - dPOCS * b.y / nrm; a.z = a.z - dPOCS * b.z / nrm; a.w = a.w - dPOCS * b.w / nrm;
        if (clamp) { a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); a.z = fmaxf(a.z, 0.f); a.w = fmaxf(a.w, 0.f); }
        x4[i] = a;
    }
}

}  // namespace ttv
