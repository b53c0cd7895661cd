// Parallel-beam back projector (voxel-driven, linear interpolation along the detector),
// x-batched, TMA-staged.  sm_100a.
//
// Arithmetic restated from the reference's ASTRA kernel devBP (A/cuda/2d/par_bp.cu:52-87):
//   fT = fX*sc - fY*ss + off ;  vol += sum_a tex(sino_a, fT) * scale_a
// with per-angle constants sc, ss, off, scale (par_bp.cu:173-179) held in __constant__ memory
// and the sinogram read through a linear-filtering texture (8-bit weights, Border addressing).
//
// Design: CTA = 32 slices (one x-block) x a 16x16 voxel tile.  For one angle the tile projects
// onto a window of <= 26 detector cells; a producer warp streams the windows of 16 angles per
// stage into shared memory with one 4-D TMA box per angle (out-of-bounds zero fill = Border).
// 8 consumer warps; a lane owns 4 slices (float4) of 8 consecutive voxels: 32 accumulators.
// Epilogues fuse the reference's separate update kernels: SIRT x = max(x + C*t, 0)
// (A/cuda/2d/sirt.cu:259-267) and the KL/EM step (T/gpu/utils/tomoengine.cpp:302-313).
#pragma once
#include "ttv_common.cuh"

namespace ttv {

constexpr int BP_TILE = 16;
constexpr int BP_VT = 8;       // voxels per lane
constexpr int BP_NW = 8;       // consumer warps
constexpr int BP_TA = 16;      // angles per stage
constexpr int BP_CW = 27;      // detector cells per angle window (26 + 1: the window starts at an even cell)
constexpr int BP_S = 2;        // stages
constexpr int BP_THREADS = (BP_NW + 1) * 32;
constexpr int BP_MAX_ANGLES = 2048;
constexpr uint32_t BP_ANGLE_BYTES = BP_CW * 128u;
constexpr uint32_t BP_STAGE_BYTES = BP_TA * BP_ANGLE_BYTES;
constexpr size_t BP_SMEM = (size_t)BP_S * BP_STAGE_BYTES + BP_S * BP_TA * sizeof(float) + 2 * BP_S * sizeof(uint64_t);

__constant__ float4 c_bp_angle[BP_MAX_ANGLES];   // {sc, ss, off, scale}

enum BpEpilogue { BP_STORE = 0, BP_SIRT_UPD = 1, BP_STEP_UPD = 2, BP_ADD = 3, BP_FUSE_UPD = 4 };

struct BpParams {
    int N, Nd, Na;
    float* vol;            // output / updated volume (x-interleaved)
    const float* C;        // [N][N] pixel weights (SIRT)
    float step;            // BP_STEP_UPD: x = max(x - step*t, 0);  BP_ADD: x += step*t
    // BP_FUSE_UPD (multi-modal fusion, T/chemistry/utils/multimodal.cpp:425-441,475-481):
    //   x = max(x - (step*t - lamH * (gamma*x^(gamma-1)) * (W*(U-M))), 0)
    const float* M = nullptr;   // modelled HAADF volume  sum_e W_e x_e^gamma
    const float* U = nullptr;   // SIRT-updated HAADF volume
    const float* W = nullptr;   // [N][N] summation weight of this element
    float gamma = 1.f, lamH = 0.f;
};

// the update that follows the back projection of one voxel (4 slices): same expressions as bp_kernel's epilogue
template <int EPI>
__device__ __forceinline__ float4 bp_apply(const BpParams& p, size_t o, float4 r, int Y, int X)
{
    const int N = p.N;
    if (EPI == BP_STORE) return r;
    const float4 x = *reinterpret_cast<const float4*>(p.vol + o);
    if (EPI == BP_SIRT_UPD) {
        const float c = p.C[(size_t)Y * N + X];
        return make_float4(fmaxf(fmaf(c, r.x, x.x), 0.f), fmaxf(fmaf(c, r.y, x.y), 0.f),
                           fmaxf(fmaf(c, r.z, x.z), 0.f), fmaxf(fmaf(c, r.w, x.w), 0.f));
    } else if (EPI == BP_STEP_UPD) {
        const float c = p.step;
        return make_float4(fmaxf(__fsub_rn(x.x, __fmul_rn(c, r.x)), 0.f), fmaxf(__fsub_rn(x.y, __fmul_rn(c, r.y)), 0.f),
                           fmaxf(__fsub_rn(x.z, __fmul_rn(c, r.z)), 0.f), fmaxf(__fsub_rn(x.w, __fmul_rn(c, r.w)), 0.f));
    } else if (EPI == BP_FUSE_UPD) {
        const float4 m = *reinterpret_cast<const float4*>(p.M + o);
        const float4 u = *reinterpret_cast<const float4*>(p.U + o);
        const float w = p.W[(size_t)Y * N + X];
        const float c = p.step, g = p.gamma, lh = p.lamH, gm1 = p.gamma - 1.0f;
        auto upd = [&](float xv, float tv, float mv, float uv) -> float {
            float h = __fmul_rn(w, __fsub_rn(uv, mv));
            if (g != 1.0f) h = __fmul_rn(__fmul_rn(g, powf(xv, gm1)), h);
            return fmaxf(__fsub_rn(xv, __fsub_rn(__fmul_rn(c, tv), __fmul_rn(lh, h))), 0.f);
        };
        return make_float4(upd(x.x, r.x, m.x, u.x), upd(x.y, r.y, m.y, u.y), upd(x.z, r.z, m.z, u.z), upd(x.w, r.w, m.w, u.w));
    } else {  // BP_ADD
        const float c = p.step;
        return make_float4(fmaf(c, r.x, x.x), fmaf(c, r.y, x.y), fmaf(c, r.z, x.z), fmaf(c, r.w, x.w));
    }
}

template <int EPI>
__global__ void __launch_bounds__(BP_THREADS, 2)
bp_kernel(const __grid_constant__ CUtensorMap tmS, const BpParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* s_C = reinterpret_cast<float*>(smem_raw + (size_t)BP_S * BP_STAGE_BYTES);   // [S][TA]
    uint64_t* full = reinterpret_cast<uint64_t*>(s_C + BP_S * BP_TA);
    uint64_t* empty = full + BP_S;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.N, Na = p.Na;
    const int x0 = blockIdx.x * BP_TILE, y0 = blockIdx.y * BP_TILE, xb = blockIdx.z;
    const int nst = (Na + BP_TA - 1) / BP_TA;

    if (tid == 0) {
        for (int s = 0; s < BP_S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], BP_NW); }
        fence_barrier_init();
    }
    __syncthreads();

    const float fX0 = (float)x0 - 0.5f * (float)N + 0.5f;
    const float fY0 = (float)y0 - 0.5f * (float)N + 0.5f;

    if (warp == BP_NW) {
        // ---------------- producer: lane l streams the window of angle a0+l ----------------
        if (lane == 0) tma_prefetch_desc(&tmS);
        for (int it = 0; it < nst; ++it) {
            const int s = it % BP_S;
            if (it >= BP_S) mbar_wait(&empty[s], ((it / BP_S) - 1) & 1);
            const int a = it * BP_TA + lane;
            if (lane < BP_TA && a < Na) {
                const float4 k = c_bp_angle[a];
                const float ax0 = fX0 * k.x, ax1 = (fX0 + (float)(BP_TILE - 1)) * k.x;
                const float ay0 = -fY0 * k.y, ay1 = -(fY0 + (float)(BP_TILE - 1)) * k.y;
                const float fmin = fminf(ax0, ax1) + fminf(ay0, ay1) + k.z;
                // even start: the parity of a window-relative cell index is the parity of the detector cell, so
                // "even cell first" means the same in bp_kernel and bpt_kernel whatever their tiles
                const int w0 = ((int)floorf(fmin - 0.5f) - 1) & ~1;
                s_C[s * BP_TA + lane] = TTV_MAGIC_C(w0);
                tma_load_4d(smem_raw + (size_t)s * BP_STAGE_BYTES + (size_t)lane * BP_ANGLE_BYTES, &tmS, &full[s], 0, w0, a, xb);
            }
            __syncwarp();
            if (lane == 0) {
                const int na = min(BP_TA, Na - it * BP_TA);
                mbar_arrive_expect_tx(&full[s], (uint32_t)na * BP_ANGLE_BYTES);
            }
        }
        return;
    }

    // ---------------- consumers ----------------
    const int lx = lane & 7, lr = lane >> 3;
    const int slot = warp * 4 + lr;                   // 0..31
    const int Y = y0 + (slot >> 1);
    const int Xb = x0 + (slot & 1) * BP_VT;
    const float fY = (float)Y - 0.5f * (float)N + 0.5f;
    float fX[BP_VT];
#pragma unroll
    for (int t = 0; t < BP_VT; ++t) fX[t] = (float)(Xb + t) - 0.5f * (float)N + 0.5f;

    Acc4 acc[BP_VT];
#pragma unroll
    for (int t = 0; t < BP_VT; ++t) acc[t].zero();
    const uint32_t smem_base = smem_u32(smem_raw) + (uint32_t)lx * 16u;

    for (int it = 0; it < nst; ++it) {
        const int s = it % BP_S;
        mbar_wait(&full[s], (it / BP_S) & 1);
        const int na = min(BP_TA, Na - it * BP_TA);
        for (int l = 0; l < na; ++l) {
            const float4 k = c_bp_angle[it * BP_TA + l];
            const float Cl = s_C[s * BP_TA + l];
            const uint32_t base = smem_base + (uint32_t)s * BP_STAGE_BYTES + (uint32_t)l * BP_ANGLE_BYTES;
            const float myss = -__fmul_rn(fY, k.y);
            const float ws = k.w * (1.0f / 512.0f);
            const float w256 = 512.0f * ws;
#pragma unroll
            for (int t = 0; t < BP_VT; ++t) {
                const float fT = __fadd_rn(__fmaf_rn(fX[t], k.x, myss), k.z);
                const float tt = __fmaf_rd(fT, 256.0f, Cl);
                const uint32_t u = (uint32_t)(__float_as_int(tt) - TTV_MAGIC_BITS);
                const uint32_t off = (u >> 9) << 7;
                const float al = __int_as_float((int)((u & 0x1FEu) | 0x4B000000u)) - 8388608.0f;
                const float w1 = al * ws;
                const float w0 = fmaf(-al, ws, w256);
                // even cell first, whichever of the pair it is: the accumulation order of the table-driven kernel
                // (ttv_bpt.cuh), so that thin and thick slabs -- 1 or 8 GPUs -- give bit-identical volumes
                const bool odd = (off & 128u) != 0u;
                const float4 vE = lds128(base + ((off + 128u) & ~128u));
                const float4 vO = lds128(base + (off | 128u));
                acc[t].fma(odd ? w1 : w0, vE);      // acc = fmaf(wO, vO, fmaf(wE, vE, acc)), 4 FFMA2
                acc[t].fma(odd ? w0 : w1, vO);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }

    // ---------------- epilogue ----------------
    if (Y < N) {
#pragma unroll
        for (int t = 0; t < BP_VT; ++t) {
            const int X = Xb + t;
            if (X < N) {
                const size_t o = (((size_t)xb * N + Y) * N + X) * TTV_XI + lx * 4;
                const float4 r = bp_apply<EPI>(p, o, acc[t].get(), Y, X);
                *reinterpret_cast<float4*>(p.vol + o) = r;
            }
        }
    }
}

}  // namespace ttv
