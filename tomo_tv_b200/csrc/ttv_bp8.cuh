// EXPERIMENTAL back projector variant: 8 slices per lane + register reuse along the
// voxel run.  NOT the default: selected with ttv_set_option("bp_variant", 8); written at the end of round 1
// without GPU access, so it has compiled but never run -- the parity tests take it through
// tests/test_gpu_experimental.py (one variant per process), which must pass before it becomes the default.
//
// Same arithmetic as bp_kernel (ttv_bp.cuh; reference devBP, A/cuda/2d/par_bp.cu:52-87): identical position,
// 8-bit weight emulation and FMA order per sample, so results are bit-identical to bp_kernel.  What changes:
//   * CTA = one x-block x a 32 (X) x 24 (Y) voxel tile, 12 consumer warps; window of <= 44 detector cells per angle,
//     8 angles per stage, 3 stages.
//   * a lane owns 8 slices (two float4) of 8 consecutive voxels along X: the index / weight arithmetic of a
//     sample is shared by 8 slices instead of 4, and 4 lanes (not 8) share one cell.
//   * consecutive voxels move the detector coordinate by |sc| <= 1 cell, so the cell pair (c, c+1) of voxel t+1
//     is the pair of voxel t or its neighbour: the pair is kept in registers and only the missing cell is
//     loaded -- on average 2 + 7|cos| cells per 8 voxels instead of 16.
//   * bank layout: a quarter warp (8 lanes) covers two cells; the lanes of odd cells read their two float4 in
//     swapped order, so the two cells of a quarter hit disjoint banks (conflict-free LDS.128).
#pragma once
#include "ttv_bp.cuh"

namespace ttv {

constexpr int BP8_TX = 32, BP8_TY = 24;
constexpr int BP8_VT = 8;       // voxels per lane
constexpr int BP8_NW = 12;      // consumer warps.  64 accumulators + 16 cell registers per lane need ~120 registers: each of the
                                // SM's 4 schedulers owns 16 K registers, so 13 warps (4+3+3+3) may use 128 each, 17 warps (5+4+4+4) only 96
constexpr int BP8_TA = 8;       // angles per stage
constexpr int BP8_CW = 44;      // detector cells per angle window: 31|sc| + 23|ss| <= 38.6, + 4 cells of margin (as BP_CW) -> 44
constexpr int BP8_S = 3;        // stages (3 x 44 KB)
constexpr int BP8_THREADS = (BP8_NW + 1) * 32;
constexpr uint32_t BP8_ANGLE_BYTES = BP8_CW * 128u;
constexpr uint32_t BP8_STAGE_BYTES = BP8_TA * BP8_ANGLE_BYTES;
constexpr size_t BP8_SMEM = (size_t)BP8_S * BP8_STAGE_BYTES + BP8_S * BP8_TA * sizeof(float) + 2 * BP8_S * sizeof(uint64_t);

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

// FORM 0: the cell pair (c, c+1) lives in registers a*, b* and is shifted by register moves when the voxel run
//         crosses a cell; operations and their order are those of bp_kernel (bit-identical results).
// FORM 1: "parity registers" -- the even cell lives in e*, the odd cell in o*, a cell that is already held is not
//         reloaded and nothing is ever moved; the weights are swapped by the parity of c instead.  The two FMAs of
//         a sample are then issued even-cell-first, i.e. in the other order whenever c is odd: same values to
//         within one rounding of the partial sum (checked against FORM 0 with a tolerance, not bit-exactly).
template <int EPI, int FORM>
__global__ void __launch_bounds__(BP8_THREADS, 1)
bp8_kernel(const __grid_constant__ CUtensorMap tmS, const BpParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* s_C = reinterpret_cast<float*>(smem_raw + (size_t)BP8_S * BP8_STAGE_BYTES);   // [S][TA]
    uint64_t* full = reinterpret_cast<uint64_t*>(s_C + BP8_S * BP8_TA);
    uint64_t* empty = full + BP8_S;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.N, Na = p.Na;
    const int x0 = blockIdx.x * BP8_TX, y0 = blockIdx.y * BP8_TY, xb = blockIdx.z;
    const int nst = (Na + BP8_TA - 1) / BP8_TA;

    if (tid == 0) {
        for (int s = 0; s < BP8_S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], BP8_NW); }
        fence_barrier_init();
    }
    __syncthreads();

    const float fX0 = (float)x0 - 0.5f * (float)N + 0.5f;
    const float fY0 = (float)y0 - 0.5f * (float)N + 0.5f;

    if (warp == BP8_NW) {
        // ---------------- producer: lane l streams the window of angle a0+l ----------------
        if (lane == 0) tma_prefetch_desc(&tmS);
        for (int it = 0; it < nst; ++it) {
            const int s = it % BP8_S;
            if (it >= BP8_S) mbar_wait(&empty[s], ((it / BP8_S) - 1) & 1);
            const int a = it * BP8_TA + lane;
            if (lane < BP8_TA && a < Na) {
                const float4 k = c_bp_angle[a];
                const float ax0 = fX0 * k.x, ax1 = (fX0 + (float)(BP8_TX - 1)) * k.x;
                const float ay0 = -fY0 * k.y, ay1 = -(fY0 + (float)(BP8_TY - 1)) * k.y;
                const float fmin = fminf(ax0, ax1) + fminf(ay0, ay1) + k.z;
                const int w0 = (int)floorf(fmin - 0.5f) - 1;
                s_C[s * BP8_TA + lane] = TTV_MAGIC_C(w0);
                tma_load_4d(smem_raw + (size_t)s * BP8_STAGE_BYTES + (size_t)lane * BP8_ANGLE_BYTES, &tmS, &full[s], 0, w0, a, xb);
            }
            __syncwarp();
            if (lane == 0) {
                const int na = min(BP8_TA, Na - it * BP8_TA);
                mbar_arrive_expect_tx(&full[s], (uint32_t)na * BP8_ANGLE_BYTES);
            }
        }
        return;
    }

    // ---------------- consumers ----------------
    const int g = lane >> 2, lx = lane & 3;
    const int slot = warp * 8 + g;                    // 0..95
    const int Y = y0 + (slot >> 2);
    const int Xb = x0 + (slot & 3) * BP8_VT;
    const uint32_t par = (uint32_t)(g & 1);
    const uint32_t o1 = (uint32_t)lx * 32u + par * 16u;             // "first" float4 of this lane inside a cell
    const uint32_t o2 = (uint32_t)lx * 32u + (par ^ 1u) * 16u;      // "second"
    const float fY = (float)Y - 0.5f * (float)N + 0.5f;
    // fX(t) = (Xb + t) - N/2 + 1/2 is a small multiple of 1/2: fXb + t is exact, one register instead of eight
    const float fXb = (float)Xb - 0.5f * (float)N + 0.5f;

    Acc4 acc1[BP8_VT], acc2[BP8_VT];
#pragma unroll
    for (int t = 0; t < BP8_VT; ++t) { acc1[t].zero(); acc2[t].zero(); }
    const uint32_t smem_base = smem_u32(smem_raw);

    for (int it = 0; it < nst; ++it) {
        const int s = it % BP8_S;
        mbar_wait(&full[s], (it / BP8_S) & 1);
        const int na = min(BP8_TA, Na - it * BP8_TA);
        for (int l = 0; l < na; ++l) {
            const float4 k = c_bp_angle[it * BP8_TA + l];
            const float Cl = s_C[s * BP8_TA + l];
            const uint32_t base = smem_base + (uint32_t)s * BP8_STAGE_BYTES + (uint32_t)l * BP8_ANGLE_BYTES;
            const float myss = -__fmul_rn(fY, k.y);
            const float ws = k.w * (1.0f / 512.0f);
            const float w256 = 512.0f * ws;
            float4 a1, a2, b1, b2;                    // FORM 0: cells c (a*) and c+1 (b*); FORM 1: even (a*) and odd (b*) cell
            uint32_t poff = 0;                        // FORM 0: offset of the held cell c
            uint32_t hE = 0xFFFFFFFFu, hO = 0xFFFFFFFFu;   // FORM 1: offsets of the held even / odd cells (none yet)
#pragma unroll
            for (int t = 0; t < BP8_VT; ++t) {
                const float fT = __fadd_rn(__fmaf_rn(__fadd_rn(fXb, (float)t), k.x, myss), k.z);
                const float tt = __fmaf_rd(fT, 256.0f, Cl);
                const uint32_t u = (uint32_t)(__float_as_int(tt) - TTV_MAGIC_BITS);
                const uint32_t off = (u >> 9) << 7;
                const float al = __int_as_float((int)((u & 0x1FEu) | 0x4B000000u)) - 8388608.0f;
                const float w1 = al * ws;
                const float w0 = fmaf(-al, ws, w256);
                if (FORM == 0) {
                    const uint32_t ad = base + off;
                    if (t == 0) {
                        a1 = lds128(ad + o1); a2 = lds128(ad + o2);
                        b1 = lds128(ad + 128u + o1); b2 = lds128(ad + 128u + o2);
                    } else if (off != poff) {
                        if (off == poff + 128u) {                    // moved one cell up: old c+1 is the new c
                            a1 = b1; a2 = b2;
                            b1 = lds128(ad + 128u + o1); b2 = lds128(ad + 128u + o2);
                        } else if (off + 128u == poff) {             // moved one cell down: old c is the new c+1
                            b1 = a1; b2 = a2;
                            a1 = lds128(ad + o1); a2 = lds128(ad + o2);
                        } else {                                     // cannot happen for |sc| <= 1; kept for safety
                            a1 = lds128(ad + o1); a2 = lds128(ad + o2);
                            b1 = lds128(ad + 128u + o1); b2 = lds128(ad + 128u + o2);
                        }
                    }
                    poff = off;
                    acc1[t].fma(w0, a1); acc1[t].fma(w1, b1);        // acc = fmaf(w1, v1, fmaf(w0, v0, acc))
                    acc2[t].fma(w0, a2); acc2[t].fma(w1, b2);
                } else {
                    // cells c and c+1 (window-relative byte offsets off, off+128): one of them is even, one odd
                    const bool codd = (off & 128u) != 0u;
                    const uint32_t offE = (off + 128u) & ~128u;      // the even one of {off, off+128}
                    const uint32_t offO = off | 128u;                // the odd one
                    if (offE != hE) { a1 = lds128(base + offE + o1); a2 = lds128(base + offE + o2); hE = offE; }
                    if (offO != hO) { b1 = lds128(base + offO + o1); b2 = lds128(base + offO + o2); hO = offO; }
                    const float wE = codd ? w1 : w0, wO = codd ? w0 : w1;
                    acc1[t].fma(wE, a1); acc1[t].fma(wO, b1);
                    acc2[t].fma(wE, a2); acc2[t].fma(wO, b2);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }

    // ---------------- epilogue ----------------
    if (Y < N) {
#pragma unroll
        for (int t = 0; t < BP8_VT; ++t) {
            const int X = Xb + t;
            if (X < N) {
                const size_t cell = (((size_t)xb * N + Y) * N + X) * TTV_XI;
                const size_t oa = cell + (o1 >> 2), ob = cell + (o2 >> 2);       // byte offsets -> float offsets
                const float4 ra = bp_apply<EPI>(p, oa, acc1[t].get(), Y, X);
                const float4 rb = bp_apply<EPI>(p, ob, acc2[t].get(), Y, X);
                *reinterpret_cast<float4*>(p.vol + oa) = ra;
                *reinterpret_cast<float4*>(p.vol + ob) = rb;
            }
        }
    }
}

}  // namespace ttv
