// EXPERIMENTAL forward projector variant: 8 slices per lane + parity-register reuse.
// NOT the default: selected with ttv_set_option("fp_variant", 8); written at the end of round 1 without GPU
// access, so it has compiled but never run (gate: tests/test_gpu_experimental.py, own process).
//
// Same plan, same CTA ray set and same per-sample arithmetic as fp_kernel (ttv_fp.cuh): 32 slices x 8 angles x
// 64 detectors per CTA, position recurrence restarted every 64 lines, 8-bit weight emulation.  What changes:
//   * 8 consumer warps instead of 16; a lane owns 8 slices (two float4) of 8 consecutive detectors, so the
//     index / weight arithmetic of a sample is shared by 8 slices and 4 lanes (not 8) share one cell;
//   * 9 warps per CTA leave 170 registers per thread (a scheduler owns 16 K registers), room for 64 accumulators;
//   * parity registers: the even cell of a sample lives in e1/e2, the odd one in o1/o2; a held cell is not loaded
//     again (neighbouring detectors are 1 .. 1.41 cells apart), the weights are swapped by the parity of the lower
//     cell.  The two taps are added even-cell-first: equal to fp_kernel within one rounding per sample;
//   * bank layout: a quarter warp covers two detector slots; lanes of odd slots read their two float4 in swapped
//     order so the two cells of a quarter hit disjoint banks.
#pragma once
#include "ttv_fp.cuh"

namespace ttv {

constexpr int FP8_NW = 8;                              // consumer warps: 8 warps x 8 slots = 64 slots = FP_AG x (FP_DG / FP_RT)
constexpr int FP8_THREADS = (FP8_NW + 1) * 32;
static_assert(FP8_NW * 8 == FP_AG * (FP_DG / FP_RT), "fp8_kernel covers the same ray set as fp_kernel");

// epilogue of one float4 (4 slices of one ray): same expressions as fp_kernel's
template <int EPI>
__device__ __forceinline__ void fp8_emit(const FpParams& p, size_t o, float4 a4, float cs, float w, float& part)
{
    const float4 ax = make_float4(a4.x * cs, a4.y * cs, a4.z * cs, a4.w * cs);
    float4 r;
    if (EPI == FP_STORE) {
        r = ax;
    } else {
        const float4 bb = *reinterpret_cast<const float4*>(p.b + o);
        if (EPI == FP_SIRT_RES) {
            r = make_float4((bb.x - ax.x) * w, (bb.y - ax.y) * w, (bb.z - ax.z) * w, (bb.w - ax.w) * w);
        } else if (EPI == FP_RESID) {
            r = make_float4(bb.x - ax.x, bb.y - ax.y, bb.z - ax.z, bb.w - ax.w);
        } else if (EPI == FP_KL) {
            const float e = p.eps;
            r = make_float4((ax.x - bb.x) / (ax.x + e), (ax.y - bb.y) / (ax.y + e),
                            (ax.z - bb.z) / (ax.z + e), (ax.w - bb.w) / (ax.w + e));
            part += (ax.x - bb.x * logf(ax.x + e)) + (ax.y - bb.y * logf(ax.y + e))
                  + (ax.z - bb.z * logf(ax.z + e)) + (ax.w - bb.w * logf(ax.w + e));
        } else {  // FP_STORE_DIST
            r = ax;
            const float dx = ax.x - bb.x, dy = ax.y - bb.y, dz = ax.z - bb.z, dw = ax.w - bb.w;
            part += dx * dx + dy * dy + dz * dz + dw * dw;
        }
    }
    *reinterpret_cast<float4*>(p.out + o) = r;
}

template <int EPI>
__global__ void __launch_bounds__(FP8_THREADS, 1)
fp8_kernel(const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmH, const FpParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int S = p.S, TR = p.TR, CW = p.CW, N = p.N;
    const uint32_t line_bytes = (uint32_t)CW * 128u;
    const uint32_t stage_bytes = line_bytes * TR;
    float* s_C = reinterpret_cast<float*>(smem_raw + (size_t)S * stage_bytes);          // [S][TR]
    uint64_t* full = reinterpret_cast<uint64_t*>(s_C + S * TR + ((S * TR) & 1));          // [S]
    uint64_t* empty = full + S;                                                           // [S]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const FpGroup g = p.grp[blockIdx.y];
    const int dg = blockIdx.x, xb = blockIdx.z;
    const int nstages_total = (N + TR - 1) / TR;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], FP8_NW); }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == FP8_NW) {
        // ---------------- producer warp: lane l streams line l of each stage (as in fp_kernel) ----------------
        const CUtensorMap* tm = g.vertical ? &tmV : &tmH;
        if (lane == 0) tma_prefetch_desc(tm);
        const int* w0tab = p.win0 + ((size_t)blockIdx.y * p.ndg + dg) * (size_t)(nstages_total * TR);
        for (int it = 0; it < nstages_total; ++it) {
            const int s = it % S;
            if (it >= S) mbar_wait(&empty[s], ((it / S) - 1) & 1);
            if (lane < TR) {
                const int line = it * TR + lane;
                const int w0 = w0tab[line];
                s_C[s * TR + lane] = TTV_MAGIC_C(w0);
                void* dst = smem_raw + (size_t)s * stage_bytes + (size_t)lane * line_bytes;
                if (g.vertical) tma_load_4d(dst, tm, &full[s], 0, w0, line, xb);
                else            tma_load_4d(dst, tm, &full[s], 0, line, w0, xb);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive_expect_tx(&full[s], stage_bytes);
        }
        return;
    }

    // ---------------- consumer warps ----------------
    const int gq = lane >> 2, lx = lane & 3;               // slot within the warp, slice group (8 slices) within the cell
    const int slot = warp * 8 + gq;                        // 0..63, same slot space as fp_kernel
    constexpr int SLOTS_PER_ANGLE = FP_DG / FP_RT;         // 8
    const int ai = slot / SLOTS_PER_ANGLE;                 // warp-uniform (a warp's 8 slots belong to one angle)
    const int ds = slot % SLOTS_PER_ANGLE;
    const bool angle_ok = ai < g.count;
    const FpAngle A = p.ang[g.first + (angle_ok ? ai : g.count - 1)];
    const int dbase = dg * FP_DG + ds * FP_RT;
    const uint32_t par = (uint32_t)(gq & 1);
    const uint32_t q1 = (uint32_t)lx * 32u + par * 16u;            // "first" float4 of this lane inside a cell (bytes)
    const uint32_t q2 = (uint32_t)lx * 32u + (par ^ 1u) * 16u;     // "second"

    const float cth = __cosf(A.theta), sth = __sinf(A.theta);
    float detStep, step, corr1;
    if (g.vertical) {
        detStep = A.detsize / cth;
        step = sth / cth;
        corr1 = (cth < 0.0f) ? -1.0f / cth : 1.0f / cth;
    } else {
        detStep = -A.detsize / sth;
        step = cth / sth;
        corr1 = (sth > 0.0f) ? 1.0f / sth : -1.0f / sth;
    }
    float m[FP_RT], fP[FP_RT];
#pragma unroll
    for (int t = 0; t < FP_RT; ++t) {
        int d = dbase + t; if (d > p.Nd - 1) d = p.Nd - 1;
        float t1 = __fadd_rn(__fadd_rn(__fmaf_rn((float)p.Nd, -0.5f, (float)d), 0.5f), -A.offset);
        m[t] = __fmul_rn(t1, detStep);
        fP[t] = 0.0f;
    }
    Acc4 acc1[FP_RT], acc2[FP_RT];
#pragma unroll
    for (int t = 0; t < FP_RT; ++t) { acc1[t].zero(); acc2[t].zero(); }

    const uint32_t smem_base = smem_u32(smem_raw);

    for (int it = 0; it < nstages_total; ++it) {
        const int s = it % S;
        mbar_wait(&full[s], (it / S) & 1);
        if (angle_ok)
        for (int l = 0; l < TR; ++l) {
            const int line = it * TR + l;
            const float Cl = s_C[s * TR + l];
            const uint32_t base = smem_base + (uint32_t)s * stage_bytes + (uint32_t)l * line_bytes;
            if ((line & 63) == 0) {                                    // g_blockSlices = 64 (par_fp.cu:49)
                const float t2 = __fadd_rn(__fmaf_rn((float)N, -0.5f, (float)line), 0.5f);
#pragma unroll
                for (int t = 0; t < FP_RT; ++t) {
                    float f = __fmaf_rn(t2, step, m[t]);
                    f = __fmaf_rn((float)N, 0.5f, f);
                    f = __fadd_rn(f, -0.5f);
                    fP[t] = __fadd_rn(f, 0.5f);
                }
            } else {
#pragma unroll
                for (int t = 0; t < FP_RT; ++t) fP[t] = __fadd_rn(fP[t], step);
            }
            float4 e1, e2, o1, o2;                                     // held even / odd cell, first / second float4
            uint32_t hE = 0xFFFFFFFFu, hO = 0xFFFFFFFFu;               // their window-relative offsets (none yet)
#pragma unroll
            for (int t = 0; t < FP_RT; ++t) {
                const float tt = __fmaf_rd(fP[t], 256.0f, Cl);
                const uint32_t u = (uint32_t)(__float_as_int(tt) - TTV_MAGIC_BITS);
                const uint32_t off = (u >> 9) << 7;
                const float w1 = __int_as_float((int)((u & 0x1FEu) | 0x4B000000u)) - 8388608.0f;   // 512*alpha
                const float w0 = 512.0f - w1;
                const bool codd = (off & 128u) != 0u;                  // cells off, off+128: one even, one odd
                const uint32_t offE = (off + 128u) & ~128u, offO = off | 128u;
                if (offE != hE) { e1 = lds128(base + offE + q1); e2 = lds128(base + offE + q2); hE = offE; }
                if (offO != hO) { o1 = lds128(base + offO + q1); o2 = lds128(base + offO + q2); hO = offO; }
                const float wE = codd ? w1 : w0, wO = codd ? w0 : w1;
                acc1[t].fma(wE, e1); acc1[t].fma(wO, o1);
                acc2[t].fma(wE, e2); acc2[t].fma(wO, o2);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }

    // ---------------- epilogue ----------------
    float part = 0.0f;
    if (angle_ok) {
        const float cs = corr1 * (1.0f / 512.0f);
#pragma unroll
        for (int t = 0; t < FP_RT; ++t) {
            const int d = dbase + t;
            if (d < p.Nd) {
                const size_t cell = (((size_t)xb * p.Na + A.orig) * p.Nd + d) * TTV_XI;
                const float w = (EPI == FP_SIRT_RES) ? p.R[(size_t)A.orig * p.Nd + d] : 0.0f;
                fp8_emit<EPI>(p, cell + (q1 >> 2), acc1[t].get(), cs, w, part);
                fp8_emit<EPI>(p, cell + (q2 >> 2), acc2[t].get(), cs, w, part);
            }
        }
    }
    if (EPI == FP_KL || EPI == FP_STORE_DIST) {
        double s = warp_sum((double)part);
        if (lane == 0) atomicAdd(p.red, s);
    }
}

}  // namespace ttv
