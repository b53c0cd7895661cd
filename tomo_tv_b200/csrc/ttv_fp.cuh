// Parallel-beam forward projector (Joseph, ray-driven), x-batched, TMA-staged.  sm_100a.
//
// Arithmetic restated from the reference's ASTRA kernels FPvertical_simple / FPhorizontal_simple
// (A/cuda/2d/par_fp.cu:53-200): per ray the sample position fP is restarted from the closed form
// every 64 lines (:49,282-286) and advanced by fSliceStep in float32 in between; the sample is a
// texture fetch with linear filtering (8-bit weights, Border addressing, A/cuda/2d/util.cu:161-179).
//
// Design (nothing like the reference's one-launch-per-64-rows-per-slice texture kernels):
//   * CTA = 32 slices (one x-block) x 8 same-orientation angles x 64 detectors = 512 rays.
//   * For each line (row for "vertical" angles, column for "horizontal" ones) the cells the 512
//     rays touch form one contiguous window of CW cells; a producer warp streams these windows
//     into a multi-stage shared-memory ring with one 4-D TMA box per line.  TMA's out-of-bounds
//     zero fill implements the Border addressing mode.  Both orientations use the same consumer
//     code; only the TMA box orientation differs (two tensor maps).
//   * 16 consumer warps; a lane owns 4 slices (float4) of 8 consecutive detectors of one angle:
//     32 accumulators.  One index/weight computation serves 4 samples; every LDS.128 is a
//     conflict-free 128-byte line per quarter warp.
//   * Epilogues fuse what the reference does in separate kernels: SIRT residual (b-Ax)*R
//     (A/cuda/2d/sirt.cu:246-257), the KL ratio and cost (T/gpu/utils/tomoengine.cpp:301-311),
//     the data-distance reduction (tomoengine.cpp:410-413).
#pragma once
#include "ttv_common.cuh"

namespace ttv {

constexpr int FP_AG = 8;      // angles per CTA
constexpr int FP_RT = 8;      // consecutive detectors per lane
constexpr int FP_NW = 16;     // consumer warps
constexpr int FP_DG = (FP_NW * 4 / FP_AG) * FP_RT;   // 64 detectors per CTA
constexpr int FP_THREADS = (FP_NW + 1) * 32;
constexpr int FP_MAX_STAGES = 8;

enum FpEpilogue { FP_STORE = 0, FP_SIRT_RES = 1, FP_KL = 2, FP_STORE_DIST = 3, FP_RESID = 4 };

struct FpAngle { float theta, detsize, offset; int orig; };
struct FpGroup { int first, count, vertical, pad; };

struct FpParams {
    const FpAngle* ang;    // angles sorted by (orientation, theta)
    const FpGroup* grp;    // groups of <= FP_AG consecutive sorted angles
    const int* win0;       // [group][detgroup][line] first staged cell
    int N, Nd, Na, ndg;
    int CW, TR, S;         // window cells per line, lines per stage, stages
    float* out;            // sinogram, x-interleaved
    const float* b;        // measured sinogram (x-interleaved) for the fused epilogues
    const float* R;        // [Na][Nd] line weights (SIRT)
    double* red;           // reduction target (KL cost / squared distance)
    float eps;
};

// what follows the projection of one ray (4 slices, already scaled): store / SIRT residual (b-Ax)*R
// (A/cuda/2d/sirt.cu:246-257) / KL ratio + cost (T/gpu/utils/tomoengine.cpp:301-311) / store + squared distance
// (tomoengine.cpp:410-413) / plain residual
template <int EPI>
__device__ __forceinline__ void fp_emit(const FpParams& p, int orig, int d, int xb, int lx, float4 ax, float& part)
{
    const size_t cell = ((size_t)xb * p.Na + orig) * p.Nd + d;
    const size_t o = cell * TTV_XI + lx * 4;
    float4 r;
    if (EPI == FP_STORE) {
        r = ax;
    } else {
        const float4 bb = *reinterpret_cast<const float4*>(p.b + o);
        if (EPI == FP_SIRT_RES) {
            const float w = p.R[(size_t)orig * p.Nd + d];
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
__global__ void __launch_bounds__(FP_THREADS, 1)
fp_kernel(const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmH, const FpParams p)
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
        for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], FP_NW); }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == FP_NW) {
        // ---------------- producer warp: lane l streams line l of each stage ----------------
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
                if (g.vertical) tma_load_4d(dst, tm, &full[s], 0, w0, line, xb);   // window along columns of row `line`
                else            tma_load_4d(dst, tm, &full[s], 0, line, w0, xb);   // window along rows of column `line`
            }
            __syncwarp();
            if (lane == 0) mbar_arrive_expect_tx(&full[s], stage_bytes);
        }
        return;
    }

    // ---------------- consumer warps ----------------
    const int lx = lane & 7, lr = lane >> 3;
    const int slot = warp * 4 + lr;
    constexpr int SLOTS_PER_ANGLE = FP_NW * 4 / FP_AG;
    const int ai = slot / SLOTS_PER_ANGLE;                 // warp-uniform
    const int ds = slot % SLOTS_PER_ANGLE;
    const bool angle_ok = ai < g.count;
    const FpAngle A = p.ang[g.first + (angle_ok ? ai : g.count - 1)];
    const int dbase = dg * FP_DG + ds * FP_RT;

    // per-angle constants, same operations as par_fp.cu:63-85 / :140-161
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
    // m[t] = (d - 0.5*Nd + 0.5 - offset) * detStep    (rounded product, see SASS note in oracle.c)
    float m[FP_RT], fP[FP_RT];
#pragma unroll
    for (int t = 0; t < FP_RT; ++t) {
        int d = dbase + t; if (d > p.Nd - 1) d = p.Nd - 1;
        float t1 = __fadd_rn(__fadd_rn(__fmaf_rn((float)p.Nd, -0.5f, (float)d), 0.5f), -A.offset);
        m[t] = __fmul_rn(t1, detStep);
        fP[t] = 0.0f;
    }
    Acc4 acc[FP_RT];
#pragma unroll
    for (int t = 0; t < FP_RT; ++t) acc[t].zero();

    const uint32_t smem_base = smem_u32(smem_raw) + (uint32_t)lx * 16u;
    const float halfN = 0.5f * (float)N;

    for (int it = 0; it < nstages_total; ++it) {
        const int s = it % S;
        mbar_wait(&full[s], (it / S) & 1);
        if (angle_ok)                 // warp-uniform: warps of a partially filled group only keep the ring moving
        for (int l = 0; l < TR; ++l) {
            const int line = it * TR + l;
            const float Cl = s_C[s * TR + l];
            const uint32_t base = smem_base + (uint32_t)s * stage_bytes + (uint32_t)l * line_bytes;
            // g_blockSlices = 64 (par_fp.cu:49): the position restarts from the closed form every 64 lines and
            // advances by one rounded add in between.  The restart is warp-uniform and rare, so it is a real
            // branch outside the sampling loop (if-converted into the loop it cost 32 predicated-off instructions
            // per 8 samples on every line: ncu showed 26.8 of 32 threads not predicated off).
            if ((line & 63) == 0) {
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
#pragma unroll
            for (int t = 0; t < FP_RT; ++t) {
                const float tt = __fmaf_rd(fP[t], 256.0f, Cl);
                const uint32_t u = (uint32_t)(__float_as_int(tt) - TTV_MAGIC_BITS);
                const uint32_t off = (u >> 9) << 7;
                const float w1 = __int_as_float((int)((u & 0x1FEu) | 0x4B000000u)) - 8388608.0f;   // 512*alpha
                const float w0 = 512.0f - w1;
                // even cell first, whichever of the pair it is: the accumulation order of the table-driven kernel
                // (ttv_fpt.cuh), so that thin and thick slabs -- 1 or 8 GPUs -- give bit-identical sinograms
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
    (void)halfN;

    // ---------------- epilogue ----------------
    float part = 0.0f;
    if (angle_ok) {
        const float cs = corr1 * (1.0f / 512.0f);
#pragma unroll
        for (int t = 0; t < FP_RT; ++t) {
            const int d = dbase + t;
            if (d < p.Nd) {
                const float4 a4 = acc[t].get();
                fp_emit<EPI>(p, A.orig, d, xb, lx, make_float4(a4.x * cs, a4.y * cs, a4.z * cs, a4.w * cs), part);
            }
        }
    }
    if (EPI == FP_KL || EPI == FP_STORE_DIST) {
        double s = warp_sum((double)part);
        if (lane == 0) atomicAdd(p.red, s);
    }
}

}  // namespace ttv
