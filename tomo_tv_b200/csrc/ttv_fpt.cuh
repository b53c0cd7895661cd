// Table-driven parallel-beam forward projector (Joseph, ray-driven): 64 slices per CTA, cells reused in
// registers, index arithmetic in helper warps.  sm_100a.  Default forward projector for slabs of >= 2 x-blocks
// (ttv_engine.cuh: Engine::fp).
//
// Arithmetic: the reference's FPvertical_simple / FPhorizontal_simple (A/cuda/2d/par_fp.cu:53-200): sample
// positions by the same float32 operations as fp_kernel (ttv_fp.cuh; restart every 64 lines, par_fp.cu:49,282-286),
// 8-bit texture weights (A/cuda/2d/util.cu:161-179).  Both kernels accumulate the two taps of a sample
// even-cell-first over ascending lines; weights here are in 1/256 instead of 1/512 (an exact power-of-two scaling
// of every intermediate), so fpt_kernel and fp_kernel are bit-identical (tests/test_gpu_table_kernels.py).
//
// Why (profiles/r2b_*): fp_kernel reads 8 B of shared memory per sample (LDS pipe 90 % busy); neighbouring
// detectors of a ray run are 1 .. 1.41 cells apart, so the upper cell of detector t is usually the lower cell of
// detector t+1 -- but deciding that per lane costs more issue slots than the LDS it saves (fp_kernel<REUSE> and
// fp8_kernel of round 1: same run time).  Here
//   * CTA = 2 x-blocks (64 slices) x 8 same-orientation angles x 64 detectors, the plan of fp_kernel (FpParams).
//     Staged cells are x-block-interleaved ([cell][xb][32], one 4-D TMA box per line over a tensor map whose
//     second dimension is the x-block): a lane's 8 slices are two LDS.128 128 B apart.
//   * a quarter warp = one run of 8 consecutive detectors; a lane owns 8 slices x 8 detectors = 64 accumulators.
//   * four helper warps (128 threads = 64 runs x 2 halves) advance the sample positions -- once per CTA instead
//     of once per 8 lanes per x-block -- and leave one 4-byte entry per (ray, line) in shared memory:
//     [even-cell index | odd-cell index | weight of the even cell (0..256)], index 0xFF = "the run holds this cell".
//     A consumer fetches the 8 entries of its run with two LDS.128 and spends 9 integer/float instructions per
//     8 samples on decoding them (fp_kernel: ~18 per 4 samples).
//   * parity registers: a run loads ~1.2 cells per detector instead of 2.
//   * setmaxnreg: helpers 40 registers, consumers 104 (the CTA pool is what .dec released).
#pragma once
#include "ttv_fp.cuh"
#include "ttv_bpt.cuh"

namespace ttv {

constexpr int FPT_XB = 2;                       // x-blocks per CTA (template parameter XB: 2 by default, 1 on slabs of fewer than 4 x-blocks)
constexpr int FPT_NW = 16;                      // consumer warps (4 runs each)
constexpr int FPT_NH = 4;                       // helper warps
constexpr int FPT_RUNS = FPT_NW * 4;            // 64 runs = 8 angles x 8 detector runs
constexpr int FPT_THREADS = (FPT_NW + FPT_NH) * 32;
constexpr uint32_t FPT_TAB = FPT_RUNS * FP_RT * 4u;     // 2 KB of entries per line
constexpr int FPT_MAX_SLOTS = 10;
static_assert(FPT_RUNS == FP_AG * (FP_DG / FP_RT), "run layout must match the plan of fp_kernel");
static_assert(FPT_NH * 32 == FPT_RUNS * 2, "one helper thread per half run");

__host__ __device__ inline uint32_t fpt_slot_bytes(int CW, int xb = FPT_XB) { return (uint32_t)CW * ((uint32_t)xb * 128u) + FPT_TAB; }

// one predicated LDS.128 (a single x-block per CTA): loaded unless idx == 0xFF
template <uint32_t CELLB>
__device__ __forceinline__ void lds_cell1_if(float4& a, uint32_t idx, uint32_t lane_base) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .u32 ad;\n\t"
        "setp.ne.u32 p, %4, 255;\n\t"
        "mad.lo.u32 ad, %4, %6, %5;\n\t"
        "@p ld.shared.v4.f32 {%0,%1,%2,%3}, [ad];\n\t}"
        : "+f"(a.x), "+f"(a.y), "+f"(a.z), "+f"(a.w)
        : "r"(idx), "r"(lane_base), "n"(CELLB));
}

// XB = 2: a lane owns 8 slices (two x-blocks) x 8 detectors.  XB = 1: 4 slices x 8 detectors -- the same kernel for thin slabs
// (a sharded volume's 64 or 96 slices per GPU), where two x-blocks per CTA would leave 148 SMs with 96 or 192 CTAs.
// the work of one CTA: detector group `dg` of angle group `gy`, x-blocks xb0 .. xb0 + XB - 1
template <int EPI, int XB>
__device__ __forceinline__ void fpt_body(const CUtensorMap& tmV, const CUtensorMap& tmH, const FpParams& p, int nxb, int NS,
                                         const int dg, const int gy, const int xb0)
{
    constexpr uint32_t FPT_CELL = XB * 128u;                 // bytes of one staged cell
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int CW = p.CW, N = p.N;
    const uint32_t data_bytes = (uint32_t)CW * (XB * 128u);
    const uint32_t slot_bytes = data_bytes + FPT_TAB;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)NS * slot_bytes);
    uint64_t* empty = full + NS;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const FpGroup g = p.grp[gy];
    const uint32_t s_base = smem_u32(smem_raw);

    if (tid == 0) {
        for (int s = 0; s < NS; ++s) { mbar_init(&full[s], 1 + FPT_NH); mbar_init(&empty[s], FPT_NW); }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp >= FPT_NW) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        // ---------------- helper warps: thread = half a run (4 rays); warp FPT_NW lane 0 also issues the TMA ----------------
        const int h = tid - FPT_NW * 32;
        const int run = h >> 1, half = h & 1;
        const int ai = run >> 3, dsr = run & 7;
        const bool angle_ok = ai < g.count;
        const FpAngle A = p.ang[g.first + (angle_ok ? ai : g.count - 1)];
        const CUtensorMap* tm = g.vertical ? &tmV : &tmH;
        const bool producer = (warp == FPT_NW) && (lane == 0);
        if (producer) tma_prefetch_desc(tm);
        const int* w0tab = p.win0 + ((size_t)gy * p.ndg + dg) * (size_t)(((N + p.TR - 1) / p.TR) * p.TR);
        // per-angle constants, same operations as fp_kernel / par_fp.cu:63-85,140-161
        const float cth = __cosf(A.theta), sth = __sinf(A.theta);
        float detStep, step;
        if (g.vertical) { detStep = A.detsize / cth; step = sth / cth; }
        else { detStep = -A.detsize / sth; step = cth / sth; }
        float m[4], fP[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            int d = dg * FP_DG + dsr * FP_RT + half * 4 + t; if (d > p.Nd - 1) d = p.Nd - 1;
            const float t1 = __fadd_rn(__fadd_rn(__fmaf_rn((float)p.Nd, -0.5f, (float)d), 0.5f), -A.offset);
            m[t] = __fmul_rn(t1, detStep);
            fP[t] = 0.0f;
        }
        for (int line = 0; line < N; ++line) {
            const int s = line % NS;
            if (line >= NS) mbar_wait(&empty[s], ((line / NS) - 1) & 1);
            const int w0 = w0tab[line];
            if (producer) {
                void* dst = smem_raw + (size_t)s * slot_bytes;
                if (g.vertical) tma_load_4d(dst, tm, &full[s], 0, xb0, w0, line);    // window along the columns of row `line`
                else            tma_load_4d(dst, tm, &full[s], 0, xb0, line, w0);    // window along the rows of column `line`
                mbar_arrive_expect_tx(&full[s], data_bytes);
            }
            const float Cl = TTV_MAGIC_C(w0);
            if ((line & 63) == 0) {
                const float t2 = __fadd_rn(__fmaf_rn((float)N, -0.5f, (float)line), 0.5f);
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    float f = __fmaf_rn(t2, step, m[t]);
                    f = __fmaf_rn((float)N, 0.5f, f);
                    f = __fadd_rn(f, -0.5f);
                    fP[t] = __fadd_rn(f, 0.5f);
                }
            } else {
#pragma unroll
                for (int t = 0; t < 4; ++t) fP[t] = __fadd_rn(fP[t], step);
            }
            int ci[4]; uint32_t a8[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const float tt = __fmaf_rd(fP[t], 256.0f, Cl);
                const uint32_t u = (uint32_t)(__float_as_int(tt) - TTV_MAGIC_BITS);
                ci[t] = min(max((int)(u >> 9), 0), CW - 2);          // inside the staged window by construction of the plan
                a8[t] = (u >> 1) & 0xFFu;                            // 256 * weight of the upper cell
            }
            const int cprev = __shfl_up_sync(0xffffffffu, ci[3], 1);   // last cell of the run's first half (lane - 1)
            uint32_t e[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int c = ci[t];
                const int cE = (c + 1) & ~1, cO = c | 1;
                const bool has_prev = (t > 0) || (half == 1);
                const int cp = (t > 0) ? ci[t > 0 ? t - 1 : 0] : cprev;
                const bool keepE = has_prev && (((cp + 1) & ~1) == cE);
                const bool keepO = has_prev && ((cp | 1) == cO);
                const uint32_t wE = (c & 1) ? a8[t] : 256u - a8[t];
                e[t] = (keepE ? 0xFFu : (uint32_t)cE) | ((keepO ? 0xFFu : (uint32_t)cO) << 8) | (wE << 16);
            }
            sts128u(s_base + (uint32_t)s * slot_bytes + data_bytes + (uint32_t)h * 16u, make_uint4(e[0], e[1], e[2], e[3]));
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[s]);
        }
        return;
    }

    // ---------------- consumer warps: quarter warp = run ----------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
    const int lx = lane & 7;
    const int run = warp * 4 + (lane >> 3);
    const int ai = run >> 3;                               // warp-uniform (8 runs per angle = 2 warps)
    const bool angle_ok = ai < g.count;
    const FpAngle A = p.ang[g.first + (angle_ok ? ai : g.count - 1)];
    const int dbase = dg * FP_DG + (run & 7) * FP_RT;

    Acc4 acc1[FP_RT], acc2[FP_RT];
#pragma unroll
    for (int t = 0; t < FP_RT; ++t) { acc1[t].zero(); acc2[t].zero(); }
    float4 E1 = make_float4(0.f, 0.f, 0.f, 0.f), E2 = E1, O1 = E1, O2 = E1;
    const uint32_t tab_off = data_bytes + (uint32_t)run * (FP_RT * 4u);

#pragma unroll 1
    for (int line = 0; line < N; ++line) {
        const int s = line % NS;
        mbar_wait(&full[s], (line / NS) & 1);
        if (angle_ok) {                                    // warps of a partially filled group only keep the ring moving
            const uint32_t slot = s_base + (uint32_t)s * slot_bytes;
            const uint4 ea = lds128u(slot + tab_off), eb = lds128u(slot + tab_off + 16u);
            const uint32_t ent[FP_RT] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y, eb.z, eb.w};
            const uint32_t lb = slot + (uint32_t)lx * 16u;
#pragma unroll
            for (int t = 0; t < FP_RT; ++t) {
                const uint32_t e = ent[t];
                if (XB == 2) { lds_cell_if<FPT_CELL>(E1, E2, entry_iE(e), lb); lds_cell_if<FPT_CELL>(O1, O2, entry_iO(e), lb); }
                else { lds_cell1_if<FPT_CELL>(E1, entry_iE(e), lb); lds_cell1_if<FPT_CELL>(O1, entry_iO(e), lb); }
                const float wE = entry_wE(e);
                const float wO = 256.0f - wE;
                acc1[t].fma(wE, E1); acc1[t].fma(wO, O1);
                if (XB == 2) { acc2[t].fma(wE, E2); acc2[t].fma(wO, O2); }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }

    // ---------------- epilogue ----------------
    float part = 0.0f;
    if (angle_ok) {
        float corr1;
        {
            const float cth = __cosf(A.theta), sth = __sinf(A.theta);
            if (g.vertical) corr1 = (cth < 0.0f) ? -1.0f / cth : 1.0f / cth;
            else corr1 = (sth > 0.0f) ? 1.0f / sth : -1.0f / sth;
        }
        const float cs = corr1 * (1.0f / 256.0f);
#pragma unroll
        for (int t = 0; t < FP_RT; ++t) {
            const int d = dbase + t;
            if (d < p.Nd) {
                if (xb0 < nxb) {
                    const float4 a4 = acc1[t].get();
                    fp_emit<EPI>(p, A.orig, d, xb0, lx, make_float4(a4.x * cs, a4.y * cs, a4.z * cs, a4.w * cs), part);
                }
                if (XB == 2 && xb0 + 1 < nxb) {
                    const float4 a4 = acc2[t].get();
                    fp_emit<EPI>(p, A.orig, d, xb0 + 1, lx, make_float4(a4.x * cs, a4.y * cs, a4.z * cs, a4.w * cs), part);
                }
            }
        }
    }
    if (EPI == FP_KL || EPI == FP_STORE_DIST) {
        double sred = warp_sum((double)part);
        if (lane == 0) atomicAdd(p.red, sred);
    }
}

template <int EPI, int XB = FPT_XB>
__global__ void __launch_bounds__(FPT_THREADS, 1)
fpt_kernel(const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmH, const FpParams p, int nxb, int NS)
{
    fpt_body<EPI, XB>(tmV, tmH, p, nxb, NS, blockIdx.x, blockIdx.y, blockIdx.z * XB);
}

// Both CTA sizes in ONE launch, against wave quantisation.  A CTA fills an SM, so a launch of n work items takes
// ceil(n / 148) rounds: 1536 items (512 slices x 180 angles x 512 detectors with 2 x-blocks per CTA) are 10.4 waves and
// run for 11 rounds, the 384 single-x-block items of a 64-slice slab 2.6 waves for 3.  Here the first `n2` items
// (linear index: detector group fastest, then angle group, then x-block pair) run with 2 x-blocks per CTA and each of the
// remaining pair items as TWO CTAs of one x-block (0.57 of the time of a pair item): the block scheduler hands out CTAs
// in index order, so the long CTAs go first and the short ones fill the last round (Engine::fp_mix_plan picks n2).
// Results do not depend on the split (the two bodies are bit-identical, tests/test_gpu_table_kernels.py).
template <int EPI>
__global__ void __launch_bounds__(FPT_THREADS, 1)
fptm_kernel(const __grid_constant__ CUtensorMap tmV2, const __grid_constant__ CUtensorMap tmH2,
            const __grid_constant__ CUtensorMap tmV1, const __grid_constant__ CUtensorMap tmH1,
            const FpParams p, int nxb, int NS2, int NS1, int ngroups, int n2)
{
    const int lin = blockIdx.x;
    const int item = lin < n2 ? lin : n2 + ((lin - n2) >> 1);
    const int dg = item % p.ndg, gy = (item / p.ndg) % ngroups, z = item / (p.ndg * ngroups);
    if (lin < n2) {
        fpt_body<EPI, 2>(tmV2, tmH2, p, nxb, NS2, dg, gy, 2 * z);
    } else {
        const int xb0 = 2 * z + ((lin - n2) & 1);
        if (xb0 >= nxb) return;                               // odd number of x-blocks: the last pair has one
        fpt_body<EPI, 1>(tmV1, tmH1, p, nxb, NS1, dg, gy, xb0);
    }
}

}  // namespace ttv
