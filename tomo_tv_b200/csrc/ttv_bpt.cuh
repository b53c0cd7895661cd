// Table-driven parallel-beam back projector: 256 slices per CTA, warp-uniform voxel runs, cells reused in
// registers.  sm_100a.  Default back projector for slabs of >= 5 x-blocks (ttv_engine.cuh: Engine::bp).
//
// Arithmetic: the reference's devBP (A/cuda/2d/par_bp.cu:52-87), sample position, 8-bit texture weights and
// per-angle scale as in bp_kernel (ttv_bp.cuh); both kernels accumulate the two taps of a sample even-cell-first
// over ascending angles and their weights are the same single roundings of the same real numbers (256*ws' == scale
// exactly), so bpt_kernel and bp_kernel are bit-identical (tests/test_gpu_table_kernels.py).
//
// Why (profiles/r2b_*): bp_kernel spends 8 B of shared-memory reads and ~5.5 instructions per sample; ncu shows
// its LDS pipe 95 % busy.  Register-reuse variants with per-lane index arithmetic halved the LDS wavefronts but
// became issue / latency bound at the same run time.  This kernel removes both costs from the hot loop:
//   * CTA = 8 x-blocks (256 slices) x a 16 x 8 voxel tile; staged detector cells are x-block-interleaved
//     ([cell][xb][32], one 4-D TMA box per angle over a tensor map whose second dimension is the x-block).
//     A consumer warp owns ONE run of 8 consecutive voxels and all 256 slices: lane l holds slice quad (l & 7)
//     of x-blocks 2*(l >> 3) and 2*(l >> 3) + 1 (two LDS.128 128 B apart), i.e. 8 slices x 8 voxels = 64
//     accumulators.  Cell index, weights and the reuse pattern are warp-uniform.
//   * One helper warp evaluates, per (angle, voxel), the sample position with the reference's float32 operations
//     ONCE per CTA (bp_kernel: once per 8 lanes per x-block, i.e. 64 times) and leaves a 4-byte entry in shared
//     memory: [even-cell index | odd-cell index | weight of the even cell (0..256)], index 0xFF = "the run holds
//     this cell".  A consumer fetches the 8 entries of its run with two LDS.128 per angle (what counts on the LDS
//     pipe is the 128 B/clk returned to the register file, so a broadcast 16-byte entry per voxel would cost as much
//     as a cell: the first version of this kernel did that and ran at 4.46 ms instead of bp_kernel's 5.57).
//   * "Parity registers": consecutive voxels of a run move the detector coordinate by |sc| <= 1 cell, so of the
//     two cells (c, c+1) of a sample one is even and one is odd and each is usually the cell held from the previous
//     voxel: a run loads 2 + 7|cos| cells instead of 16 (2.5x fewer LDS wavefronts on average).
//   * setmaxnreg: the helper warpgroup drops to 40 registers, the four consumer warpgroups grow to 104 (640 x 96
//     launch registers = 512 x 104 + 128 x 40 + slack: setmaxnreg.inc can only take what .dec released).
#pragma once
#include "ttv_bp.cuh"

namespace ttv {

constexpr int BPT_VT = 8;                       // voxels per run
constexpr int BPT_NW = 16;                      // consumer warps
constexpr int BPT_NS = 8;                       // angle slots in flight
constexpr int BPT_THREADS = (BPT_NW + 4) * 32;

// XB = x-blocks per CTA.  8: a warp is one run (all 32 lanes = 256 slices), tile 16 x 8 voxels, one helper warp.
//                        2: a quarter warp is one run (8 lanes = 64 slices), tile 32 x 16 voxels, four helper warps --
//                           the same kernel for the thin slabs of a sharded volume (64 or 128 slices per GPU).
template <int XB> struct BptCfg {
    static constexpr int LS = 4 * XB;                               // lanes per run
    static constexpr int RW = 32 / LS;                              // runs per warp
    static constexpr int RUNS = BPT_NW * RW;                        // runs per tile
    static constexpr int RX = (XB == 8) ? 2 : 4;                    // runs along X
    static constexpr int TX = RX * BPT_VT, TY = RUNS / RX;          // tile
    // detector cells per angle window: (TX-1)|sc| + (TY-1)|ss| <= hypot(TX-1, TY-1), + 2 cells of a sample, + one spare at
    // either end, + 1 for the even start: 16 x 8 -> 22, 32 x 16 -> 40
    static constexpr int CW = (XB == 8) ? 22 : 40;
    static constexpr uint32_t CELL = XB * 128u;                     // bytes of one staged cell
    static constexpr uint32_t DATA = CW * CELL;
    static constexpr uint32_t TAB = RUNS * BPT_VT * 4u;
    static constexpr uint32_t SLOT = DATA + TAB;
    static constexpr int NH = (RUNS * BPT_VT) / 128 > 0 ? ((RUNS * BPT_VT) / 128 > 4 ? 4 : (RUNS * BPT_VT) / 128) : 1;   // helper warps (4 entries per lane)
    static constexpr size_t SMEM = (size_t)BPT_NS * SLOT + 2 * BPT_NS * sizeof(uint64_t);
};

// two predicated LDS.128 (x-blocks 2q and 2q+1 of cell `idx`, CELLB bytes per cell): loaded unless idx == 0xFF
// ("the run holds this cell"); the registers keep their values otherwise
template <uint32_t CELLB>
__device__ __forceinline__ void lds_cell_if(float4& a, float4& b, uint32_t idx, uint32_t lane_base) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .u32 ad;\n\t"
        "setp.ne.u32 p, %8, 255;\n\t"
        "mad.lo.u32 ad, %8, %10, %9;\n\t"
        "@p ld.shared.v4.f32 {%0,%1,%2,%3}, [ad];\n\t"
        "@p ld.shared.v4.f32 {%4,%5,%6,%7}, [ad+128];\n\t}"
        : "+f"(a.x), "+f"(a.y), "+f"(a.z), "+f"(a.w), "+f"(b.x), "+f"(b.y), "+f"(b.z), "+f"(b.w)
        : "r"(idx), "r"(lane_base), "n"(CELLB));
}
// 32-bit table entry [even-cell index : 8 | odd-cell index : 8 | 256 * weight of the even cell : 16]
__device__ __forceinline__ uint32_t entry_iE(uint32_t e) { return e & 0xFFu; }
__device__ __forceinline__ uint32_t entry_iO(uint32_t e) { return __byte_perm(e, 0u, 0x4441u); }
__device__ __forceinline__ float entry_wE(uint32_t e) {          // bytes 2,3 of e under the exponent of 2^23, minus 2^23
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(e), "r"(0x4B000000u));
    return __uint_as_float(r) - 8388608.0f;
}
__device__ __forceinline__ uint4 lds128u(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts128u(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts32u(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

template <int EPI, int XB>
__global__ void __launch_bounds__(BPT_THREADS, 1)
bpt_kernel(const __grid_constant__ CUtensorMap tmS, const BpParams p, int nxb)
{
    using K = BptCfg<XB>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)BPT_NS * K::SLOT);
    uint64_t* empty = full + BPT_NS;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.N, Na = p.Na;
    const int x0 = blockIdx.x * K::TX, y0 = blockIdx.y * K::TY, xb0 = blockIdx.z * XB;

    if (tid == 0) {
        for (int s = 0; s < BPT_NS; ++s) { mbar_init(&full[s], K::NH); mbar_init(&empty[s], BPT_NW); }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp >= BPT_NW) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        const int hw = warp - BPT_NW;
        if (hw >= K::NH) return;
        // ---------------- helper warps: table builders; the first one is also the TMA producer ----------------
        if (hw == 0 && lane == 0) tma_prefetch_desc(&tmS);
        const float fX0 = (float)x0 - 0.5f * (float)N + 0.5f;
        const float fY0 = (float)y0 - 0.5f * (float)N + 0.5f;
        const uint32_t s_base = smem_u32(smem_raw);
        for (int a = 0; a < Na; ++a) {
            const int s = a % BPT_NS;
            if (a >= BPT_NS) mbar_wait(&empty[s], ((a / BPT_NS) - 1) & 1);
            const float4 k = c_bp_angle[a];
            const float ax0 = fX0 * k.x, ax1 = (fX0 + (float)(K::TX - 1)) * k.x;
            const float ay0 = -fY0 * k.y, ay1 = -(fY0 + (float)(K::TY - 1)) * k.y;
            const float fmin = fminf(ax0, ax1) + fminf(ay0, ay1) + k.z;
            const int w0 = ((int)floorf(fmin - 0.5f) - 1) & ~1;          // even start: window-relative parity = detector-cell parity (as bp_kernel)
            if (hw == 0 && lane == 0) tma_load_4d(smem_raw + (size_t)s * K::SLOT, &tmS, &full[s], 0, xb0, w0, a);
            const float Cl = TTV_MAGIC_C(w0);
#pragma unroll
            for (int i = 0; i < (K::RUNS * BPT_VT) / (32 * K::NH); ++i) {
                const int v = (i * K::NH + hw) * 32 + lane, run = v >> 3, t = v & 7;
                const float fX = (float)(x0 + (run % K::RX) * BPT_VT + t) - 0.5f * (float)N + 0.5f;
                const float fY = (float)(y0 + (run / K::RX)) - 0.5f * (float)N + 0.5f;
                const float myss = -__fmul_rn(fY, k.y);
                const float fT = __fadd_rn(__fmaf_rn(fX, k.x, myss), k.z);
                const float tt = __fmaf_rd(fT, 256.0f, Cl);
                const uint32_t u = (uint32_t)(__float_as_int(tt) - TTV_MAGIC_BITS);
                int ci = (int)(u >> 9);
                ci = min(max(ci, 0), K::CW - 2);                        // never leaves the window (margin in CW); keeps a bad geometry from faulting
                const uint32_t a8 = (u >> 1) & 0xFFu;                   // 256 * weight of the upper cell
                const int cE = (ci + 1) & ~1, cO = ci | 1;
                const int cip = __shfl_up_sync(0xffffffffu, ci, 1);
                const bool first = (t == 0);
                const bool keepE = !first && (((cip + 1) & ~1) == cE);
                const bool keepO = !first && ((cip | 1) == cO);
                const uint32_t wE = (ci & 1) ? a8 : 256u - a8;
                sts32u(s_base + (uint32_t)s * K::SLOT + K::DATA + (uint32_t)v * 4u,
                       (keepE ? 0xFFu : (uint32_t)cE) | ((keepO ? 0xFFu : (uint32_t)cO) << 8) | (wE << 16));
            }
            __syncwarp();
            if (lane == 0) {
                if (hw == 0) mbar_arrive_expect_tx(&full[s], K::DATA);
                else mbar_arrive(&full[s]);
            }
        }
        return;
    }

    // ---------------- consumer warps: run = warp (XB = 8) or quarter warp (XB = 2) ----------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
    const int lx = lane % K::LS;
    const int run = warp * K::RW + lane / K::LS;
    const int Y = y0 + run / K::RX;
    const int Xb = x0 + (run % K::RX) * BPT_VT;
    const int xbA = xb0 + (lx >> 3) * 2;
    const uint32_t lane_off = (uint32_t)(lx >> 3) * 256u + (uint32_t)(lx & 7) * 16u;

    Acc4 acc1[BPT_VT], acc2[BPT_VT];
#pragma unroll
    for (int t = 0; t < BPT_VT; ++t) { acc1[t].zero(); acc2[t].zero(); }
    const uint32_t s_base = smem_u32(smem_raw);
    float4 E1 = make_float4(0.f, 0.f, 0.f, 0.f), E2 = E1, O1 = E1, O2 = E1;

#pragma unroll 1
    for (int a = 0; a < Na; ++a) {
        const int s = a % BPT_NS;
        mbar_wait(&full[s], (a / BPT_NS) & 1);
        const uint32_t slot = s_base + (uint32_t)s * K::SLOT;
        const uint32_t tab = slot + K::DATA + (uint32_t)run * (BPT_VT * 4u);
        const uint32_t lb = slot + lane_off;
        const uint4 ea = lds128u(tab), eb = lds128u(tab + 16u);
        const uint32_t ent[BPT_VT] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y, eb.z, eb.w};
        const float sc256 = c_bp_angle[a].w;                 // per-angle scale (par_bp.cu:178); weights are in 1/256
        const float ws = sc256 * (1.0f / 256.0f);
#pragma unroll
        for (int t = 0; t < BPT_VT; ++t) {
            const uint32_t e = ent[t];
            lds_cell_if<K::CELL>(E1, E2, entry_iE(e), lb);
            lds_cell_if<K::CELL>(O1, O2, entry_iO(e), lb);
            const float w8 = entry_wE(e);
            const float wE = w8 * ws, wO = fmaf(-w8, ws, sc256);
            acc1[t].fma(wE, E1); acc1[t].fma(wO, O1);
            acc2[t].fma(wE, E2); acc2[t].fma(wO, O2);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }

    // ---------------- epilogue ----------------
    if (Y < N) {
#pragma unroll
        for (int t = 0; t < BPT_VT; ++t) {
            const int X = Xb + t;
            if (X < N) {
                const size_t oa = (((size_t)xbA * N + Y) * N + X) * TTV_XI + (size_t)(lx & 7) * 4;
                const size_t ob = oa + (size_t)N * N * TTV_XI;
                if (xbA < nxb) *reinterpret_cast<float4*>(p.vol + oa) = bp_apply<EPI>(p, oa, acc1[t].get(), Y, X);
                if (xbA + 1 < nxb) *reinterpret_cast<float4*>(p.vol + ob) = bp_apply<EPI>(p, ob, acc2[t].get(), Y, X);
            }
        }
    }
}

}  // namespace ttv
