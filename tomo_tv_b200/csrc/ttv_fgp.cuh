// TV-FGP sweep, tiled + TMA-staged: one pass reads in,P1,P2,P3 once and writes P1,P2,P3 once.  sm_100a.
//
// Arithmetic: the reference's Obj_func3D / nonneg / Grad_func3D / Proj_func3D kernels fused
// (T/gpu/utils/regularizers/tv_fgp.cu:44-115,143-154; host loop :192-281):
//   D(p)  = max(0, in(p) - lambda*((P1+P2+P3)(p) - P1(x-1) - P2(j-1) - P3(k-1)))   zero below the low edge
//   Pn(p) = proj( P(p) + multip*(D(p) - D(p+e)) )                                  zero difference at the high edge
//
// Design (the reference runs 4 uncoalesced launches + 4 host syncs per sweep):
//   * CTA = one x-block (32 slices) x a k-tile of 31 cells x a j-segment, marching along j.
//   * A producer warp streams RJ=2 rows per stage into a 4-stage shared-memory ring with TMA:
//     four 128-byte-cell boxes (in,P1,P2,P3; k0-1 .. k0+31) plus five 16-byte boxes that carry
//     the x-neighbours of the block edge (slice 0 of x-block xb+1 for in,P1,P2,P3; slice 31 of
//     x-block xb-1 for P1).  TMA out-of-bounds zero fill implements every low/high edge rule.
//   * 16 compute warps; a warp owns one cell (32 slices = one 128-byte line) per task, so the
//     x-1 / x+1 neighbours are warp shuffles.  Per stage: D for the 2 new rows (32 cells each,
//     incl. the k+1 halo column, + the x+1 halo lane) -> shared D ring; one named barrier;
//     then Pn for the 2 rows lagging by one (they need D(j+1)).  Nothing is re-read from HBM:
//     k-halo columns and x-halo sectors are L2 hits from the neighbouring CTAs (x-block is the
//     fastest grid index so the sharers run concurrently).
//   * MODE 1 (first sweep, P == 0) stages only `in`; MODE 2 (last sweep) writes D instead of Pn.
#pragma once
#include "ttv_common.cuh"
#include "ttv_tv.cuh"

namespace ttv {

constexpr int FGP_TK = 31;                 // output cells per k-tile
constexpr int FGP_CW = FGP_TK + 2;         // staged cells per row: k0-1 .. k0+TK
constexpr int FGP_RJ = 2;                  // rows per stage
constexpr int FGP_NW = 16;                 // compute warps
constexpr int FGP_NS = 4;                  // stages
constexpr int FGP_DR = 8;                  // D ring rows
constexpr int FGP_THREADS = (FGP_NW + 1) * 32;
constexpr uint32_t FGP_MAIN_ARR = FGP_RJ * FGP_CW * 128u;                 // bytes of one staged array
constexpr uint32_t FGP_HALO_BOX = FGP_RJ * FGP_CW * 16u;                  // bytes of one 16-byte-cell box
constexpr uint32_t FGP_HALO_ARR = (FGP_HALO_BOX + 127u) / 128u * 128u;    // padded to the TMA alignment
constexpr uint32_t FGP_STAGE = 4 * FGP_MAIN_ARR + 5 * FGP_HALO_ARR;
constexpr size_t FGP_SMEM = (size_t)FGP_NS * FGP_STAGE + (size_t)FGP_DR * 32 * 128 + (size_t)FGP_DR * 32 * 4 + 2 * FGP_NS * 8;

enum FgpMode { FGP_SWEEP = 0, FGP_FIRST = 1, FGP_FINAL = 2 };

struct FgpMaps { CUtensorMap in, p1, p2, p3, inh, p1h, p2h, p3h; };

template <int MODE>
__global__ void __launch_bounds__(FGP_THREADS, 1)
fgp_tile_kernel(const __grid_constant__ FgpMaps tm, const FgpParams p, float* __restrict__ out, int seg)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* Dsm = reinterpret_cast<float*>(smem_raw + (size_t)FGP_NS * FGP_STAGE);     // [DR][32 cells][32 lanes]
    float* Dxp = Dsm + FGP_DR * 32 * 32;                                               // [DR][32 cells]
    uint64_t* full = reinterpret_cast<uint64_t*>(Dxp + FGP_DR * 32);
    uint64_t* empty = full + FGP_NS;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.d.N;
    const int xb = blockIdx.x;
    const int k0 = blockIdx.y * FGP_TK;
    const int ja = blockIdx.z * seg, jb = min(ja + seg, N);
    const int nst = (jb - ja + 2 + FGP_RJ - 1) / FGP_RJ;

    if (tid == 0) {
        for (int s = 0; s < FGP_NS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], FGP_NW); }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == FGP_NW) {
        // ---------------- producer ----------------
        const bool lastb = (xb == p.d.nxb - 1);
        for (int it = 0; it < nst; ++it) {
            const int buf = it % FGP_NS;
            if (it >= FGP_NS) mbar_wait(&empty[buf], ((it / FGP_NS) - 1) & 1);
            unsigned char* sb = smem_raw + (size_t)buf * FGP_STAGE;
            const int row = ja - 1 + it * FGP_RJ;
            if (row >= N) {                                     // rows past the volume: nothing to stage (never read for valid output)
                if (lane == 0) mbar_arrive(&full[buf]);
                continue;
            }
            // box ids: 0..3 main in,P1,P2,P3 ; 4..7 x+1 halo in,P1,P2,P3 ; 8 x-1 halo P1
            bool want = false;
            if (lane < 9) {
                if (MODE == FGP_FIRST) want = (lane == 0) || (lane == 4 && !lastb);
                else if (MODE == FGP_FINAL) want = (lane < 4) || (lane == 8 && xb > 0);
                else want = (lane < 4) || (lane < 8 && !lastb) || (lane == 8 && xb > 0);
            }
            if (want) {
                const int id = lane;
                if (id < 4) {
                    const CUtensorMap* m = id == 0 ? &tm.in : id == 1 ? &tm.p1 : id == 2 ? &tm.p2 : &tm.p3;
                    tma_load_4d(sb + (size_t)id * FGP_MAIN_ARR, m, &full[buf], 0, k0 - 1, row, xb);
                } else if (id < 8) {
                    const int a = id - 4;
                    const CUtensorMap* m = a == 0 ? &tm.inh : a == 1 ? &tm.p1h : a == 2 ? &tm.p2h : &tm.p3h;
                    tma_load_4d(sb + 4 * FGP_MAIN_ARR + (size_t)a * FGP_HALO_ARR, m, &full[buf], 0, k0 - 1, row, xb + 1);
                } else {
                    tma_load_4d(sb + 4 * FGP_MAIN_ARR + 4 * FGP_HALO_ARR, &tm.p1h, &full[buf], 28, k0 - 1, row, xb - 1);
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, want);
            if (lane == 0) {
                const uint32_t bytes = (uint32_t)__popc(m & 0xFu) * FGP_MAIN_ARR + (uint32_t)__popc(m & 0x1F0u) * FGP_HALO_BOX;
                mbar_arrive_expect_tx(&full[buf], bytes);
            }
        }
        return;
    }

    // ---------------- compute warps ----------------
    const int nx = p.d.nx;
    const bool lastblk = (xb == p.d.nxb - 1);
    const bool valid = xb * 32 + lane < nx;
    const int edge_lane = lastblk ? ((nx - 1) & 31) : 31;
    const bool has_xp = !lastblk || (p.hr_in != nullptr);
    const float lambda = p.lambda, multip = p.multip;
    const int64_t cpb = (int64_t)N * N;

    auto mainp = [&](const unsigned char* sb, int a, int r, int cw) -> const float* {
        return reinterpret_cast<const float*>(sb + (size_t)a * FGP_MAIN_ARR) + (r * FGP_CW + cw) * 32;
    };
    auto halop = [&](const unsigned char* sb, int a, int r, int cw) -> const float* {
        return reinterpret_cast<const float*>(sb + 4 * FGP_MAIN_ARR + (size_t)a * FGP_HALO_ARR) + (r * FGP_CW + cw) * 4;
    };

    for (int it = 0; it < nst; ++it) {
        const int buf = it % FGP_NS;
        mbar_wait(&full[buf], (it / FGP_NS) & 1);
        const unsigned char* sb = smem_raw + (size_t)buf * FGP_STAGE;
        const unsigned char* sp = smem_raw + (size_t)((it + FGP_NS - 1) % FGP_NS) * FGP_STAGE;   // previous stage
        const int rowbase = ja - 1 + it * FGP_RJ;

        // ---- D of the new rows: cells c = 0..31 <-> k = k0 + c (c == 31 is the k+1 halo column)
#pragma unroll
        for (int i = 0; i < FGP_RJ * 32 / FGP_NW; ++i) {
            const int t = warp + FGP_NW * i;
            const int r = t >> 5, c = t & 31;
            const int j = rowbase + r;
            if (j < ja) continue;                               // warp-uniform
            const int k = k0 + c;
            const float inc = mainp(sb, 0, r, c + 1)[lane];
            float D;
            if (MODE == FGP_FIRST) {
                D = fmaxf(inc, 0.0f);
            } else {
                const float p1 = mainp(sb, 1, r, c + 1)[lane];
                const float p2 = mainp(sb, 2, r, c + 1)[lane];
                const float p3 = mainp(sb, 3, r, c + 1)[lane];
                const float v2 = (r > 0) ? mainp(sb, 2, r - 1, c + 1)[lane] : mainp(sp, 2, FGP_RJ - 1, c + 1)[lane];
                const float v3 = mainp(sb, 3, r, c)[lane];
                float v1 = __shfl_up_sync(0xffffffffu, p1, 1);
                if (lane == 0) {
                    if (xb > 0) v1 = halop(sb, 4, r, c + 1)[3];
                    else v1 = (p.hl_P1 && j < N && k < N) ? p.hl_P1[(int64_t)j * N + k] : 0.0f;
                }
                const float s = ((((p1 + p2) + p3) - v1) - v2) - v3;
                D = fmaxf(__fmaf_rn(-lambda, s, inc), 0.0f);
            }
            if (MODE == FGP_FINAL) {
                if (c < FGP_TK && k < N && j < jb) out[((int64_t)xb * cpb + (int64_t)j * N + k) * 32 + lane] = valid ? D : 0.0f;
            } else {
                Dsm[((j & (FGP_DR - 1)) * 32 + c) * 32 + lane] = D;
            }
        }
        if (MODE != FGP_FINAL) {
            // ---- D of slice x+1 for the edge lane: lane <-> cell c of row r = warp
            if (warp < FGP_RJ && has_xp) {
                const int r = warp, c = lane;
                const int j = rowbase + r, k = k0 + c;
                if (j >= ja) {
                    float D = 0.0f;
                    if (!lastblk) {
                        const float inn = halop(sb, 0, r, c + 1)[0];
                        if (MODE == FGP_FIRST) D = fmaxf(inn, 0.0f);
                        else {
                            const float q1 = halop(sb, 1, r, c + 1)[0], q2 = halop(sb, 2, r, c + 1)[0], q3 = halop(sb, 3, r, c + 1)[0];
                            const float v2 = (r > 0) ? halop(sb, 2, r - 1, c + 1)[0] : halop(sp, 2, FGP_RJ - 1, c + 1)[0];
                            const float v3 = halop(sb, 3, r, c)[0];
                            const float v1 = mainp(sb, 1, r, c + 1)[31];
                            const float s = ((((q1 + q2) + q3) - v1) - v2) - v3;
                            D = fmaxf(__fmaf_rn(-lambda, s, inn), 0.0f);
                        }
                    } else if (j < N && k < N) {
                        // right halo planes of the neighbouring rank (slab decomposition)
                        const int64_t jk = (int64_t)j * N + k;
                        const float inn = p.hr_in[jk];
                        if (MODE == FGP_FIRST) D = fmaxf(inn, 0.0f);
                        else {
                            const float q1 = p.hr_P1[jk], q2 = p.hr_P2[jk], q3 = p.hr_P3[jk];
                            const float v2 = j > 0 ? p.hr_P2[jk - N] : 0.0f;
                            const float v3 = k > 0 ? p.hr_P3[jk - 1] : 0.0f;
                            const float v1 = mainp(sb, 1, r, c + 1)[edge_lane];
                            const float s = ((((q1 + q2) + q3) - v1) - v2) - v3;
                            D = fmaxf(__fmaf_rn(-lambda, s, inn), 0.0f);
                        }
                    }
                    Dxp[(j & (FGP_DR - 1)) * 32 + c] = D;
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(FGP_NW * 32) : "memory");

            // ---- Pn of the rows lagging by one: jq = rowbase-1 .. rowbase+RJ-2
#pragma unroll
            for (int i = 0; i < (FGP_RJ * FGP_TK + FGP_NW - 1) / FGP_NW; ++i) {
                const int t = warp + FGP_NW * i;
                if (t >= FGP_RJ * FGP_TK) break;
                const int rq = t / FGP_TK, c = t - rq * FGP_TK;
                const int j = rowbase - 1 + rq, k = k0 + c;
                if (j < ja || j >= jb || k >= N) continue;      // warp-uniform
                const int slot = (j & (FGP_DR - 1)) * 32;
                const float Dc = Dsm[(slot + c) * 32 + lane];
                const float Dk = Dsm[(slot + c + 1) * 32 + lane];
                const float Dj = Dsm[((((j + 1) & (FGP_DR - 1)) * 32) + c) * 32 + lane];
                float Dx = __shfl_down_sync(0xffffffffu, Dc, 1);
                if (lane == edge_lane) Dx = Dxp[slot + c];
                float p1 = 0.f, p2 = 0.f, p3 = 0.f;
                if (MODE != FGP_FIRST) {
                    const unsigned char* s2 = rq == 0 ? sp : sb;
                    const int r2 = rq == 0 ? FGP_RJ - 1 : rq - 1;
                    p1 = mainp(s2, 1, r2, c + 1)[lane]; p2 = mainp(s2, 2, r2, c + 1)[lane]; p3 = mainp(s2, 3, r2, c + 1)[lane];
                }
                const float val1 = (valid && (lane != edge_lane || has_xp)) ? Dc - Dx : 0.0f;
                const float val2 = (j < N - 1) ? Dc - Dj : 0.0f;
                const float val3 = (k < N - 1) ? Dc - Dk : 0.0f;
                float q1 = __fmaf_rn(multip, val1, p1);
                float q2 = __fmaf_rn(multip, val2, p2);
                float q3 = __fmaf_rn(multip, val3, p3);
                const float den = __fmaf_rn(q3, q3, __fmaf_rn(q2, q2, __fmul_rn(q1, q1)));
                if (den > 1.0f) {
                    const float sq = __fdiv_rn(1.0f, __fsqrt_rn(den));
                    q1 *= sq; q2 *= sq; q3 *= sq;
                }
                if (!valid) { q1 = q2 = q3 = 0.f; }
                const int64_t o = ((int64_t)xb * cpb + (int64_t)j * N + k) * 32 + lane;
                p.Q1[o] = q1; p.Q2[o] = q2; p.Q3[o] = q3;
            }
        }
        __syncwarp();
        if (it >= 1 && lane == 0) mbar_arrive(&empty[(it + FGP_NS - 1) % FGP_NS]);
    }
}

}  // namespace ttv
