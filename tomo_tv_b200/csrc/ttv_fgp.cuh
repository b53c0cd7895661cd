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
//   * 16 compute warps; a warp owns two columns (k0+w, k0+w+16) and marches down j with its
//     lanes = the 32 slices of a cell (one 128-byte line), so x-1 / x+1 neighbours are warp
//     shuffles, P2(j-1) / D(j+1) are register carries, and only P3(k-1) / D(k+1) come from the
//     neighbouring warp through shared memory (D ring).  Per stage: D for the 2 new rows, one
//     named barrier, then Pn for the 2 rows lagging by one (they need D(j+1)); straight-line
//     code, predicated stores.  The producer warp also evaluates D of slice x+1 for the block's
//     edge lane from the 16-byte boxes.  Nothing is re-read from HBM: k-halo columns and
//     x-halo sectors are L2 hits from the neighbouring CTAs (x-block is the fastest grid index
//     so the sharers run concurrently).
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
constexpr int FGP_PB = FGP_CW + 3;                 // compact ghost-plane box: 36 floats = 144 B per row
constexpr uint32_t FGP_PLANE_BOX = FGP_RJ * FGP_PB * 4u;
constexpr size_t FGP_SMEM = (size_t)FGP_NS * FGP_STAGE + (size_t)FGP_DR * 32 * 128 + (size_t)FGP_DR * 32 * 4 + 128 + 2048 + FGP_NS * 8;

enum FgpMode { FGP_SWEEP = 0, FGP_FIRST = 1, FGP_FINAL = 2 };

// main boxes, 16-byte halo boxes over the same arrays, and -- for a slab of a sharded volume -- the
// ghost cells [j][k][4] received from the neighbouring ranks (gr*: right neighbour's slice 0 in
// element 0; gl1: left neighbour's last slice of P1 in element 3).  has_gr / has_gl say whether they exist.
struct FgpMaps { CUtensorMap in, p1, p2, p3, inh, p1h, p2h, p3h, grin, gr1, gr2, gr3, gl1; };

// SHARD: the volume is a slab of a sharded one (ghost maps, peer stores); compiled out on a single GPU
template <int MODE, bool SHARD>
__global__ void __launch_bounds__(FGP_THREADS, 1)
fgp_tile_kernel(const __grid_constant__ FgpMaps tm, const FgpParams p, float* __restrict__ out, int seg)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t s_stage = smem_u32(smem_raw);
    const uint32_t s_D = s_stage + FGP_NS * FGP_STAGE;             // [DR][32 cells][32 lanes]
    const uint32_t s_X = s_D + FGP_DR * 32 * 128;                  // [DR][32 cells] D of slice x+1 (+128 B pad)
    const uint32_t s_E = s_X + FGP_DR * 32 * 4 + 128;              // [2 parities][4 planes][2 rows][32 cells] boundary-slice values for the neighbours
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)FGP_NS * FGP_STAGE + (size_t)FGP_DR * 32 * 128 + (size_t)FGP_DR * 32 * 4 + 128 + 2048);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.d.N;
    const int xb = blockIdx.x;
    const int k0 = blockIdx.y * FGP_TK;
    const int ja = blockIdx.z * seg, jb = min(ja + seg, N);
    const int nst = (jb - ja + 2 + FGP_RJ - 1) / FGP_RJ;
    const int nx = p.d.nx;
    const bool lastblk = (xb == p.d.nxb - 1);
    const int edge_lane = lastblk ? ((nx - 1) & 31) : 31;
    const bool has_xp = !lastblk || (SHARD && p.hr_in != nullptr);     // a slice x+1 exists for the block's edge lane
    const bool has_xm = xb > 0 || (SHARD && p.hl_P1 != nullptr);       // a slice x-1 exists for lane 0
    const float lambda = p.lambda, multip = p.multip;
    // sharded volume: ghost layout of the P slices (16-byte cells from NCCL, or compact planes written by the peer)
    const bool cR = SHARD && lastblk && p.compact, cL = SHARD && (xb == 0) && p.compact;
    // (a TMA box must start 16-byte aligned in the innermost dimension: the compact box starts at ka <= k0-1)
    const int ka = (k0 - 1) & ~3;
    const uint32_t rsR = cR ? FGP_PB * 4u : FGP_CW * 16u, csR = cR ? 4u : 16u;             // x+1 side: row / cell stride
    const uint32_t aoR = cR ? (uint32_t)(k0 - 1 - ka) * 4u : 0u;
    const uint32_t rsL = cL ? FGP_PB * 4u : FGP_CW * 16u, csL = cL ? 4u : 16u, eoL = cL ? (uint32_t)(k0 - 1 - ka) * 4u : 12u;
    const bool send_r = SHARD && (MODE != FGP_FINAL) && lastblk && (p.rm_gl1 != nullptr);           // my last slice -> right neighbour
    const bool send_l = SHARD && (MODE != FGP_FINAL) && (xb == 0) && (p.rm_gr1 != nullptr);         // my first slice -> left neighbour

    if (tid == 0) {
        for (int s = 0; s < FGP_NS; ++s) mbar_init(&full[s], 1);
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == FGP_NW) {
        // ---------------- producer + x+1 halo warp ----------------
        auto issue = [&](int s) {
            if (s >= nst) return;
            const int buf = s % FGP_NS;
            unsigned char* sb = smem_raw + (size_t)buf * FGP_STAGE;
            const int row = ja - 1 + s * FGP_RJ;
            if (row >= N) {                                     // rows past the volume: nothing to stage (never read for valid output)
                if (lane == 0) mbar_arrive(&full[buf]);
                return;
            }
            // box ids: 0..3 main in,P1,P2,P3 ; 4..7 x+1 halo in,P1,P2,P3 ; 8 x-1 halo P1
            bool want = false;
            if (lane < 9) {
                if (MODE == FGP_FIRST) want = (lane == 0) || (lane == 4 && has_xp);
                else if (MODE == FGP_FINAL) want = (lane < 4) || (lane == 8 && has_xm);
                else want = (lane < 4) || (lane < 8 && has_xp) || (lane == 8 && has_xm);
            }
            if (want) {
                const int id = lane;
                if (id < 4) {
                    const CUtensorMap* m = id == 0 ? &tm.in : id == 1 ? &tm.p1 : id == 2 ? &tm.p2 : &tm.p3;
                    tma_load_4d(sb + (size_t)id * FGP_MAIN_ARR, m, &full[buf], 0, k0 - 1, row, xb);
                } else if (id < 8) {
                    const int a = id - 4;
                    if (!lastblk) {
                        const CUtensorMap* m = a == 0 ? &tm.inh : a == 1 ? &tm.p1h : a == 2 ? &tm.p2h : &tm.p3h;
                        tma_load_4d(sb + 4 * FGP_MAIN_ARR + (size_t)a * FGP_HALO_ARR, m, &full[buf], 0, k0 - 1, row, xb + 1);
                    } else {
                        const CUtensorMap* m = a == 0 ? &tm.grin : a == 1 ? &tm.gr1 : a == 2 ? &tm.gr2 : &tm.gr3;
                        if (a > 0 && cR) tma_load_4d(sb + 4 * FGP_MAIN_ARR + (size_t)a * FGP_HALO_ARR, m, &full[buf], ka, row, 0, 0);
                        else tma_load_4d(sb + 4 * FGP_MAIN_ARR + (size_t)a * FGP_HALO_ARR, m, &full[buf], 0, k0 - 1, row, 0);
                    }
                } else if (xb > 0) {
                    tma_load_4d(sb + 4 * FGP_MAIN_ARR + 4 * FGP_HALO_ARR, &tm.p1h, &full[buf], 28, k0 - 1, row, xb - 1);
                } else if (cL) {
                    tma_load_4d(sb + 4 * FGP_MAIN_ARR + 4 * FGP_HALO_ARR, &tm.gl1, &full[buf], ka, row, 0, 0);
                } else {
                    tma_load_4d(sb + 4 * FGP_MAIN_ARR + 4 * FGP_HALO_ARR, &tm.gl1, &full[buf], 0, k0 - 1, row, 0);
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, want);
            if (lane == 0) {
                uint32_t bytes = (uint32_t)__popc(m & 0xFu) * FGP_MAIN_ARR + (uint32_t)__popc(m & 0x10u) * FGP_HALO_BOX;
                bytes += (uint32_t)__popc(m & 0xE0u) * (cR ? FGP_PLANE_BOX : FGP_HALO_BOX);
                bytes += (uint32_t)__popc(m & 0x100u) * (cL ? FGP_PLANE_BOX : FGP_HALO_BOX);
                mbar_arrive_expect_tx(&full[buf], bytes);
            }
        };
        // coalesced store of the boundary-slice values the compute warps left in shared memory for stage s
        // (rows rowbase-1 and rowbase, 31 cells each) into the neighbours' compact planes
        auto flush = [&](int s) {
            const int rowb = ja - 1 + s * FGP_RJ;
            const int k = k0 + lane;
            const uint32_t eb = s_E + (uint32_t)(s & 1) * 1024u + (uint32_t)lane * 4u;
#pragma unroll
            for (int rq = 0; rq < 2; ++rq) {
                const int j = rowb - 1 + rq;
                if (lane < FGP_TK && k < N && j >= ja && j < jb) {
                    const int64_t o = (int64_t)j * p.pitch + k;
                    if (send_r) p.rm_gl1[o] = lds32(eb + (uint32_t)(0 * 2 + rq) * 128u);
                    if (send_l) {
                        p.rm_gr1[o] = lds32(eb + (uint32_t)(1 * 2 + rq) * 128u);
                        p.rm_gr2[o] = lds32(eb + (uint32_t)(2 * 2 + rq) * 128u);
                        p.rm_gr3[o] = lds32(eb + (uint32_t)(3 * 2 + rq) * 128u);
                    }
                }
            }
        };
        for (int s = 0; s < FGP_NS - 1; ++s) issue(s);
        float hq2 = 0.0f;                                       // P2 of slice x+1 on the previous row
        for (int it = 0; it < nst; ++it) {
            // the buffer of stage it-1 was released by the barrier that closed iteration it-1
            issue(it + FGP_NS - 1);
            if (MODE != FGP_FINAL && has_xp) {
                const int buf = it % FGP_NS;
                mbar_wait(&full[buf], (it / FGP_NS) & 1);
                const uint32_t sb = s_stage + buf * FGP_STAGE;
                const int rowbase = ja - 1 + it * FGP_RJ;
#pragma unroll
                for (int r = 0; r < FGP_RJ; ++r) {
                    const int j = rowbase + r;
                    float D;
                    const float inn = lds32(sb + 4 * FGP_MAIN_ARR + (uint32_t)(r * FGP_CW + lane + 1) * 16u);
                    if (MODE == FGP_FIRST) D = fmaxf(inn, 0.0f);
                    else {
                        const uint32_t hb = sb + 4 * FGP_MAIN_ARR + (uint32_t)r * rsR + (uint32_t)(lane + 1) * csR + aoR;
                        const float q1 = lds32(hb + FGP_HALO_ARR), q2 = lds32(hb + 2 * FGP_HALO_ARR), q3 = lds32(hb + 3 * FGP_HALO_ARR);
                        const float v3 = lds32(hb + 3 * FGP_HALO_ARR - csR);
                        const float v1 = lds32(sb + FGP_MAIN_ARR + (uint32_t)(r * FGP_CW + lane + 1) * 128u + (uint32_t)edge_lane * 4u);
                        const float sum = ((((q1 + q2) + q3) - v1) - hq2) - v3;
                        D = fmaxf(__fmaf_rn(-lambda, sum, inn), 0.0f);
                        hq2 = q2;
                    }
                    sts32(s_X + (uint32_t)(((j & (FGP_DR - 1)) * 32 + lane) * 4), D);
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(FGP_THREADS) : "memory");
            if ((send_r || send_l) && it >= 1) flush(it - 1);           // stage it-1's Pn rows are complete
        }
        if (send_r || send_l) {
            asm volatile("bar.sync 1, %0;" ::"n"(FGP_THREADS) : "memory");
            flush(nst - 1);
            // every store of this CTA is performed system-wide before the CTA is counted; the last CTA of
            // each side publishes the epoch on the neighbour (release) -- its wait kernel acquires it
            __threadfence_system();
            __syncwarp();
            if (lane == 0) {
                const unsigned expect = gridDim.y * gridDim.z;
                if (send_r) {
                    if (atomicAdd(&p.cnt[0], 1u) == expect - 1) {
                        p.cnt[0] = 0; __threadfence_system();
                        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.rm_flagR), "r"(p.epoch) : "memory");
                    }
                }
                if (send_l) {
                    if (atomicAdd(&p.cnt[1], 1u) == expect - 1) {
                        p.cnt[1] = 0; __threadfence_system();
                        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.rm_flagL), "r"(p.epoch) : "memory");
                    }
                }
            }
        }
        return;
    }

    // ---------------- compute warps: columns c = warp and warp + 16 of the k-tile ----------------
    const bool valid = xb * 32 + lane < nx;
    const bool x_ok = valid && (lane != edge_lane || has_xp);   // the x-difference exists for this lane
    const uint32_t tc = (uint32_t)(warp + 1) * 128u + (uint32_t)lane * 4u;
    bool pk[2];
    pk[0] = (k0 + warp < N);
    pk[1] = (warp + 16 < FGP_TK) && (k0 + warp + 16 < N);
    bool kin[2];                                                // k < N-1: the k-difference exists
    kin[0] = (k0 + warp < N - 1);
    kin[1] = (k0 + warp + 16 < N - 1);
    // element offset of (row rowbase-1, column k0+warp, this lane); advanced by RJ rows per stage
    int64_t orow = (((int64_t)xb * N + (ja - 2)) * N + (k0 + warp)) * 32 + lane;
    const int64_t rstride = (int64_t)N * 32;

    float cP1[2] = {0.f, 0.f}, cP2[2] = {0.f, 0.f}, cP3[2] = {0.f, 0.f}, cD[2] = {0.f, 0.f};   // last row of the previous stage

    for (int it = 0; it < nst; ++it) {
        const int buf = it % FGP_NS;
        mbar_wait(&full[buf], (it / FGP_NS) & 1);
        const uint32_t sb = s_stage + buf * FGP_STAGE + tc;
        const int rowbase = ja - 1 + it * FGP_RJ;
        const uint32_t slotm = (uint32_t)((rowbase - 1) & (FGP_DR - 1)), slot0 = (uint32_t)(rowbase & (FGP_DR - 1)),
                       slot1 = (uint32_t)((rowbase + 1) & (FGP_DR - 1));

        float P1[2][2], P2[2][2], P3[2][2], D[2][2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
#pragma unroll
            for (int r = 0; r < FGP_RJ; ++r) {
                const uint32_t a = sb + (uint32_t)(h * 16 + r * FGP_CW) * 128u;
                const float inc = lds32(a);
                if (MODE == FGP_FIRST) {
                    P1[h][r] = P2[h][r] = P3[h][r] = 0.0f;
                    D[h][r] = fmaxf(inc, 0.0f);
                } else {
                    const float p1 = lds32(a + FGP_MAIN_ARR), p2 = lds32(a + 2 * FGP_MAIN_ARR), p3 = lds32(a + 3 * FGP_MAIN_ARR);
                    const float v3 = lds32(a + 3 * FGP_MAIN_ARR - 128u);
                    // slice 31 of x-block xb-1 / the left ghost cell (broadcast read; only lane 0 uses it)
                    const float hv = lds32(s_stage + buf * FGP_STAGE + 4 * FGP_MAIN_ARR + 4 * FGP_HALO_ARR
                                           + (uint32_t)r * rsL + (uint32_t)(warp + 1 + h * 16) * csL + eoL);
                    float v1 = __shfl_up_sync(0xffffffffu, p1, 1);
                    if (lane == 0) v1 = has_xm ? hv : 0.0f;
                    const float v2 = (r == 0) ? cP2[h] : P2[h][0];
                    const float sum = ((((p1 + p2) + p3) - v1) - v2) - v3;
                    P1[h][r] = p1; P2[h][r] = p2; P3[h][r] = p3;
                    D[h][r] = fmaxf(__fmaf_rn(-lambda, sum, inc), 0.0f);
                }
            }
        }
        if (MODE == FGP_FINAL) {
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int r = 0; r < FGP_RJ; ++r) {
                    const int j = rowbase + r;
                    if (pk[h] && j >= ja && j < jb) out[orow + (1 + r) * rstride + h * 512] = valid ? D[h][r] : 0.0f;
                }
            asm volatile("bar.sync 1, %0;" ::"n"(FGP_THREADS) : "memory");
        } else {
            const uint32_t dcol = (uint32_t)warp * 128u + (uint32_t)lane * 4u;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                sts32(s_D + (slot0 * 32 + h * 16) * 128u + dcol, D[h][0]);
                sts32(s_D + (slot1 * 32 + h * 16) * 128u + dcol, D[h][1]);
            }
            asm volatile("bar.sync 1, %0;" ::"n"(FGP_THREADS) : "memory");

            // ---- Pn of rows rowbase-1 (rq = 0, carried registers) and rowbase (rq = 1)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
#pragma unroll
                for (int rq = 0; rq < 2; ++rq) {
                    const int j = rowbase - 1 + rq;
                    const uint32_t slot = rq == 0 ? slotm : slot0;
                    const float Dc = rq == 0 ? cD[h] : D[h][0];
                    const float Dj = rq == 0 ? D[h][0] : D[h][1];
                    const float Dk = lds32(s_D + (slot * 32 + h * 16 + 1) * 128u + dcol);
                    const float Dxe = lds32(s_X + (slot * 32 + h * 16 + (uint32_t)warp) * 4u);
                    float Dx = __shfl_down_sync(0xffffffffu, Dc, 1);
                    if (lane == edge_lane) Dx = Dxe;
                    const float p1 = rq == 0 ? cP1[h] : P1[h][0], p2 = rq == 0 ? cP2[h] : P2[h][0], p3 = rq == 0 ? cP3[h] : P3[h][0];
                    const float val1 = x_ok ? Dc - Dx : 0.0f;
                    const float val2 = (j < N - 1) ? Dc - Dj : 0.0f;
                    const float val3 = kin[h] ? Dc - Dk : 0.0f;
                    float q1 = __fmaf_rn(multip, val1, p1);
                    float q2 = __fmaf_rn(multip, val2, p2);
                    float q3 = __fmaf_rn(multip, val3, p3);
                    const float den = __fmaf_rn(q3, q3, __fmaf_rn(q2, q2, __fmul_rn(q1, q1)));
                    if (den > 1.0f) {
                        const float sq = __fdiv_rn(1.0f, __fsqrt_rn(den));
                        q1 *= sq; q2 *= sq; q3 *= sq;
                    }
                    if (!valid) { q1 = q2 = q3 = 0.f; }
                    if (pk[h] && j >= ja && j < jb) {
                        const int64_t o = orow + rq * rstride + h * 512;
                        p.Q1[o] = q1; p.Q2[o] = q2; p.Q3[o] = q3;
                    }
                    if (send_r || send_l) {
                        const uint32_t eb = s_E + (uint32_t)(it & 1) * 1024u + (uint32_t)rq * 128u + (uint32_t)(warp + h * 16) * 4u;
                        if (send_r && lane == edge_lane) sts32(eb, q1);
                        if (send_l && lane == 0) { sts32(eb + 256u, q1); sts32(eb + 512u, q2); sts32(eb + 768u, q3); }
                    }
                }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) { cP1[h] = P1[h][1]; cP2[h] = P2[h][1]; cP3[h] = P3[h][1]; cD[h] = D[h][1]; }
        }
        if (MODE == FGP_FINAL) {
#pragma unroll
            for (int h = 0; h < 2; ++h) cP2[h] = P2[h][1];
        }
        orow += FGP_RJ * rstride;
    }
    if (send_r || send_l) asm volatile("bar.sync 1, %0;" ::"n"(FGP_THREADS) : "memory");
}

}  // namespace ttv
