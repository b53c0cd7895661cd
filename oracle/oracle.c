/*
 * oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C (float32) CPU restatement of the arithmetic of tomo_TV's GPU iterative-reconstruction
 * hot path.  It exists only so that tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs can check and time the CUDA path against something that runs without a
 * GPU.  The product package (tomo_tv_b200) never imports, links or executes anything from oracle/.
 *
 * Parity status: PINNED against (a) ASTRA's own analytic rectangle known-answer model
 * (thirdparty/astra-toolbox/tests/python/functional/test_line2d.py:476-510, tolerance 2e-2 with 8-bit
 * weights) in tests/test_oracle_kat.py and (b) outputs of the real reference extension
 * (oracle/_ref, built by oracle/build_ref.sh) run on a B200, stored under tests/golden/.
 *
 * Every function cites the reference file:line it restates.  Paths are relative to
 * /root/reference; A/ = thirdparty/astra-toolbox, T/ = tomofusion.
 *
 * Conventions: one 2-D slice is vol[row j][col k] (H rows, W cols); texture x <-> column,
 * texture y <-> row.  A sinogram is sino[angle][det].  3-D volumes are [x][j][k] row-major
 * (T/gpu/utils/container/Matrix3D.cpp:20-23), sinogram stacks [x][angle*Nd + det]
 * (T/gpu/reconstructor.py:54-56).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

typedef struct {
    /* forward-projector constants: getParParameters, A/src/GeometryUtil2D.cpp:107-130 */
    float angle, detsize, offset;
    /* back-projector constants: transferConstants, A/cuda/2d/par_bp.cu:173-179 */
    float sc, ss, off, scale;
    /* ray direction (float32) and orientation flag, A/cuda/2d/par_fp.cu:271 */
    float rayX, rayY;
    int32_t vertical;
} orc_angle_t;

/* ------------------------------------------------------------------------------------------
 * Geometry.  genParProjections (A/src/GeometryUtil2D.cpp:35-74): base vectors rotated by the
 * float32 angle in double arithmetic, stored as float32.  Volume window is [-N/2,N/2]^2 with unit
 * pixels so convertAstraGeometry_internal (A/cuda/2d/astra.cu:281-309) is the identity
 * (dx=dy=0, factor=1, fOutputScale=1).
 * ------------------------------------------------------------------------------------------ */
ORC_API void orc_geometry(int nang, int ndet, const float* angles, orc_angle_t* out)
{
    for (int i = 0; i < nang; ++i) {
        /* CProjectionGeometry2D::_check (A/src/ProjectionGeometry2D.cpp:98-102) wraps every angle
         * into [0, 2pi) in float32 arithmetic with the float32 constant PI (A/include/astra/Globals.h:108);
         * for negative tilt angles this perturbs the angle by ~2e-7 rad, which is visible at 1e-5. */
        const float twopi = 2 * 3.14159265358979323846264338328f;
        float af = angles[i];
        while (twopi <= af) af -= twopi;
        while (af < 0) af += twopi;
        double al = (double)af;
        double c = cos(al), s = sin(al);
        float bDetSX = (float)(ndet * 1.0 * -0.5f);
        /* ROTATE0: X' = X cos - Y sin ; Y' = X sin + Y cos */
        float rayX = (float)(0.0f * c - 1.0f * s);
        float rayY = (float)(0.0f * s + 1.0f * c);
        float detSX = (float)(bDetSX * c - 0.0f * s);
        float detSY = (float)(bDetSX * s + 0.0f * c);
        float detUX = (float)(1.0f * c - 0.0f * s);
        float detUY = (float)(1.0f * s + 0.0f * c);
        orc_angle_t* g = &out[i];
        g->rayX = rayX; g->rayY = rayY;
        g->vertical = (fabsf(rayX) <= fabsf(rayY)) ? 1 : 0;
        /* getParParameters */
        {
            double ux = detUX, uy = detUY;
            /* the denominator is float * float + float * float: evaluated in float32 (GeometryUtil2D.cpp:113) */
            volatile float rx2 = rayX * rayX, ry2 = rayY * rayY;
            volatile float den = rx2 + ry2;
            double t = (ux * rayX + uy * rayY) / (double)den;
            ux -= t * rayX; uy -= t * rayY;
            double ang = atan2(uy, ux);
            double norm2 = uy * uy + ux * ux;
            g->angle = (float)ang;
            g->detsize = (float)sqrt(norm2);
            g->offset = (float)(-0.5 * ndet - ((double)detSY * uy + (double)detSX * ux) / norm2);
        }
        /* par_bp.cu:173-179 */
        {
            /* `double d = fDetUX * fRayY - fDetUY * fRayX` etc.: the operands are float members, so products and
             * differences are rounded to float32 before the promotion (par_bp.cu:174-178; the reference's x86-64
             * build has no FMA).  Measured: with double products 55 voxels of golden B/kl_bp flip a weight by
             * 1/256; with float products none do. */
            volatile float p1 = detUX * rayY, p2 = detUY * rayX;
            volatile float df = p1 - p2;
            double d = (double)df;
            volatile float q1 = detSY * rayX, q2 = detSX * rayY;
            volatile float nf = q1 - q2;
            volatile float r1 = rayX * rayX, r2 = rayY * rayY;
            volatile float rs = r1 + r2;
            g->sc = (float)(rayY / d);
            g->ss = (float)(-rayX / d);
            g->off = (float)((double)nf / d);
            g->scale = (float)((double)sqrtf(rs) / fabs(d));
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Texture model (A/cuda/2d/util.cu:161-179 and :198-223: cudaFilterModeLinear, unnormalised
 * coordinates).  CUDA linear filtering: xB = x-0.5, i=floor(xB), alpha=frac(xB) held in 9-bit
 * fixed point with 8 fractional bits.  quant: 0 = exact weights, 1 = round-to-nearest 1/256
 * (what ASTRA's own test assumes, test_line2d.py:492-503; confirmed on B200 by tests/probe).
 * In every call site of the hot path the other coordinate is an exact texel centre, so the
 * fetch is one-dimensional.  mode: 0 = Border (zeros outside), 1 = Clamp (edge replicate).
 * ------------------------------------------------------------------------------------------ */
static inline void tex_coord(float x, int quant, int* i, float* a)
{
    if (quant) {
        /* coordinate -> fixed point with 8 fractional bits, round HALF UP (floor(256x + 0.5)): measured
         * against the real reference on B200 -- exact ties (e.g. 256*frac = 60.5, 118.5 in
         * tests/golden B_fp) go up, not to even.  Exact in float32 for |x| < 2^14. */
        int q = (int)floorf(x * 256.0f + 0.5f) - 128;
        *i = q >> 8;                    /* arithmetic shift = floor division */
        *a = (float)(q & 255) * (1.0f / 256.0f);
    } else {
        float xb = x - 0.5f, fl = floorf(xb);
        *i = (int)fl; *a = xb - fl;
    }
}
static inline float tex1d(const float* row, int n, float x, int mode, int quant)
{
    int i; float a;
    tex_coord(x, quant, &i, &a);
    float v0, v1;
    if (mode == 0) {
        v0 = (i >= 0 && i < n) ? row[i] : 0.0f;
        v1 = (i + 1 >= 0 && i + 1 < n) ? row[i + 1] : 0.0f;
    } else {
        int i0 = i < 0 ? 0 : (i > n - 1 ? n - 1 : i);
        int i1 = i + 1 < 0 ? 0 : (i + 1 > n - 1 ? n - 1 : i + 1);
        v0 = row[i0]; v1 = row[i1];
    }
    return (1.0f - a) * v0 + a * v1;
}

/* ------------------------------------------------------------------------------------------
 * Forward projection of ONE slice: sino += outscale * A vol.   A/cuda/2d/par_fp.cu:53-200
 * (kernels), :234-300 (chunking in 64-slice blocks, each block's partial sum is scaled and
 * added to global memory separately).  cosv/sinv: per-angle cos/sin of g[a].angle as float32;
 * the reference uses the __cosf/__sinf intrinsics (par_fp.cu:64-65,141-142) -- the caller may
 * pass the device-evaluated values; if NULL cosf/sinf are used.
 * ------------------------------------------------------------------------------------------ */
static void fp_slice(const float* vol, float* sino, int W, int H, int Na, int Nd,
                     const orc_angle_t* g, const float* cosv, const float* sinv,
                     float outscale, int quant, int a_begin, int a_end)
{
    const int CH = 64; /* g_blockSlices, par_fp.cu:49 */
    for (int a = a_begin; a < a_end; ++a) {
        float th = g[a].angle;
        float c = cosv ? cosv[a] : cosf(th);
        float s = sinv ? sinv[a] : sinf(th);
        float* out = sino + (size_t)a * Nd;
        if (g[a].vertical) {
            /* FPvertical_simple, par_fp.cu:130-200: step over rows, interpolate along columns */
            float detStep = g[a].detsize / c;
            float step = s / c;
            float corr = (c < 0.0f) ? -outscale / c : outscale / c;
            for (int d = 0; d < Nd; ++d) {
                for (int r0 = 0; r0 < H; r0 += CH) {
                    /* nvcc contracts the reference expression as m = t1*detStep (rounded), then
                     * fma(t2, step, m), fma(W, 0.5, .), -0.5, +0.5 (SASS of par_fp.cu, sm_100) */
                    float fP = fmaf((float)r0 - 0.5f * H + 0.5f, step,
                                    ((float)d - 0.5f * Nd + 0.5f - g[a].offset) * detStep);
                    fP = fmaf((float)W, 0.5f, fP); fP = fP - 0.5f; fP = fP + 0.5f;
                    int r1 = r0 + CH < H ? r0 + CH : H;
                    float acc = 0.0f;
                    for (int r = r0; r < r1; ++r) {
                        acc += tex1d(vol + (size_t)r * W, W, fP, 0, quant);
                        fP += step;
                    }
                    out[d] = fmaf(acc, corr, out[d]);
                }
            }
        } else {
            /* FPhorizontal_simple, par_fp.cu:53-125: step over columns, interpolate along rows */
            float detStep = -g[a].detsize / s;
            float step = c / s;
            float corr = (s > 0.0f) ? outscale / s : -outscale / s;
            for (int d = 0; d < Nd; ++d) {
                for (int c0 = 0; c0 < W; c0 += CH) {
                    float fP = fmaf((float)c0 - 0.5f * W + 0.5f, step,
                                    ((float)d - 0.5f * Nd + 0.5f - g[a].offset) * detStep);
                    fP = fmaf((float)H, 0.5f, fP); fP = fP - 0.5f; fP = fP + 0.5f;
                    int c1 = c0 + CH < W ? c0 + CH : W;
                    float acc = 0.0f;
                    for (int cc = c0; cc < c1; ++cc) {
                        /* 1-D fetch down column cc */
                        int j; float al;
                        tex_coord(fP, quant, &j, &al);
                        float v0 = (j >= 0 && j < H) ? vol[(size_t)j * W + cc] : 0.0f;
                        float v1 = (j + 1 >= 0 && j + 1 < H) ? vol[(size_t)(j + 1) * W + cc] : 0.0f;
                        acc += (1.0f - al) * v0 + al * v1;
                        fP += step;
                    }
                    out[d] = fmaf(acc, corr, out[d]);
                }
            }
        }
    }
}

ORC_API void orc_fp(const float* vol, float* sino, int W, int H, int Na, int Nd,
                    const orc_angle_t* g, const float* cosv, const float* sinv,
                    float outscale, int quant)
{
    fp_slice(vol, sino, W, H, Na, Nd, g, cosv, sinv, outscale, quant, 0, Na);
}

/* single-angle FP used by SART (A/cuda/2d/sart.cu:256-271): out[Nd] += outscale * A_a vol */
ORC_API void orc_fp_angle(const float* vol, float* out, int W, int H, int Na, int Nd,
                          const orc_angle_t* g, const float* cosv, const float* sinv,
                          float outscale, int quant, int a)
{
    fp_slice(vol, out - (size_t)a * Nd, W, H, Na, Nd, g, cosv, sinv, outscale, quant, a, a + 1);
}

/* ------------------------------------------------------------------------------------------
 * Back projection of ONE slice: vol += outscale * B sino.  devBP, A/cuda/2d/par_bp.cu:52-87;
 * 16 angles per launch (:209-215), each launch's partial sum added to global memory.
 * ------------------------------------------------------------------------------------------ */
ORC_API void orc_bp(float* vol, const float* sino, int W, int H, int Na, int Nd,
                    const orc_angle_t* g, float outscale, int quant)
{
    const int CH = 16; /* g_anglesPerBlock, par_bp.cu:40 */
    for (int Y = 0; Y < H; ++Y) {
        float fY = (float)Y - 0.5f * H + 0.5f;
        for (int X = 0; X < W; ++X) {
            float fX = (float)X - 0.5f * W + 0.5f;
            for (int a0 = 0; a0 < Na; a0 += CH) {
                int a1 = a0 + CH < Na ? a0 + CH : Na;
                float acc = 0.0f;
                for (int a = a0; a < a1; ++a) {
                    /* SASS of devBP: m = fY*ss (rounded); fma(fX, sc, -m); + off; fma(tex, scale, acc) */
                    float fT = fmaf(fX, g[a].sc, -(fY * g[a].ss)) + g[a].off;
                    acc = fmaf(tex1d(sino + (size_t)a * Nd, Nd, fT, 0, quant), g[a].scale, acc);
                }
                vol[(size_t)Y * W + X] = fmaf(acc, outscale, vol[(size_t)Y * W + X]);
            }
        }
    }
}

/* devBP_SART, par_bp.cu:139-159 + BP_SART :260-297: one angle, Clamp addressing, no per-angle
 * scale.  proj is the single detector row. */
ORC_API void orc_bp_sart(float* vol, const float* proj, int W, int H, int Nd,
                         const orc_angle_t* ga, float outscale, int quant)
{
    for (int Y = 0; Y < H; ++Y) {
        float fY = (float)Y - 0.5f * H + 0.5f;
        for (int X = 0; X < W; ++X) {
            float fX = (float)X - 0.5f * W + 0.5f;
            float fT = fmaf(fX, ga->sc, -(fY * ga->ss)) + ga->off;
            vol[(size_t)Y * W + X] = fmaf(tex1d(proj, Nd, fT, 1, quant), outscale, vol[(size_t)Y * W + X]);
        }
    }
}

/* opInvert, A/cuda/2d/arith.cu:94-101 */
static inline float inv_pos(float x) { return x > 0.000001f ? 1.0f / x : 0.0f; }

/* SIRT::precomputeWeights, A/cuda/2d/sirt.cu:112-147 (no masks, relaxation 1):
 * R = inv(FP(1)) [Na*Nd], C = inv(BP(1)) [H*W]. */
ORC_API void orc_sirt_weights(float* R, float* C, int W, int H, int Na, int Nd,
                              const orc_angle_t* g, const float* cosv, const float* sinv, int quant)
{
    size_t nv = (size_t)W * H, ns = (size_t)Na * Nd;
    float* ones_v = (float*)malloc(nv * sizeof(float));
    float* ones_s = (float*)malloc(ns * sizeof(float));
    for (size_t i = 0; i < nv; ++i) ones_v[i] = 1.0f;
    for (size_t i = 0; i < ns; ++i) ones_s[i] = 1.0f;
    memset(R, 0, ns * sizeof(float));
    orc_fp(ones_v, R, W, H, Na, Nd, g, cosv, sinv, 1.0f, quant);
    for (size_t i = 0; i < ns; ++i) R[i] = inv_pos(R[i]);
    memset(C, 0, nv * sizeof(float));
    orc_bp(C, ones_s, W, H, Na, Nd, g, 1.0f, quant);
    for (size_t i = 0; i < nv; ++i) C[i] = inv_pos(C[i]);
    free(ones_v); free(ones_s);
}

/* SIRT::iterate, A/cuda/2d/sirt.cu:237-277, with setConstraints(min=0)
 * (T/gpu/utils/tomoengine.cpp:184).  One slice, n iterations, in place. */
ORC_API void orc_sirt_slice(float* vol, const float* b, const float* R, const float* C,
                            int W, int H, int Na, int Nd, const orc_angle_t* g,
                            const float* cosv, const float* sinv, int niter, int quant)
{
    size_t nv = (size_t)W * H, ns = (size_t)Na * Nd;
    float* p = (float*)malloc(ns * sizeof(float));
    float* t = (float*)malloc(nv * sizeof(float));
    for (int it = 0; it < niter; ++it) {
        memcpy(p, b, ns * sizeof(float));
        orc_fp(vol, p, W, H, Na, Nd, g, cosv, sinv, -1.0f, quant);
        for (size_t i = 0; i < ns; ++i) p[i] *= R[i];
        memset(t, 0, nv * sizeof(float));
        orc_bp(t, p, W, H, Na, Nd, g, 1.0f, quant);
        for (size_t i = 0; i < nv; ++i) {
            float x = vol[i] + C[i] * t[i];        /* opAddMul */
            vol[i] = x < 0.0f ? 0.0f : x;          /* opClampMin 0 */
        }
    }
    free(p); free(t);
}

/* SART::iterate, A/cuda/2d/sart.cu:175-227; order[] = projection order
 * (A/src/CudaSartAlgorithm.cpp:109-136); it0 = persistent iteration counter. */
ORC_API void orc_sart_slice(float* vol, const float* b, const float* R, int W, int H, int Na, int Nd,
                            const orc_angle_t* g, const float* cosv, const float* sinv,
                            const int32_t* order, int it0, int nsub, float relax, int quant)
{
    size_t nv = (size_t)W * H;
    float* p = (float*)malloc((size_t)Nd * sizeof(float));
    for (int it = it0; it < it0 + nsub; ++it) {
        int a = order ? order[it % Na] : it % Na;
        memcpy(p, b + (size_t)a * Nd, (size_t)Nd * sizeof(float));
        orc_fp_angle(vol, p, W, H, Na, Nd, g, cosv, sinv, -1.0f, quant, a);
        for (int d = 0; d < Nd; ++d) p[d] *= R[(size_t)a * Nd + d];   /* MUL_SART */
        orc_bp_sart(vol, p, W, H, Nd, &g[a], relax, quant);
        for (size_t i = 0; i < nv; ++i) if (vol[i] < 0.0f) vol[i] = 0.0f;
    }
    free(p);
}

/* ------------------------------------------------------------------------------------------
 * 3-D drivers (slice loops of T/gpu/utils/tomoengine.cpp).  OpenMP over slices is only a
 * speed-up of the checker; arithmetic per slice is unchanged.
 * ------------------------------------------------------------------------------------------ */
/* tomoengine::forwardProjection, tomoengine.cpp:416-427: g = FP(recon), fresh zeroed sinogram */
ORC_API void orc_fp3d(const float* vol, float* sino, int Nx, int N, int Na, const orc_angle_t* g,
                      const float* cosv, const float* sinv, int quant)
{
    size_t nv = (size_t)N * N, ns = (size_t)Na * N;
#pragma omp parallel for schedule(dynamic)
    for (int x = 0; x < Nx; ++x) {
        memset(sino + x * ns, 0, ns * sizeof(float));
        orc_fp(vol + x * nv, sino + x * ns, N, N, Na, N, g, cosv, sinv, 1.0f, quant);
    }
}
/* tomoengine::back_projection, tomoengine.cpp:279-291 applied per slice (fresh zeroed volume) */
ORC_API void orc_bp3d(float* vol, const float* sino, int Nx, int N, int Na, const orc_angle_t* g, int quant)
{
    size_t nv = (size_t)N * N, ns = (size_t)Na * N;
#pragma omp parallel for schedule(dynamic)
    for (int x = 0; x < Nx; ++x) {
        memset(vol + x * nv, 0, nv * sizeof(float));
        orc_bp(vol + x * nv, sino + x * ns, N, N, Na, N, g, 1.0f, quant);
    }
}
/* tomoengine::SIRT, tomoengine.cpp:189-205 */
ORC_API void orc_sirt3d(float* vol, const float* b, const float* R, const float* C, int Nx, int N, int Na,
                        const orc_angle_t* g, const float* cosv, const float* sinv, int niter, int quant)
{
    size_t nv = (size_t)N * N, ns = (size_t)Na * N;
#pragma omp parallel for schedule(dynamic)
    for (int x = 0; x < Nx; ++x)
        orc_sirt_slice(vol + x * nv, b + x * ns, R, C, N, N, Na, N, g, cosv, sinv, niter, quant);
}
/* tomoengine::SART, tomoengine.cpp:162-179: Na*nIter sub-iterations per slice; the ASTRA object is
 * shared by all slices and its counter advances by Na*nIter per slice, so every slice starts at
 * order[it0 % Na] with it0 a multiple of Na -> it0 = 0 is equivalent. */
ORC_API void orc_sart3d(float* vol, const float* b, const float* R, int Nx, int N, int Na,
                        const orc_angle_t* g, const float* cosv, const float* sinv,
                        const int32_t* order, int niter, float relax, int quant)
{
    size_t nv = (size_t)N * N, ns = (size_t)Na * N;
#pragma omp parallel for schedule(dynamic)
    for (int x = 0; x < Nx; ++x)
        orc_sart_slice(vol + x * nv, b + x * ns, R, N, N, Na, N, g, cosv, sinv, order, 0, Na * niter, relax, quant);
}

/* L = max(BP(FP(1))), tomoengine.cpp:241-242, 370-371 */
ORC_API float orc_lipschitz(int N, int Na, const orc_angle_t* g, const float* cosv, const float* sinv, int quant)
{
    size_t nv = (size_t)N * N, ns = (size_t)Na * N;
    float* v = (float*)malloc(nv * sizeof(float));
    float* s = (float*)calloc(ns, sizeof(float));
    for (size_t i = 0; i < nv; ++i) v[i] = 1.0f;
    orc_fp(v, s, N, N, Na, N, g, cosv, sinv, 1.0f, quant);
    memset(v, 0, nv * sizeof(float));
    orc_bp(v, s, N, N, Na, N, g, 1.0f, quant);
    float m = v[0];
    for (size_t i = 1; i < nv; ++i) if (v[i] > m) m = v[i];
    free(v); free(s);
    return m;
}

/* tomoengine::poisson_ML, tomoengine.cpp:293-315: per slice Ax = FP(x);
 * x -= (lambda/L) * BP((Ax-b)/(Ax+eps)); cost += sum(Ax - b*log(Ax+eps)); then positivity.
 * Returns the cost (float32 accumulation per slice as Eigen .sum(), summed over slices). */
ORC_API float orc_kl3d(float* vol, const float* b, int Nx, int N, int Na, const orc_angle_t* g,
                       const float* cosv, const float* sinv, float lambda, float L, int quant)
{
    size_t nv = (size_t)N * N, ns = (size_t)Na * N;
    const float eps = 1e-1f;
    double cost = 0.0;
    float step = lambda / L;
#pragma omp parallel for schedule(dynamic) reduction(+:cost)
    for (int x = 0; x < Nx; ++x) {
        float* Ax = (float*)calloc(ns, sizeof(float));
        float* u = (float*)calloc(nv, sizeof(float));
        float* xs = vol + x * nv;
        const float* bs = b + x * ns;
        orc_fp(xs, Ax, N, N, Na, N, g, cosv, sinv, 1.0f, quant);
        double c = 0.0;
        for (size_t i = 0; i < ns; ++i) {
            c += (double)(Ax[i] - bs[i] * logf(Ax[i] + eps));
            Ax[i] = (Ax[i] - bs[i]) / (Ax[i] + eps);
        }
        orc_bp(u, Ax, N, N, Na, N, g, 1.0f, quant);
        for (size_t i = 0; i < nv; ++i) {
            float v = xs[i] - step * u[i];
            xs[i] = v < 0.0f ? 0.0f : v;      /* recon.positivity(), tomoengine.cpp:313 */
        }
        cost += c;
        free(Ax); free(u);
    }
    return (float)cost;
}

/* ------------------------------------------------------------------------------------------
 * TV value: tv_3D_kernel, T/gpu/utils/regularizers/tv_gd.cu:27-47 (periodic forward
 * differences, eps=1e-6); summed (thrust::reduce, fp32; here double accumulation).
 * ------------------------------------------------------------------------------------------ */
#define IDX(i, j, k) (((size_t)(i) * ny + (j)) * nz + (k))
ORC_API double orc_tv(const float* v, int nx, int ny, int nz)
{
    const float eps = 1e-6f;
    double tot = 0.0;
#pragma omp parallel for reduction(+:tot)
    for (int i = 0; i < nx; ++i)
        for (int j = 0; j < ny; ++j)
            for (int k = 0; k < nz; ++k) {
                float c = v[IDX(i, j, k)];
                float dx = c - v[IDX((i + 1) % nx, j, k)];
                float dy = c - v[IDX(i, (j + 1) % ny, k)];
                float dz = c - v[IDX(i, j, (k + 1) % nz)];
                tot += (double)sqrtf(eps + dx * dx + dy * dy + dz * dz);
            }
    return tot;
}

/* out-of-range reads of the reference's gradient kernel are undefined; we define them as 0 */
static inline float rd(const float* v, long idx, long n) { return (idx >= 0 && idx < n) ? v[idx] : 0.0f; }

/* tv_gradient_3D_kernel, tv_gd.cu:52-106 -- including its index quirks:
 *   ip_jm uses (i+nx)%nx (== i), and (k-1-nz)%nz is evaluated with C truncation (negative). */
ORC_API void orc_tv_gradient(const float* r, float* g, int nx, int ny, int nz)
{
    const float eps = 1e-6f;
    long n = (long)nx * ny * nz;
#pragma omp parallel for
    for (int i = 0; i < nx; ++i)
        for (int j = 0; j < ny; ++j)
            for (int k = 0; k < nz; ++k) {
                long P = (long)ny * nz;
                long ijk = P * i + (long)nz * j + k;
                long ip = P * ((i + 1) % nx) + (long)nz * j + k;
                long im = P * ((i - 1 + nx) % nx) + (long)nz * j + k;
                long jp = P * i + (long)nz * ((j + 1) % ny) + k;
                long jm = P * i + (long)nz * ((j - 1 + ny) % ny) + k;
                long kp = P * i + (long)nz * j + ((k + 1) % nz);
                long km = P * i + (long)nz * j + ((k - 1 + nz) % nz);
                long im_jp = P * ((i - 1 + nx) % nx) + (long)nz * ((j + 1) % ny) + k;
                long ip_jm = P * ((i + nx) % nx) + (long)nz * ((j - 1 + ny) % ny) + k;
                long jm_kp = P * i + (long)nz * ((j - 1 + ny) % ny) + ((k + 1) % nz);
                long jp_km = P * i + (long)nz * ((j + 1) % ny) + ((k - 1 - nz) % nz);
                long im_kp = P * ((i - 1 + nx) % nx) + (long)nz * j + ((k + 1) % nz);
                long ip_km = P * ((i + 1) % nx) + (long)nz * j + ((k - 1 - nz) % nz);
                float c = r[ijk];
                float xip = r[ip], xim = r[im], xjp = r[jp], xjm = r[jm], xkp = r[kp], xkm = r[km];
                float v1n = (float)(3.0 * (double)c - (double)xip - (double)xjp - (double)xkp);
                float v1d = sqrtf(eps + (c - xip) * (c - xip) + (c - xjp) * (c - xjp) + (c - xkp) * (c - xkp));
                float v2n = c - xim;
                float a1 = r[im_jp], a2 = r[im_kp];
                float v2d = sqrtf(eps + (xim - c) * (xim - c) + (xim - a1) * (xim - a1) + (xim - a2) * (xim - a2));
                float v3n = c - xjm;
                float b1 = r[ip_jm], b2 = r[jm_kp];
                float v3d = sqrtf(eps + (xjm - b1) * (xjm - b1) + (xjm - c) * (xjm - c) + (xjm - b2) * (xjm - b2));
                float v4n = c - xkm;
                float c1 = rd(r, ip_km, n), c2 = rd(r, jp_km, n);
                float v4d = sqrtf(eps + (xkm - c1) * (xkm - c1) + (xkm - c2) * (xkm - c2) + (xkm - c) * (xkm - c));
                g[ijk] = v1n / v1d + v2n / v2d + v3n / v3d + v4n / v4d;
            }
}

/* cuda_tv_gd_3D, tv_gd.cu:141-218: returns TV of the INPUT; ng x { g = grad; x -= dPOCS*g/||g|| };
 * then positivity. */
ORC_API float orc_tv_gd(float* x, int ng, float dPOCS, int nx, int ny, int nz)
{
    size_t n = (size_t)nx * ny * nz;
    float tv0 = (float)orc_tv(x, nx, ny, nz);
    float* g = (float*)malloc(n * sizeof(float));
    for (int it = 0; it < ng; ++it) {
        orc_tv_gradient(x, g, nx, ny, nz);
        double s = 0.0;
#pragma omp parallel for reduction(+:s)
        for (long q = 0; q < (long)n; ++q) s += (double)(g[q] * g[q]);
        float nrm = sqrtf((float)s);
#pragma omp parallel for
        for (long q = 0; q < (long)n; ++q) x[q] -= dPOCS * g[q] / nrm;
    }
#pragma omp parallel for
    for (long q = 0; q < (long)n; ++q) if (x[q] < 0.0f) x[q] = 0.0f;
    free(g);
    return tv0;
}

/* cuda_tv_fgp_3D, T/gpu/utils/regularizers/tv_fgp.cu:192-281 (kernels :44-115):
 * returns TV(input); P=0; iter x { D = in - lambda*div(P) ; D=max(D,0) ; P += grad(D)/(26 lambda) ;
 * project P onto the unit ball }; the volume is replaced by the LAST D (:272).  iter == 0 returns
 * the TV map in place of the volume (:232,272). */
ORC_API float orc_tv_fgp(float* vol, int iter, float lambda, int nx, int ny, int nz)
{
    size_t n = (size_t)nx * ny * nz;
    float* D = (float*)calloc(n, sizeof(float));
    float* P1 = (float*)calloc(n, sizeof(float));
    float* P2 = (float*)calloc(n, sizeof(float));
    float* P3 = (float*)calloc(n, sizeof(float));
    const float eps = 1e-6f;
    double tot = 0.0;
#pragma omp parallel for reduction(+:tot)
    for (int i = 0; i < nx; ++i)
        for (int j = 0; j < ny; ++j)
            for (int k = 0; k < nz; ++k) {
                float c = vol[IDX(i, j, k)];
                float dx = c - vol[IDX((i + 1) % nx, j, k)];
                float dy = c - vol[IDX(i, (j + 1) % ny, k)];
                float dz = c - vol[IDX(i, j, (k + 1) % nz)];
                float t = sqrtf(eps + dx * dx + dy * dy + dz * dz);
                D[IDX(i, j, k)] = t;
                tot += (double)t;
            }
    float multip = 1.0f / (26.0f * lambda);
    for (int it = 0; it < iter; ++it) {
#pragma omp parallel for
        for (int i = 0; i < nx; ++i)
            for (int j = 0; j < ny; ++j)
                for (int k = 0; k < nz; ++k) {
                    size_t q = IDX(i, j, k);
                    float v1 = i <= 0 ? 0.0f : P1[IDX(i - 1, j, k)];
                    float v2 = j <= 0 ? 0.0f : P2[IDX(i, j - 1, k)];
                    float v3 = k <= 0 ? 0.0f : P3[IDX(i, j, k - 1)];
                    float d = vol[q] - lambda * (P1[q] + P2[q] + P3[q] - v1 - v2 - v3);
                    D[q] = d < 0.0f ? 0.0f : d;
                }
#pragma omp parallel for
        for (int i = 0; i < nx; ++i)
            for (int j = 0; j < ny; ++j)
                for (int k = 0; k < nz; ++k) {
                    size_t q = IDX(i, j, k);
                    float v1 = i >= nx - 1 ? 0.0f : D[q] - D[IDX(i + 1, j, k)];
                    float v2 = j >= ny - 1 ? 0.0f : D[q] - D[IDX(i, j + 1, k)];
                    float v3 = k >= nz - 1 ? 0.0f : D[q] - D[IDX(i, j, k + 1)];
                    float p1 = P1[q] + multip * v1, p2 = P2[q] + multip * v2, p3 = P3[q] + multip * v3;
                    float den = p1 * p1 + p2 * p2 + p3 * p3;
                    if (den > 1.0f) {
                        float sq = 1.0f / sqrtf(den);
                        p1 *= sq; p2 *= sq; p3 *= sq;
                    }
                    P1[q] = p1; P2[q] = p2; P3[q] = p3;
                }
    }
    memcpy(vol, D, n * sizeof(float));
    free(D); free(P1); free(P2); free(P3);
    return (float)tot;
}

/* nesterov_momentum_kernel, T/gpu/utils/container/matrix_ops.cu:77-86 +
 * tomoengine::fista_nesterov_momentum, tomoengine.cpp:381-384 */
ORC_API void orc_momentum(float* recon, float* yk, float* recon_old, float beta, size_t n)
{
    memcpy(recon, yk, n * sizeof(float));
#pragma omp parallel for
    for (long q = 0; q < (long)n; ++q) yk[q] = recon[q] + beta * (recon[q] - recon_old[q]);
    memcpy(recon_old, recon, n * sizeof(float));
}

/* cuda_soft_thresholding, matrix_ops.cu:64-75 */
ORC_API void orc_soft_threshold(float* v, float lambda, size_t n)
{
    for (size_t q = 0; q < n; ++q) {
        float a = fabsf(v[q]);
        v[q] = (float)(signbit(lambda - a) ? 1 : 0) * copysignf(a - lambda, v[q]);
    }
}
ORC_API void orc_positivity(float* v, size_t n) { for (size_t q = 0; q < n; ++q) if (v[q] < 0.0f) v[q] = 0.0f; }

/* cuda_euclidean_dist / cuda_rmse, matrix_ops.cu:174-264: sum (a-b)^2 */
ORC_API double orc_sqdist(const float* a, const float* b, size_t n)
{
    double s = 0.0;
#pragma omp parallel for reduction(+:s)
    for (long q = 0; q < (long)n; ++q) { float d = a[q] - b[q]; s += (double)(d * d); }
    return s;
}
ORC_API double orc_sum(const float* a, size_t n) { double s = 0; for (size_t q = 0; q < n; ++q) s += a[q]; return s; }
ORC_API double orc_l1(const float* a, size_t n) { double s = 0; for (size_t q = 0; q < n; ++q) s += fabsf(a[q]); return s; }
ORC_API double orc_sumsq(const float* a, size_t n) { double s = 0; for (size_t q = 0; q < n; ++q) s += (double)(a[q] * a[q]); return s; }

ORC_API int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
