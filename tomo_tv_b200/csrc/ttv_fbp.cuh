// CGLS and FBP/WBP, batched over all slices of the slab.  sm_100a.
//
// CGLS  -- A/cuda/2d/cgls.cu:112-176 driven per slice by tomoengine::CGLS (T/gpu/utils/tomoengine.cpp:214-229):
//          every slice is its own conjugate-gradient run, so gamma, <w,w>, alpha and beta are PER-SLICE
//          scalars.  They live in device arrays indexed by slice; the vector updates read them, no host
//          round trip inside the iteration.
// FBP   -- A/cuda/2d/fbp.cu:105-190 + fft.cu:251-375 + A/src/Filters.cpp:40-420 driven by tomoengine::FBP
//          (tomoengine.cpp:330-347): pad each detector row to nextpow2(2N), cuFFT R2C, multiply with the real
//          filter, C2R, divide by the padded length, back-project with (pi/2)/Nangles.
#pragma once
#include "ttv_engine.cuh"
#include <cufft.h>

namespace ttv {

// out[x] (+)= sum over cells of a[x]^2 : per-slice squared norm of an x-interleaved array with `cells` cells per x-block
__global__ void __launch_bounds__(256) slice_sumsq_kernel(const float4* __restrict__ a, int64_t cells, double* __restrict__ out)
{
    const int xb = blockIdx.y, l8 = threadIdx.x & 7;
    const float4* base = a + (int64_t)xb * cells * 8;
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int64_t c = (int64_t)blockIdx.x * 32 + (threadIdx.x >> 3); c < cells; c += (int64_t)gridDim.x * 32) {
        const float4 v = base[c * 8 + l8];
        s0 += (double)(v.x * v.x); s1 += (double)(v.y * v.y); s2 += (double)(v.z * v.z); s3 += (double)(v.w * v.w);
    }
    // lanes with equal l8 hold the same 4 slices: reduce over the 4 quarter-warps, then over the 8 warps via atomics
#pragma unroll
    for (int o = 8; o < 32; o <<= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o); s3 += __shfl_xor_sync(0xffffffffu, s3, o);
    }
    if ((threadIdx.x & 31) < 8) {
        double* o = out + xb * 32 + l8 * 4;
        atomicAdd(o, s0); atomicAdd(o + 1, s1); atomicAdd(o + 2, s2); atomicAdd(o + 3, s3);
    }
}

enum CgOp { CG_ALPHA = 0, CG_BETA = 1 };
// CG_ALPHA: alpha[x] = gamma[x] / ww[x]                       (cgls.cu:150-151)
// CG_BETA : beta[x] = (1/gamma[x]) * gnew[x]; gamma[x] = gnew[x]   (cgls.cu:166-168)
__global__ void cg_scalar_kernel(int op, double* gamma, const double* other, float* coef, int nx, int npad)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= npad) return;
    float c = 0.0f;
    if (x < nx) {
        const float g = (float)gamma[x], o = (float)other[x];
        if (op == CG_ALPHA) c = g / o;
        else { c = (1.0f / g) * o; gamma[x] = other[x]; }
    }
    coef[x] = c;
}
// MODE 0: y += s*c[x]*v  (opAddScaled, arith.cu)      MODE 1: y = v + c[x]*y  (opScaleAndAdd)
template <int MODE>
__global__ void __launch_bounds__(256) cg_axpy_kernel(float4* __restrict__ y, const float4* __restrict__ v, const float* __restrict__ coef, float s,
                                                      int64_t cells, int nxb)
{
    const int64_t n4 = cells * 8 * nxb;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int xb = (int)(i / (cells * 8));
        const float4 c = *reinterpret_cast<const float4*>(coef + xb * 32 + (int)(i & 7) * 4);
        float4 a = y[i]; const float4 b = v[i];
        if (MODE == 0) { a.x = fmaf(b.x, s * c.x, a.x); a.y = fmaf(b.y, s * c.y, a.y); a.z = fmaf(b.z, s * c.z, a.z); a.w = fmaf(b.w, s * c.w, a.w); }
        else { a.x = fmaf(a.x, c.x, b.x); a.y = fmaf(a.y, c.y, b.y); a.z = fmaf(a.z, c.z, b.z); a.w = fmaf(a.w, c.w, b.w); }
        y[i] = a;
    }
}

inline void cgls_run(Engine& e, int niter) {
    e.need_angles();
    const size_t nv = e.vol_elems(), ns = e.sino_elems();
    const int npad = e.nxb * 32;
    if (!e.r.ok()) e.r.alloc(ns);
    if (!e.P[0].ok()) e.P[0].alloc(nv);
    if (!e.P[1].ok()) e.P[1].alloc(nv);
    DBuf sc; sc.alloc((size_t)npad * 7);                         // 3 double arrays + 1 float array
    double* gamma = reinterpret_cast<double*>(sc.p); double* ww = gamma + npad; double* gnew = ww + npad;
    float* coef = sc.p + (size_t)npad * 6;
    float* x = e.recon.p; float* r = e.r.p; float* w = e.g.p; float* p = e.P[0].p; float* z = e.P[1].p;
    const int64_t vcells = (int64_t)e.N * e.N, scells = (int64_t)e.Na * e.N;
    auto sumsq = [&](const float* a, int64_t cells, double* out) {
        TTV_CUDA(cudaMemsetAsync(out, 0, npad * sizeof(double), e.st));
        dim3 g((unsigned)std::min<int64_t>((cells + 31) / 32, 64), e.nxb);
        slice_sumsq_kernel<<<g, 256, 0, e.st>>>(reinterpret_cast<const float4*>(a), cells, out); ++e.launches;
    };
    const int eb = ew_blocks(nv / 4), sb = ew_blocks(ns / 4);
    // slice initialisation (cgls.cu:114-141): r = b - A x ; p = A' r ; gamma = <p,p>
    e.fp<FP_RESID>(x, r, e.b.p, nullptr);
    e.bp<BP_STORE>(r, p);
    sumsq(p, vcells, gamma);
    for (int it = 0; it < niter; ++it) {
        e.fp<FP_STORE>(p, w, nullptr, nullptr);                                                   // w = A p
        sumsq(w, scells, ww);
        cg_scalar_kernel<<<(npad + 127) / 128, 128, 0, e.st>>>(CG_ALPHA, gamma, ww, coef, e.nx, npad);
        cg_axpy_kernel<0><<<eb, 256, 0, e.st>>>(reinterpret_cast<float4*>(x), reinterpret_cast<const float4*>(p), coef, 1.0f, vcells, e.nxb);   // x += alpha p
        cg_axpy_kernel<0><<<sb, 256, 0, e.st>>>(reinterpret_cast<float4*>(r), reinterpret_cast<const float4*>(w), coef, -1.0f, scells, e.nxb);  // r -= alpha w
        e.bp<BP_STORE>(r, z);                                                                     // z = A' r
        sumsq(z, vcells, gnew);
        cg_scalar_kernel<<<(npad + 127) / 128, 128, 0, e.st>>>(CG_BETA, gamma, gnew, coef, e.nx, npad);
        cg_axpy_kernel<1><<<eb, 256, 0, e.st>>>(reinterpret_cast<float4*>(p), reinterpret_cast<const float4*>(z), coef, 1.0f, vcells, e.nxb);   // p = z + beta p
        e.launches += 5;
    }
    ew_kernel<EW_POS><<<eb, 256, 0, e.st>>>(reinterpret_cast<float4*>(x), 0.f, nv / 4); ++e.launches;      // recon.positivity(), tomoengine.cpp:228
    TTV_LAUNCH_CHECK();
    e.sync();
    sc.release();
}

// -----------------------------------------------------------------------------------------------
// FBP filter in the Fourier domain: astra::genFilter (A/src/Filters.cpp:40-420) with the defaults the
// reference uses (m_fD = 1, m_fParameter = -1, A/include/astra/Filters.h:79).  P = padded length,
// F = P/2+1 bins.  The ramp is built in the spatial domain and transformed (Filters.cpp:53-80; a
// double-precision DFT here instead of the float32 Ooura FFT).
// -----------------------------------------------------------------------------------------------
inline bool fbp_filter(const std::string& name_in, int P, std::vector<float>& filt) {
    std::string name = name_in;
    for (auto& ch : name) ch = (char)std::tolower((unsigned char)ch);
    const int F = P / 2 + 1;
    const float PI = 3.14159265358979323846264338328f;
    filt.assign(F, 0.0f);
    if (name == "none") return false;                     // no filtering at all (fbp.cu:157)
    std::vector<float> h(P, 0.0f);
    for (int i = 0; i < P; ++i)
        if (i & 1) { int j = i; if (2 * j > P) j = P - j; const float f = PI * (float)j; h[i] = -1.0f / (f * f); }
    h[0] = 0.25f;
    std::vector<float> w(F);
    {
        // real even sequence -> real spectrum; cos table by index (i*k mod P)
        std::vector<double> ct(P);
        for (int i = 0; i < P; ++i) ct[i] = std::cos(2.0 * M_PI * (double)i / (double)P);
        for (int k = 0; k < F; ++k) {
            double s = 0.0;
            for (int i = 0; i < P; ++i) if (h[i] != 0.0f) s += (double)h[i] * ct[(int)(((int64_t)i * k) % P)];
            filt[k] = 2.0f * (float)s;
            w[k] = PI * 2.0f * ((float)k / (float)P);
        }
    }
    const float d = 1.0f; const float N1 = (float)(F - 1);
    auto cos4 = [&](int k, float a0, float a1, float a2, float a3, float a4) {
        const float base = PI * (float)k / ((float)F - 1.0f);
        return a0 - a1 * cosf(2.0f * base) + a2 * cosf(4.0f * base) - a3 * cosf(6.0f * base) + a4 * cosf(8.0f * base);
    };
    for (int k = 1; k < F; ++k) {
        const float n = (float)k;
        float v = 1.0f;
        if (name == "ram-lak") v = 1.0f;
        else if (name == "shepp-logan") v = sinf(w[k] / 2.0f / d) / (w[k] / 2.0f / d);
        else if (name == "cosine") v = cosf(w[k] / 2.0f / d);
        else if (name == "hamming") v = 0.54f + 0.46f * cosf(w[k] / d);
        else if (name == "hann") v = (1.0f + cosf(w[k] / d)) / 2.0f;
        else if (name == "tukey") {
            const float a = 0.5f, half = (float)F / 2.0f, en = a * half, den = (1.0f - a) * half;
            v = (half - en <= n && n <= half + en) ? 1.0f : 0.5f * (1.0f + cosf(PI * (n - en) / den));
        } else if (name == "lanczos") {
            const float s = PI * (2.0f * n / N1 - 1.0f);
            v = fabsf(s) > 0.001f ? (float)(std::sin((double)s) / (double)s) : 1.0f;
        } else if (name == "triangular") v = 2.0f / N1 * (N1 / 2.0f - fabsf(n - N1 / 2.0f));
        else if (name == "gaussian") {
            const float q = ((float)F - 1.0f) / 2.0f, t = (n - q) / (0.4f * q);
            v = expf(-0.5f * t * t);
        } else if (name == "barlett-hann") v = 0.62f - 0.48f * fabsf(n / N1 - 0.5f) - 0.38f * cosf(2.0f * PI * n / N1);
        else if (name == "blackman") {
            const float al = 0.16f;
            v = (1.0f - al) / 2.0f - 0.5f * cosf(2.0f * PI * 0.5f * n / N1) + (al / 2.0f) * cosf(4.0f * PI * 0.5f * n / N1);
        } else if (name == "nuttall") v = cos4(k, 0.355768f, 0.487396f, 0.144232f, 0.012604f, 0.0f);
        else if (name == "blackman-harris") v = cos4(k, 0.35875f, 0.48829f, 0.14128f, 0.01168f, 0.0f);
        else if (name == "blackman-nuttall") v = cos4(k, 0.3635819f, 0.4891775f, 0.1365995f, 0.0106411f, 0.0f);
        else if (name == "flat-top") v = cos4(k, 1.0f, 1.93f, 1.29f, 0.388f, 0.032f);
        else if (name == "kaiser") {
            const float pa = PI * 3.0f, sq = 2.0f * n / N1 - 1.0f;
            v = (float)j0((double)(pa * (float)std::sqrt((double)(1.0f - sq * sq)))) / (float)j0((double)pa);
        } else if (name == "parzen") {
            const float q = n / N1;
            v = q <= 0.5f ? 1.0f - 6.0f * q * q * (1.0f - q) : 2.0f * (1.0f - q) * (1.0f - q) * (1.0f - q);
        } else throw CudaError("FBP: unknown filter '" + name_in + "'");
        filt[k] *= v;
    }
    for (int k = 0; k < F; ++k) if (w[k] > PI * d) filt[k] = 0.0f;
    return true;
}

// rows [rows][N] -> zero-padded [rows][P]
__global__ void __launch_bounds__(256) fbp_pad_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t rows, int N, int P)
{
    const int64_t n = rows * P;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / P; const int d = (int)(i - r * P);
        dst[i] = d < N ? src[r * N + d] : 0.0f;
    }
}
// spectrum *= filter (real), fft.cu:73-97
__global__ void __launch_bounds__(256) fbp_filter_kernel(float2* __restrict__ spec, const float* __restrict__ filt, int64_t rows, int F)
{
    const int64_t n = rows * F;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float c = filt[(int)(i % F)];
        float2 v = spec[i];
        const float a = v.x, b = v.y;
        v.x = a * c - b * 0.0f; v.y = a * 0.0f + c * b;
        spec[i] = v;
    }
}
// padded [rows][P] / P -> [rows][N]   (rescaleInverseFourier, fft.cu:97-113)
__global__ void __launch_bounds__(256) fbp_unpad_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t rows, int N, int P)
{
    const int64_t n = rows * N;
    const float fp = (float)P;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / N; const int d = (int)(i - r * N);
        dst[i] = src[r * P + d] / fp;
    }
}

inline void cufft_check(cufftResult r, const char* what) {
    if (r != CUFFT_SUCCESS) throw CudaError(std::string(what) + " failed (cuFFT code " + std::to_string((int)r) + ")");
}

inline void fbp_run(Engine& e, bool positivity) {
    e.need_angles();
    if (e.fbpFilter.empty()) throw CudaError("FBP called before initialize_FBP");
    const int N = e.N, Na = e.Na;
    int P = 1; while (P < 2 * N) P <<= 1;                         // calcNextPowerOfTwo(2*dets), fbp.cu:159
    const int F = P / 2 + 1;
    std::vector<float> hf;
    const bool filtered = fbp_filter(e.fbpFilter, P, hf);
    if (!e.r.ok()) e.r.alloc(e.sino_elems());
    const float* sino = e.b.p;
    if (filtered) {
        // x-blocks are filtered in chunks: unpack to plain rows, pad, R2C, filter, C2R, rescale, repack
        const int64_t M = (int64_t)Na * N;
        const int chunk = std::max(1, std::min(e.nxb, (int)((size_t)(256u << 20) / ((size_t)32 * Na * P * sizeof(float)) + 1)));
        const int64_t rows_max = (int64_t)chunk * 32 * Na;
        DBuf plain, pad, spec, dfilt;
        plain.alloc((size_t)rows_max * N); pad.alloc((size_t)rows_max * P); spec.alloc((size_t)rows_max * F * 2); dfilt.alloc(F);
        TTV_CUDA(cudaMemcpyAsync(dfilt.p, hf.data(), F * sizeof(float), cudaMemcpyHostToDevice, e.st));
        for (int xb0 = 0; xb0 < e.nxb; xb0 += chunk) {
            const int nb = std::min(chunk, e.nxb - xb0);
            const int64_t rows = (int64_t)nb * 32 * Na;       // padding slices are zero rows: filtered zeros stay zero
            dim3 g((unsigned)((M + 31) / 32), nb);
            unpack_kernel<<<g, 256, 0, e.st>>>(e.b.p + (size_t)xb0 * M * 32, plain.p, nb * 32, M, 0, nb * 32);
            fbp_pad_kernel<<<ew_blocks((size_t)rows * P / 4), 256, 0, e.st>>>(plain.p, pad.p, rows, N, P);
            cufftHandle fwd, inv;
            int n1[1] = {P};
            cufft_check(cufftPlanMany(&fwd, 1, n1, nullptr, 1, P, nullptr, 1, F, CUFFT_R2C, (int)rows), "cufftPlanMany(R2C)");
            cufft_check(cufftPlanMany(&inv, 1, n1, nullptr, 1, F, nullptr, 1, P, CUFFT_C2R, (int)rows), "cufftPlanMany(C2R)");
            cufft_check(cufftSetStream(fwd, e.st), "cufftSetStream"); cufft_check(cufftSetStream(inv, e.st), "cufftSetStream");
            cufft_check(cufftExecR2C(fwd, pad.p, reinterpret_cast<cufftComplex*>(spec.p)), "cufftExecR2C");
            fbp_filter_kernel<<<ew_blocks((size_t)rows * F / 4), 256, 0, e.st>>>(reinterpret_cast<float2*>(spec.p), dfilt.p, rows, F);
            cufft_check(cufftExecC2R(inv, reinterpret_cast<cufftComplex*>(spec.p), pad.p), "cufftExecC2R");
            fbp_unpad_kernel<<<ew_blocks((size_t)rows * N / 4), 256, 0, e.st>>>(pad.p, plain.p, rows, N, P);
            pack_kernel<<<g, 256, 0, e.st>>>(plain.p, e.r.p + (size_t)xb0 * M * 32, nb * 32, M, 0, nb * 32);
            e.launches += 5;
            TTV_LAUNCH_CHECK();
            e.sync();
            cufftDestroy(fwd); cufftDestroy(inv);
        }
        plain.release(); pad.release(); spec.release(); dfilt.release();
        sino = e.r.p;
    }
    // zeroVolumeData + BP with fOutputScale = (pi/2)/Nangles   (fbp.cu:106,179-181): vol = 0 + t*scale
    TTV_CUDA(cudaMemsetAsync(e.recon.p, 0, e.vol_elems() * sizeof(float), e.st));
    const float scale = (float)((M_PI / 2.0f) / (float)Na);
    e.bp<BP_ADD>(sino, e.recon.p, scale);
    if (positivity) { ew_kernel<EW_POS><<<ew_blocks(e.vol_elems() / 4), 256, 0, e.st>>>(reinterpret_cast<float4*>(e.recon.p), 0.f, e.vol_elems() / 4); ++e.launches; }
    TTV_LAUNCH_CHECK();
    e.sync();
}

}  // namespace ttv
