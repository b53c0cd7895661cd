// tomo_tv_b200 -- common definitions for the sm_100a kernels.
//
// Device data layout ("x-interleaved"): the tilt axis x (slices) is the fastest index in HBM.
//   volume  element (x, j, k)  at  ((xb*N + j)*N + k)*32 + xi        xb = x/32, xi = x%32
//   sinogram element (x, a, d) at  ((xb*Na + a)*Nd + d)*32 + xi
// One "cell" = 32 consecutive slices of one (j,k) voxel / one (a,d) ray = 128 B = one cache
// line = one shared-memory bank row.  The interpolation indices and weights of both projectors
// depend only on (angle, detector, row) and never on x, so a warp maps 8 lanes x float4 onto
// one cell: every shared/global access is a conflict-free, fully coalesced 128-byte line and a
// single index computation serves 4 samples per lane.  Slices beyond Nx (padding of the last
// block) are kept identically zero by every kernel.
//
// The host-facing layout of the reference ([x][j][k] volumes, Matrix3D.cpp:20-23; [x][a*Nd+d]
// sinograms, reconstructor.py:54-56) only exists at the C-ABI boundary (ttv_layout.cuh).
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <cstdint>
#include <cstdio>
#include <cmath>
#include <stdexcept>
#include <string>

#define TTV_XI 32  // slices per cell

namespace ttv {

struct CudaError : std::runtime_error { using std::runtime_error::runtime_error; };

inline void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
    if (e != cudaSuccess) {
        char buf[512];
        snprintf(buf, sizeof buf, "%s failed: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
        throw CudaError(buf);
    }
}
#define TTV_CUDA(x) ::ttv::cuda_check((x), #x, __FILE__, __LINE__)
#define TTV_LAUNCH_CHECK() ::ttv::cuda_check(cudaGetLastError(), "kernel launch", __FILE__, __LINE__)

// Per-angle geometry constants, computed on the host exactly as ASTRA does
// (A/src/GeometryUtil2D.cpp:35-74,107-130 and A/cuda/2d/par_bp.cu:173-179).
struct AngleGeom {
    float theta, detsize, offset;   // forward projector
    float sc, ss, off, scale;       // back projector
    float rayX, rayY;
    int vertical;                   // |rayX| <= |rayY|  (A/cuda/2d/par_fp.cu:271)
};

__host__ __device__ inline int64_t vol_cell(int xb, int j, int k, int N) { return ((int64_t)xb * N + j) * N + k; }

// ---------------------------------------------------------------------------------------------
// PTX helpers: mbarrier + TMA (cp.async.bulk.tensor), sm_100a.
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 4-D tiled TMA load: global (tensor map) -> shared, completion on an mbarrier.
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}
// Four fp32 accumulators held as two packed pairs for sm_100a's FFMA2 (fma.rn.f32x2: two IEEE-754 fused
// multiply-adds per instruction, bit-identical to two fmaf; the scalar weight is broadcast by the instruction
// itself, `FFMA2 R, Rw.F32, Rv.F32x2, Racc.F32x2`).  The projectors issue 4 FFMA2 instead of 8 FFMA per
// bilinear sample group, which frees issue slots next to the LDS.128 stream that bounds them.
struct Acc4 {
    unsigned long long lo, hi;          // {x, y}, {z, w}
    __device__ __forceinline__ void zero() { lo = 0ull; hi = 0ull; }
    // this += w * v   (per component: fmaf(w, v, acc))
    __device__ __forceinline__ void fma(float w, const float4& v) {
        asm("{\n\t.reg .b64 ww, vv;\n\t"
            "mov.b64 ww, {%2, %2};\n\t"
            "mov.b64 vv, {%3, %4};\n\t"
            "fma.rn.f32x2 %0, ww, vv, %0;\n\t"
            "mov.b64 vv, {%5, %6};\n\t"
            "fma.rn.f32x2 %1, ww, vv, %1;\n\t}"
            : "+l"(lo), "+l"(hi) : "f"(w), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w));
    }
    __device__ __forceinline__ float4 get() const {
        float4 r;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(lo));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(r.z), "=f"(r.w) : "l"(hi));
        return r;
    }
};

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ float lds32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ void sts32(uint32_t addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

// Texture-unit interpolation model (cudaFilterModeLinear, A/cuda/2d/util.cu:161-179,198-223):
// the coordinate is converted to fixed point with 8 fractional bits rounding HALF UP (floor(256 x + 0.5);
// measured on B200 with scripts/probe/tex_probe.cu: 0 mismatches over 4.9 M coordinates including exact ties,
// round-half-even fails on 7 %), 0.5 is subtracted, the integer part selects the texel pair and the
// 8-bit fraction is the weight.
// t = fma.rd(x, 256, C) with C = 1.5*2^22 + 0.5 - 128 - 256*w0 (floats in [2^22,2^23) have ulp 0.5, the
// round-down FMA floors to it) leaves  u = floor(2*(256 x + 0.5 - 128 - 256 w0))  in the low mantissa bits:
// u>>9 = texel index relative to the staged window start w0, u & 0x1FE = 2*256*alpha (weights in 1/512).
#define TTV_MAGIC 6291456.0f        // 1.5 * 2^22
#define TTV_MAGIC_BITS 0x4AC00000
#define TTV_MAGIC_C(w0) ((TTV_MAGIC - 127.5f) - 256.0f * (float)(w0))

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
#endif

}  // namespace ttv
