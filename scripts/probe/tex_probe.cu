// Measures how the texture unit quantises an unnormalised linear-filter coordinate (the model the oracle and the
// projector kernels emulate).  Texture row: texel i holds the value i, so a fetch returns (i + alpha).
#include <cuda_runtime.h>
#include <cstdio>
#include <cmath>
#include <vector>
#include <cstdint>
#include <cstring>
__global__ void fetch(cudaTextureObject_t tex, const float* x, float* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex2D<float>(tex, x[i], 0.5f);
}
int main() {
    const int W = 2048;
    float* d_img; size_t pitch;
    cudaMallocPitch(&d_img, &pitch, W * sizeof(float), 2);
    std::vector<float> row(W); for (int i = 0; i < W; ++i) row[i] = (float)i;
    cudaMemcpy2D(d_img, pitch, row.data(), W * sizeof(float), W * sizeof(float), 1, cudaMemcpyHostToDevice);
    cudaMemcpy2D((char*)d_img + pitch, pitch, row.data(), W * sizeof(float), W * sizeof(float), 1, cudaMemcpyHostToDevice);
    cudaResourceDesc rd; memset(&rd, 0, sizeof rd);
    rd.resType = cudaResourceTypePitch2D; rd.res.pitch2D.devPtr = d_img; rd.res.pitch2D.width = W; rd.res.pitch2D.height = 2;
    rd.res.pitch2D.pitchInBytes = pitch; rd.res.pitch2D.desc = cudaCreateChannelDesc<float>();
    cudaTextureDesc td; memset(&td, 0, sizeof td);
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeBorder; td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
    cudaTextureObject_t tex; cudaCreateTextureObject(&tex, &rd, &td, nullptr);
    // coordinates: random floats in several magnitude bands, plus exact ties and their neighbours
    std::vector<float> xs;
    uint64_t s = 88172645463325252ull;
    auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (double)(s >> 11) / 9007199254740992.0; };
    const float bands[] = {2.f, 8.f, 40.f, 100.f, 300.f, 700.f, 1500.f};
    for (float b : bands) for (int i = 0; i < 400000; ++i) xs.push_back((float)(1.0 + rnd() * (b - 1.0)));
    for (float b : bands) for (int i = 0; i < 100000; ++i) {
        float t = (float)((floor(rnd() * (b - 1.0) * 256.0) + 0.5) / 256.0 + 1.0);   // exact tie of 256x
        xs.push_back(t); xs.push_back(nextafterf(t, 0.f)); xs.push_back(nextafterf(t, 1e9f));
    }
    const int n = (int)xs.size();
    float *dx, *dout; cudaMalloc(&dx, n * 4); cudaMalloc(&dout, n * 4);
    cudaMemcpy(dx, xs.data(), n * 4, cudaMemcpyHostToDevice);
    fetch<<<(n + 255) / 256, 256>>>(tex, dx, dout, n);
    std::vector<float> out(n); cudaMemcpy(out.data(), dout, n * 4, cudaMemcpyDeviceToHost);
    printf("err %s n %d\n", cudaGetErrorString(cudaGetLastError()), n);
    long bad[6] = {0}; int shown = 0;
    for (int i = 0; i < n; ++i) {
        const float x = xs[i];
        const long hw = lround((double)out[i] * 256.0);                        // (i + alpha) * 256
        const long m0 = (long)floor((double)x * 256.0 + 0.5) - 128;            // round half up, exact
        const long m1 = (long)nearbyint((double)x * 256.0) - 128;              // round half even, exact
        const float xb = x - 0.5f;
        const long m2 = (long)floor((double)xb * 256.0 + 0.5);                 // float subtract first, half up
        const long m3 = (long)nearbyint((double)xb * 256.0);                   // float subtract first, half even
        const long m4 = (long)floor((double)x * 256.0) - 128;                  // truncate
        const long m5 = (long)ceil((double)x * 256.0 - 0.5) - 128;             // round half down
        const long m[6] = {m0, m1, m2, m3, m4, m5};
        for (int k = 0; k < 6; ++k) if (m[k] != hw) ++bad[k];
        if (m0 != hw && shown < 25) { printf("x=%.9g (0x%08x) 256x=%.6f hw=%ld m0=%ld\n", x, *(const unsigned*)&x, (double)x * 256.0, hw + 128, m0 + 128); ++shown; }
    }
    printf("mismatches: half-up %ld  half-even %ld  fsub+half-up %ld  fsub+half-even %ld  trunc %ld  half-down %ld\n", bad[0], bad[1], bad[2], bad[3], bad[4], bad[5]);
    return 0;
}
