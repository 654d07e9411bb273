// dwconv.cu — depthwise 3x3 convolution (groups == channels) forward, data gradient and weight gradient.
//
// Reference site: nn.Conv2d(planes, planes, 3, stride, padding=1, groups=planes, bias=False),
// models/mobilenetv2.py:20 (and its autograd, main.py:104). 9 MACs per output element (~2 FLOP/B): this is
// HBM-bound work for the CUDA cores, judged on GB/s, not something to reshape into a GEMM.
//
// Layout: NHWC bf16 activations, 16-byte vectors of 8 channels per thread; fp32 master weights [C][3][3] read
// directly (72 values per thread, held in registers across all pixels the thread handles).
//   algorithmic bytes per output element: fprop 2 B read (x, neighbours come from L1/L2) + 2 B written;
//   dgrad the same; wgrad 4 B read.
// Thread layout (as bn_bwd_reduce): blockDim 256; thread t owns channel vector t % (C/8) and walks pixels.
// fprop also reduces per-channel sum / sum of squares of its fp32 results for the following BatchNorm
// (one partial row per block, same format as the conv epilogue's).
#include "b2c_elementwise.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    f[0] = bf16lo(u.x); f[1] = bf16hi(u.x); f[2] = bf16lo(u.y); f[3] = bf16hi(u.y);
    f[4] = bf16lo(u.z); f[5] = bf16hi(u.z); f[6] = bf16lo(u.w); f[7] = bf16hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 u;
    u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
    u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
    return u;
}

// weights of 8 consecutive channels: w[t][e] = master[(c0 + e) * 9 + t], zero for channels >= c_valid
__device__ __forceinline__ void load_w(const float* __restrict__ wt, int c0, int c_valid, float (&w)[9][8]) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
#pragma unroll
        for (int t = 0; t < 9; ++t) w[t][e] = (c0 + e < c_valid) ? __ldg(wt + (c0 + e) * 9 + t) : 0.f;
    }
}

// y[n, ho, wo, c] = sum_{r,s} x[n, ho*stride + r - 1, wo*stride + s - 1, c] * w[c, r, s]
__global__ void __launch_bounds__(256)
dw_fprop_kernel(const uint4* __restrict__ x, const float* __restrict__ wt, uint4* __restrict__ y, int N, int H, int W,
                int C, int c_valid, int stride, float* __restrict__ part_sum, float* __restrict__ part_sq) {
    pdl_trigger_ew();
    pdl_wait();
    extern __shared__ float red[];  // [rpi][2][C]
    const int cvec = C >> 3;
    const int rpi = 256 / cvec;
    const int cv = threadIdx.x % cvec, rl = threadIdx.x / cvec;
    const bool active = rl < rpi;
    const int Ho = H / stride, Wo = W / stride;
    const long long pixels = static_cast<long long>(N) * Ho * Wo;
    float w[9][8];
    load_w(wt, cv * 8, c_valid, w);
    float s1[8], s2[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { s1[e] = 0.f; s2[e] = 0.f; }
    for (long long p = static_cast<long long>(blockIdx.x) * rpi + rl; active && p < pixels;
         p += static_cast<long long>(gridDim.x) * rpi) {
        const int wo = static_cast<int>(p % Wo);
        const int ho = static_cast<int>((p / Wo) % Ho);
        const long long n = p / (static_cast<long long>(Wo) * Ho);
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int hi = ho * stride + r - 1;
            if (hi < 0 || hi >= H) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int wi = wo * stride + s - 1;
                if (wi < 0 || wi >= W) continue;
                float v[8];
                unpack8(__ldg(x + ((n * H + hi) * W + wi) * cvec + cv), v);
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] = fmaf(v[e], w[r * 3 + s][e], acc[e]);
            }
        }
        y[p * cvec + cv] = pack8(acc);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            s1[e] += acc[e];
            s2[e] = fmaf(acc[e], acc[e], s2[e]);
        }
    }
    if (part_sum == nullptr) return;
    if (active) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            red[(rl * 2 + 0) * C + cv * 8 + e] = s1[e];
            red[(rl * 2 + 1) * C + cv * 8 + e] = s2[e];
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * C; i += 256) {
        float s = 0.f;
        for (int k = 0; k < rpi; ++k) s += red[k * 2 * C + i];
        if (i < C) part_sum[static_cast<size_t>(blockIdx.x) * C + i] = s;
        else part_sq[static_cast<size_t>(blockIdx.x) * C + (i - C)] = s;
    }
}

// dx[n, h, w, c] = sum_{r,s : (h+1-r) % stride == 0, (w+1-s) % stride == 0} dy[n, (h+1-r)/stride, (w+1-s)/stride, c] * w[c, r, s]
__global__ void __launch_bounds__(256)
dw_dgrad_kernel(const uint4* __restrict__ dy, const float* __restrict__ wt, uint4* __restrict__ dx, int N, int H, int W,
                int C, int c_valid, int stride) {
    pdl_trigger_ew();
    pdl_wait();
    const int cvec = C >> 3;
    const int rpi = 256 / cvec;
    const int cv = threadIdx.x % cvec, rl = threadIdx.x / cvec;
    if (rl >= rpi) return;
    const int Ho = H / stride, Wo = W / stride;
    const long long pixels = static_cast<long long>(N) * H * W;
    float w[9][8];
    load_w(wt, cv * 8, c_valid, w);
    for (long long p = static_cast<long long>(blockIdx.x) * rpi + rl; p < pixels;
         p += static_cast<long long>(gridDim.x) * rpi) {
        const int wi = static_cast<int>(p % W);
        const int hi = static_cast<int>((p / W) % H);
        const long long n = p / (static_cast<long long>(W) * H);
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int th = hi + 1 - r;
            if (th < 0 || (th % stride) != 0) continue;
            const int ho = th / stride;
            if (ho >= Ho) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int tw = wi + 1 - s;
                if (tw < 0 || (tw % stride) != 0) continue;
                const int wo = tw / stride;
                if (wo >= Wo) continue;
                float v[8];
                unpack8(__ldg(dy + ((n * Ho + ho) * Wo + wo) * cvec + cv), v);
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] = fmaf(v[e], w[r * 3 + s][e], acc[e]);
            }
        }
        dx[p * cvec + cv] = pack8(acc);
    }
}

// dw[c, r, s] += sum_{n, ho, wo} x[n, ho*stride + r - 1, wo*stride + s - 1, c] * dy[n, ho, wo, c]
__global__ void __launch_bounds__(256)
dw_wgrad_kernel(const uint4* __restrict__ x, const uint4* __restrict__ dy, float* __restrict__ dw, int N, int H, int W,
                int C, int c_valid, int stride) {
    pdl_trigger_ew();
    pdl_wait();
    extern __shared__ float red[];  // [rpi][9][C]
    const int cvec = C >> 3;
    const int rpi = 256 / cvec;
    const int cv = threadIdx.x % cvec, rl = threadIdx.x / cvec;
    const bool active = rl < rpi;
    const int Ho = H / stride, Wo = W / stride;
    const long long pixels = static_cast<long long>(N) * Ho * Wo;
    float acc[9][8];
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[t][e] = 0.f;
    for (long long p = static_cast<long long>(blockIdx.x) * rpi + rl; active && p < pixels;
         p += static_cast<long long>(gridDim.x) * rpi) {
        const int wo = static_cast<int>(p % Wo);
        const int ho = static_cast<int>((p / Wo) % Ho);
        const long long n = p / (static_cast<long long>(Wo) * Ho);
        float g[8];
        unpack8(__ldg(dy + p * cvec + cv), g);
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int hi = ho * stride + r - 1;
            if (hi < 0 || hi >= H) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int wi = wo * stride + s - 1;
                if (wi < 0 || wi >= W) continue;
                float v[8];
                unpack8(__ldg(x + ((n * H + hi) * W + wi) * cvec + cv), v);
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[r * 3 + s][e] = fmaf(v[e], g[e], acc[r * 3 + s][e]);
            }
        }
    }
    if (active) {
#pragma unroll
        for (int t = 0; t < 9; ++t)
#pragma unroll
            for (int e = 0; e < 8; ++e) red[(rl * 9 + t) * C + cv * 8 + e] = acc[t][e];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 9 * C; i += 256) {
        const int t = i / C, c = i - t * C;
        if (c >= c_valid) continue;
        float s = 0.f;
        for (int k = 0; k < rpi; ++k) s += red[(k * 9 + t) * C + c];
        atomicAdd(dw + c * 9 + t, s);
    }
}

}  // namespace

int dw_fprop_partial_rows(long long out_pixels, int C) {
    const int rpi = 256 / (C >> 3);
    long long blocks = (out_pixels + rpi - 1) / rpi;
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

cudaError_t launch_dw_fprop(const void* x, const float* w, void* y, int N, int H, int W, int C, int c_valid, int stride,
                            float* part_sum, float* part_sq, int* stat_rows, cudaStream_t s) {
    const int cvec = C >> 3;
    if (cvec < 1 || cvec > 256) return cudaErrorInvalidValue;
    const int rpi = 256 / cvec;
    const int blocks = dw_fprop_partial_rows(static_cast<long long>(N) * (H / stride) * (W / stride), C);
    if (stat_rows != nullptr) *stat_rows = blocks;
    const size_t smem = part_sum != nullptr ? static_cast<size_t>(rpi) * 2 * C * sizeof(float) : 0;
    B2C_LAUNCH(dw_fprop_kernel, blocks, 256, smem, s, static_cast<const uint4*>(x), w, static_cast<uint4*>(y), N, H, W, C,
                                              c_valid, stride, part_sum, part_sq);
    return cudaGetLastError();
}

cudaError_t launch_dw_dgrad(const void* dy, const float* w, void* dx, int N, int H, int W, int C, int c_valid,
                            int stride, cudaStream_t s) {
    const int cvec = C >> 3;
    if (cvec < 1 || cvec > 256) return cudaErrorInvalidValue;
    const int blocks = dw_fprop_partial_rows(static_cast<long long>(N) * H * W, C);
    B2C_LAUNCH(dw_dgrad_kernel, blocks * 2 > 148 * 8 ? 148 * 8 : blocks * 2, 256, 0, s, 
        static_cast<const uint4*>(dy), w, static_cast<uint4*>(dx), N, H, W, C, c_valid, stride);
    return cudaGetLastError();
}

cudaError_t launch_dw_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int c_valid,
                            int stride, cudaStream_t s) {
    const int cvec = C >> 3;
    if (cvec < 1 || cvec > 256) return cudaErrorInvalidValue;
    const int rpi = 256 / cvec;
    const size_t smem = static_cast<size_t>(rpi) * 9 * C * sizeof(float);      // 72 KB at most
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(dw_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int blocks = dw_fprop_partial_rows(static_cast<long long>(N) * (H / stride) * (W / stride), C);
    B2C_LAUNCH(dw_wgrad_kernel, blocks > 296 ? 296 : blocks, 256, smem, s, static_cast<const uint4*>(x),
                                                                   static_cast<const uint4*>(dy), dw, N, H, W, C,
                                                                   c_valid, stride);
    return cudaGetLastError();
}

}  // namespace b2c
