// pool3.cu — nn.MaxPool2d(3, stride, padding=1) forward / backward (models/googlenet.py:40,67) and the channel-slice
// copy that materialises a concatenation where one is needed (the head of GoogLeNet, googlenet.py:96-98).
// NHWC bf16, one thread per (pixel, 8-channel vector); HBM-bound (the 3x3 neighbourhoods overlap in L1 / L2).
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

// y = max over the 3x3 window (padding excluded, as -inf); arg = window position r*3+s of the FIRST maximum in scan
// order (torch's rule: a later value replaces the maximum only if strictly greater), kept for the backward pass.
__global__ void __launch_bounds__(256)
maxpool3_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, uint2* __restrict__ arg, long long nvec,
                    int H, int W, int Ho, int Wo, int cvec, int stride) {
    pdl_trigger_ew();
    pdl_wait();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int cv = static_cast<int>(i % cvec);
        long long pix = i / cvec;
        const int ox = static_cast<int>(pix % Wo);
        pix /= Wo;
        const int oy = static_cast<int>(pix % Ho);
        const long long n = pix / Ho;
        float best[8];
        int pos[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { best[k] = -INFINITY; pos[k] = 0; }
        bool first = true;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int iy = oy * stride - 1 + r;
            if (iy < 0 || iy >= H) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int ix = ox * stride - 1 + s;
                if (ix < 0 || ix >= W) continue;
                float f[8];
                unpack8(x[((n * H + iy) * W + ix) * cvec + cv], f);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    if (first || f[k] > best[k]) { best[k] = f[k]; pos[k] = r * 3 + s; }
                }
                first = false;
            }
        }
        y[i] = pack8(best);
        uint2 a;
        a.x = pos[0] | (pos[1] << 8) | (pos[2] << 16) | (pos[3] << 24);
        a.y = pos[4] | (pos[5] << 8) | (pos[6] << 16) | (pos[7] << 24);
        arg[i] = a;
    }
}

// dx = (add) + sum over the windows that contain this input pixel of dy[window] where the window's argmax is this pixel.
__global__ void __launch_bounds__(256)
maxpool3_bwd_kernel(const uint4* __restrict__ dy, const uint2* __restrict__ arg, const uint4* __restrict__ add,
                    uint4* __restrict__ dx, long long nvec, int H, int W, int Ho, int Wo, int cvec, int stride) {
    pdl_trigger_ew();
    pdl_wait();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int cv = static_cast<int>(i % cvec);
        long long pix = i / cvec;
        const int ix = static_cast<int>(pix % W);
        pix /= W;
        const int iy = static_cast<int>(pix % H);
        const long long n = pix / H;
        float g[8];
        if (add != nullptr) unpack8(add[i], g);
        else {
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = 0.f;
        }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int ty = iy + 1 - r;                       // = oy * stride
            if (ty < 0 || (ty % stride) != 0) continue;
            const int oy = ty / stride;
            if (oy >= Ho) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int tx = ix + 1 - s;
                if (tx < 0 || (tx % stride) != 0) continue;
                const int ox = tx / stride;
                if (ox >= Wo) continue;
                const long long o = ((n * Ho + oy) * Wo + ox) * cvec + cv;
                const uint2 a = arg[o];
                float f[8];
                unpack8(dy[o], f);
                const unsigned int me = static_cast<unsigned int>(r * 3 + s);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (((a.x >> (8 * k)) & 0xffu) == me) g[k] += f[k];
                    if (((a.y >> (8 * k)) & 0xffu) == me) g[4 + k] += f[4 + k];
                }
            }
        }
        dx[i] = pack8(g);
    }
}

// dst[row][dst_off + c] = src[row][src_off + c], c < channels; all offsets / pitches / channels multiples of 8.
__global__ void __launch_bounds__(256)
copy_channels_kernel(const uint4* __restrict__ src, int src_ld, int src_off, uint4* __restrict__ dst, int dst_ld,
                     int dst_off, long long rows, int cvec) {
    pdl_trigger_ew();
    pdl_wait();
    const long long nvec = rows * cvec;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long row = i / cvec;
        const int cv = static_cast<int>(i - row * cvec);
        dst[row * dst_ld + dst_off + cv] = src[row * src_ld + src_off + cv];
    }
}

inline int ew_grid(long long nvec) {
    long long g = (nvec + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    if (g < 1) g = 1;
    return static_cast<int>(g);
}

}  // namespace

cudaError_t launch_maxpool3_fwd(const void* x, void* y, void* arg, int n, int h, int w, int C, int stride,
                                cudaStream_t s) {
    if ((C & 7) || (stride != 1 && stride != 2) || (h % stride) || (w % stride)) return cudaErrorInvalidValue;
    const int ho = h / stride, wo = w / stride;
    const long long nvec = static_cast<long long>(n) * ho * wo * (C / 8);
    if (nvec == 0) return cudaSuccess;
    B2C_LAUNCH(maxpool3_fwd_kernel, ew_grid(nvec), 256, 0, s, static_cast<const uint4*>(x), static_cast<uint4*>(y),
               static_cast<uint2*>(arg), nvec, h, w, ho, wo, C / 8, stride);
    return cudaGetLastError();
}

cudaError_t launch_maxpool3_bwd(const void* dy, const void* arg, const void* add, void* dx, int n, int h, int w, int C,
                                int stride, cudaStream_t s) {
    if ((C & 7) || (stride != 1 && stride != 2) || (h % stride) || (w % stride)) return cudaErrorInvalidValue;
    const long long nvec = static_cast<long long>(n) * h * w * (C / 8);
    if (nvec == 0) return cudaSuccess;
    B2C_LAUNCH(maxpool3_bwd_kernel, ew_grid(nvec), 256, 0, s, static_cast<const uint4*>(dy),
               static_cast<const uint2*>(arg), static_cast<const uint4*>(add), static_cast<uint4*>(dx), nvec, h, w,
               h / stride, w / stride, C / 8, stride);
    return cudaGetLastError();
}

cudaError_t launch_copy_channels(const void* src, int src_ld, int src_off, void* dst, int dst_ld, int dst_off,
                                 long long rows, int channels, cudaStream_t s) {
    if ((src_ld | src_off | dst_ld | dst_off | channels) & 7) return cudaErrorInvalidValue;
    if (rows == 0 || channels == 0) return cudaSuccess;
    B2C_LAUNCH(copy_channels_kernel, ew_grid(rows * (channels / 8)), 256, 0, s, static_cast<const uint4*>(src),
               src_ld / 8, src_off / 8, static_cast<uint4*>(dst), dst_ld / 8, dst_off / 8, rows, channels / 8);
    return cudaGetLastError();
}

}  // namespace b2c
