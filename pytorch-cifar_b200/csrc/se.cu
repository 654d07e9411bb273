// se.cu — squeeze-and-excitation (models/senet.py:27-41,64-76; models/regnet.py:12-26):
//     w = sigmoid(fc2(relu(fc1(avg_pool(y)))));  out = y * w (+ shortcut)
// over NHWC bf16 activations. The two full-tensor passes (pool, scale) and their backward counterparts are HBM-bound;
// the per-image MLP is a few thousand multiply-adds per image and runs one block per image in fp32.
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

// out[n][c] = mult * sum_hw a[n][hw][c] * (b ? b[n][hw][c] : 1).   Squeeze: a = y, mult = 1/HW (F.avg_pool2d(out,
// out.size(2)), senet.py:68); backward of the excitation: a = dout, b = y, mult = 1 (d out / d w summed over the map).
// One block per (image, 64-channel slab): 8 channel vectors x 32 row lanes, rows strided over the lanes.
__global__ void __launch_bounds__(256)
se_pool_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int hw, int cvec, float mult,
               float* __restrict__ out) {
    pdl_trigger_ew();
    pdl_wait();
    __shared__ float red[32][65];
    const int n = blockIdx.x, slab = blockIdx.y;
    const int v = threadIdx.x & 7, r0 = threadIdx.x >> 3;
    const int cv = slab * 8 + v;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    if (cv < cvec) {
        const size_t base = static_cast<size_t>(n) * hw * cvec + cv;
        for (int r = r0; r < hw; r += 32) {
            float fa[8];
            unpack8(a[base + static_cast<size_t>(r) * cvec], fa);
            if (b != nullptr) {
                float fb[8];
                unpack8(b[base + static_cast<size_t>(r) * cvec], fb);
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] = fmaf(fa[k], fb[k], acc[k]);
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] += fa[k];
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) red[r0][v * 8 + k] = acc[k];
    __syncthreads();
    if (threadIdx.x < 64) {
        float t = 0.f;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) t += red[i][threadIdx.x];
        const int c = slab * 64 + threadIdx.x;
        if (c < cvec * 8) out[static_cast<size_t>(n) * cvec * 8 + c] = t * mult;
    }
}

// Per image: h = relu(W1 p + b1) (R values), s = sigmoid(W2 h + b2) (C values). W1 [R][C], W2 [C][R] — the 1x1
// nn.Conv2d weights of senet.py:37-38 — bias vectors b1 [R], b2 [C]; p, s have pitch ld (the padded channel layout), h
// pitch R. One block per image.
__global__ void __launch_bounds__(256)
se_fc_kernel(const float* __restrict__ p, int C, int R, int ld, const float* __restrict__ W1,
             const float* __restrict__ b1, const float* __restrict__ W2, const float* __restrict__ b2,
             float* __restrict__ h_out, float* __restrict__ s_out) {
    pdl_trigger_tiny();
    pdl_wait();
    extern __shared__ float sm[];          // p [C] | h [R]
    float* sp = sm;
    float* sh = sm + C;
    const int n = blockIdx.x;
    for (int c = threadIdx.x; c < C; c += blockDim.x) sp[c] = p[static_cast<size_t>(n) * ld + c];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int j = warp; j < R; j += nw) {
        float acc = 0.f;
        for (int c = lane; c < C; c += 32) acc = fmaf(__ldg(W1 + static_cast<size_t>(j) * C + c), sp[c], acc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) {
            const float hv = fmaxf(acc + b1[j], 0.f);
            sh[j] = hv;
            h_out[static_cast<size_t>(n) * R + j] = hv;
        }
    }
    __syncthreads();
    // one warp per channel, lanes over the hidden units: row c of W2 is read as one coalesced run (a thread per
    // channel would touch 32 different lines per load: measured 52 us per launch for 512 images x 384 channels x 96)
    for (int c = warp; c < ld; c += nw) {
        float acc = 0.f;
        if (c < C)
            for (int j = lane; j < R; j += 32) acc = fmaf(__ldg(W2 + static_cast<size_t>(c) * R + j), sh[j], acc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        // padded channels: scale 0 (their activations are zero anyway)
        if (lane == 0) s_out[static_cast<size_t>(n) * ld + c] = c < C ? 1.f / (1.f + expf(-(acc + b2[c]))) : 0.f;
    }
}

// out = y * s[n][c] (+ add): the excitation `out = out * w` and the block's `out += shortcut`, senet.py:72-74.
__global__ void __launch_bounds__(256)
se_scale_kernel(const uint4* __restrict__ y, const float* __restrict__ s, const uint4* __restrict__ add,
                uint4* __restrict__ out, long long nvec, int hw, int cvec) {
    pdl_trigger_ew();
    pdl_wait();
    const long long per_img = static_cast<long long>(hw) * cvec;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long n = i / per_img;
        const int cv = static_cast<int>(i % cvec);
        const float4 s0 = *reinterpret_cast<const float4*>(s + (n * cvec + cv) * 8);
        const float4 s1 = *reinterpret_cast<const float4*>(s + (n * cvec + cv) * 8 + 4);
        float f[8];
        unpack8(y[i], f);
        f[0] *= s0.x; f[1] *= s0.y; f[2] *= s0.z; f[3] *= s0.w;
        f[4] *= s1.x; f[5] *= s1.y; f[6] *= s1.z; f[7] *= s1.w;
        if (add != nullptr) {
            float g[8];
            unpack8(add[i], g);
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] += g[k];
        }
        out[i] = pack8(f);
    }
}

// dy = dout * s[n][c] + dp[n][c] * inv_hw: gradient w.r.t. the SE block's input (through the scaling and through the
// average pool that fed the MLP).
__global__ void __launch_bounds__(256)
se_bwd_apply_kernel(const uint4* __restrict__ dout, const float* __restrict__ s, const float* __restrict__ dp,
                    uint4* __restrict__ dy, long long nvec, int hw, int cvec, float inv_hw) {
    pdl_trigger_ew();
    pdl_wait();
    const long long per_img = static_cast<long long>(hw) * cvec;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long n = i / per_img;
        const int cv = static_cast<int>(i % cvec);
        const float* sp = s + (n * cvec + cv) * 8;
        const float* dpp = dp + (n * cvec + cv) * 8;
        float f[8];
        unpack8(dout[i], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sp[k], dpp[k] * inv_hw);
        dy[i] = pack8(f);
    }
}

// Backward of the per-image MLP. Given ds[n][c] = d loss / d s: dz2 = ds * s (1 - s); dh = W2^T dz2 masked by h > 0
// (= dz1); dp = W1^T dz1. dz2 overwrites ds, dz1 is written to dz1_out: the parameter-gradient kernel reduces them
// over the batch.
__global__ void __launch_bounds__(256)
se_fc_bwd_kernel(float* __restrict__ ds, const float* __restrict__ s, const float* __restrict__ h, int C, int R,
                 int ld, const float* __restrict__ W1, const float* __restrict__ W2, float* __restrict__ dz1_out,
                 float* __restrict__ dp) {
    pdl_trigger_tiny();
    pdl_wait();
    extern __shared__ float sm[];          // dz2 [C] | dz1 [R]
    float* z2 = sm;
    float* z1 = sm + C;
    const int n = blockIdx.x;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float sv = s[static_cast<size_t>(n) * ld + c];
        const float v = ds[static_cast<size_t>(n) * ld + c] * sv * (1.f - sv);
        z2[c] = v;
        ds[static_cast<size_t>(n) * ld + c] = v;
    }
    __syncthreads();
    // dh[j] = sum_c W2[c][j] dz2[c]: threads run along j (coalesced rows of W2); when R is small, the channel range is
    // split over 256 / Rp thread groups whose partial sums meet in shared memory
    __shared__ float part[256];
    const int Rp = (R + 31) & ~31;
    if (Rp <= 256) {
        const int G = 256 / Rp;
        const int jj = threadIdx.x % Rp, g = threadIdx.x / Rp;
        float acc = 0.f;
        if (g < G && jj < R)
            for (int c = g; c < C; c += G) acc = fmaf(__ldg(W2 + static_cast<size_t>(c) * R + jj), z2[c], acc);
        part[threadIdx.x] = acc;
        __syncthreads();
        if (threadIdx.x < R) {
            float t = 0.f;
            for (int q = 0; q < G; ++q) t += part[q * Rp + threadIdx.x];
            const float v = h[static_cast<size_t>(n) * R + threadIdx.x] > 0.f ? t : 0.f;
            z1[threadIdx.x] = v;
            dz1_out[static_cast<size_t>(n) * R + threadIdx.x] = v;
        }
    } else {
        for (int j = threadIdx.x; j < R; j += blockDim.x) {
            float acc = 0.f;
            for (int c = 0; c < C; ++c) acc = fmaf(__ldg(W2 + static_cast<size_t>(c) * R + j), z2[c], acc);
            const float v = h[static_cast<size_t>(n) * R + j] > 0.f ? acc : 0.f;
            z1[j] = v;
            dz1_out[static_cast<size_t>(n) * R + j] = v;
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < ld; c += blockDim.x) {
        float acc = 0.f;
        if (c < C)
            for (int j = 0; j < R; ++j) acc = fmaf(__ldg(W1 + static_cast<size_t>(j) * C + c), z1[j], acc);
        dp[static_cast<size_t>(n) * ld + c] = acc;
    }
}

// Parameter gradients of the MLP, reduced over the batch (accumulating, like every weight-gradient kernel here):
//   dW2[c][j] += sum_n dz2[n][c] h[n][j]   db2[c] += sum_n dz2[n][c]
//   dW1[j][c] += sum_n dz1[n][j] p[n][c]   db1[j] += sum_n dz1[n][j]
// grid = (ceil(C / 32), R): block (cx, j) owns 32 channels of column j of W2 / row j of W1; 32 batch lanes per channel.
__global__ void __launch_bounds__(1024)
se_param_grad_kernel(const float* __restrict__ dz2, const float* __restrict__ dz1, const float* __restrict__ p,
                     const float* __restrict__ h, int N, int C, int R, int ld, float* __restrict__ dW1,
                     float* __restrict__ db1, float* __restrict__ dW2, float* __restrict__ db2) {
    pdl_trigger_tiny();
    pdl_wait();
    __shared__ float red[4][32][33];
    const int j = blockIdx.y;
    const int cl = threadIdx.x & 31, nl = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cl;
    const bool valid = c < C;
    float a2 = 0.f, ab2 = 0.f, a1 = 0.f, ab1 = 0.f;
    for (int n = nl; n < N; n += 32) {
        const float z2 = valid ? dz2[static_cast<size_t>(n) * ld + c] : 0.f;
        const float pv = valid ? p[static_cast<size_t>(n) * ld + c] : 0.f;
        const float hv = h[static_cast<size_t>(n) * R + j];
        const float z1 = dz1[static_cast<size_t>(n) * R + j];
        a2 = fmaf(z2, hv, a2);
        ab2 += z2;
        a1 = fmaf(z1, pv, a1);
        ab1 += z1;
    }
    red[0][nl][cl] = a2; red[1][nl][cl] = ab2; red[2][nl][cl] = a1; red[3][nl][cl] = ab1;
    __syncthreads();
    if (nl < 4) {
        float t = 0.f;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) t += red[nl][i][cl];
        if (nl == 0 && valid) dW2[static_cast<size_t>(c) * R + j] += t;
        if (nl == 1 && valid && j == 0) db2[c] += t;
        if (nl == 2 && valid) dW1[static_cast<size_t>(j) * C + c] += t;
        if (nl == 3 && cl == 0 && blockIdx.x == 0) db1[j] += t;       // every channel lane holds the same sum
    }
}

inline int ew_grid(long long nvec) {
    long long g = (nvec + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    if (g < 1) g = 1;
    return static_cast<int>(g);
}

}  // namespace

cudaError_t launch_se_pool(const void* a, const void* b, int N, int hw, int ld, float mult, float* out,
                           cudaStream_t s) {
    if (ld & 7) return cudaErrorInvalidValue;
    dim3 grid(N, (ld + 63) / 64);
    B2C_LAUNCH(se_pool_kernel, grid, 256, 0, s, static_cast<const uint4*>(a), static_cast<const uint4*>(b), hw, ld / 8,
               mult, out);
    return cudaGetLastError();
}

cudaError_t launch_se_fc(const float* p, int N, int C, int R, int ld, const float* W1, const float* b1,
                         const float* W2, const float* b2, float* h, float* sc, cudaStream_t s) {
    B2C_LAUNCH(se_fc_kernel, N, 256, (C + R) * sizeof(float), s, p, C, R, ld, W1, b1, W2, b2, h, sc);
    return cudaGetLastError();
}

cudaError_t launch_se_scale(const void* y, const float* sc, const void* add, void* out, int N, int hw, int ld,
                            cudaStream_t s) {
    if (ld & 7) return cudaErrorInvalidValue;
    const long long nvec = static_cast<long long>(N) * hw * (ld / 8);
    B2C_LAUNCH(se_scale_kernel, ew_grid(nvec), 256, 0, s, static_cast<const uint4*>(y), sc,
               static_cast<const uint4*>(add), static_cast<uint4*>(out), nvec, hw, ld / 8);
    return cudaGetLastError();
}

cudaError_t launch_se_bwd(const void* dout, const void* y, float* ds, const float* sc, const float* h, const float* p,
                          int N, int hw, int C, int R, int ld, const float* W1, const float* W2, float* dz1, float* dp,
                          float* dW1, float* db1, float* dW2, float* db2, void* dy, cudaStream_t s) {
    if (ld & 7) return cudaErrorInvalidValue;
    dim3 pgrid(N, (ld + 63) / 64);
    B2C_LAUNCH(se_pool_kernel, pgrid, 256, 0, s, static_cast<const uint4*>(dout), static_cast<const uint4*>(y), hw,
               ld / 8, 1.f, ds);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    B2C_LAUNCH(se_fc_bwd_kernel, N, 256, (C + R) * sizeof(float), s, ds, sc, h, C, R, ld, W1, W2, dz1, dp);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    dim3 ggrid((C + 31) / 32, R);
    B2C_LAUNCH(se_param_grad_kernel, ggrid, 1024, 0, s, static_cast<const float*>(ds), static_cast<const float*>(dz1), p,
               h, N, C, R, ld, dW1, db1, dW2, db2);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const long long nvec = static_cast<long long>(N) * hw * (ld / 8);
    B2C_LAUNCH(se_bwd_apply_kernel, ew_grid(nvec), 256, 0, s, static_cast<const uint4*>(dout), sc,
               static_cast<const float*>(dp), static_cast<uint4*>(dy), nvec, hw, ld / 8, 1.f / static_cast<float>(hw));
    return cudaGetLastError();
}

}  // namespace b2c
