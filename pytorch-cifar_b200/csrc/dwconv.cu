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
                int C, int c_valid, int stride, float* __restrict__ part_sum, float* __restrict__ part_sq, int stat_acc) {
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
        if (stat_acc) atomicAdd(i < C ? part_sum + i : part_sq + (i - C), s);      // one accumulated row
        else if (i < C) part_sum[static_cast<size_t>(blockIdx.x) * C + i] = s;
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

// ------------------------------------------------------------------------------------------------
// Stride-1 variants: a thread owns 4 channels of one output column and walks a strip of T consecutive rows with a
// 3x3 window held in registers — every input vector is loaded and converted to fp32 once per thread (3 loads per
// output instead of 9, and the horizontal neighbours come from L1). The per-pixel kernels above issue ~250
// instructions per 8-channel output and are instruction-bound at a third of the HBM roofline.
//   MODE 0: fprop (+ BatchNorm partial sums)   MODE 1: dgrad (the same correlation with the weights flipped)
//   MODE 2: wgrad (window of x, one dy vector per row, 36 accumulators, block reduction + fp32 atomics)
// Work item = (image, strip, column); thread t owns channel quad t % (C/4); blockDim 256.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void unpack4(const uint2& u, float (&f)[4]) {
    f[0] = bf16lo(u.x); f[1] = bf16hi(u.x); f[2] = bf16lo(u.y); f[3] = bf16hi(u.y);
}

// XF (MODE 0 / 2): `in` is the raw output y of the preceding conv and the kernel applies that conv's BatchNorm + ReLU
// while loading, x = relu(fma(y, in_scale[c], in_shift[c])) — `out = F.relu(self.bn1(self.conv1(x)))` feeding
// `self.conv2` (depthwise), models/mobilenetv2.py:32-33 — so the activation between the two is never written or
// re-read. Zero padding is applied AFTER the transform (a padded pixel is 0, not relu(shift)).
template <int MODE, int T, bool XF = false>
__global__ void __launch_bounds__(256)
dw_strip_kernel(const uint2* __restrict__ in, const uint2* __restrict__ in2, const float* __restrict__ wt,
                uint2* __restrict__ out, float* __restrict__ dw, int N, int H, int W, int C, int c_valid,
                float* __restrict__ part_sum, float* __restrict__ part_sq, int stat_acc,
                const float* __restrict__ in_scale = nullptr, const float* __restrict__ in_shift = nullptr) {
    pdl_trigger_ew();
    pdl_wait();
    extern __shared__ float red[];   // MODE 0: [rpi][2][C]; MODE 2: [rpi][9][C]
    const int cq_n = C >> 2;
    const int rpi = 256 / cq_n;
    const int cq = threadIdx.x % cq_n, rl = threadIdx.x / cq_n;
    const bool active = rl < rpi;
    const int strips = H / T;
    const long long items = static_cast<long long>(N) * strips * W;

    float w[9][4];       // MODE 0/1: weights (flipped for dgrad); MODE 2: accumulators
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (MODE == 2) w[t][e] = 0.f;
            else w[t][e] = (cq * 4 + e < c_valid) ? __ldg(wt + (cq * 4 + e) * 9 + (MODE == 1 ? 8 - t : t)) : 0.f;
        }
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    const uint2 zero = make_uint2(0u, 0u);
    float xsc[4] = {1.f, 1.f, 1.f, 1.f}, xsh[4] = {0.f, 0.f, 0.f, 0.f};
    if (XF) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            xsc[e] = __ldg(in_scale + cq * 4 + e);
            xsh[e] = __ldg(in_shift + cq * 4 + e);
        }
    }

    for (long long it = static_cast<long long>(blockIdx.x) * rpi + rl; active && it < items;
         it += static_cast<long long>(gridDim.x) * rpi) {
        const int wo = static_cast<int>(it % W);
        const int st = static_cast<int>((it / W) % strips);
        const long long n = it / (static_cast<long long>(W) * strips);
        const int h0 = st * T;
        const bool left = wo > 0, right = wo + 1 < W;
        // row pointer arithmetic in units of uint2 (4 channels)
        const uint2* base = in + ((n * H) * W + wo) * cq_n + cq;
        const long long row_pitch = static_cast<long long>(W) * cq_n;
        bool rok = false;        // the row held in `raw` lies inside the image
        auto load_row = [&](int h, uint2 (&r)[3]) {
            rok = h >= 0 && h < H;
            if (!rok) { r[0] = zero; r[1] = zero; r[2] = zero; return; }
            const uint2* p = base + h * row_pitch;
            r[0] = left ? __ldg(p - cq_n) : zero;
            r[1] = __ldg(p);
            r[2] = right ? __ldg(p + cq_n) : zero;
        };
        // raw row -> fp32 window slot; with XF the preceding BatchNorm + ReLU is applied to pixels inside the image
        auto unpack_row = [&](const uint2 (&r)[3], float (&dst)[3][4]) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                unpack4(r[c], dst[c]);
                if (XF) {
                    const bool ok = rok && (c == 1 || (c == 0 ? left : right));
#pragma unroll
                    for (int e = 0; e < 4; ++e) dst[c][e] = ok ? fmaxf(fmaf(dst[c][e], xsc[e], xsh[e]), 0.f) : 0.f;
                }
            }
        };
        float win[3][3][4];      // [row slot][col][channel]; slot = input row modulo 3 (static after unrolling)
        uint2 raw[3];
        load_row(h0 - 1, raw);
        unpack_row(raw, win[0]);
        load_row(h0, raw);
        unpack_row(raw, win[1]);
        load_row(h0 + 1, raw);               // prefetched one row ahead of its use
#pragma unroll
        for (int i = 0; i < T; ++i) {
            // rows h0+i-1, h0+i, h0+i+1 live in slots i%3, (i+1)%3, (i+2)%3
            unpack_row(raw, win[(i + 2) % 3]);
            if (i + 1 < T) load_row(h0 + i + 2, raw);
            const long long opix = ((n * H + h0 + i) * W + wo) * cq_n + cq;
            if (MODE == 2) {
                float g[4];
                unpack4(__ldg(in2 + opix), g);
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int c = 0; c < 3; ++c)
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            w[r * 3 + c][e] = fmaf(win[(i + r) % 3][c][e], g[e], w[r * 3 + c][e]);
            } else {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int c = 0; c < 3; ++c)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[e] = fmaf(win[(i + r) % 3][c][e], w[r * 3 + c][e], acc[e]);
                uint2 o;
                o.x = pack_bf16x2(acc[0], acc[1]);
                o.y = pack_bf16x2(acc[2], acc[3]);
                out[opix] = o;
                if (MODE == 0) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        s1[e] += acc[e];
                        s2[e] = fmaf(acc[e], acc[e], s2[e]);
                    }
                }
            }
        }
    }
    if (MODE == 0) {
        if (part_sum == nullptr) return;
        if (active) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                red[(rl * 2 + 0) * C + cq * 4 + e] = s1[e];
                red[(rl * 2 + 1) * C + cq * 4 + e] = s2[e];
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < 2 * C; i += 256) {
            float s = 0.f;
            for (int k = 0; k < rpi; ++k) s += red[k * 2 * C + i];
            if (stat_acc) atomicAdd(i < C ? part_sum + i : part_sq + (i - C), s);      // one accumulated row
            else if (i < C) part_sum[static_cast<size_t>(blockIdx.x) * C + i] = s;
            else part_sq[static_cast<size_t>(blockIdx.x) * C + (i - C)] = s;
        }
    } else if (MODE == 2) {
        if (active) {
#pragma unroll
            for (int t = 0; t < 9; ++t)
#pragma unroll
                for (int e = 0; e < 4; ++e) red[(rl * 9 + t) * C + cq * 4 + e] = w[t][e];
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
}

// ------------------------------------------------------------------------------------------------
// Stride-2 variants of the strip kernels (models/mobilenetv2.py:20 with stride 2: three layers).
// fprop / wgrad: a thread owns 4 channels of one OUTPUT column and walks T output rows; output row ho reads input
// rows 2ho-1 .. 2ho+1, the last of which is the next output row's first: 6 loads per output instead of 9.
//   MODE 0: fprop (+ BatchNorm partial sums)   MODE 2: wgrad
// ------------------------------------------------------------------------------------------------
template <int MODE, int T>
__global__ void __launch_bounds__(256)
dw_strip2_kernel(const uint2* __restrict__ in, const uint2* __restrict__ in2, const float* __restrict__ wt,
                 uint2* __restrict__ out, float* __restrict__ dw, int N, int H, int W, int C, int c_valid,
                 float* __restrict__ part_sum, float* __restrict__ part_sq, int stat_acc) {
    pdl_trigger_ew();
    pdl_wait();
    extern __shared__ float red[];   // MODE 0: [rpi][2][C]; MODE 2: [rpi][9][C]
    const int cq_n = C >> 2;
    const int rpi = 256 / cq_n;
    const int cq = threadIdx.x % cq_n, rl = threadIdx.x / cq_n;
    const bool active = rl < rpi;
    const int Ho = H >> 1, Wo = W >> 1;
    const int strips = Ho / T;
    const long long items = static_cast<long long>(N) * strips * Wo;

    float w[9][4];       // MODE 0: weights; MODE 2: accumulators
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (MODE == 2) w[t][e] = 0.f;
            else w[t][e] = (cq * 4 + e < c_valid) ? __ldg(wt + (cq * 4 + e) * 9 + t) : 0.f;
        }
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    const uint2 zero = make_uint2(0u, 0u);

    for (long long it = static_cast<long long>(blockIdx.x) * rpi + rl; active && it < items;
         it += static_cast<long long>(gridDim.x) * rpi) {
        const int wo = static_cast<int>(it % Wo);
        const int st = static_cast<int>((it / Wo) % strips);
        const long long n = it / (static_cast<long long>(Wo) * strips);
        const int ho0 = st * T;
        const bool left = wo > 0;
        const uint2* base = in + ((n * H) * W + 2 * wo) * cq_n + cq;     // input pixel (row 0, col 2*wo)
        const long long row_pitch = static_cast<long long>(W) * cq_n;
        auto load_row = [&](int h, float (&r)[3][4]) {
            uint2 a = zero, b = zero, c = zero;
            if (h >= 0) {                        // h <= H-1 always (H even, h <= 2*Ho-1)
                const uint2* p = base + h * row_pitch;
                if (left) a = __ldg(p - cq_n);
                b = __ldg(p);
                c = __ldg(p + cq_n);             // column 2*wo+1 <= W-1
            }
            unpack4(a, r[0]); unpack4(b, r[1]); unpack4(c, r[2]);
        };
        float win[3][3][4];                      // rows 2ho-1, 2ho, 2ho+1
        load_row(2 * ho0 - 1, win[0]);
#pragma unroll
        for (int i = 0; i < T; ++i) {
            const int ho = ho0 + i;
            load_row(2 * ho, win[1]);
            load_row(2 * ho + 1, win[2]);
            const long long opix = ((n * Ho + ho) * Wo + wo) * cq_n + cq;
            if (MODE == 2) {
                float g[4];
                unpack4(__ldg(in2 + opix), g);
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int c = 0; c < 3; ++c)
#pragma unroll
                        for (int e = 0; e < 4; ++e) w[r * 3 + c][e] = fmaf(win[r][c][e], g[e], w[r * 3 + c][e]);
            } else {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int c = 0; c < 3; ++c)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[e] = fmaf(win[r][c][e], w[r * 3 + c][e], acc[e]);
                uint2 o;
                o.x = pack_bf16x2(acc[0], acc[1]);
                o.y = pack_bf16x2(acc[2], acc[3]);
                out[opix] = o;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    s1[e] += acc[e];
                    s2[e] = fmaf(acc[e], acc[e], s2[e]);
                }
            }
            // input row 2ho+1 is the next output row's top row
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int e = 0; e < 4; ++e) win[0][c][e] = win[2][c][e];
        }
    }
    if (MODE == 0) {
        if (part_sum == nullptr) return;
        if (active) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                red[(rl * 2 + 0) * C + cq * 4 + e] = s1[e];
                red[(rl * 2 + 1) * C + cq * 4 + e] = s2[e];
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < 2 * C; i += 256) {
            float s = 0.f;
            for (int k = 0; k < rpi; ++k) s += red[k * 2 * C + i];
            if (stat_acc) atomicAdd(i < C ? part_sum + i : part_sq + (i - C), s);      // one accumulated row
            else if (i < C) part_sum[static_cast<size_t>(blockIdx.x) * C + i] = s;
            else part_sq[static_cast<size_t>(blockIdx.x) * C + (i - C)] = s;
        }
    } else {
        if (active) {
#pragma unroll
            for (int t = 0; t < 9; ++t)
#pragma unroll
                for (int e = 0; e < 4; ++e) red[(rl * 9 + t) * C + cq * 4 + e] = w[t][e];
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
}

// Stride-2 dgrad: a thread owns 4 channels of one 2x2 block column of dx and walks T block rows. Block (i, j) =
// dx[2i..2i+1][2j..2j+1] reads dy[i..i+1][j..j+1] — the 1 / 2 / 2 / 4 taps that reach each parity phase, 9 MACs
// per block and channel, no divisions or modulo; dy row i+1 is the next block row's row i (2 loads per 4 outputs).
template <int T>
__global__ void __launch_bounds__(256)
dw_strip2_dgrad_kernel(const uint2* __restrict__ dy, const float* __restrict__ wt, uint2* __restrict__ dx, int N,
                       int H, int W, int C, int c_valid) {
    pdl_trigger_ew();
    pdl_wait();
    const int cq_n = C >> 2;
    const int rpi = 256 / cq_n;
    const int cq = threadIdx.x % cq_n, rl = threadIdx.x / cq_n;
    if (rl >= rpi) return;
    const int Ho = H >> 1, Wo = W >> 1;
    const int strips = Ho / T;
    const long long items = static_cast<long long>(N) * strips * Wo;
    float w[9][4];
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e) w[t][e] = (cq * 4 + e < c_valid) ? __ldg(wt + (cq * 4 + e) * 9 + t) : 0.f;
    const uint2 zero = make_uint2(0u, 0u);
    for (long long it = static_cast<long long>(blockIdx.x) * rpi + rl; it < items;
         it += static_cast<long long>(gridDim.x) * rpi) {
        const int j = static_cast<int>(it % Wo);
        const int st = static_cast<int>((it / Wo) % strips);
        const long long n = it / (static_cast<long long>(Wo) * strips);
        const int i0 = st * T;
        const bool right = j + 1 < Wo;
        const uint2* base = dy + ((n * Ho) * Wo + j) * cq_n + cq;
        const long long row_pitch = static_cast<long long>(Wo) * cq_n;
        auto load_row = [&](int i, float (&a)[4], float (&b)[4]) {      // dy[i][j], dy[i][j+1]
            uint2 ua = zero, ub = zero;
            if (i < Ho) {
                const uint2* p = base + i * row_pitch;
                ua = __ldg(p);
                if (right) ub = __ldg(p + cq_n);
            }
            unpack4(ua, a); unpack4(ub, b);
        };
        float a0[4], b0[4], a1[4], b1[4];        // rows i (a0, b0) and i+1 (a1, b1); a = column j, b = column j+1
        load_row(i0, a0, b0);
#pragma unroll
        for (int k = 0; k < T; ++k) {
            const int i = i0 + k;
            load_row(i + 1, a1, b1);
            float o00[4], o01[4], o10[4], o11[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                o00[e] = a0[e] * w[4][e];                                                   // (r, s) = (1, 1)
                o01[e] = fmaf(b0[e], w[3][e], a0[e] * w[5][e]);                             // (1, 0), (1, 2)
                o10[e] = fmaf(a1[e], w[1][e], a0[e] * w[7][e]);                             // (0, 1), (2, 1)
                o11[e] = fmaf(b1[e], w[0][e], fmaf(a1[e], w[2][e], fmaf(b0[e], w[6][e], a0[e] * w[8][e])));
            }
            uint2* q = dx + ((n * H + 2 * i) * W + 2 * j) * cq_n + cq;
            uint2 u;
            u.x = pack_bf16x2(o00[0], o00[1]); u.y = pack_bf16x2(o00[2], o00[3]); q[0] = u;
            u.x = pack_bf16x2(o01[0], o01[1]); u.y = pack_bf16x2(o01[2], o01[3]); q[cq_n] = u;
            q += static_cast<long long>(W) * cq_n;
            u.x = pack_bf16x2(o10[0], o10[1]); u.y = pack_bf16x2(o10[2], o10[3]); q[0] = u;
            u.x = pack_bf16x2(o11[0], o11[1]); u.y = pack_bf16x2(o11[2], o11[3]); q[cq_n] = u;
#pragma unroll
            for (int e = 0; e < 4; ++e) { a0[e] = a1[e]; b0[e] = b1[e]; }
        }
    }
}

// stride-1 strip kernels: T = 8 rows (images of 8 pixels and more) or 4
inline bool dw_strip_ok(int H, int W, int C, int stride) {
    return stride == 1 && (C & 3) == 0 && (C >> 2) <= 256 && H >= 4 && (H % 4) == 0 && W >= 2;
}

inline bool dw_strip2_ok(int H, int W, int C, int stride) {
    return stride == 2 && (C & 3) == 0 && (C >> 2) <= 256 && !(H & 1) && !(W & 1) && (H >> 1) >= 4 && ((H >> 1) % 4) == 0;
}

// MODE 0 fprop, 1 dgrad, 2 wgrad (stride 2)
template <int MODE>
cudaError_t launch_dw_strip2(const void* in, const void* in2, const float* w, void* out, float* dw, int N, int H, int W,
                             int C, int c_valid, float* part_sum, float* part_sq, int* stat_rows, cudaStream_t s, int stat_acc = 0) {
    const int cq_n = C >> 2;
    const int rpi = 256 / cq_n;
    const int Ho = H >> 1, Wo = W >> 1;
    const int T = (Ho % 8) == 0 ? 8 : 4;
    const long long items = static_cast<long long>(N) * (Ho / T) * Wo;
    long long blocks = (items + rpi - 1) / rpi;
    const int cap = MODE == 2 ? 296 : 148 * 4;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (stat_rows != nullptr) *stat_rows = stat_acc ? 1 : static_cast<int>(blocks);
    const int grid = static_cast<int>(blocks);
    if (MODE == 1) {
        if (T == 8)
            B2C_LAUNCH(dw_strip2_dgrad_kernel<8>, grid, 256, 0, s, static_cast<const uint2*>(in), w,
                       static_cast<uint2*>(out), N, H, W, C, c_valid);
        else
            B2C_LAUNCH(dw_strip2_dgrad_kernel<4>, grid, 256, 0, s, static_cast<const uint2*>(in), w,
                       static_cast<uint2*>(out), N, H, W, C, c_valid);
        return cudaGetLastError();
    }
    constexpr int M = MODE == 1 ? 0 : MODE;
    size_t smem = 0;
    if (M == 0 && part_sum != nullptr) smem = static_cast<size_t>(rpi) * 2 * C * sizeof(float);
    if (M == 2) smem = static_cast<size_t>(rpi) * 9 * C * sizeof(float);
    auto k8 = dw_strip2_kernel<M, 8>;
    auto k4 = dw_strip2_kernel<M, 4>;
    if (M == 2) {
        static bool attr_done = false;
        if (!attr_done) {
            cudaError_t e = cudaFuncSetAttribute(k8, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            if (e != cudaSuccess) return e;
            attr_done = true;
        }
    }
    if (T == 8)
        B2C_LAUNCH(k8, grid, 256, smem, s, static_cast<const uint2*>(in), static_cast<const uint2*>(in2), w,
                   static_cast<uint2*>(out), dw, N, H, W, C, c_valid, part_sum, part_sq, stat_acc);
    else
        B2C_LAUNCH(k4, grid, 256, smem, s, static_cast<const uint2*>(in), static_cast<const uint2*>(in2), w,
                   static_cast<uint2*>(out), dw, N, H, W, C, c_valid, part_sum, part_sq, stat_acc);
    return cudaGetLastError();
}

template <int MODE>
cudaError_t launch_dw_strip(const void* in, const void* in2, const float* w, void* out, float* dw, int N, int H, int W,
                            int C, int c_valid, float* part_sum, float* part_sq, int* stat_rows, cudaStream_t s, int stat_acc = 0,
                            const float* in_scale = nullptr, const float* in_shift = nullptr) {
    const int cq_n = C >> 2;
    const int rpi = 256 / cq_n;
    const int T = (H % 8) == 0 ? 8 : 4;
    const long long items = static_cast<long long>(N) * (H / T) * W;
    long long blocks = (items + rpi - 1) / rpi;
    const int cap = MODE == 2 ? 296 : 148 * 4;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (stat_rows != nullptr) *stat_rows = stat_acc ? 1 : static_cast<int>(blocks);
    size_t smem = 0;
    if (MODE == 0 && part_sum != nullptr) smem = static_cast<size_t>(rpi) * 2 * C * sizeof(float);
    if (MODE == 2) smem = static_cast<size_t>(rpi) * 9 * C * sizeof(float);
    auto k8 = dw_strip_kernel<MODE, 8>;
    auto k4 = dw_strip_kernel<MODE, 4>;
    constexpr bool kCanXf = MODE != 1;
    auto x8 = dw_strip_kernel<MODE, 8, kCanXf>;       // input transformed on load (fprop / wgrad only)
    auto x4 = dw_strip_kernel<MODE, 4, kCanXf>;
    const bool xf = in_scale != nullptr;
    if (xf && (!kCanXf || in_shift == nullptr)) return cudaErrorInvalidValue;
    if (MODE == 2) {
        static bool attr_done = false;
        if (!attr_done) {
            cudaError_t e = cudaFuncSetAttribute(k8, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(x8, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            if (e == cudaSuccess) e = cudaFuncSetAttribute(x4, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            if (e != cudaSuccess) return e;
            attr_done = true;
        }
    }
    auto kern = T == 8 ? (xf ? x8 : k8) : (xf ? x4 : k4);
    B2C_LAUNCH(kern, static_cast<int>(blocks), 256, smem, s, static_cast<const uint2*>(in),
               static_cast<const uint2*>(in2), w, static_cast<uint2*>(out), dw, N, H, W, C, c_valid, part_sum,
               part_sq, stat_acc, in_scale, in_shift);
    return cudaGetLastError();
}

}  // namespace

bool dw_input_transform_ok(int H, int W, int C, int stride) { return dw_strip_ok(H, W, C, stride); }

cudaError_t launch_dw_fprop_xf(const void* y_prev, const float* in_scale, const float* in_shift, const float* w, void* y,
                               int N, int H, int W, int C, int c_valid, float* part_sum, float* part_sq, int* stat_rows,
                               cudaStream_t s, int stat_acc) {
    if (!dw_strip_ok(H, W, C, 1)) return cudaErrorInvalidValue;
    return launch_dw_strip<0>(y_prev, nullptr, w, y, nullptr, N, H, W, C, c_valid, part_sum, part_sq, stat_rows, s,
                              stat_acc, in_scale, in_shift);
}

cudaError_t launch_dw_wgrad_xf(const void* y_prev, const float* in_scale, const float* in_shift, const void* dy,
                               float* dw, int N, int H, int W, int C, int c_valid, cudaStream_t s) {
    if (!dw_strip_ok(H, W, C, 1)) return cudaErrorInvalidValue;
    return launch_dw_strip<2>(y_prev, dy, nullptr, nullptr, dw, N, H, W, C, c_valid, nullptr, nullptr, nullptr, s, 0,
                              in_scale, in_shift);
}

int dw_fprop_partial_rows(long long out_pixels, int C) {
    const int rpi = 256 / (C >> 3);
    long long blocks = (out_pixels + rpi - 1) / rpi;
    if (blocks > 148 * 4) blocks = 148 * 4;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

cudaError_t launch_dw_fprop(const void* x, const float* w, void* y, int N, int H, int W, int C, int c_valid, int stride,
                            float* part_sum, float* part_sq, int* stat_rows, cudaStream_t s, int stat_acc) {
    const int cvec = C >> 3;
    if (cvec < 1 || cvec > 256) return cudaErrorInvalidValue;
    if (dw_strip_ok(H, W, C, stride))
        return launch_dw_strip<0>(x, nullptr, w, y, nullptr, N, H, W, C, c_valid, part_sum, part_sq, stat_rows, s, stat_acc);
    if (dw_strip2_ok(H, W, C, stride))
        return launch_dw_strip2<0>(x, nullptr, w, y, nullptr, N, H, W, C, c_valid, part_sum, part_sq, stat_rows, s, stat_acc);
    const int rpi = 256 / cvec;
    const int blocks = dw_fprop_partial_rows(static_cast<long long>(N) * (H / stride) * (W / stride), C);
    if (stat_rows != nullptr) *stat_rows = stat_acc ? 1 : blocks;
    const size_t smem = part_sum != nullptr ? static_cast<size_t>(rpi) * 2 * C * sizeof(float) : 0;
    B2C_LAUNCH(dw_fprop_kernel, blocks, 256, smem, s, static_cast<const uint4*>(x), w, static_cast<uint4*>(y), N, H, W, C,
                                              c_valid, stride, part_sum, part_sq, stat_acc);
    return cudaGetLastError();
}

cudaError_t launch_dw_dgrad(const void* dy, const float* w, void* dx, int N, int H, int W, int C, int c_valid,
                            int stride, cudaStream_t s) {
    const int cvec = C >> 3;
    if (cvec < 1 || cvec > 256) return cudaErrorInvalidValue;
    if (dw_strip_ok(H, W, C, stride))
        return launch_dw_strip<1>(dy, nullptr, w, dx, nullptr, N, H, W, C, c_valid, nullptr, nullptr, nullptr, s);
    if (dw_strip2_ok(H, W, C, stride))
        return launch_dw_strip2<1>(dy, nullptr, w, dx, nullptr, N, H, W, C, c_valid, nullptr, nullptr, nullptr, s);
    const int blocks = dw_fprop_partial_rows(static_cast<long long>(N) * H * W, C);
    B2C_LAUNCH(dw_dgrad_kernel, blocks * 2 > 148 * 8 ? 148 * 8 : blocks * 2, 256, 0, s, 
        static_cast<const uint4*>(dy), w, static_cast<uint4*>(dx), N, H, W, C, c_valid, stride);
    return cudaGetLastError();
}

cudaError_t launch_dw_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int c_valid,
                            int stride, cudaStream_t s) {
    const int cvec = C >> 3;
    if (cvec < 1 || cvec > 256) return cudaErrorInvalidValue;
    if (dw_strip_ok(H, W, C, stride))
        return launch_dw_strip<2>(x, dy, nullptr, nullptr, dw, N, H, W, C, c_valid, nullptr, nullptr, nullptr, s);
    if (dw_strip2_ok(H, W, C, stride))
        return launch_dw_strip2<2>(x, dy, nullptr, nullptr, dw, N, H, W, C, c_valid, nullptr, nullptr, nullptr, s);
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
