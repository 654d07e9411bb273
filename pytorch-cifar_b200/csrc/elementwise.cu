// elementwise.cu — the HBM-bound kernels of the training step (NHWC bf16 activations, fp32 statistics):
// BatchNorm finalize / apply / backward, the pooling+linear+cross-entropy head, the input
// augmentation, fused SGD and the weight re-layout. Each kernel states the reference call site it
// replaces; all of them are judged against HBM bandwidth, not the tensor pipe.
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
__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
// Per-channel parameters re-read every iteration (L1 hit) instead of living in registers: `volatile` keeps the
// compiler from hoisting the load out of the row loop, which would cost 16 registers and a resident block per SM.
__device__ __forceinline__ float4 ld_param4(const float* p) {
    float4 r;
    asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
// ReLU mask recomputed from the conv output: act = relu(fma(y, scale, shift)) > 0  <=>  fma(y, scale, shift) > 0.
__device__ __forceinline__ void mask_from_y8(float (&d)[8], const float (&yy)[8], const float* scale,
                                             const float* shift) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const float4 sc = ld_param4(scale + 4 * h), sh = ld_param4(shift + 4 * h);
        d[4 * h + 0] = fmaf(yy[4 * h + 0], sc.x, sh.x) > 0.f ? d[4 * h + 0] : 0.f;
        d[4 * h + 1] = fmaf(yy[4 * h + 1], sc.y, sh.y) > 0.f ? d[4 * h + 1] : 0.f;
        d[4 * h + 2] = fmaf(yy[4 * h + 2], sc.z, sh.z) > 0.f ? d[4 * h + 2] : 0.f;
        d[4 * h + 3] = fmaf(yy[4 * h + 3], sc.w, sh.w) > 0.f ? d[4 * h + 3] : 0.f;
    }
}

// ------------------------------------------------------------------------------------------------
// BatchNorm finalize (training statistics). Reference: nn.BatchNorm2d in train mode,
// models/resnet.py:25,28,35,104 -> F.batch_norm: biased variance for normalisation, unbiased for the
// running estimate, momentum 0.1, eps 1e-5. One block per 32 channels; its 32 warps stride over the per-CTA
// partial rows the conv epilogue wrote and combine in fp64 (colsum2_block).
// ------------------------------------------------------------------------------------------------
// Column sums of two [rows][stride] fp32 matrices for the 32 channels of this block, in fp64. 32 warps walk the
// rows (one coalesced 128-byte read per warp and row, four rows in flight), then combine through shared memory.
// Latency matters more than bandwidth here: the inputs are a few hundred KB but the kernel sits on the critical path
// between two full-tensor passes, 40 times per step.
constexpr int kFinWarps = 32;
__device__ __forceinline__ void colsum2_block(const float* __restrict__ a, const float* __restrict__ b, size_t stride,
                                              int rows, bool valid, double& s1, double& s2) {
    __shared__ double sm[2][kFinWarps][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    double d1 = 0.0, d2 = 0.0;
    if (valid) {
        int r = w;
        for (; r + 3 * kFinWarps < rows; r += 4 * kFinWarps) {
            float x[4], z[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                x[u] = a[static_cast<size_t>(r + u * kFinWarps) * stride];
                z[u] = b[static_cast<size_t>(r + u * kFinWarps) * stride];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) { d1 += static_cast<double>(x[u]); d2 += static_cast<double>(z[u]); }
        }
        for (; r < rows; r += kFinWarps) {
            d1 += static_cast<double>(a[static_cast<size_t>(r) * stride]);
            d2 += static_cast<double>(b[static_cast<size_t>(r) * stride]);
        }
    }
    sm[0][w][lane] = d1;
    sm[1][w][lane] = d2;
    __syncthreads();
    if (w < 2) {                       // warp 0 folds the first matrix, warp 1 the second
        double t = 0.0;
#pragma unroll 8
        for (int k = 0; k < kFinWarps; ++k) t += sm[w][k][lane];
        sm[w][0][lane] = t;
    }
    __syncthreads();
    s1 = sm[0][0][lane];
    s2 = sm[1][0][lane];
}

__global__ void __launch_bounds__(kFinWarps * 32)
bn_finalize_kernel(const float* __restrict__ part_sum, const float* __restrict__ part_sq, int rows,
                                   int C, float count, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, const float* __restrict__ conv_bias, float eps,
                                   float momentum,
                                   float* __restrict__ running_mean, float* __restrict__ running_var,
                                   float* __restrict__ scale, float* __restrict__ shift,
                                   float* __restrict__ mean_out, float* __restrict__ rstd_out) {
    pdl_trigger_tiny();
    pdl_wait();
    const int warp = blockIdx.x * 32 + (threadIdx.x & 31);      // channel (name kept from the warp-per-channel form)
    double s1, s2;
    colsum2_block(part_sum + warp, part_sq + warp, static_cast<size_t>(C), rows, warp < C, s1, s2);
    if (warp >= C) return;
    const int lane = threadIdx.x >> 5;                            // 0 for the warp that writes the results
    if (lane == 0) {
        const double n = static_cast<double>(count);
        const double mean = s1 / n;
        double var = s2 / n - mean * mean;
        if (var < 0.0) var = 0.0;
        const float rstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
        const float g = gamma[warp];
        const float sc = g * rstd;
        scale[warp] = sc;
        shift[warp] = beta[warp] - static_cast<float>(mean) * sc;
        mean_out[warp] = static_cast<float>(mean);
        rstd_out[warp] = rstd;
        if (running_mean != nullptr) {
            const double unbiased = n > 1.0 ? var * n / (n - 1.0) : var;
            const float bias = conv_bias != nullptr ? conv_bias[warp] : 0.f;   // reference tracks mean(conv + bias)
            running_mean[warp] = (1.f - momentum) * running_mean[warp] + momentum * (static_cast<float>(mean) + bias);
            running_var[warp] = (1.f - momentum) * running_var[warp] + momentum * static_cast<float>(unbiased);
        }
    }
}

// Eval-mode BatchNorm: fold the running statistics into scale / shift (main.py:116-134 `net.eval()`).
__global__ void bn_eval_scale_shift_kernel(int C, const float* __restrict__ gamma, const float* __restrict__ beta,
                                           const float* __restrict__ conv_bias,
                                           const float* __restrict__ running_mean,
                                           const float* __restrict__ running_var, float eps,
                                           float* __restrict__ scale, float* __restrict__ shift) {
    pdl_trigger_tiny();
    pdl_wait();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float sc = gamma[c] / sqrtf(running_var[c] + eps);
    scale[c] = sc;
    const float bias = conv_bias != nullptr ? conv_bias[c] : 0.f;
    shift[c] = beta[c] - (running_mean[c] - bias) * sc;
}

// ------------------------------------------------------------------------------------------------
// BN apply + residual + ReLU. Reference: `F.relu(self.bn1(...))`, `out += self.shortcut(x)`,
// `F.relu(out)` models/resnet.py:47-50. 16-byte vectors of 8 channels; C % 8 == 0.
//   out = act( y*scale + shift  [+ res]  or  [+ res*res_scale + res_shift] )
// ------------------------------------------------------------------------------------------------
template <bool kUniform, bool kResAffine>
__global__ void __launch_bounds__(256)
bn_act_apply_kernel(const uint4* __restrict__ y, const float* __restrict__ scale, const float* __restrict__ shift,
                    const uint4* __restrict__ res, const float* __restrict__ res_scale,
                    const float* __restrict__ res_shift, uint4* __restrict__ out, long long nvec, int C, int relu,
                    uint8_t* __restrict__ relu_bits) {
    pdl_trigger_ew();
    pdl_wait();
    // kUniform: the grid stride is a multiple of C/8, so a thread sees the same 8 channels in every iteration and
    // keeps their scale / shift in registers (no 64-bit modulo, no per-element parameter loads in the loop); two
    // grid strides are then in flight per iteration (all loads issued before the first use).
    const int cvec = C >> 3;
    const long long start = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    float sc[8], sh[8], rs[8], rb[8];
    auto load_params = [&](int c0) {
        const float4 sa = __ldg(reinterpret_cast<const float4*>(scale + c0));
        const float4 sb = __ldg(reinterpret_cast<const float4*>(scale + c0 + 4));
        const float4 ha = __ldg(reinterpret_cast<const float4*>(shift + c0));
        const float4 hb = __ldg(reinterpret_cast<const float4*>(shift + c0 + 4));
        sc[0] = sa.x; sc[1] = sa.y; sc[2] = sa.z; sc[3] = sa.w; sc[4] = sb.x; sc[5] = sb.y; sc[6] = sb.z; sc[7] = sb.w;
        sh[0] = ha.x; sh[1] = ha.y; sh[2] = ha.z; sh[3] = ha.w; sh[4] = hb.x; sh[5] = hb.y; sh[6] = hb.z; sh[7] = hb.w;
        if (kResAffine) {
            const float4 ra = __ldg(reinterpret_cast<const float4*>(res_scale + c0));
            const float4 rc = __ldg(reinterpret_cast<const float4*>(res_scale + c0 + 4));
            const float4 ta = __ldg(reinterpret_cast<const float4*>(res_shift + c0));
            const float4 tb = __ldg(reinterpret_cast<const float4*>(res_shift + c0 + 4));
            rs[0] = ra.x; rs[1] = ra.y; rs[2] = ra.z; rs[3] = ra.w; rs[4] = rc.x; rs[5] = rc.y; rs[6] = rc.z; rs[7] = rc.w;
            rb[0] = ta.x; rb[1] = ta.y; rb[2] = ta.z; rb[3] = ta.w; rb[4] = tb.x; rb[5] = tb.y; rb[6] = tb.z; rb[7] = tb.w;
        }
    };
    auto finish = [&](const uint4 yraw, const uint4 rraw, long long v) {
        float f[8];
        unpack8(yraw, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = fmaf(f[e], sc[e], sh[e]);
        if (res != nullptr) {
            float r[8];
            unpack8(rraw, r);
            if (kResAffine) {
#pragma unroll
                for (int e = 0; e < 8; ++e) r[e] = fmaf(r[e], rs[e], rb[e]);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] += r[e];
        }
        if (relu) {
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
        }
        out[v] = pack8(f);
        if (relu_bits != nullptr) {      // one byte per 8 channels: bit e = [out[e] > 0], for the backward pass
            unsigned bits = 0;
#pragma unroll
            for (int e = 0; e < 8; ++e) bits |= (f[e] > 0.f ? 1u : 0u) << e;
            relu_bits[v] = static_cast<uint8_t>(bits);
        }
    };
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    if (kUniform) {
        load_params(static_cast<int>(start % cvec) << 3);
        for (long long v = start; v < nvec; v += 2 * step) {
            const long long v2 = v + step;
            const bool two = v2 < nvec;
            const uint4 y0 = ld_stream(y + v);
            const uint4 y1 = two ? ld_stream(y + v2) : zero;
            const uint4 r0 = res != nullptr ? ld_stream(res + v) : zero;
            const uint4 r1 = (res != nullptr && two) ? ld_stream(res + v2) : zero;
            finish(y0, r0, v);
            if (two) finish(y1, r1, v2);
        }
    } else {
        for (long long v = start; v < nvec; v += step) {
            load_params(static_cast<int>(v % cvec) << 3);
            finish(ld_stream(y + v), res != nullptr ? ld_stream(res + v) : zero, v);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// The same pass with the BatchNorm finalize folded into its prologue (b2c_bn_act_apply_stats): the producing conv
// ACCUMULATED its per-channel sum / sum of squares into one row (fp32 atomics, B2cConvDesc.flags bit 9), so scale and
// shift are a few operations per channel. Every block derives them for all C channels into shared memory (2*C floats
// read from L2 per block); block 0 also publishes scale / shift / mean / rstd and updates the running statistics —
// what bn_finalize_kernel did as a launch of its own, 20 times per ResNet-18 forward.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void bn_stats_to_smem(const BnStatArgs& st, int C, float* s_scale, float* s_shift,
                                                 bool publish) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const double n = static_cast<double>(st.count);
        const double mean = static_cast<double>(st.sum[c]) / n;
        double var = static_cast<double>(st.sumsq[c]) / n - mean * mean;
        if (var < 0.0) var = 0.0;
        const float rstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(st.eps)));
        const float sc = st.gamma[c] * rstd;
        const float sh = st.beta[c] - static_cast<float>(mean) * sc;
        s_scale[c] = sc;
        s_shift[c] = sh;
        if (publish) {
            st.scale[c] = sc;
            st.shift[c] = sh;
            st.save_mean[c] = static_cast<float>(mean);
            st.save_rstd[c] = rstd;
            if (st.running_mean != nullptr) {
                const double unbiased = n > 1.0 ? var * n / (n - 1.0) : var;
                const float bias = st.conv_bias != nullptr ? st.conv_bias[c] : 0.f;   // reference tracks mean(conv + bias)
                st.running_mean[c] = (1.f - st.momentum) * st.running_mean[c] + st.momentum * (static_cast<float>(mean) + bias);
                st.running_var[c] = (1.f - st.momentum) * st.running_var[c] + st.momentum * static_cast<float>(unbiased);
            }
        }
    }
}

template <bool kResAffine>
__global__ void __launch_bounds__(256)
bn_act_apply_stats_kernel(const uint4* __restrict__ y, const BnStatArgs st, const uint4* __restrict__ res,
                          const BnStatArgs rst, uint4* __restrict__ out, long long nvec, int C, int relu,
                          uint8_t* __restrict__ relu_bits) {
    pdl_trigger_ew();
    pdl_wait();
    extern __shared__ float s_par[];      // [scale | shift | res_scale | res_shift], C floats each
    float* s_scale = s_par;
    float* s_shift = s_par + C;
    bn_stats_to_smem(st, C, s_scale, s_shift, blockIdx.x == 0);
    if (kResAffine) bn_stats_to_smem(rst, C, s_par + 2 * C, s_par + 3 * C, blockIdx.x == 0);
    __syncthreads();
    const int cvec = C >> 3;
    const long long start = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    // the launcher makes the grid stride a multiple of C/8: a thread keeps its 8 channels' parameters in registers
    const int c0 = static_cast<int>(start % cvec) << 3;
    float sc[8], sh[8], rs[8], rb[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        sc[e] = s_scale[c0 + e];
        sh[e] = s_shift[c0 + e];
        rs[e] = kResAffine ? s_par[2 * C + c0 + e] : 1.f;
        rb[e] = kResAffine ? s_par[3 * C + c0 + e] : 0.f;
    }
    auto finish = [&](const uint4 yraw, const uint4 rraw, long long v) {
        float f[8];
        unpack8(yraw, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = fmaf(f[e], sc[e], sh[e]);
        if (res != nullptr) {
            float r[8];
            unpack8(rraw, r);
            if (kResAffine) {
#pragma unroll
                for (int e = 0; e < 8; ++e) r[e] = fmaf(r[e], rs[e], rb[e]);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] += r[e];
        }
        if (relu) {
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
        }
        out[v] = pack8(f);
        if (relu_bits != nullptr) {
            unsigned bits = 0;
#pragma unroll
            for (int e = 0; e < 8; ++e) bits |= (f[e] > 0.f ? 1u : 0u) << e;
            relu_bits[v] = static_cast<uint8_t>(bits);
        }
    };
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    for (long long v = start; v < nvec; v += 2 * step) {
        const long long v2 = v + step;
        const bool two = v2 < nvec;
        const uint4 y0 = ld_stream(y + v);
        const uint4 y1 = two ? ld_stream(y + v2) : zero;
        const uint4 r0 = res != nullptr ? ld_stream(res + v) : zero;
        const uint4 r1 = (res != nullptr && two) ? ld_stream(res + v2) : zero;
        finish(y0, r0, v);
        if (two) finish(y1, r1, v2);
    }
}

// ------------------------------------------------------------------------------------------------
// BN(+ReLU) backward, pass 1: per-channel sums of dz and dz*y, where dz = dout * [act > 0].
// Reference: autograd of F.relu + nn.BatchNorm2d (cudnn_batch_norm_backward), main.py:104.
// If `dz_out` is given, the masked gradient is written back (so later passes and the shortcut
// branch read dz directly). Each block writes one partial row [2][C]; bn_bwd_finalize combines them.
// Thread layout: blockDim = 256; thread t owns channel vector (t % cvec) and walks rows.
// ------------------------------------------------------------------------------------------------
// kMask: 0 = mask from `act` (or none), 1 = recomputed from y, 2 = from the forward pass's bit mask
template <int kMask>
__global__ void __launch_bounds__(256)
bn_bwd_reduce_kernel(const uint4* __restrict__ dout, const uint4* __restrict__ act, const uint4* __restrict__ y,
                     uint4* __restrict__ dz_out, long long rows, int C, float* __restrict__ partial,
                     const float* __restrict__ relu_scale, const float* __restrict__ relu_shift,
                     const uint8_t* __restrict__ relu_bits, int accumulate) {
    pdl_trigger_ew();
    pdl_wait();
    extern __shared__ float red[];  // [rows_per_iter][2][C]
    const int cvec = C >> 3;
    const int rpi = 256 / cvec;     // rows handled per block iteration (cvec <= 256); threads past rpi*cvec idle
    const int cv = threadIdx.x % cvec;
    const int rl = threadIdx.x / cvec;
    const bool active = rl < rpi;
    float s1[8], s2[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { s1[e] = 0.f; s2[e] = 0.f; }
    // mask from y (the forward's own expression): neither the activation nor a materialised dz is read or written
    // for plain BN+ReLU units
    for (long long r = static_cast<long long>(blockIdx.x) * rpi + rl; active && r < rows;
         r += static_cast<long long>(gridDim.x) * rpi) {
        const long long v = r * cvec + cv;
        float d[8], yy[8];
        unpack8(ld_stream(dout + v), d);
        unpack8(ld_stream(y + v), yy);
        if (kMask == 1) mask_from_y8(d, yy, relu_scale + cv * 8, relu_shift + cv * 8);
        if (kMask == 2) {
            const unsigned bits = relu_bits[v];
#pragma unroll
            for (int e = 0; e < 8; ++e) d[e] = (bits >> e) & 1u ? d[e] : 0.f;
            if (dz_out != nullptr) dz_out[v] = pack8(d);
        }
        if (kMask == 0 && act != nullptr) {
            float a[8];
            unpack8(ld_stream(act + v), a);
#pragma unroll
            for (int e = 0; e < 8; ++e) d[e] = a[e] > 0.f ? d[e] : 0.f;
            if (dz_out != nullptr) dz_out[v] = pack8(d);
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            s1[e] += d[e];
            s2[e] = fmaf(d[e], yy[e], s2[e]);
        }
    }
    // block reduce over the rl dimension
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
        if (accumulate) atomicAdd(partial + i, s);      // one row [2][C], zero-filled by the caller
        else partial[static_cast<size_t>(blockIdx.x) * 2 * C + i] = s;
    }
}

// Pass 2 (tiny): combine partial rows, emit dgamma / dbeta and the per-channel coefficients of
//   dy = A*dz + B*y + Cc   (= gamma*rstd*(dz - dbeta/n - xhat*dgamma/n), xhat = (y-mean)*rstd).
__global__ void __launch_bounds__(kFinWarps * 32)
bn_bwd_finalize_kernel(const float* __restrict__ partial, int prow, int C, float count,
                                       const float* __restrict__ gamma, const float* __restrict__ mean,
                                       const float* __restrict__ rstd, float* __restrict__ dgamma,
                                       float* __restrict__ dbeta, float* __restrict__ coefA,
                                       float* __restrict__ coefB, float* __restrict__ coefC) {
    pdl_trigger_tiny();
    pdl_wait();
    const int warp = blockIdx.x * 32 + (threadIdx.x & 31);      // channel
    double s1, s2;
    colsum2_block(partial + warp, partial + C + warp, static_cast<size_t>(2) * C, prow, warp < C, s1, s2);
    if (warp >= C) return;
    const int lane = threadIdx.x >> 5;
    if (lane == 0) {
        const double mu = mean[warp], rs = rstd[warp], g = gamma[warp], n = count;
        const double db = s1;
        const double dg = rs * (s2 - mu * s1);
        dgamma[warp] = static_cast<float>(dg);
        dbeta[warp] = static_cast<float>(db);
        const double A = g * rs;
        const double B = -g * rs * rs * dg / n;
        const double Cc = -g * rs * db / n + g * rs * rs * mu * dg / n;
        coefA[warp] = static_cast<float>(A);
        coefB[warp] = static_cast<float>(B);
        coefC[warp] = static_cast<float>(Cc);
    }
}

// Pass 3: dy = A*dz + B*y + Cc (+ add) (bf16 out). kUniform as in bn_act_apply_kernel.
template <bool kUniform, bool kMaskY>
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const uint4* __restrict__ dz, const uint4* __restrict__ y, const float* __restrict__ coefA,
                    const float* __restrict__ coefB, const float* __restrict__ coefC, const uint4* __restrict__ add,
                    uint4* __restrict__ dy, long long nvec, int C, const float* __restrict__ relu_scale,
                    const float* __restrict__ relu_shift) {
    pdl_trigger_ew();
    pdl_wait();
    const int cvec = C >> 3;
    const long long start = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    float ca[8], cb[8], cc[8];
    int cur_c0 = 0;
    auto load_params = [&](int c0) {
        cur_c0 = c0;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float4 a4 = __ldg(reinterpret_cast<const float4*>(coefA + c0 + 4 * h));
            const float4 b4 = __ldg(reinterpret_cast<const float4*>(coefB + c0 + 4 * h));
            const float4 c4 = __ldg(reinterpret_cast<const float4*>(coefC + c0 + 4 * h));
            ca[4 * h] = a4.x; ca[4 * h + 1] = a4.y; ca[4 * h + 2] = a4.z; ca[4 * h + 3] = a4.w;
            cb[4 * h] = b4.x; cb[4 * h + 1] = b4.y; cb[4 * h + 2] = b4.z; cb[4 * h + 3] = b4.w;
            cc[4 * h] = c4.x; cc[4 * h + 1] = c4.y; cc[4 * h + 2] = c4.z; cc[4 * h + 3] = c4.w;
        }
    };
    if (kUniform) load_params(static_cast<int>(start % cvec) << 3);
    for (long long v = start; v < nvec; v += step) {
        if (!kUniform) load_params(static_cast<int>(v % cvec) << 3);
        float d[8], yy[8], o[8];
        unpack8(ld_stream(dz + v), d);
        unpack8(ld_stream(y + v), yy);
        // dz was not materialised: apply the ReLU mask here (same expression as the reduce pass)
        if (kMaskY) mask_from_y8(d, yy, relu_scale + cur_c0, relu_shift + cur_c0);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = fmaf(ca[e], d[e], fmaf(cb[e], yy[e], cc[e]));
        if (add != nullptr) {
            float a[8];
            unpack8(ld_stream(add + v), a);
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] += a[e];
        }
        dy[v] = pack8(o);
    }
}

// Pass 3 with pass 2 folded into its prologue (b2c_bn_bwd_apply_stats): `sums` is ONE accumulated row [2][C]
// (sum dz | sum dz*y). Every block derives A, B, Cc for all channels into shared memory; block 0 writes dgamma / dbeta.
__global__ void __launch_bounds__(256)
bn_bwd_apply_stats_kernel(const uint4* __restrict__ dz, const uint4* __restrict__ y, const float* __restrict__ sums,
                          float count, const float* __restrict__ gamma, const float* __restrict__ mean,
                          const float* __restrict__ rstd, float* __restrict__ dgamma, float* __restrict__ dbeta,
                          const uint4* __restrict__ add, uint4* __restrict__ dy, long long nvec, int C,
                          const float* __restrict__ relu_scale, const float* __restrict__ relu_shift) {
    pdl_trigger_ew();
    pdl_wait();
    extern __shared__ float s_coef[];     // [A | B | Cc], C floats each
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const double s1 = sums[c], s2 = sums[C + c];
        const double mu = mean[c], rs = rstd[c], g = gamma[c], n = count;
        const double dg = rs * (s2 - mu * s1);
        if (blockIdx.x == 0) {
            dgamma[c] = static_cast<float>(dg);
            dbeta[c] = static_cast<float>(s1);
        }
        s_coef[c] = static_cast<float>(g * rs);
        s_coef[C + c] = static_cast<float>(-g * rs * rs * dg / n);
        s_coef[2 * C + c] = static_cast<float>(-g * rs * s1 / n + g * rs * rs * mu * dg / n);
    }
    __syncthreads();
    const int cvec = C >> 3;
    const long long start = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
    const int c0 = static_cast<int>(start % cvec) << 3;      // grid stride is a multiple of C/8 (launcher)
    float ca[8], cb[8], cc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        ca[e] = s_coef[c0 + e];
        cb[e] = s_coef[C + c0 + e];
        cc[e] = s_coef[2 * C + c0 + e];
    }
    for (long long v = start; v < nvec; v += step) {
        float d[8], yy[8], o[8];
        unpack8(ld_stream(dz + v), d);
        unpack8(ld_stream(y + v), yy);
        if (relu_scale != nullptr) mask_from_y8(d, yy, relu_scale + c0, relu_shift + c0);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = fmaf(ca[e], d[e], fmaf(cb[e], yy[e], cc[e]));
        if (add != nullptr) {
            float a[8];
            unpack8(ld_stream(add + v), a);
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] += a[e];
        }
        dy[v] = pack8(o);
    }
}

// ------------------------------------------------------------------------------------------------
// nn.MaxPool2d(2, 2) forward / backward on NHWC bf16 (models/vgg.py:31). One thread = 8 channels of one
// output pixel. Backward routes the gradient to the FIRST maximum of the window in (row, col) scan order,
// which is what torch's max_pool2d_with_indices records.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, long long nvec_out, int Ho, int Wo, int C) {
    pdl_trigger_ew();
    pdl_wait();
    const int cvec = C >> 3;
    for (long long v = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; v < nvec_out;
         v += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int cv = static_cast<int>(v % cvec);
        long long pix = v / cvec;
        const int wo = static_cast<int>(pix % Wo);
        pix /= Wo;
        const int ho = static_cast<int>(pix % Ho);
        const long long n = pix / Ho;
        const long long W = 2LL * Wo;
        const long long base = ((n * 2 * Ho + 2 * ho) * W + 2 * wo) * cvec + cv;
        float a[8], b[8], c[8], d[8], m[8];
        unpack8(ld_stream(x + base), a);
        unpack8(ld_stream(x + base + cvec), b);
        unpack8(ld_stream(x + base + W * cvec), c);
        unpack8(ld_stream(x + base + W * cvec + cvec), d);
#pragma unroll
        for (int e = 0; e < 8; ++e) m[e] = fmaxf(fmaxf(a[e], b[e]), fmaxf(c[e], d[e]));
        y[v] = pack8(m);
    }
}

__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const uint4* __restrict__ x, const uint4* __restrict__ y, const uint4* __restrict__ dy,
                   uint4* __restrict__ dx, long long nvec_out, int Ho, int Wo, int C) {
    pdl_trigger_ew();
    pdl_wait();
    const int cvec = C >> 3;
    for (long long v = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; v < nvec_out;
         v += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int cv = static_cast<int>(v % cvec);
        long long pix = v / cvec;
        const int wo = static_cast<int>(pix % Wo);
        pix /= Wo;
        const int ho = static_cast<int>(pix % Ho);
        const long long n = pix / Ho;
        const long long W = 2LL * Wo;
        const long long base = ((n * 2 * Ho + 2 * ho) * W + 2 * wo) * cvec + cv;
        float a[8], b[8], c[8], d[8], m[8], g[8];
        unpack8(ld_stream(x + base), a);
        unpack8(ld_stream(x + base + cvec), b);
        unpack8(ld_stream(x + base + W * cvec), c);
        unpack8(ld_stream(x + base + W * cvec + cvec), d);
        unpack8(ld_stream(y + v), m);
        unpack8(ld_stream(dy + v), g);
        float oa[8], ob[8], oc[8], od[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const bool ha = a[e] == m[e];
            const bool hb = !ha && b[e] == m[e];
            const bool hc = !ha && !hb && c[e] == m[e];
            const bool hd = !ha && !hb && !hc;
            oa[e] = ha ? g[e] : 0.f;
            ob[e] = hb ? g[e] : 0.f;
            oc[e] = hc ? g[e] : 0.f;
            od[e] = hd ? g[e] : 0.f;
        }
        dx[base] = pack8(oa);
        dx[base + cvec] = pack8(ob);
        dx[base + W * cvec] = pack8(oc);
        dx[base + W * cvec + cvec] = pack8(od);
    }
}

// ------------------------------------------------------------------------------------------------
// Head forward. Reference: F.avg_pool2d(out, 4) -> view -> nn.Linear(512, 10), models/resnet.py:137-139.
// One block per image; thread c owns channel c (blockDim.x == C).
// ------------------------------------------------------------------------------------------------
// ld = channel pitch of act (>= C: layouts padded to 64 channels, e.g. RegNetX's 368-channel final map).
__global__ void head_fwd_kernel(const __nv_bfloat16* __restrict__ act, int hw, int C, int ld, const float* __restrict__ W,
                                const float* __restrict__ bias, int K, float* __restrict__ pooled,
                                float* __restrict__ logits) {
    pdl_trigger_tiny();
    pdl_wait();
    extern __shared__ float sp[];  // [C]
    const int n = blockIdx.x;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float s = 0.f;
        const __nv_bfloat16* a = act + static_cast<size_t>(n) * hw * ld + c;
        for (int i = 0; i < hw; ++i) s += __bfloat162float(a[static_cast<size_t>(i) * ld]);
        s *= 1.f / static_cast<float>(hw);
        sp[c] = s;
        pooled[static_cast<size_t>(n) * C + c] = s;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int k = warp; k < K; k += nwarps) {
        float acc = 0.f;
        for (int j = lane; j < C; j += 32) acc = fmaf(sp[j], __ldg(W + static_cast<size_t>(k) * C + j), acc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) logits[static_cast<size_t>(n) * K + k] = acc + bias[k];
    }
}

// Cross-entropy (mean reduction) and its gradient. Reference: nn.CrossEntropyLoss(), main.py:86,103.
// One thread per image; a single block reduces the loss (N <= a few thousand).
__global__ void ce_loss_kernel(const float* __restrict__ logits, const long long* __restrict__ target, int N, int K,
                               float* __restrict__ dlogits, float* __restrict__ loss_out,
                               int* __restrict__ correct_out) {
    pdl_trigger_tiny();
    pdl_wait();
    __shared__ float sl[32];
    __shared__ int sc[32];
    float local = 0.f;
    int corr = 0;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        const float* l = logits + static_cast<size_t>(n) * K;
        float m = l[0];
        int arg = 0;
        for (int k = 1; k < K; ++k) {
            if (l[k] > m) { m = l[k]; arg = k; }
        }
        float se = 0.f;
        for (int k = 0; k < K; ++k) se += expf(l[k] - m);
        const float lse = m + logf(se);
        const int t = static_cast<int>(target[n]);
        local += lse - l[t];
        corr += (arg == t);
        const float invn = 1.f / static_cast<float>(N);
        for (int k = 0; k < K; ++k) {
            const float pk = expf(l[k] - lse);
            dlogits[static_cast<size_t>(n) * K + k] = (pk - (k == t ? 1.f : 0.f)) * invn;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        local += __shfl_xor_sync(0xffffffffu, local, o);
        corr += __shfl_xor_sync(0xffffffffu, corr, o);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sl[warp] = local; sc[warp] = corr; }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        local = lane < nw ? sl[lane] : 0.f;
        corr = lane < nw ? sc[lane] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            local += __shfl_xor_sync(0xffffffffu, local, o);
            corr += __shfl_xor_sync(0xffffffffu, corr, o);
        }
        if (lane == 0) {
            loss_out[0] = local / static_cast<float>(N);
            if (correct_out != nullptr) correct_out[0] = corr;
        }
    }
}

// Head backward, part 1: gradient w.r.t. the last activation (avg-pool + linear backward), bf16 NHWC.
__global__ void head_bwd_act_kernel(const float* __restrict__ dlogits, const float* __restrict__ W, int K, int hw,
                                    int C, int ld, __nv_bfloat16* __restrict__ dact) {
    pdl_trigger_tiny();
    pdl_wait();
    const int n = blockIdx.x;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float g = 0.f;
        for (int k = 0; k < K; ++k)
            g = fmaf(dlogits[static_cast<size_t>(n) * K + k], __ldg(W + static_cast<size_t>(k) * C + c), g);
        g *= 1.f / static_cast<float>(hw);
        const __nv_bfloat16 b = __float2bfloat16(g);
        __nv_bfloat16* d = dact + static_cast<size_t>(n) * hw * ld + c;
        for (int i = 0; i < hw; ++i) d[static_cast<size_t>(i) * ld] = b;
    }
}
// Head backward, part 2: dW[k, c] = sum_n dlogits[n,k] * pooled[n,c]; db[k] = sum_n dlogits[n,k].
// grid = (C/32, K); block = 32 channels x 32 batch lanes (four samples in flight per lane), smem fold over the lanes.
__global__ void __launch_bounds__(1024)
head_bwd_param_kernel(const float* __restrict__ dlogits, const float* __restrict__ pooled, int N,
                      int K, int C, float* __restrict__ dW, float* __restrict__ db) {
    pdl_trigger_tiny();
    pdl_wait();
    __shared__ float red[32][33];
    const int k = blockIdx.y;
    const int cl = threadIdx.x & 31, nl = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cl;
    const bool valid = c < C;
    const float* pc = pooled + (valid ? c : 0);
    float s = 0.f, sb = 0.f;
    int n = nl;
    for (; n + 96 < N; n += 128) {
        float d[4], q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            d[u] = __ldg(dlogits + static_cast<size_t>(n + 32 * u) * K + k);
            q[u] = __ldg(pc + static_cast<size_t>(n + 32 * u) * C);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) { s = fmaf(d[u], q[u], s); sb += d[u]; }
    }
    for (; n < N; n += 32) {
        const float d = __ldg(dlogits + static_cast<size_t>(n) * K + k);
        s = fmaf(d, __ldg(pc + static_cast<size_t>(n) * C), s);
        sb += d;
    }
    red[nl][cl] = s;
    if (cl == 0) red[nl][32] = sb;         // every lane of a warp holds the same sb
    __syncthreads();
    if (nl == 0) {
        float t = 0.f;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) t += red[i][cl];
        if (valid) dW[static_cast<size_t>(k) * C + c] = t;
    } else if (nl == 1 && cl == 0 && blockIdx.x == 0) {
        float t = 0.f;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) t += red[i][32];
        db[k] = t;
    }
}

// ------------------------------------------------------------------------------------------------
// Input pipeline. Reference: transforms.Compose([RandomCrop(32, padding=4), RandomHorizontalFlip(),
// ToTensor(), Normalize(mean, std)]) main.py:30-35, per sample in DataLoader workers.
// The uint8 dataset [Nd,32,32,3] lives in HBM; (i, j, flip) per sample come from the host-side
// MT19937 restatement (bit-exact with torchvision's draws). Pixel math is the reference's:
// pad(0) on uint8 -> crop -> flip -> x/255 -> (x-mean)/std in fp32.
//
// augment_sample(): value of augmented pixel (y, x, c), y/x in [0,32).
// ------------------------------------------------------------------------------------------------
struct AugConst {
    float mean[3];
    float std[3];
};

__device__ __forceinline__ float aug_value(const uint8_t* simg, int y, int x, int c, int ci, int cj, int flip,
                                           const AugConst& k) {
    const int xs = flip ? 31 - x : x;
    const int oy = y + ci - 4;
    const int ox = xs + cj - 4;
    unsigned int u = 0;
    if (oy >= 0 && oy < 32 && ox >= 0 && ox < 32) u = simg[(oy * 32 + ox) * 3 + c];
    const float f = __fdiv_rn(static_cast<float>(u), 255.f);   // ToTensor: .div(255)
    return __fdiv_rn(__fsub_rn(f, k.mean[c]), k.std[c]);       // Normalize: sub_(mean).div_(std)
}

// fp32 NCHW output — the reference's tensor, for parity checks and the drop-in path.
__global__ void __launch_bounds__(256)
augment_nchw_kernel(const uint8_t* __restrict__ data, const long long* __restrict__ idx,
                    const int8_t* __restrict__ ijf, float* __restrict__ out, AugConst k) {
    pdl_trigger_ew();
    pdl_wait();
    __shared__ __align__(16) uint8_t simg[3072];
    const int n = blockIdx.x;
    const uint4* src = reinterpret_cast<const uint4*>(data + static_cast<size_t>(idx[n]) * 3072);
    if (threadIdx.x < 192) reinterpret_cast<uint4*>(simg)[threadIdx.x] = __ldg(src + threadIdx.x);
    __syncthreads();
    const int ci = ijf[n * 3 + 0], cj = ijf[n * 3 + 1], flip = ijf[n * 3 + 2];
    float* o = out + static_cast<size_t>(n) * 3072;
    for (int e = threadIdx.x; e < 3072; e += 256) {
        const int c = e >> 10, y = (e >> 5) & 31, x = e & 31;
        o[e] = aug_value(simg, y, x, c, ci, cj, flip, k);
    }
}

// Fused fast path: augmented image written directly as the stem conv's im2col operand,
// [N,32,32,64] bf16 with k = (r*3+s)*3 + c for k < 27 and zeros above (conv zero padding applied).
// kTaps = 1: the operand of a 1x1 stem conv (models/resnext.py:47), k = c for k < 3.
template <int kTaps>
__device__ __forceinline__ __nv_bfloat16 stem_operand_value(const __nv_bfloat16* saug, int y, int x, int kk) {
    if (kTaps == 1) return kk < 3 ? saug[((y + 1) * 34 + (x + 1)) * 3 + kk] : __float2bfloat16(0.f);
    if (kk >= 27) return __float2bfloat16(0.f);
    const int tap = kk / 3, c = kk - tap * 3;
    const int r = tap / 3, s = tap - r * 3;
    return saug[((y + r) * 34 + (x + s)) * 3 + c];
}

template <int kTaps>
__global__ void __launch_bounds__(256)
augment_im2col_kernel(const uint8_t* __restrict__ data, const long long* __restrict__ idx,
                      const int8_t* __restrict__ ijf, uint4* __restrict__ out, AugConst k) {
    pdl_trigger_ew();
    pdl_wait();
    __shared__ __align__(16) uint8_t simg[3072];
    __shared__ __nv_bfloat16 saug[34 * 34 * 3];  // augmented image with the conv's 1-px zero border
    const int n = blockIdx.x;
    const uint4* src = reinterpret_cast<const uint4*>(data + static_cast<size_t>(idx[n]) * 3072);
    if (threadIdx.x < 192) reinterpret_cast<uint4*>(simg)[threadIdx.x] = __ldg(src + threadIdx.x);
    __syncthreads();
    const int ci = ijf[n * 3 + 0], cj = ijf[n * 3 + 1], flip = ijf[n * 3 + 2];
    for (int e = threadIdx.x; e < 34 * 34 * 3; e += 256) {
        const int c = e % 3, x = (e / 3) % 34 - 1, y = e / (3 * 34) - 1;
        float v = 0.f;
        if (y >= 0 && y < 32 && x >= 0 && x < 32) v = aug_value(simg, y, x, c, ci, cj, flip, k);
        saug[e] = __float2bfloat16(v);
    }
    __syncthreads();
    // 1024 pixels x 8 vectors of 8 bf16
    uint4* o = out + static_cast<size_t>(n) * 1024 * 8;
    for (int e = threadIdx.x; e < 1024 * 8; e += 256) {
        const int pix = e >> 3, vec = e & 7;
        const int y = pix >> 5, x = pix & 31;
        __nv_bfloat16 h[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) h[q] = stem_operand_value<kTaps>(saug, y, x, vec * 8 + q);
        o[e] = *reinterpret_cast<uint4*>(h);
    }
}

// Drop-in path: fp32 NCHW network input (what the reference's DataLoader yields) -> stem im2col operand.
template <int kTaps>
__global__ void __launch_bounds__(256)
stem_im2col_kernel(const float* __restrict__ x, uint4* __restrict__ out) {
    pdl_trigger_ew();
    pdl_wait();
    __shared__ __nv_bfloat16 saug[34 * 34 * 3];
    const int n = blockIdx.x;
    const float* xi = x + static_cast<size_t>(n) * 3072;
    for (int e = threadIdx.x; e < 34 * 34 * 3; e += 256) {
        const int c = e % 3, xx = (e / 3) % 34 - 1, y = e / (3 * 34) - 1;
        float v = 0.f;
        if (y >= 0 && y < 32 && xx >= 0 && xx < 32) v = xi[c * 1024 + y * 32 + xx];
        saug[e] = __float2bfloat16(v);
    }
    __syncthreads();
    uint4* o = out + static_cast<size_t>(n) * 1024 * 8;
    for (int e = threadIdx.x; e < 1024 * 8; e += 256) {
        const int pix = e >> 3, vec = e & 7;
        const int y = pix >> 5, xx = pix & 31;
        __nv_bfloat16 h[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) h[q] = stem_operand_value<kTaps>(saug, y, xx, vec * 8 + q);
        o[e] = *reinterpret_cast<uint4*>(h);
    }
}

// ------------------------------------------------------------------------------------------------
// Grouped conv (nn.Conv2d(..., groups=g), models/resnext.py:19). The conv runs as a dense conv whose bf16 operands are
// block-diagonal (one repack entry per group fills its diagonal block, the rest stays zero), so the weight-gradient
// kernels produce a dense [tap][cin][cout] fp32 matrix; only its diagonal blocks are gradients of real parameters:
//   dw[o][ci][tap] += dense[tap][(o / cout_g) * cin_g + ci][o]        (dw in the reference's [cout][cin/g][kh][kw])
// One thread per diagonal element, in the dense matrix's order (coalesced reads, scattered 4-byte writes: the whole
// gradient is at most a few MB).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
grouped_wgrad_extract_kernel(const float* __restrict__ dense, float* __restrict__ dw, int cout, int cin, int taps,
                             int groups) {
    pdl_trigger_ew();
    pdl_wait();
    const int cout_g = cout / groups, cin_g = cin / groups;
    const long long n = static_cast<long long>(taps) * cin * cout_g;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int ol = static_cast<int>(i % cout_g);
        const long long r = i / cout_g;
        const int c = static_cast<int>(r % cin), t = static_cast<int>(r / cin);
        const int g = c / cin_g, o = g * cout_g + ol;
        dw[(static_cast<long long>(o) * cin_g + (c - g * cin_g)) * taps + t] +=
            dense[(static_cast<long long>(t) * cin + c) * cout + o];
    }
}

// ------------------------------------------------------------------------------------------------
// Fused SGD. Reference: optim.SGD(lr, momentum=0.9, weight_decay=5e-4).step(), main.py:87-88,105:
//   d = g + wd*p;  buf = d (first step) | momentum*buf + d;  p -= lr*buf      (dampening 0, no Nesterov)
// over one flat fp32 arena holding every parameter (BN gamma/beta and biases included).
// hyper = {lr, momentum, weight_decay, grad_scale (1/world for DDP mean), first_step} on the device,
// so a captured CUDA graph replays with new values.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sgd_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ buf, long long nvec,
           const float* __restrict__ hyper) {
    pdl_trigger_ew();
    pdl_wait();
    const float lr = hyper[0], mom = hyper[1], wd = hyper[2], gs = hyper[3];
    const bool first = hyper[4] != 0.f;
    for (long long v = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; v < nvec;
         v += static_cast<long long>(gridDim.x) * blockDim.x) {
        float4 pp = p[v];
        const float4 gg = g[v];
        float4 bb = first ? make_float4(0.f, 0.f, 0.f, 0.f) : buf[v];
        const float dx = fmaf(wd, pp.x, gg.x * gs), dy = fmaf(wd, pp.y, gg.y * gs);
        const float dz = fmaf(wd, pp.z, gg.z * gs), dw = fmaf(wd, pp.w, gg.w * gs);
        if (first) { bb.x = dx; bb.y = dy; bb.z = dz; bb.w = dw; }
        else {
            bb.x = fmaf(mom, bb.x, dx); bb.y = fmaf(mom, bb.y, dy);
            bb.z = fmaf(mom, bb.z, dz); bb.w = fmaf(mom, bb.w, dw);
        }
        pp.x = fmaf(-lr, bb.x, pp.x); pp.y = fmaf(-lr, bb.y, pp.y);
        pp.z = fmaf(-lr, bb.z, pp.z); pp.w = fmaf(-lr, bb.w, pp.w);
        buf[v] = bb;
        p[v] = pp;
    }
}

// ------------------------------------------------------------------------------------------------
// Weight re-layout: fp32 OIHW master weights (the reference's state_dict layout, models/resnet.py:23-27)
// -> the two bf16 operand layouts the conv kernels read:
//    fprop B  [tap][cout][cin_pad]      dgrad B  [tap][cin][cout]
// One launch for every conv of the network: blockIdx.y = tensor index into a device table.
// kflat != 0 (stem): fprop B is [cout][64] with k = tap*cin + c (matches the im2col operand), no dgrad copy.
// ------------------------------------------------------------------------------------------------
// Work of one table entry in tiles, the unit one block processes per iteration: kflat 1 (stem) = 256 elements,
// kflat 2 = one 32 (cin) x 64 (cout) tile of one tap, kflat 0 = one 32 x 32 tile with all its taps.
__device__ __forceinline__ int repack_entry_tiles(const RepackEntry& e) {
    if (e.kflat == 1) return (e.cout * e.cin * e.taps + 255) >> 8;
    if (e.kflat == 2) return e.taps * ((e.cin + 31) >> 5) * ((e.cout + 63) >> 6);
    return ((e.cin + 31) >> 5) * ((e.cout + 31) >> 5);
}

// The tiles of ALL entries form one list that the grid strides over, so a block's share of the work does not depend
// on which tensor it belongs to (a fixed number of blocks per tensor left the four 2.4 M-element layer-4 weights of
// ResNet-18, 75 % of all bytes, to 4 x 128 blocks running 18 tiles each in series: 65 us for 89 MB).
__global__ void __launch_bounds__(256)
repack_kernel(const float* __restrict__ master, __nv_bfloat16* __restrict__ wf, __nv_bfloat16* __restrict__ wd,
              const RepackEntry* __restrict__ table, int entries) {
    pdl_trigger_ew();
    pdl_wait();
    __shared__ float tile[32][32 * 9 + 1];
    // first tile of every entry, once per block (a dependent chain of global loads per entry otherwise)
    constexpr int kMaxEntries = 1024;
    __shared__ int s_first[kMaxEntries + 1];
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int i = 0; i < entries; ++i) {
            s_first[i] = acc;
            acc += repack_entry_tiles(table[i]);
        }
        s_first[entries] = acc;
    }
    __syncthreads();
    const int total = s_first[entries];
    int ei = 0;
    for (int g = blockIdx.x; g < total; g += gridDim.x) {
        while (g >= s_first[ei + 1]) ++ei;
        const RepackEntry e = table[ei];
        const int first = s_first[ei];
        const int tidx = g - first;
        const int taps = e.taps;
        if (e.kflat == 1) {
            // stem: [cout][cin*taps] fp32 -> [cout][cin_pad] bf16 with k = tap*cin + c
            const int row = e.cin * taps;
            const int i = tidx * 256 + threadIdx.x;
            if (i < e.cout * row) {
                const int o = i / row, r = i - o * row;
                const int c = r / taps, t = r - c * taps;
                wf[e.wf_off + static_cast<long long>(o) * e.cin_pad + t * e.cin + c] = __float2bfloat16(master[e.src_off + i]);
            }
            continue;
        }
        const int cout_pad = e.pad_ > 0 ? e.pad_ : e.cout;      // pad_ carries cout_pad (0: no padding)
        if (e.kflat == 2) {
            // master is [tap][cin][cout] fp32 (the engine's internal layout of conv weights, which the wgrad kernels
            // write with coalesced vector reductions): the dgrad operand is a straight bf16 conversion, the fprop
            // operand a 32x32 transpose per tap. Channel counts are even; pairs go out as bf16x2.
            // Tile = 32 (cin) x 64 (cout): 16-byte loads of whole 256-byte rows, 8-byte stores of both operands.
            float (*t2)[65] = reinterpret_cast<float (*)[65]>(&tile[0][0]);
            const int ct = (e.cin + 31) >> 5, ot = (e.cout + 63) >> 6;
            const int t = tidx / (ct * ot);
            const int rem = tidx - t * ct * ot;
            const int c0 = (rem / ot) << 5, o0 = (rem % ot) << 6;
            const int nc = min(32, e.cin - c0), no = min(64, e.cout - o0);
            const bool vec = !(e.cout & 3) && !(e.cin & 3) && !(cout_pad & 3) && !(e.cin_pad & 3) && !(e.src_off & 3) &&
                             !(e.wf_off & 3) && !((e.wd_off < 0 ? 0 : e.wd_off) & 3);
            __syncthreads();
            for (int i = threadIdx.x; i < 32 * 16; i += 256) {       // 32 rows x 16 float4
                const int a = i >> 4, b = (i & 15) << 2;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                const float* src = master + e.src_off + (static_cast<long long>(t) * e.cin + c0 + a) * e.cout + o0 + b;
                if (a < nc) {
                    if (vec && b + 3 < no) v = *reinterpret_cast<const float4*>(src);
                    else {
                        if (b < no) v.x = src[0];
                        if (b + 1 < no) v.y = src[1];
                        if (b + 2 < no) v.z = src[2];
                        if (b + 3 < no) v.w = src[3];
                    }
                    // dgrad operand [tap][cin_pad][cout_pad]: the same row, converted
                    if (e.wd_off >= 0) {
                        __nv_bfloat16* dst = wd + e.wd_off + (static_cast<long long>(t) * e.cin_pad + c0 + a) * cout_pad + o0 + b;
                        if (vec && b + 3 < no) {
                            uint2 u;
                            u.x = pack_bf16x2(v.x, v.y);
                            u.y = pack_bf16x2(v.z, v.w);
                            *reinterpret_cast<uint2*>(dst) = u;
                        } else {
                            if (b < no) dst[0] = __float2bfloat16(v.x);
                            if (b + 1 < no) dst[1] = __float2bfloat16(v.y);
                            if (b + 2 < no) dst[2] = __float2bfloat16(v.z);
                            if (b + 3 < no) dst[3] = __float2bfloat16(v.w);
                        }
                    }
                }
                t2[a][b] = v.x; t2[a][b + 1] = v.y; t2[a][b + 2] = v.z; t2[a][b + 3] = v.w;
            }
            __syncthreads();
            // fprop operand [tap][cout_pad][cin_pad]: row o holds the tile's 32 cin values, 4 per thread
            for (int i = threadIdx.x; i < 64 * 8; i += 256) {
                const int a = i >> 3, b = (i & 7) << 2;            // a: cout within the tile, b: cin within the tile
                if (a >= no || b >= nc) continue;
                __nv_bfloat16* dst = wf + e.wf_off + (static_cast<long long>(t) * cout_pad + o0 + a) * e.cin_pad + c0 + b;
                if (vec && b + 3 < nc) {
                    uint2 u;
                    u.x = pack_bf16x2(t2[b][a], t2[b + 1][a]);
                    u.y = pack_bf16x2(t2[b + 2][a], t2[b + 3][a]);
                    *reinterpret_cast<uint2*>(dst) = u;
                } else {
                    for (int k = 0; k < 4 && b + k < nc; ++k) dst[k] = __float2bfloat16(t2[b + k][a]);
                }
            }
            continue;
        }
        // kflat 0: OIHW master. One block = a 32 (cout) x 32 (cin) tile, all taps: coalesced fp32 reads of the OIHW
        // rows, transposes through shared memory, coalesced bf16 writes of both operand layouts.
        const int ctiles = (e.cin + 31) >> 5;
        const int o0 = (tidx / ctiles) << 5, c0 = (tidx % ctiles) << 5;
        const int no = min(32, e.cout - o0), nc = min(32, e.cin - c0);
        const int w = nc * taps;                                  // contiguous floats per cout row of the tile
        __syncthreads();
        for (int i = threadIdx.x; i < no * w; i += 256) {
            const int o = i / w, r = i - o * w;
            tile[o][r] = master[e.src_off + (static_cast<long long>(o0 + o) * e.cin + c0) * taps + r];
        }
        __syncthreads();
        // two adjacent channels per thread: one 4-byte bf16x2 store per operand layout (channel counts are even)
        const bool even = !((nc | no | cout_pad | e.cin_pad) & 1) && !((e.wf_off | (e.wd_off < 0 ? 0 : e.wd_off)) & 1);
        for (int i = threadIdx.x; i < taps * 32 * 16; i += 256) {
            const int t = i >> 9, a = (i >> 4) & 31, b = (i & 15) * 2;
            // fprop operand [tap][cout_pad][cin_pad]: b runs over cin (contiguous)
            if (a < no && b < nc) {
                __nv_bfloat16* dst = wf + e.wf_off + (static_cast<long long>(t) * cout_pad + o0 + a) * e.cin_pad + c0 + b;
                if (even) {
                    *reinterpret_cast<__nv_bfloat162*>(dst) =
                        __floats2bfloat162_rn(tile[a][b * taps + t], tile[a][(b + 1) * taps + t]);
                } else {
                    dst[0] = __float2bfloat16(tile[a][b * taps + t]);
                    if (b + 1 < nc) dst[1] = __float2bfloat16(tile[a][(b + 1) * taps + t]);
                }
            }
            // dgrad operand [tap][cin_pad][cout_pad]: b runs over cout (contiguous)
            if (e.wd_off >= 0 && a < nc && b < no) {
                __nv_bfloat16* dst = wd + e.wd_off + (static_cast<long long>(t) * e.cin_pad + c0 + a) * cout_pad + o0 + b;
                if (even) {
                    *reinterpret_cast<__nv_bfloat162*>(dst) =
                        __floats2bfloat162_rn(tile[b][a * taps + t], tile[b + 1][a * taps + t]);
                } else {
                    dst[0] = __float2bfloat16(tile[b][a * taps + t]);
                    if (b + 1 < no) dst[1] = __float2bfloat16(tile[b + 1][a * taps + t]);
                }
            }
        }
    }
}

inline int grid_for(long long work_items, int block, int max_blocks) {
    long long g = (work_items + block - 1) / block;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return static_cast<int>(g);
}

}  // namespace

// ---------------------------------------- launchers ---------------------------------------------
// Grid whose stride (grid * 256 threads) is a multiple of C/8 vectors, so a thread meets the same 8 channels in every
// iteration: round the preferred grid down to a multiple of cvec / gcd(cvec, 256).
static int uniform_grid(long long nvec, int cvec) {
    int g = grid_for(nvec, 256, 148 * 8);
    int a = cvec, b = 256;
    while (b != 0) { const int t = a % b; a = b; b = t; }
    const int unit = cvec / a;
    if (g >= unit) g -= g % unit;
    else g = unit;
    return g;
}

cudaError_t launch_bn_finalize(const float* part_sum, const float* part_sq, int rows, int C, float count,
                               const float* gamma, const float* beta, const float* conv_bias, float eps,
                               float momentum, float* running_mean, float* running_var, float* scale, float* shift,
                               float* mean, float* rstd, cudaStream_t s) {
    B2C_LAUNCH(bn_finalize_kernel, (C + 31) / 32, kFinWarps * 32, 0, s, part_sum, part_sq, rows, C, count, gamma, beta, conv_bias, eps, momentum,
                                                  running_mean, running_var, scale, shift, mean, rstd);
    return cudaGetLastError();
}

cudaError_t launch_bn_eval_scale_shift(int C, const float* gamma, const float* beta, const float* conv_bias,
                                       const float* rm, const float* rv, float eps, float* scale, float* shift,
                                       cudaStream_t s) {
    B2C_LAUNCH(bn_eval_scale_shift_kernel, (C + 127) / 128, 128, 0, s, C, gamma, beta, conv_bias, rm, rv, eps, scale, shift);
    return cudaGetLastError();
}

cudaError_t launch_bn_act_apply(const void* y, const float* scale, const float* shift, const void* res,
                                const float* res_scale, const float* res_shift, void* out, long long rows, int C,
                                int relu, uint8_t* relu_bits, cudaStream_t s) {
    const long long nvec = rows * (C >> 3);
    // a grid whose stride is a multiple of C/8: every channel count takes the register-resident (kUniform) variant
    // (MobileNetV2's 96 / 144 / 192 / 576 / 960-channel layers fell to the per-iteration parameter loads otherwise)
    const int grid = uniform_grid(nvec, C >> 3);
    const bool uniform = (static_cast<long long>(grid) * 256) % (C >> 3) == 0;
    const bool affine = res != nullptr && res_scale != nullptr;
    auto go = [&](auto kern) {
        B2C_LAUNCH(kern, grid, 256, 0, s, static_cast<const uint4*>(y), scale, shift, static_cast<const uint4*>(res), res_scale,
                                  res_shift, static_cast<uint4*>(out), nvec, C, relu, relu_bits);
    };
    if (uniform && affine) go(bn_act_apply_kernel<true, true>);
    else if (uniform) go(bn_act_apply_kernel<true, false>);
    else if (affine) go(bn_act_apply_kernel<false, true>);
    else go(bn_act_apply_kernel<false, false>);
    return cudaGetLastError();
}

// Upper bound of the partial rows bn_bwd_reduce writes (= its grid): one constant shared by the grid cap below and by
// the fixed scratch of the no-allocation entry point, so raising one cannot silently overrun the other.
constexpr int kBnBwdMaxPartialRows = 148 * 4;

int bn_bwd_partial_rows(long long rows, int C) {
    const int rpi = 256 / (C >> 3);
    long long blocks = (rows + rpi - 1) / rpi;
    if (blocks > kBnBwdMaxPartialRows) blocks = kBnBwdMaxPartialRows;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

cudaError_t launch_bn_bwd_reduce(const void* dout, const void* act, const void* y, void* dz_out, long long rows,
                                 int C, float* partial, const float* relu_scale, const float* relu_shift,
                                 const uint8_t* relu_bits, cudaStream_t s, int accumulate) {
    const int cvec = C >> 3;
    if (cvec < 1 || cvec > 256) return cudaErrorInvalidValue;
    const int rpi = 256 / cvec;
    const int blocks = bn_bwd_partial_rows(rows, C);
    const size_t smem = static_cast<size_t>(rpi) * 2 * C * sizeof(float);
    if (relu_bits != nullptr)
        B2C_LAUNCH(bn_bwd_reduce_kernel<2>, blocks, 256, smem, s, static_cast<const uint4*>(dout), nullptr,
                                                         static_cast<const uint4*>(y), static_cast<uint4*>(dz_out),
                                                         rows, C, partial, nullptr, nullptr, relu_bits, accumulate);
    else if (act == nullptr && relu_scale != nullptr)
        B2C_LAUNCH(bn_bwd_reduce_kernel<1>, blocks, 256, smem, s, static_cast<const uint4*>(dout), nullptr,
                                                         static_cast<const uint4*>(y), nullptr, rows, C, partial,
                                                         relu_scale, relu_shift, nullptr, accumulate);
    else
        B2C_LAUNCH(bn_bwd_reduce_kernel<0>, blocks, 256, smem, s, static_cast<const uint4*>(dout),
                                                         static_cast<const uint4*>(act),
                                                         static_cast<const uint4*>(y), static_cast<uint4*>(dz_out),
                                                         rows, C, partial, nullptr, nullptr, nullptr, accumulate);
    return cudaGetLastError();
}

cudaError_t launch_bn_bwd_finalize(const float* partial, int prow, int C, float count, const float* gamma,
                                   const float* mean, const float* rstd, float* dgamma, float* dbeta, float* coefA,
                                   float* coefB, float* coefC, cudaStream_t s) {
    B2C_LAUNCH(bn_bwd_finalize_kernel, (C + 31) / 32, kFinWarps * 32, 0, s, partial, prow, C, count, gamma, mean, rstd, dgamma, dbeta,
                                                      coefA, coefB, coefC);
    return cudaGetLastError();
}

cudaError_t launch_bn_bwd_apply(const void* dz, const void* y, const float* coefA, const float* coefB,
                                const float* coefC, const void* add, void* dy, long long rows, int C,
                                const float* relu_scale, const float* relu_shift, cudaStream_t s) {
    const long long nvec = rows * (C >> 3);
    const int grid = uniform_grid(nvec, C >> 3);
    const bool uniform = (static_cast<long long>(grid) * 256) % (C >> 3) == 0;
    const bool masky = relu_scale != nullptr;
    auto go = [&](auto kern) {
        B2C_LAUNCH(kern, grid, 256, 0, s, static_cast<const uint4*>(dz), static_cast<const uint4*>(y), coefA, coefB, coefC,
                                  static_cast<const uint4*>(add), static_cast<uint4*>(dy), nvec, C, relu_scale,
                                  relu_shift);
    };
    if (uniform && masky) go(bn_bwd_apply_kernel<true, true>);
    else if (uniform) go(bn_bwd_apply_kernel<true, false>);
    else if (masky) go(bn_bwd_apply_kernel<false, true>);
    else go(bn_bwd_apply_kernel<false, false>);
    return cudaGetLastError();
}

cudaError_t launch_bn_act_apply_stats(const void* y, const BnStatArgs& st, const void* res, const BnStatArgs* rst,
                                      void* out, long long rows, int C, int relu, uint8_t* relu_bits, cudaStream_t s) {
    const long long nvec = rows * (C >> 3);
    const int grid = uniform_grid(nvec, C >> 3);
    const bool affine = res != nullptr && rst != nullptr;
    const size_t smem = static_cast<size_t>(affine ? 4 : 2) * C * sizeof(float);
    const BnStatArgs none = {};
    if (affine)
        B2C_LAUNCH(bn_act_apply_stats_kernel<true>, grid, 256, smem, s, static_cast<const uint4*>(y), st,
                   static_cast<const uint4*>(res), *rst, static_cast<uint4*>(out), nvec, C, relu, relu_bits);
    else
        B2C_LAUNCH(bn_act_apply_stats_kernel<false>, grid, 256, smem, s, static_cast<const uint4*>(y), st,
                   static_cast<const uint4*>(res), none, static_cast<uint4*>(out), nvec, C, relu, relu_bits);
    return cudaGetLastError();
}

cudaError_t launch_bn_bwd_apply_stats(const void* dz, const void* y, const float* sums, float count,
                                      const float* gamma, const float* mean, const float* rstd, float* dgamma,
                                      float* dbeta, const void* add, void* dy, long long rows, int C,
                                      const float* relu_scale, const float* relu_shift, cudaStream_t s) {
    const long long nvec = rows * (C >> 3);
    const int grid = uniform_grid(nvec, C >> 3);
    B2C_LAUNCH(bn_bwd_apply_stats_kernel, grid, 256, static_cast<size_t>(3) * C * sizeof(float), s,
               static_cast<const uint4*>(dz), static_cast<const uint4*>(y), sums, count, gamma, mean, rstd, dgamma,
               dbeta, static_cast<const uint4*>(add), static_cast<uint4*>(dy), nvec, C, relu_scale, relu_shift);
    return cudaGetLastError();
}

cudaError_t launch_maxpool_fwd(const void* x, void* y, int n, int h, int w, int C, cudaStream_t s) {
    const long long nvec = static_cast<long long>(n) * (h / 2) * (w / 2) * (C >> 3);
    B2C_LAUNCH(maxpool_fwd_kernel, grid_for(nvec, 256, 148 * 8), 256, 0, s, static_cast<const uint4*>(x),
                                                                   static_cast<uint4*>(y), nvec, h / 2, w / 2, C);
    return cudaGetLastError();
}

cudaError_t launch_maxpool_bwd(const void* x, const void* y, const void* dy, void* dx, int n, int h, int w, int C,
                               cudaStream_t s) {
    const long long nvec = static_cast<long long>(n) * (h / 2) * (w / 2) * (C >> 3);
    B2C_LAUNCH(maxpool_bwd_kernel, grid_for(nvec, 256, 148 * 8), 256, 0, s, 
        static_cast<const uint4*>(x), static_cast<const uint4*>(y), static_cast<const uint4*>(dy),
        static_cast<uint4*>(dx), nvec, h / 2, w / 2, C);
    return cudaGetLastError();
}

cudaError_t launch_head_fwd(const void* act, int N, int hw, int C, const float* W, const float* bias, int K,
                            float* pooled, float* logits, cudaStream_t s, int ld) {
    if (ld == 0) ld = C;
    if (C > 4096 || C < 1 || ld < C) return cudaErrorInvalidValue;
    const int threads = C < 512 ? (C + 31) / 32 * 32 : 512;
    B2C_LAUNCH(head_fwd_kernel, N, threads, C * sizeof(float), s, static_cast<const __nv_bfloat16*>(act), hw, C, ld, W, bias, K, pooled,
                                                    logits);
    return cudaGetLastError();
}

cudaError_t launch_ce_loss(const float* logits, const long long* target, int N, int K, float* dlogits, float* loss,
                           int* correct, cudaStream_t s) {
    B2C_LAUNCH(ce_loss_kernel, 1, 1024, 0, s, logits, target, N, K, dlogits, loss, correct);
    return cudaGetLastError();
}

cudaError_t launch_head_bwd(const float* dlogits, const float* pooled, const float* W, int N, int hw, int C, int K,
                            void* dact, float* dW, float* db, cudaStream_t s, int ld) {
    if (ld == 0) ld = C;
    if (C > 4096 || C < 1 || ld < C) return cudaErrorInvalidValue;
    const int threads = C < 512 ? (C + 31) / 32 * 32 : 512;
    B2C_LAUNCH(head_bwd_act_kernel, N, threads, 0, s, dlogits, W, K, hw, C, ld, static_cast<__nv_bfloat16*>(dact));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    dim3 grid((C + 31) / 32, K);
    B2C_LAUNCH(head_bwd_param_kernel, grid, 1024, 0, s, dlogits, pooled, N, K, C, dW, db);
    return cudaGetLastError();
}

static AugConst make_aug(const float* mean, const float* stdv) {
    AugConst k;
    for (int i = 0; i < 3; ++i) { k.mean[i] = mean[i]; k.std[i] = stdv[i]; }
    return k;
}

cudaError_t launch_augment_nchw(const uint8_t* data, const long long* idx, const int8_t* ijf, float* out, int N,
                                const float* mean, const float* stdv, cudaStream_t s) {
    if (N == 0) return cudaSuccess;
    B2C_LAUNCH(augment_nchw_kernel, N, 256, 0, s, data, idx, ijf, out, make_aug(mean, stdv));
    return cudaGetLastError();
}

cudaError_t launch_augment_im2col(const uint8_t* data, const long long* idx, const int8_t* ijf, void* out, int N,
                                  const float* mean, const float* stdv, cudaStream_t s, int taps) {
    if (N == 0) return cudaSuccess;
    if (taps == 1) {
        B2C_LAUNCH(augment_im2col_kernel<1>, N, 256, 0, s, data, idx, ijf, static_cast<uint4*>(out), make_aug(mean, stdv));
    } else {
        B2C_LAUNCH(augment_im2col_kernel<9>, N, 256, 0, s, data, idx, ijf, static_cast<uint4*>(out), make_aug(mean, stdv));
    }
    return cudaGetLastError();
}

cudaError_t launch_stem_im2col(const float* x, void* out, int N, cudaStream_t s, int taps) {
    if (N == 0) return cudaSuccess;
    if (taps == 1) {
        B2C_LAUNCH(stem_im2col_kernel<1>, N, 256, 0, s, x, static_cast<uint4*>(out));
    } else {
        B2C_LAUNCH(stem_im2col_kernel<9>, N, 256, 0, s, x, static_cast<uint4*>(out));
    }
    return cudaGetLastError();
}

cudaError_t launch_grouped_wgrad_extract(const float* dense, float* dw, int cout, int cin, int taps, int groups,
                                         cudaStream_t s) {
    const long long n = static_cast<long long>(taps) * cin * (cout / groups);
    if (n == 0) return cudaSuccess;
    B2C_LAUNCH(grouped_wgrad_extract_kernel, grid_for(n, 256, 148 * 8), 256, 0, s, dense, dw, cout, cin, taps, groups);
    return cudaGetLastError();
}

cudaError_t launch_sgd(float* p, const float* g, float* buf, long long n, const float* hyper, cudaStream_t s) {
    if (n & 3) return cudaErrorInvalidValue;
    const long long nvec = n >> 2;
    B2C_LAUNCH(sgd_kernel, grid_for(nvec, 256, 148 * 8), 256, 0, s, reinterpret_cast<float4*>(p),
                                                          reinterpret_cast<const float4*>(g),
                                                          reinterpret_cast<float4*>(buf), nvec, hyper);
    return cudaGetLastError();
}

cudaError_t launch_repack(const float* master, void* wf, void* wd, const RepackEntry* table_dev, int entries,
                          cudaStream_t s) {
    if (entries == 0) return cudaSuccess;
    // the table lives on the device, so the tile total is not known here: a fixed grid of 8 blocks per SM strides over
    // the global tile list (ResNet-18: 10.9 k tiles)
    if (entries > 1024) return cudaErrorInvalidValue;      // kMaxEntries of the kernel's shared table index
    B2C_LAUNCH(repack_kernel, 148 * 8, 256, 0, s, master, static_cast<__nv_bfloat16*>(wf), static_cast<__nv_bfloat16*>(wd),
                                       table_dev, entries);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Allocation-free BatchNorm-backward statistics (b2c_bn_bwd_stats_noalloc in include/b2c.h).
//
// Same two kernels as launch_bn_bwd_reduce + launch_bn_bwd_finalize, but the partial-row scratch between
// them is not a caller buffer: it is the fixed array below, which lives in every GPU that loads this
// library (2.3 MiB each) for the life of the process. That serves callers that must not, or cannot, call
// cudaMalloc around the launch: a host application whose device memory is entirely owned by another
// allocator or a fixed budget, a real-time loop where cudaMalloc's implicit device synchronisation and
// unbounded latency are not acceptable, or a plug-in that is only handed input / output pointers.
//
// kNoAllocMaxChannels bounds the channel count; rows are unbounded (the reduce kernel's grid is capped at
// 592 blocks, one partial row each). Every element the finalize kernel reads is written by the reduce
// kernel of the SAME call, so nothing in the array has to start at zero and stale contents from the
// previous call are never observed. Two calls that overlap on the GPU (different streams, or graphs
// replayed concurrently) share the array and race; callers serialise them.
// ------------------------------------------------------------------------------------------------
constexpr int kNoAllocMaxChannels = 512;
constexpr int kNoAllocMaxRows = kBnBwdMaxPartialRows;
static __device__ __align__(256) float g_noalloc_partial[kNoAllocMaxRows * 2 * kNoAllocMaxChannels];

int bn_bwd_noalloc_max_channels() { return kNoAllocMaxChannels; }

// In host code `g_noalloc_partial` is only a handle for the runtime, not a device address: the address is
// looked up once per device, with that device current, and cached. The first call on a device performs
// this look-up (it may initialise the module on that device), so it should happen before any stream
// capture begins.
static cudaError_t noalloc_scratch(float** out) {
    constexpr int kMaxDevices = 64;
    static float* cache[kMaxDevices] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    float* p = __atomic_load_n(&cache[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_partial);
        if (e != cudaSuccess) return e;
        if (p == nullptr) return cudaErrorInvalidSymbol;
        __atomic_store_n(&cache[dev], p, __ATOMIC_RELEASE);   // same value from every thread: benign if repeated
    }
    *out = p;
    return cudaSuccess;
}

cudaError_t launch_bn_bwd_stats_noalloc(const void* dout, const void* act, const void* y, void* dz_out,
                                        long long rows, int C, float count, const float* gamma, const float* mean,
                                        const float* rstd, float* dgamma, float* dbeta, float* coefA, float* coefB,
                                        float* coefC, cudaStream_t s) {
    if (C > kNoAllocMaxChannels) return cudaErrorInvalidValue;   // api layer reports B2C_ERR_TOO_LARGE first
    float* scratch = nullptr;
    cudaError_t e = noalloc_scratch(&scratch);
    if (e != cudaSuccess) return e;
    const int prow = bn_bwd_partial_rows(rows, C);               // capped at kBnBwdMaxPartialRows == kNoAllocMaxRows
    if (prow > kNoAllocMaxRows) return cudaErrorInvalidValue;    // never with the shared constant; guards a future edit
    e = launch_bn_bwd_reduce(dout, act, y, dz_out, rows, C, scratch, nullptr, nullptr, nullptr, s, 0);
    if (e != cudaSuccess) return e;
    return launch_bn_bwd_finalize(scratch, prow, C, count, gamma, mean, rstd, dgamma, dbeta, coefA, coefB, coefC, s);
}

}  // namespace b2c
