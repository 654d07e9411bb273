// b2c_kernels.h — internal (C++) launch interface between the C-ABI layer (api.cu) and the
// kernel translation units. Not part of the public boundary; see include/b2c.h for that.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace b2c {

constexpr int kMaxTaps = 12;

// One "tap GEMM": D[pixel, cout] = sum_t A_t[pixel + shift_t, :] . B_t[cout, :]^T
// covers conv fprop (3x3 / 1x1, stride 1 / 2) and dgrad (stride 1, and stride 2 by output phase).
struct TapGemmParams {
    int m_total;      // valid GEMM rows (pixels of the GEMM grid = n * gH * gW)
    int num_m_tiles;  // ceil(m_total / 128)
    int num_n_tiles;  // cout / BN
    int k_chunks;     // 64-channel chunks of the reduction per tap
    int num_taps;
    int gH, gW;       // pixel grid of the GEMM rows (j fastest, then i, then image)
    int in_stride;    // A coordinate = grid coordinate * in_stride + tap offset
    int8_t tap_dh[kMaxTaps];
    int8_t tap_dw[kMaxTaps];
    uint8_t tap_a[kMaxTaps];   // which A tensor map the tap reads (0..3; the CTA-pair kernels take 0 or 1)
    int tap_brow[kMaxTaps];    // first row of the tap's weight slab in the B matrix
    uint8_t tap_kc[kMaxTaps];  // 64-channel chunks of THIS tap's reduction; 0 = k_chunks (single-CTA kernel only: the
                               // 1x1 conv over a concatenation of tensors of different widths)
    // Output phases (stride-2 dgrad): a tile belongs to one phase = a sub-range of the tap list and an output
    // pixel offset. num_phases == 1 for everything else (phase_begin = {0, num_taps}).
    int num_phases;
    uint8_t phase_begin[5];
    uint8_t phase_py[4], phase_px[4];
    int a_split_w;             // A maps are 5-D with the W axis split into (w/2, parity): stride-2 fprop
    // epilogue
    __nv_bfloat16* out;
    const __nv_bfloat16* add_src;  // optional, same addressing as out
    int out_H, out_W, out_ld;      // output tensor pixel dims and channel pitch
    int sy, sx;                    // output pixel = (i * sy + phase_py, j * sx + phase_px)
    float* part_sum;               // optional [stat_rows][cout] per-CTA column sums (stat_rows <= 148)
    float* part_sq;                // optional [stat_rows][cout] per-CTA column sums of squares
    int cout;
    // Fused BatchNorm backward (dgrad feeding relu(bn(y))): when bwd_y is set the epilogue masks the result with
    // fma(y, bwd_scale, bwd_shift) > 0 before the store and the two statistics become sum(dz), sum(dz * y).
    const __nv_bfloat16* bwd_y;
    const float* bwd_scale;
    const float* bwd_shift;
    int stat_ld;                   // row pitch of part_sum / part_sq (floats)
};

// wgrad: dW[cout, cin, tap] += sum_pixels dY[pixel, cout] * X[pixel + shift_tap, cin]
// D rows = two 64-channel (tap, cin-block) slabs, D cols = cout tile, reduction = pixels.
struct WgradParams {
    int k_pixels;         // pixels per k-step: 64 or 128
    int k_chunks_total;   // ceil(pixels / k_pixels)
    int splits;           // split-K factor; CTA handles chunks [s*per, min((s+1)*per, total))
    int chunks_per_split;
    int num_m_blocks;     // ceil(num_slabs / 2)
    int num_n_tiles;      // cout / BN
    int num_slabs;        // taps * cin/64
    int cin_blocks;       // cin / 64
    int gH, gW;           // dY pixel grid
    int in_stride;
    int debug;            // measurement switches (tools/ only): 8 plain stores instead of atomics, 16 no stores
    int x_split_w;        // X map is 5-D with the W axis split into (parity, w/2): stride-2 convs
    int8_t tap_dh[kMaxTaps];
    int8_t tap_dw[kMaxTaps];
    int num_taps;
    int cin, cout;
    int cin_valid;        // real input channels (stem: 27 of 64)
    int cout_valid;       // real output channels (channel-padded layouts)
    int kflat_cin;        // >0: the cin axis is k = tap*kflat_cin + c (im2col'ed stem); address from (c, tap)
    float* dw;            // fp32 gradient, element (o, c, t) at o*o_stride + c*c_stride + t*t_stride
    long long o_stride, c_stride, t_stride;
};

// 3x3 stride-1 conv (fprop / dgrad) with activation-halo reuse, see conv_halo.cu.
struct HaloParams {
    int N, H, W;          // images, height, width (input dims == output dims)
    int num_m_tiles, num_n_tiles, k_chunks;
    int mode;             // 0: 16x8 tile of one image (W >= 16); 1: two whole 8x8 images (W == 8)
    int tiles_h, tiles_w; // mode 0: tiles per image
    int flip;             // dgrad: window position (a, b) uses weight tap (2-a, 2-b)
    int b_rows_per_tap;   // rows of the weight matrix per tap (= GEMM N total)
    int b_resident;       // set by the launcher: all weight slabs fit the ring and are loaded once per CTA
    int debug;            // measurement switches (tools/ only): 2 no stores, 4 no stats, 8 no MMA issue, 16 no TMEM read
    __nv_bfloat16* out;
    const __nv_bfloat16* add_src;
    int out_ld;
    float* part_sum;
    float* part_sq;
    int cout;
    // Fused BatchNorm backward (dgrad feeding relu(bn(y))): when bwd_y is set the epilogue masks the result with
    // fma(y, bwd_scale, bwd_shift) > 0 before the store and the two statistics become sum(dz), sum(dz * y).
    const __nv_bfloat16* bwd_y;
    const float* bwd_scale;
    const float* bwd_shift;
    int stat_ld;                   // row pitch of part_sum / part_sq (floats)
};

cudaError_t launch_conv_halo(int bn, const CUtensorMap& a, const CUtensorMap& b, const HaloParams& p, int max_ctas,
                             int* stat_rows, cudaStream_t stream);
cudaError_t launch_tap_gemm(int bn, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                            const TapGemmParams& p, int max_ctas, int* stat_rows, cudaStream_t stream);
// up to four A tensor maps (tap_a[t] in 0..3): the 1x1 conv over a concatenation of up to four tensors
cudaError_t launch_tap_gemm4(int bn, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& a2,
                             const CUtensorMap& a3, const CUtensorMap& b, const TapGemmParams& p, int max_ctas,
                             int* stat_rows, cudaStream_t stream);
cudaError_t launch_conv_halo2(int bn, const CUtensorMap& a, const CUtensorMap& b, const HaloParams& p, int max_ctas,
                              int* stat_rows, cudaStream_t stream);   // CTA-pair version (conv_halo2.cu), b box = bn/2 rows
// CTA-pair (tcgen05.mma.cta_group::2) version of launch_tap_gemm; b has a box of bn/2 rows (conv_gemm2.cu)
cudaError_t launch_tap_gemm2(int bn, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                             const TapGemmParams& p, int max_ctas, int* stat_rows, cudaStream_t stream);
// 3x3 stride-1 wgrad with halo reuse (conv_wgrad_halo.cu)
struct WgradHaloParams {
    int cin_slabs, cout_slabs;   // 64-channel slabs
    int splits, tiles_per_split; // split-K over 16x8-pixel tiles
    int num_tiles, tiles_h, tiles_w;
    float* partial;              // [cin_slabs*cout_slabs][splits][5][128][64] fp32
    int tco;                     // dW layout: 0 = OIHW, 1 = [tap][cin][cout]
};
// 3x3 stride-1 wgrad, one kernel row of taps per CTA (conv_wgrad_row.cu); dW in [tap][cin][cout] order
struct WgradRowParams {
    int cin_pairs, cout_slabs;   // 128-channel units
    int splits, steps_per_split; // split-K over 64-pixel k-steps
    int num_steps;
    int cin, cout;
    float* dw;
};
cudaError_t launch_wgrad_row(const CUtensorMap& x, const CUtensorMap& dy, const WgradRowParams& p, int w16,
                             cudaStream_t stream);
// CTA-pair version (conv_wgrad_row2.cu); the dY map's box holds ONE 64-channel slab
cudaError_t launch_wgrad_row2(const CUtensorMap& x, const CUtensorMap& dy, const WgradRowParams& p, int w16,
                              cudaStream_t stream);
size_t wgrad_halo_scratch_floats(int cin, int cout, int splits);
cudaError_t launch_wgrad_halo(const CUtensorMap& x, const CUtensorMap& dy, const WgradHaloParams& p, float* dw,
                              cudaStream_t stream);
cudaError_t launch_wgrad(int bn, const CUtensorMap& x, const CUtensorMap& dy, const WgradParams& p,
                         cudaStream_t stream);

}  // namespace b2c
