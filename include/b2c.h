/* b2c.h — C ABI of libb2c.so, the B200-native kernels behind the training hot path of
 * aqualovers/pytorch-cifar (conv + BatchNorm + ReLU forward/backward, input augmentation, SGD).
 *
 * The reference has no FFI of its own: every arithmetic step on this path is a call from
 * models/*.py / main.py into torch / torchvision. Each entry point below therefore
 * cites the reference call site (file:line under the reference tree) whose library kernel it replaces.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is DEVICE memory unless it says "host";
 *   - the caller owns every buffer, including scratch ("part_*", "partial"); nothing here allocates;
 *   - `stream` is a cudaStream_t passed as void*; all work is asynchronous on it;
 *   - activations are NHWC bf16; statistics, master weights, gradients of weights are fp32;
 *   - return 0 on success; a negative value otherwise: -(cudaError_t) for CUDA runtime errors,
 *     B2C_ERR_* below for argument / capability errors. Nothing throws, nothing falls back to a CPU or
 *     library path: an unsupported shape is an error.
 */
#ifndef B2C_H_
#define B2C_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B2C_API __attribute__((visibility("default")))
#else
#define B2C_API
#endif

#define B2C_ERR_UNSUPPORTED (-10001) /* shape / option outside what the kernels implement */
#define B2C_ERR_DRIVER (-10002)      /* cuTensorMapEncodeTiled unavailable or failed */
#define B2C_ERR_ARG (-10003)         /* null / misaligned / inconsistent argument */
#define B2C_ERR_TOO_LARGE (-10004)   /* size beyond the built-in scratch of a *_noalloc entry point */

B2C_API int b2c_version(void);
/* Human-readable text for a return code of this library (static storage). */
B2C_API const char* b2c_error_string(int code);

/* One convolution as the reference constructs it: nn.Conv2d(cin, cout, ksize, stride, padding=ksize/2,
 * bias=False), models/resnet.py:23-27 (3x3) and :33-34 (1x1 shortcut). */
typedef struct B2cConvDesc {
    int n, h, w;     /* input batch / height / width; tensor is [n, h, w, cin] bf16 */
    int cin, cout;   /* channel counts as laid out in memory, multiples of 64 */
    int ksize;       /* 1 or 3 */
    int stride;      /* 1 or 2 */
    int cin_valid;   /* wgrad only: real input channels (<= cin); 0 means cin */
    int cout_valid;  /* wgrad only: real output channels (<= cout); 0 means cout */
    int kflat_cin;   /* wgrad only: >0 for the im2col'ed stem, whose `cin` axis is k = tap*kflat_cin + c */
    int block_n;     /* GEMM N tile (64/128/256); 0 = choose */
    int max_ctas;    /* persistent grid cap; 0 = 148 (one CTA per SM) */
    int flags;       /* bit 0: force the per-tap kernels where a halo-reuse kernel would be chosen;
                        bit 1: wgrad: take the halo-reuse kernel for any qualifying shape, not just 64x64 channels;
                        bit 2: stride 2: address x with a TMA traversal stride instead of the split-W map
                               (measurement aid: same result, twice the L2->SM operand bytes);
                        bits 3,4: wgrad measurement switches (plain stores / no stores instead of reductions);
                        bit 5: wgrad: dw is laid out [tap][cin_valid][cout_valid] instead of OIHW, which lets the
                               epilogue reduce 4 rows x 128 contiguous bytes per instruction;
                        bit 6: wgrad: keep the per-tap kernel where the row kernel (conv_wgrad_row.cu) would be
                               chosen (measurement aid);
                        bit 7: fprop / dgrad (stride 2 and the fused shortcut form included): CTA-pair kernels
                               (tcgen05.mma.cta_group::2): conv_halo2.cu where
                               the halo kernel would run, conv_gemm2.cu otherwise (or with bit 0); wgrad: the pair
                               row kernel (conv_wgrad_row2.cu) where cin is a multiple of 256;
                        bit 8: wgrad: the pair row kernel for any cin the row kernel takes (test aid);
                        bit 9: fprop / dgrad_bn: the per-channel statistics are ACCUMULATED (fp32 atomics) into ONE row
                               — part_sum[cout], part_sq[cout], resp. partial[2][cin] — that the caller zero-filled,
                               instead of one row per CTA; *stat_rows = 1. The consumers are b2c_bn_act_apply_stats /
                               b2c_bn_bwd_apply_stats, which finalize the statistics themselves (no finalize launch).
                               Sums of ~148 per-CTA partials in arrival order: run-to-run differences of a few ulp. */
    int x_ld;        /* channel pitch (elements, multiple of 8) of the input-side tensors x / dx; 0 = cin */
    int y_ld;        /* channel pitch of the output-side tensors y / dy / the fprop's add_src; 0 = cout. With a pitch
                        wider than the channel count the pointers address a CHANNEL SLICE of a wider tensor: one group of
                        nn.Conv2d(C, C, 3, groups=g) (models/resnext.py:19) is the conv {cin = cout = C/g, x_ld = y_ld = C}
                        on x + j*C/g, y + j*C/g. Per-CTA statistic rows keep pitch cout (use flags bit 9 and offset
                        part_sum / part_sq to accumulate a slice of a wider BatchNorm); not for b2c_conv_dgrad_bn,
                        b2c_conv1x1_cat*_fprop or a 1x1 stride-2 dgrad without add_src. */
} B2cConvDesc;

/* y[n,ho,wo,cout] = conv(x, w). Replaces nn.Conv2d.forward -> aten::convolution (cuDNN)
 * at models/resnet.py:47-48,102-103,132. `w_fprop` is bf16 [tap][cout][cin] (see b2c_repack_weights).
 * add_src (optional, bf16, y's shape) is added before the store: the pre-activation residual
 * `out += shortcut` of models/preact_resnet.py:33. If part_sum/part_sq are non-null the epilogue also writes
 * per-CTA per-channel sums of y and y^2 (after the add; fp32, rows*cout each, rows <= 148 returned in
 * *stat_rows (host int)) for b2c_bn_finalize. */
B2C_API int b2c_conv_fprop(const B2cConvDesc* d, const void* x, const void* w_fprop, void* y, const void* add_src,
                   float* part_sum, float* part_sq, int* stat_rows, void* stream);

/* y = conv1x1(cat([x0, x1], channels), w): the `Root` of the reference's DLA trees, `torch.cat(xs, 1)` followed by a
 * 1x1 nn.Conv2d(2*cin, cout), models/dla_simple.py:44-56 — without materialising the concatenation: the two sources
 * are read through their own tensor maps as the two halves of the GEMM's K axis. d describes ONE source (cin channels
 * each, ksize 1, stride 1); `w_fprop2` is bf16 [2][cout][cin] (slab t multiplies x_t), which is what
 * b2c_repack_weights writes for a [tap=2][cin][cout] master. Statistics as in b2c_conv_fprop. The gradients need no
 * entry point of their own: dx_t = b2c_conv_dgrad with slab t of the [2][cin][cout] operand, dW slab t =
 * b2c_conv_wgrad(x_t, dy). */
B2C_API int b2c_conv1x1_cat_fprop(const B2cConvDesc* d, const void* x0, const void* x1, const void* w_fprop2, void* y,
                          float* part_sum, float* part_sq, int* stat_rows, void* stream);

/* The same over 2..4 sources (xs: HOST array of n_src device pointers): `Root((level + 2) * out_channels, ...)` of
 * a level-2 tree, models/dla.py:61,72-84, concatenates four tensors. w_fprop_n is bf16 [n_src][cout][cin]. */
B2C_API int b2c_conv1x1_catn_fprop(const B2cConvDesc* d, const void* const* xs, int n_src, const void* w_fprop_n, void* y,
                           float* part_sum, float* part_sq, int* stat_rows, void* stream);

/* The same over sources of DIFFERENT widths (cins: HOST array of n_src channel counts, each a multiple of 64 and at
 * most d->cin): every 1x1 conv of an Inception module reads torch.cat([y1, y2, y3, y4], 1) of the previous one,
 * models/googlenet.py:52-57. w_fprop_n is bf16 [n_src][cout][d->cin]: slab t holds source t's weights in its first
 * cins[t] columns. */
B2C_API int b2c_conv1x1_catr_fprop(const B2cConvDesc* d, const void* const* xs, const int* cins, int n_src,
                           const void* w_fprop_n, void* y, float* part_sum, float* part_sq, int* stat_rows,
                           void* stream);

/* dx[n,h,w,cin] = conv_transpose(dy, w) (+ add_src). Replaces cudnn_convolution_backward_input reached
 * from loss.backward(), main.py:104. `w_dgrad` is bf16 [tap][cin][cout]. add_src (optional, bf16, dx's
 * shape) is added in the epilogue; for ksize==1 && stride==2 it must be NULL (dx is zero-filled first)
 * or equal to dx (accumulate in place). */
B2C_API int b2c_conv_dgrad(const B2cConvDesc* d, const void* dy, const void* w_dgrad, void* dx, const void* add_src,
                   void* stream);

/* b2c_conv_dgrad whose result is the gradient w.r.t. relu(bn(y_bn)) — `out = F.relu(self.bn1(self.conv1(x)))`,
 * models/resnet.py:47 — with that ReLU/BatchNorm backward's first pass fused into the epilogue:
 *   dz = (conv_transpose(dy, w) [+ add_src]) * [fma(y_bn, scale, shift) > 0]      (written to `dz`, bf16)
 *   partial[r][0][c] = sum over the pixels CTA r handled of dz,  partial[r][1][c] = ... of dz * y_bn   (fp32)
 * *stat_rows (host int) receives the number of rows r (<= 148); pass `partial` and it to b2c_bn_bwd_finalize, then
 * b2c_bn_bwd_apply(dz, y_bn, ...). Replaces b2c_conv_dgrad + b2c_bn_bwd_reduce: one full read of dz and y less.
 * scale/shift are the forward b2c_bn_finalize outputs of that BatchNorm. Not for ksize==1 && stride==2. */
B2C_API int b2c_conv_dgrad_bn(const B2cConvDesc* d, const void* dy, const void* w_dgrad, void* dz, const void* add_src,
                      const void* y_bn, const float* scale, const float* shift, float* partial, int* stat_rows,
                      void* stream);

/* dx = dgrad3x3s2(dy, w[0..8]) + dgrad1x1s2(dy_shortcut, w[9]) (+ add_src) in one launch: the backward of a
 * down-sampling block's fan-out `conv1(x)` / `shortcut(x)`, models/resnet.py:23,33-34,47-49 and
 * models/preact_resnet.py:27-31 (autograd sums the two input gradients, main.py:104). d describes the 3x3
 * stride-2 conv; the 1x1 stride-2 shortcut conv has the same cin/cout. `w_dgrad10` is bf16 [10][cin][cout]: the
 * 3x3's nine tap slabs followed by the shortcut's single slab. */
B2C_API int b2c_conv_dgrad_with_shortcut(const B2cConvDesc* d, const void* dy, const void* dy_shortcut,
                                 const void* w_dgrad10, void* dx, const void* add_src, void* stream);

/* dw[cout_valid,cin_valid,k,k] (fp32, OIHW = the reference's .grad layout) += sum over pixels of x * dy.
 * Replaces cudnn_convolution_backward_weight, main.py:104. Accumulates with fp32 atomics: the caller
 * zero-fills dw (optimizer.zero_grad(), main.py:101). */
B2C_API int b2c_conv_wgrad(const B2cConvDesc* d, const void* x, const void* dy, float* dw, float* scratch,
                   long long scratch_floats, void* stream);
/* scratch (optional): with at least 40960 floats (ideally 148*40960 = 24 MB) 3x3 stride-1 convs on 16- or
 * 32-pixel images take the halo-reuse kernel, which writes split-K partial tiles there and sums them into dw
 * with a second small kernel (no atomics, run-to-run deterministic). Without it, or for other shapes, split-K
 * partials are combined with fp32 atomics directly in dw. */

/* Depthwise 3x3 convolution, nn.Conv2d(C, C, 3, stride, padding=1, groups=C, bias=False), models/mobilenetv2.py:20,
 * and its two gradients (main.py:104). x [n,h,w,channels] bf16, weight = the fp32 master tensor [c_valid,1,3,3];
 * channels is the padded layout width (multiple of 8), channels >= c_valid are treated as zero weights.
 * fprop optionally writes BatchNorm partial sums like b2c_conv_fprop, one row per block:
 * b2c_dwconv_stat_rows(out_pixels, channels) rows (<= 592). wgrad accumulates with fp32 atomics into dw. */
B2C_API int b2c_dwconv_stat_rows(long long out_pixels, int channels);
B2C_API int b2c_dwconv_fprop(const void* x, const float* weight, void* y, int n, int h, int w, int channels, int c_valid,
                     int stride, float* part_sum, float* part_sq, int* stat_rows, void* stream);
/* b2c_dwconv_fprop whose BatchNorm sums are accumulated (fp32 atomics) into ONE zero-filled row sum[channels] /
 * sumsq[channels], for b2c_bn_act_apply_stats. */
B2C_API int b2c_dwconv_fprop_acc(const void* x, const float* weight, void* y, int n, int h, int w, int channels, int c_valid,
                         int stride, float* sum, float* sumsq, void* stream);
/* The depthwise conv fed straight from the conv before it: `out = F.relu(self.bn1(self.conv1(x)))` followed by
 * `self.conv2` (groups = planes), models/mobilenetv2.py:32-33. y_prev is conv1's RAW output; the kernel applies that
 * BatchNorm + ReLU while loading, x = relu(y_prev * in_scale[c] + in_shift[c]) (scale / shift from b2c_bn_finalize), and
 * pads with zeros AFTER the transform, so the activation between the two convs is never written or re-read — the BN
 * apply lives in its consumer. Stride 1 only, shapes for which b2c_dwconv_input_transform_ok returns 1 (else
 * B2C_ERR_UNSUPPORTED: use b2c_bn_act_apply + b2c_dwconv_fprop). accumulate != 0: statistics as in b2c_dwconv_fprop_acc.
 * The matching weight gradient reads the same y_prev the same way; the data gradient needs no variant. */
B2C_API int b2c_dwconv_input_transform_ok(int h, int w, int channels, int stride);
B2C_API int b2c_dwconv_fprop_bnrelu(const void* y_prev, const float* in_scale, const float* in_shift, const float* weight,
                            void* y, int n, int h, int w, int channels, int c_valid, float* part_sum, float* part_sq,
                            int* stat_rows, int accumulate, void* stream);
B2C_API int b2c_dwconv_wgrad_bnrelu(const void* y_prev, const float* in_scale, const float* in_shift, const void* dy,
                            float* dweight, int n, int h, int w, int channels, int c_valid, void* stream);
B2C_API int b2c_dwconv_dgrad(const void* dy, const float* weight, void* dx, int n, int h, int w, int channels, int c_valid,
                     int stride, void* stream);
B2C_API int b2c_dwconv_wgrad(const void* x, const void* dy, float* dweight, int n, int h, int w, int channels, int c_valid,
                     int stride, void* stream);

/* Training-mode BatchNorm statistics -> per-channel scale/shift, saved mean/rstd, running-stat update.
 * Replaces the statistics half of F.batch_norm for nn.BatchNorm2d, models/resnet.py:25,28,35,104.
 * running_mean/var may be NULL (no update). count = n*h*w. conv_bias (optional, [channels]): bias of the
 * preceding conv (models/vgg.py:33). The kernels never add it to the stored conv output — BatchNorm cancels
 * it — but the reference's running_mean tracks mean(conv)+bias, so it is added there to keep checkpoints
 * interchangeable. */
B2C_API int b2c_bn_finalize(const float* part_sum, const float* part_sq, int rows, int channels, float count,
                    const float* gamma, const float* beta, const float* conv_bias, float eps, float momentum,
                    float* running_mean, float* running_var, float* scale, float* shift, float* save_mean,
                    float* save_rstd, void* stream);

/* Eval-mode BatchNorm: scale/shift from running statistics (net.eval(), main.py:118); conv_bias as above. */
B2C_API int b2c_bn_eval_scale_shift(int channels, const float* gamma, const float* beta, const float* conv_bias,
                            const float* running_mean, const float* running_var, float eps, float* scale,
                            float* shift, void* stream);

/* out = act(y*scale + shift + residual'), residual' = res (identity shortcut) or res*res_scale+res_shift
 * (conv shortcut's own BatchNorm). Replaces F.relu(bn(..)), `out += self.shortcut(x)`, F.relu(out),
 * models/resnet.py:47-50. rows = n*h*w; tensors are [rows, channels] bf16. relu_bits (optional,
 * rows*channels/8 bytes): bit e of byte i = [out[8*i + e] > 0], the ReLU mask for b2c_bn_bwd_reduce. */
B2C_API int b2c_bn_act_apply(const void* y, const float* scale, const float* shift, const void* res,
                     const float* res_scale, const float* res_shift, void* out, long long rows, int channels,
                     int relu, unsigned char* relu_bits, void* stream);

/* One training-mode BatchNorm whose batch statistics arrive as ONE accumulated row (B2cConvDesc.flags bit 9,
 * b2c_dwconv_fprop_acc): everything b2c_bn_finalize takes and produces. sum/sumsq/gamma/beta/conv_bias are inputs,
 * [channels] fp32; running_mean/var (both or neither) are updated in place; scale/shift/save_mean/save_rstd are
 * written for the backward pass. count = n*h*w. */
typedef struct B2cBnStats {
    const float* sum;
    const float* sumsq;
    const float* gamma;
    const float* beta;
    const float* conv_bias; /* optional, see b2c_bn_finalize */
    float* running_mean;
    float* running_var;
    float* scale;
    float* shift;
    float* save_mean;
    float* save_rstd;
    float count, eps, momentum;
} B2cBnStats;

/* b2c_bn_finalize + b2c_bn_act_apply in ONE launch — F.relu(self.bn1(self.conv1(x))) and
 * F.relu(self.bn2(..) + self.shortcut(x)), models/resnet.py:47-50, given the conv's accumulated statistics: every
 * block derives scale / shift for all channels in its prologue (2*channels floats from L2), block 0 publishes them
 * and updates the running statistics. res_bn (optional, needs res): the shortcut conv's own BatchNorm
 * (models/resnet.py:33-35), finalized the same way; res without res_bn is an identity shortcut. Saves the 20 + 20
 * finalize launches of a ResNet-18 step together with b2c_bn_bwd_apply_stats. channels <= 4096. */
B2C_API int b2c_bn_act_apply_stats(const void* y, const B2cBnStats* bn, const void* res, const B2cBnStats* res_bn,
                           void* out, long long rows, int channels, int relu, unsigned char* relu_bits, void* stream);

/* b2c_bn_bwd_reduce that ACCUMULATES (fp32 atomics) into one zero-filled row sums[2][channels] (sum dz | sum dz*y)
 * instead of writing one row per block; same masking options. For b2c_bn_bwd_apply_stats. */
B2C_API int b2c_bn_bwd_reduce_acc(const void* dout, const void* act, const void* y, void* dz_out, long long rows,
                          int channels, float* sums, const float* relu_scale, const float* relu_shift,
                          const unsigned char* relu_bits, void* stream);

/* b2c_bn_bwd_finalize + b2c_bn_bwd_apply in ONE launch, from one accumulated row sums[2][channels]: dgamma / dbeta
 * are written by block 0, dy = A*dz + B*y + C (+ add) by everyone (cudnn_batch_norm_backward, main.py:104). */
B2C_API int b2c_bn_bwd_apply_stats(const void* dz, const void* y, const float* sums, float count, const float* gamma,
                           const float* save_mean, const float* save_rstd, float* dgamma, float* dbeta,
                           const void* add, void* dy, long long rows, int channels, const float* relu_scale,
                           const float* relu_shift, void* stream);

/* Number of partial rows b2c_bn_bwd_reduce writes for (rows, channels): size `partial` as
 * rows_returned * 2 * channels floats. Host-only arithmetic. */
B2C_API int b2c_bn_bwd_partial_rows(long long rows, int channels);

/* BatchNorm(+ReLU) backward, reduction pass: per-channel sums of dz and dz*y with dz = dout*[act>0]
 * (act NULL: dz = dout). If dz_out is non-NULL the masked gradient is stored there (may alias dout).
 * Replaces the reductions of cudnn_batch_norm_backward + relu backward, main.py:104.
 * relu_scale/relu_shift (optional, with act == NULL): the unit is a plain F.relu(bn(y)) whose forward scale/shift
 * these are; the mask is then recomputed as fma(y, scale, shift) > 0 and neither act nor dz touch memory
 * (pass the same pair to b2c_bn_bwd_apply, which then takes dout in place of dz).
 * relu_bits (optional, instead of act and relu_scale/shift): the bit mask b2c_bn_act_apply wrote — an eighth of
 * a byte per element instead of re-reading the activation. */
B2C_API int b2c_bn_bwd_reduce(const void* dout, const void* act, const void* y, void* dz_out, long long rows,
                      int channels, float* partial, const float* relu_scale, const float* relu_shift,
                      const unsigned char* relu_bits, void* stream);

/* Combine partial rows: dgamma, dbeta (fp32, written) and the coefficients of dy = A*dz + B*y + C. */
B2C_API int b2c_bn_bwd_finalize(const float* partial, int partial_rows, int channels, float count, const float* gamma,
                        const float* save_mean, const float* save_rstd, float* dgamma, float* dbeta, float* coef_a,
                        float* coef_b, float* coef_c, void* stream);

/* (No counterpart in the reference: this entry point exists for the embedding application, not for parity.)
 * b2c_bn_bwd_reduce + b2c_bn_bwd_finalize in ONE call that allocates nothing and takes no scratch pointer:
 * the partial rows between the two kernels live in a fixed array inside the library (2.3 MiB, resident on
 * every GPU that loads libb2c.so until the process exits). Use it where a cudaMalloc is not an option around
 * the call: device memory fully owned by another allocator or a fixed budget, a latency-bound loop that
 * cannot absorb cudaMalloc's device synchronisation, or a plug-in that only ever receives input/output
 * pointers. Inputs and outputs are the caller's, exactly as for the two-call form, and the results are the
 * same. Limits and rules:
 *   - channels <= B2C_NOALLOC_MAX_CHANNELS, otherwise B2C_ERR_TOO_LARGE (rows are not limited);
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress, once per
 *     device, with that device current); make that first call before starting any stream capture;
 *   - all calls on a device share the one array. A call never reads what an earlier call left there
 *     (everything it reads it has written first, so nothing needs clearing), but two calls that overlap on
 *     the GPU (two streams, concurrent graph replays) race: serialise them. */
#define B2C_NOALLOC_MAX_CHANNELS 512
B2C_API int b2c_bn_bwd_stats_noalloc(const void* dout, const void* act, const void* y, void* dz_out, long long rows,
                             int channels, float count, const float* gamma, const float* save_mean,
                             const float* save_rstd, float* dgamma, float* dbeta, float* coef_a, float* coef_b,
                             float* coef_c, void* stream);

/* dy = A*dz + B*y + C per channel (+ add, optional bf16 tensor of dy's shape: the gradient that bypasses
 * the BatchNorm through an identity shortcut, models/preact_resnet.py:29-33). bf16 out. */
B2C_API int b2c_bn_bwd_apply(const void* dz, const void* y, const float* coef_a, const float* coef_b, const float* coef_c,
                     const void* add, void* dy, long long rows, int channels, const float* relu_scale,
                     const float* relu_shift, void* stream);

/* nn.MaxPool2d(kernel_size=2, stride=2), models/vgg.py:31, on NHWC bf16 [n,h,w,c] -> [n,h/2,w/2,c]. */
B2C_API int b2c_maxpool2x2_fwd(const void* x, void* y, int n, int h, int w, int channels, void* stream);
/* Its backward: dx[n,h,w,c] = dy of the window if x is the window's first maximum (torch's tie rule), else 0. */
B2C_API int b2c_maxpool2x2_bwd(const void* x, const void* y, const void* dy, void* dx, int n, int h, int w, int channels,
                       void* stream);

/* nn.MaxPool2d(3, stride, padding=1), models/googlenet.py:40 (stride 1, inside an Inception module) and :67 (stride 2,
 * between stages), on NHWC bf16 [n,h,w,c] -> [n,h/stride,w/stride,c]; argmax (uint8, y's shape) receives the window
 * position r*3+s of the first maximum in scan order (torch's tie rule) for the backward pass. */
B2C_API int b2c_maxpool3x3_fwd(const void* x, void* y, void* argmax, int n, int h, int w, int channels, int stride,
                       void* stream);
/* dx[n,h,w,c] = (add, optional: another consumer's gradient of x) + the sum of dy over the windows whose maximum is
 * this pixel. h, w are x's. */
B2C_API int b2c_maxpool3x3_bwd(const void* dy, const void* argmax, const void* add, void* dx, int n, int h, int w, int channels,
                       int stride, void* stream);
/* dst[row][dst_off + c] = src[row][src_off + c] for c < channels (bf16; pitches, offsets and channels multiples of
 * 8): torch.cat / its backward where a concatenation has to exist in memory — the input of GoogLeNet's
 * avgpool + linear head, models/googlenet.py:96-98. */
B2C_API int b2c_copy_channels(const void* src, int src_ld, int src_off, void* dst, int dst_ld, int dst_off, long long rows,
                      int channels, void* stream);

/* Head: F.avg_pool2d(out, 4) -> view -> nn.Linear, models/resnet.py:137-139. act is [n, hw, channels]
 * bf16; pooled [n, channels] and logits [n, classes] are fp32 outputs. */
B2C_API int b2c_head_fwd(const void* act, int n, int hw, int channels, const float* weight, const float* bias,
                 int classes, float* pooled, float* logits, void* stream);

/* The same head over an activation whose channel pitch act_ld exceeds the channel count (layouts padded to 64
 * channels): F.adaptive_avg_pool2d(out, (1, 1)) -> view -> nn.Linear(368, 10) of models/regnet.py:104-106. Any
 * channels <= 4096; pooled / weight / dweight keep the true width, channels >= `channels` of dact are left untouched. */
B2C_API int b2c_head_fwd_ld(const void* act, int n, int hw, int channels, int act_ld, const float* weight, const float* bias,
                    int classes, float* pooled, float* logits, void* stream);
B2C_API int b2c_head_bwd_ld(const float* dlogits, const float* pooled, const float* weight, int n, int hw, int channels,
                    int act_ld, int classes, void* dact, float* dweight, float* dbias, void* stream);

/* Squeeze-and-excitation, models/senet.py:67-74 (and models/regnet.py:21-25):
 *     w = F.avg_pool2d(out, out.size(2)); w = F.relu(fc1(w)); w = F.sigmoid(fc2(w)); out = out * w; out += shortcut
 * y, add (optional: the shortcut) and out are bf16 [n, hw, ld] with `channels` <= ld real channels (ld % 8 == 0; the
 * padded ones are zero and get scale 0); w1 [reduced][channels], b1 [reduced], w2 [channels][reduced], b2 [channels]
 * are the fp32 parameters of the two 1x1 convs. Outputs kept for the backward pass: pooled [n][ld], hidden [n][reduced]
 * (after the ReLU), scale [n][ld] (after the sigmoid), all fp32. Three launches: pool, per-image MLP, scale (+ add). */
B2C_API int b2c_se_fwd(const void* y, int n, int hw, int channels, int ld, int reduced, const float* w1, const float* b1,
               const float* w2, const float* b2, float* pooled, float* hidden, float* scale, const void* add, void* out,
               void* stream);
/* Its backward, given dout = d loss / d out: dy = dout * scale + (gradient through the pool) (bf16 [n, hw, ld]);
 * dw1 / db1 / dw2 / db2 are ACCUMULATED (the caller zero-fills them with the other gradients). The gradient of the
 * shortcut is dout itself. Scratch (caller-owned, fp32): ds [n][ld], dz1 [n][reduced], dp [n][ld]. */
B2C_API int b2c_se_bwd(const void* dout, const void* y, int n, int hw, int channels, int ld, int reduced, const float* w1,
               const float* w2, const float* pooled, const float* hidden, const float* scale, float* ds_scratch,
               float* dz1_scratch, float* dp_scratch, float* dw1, float* db1, float* dw2, float* db2, void* dy,
               void* stream);

/* nn.CrossEntropyLoss() (mean), main.py:86,103: loss[0], dlogits = (softmax - onehot)/n, and the number of
 * top-1 hits (main.py:108-110; may be NULL). target is int64. n <= 65536. */
B2C_API int b2c_ce_loss(const float* logits, const long long* target, int n, int classes, float* dlogits, float* loss,
                int* correct, void* stream);

/* Backward of the head given dlogits: dact (bf16 [n,hw,channels]), dweight [classes,channels], dbias. */
B2C_API int b2c_head_bwd(const float* dlogits, const float* pooled, const float* weight, int n, int hw, int channels,
                 int classes, void* dact, float* dweight, float* dbias, void* stream);

/* Input pipeline of main.py:30-35 on the GPU: RandomCrop(32, padding=4) + RandomHorizontalFlip + ToTensor
 * + Normalize over a uint8 dataset [*,32,32,3] resident in HBM. idx[n] (int64) selects the sample,
 * ijf[n*3..] (int8) = crop row i, crop col j (0..8), flip (0/1) drawn on the host exactly as torchvision
 * draws them. mean/std are HOST pointers to 3 floats. */
B2C_API int b2c_augment_nchw(const uint8_t* data, const long long* idx, const int8_t* ijf, float* out_nchw_f32, int n,
                     const float* mean_host, const float* std_host, void* stream);
/* Same pixels, written as the stem conv's GEMM operand: [n,32,32,64] bf16, k = (r*3+s)*3+c for k<27. */
B2C_API int b2c_augment_im2col(const uint8_t* data, const long long* idx, const int8_t* ijf, void* out_im2col_bf16,
                       int n, const float* mean_host, const float* std_host, void* stream);
/* fp32 NCHW network input (what the reference DataLoader yields, main.py:99) -> stem GEMM operand. */
B2C_API int b2c_stem_im2col(const float* x_nchw_f32, void* out_im2col_bf16, int n, void* stream);

/* The same two producers for a 1x1 stem conv (nn.Conv2d(3, 64, kernel_size=1), models/resnext.py:47): the operand is
 * the image itself, [n,32,32,64] bf16 with channel c < 3 = the pixel and zeros above. */
B2C_API int b2c_augment_nhwc64(const uint8_t* data, const long long* idx, const int8_t* ijf, void* out_nhwc64_bf16, int n,
                       const float* mean_host, const float* std_host, void* stream);
B2C_API int b2c_stem_nhwc64(const float* x_nchw_f32, void* out_nhwc64_bf16, int n, void* stream);

/* Weight gradient of a grouped conv — nn.Conv2d(C, C, 3, groups=g), models/resnext.py:19 — out of the dense one.
 * A grouped conv runs here as a dense conv over block-diagonal bf16 operands (b2c_repack_weights with one table entry
 * per group: the entry's operand offsets select the group's diagonal block, its cin_pad / cout_pad are the dense
 * pitches; everything off the diagonal stays zero), so b2c_conv_wgrad (flags bit 5) yields a dense
 * dense_tco[tap][cin][cout] fp32 matrix. This adds its diagonal blocks to the parameter's gradient in the reference's
 * layout: dw_oihw[o][ci][tap] += dense_tco[tap][(o / (cout/g)) * (cin/g) + ci][o]  (autograd of
 * aten::convolution with groups, main.py:104). The caller zero-fills dense_tco before the wgrad launch. */
B2C_API int b2c_grouped_wgrad_extract(const float* dense_tco, float* dw_oihw, int cout, int cin, int taps, int groups,
                              void* stream);

/* optim.SGD(...).step(), main.py:87-88,105, over one flat fp32 arena (n % 4 == 0, 16-byte aligned).
 * hyper is a DEVICE array {lr, momentum, weight_decay, grad_scale, first_step(0/1)}. */
B2C_API int b2c_sgd_step(float* param, const float* grad, float* momentum_buf, long long n, const float* hyper,
                 void* stream);

/* One conv weight tensor in the fp32 master arena and where its bf16 operand copies go. */
typedef struct B2cRepackEntry {
    long long src_off; /* element offset of the OIHW fp32 tensor */
    long long wf_off;  /* element offset of the fprop operand [tap][cout][cin_pad] */
    long long wd_off;  /* element offset of the dgrad operand [tap][cin][cout]; <0: none */
    int cout, cin, taps, cin_pad;
    int kflat;         /* 0: master OIHW; 1: stem, fprop operand is [cout][cin_pad], k = tap*cin + c;
                          2: master is [tap][cin][cout] (the layout flags bit 5 of b2c_conv_wgrad writes) */
    int pad_;          /* cout_pad: rows per tap of the fprop operand / row pitch of the dgrad operand; 0 = cout */
} B2cRepackEntry;

/* fp32 OIHW master weights (the reference's state_dict layout) -> bf16 GEMM operands, all convs in
 * one launch. `table` is a DEVICE array of `entries` records (at most 1024). */
B2C_API int b2c_repack_weights(const float* master, void* w_fprop_arena, void* w_dgrad_arena, const B2cRepackEntry* table,
                       int entries, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2C_H_ */
